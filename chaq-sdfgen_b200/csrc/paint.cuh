// paint.cuh -- pass 2c of the banded path: evaluation of the finished hull as plan + paint per 32-row block
// (paint_core.h). Replaces the fill loop of dist_transform_1d (reference openmp/df.c:83-96, incl. the sqrtf at
// df.c:95) for the inside and the outside field of openmp/sdfgen.c:277-289, and fuses transform_float_sub
// (sdfgen.c:98-106) and transform_float_to_byte (sdfgen.c:75-96) as an exact integer threshold look-up (tables.h).
//
//   hull_paint_kernel  CTA = kEmitWarps warps = one strip x one run of output rows, lane = column. Per block each
//                      lane first lists the owners of its 32 rows (data dependent, a few iterations), then all
//                      lanes paint the rows in step: owner from the list, parabola, in-column class distance, min,
//                      threshold look-up, byte store (a warp's 32 bytes are one full sector). Blocks in which no
//                      lane has anything within reach are written as 16-byte pieces.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "hull.cuh"
#include "paint_core.h"

namespace chaq {

constexpr int kPaintSegWords = 33;  // owners of one block per lane: at most one per row, plus the one carried in

__host__ __device__ inline size_t hull_paint_smem() { return kTableBytes + (size_t)kEmitWarps * kPaintSegWords * 128; }

template <bool WIDE>
struct HullPaintCtx {
    const uint32_t* hcol;    // hull word of block 0 of this column (stride wc per block)
    const uint32_t* clst;    // transposed class word of block 0 of this column (stride wc per block)
    const uint8_t* rec_lane; // g8 of buffer row 0 of this column
    const uint8_t* flag0;    // flag of block 0 of this column group (stride ncg per block)
    uint32_t* seg;           // owner list, already offset by lane (stride 32 words)
    unsigned rs32;
    int wc, ncg, nblk, lane;
    // epilogue: level tracking against the thresholds of tables.h
    const int32_t* thr;
    const int32_t* qthr;
    uint8_t* optr;
    long long ostride;
    bool store_ok;
    bool wide_ok;            // whole blocks may be written as 16-byte pieces (row pitch and base 16-byte aligned)
    int qcls, qm, qlo, qhi, qbyte, qsign;

    __device__ __forceinline__ uint32_t h_ld(int b) const { return b < nblk ? __ldg(hcol + (long long)b * wc) : 0u; }
    __device__ __forceinline__ uint32_t g_row(int r) const { return g_load<WIDE>(rec_lane + (unsigned)r * rs32); }
    __device__ __forceinline__ uint32_t cls_t(int b) const {
        const uint32_t w = __ldg(clst + (long long)(b < nblk ? b : nblk - 1) * wc);
        return b < nblk ? w : (uint32_t)((int32_t)w >> 31);  // past the buffer: the class of the last row (bit 31 of the last word)
    }
    __device__ __forceinline__ int q_flag(int blk) const {
        return (blk < 0 || blk >= nblk) ? 3 : (int)__ldg(flag0 + (long long)blk * ncg);
    }
    __device__ __forceinline__ void seg_st(int k, uint32_t v) { seg[k * 32] = v; }
    __device__ __forceinline__ uint32_t seg_ld(int k) const { return seg[k * 32]; }
    __device__ __forceinline__ bool any(bool b) const { return __any_sync(0xffffffffu, b); }
    __device__ __forceinline__ void prefetch_block(int b) const {
        if (b < nblk) {
            prefetch_l1(hcol + (long long)b * wc);
            prefetch_l1(clst + (long long)b * wc);
        }
    }
    __device__ __forceinline__ void prefetch_g(int r) const { prefetch_l1(rec_lane + (unsigned)r * rs32); }

    __device__ __forceinline__ void requant(int cls, int d2c) {
        if (cls != qcls) {
            qcls = cls;
            qthr = thr + cls * 260;
            qsign = cls ? 1 : -1;
            qm = level_of(qthr, d2c);
        } else {
            while (d2c >= qthr[qm + 1]) ++qm;
            while (d2c < qthr[qm]) --qm;
        }
        qlo = qthr[qm];
        qhi = qthr[qm + 1];
        qbyte = cls ? qm : 255 - qm;
    }
    // level of the row's clamped squared distance: the level of the row above, moved by at most one step in the
    // common case; anything else (class change, first row, small spreads) searches the table
    __device__ __forceinline__ void out(int cls, int d2c) {
        const int step = (d2c >= qhi ? 1 : 0) - (d2c < qlo ? 1 : 0);
        if (step != 0) {
            qm += step;
            qbyte += step * qsign;
            qlo = qthr[qm];
            qhi = qthr[qm + 1];
        }
        if (CQ_UNLIKELY(d2c >= qhi || d2c < qlo || cls != qcls)) requant(cls, d2c);
        if (store_ok) *optr = (uint8_t)qbyte;
        optr += ostride;
    }
    // n rows of one saturated value per column. Whole blocks go out as two 16-byte stores per lane.
    __device__ __forceinline__ void out_run(int cls, int d2c, int n) {
        out(cls, d2c);
        const uint32_t b = (uint32_t)qbyte & 0xFFu;
        if (n == 32 && wide_ok) {
            // word q of a row = bytes of lanes 4q..4q+3; lane l writes words 4*(l&1)..+3 of rows l>>1 and (l>>1)+16
            const uint32_t b01 = b | (__shfl_down_sync(0xffffffffu, b, 1) << 8);
            const uint32_t wq = b01 | (__shfl_down_sync(0xffffffffu, b01, 2) << 16);  // valid in lanes 4q
            const int h4 = (lane & 1) * 16;
            uint4 v;
            v.x = __shfl_sync(0xffffffffu, wq, h4);
            v.y = __shfl_sync(0xffffffffu, wq, h4 + 4);
            v.z = __shfl_sync(0xffffffffu, wq, h4 + 8);
            v.w = __shfl_sync(0xffffffffu, wq, h4 + 12);
            uint8_t* row0 = optr - ostride - lane;  // out() has advanced one row; group's first column
            uint8_t* dst = row0 + (long long)(lane >> 1) * ostride + h4;
            *reinterpret_cast<uint4*>(dst) = v;
            *reinterpret_cast<uint4*>(dst + 16 * ostride) = v;
            optr += 31 * ostride;
            return;
        }
        for (int i = 1; i < n; ++i) {
            if (store_ok) *optr = (uint8_t)b;
            optr += ostride;
        }
    }
};

template <bool WIDE>
__global__ void __launch_bounds__(kEmitWarps * 32, 8) hull_paint_kernel(HullParams p) {
    extern __shared__ __align__(16) uint8_t s_hull[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    int32_t* s_thr = reinterpret_cast<int32_t*>(s_hull);
    for (int i = threadIdx.x; i < 2 * 260; i += kEmitWarps * 32) s_thr[i] = (&p.table->thr[0][0])[i];
    __syncthreads();
    long long id = blockIdx.x;
    const int run = (int)(id % p.nruns_e); id /= p.nruns_e;
    const int strip = (int)(id % p.nstrips_e);
    const int img = (int)(id / p.nstrips_e);
    const int cg = p.cg0 + strip * kEmitWarps + warp;
    if (cg >= p.cg1) return;
    // runs are cut at multiples of 32 buffer rows so that whole blocks can be written at once
    const int bfirst = p.out_begin >> 5;
    int y0 = (bfirst + run * p.eb) * 32, y1 = y0 + p.eb * 32;
    y0 = y0 > p.out_begin ? y0 : p.out_begin;
    y1 = y1 < p.out_end ? y1 : p.out_end;
    if (y0 >= y1) return;
    HullPaintCtx<WIDE> c;
    const int x = cg * 32 + lane;
    c.hcol = p.hull + (long long)img * p.nblk * p.wc + x;
    c.clst = p.clsT + (long long)img * p.nblk * p.wc + x;
    c.rec_lane = p.inter + (long long)img * p.buf_rows * p.rs + x * (int)sizeof(typename GVal<WIDE>::T);
    c.flag0 = p.flags + (long long)img * p.nblk * p.ncg + cg;
    c.seg = reinterpret_cast<uint32_t*>(s_hull + kTableBytes + (size_t)warp * kPaintSegWords * 128) + lane;
    c.rs32 = (unsigned)p.rs; c.wc = p.wc; c.ncg = p.ncg; c.nblk = p.nblk; c.lane = lane;
    c.thr = s_thr; c.ostride = p.w; c.store_ok = x < p.w;
    const int out_rows = p.out_end - p.out_begin;
    c.optr = p.out + ((long long)img * out_rows + (y0 - p.out_begin)) * (long long)p.w + (x < p.w ? x : 0);
    // 16-byte stores need the whole group inside the image and every row of the group 16-byte aligned
    c.wide_ok = cg * 32 + 32 <= p.w && (p.w & 15) == 0 && (reinterpret_cast<uintptr_t>(p.out) & 15) == 0;
    c.qcls = -1; c.qm = 1; c.qlo = 0; c.qhi = 0x7fffffff; c.qbyte = 0; c.qsign = 1; c.qthr = s_thr;
    int lag = p.L;
    asm volatile("" : "+r"(lag));
    HullPaint<HullPaintCtx<WIDE>> hp(c, lag, p.c2);
    hp.run(y0, y1);
}

}  // namespace chaq
