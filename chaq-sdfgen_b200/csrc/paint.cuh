// paint.cuh -- pass 2c of the banded path as plan + paint per 32-row block (paint_core.h). Same inputs, outputs and
// grid as hull_emit_kernel (hull.cuh); replaces the row-by-row cursor walk of openmp/df.c:83-96.
//
//   hull_paint_kernel  CTA = kEmitWarps warps = one strip x one run of output rows, lane = column. Per block each
//                      lane first lists the owners of its 32 rows (data dependent, a few iterations), then all
//                      lanes paint the rows in step: owner from the list, parabola, in-column class distance, min,
//                      threshold look-up (tables.h), byte store (a warp's 32 bytes are one full sector).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "hull.cuh"
#include "paint_core.h"

namespace chaq {

constexpr int kPaintSegWords = 33;  // owners of one block per lane: at most one per row, plus the one carried in

__host__ __device__ inline size_t hull_paint_smem() {
    return kSweepTableBytes + (size_t)kEmitWarps * kPaintSegWords * 128;
}

struct HullPaintCtx : HullEmitCtx {
    uint32_t* seg;           // owner list, already offset by lane (stride 32 words)
    const uint32_t* clst;    // transposed class word of block 0 of this column (stride wc per block)
    bool wide_ok;            // whole blocks may be written as 16-byte pieces (row pitch and base 16-byte aligned)
    __device__ __forceinline__ void seg_st(int k, uint32_t v) { seg[k * 32] = v; }
    __device__ __forceinline__ uint32_t seg_ld(int k) const { return seg[k * 32]; }
    __device__ __forceinline__ uint32_t cls_t(int b) const {
        const uint32_t w = __ldg(clst + (long long)(b < nblk ? b : nblk - 1) * wc);
        return b < nblk ? w : (uint32_t)((int32_t)w >> 31);  // past the buffer: the class of the last row (bit 31 of the last word)
    }
    __device__ __forceinline__ bool any(bool b) const { return __any_sync(0xffffffffu, b); }
    static __device__ __forceinline__ void pf(const void* p) { asm volatile("prefetch.global.L1 [%0];" ::"l"(p)); }
    __device__ __forceinline__ void prefetch_block(int b) const {
        if (b < nblk) {
            pf(hcol + (long long)b * wc);
            pf(clst + (long long)b * wc);
            if (lane == 0) pf(flag0 + (long long)b * ncg);
        }
    }
    __device__ __forceinline__ void prefetch_g(int r) const { pf(rec_lane + (unsigned)r * rs32); }
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

#ifndef CQ_PAINT_MINB
#define CQ_PAINT_MINB 8
#endif
__global__ void __launch_bounds__(kEmitWarps * 32, CQ_PAINT_MINB) hull_paint_kernel(HullParams p) {
    extern __shared__ __align__(16) uint8_t s_hull[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    int32_t* s_thr = reinterpret_cast<int32_t*>(s_hull);
    for (int i = threadIdx.x; i < 2 * 260; i += kEmitWarps * 32) s_thr[i] = (&p.table->thr[0][0])[i];
    __syncthreads();
    long long id = blockIdx.x;
    const int run = (int)(id % p.nruns_e); id /= p.nruns_e;
    const int strip = (int)(id % p.nstrips_e);
    const int img = (int)(id / p.nstrips_e);
    const int cg = strip * kEmitWarps + warp;
    if (cg >= p.ncg) return;
    const int bfirst = p.out_begin >> 5;
    int y0 = (bfirst + run * p.eb) * 32, y1 = y0 + p.eb * 32;
    y0 = y0 > p.out_begin ? y0 : p.out_begin;
    y1 = y1 < p.out_end ? y1 : p.out_end;
    if (y0 >= y1) return;
    HullPaintCtx c;
    const int x = cg * 32 + lane;
    c.hcol = p.hull + (long long)img * p.nblk * p.wc + x;
    const uint8_t* recbase = p.inter + (long long)img * p.buf_rows * p.rs;
    c.rec_lane = recbase + x; c.cls0 = recbase + p.wp + cg * 4;
    c.flag0 = p.flags + (long long)img * p.nblk * p.ncg + cg;
    c.seg = reinterpret_cast<uint32_t*>(s_hull + kSweepTableBytes + (size_t)warp * kPaintSegWords * 128) + lane;
    c.clst = p.clsT + (long long)img * p.nblk * p.wc + x;
    c.cring = nullptr;
    c.rs = p.rs; c.rs32 = (unsigned)p.rs; c.wc = p.wc; c.ncg = p.ncg; c.nrows = p.buf_rows; c.nblk = p.nblk; c.lane = lane;
    c.thr = s_thr; c.ostride = p.w; c.store_ok = x < p.w;
    const int out_rows = p.out_end - p.out_begin;
    c.optr = p.out + ((long long)img * out_rows + (y0 - p.out_begin)) * (long long)p.w + (x < p.w ? x : 0);
    // 16-byte stores need the whole group inside the image and every row of the group 16-byte aligned
    c.wide_ok = cg * 32 + 32 <= p.w && (p.w & 15) == 0 && (reinterpret_cast<uintptr_t>(p.out) & 15) == 0;
    c.qcls = -1; c.qm = 1; c.qlo = 0; c.qhi = 0x7fffffff; c.qbyte = 0; c.qsign = 1; c.qthr = s_thr;
    int lag = p.L;
    asm volatile("" : "+r"(lag));
    HullPaint<HullPaintCtx> hp(c, lag, p.c2);
    hp.run(y0, y1);
}

}  // namespace chaq
