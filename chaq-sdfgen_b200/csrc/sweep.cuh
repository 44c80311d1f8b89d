// sweep.cuh -- pass 2, banded fast path (spread <= 256): column lower envelope for both fields + fused epilogue.
//
// Replaces dist_transform_axis #2 (reference openmp/df.c:133 -> df.c:29-97, incl. the sqrtf at df.c:95) for the
// inside and the outside field of openmp/sdfgen.c:277-289 in ONE launch, and fuses transform_float_sub
// (sdfgen.c:98-106) and transform_float_to_byte (sdfgen.c:75-96) as an exact integer threshold look-up
// (tables.h). The per-column algorithm is sweep_core.h.
//
// Intermediate ("banded" format, written by rowscan8.cuh): one record per image row,
//     [ Wp bytes g8 | Wp/8 bytes class bits ]      Wp = W rounded up to 128, record stride RS = Wp + Wp/8
// Mapping: a warp owns 32 adjacent columns (lane = column) and walks down a run of rows; every row access of
// the warp is one 32-byte piece of the record, moved by cp.async (16-byte copies, one block of 32 rows per
// group, one block ahead) straight into a shared-memory ring that holds the last 2L+64 rows -- everything the
// envelope ever re-reads. Class bits arrive as one 32-bit word per row and are transposed in the warp
// (5 shuffle stages) into one word per column. Nothing is transposed through global memory.
//
// Work split: the (image, column group, row) space is linearised and cut into equal runs, one per resident
// warp, so a launch is a single wave on 148 SMs whatever the image shape; a run that crosses a column-group
// boundary is finished as two pieces. Each piece re-reads L rows of context above and below itself.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "sweep_core.h"
#include "tables.h"

namespace chaq {

struct SweepParams {
    const uint8_t* inter;    // banded intermediate, [n_img][buf_rows] records
    uint8_t* out;            // [n_img][out_end-out_begin][w]
    const LevelTable* table;
    long long rs;            // record stride in bytes
    int wp;                  // padded width (bytes of g8 per record)
    int w, buf_rows;
    int out_begin, out_end;  // rows of the buffer to produce
    int ncg;                 // column groups per image = ceil(w/32)
    long long total_rows;    // n_img * ncg * (out_end-out_begin): the linearised work
    long long rows_per_warp; // run length per warp
    int nb, L, c2;           // ring blocks, lag, clamp
};

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void cp_async16(uint32_t dst, const void* src) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;\n" ::"r"(dst), "l"(src));
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;\n" ::); }
__device__ __forceinline__ void cp_async_wait1() { asm volatile("cp.async.wait_group 1;\n" ::: "memory"); }

// 32x32 bit transpose across a warp: lane l holds row l; afterwards lane c holds column c (bit l = row l)
__device__ __forceinline__ uint32_t warp_transpose32(uint32_t x, int lane) {
#pragma unroll
    for (int k = 16; k >= 1; k >>= 1) {
        const uint32_t mask = k == 16 ? 0x0000FFFFu : k == 8 ? 0x00FF00FFu : k == 4 ? 0x0F0F0F0Fu : k == 2 ? 0x33333333u : 0x55555555u;
        const uint32_t y = __shfl_xor_sync(0xffffffffu, x, k);
        x = (lane & k) ? ((x & ~mask) | ((y >> k) & mask)) : ((x & mask) | ((y & mask) << k));
    }
    return x;
}

struct SweepCtx {
    // this warp's rings in shared memory
    uint8_t* gring;      // [nb][32 rows][32 lanes]
    uint32_t* hring;     // [nb][32 lanes] + lane
    uint32_t* cring;
    uint32_t gring_s;    // shared-space address of gring
    int nb, lane;
    // current piece
    const uint8_t* rec0;   // record of relative row 0, at this warp's first column (16-byte aligned)
    const uint32_t* cls0;  // class word of relative row 0 for this column group
    long long rs;
    int nrows;             // relative rows that exist
    int nblk;              // blocks that will be pushed
    uint32_t craw_next;    // class word of row (blk+1)*32+lane, loaded one block ahead
    uint32_t craw;
    // epilogue
    const int32_t* thr;    // LevelTable::thr
    uint8_t* o8;           // output of relative row 0 at this lane's column (nullptr: column outside the image)
    long long ostride;
    int qcls, qm, qlo, qhi;

    __device__ __forceinline__ void issue(int blk, int slot) {
        if (blk < nblk) {
#pragma unroll
            for (int i = 0; i < 2; ++i) {
                const int q = lane + 32 * i, row = q >> 1, half = q & 1;
                const int r = blk * 32 + row;
                if (r < nrows) cp_async16(gring_s + slot * 1024 + row * 32 + half * 16, rec0 + (long long)r * rs + half * 16);
            }
        }
        cp_async_commit();
    }
    __device__ __forceinline__ uint32_t ldg_cls(int blk) {
        int r = blk * 32 + lane;
        r = r < nrows ? r : nrows - 1;
        return __ldg(reinterpret_cast<const uint32_t*>(reinterpret_cast<const uint8_t*>(cls0) + (long long)r * rs));
    }
    __device__ __forceinline__ void block_begin(int blk, int slot) {
        __syncwarp();  // every lane is done with the block whose slot the next prefetch overwrites
        if (blk == 0) {
            issue(0, slot);
            craw_next = ldg_cls(0);
        }
        issue(blk + 1, slot + 1 == nb ? 0 : slot + 1);
        cp_async_wait1();
        __syncwarp();
        craw = craw_next;
        craw_next = ldg_cls(blk + 1);
    }
    __device__ __forceinline__ uint32_t cls_fetch(int) { return warp_transpose32(craw, lane); }
    __device__ __forceinline__ uint32_t ld_g(int slot, int j) const { return gring[slot * 1024 + j * 32 + lane]; }
    __device__ __forceinline__ uint32_t hull_ld(int s) const { return hring[s * 32]; }
    __device__ __forceinline__ void hull_st(int s, uint32_t v) { hring[s * 32] = v; }
    __device__ __forceinline__ uint32_t cls_ld(int s) const { return cring[s * 32]; }
    __device__ __forceinline__ void cls_st(int s, uint32_t v) { cring[s * 32] = v; }
    __device__ __forceinline__ void out(int rr, int cls, int d2c) {
        if (cls != qcls) {
            qcls = cls;
            qm = level_of(thr + cls * 260, d2c);
            qlo = __ldg(thr + cls * 260 + qm);
            qhi = __ldg(thr + cls * 260 + qm + 1);
        } else {
            while (d2c >= qhi) { ++qm; qlo = qhi; qhi = __ldg(thr + cls * 260 + qm + 1); }
            while (d2c < qlo) { --qm; qhi = qlo; qlo = __ldg(thr + cls * 260 + qm); }
        }
        if (o8) o8[(long long)rr * ostride] = (uint8_t)(cls ? qm : 255 - qm);
    }
};

template <int WPC>
__global__ void __launch_bounds__(WPC * 32) sweep_kernel(SweepParams p) {
    extern __shared__ __align__(16) uint8_t s_sweep[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const size_t per_warp = (size_t)p.nb * (1024 + 128 + 128);
    uint8_t* base = s_sweep + (size_t)warp * per_warp;

    SweepCtx c;
    c.gring = base;
    c.hring = reinterpret_cast<uint32_t*>(base + (size_t)p.nb * 1024) + lane;
    c.cring = c.hring + p.nb * 32;
    c.gring_s = smem_u32(base);
    c.nb = p.nb; c.lane = lane; c.rs = p.rs;
    c.thr = &p.table->thr[0][0];
    c.ostride = p.w;

    const int out_rows = p.out_end - p.out_begin;
    const long long slot = (long long)blockIdx.x * WPC + warp;
    long long lin = slot * p.rows_per_warp;
    long long lin_end = lin + p.rows_per_warp;
    if (lin_end > p.total_rows) lin_end = p.total_rows;
    while (lin < lin_end) {
        const long long piece = lin / out_rows;         // (image, column group)
        const int ya = (int)(lin - piece * out_rows);   // first output row (relative to out_begin)
        long long rest = lin_end - lin;
        const int yb = (int)(rest < (long long)(out_rows - ya) ? ya + rest : out_rows);
        const int img = (int)(piece / p.ncg), cg = (int)(piece - (long long)img * p.ncg);
        const int y0 = p.out_begin + ya, y1 = p.out_begin + yb;  // buffer rows
        const int row0 = (y0 - p.L > 0 ? y0 - p.L : 0) & ~31;
        const uint8_t* recbase = p.inter + ((long long)img * p.buf_rows + row0) * p.rs;
        c.rec0 = recbase + cg * 32;
        c.cls0 = reinterpret_cast<const uint32_t*>(recbase + p.wp) + cg;
        c.nrows = p.buf_rows - row0;
        const int tend = (c.nrows < (y1 - row0) + p.L) ? c.nrows : (y1 - row0) + p.L;
        c.nblk = (tend + 31) >> 5;
        const int x = cg * 32 + lane;
        c.o8 = x < p.w ? p.out + ((long long)img * out_rows + (row0 - p.out_begin)) * (long long)p.w + x : nullptr;
        c.qcls = -1; c.qm = 0; c.qlo = 0; c.qhi = 0;
        ColumnSweep<SweepCtx> sw(c, p.nb, p.L, p.c2);
        sw.run(c.nrows, y0 - row0, y1 - row0);
        __syncwarp();
        lin += yb - ya;
    }
}

}  // namespace chaq
