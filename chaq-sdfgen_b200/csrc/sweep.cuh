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
// Work split: a CTA of kSweepWarps warps owns a strip of 32*kSweepWarps adjacent columns, so its warps read
// the same 128-byte lines of the record at about the same time. The (image, strip, row) space is linearised
// and cut into runs of equal estimated cost (sweep_plan.cuh), one per resident CTA, so a launch is a single
// wave on 148 SMs whatever the image shape; a run that crosses a strip boundary is finished as two pieces.
// Each piece re-reads L rows of context above and below itself. Blocks of 32 rows in which no column of the
// warp has anything within reach are skipped (sweep_core.h, note_quiet / emit_quiet_block).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "sweep_core.h"
#include "tables.h"

namespace chaq {

constexpr int kSweepWarps = 4;

struct SweepParams {
    const uint8_t* inter;    // banded intermediate, [n_img][buf_rows] records
    uint8_t* out;            // [n_img][out_end-out_begin][w]
    const LevelTable* table;
    long long rs;            // record stride in bytes
    int wp;                  // padded width (bytes of g8 per record)
    int w, buf_rows;
    int out_begin, out_end;  // rows of the buffer to produce
    int ncg;                 // column groups (32 columns) per image = ceil(w/32)
    int nstrips;             // strips per image = ceil(ncg / kSweepWarps)
    long long total_rows;    // n_img * nstrips * (out_end-out_begin): the linearised work
    long long rows_per_cta;  // run length when there are no cut points
    const long long* cuts;   // [gridDim.x + 1] cut points of the linearised work (sweep_plan.cuh), or null
    int nb, L, c2;           // ring blocks, lag, clamp
};

constexpr int kSweepTableBytes = 2176;  // 2*260 int32 thresholds, padded to 128
constexpr int kSweepStageBlocks = 2;    // GMODE: blocks of g8 staged in shared memory on the push side only
// Two ways to keep the g8 rows that the envelope re-reads:
//   GMODE = false: all 2L+64 rows of the window in the shared-memory ring (no global re-read at all);
//   GMODE = true : only the block being pushed (and the one being prefetched) in shared memory, older rows are
//                  re-read through L1/L2 -- 3x less shared memory per warp, so 3x more warps hide that latency.
__host__ __device__ inline size_t sweep_warp_smem(int nb, bool gmode) {
    return (size_t)(gmode ? kSweepStageBlocks : nb) * 1024 + (size_t)nb * 256;
}
__host__ __device__ inline size_t sweep_smem_bytes(int nb, bool gmode) {
    return kSweepTableBytes + (size_t)kSweepWarps * sweep_warp_smem(nb, gmode);
}

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

template <bool GMODE>
struct SweepCtx {
    // this warp's rings in shared memory
    const uint8_t* gring;  // [nb][32 rows][32 lanes], already offset by lane
    const uint8_t* gpush;  // row 0 of the block being pushed, at this lane
    uint32_t* hring;       // [nb][32 lanes], already offset by lane
    uint32_t* cring;
    uint32_t gring_s;      // shared-space address of the ring (lane 0)
    int nb, lane;
    int gslot;             // GMODE: staging slot of the block being pushed
    int cur_blk;           // block being pushed
    // current piece
    const uint8_t* rec0;   // record of relative row 0, at this warp's first column (16-byte aligned)
    const uint8_t* cls0;   // class word of relative row 0 for this column group
    long long rs;
    const uint8_t* rec_lane;  // rec0 + lane
    unsigned rs32;         // record stride (a piece spans < 2^31 bytes)
    int nrows;             // relative rows that exist
    int nblk;              // blocks that will be pushed
    uint32_t craw_next;    // class word of row (blk+1)*32+lane, loaded one block ahead
    uint32_t craw;
    // epilogue
    const int32_t* thr;    // shared-memory copy of LevelTable::thr
    uint8_t* optr;         // where the next result goes
    bool store_ok;         // false: this lane's column is outside the image
    long long ostride;
    int qcls, qm, qlo, qhi, qbyte, qsign;
    const int32_t* qthr;

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
        return __ldg(reinterpret_cast<const uint32_t*>(cls0 + (long long)r * rs));
    }
    __device__ __forceinline__ void block_begin(int blk, int slot) {
        __syncwarp();  // every lane is done with the block whose slot the next prefetch overwrites
        if (GMODE) {   // own small ring: slot = blk mod kSweepStageBlocks
            if (blk == 0) gslot = 0; else gslot = gslot + 1 == kSweepStageBlocks ? 0 : gslot + 1;
            slot = gslot;
        }
        const int ring_blocks = GMODE ? kSweepStageBlocks : nb;
        if (blk == 0) {
            issue(0, slot);
            craw_next = ldg_cls(0);
        }
        issue(blk + 1, slot + 1 == ring_blocks ? 0 : slot + 1);
        cp_async_wait1();
        __syncwarp();
        craw = craw_next;
        craw_next = ldg_cls(blk + 1);
        gpush = gring + slot * 1024;
        cur_blk = blk;
    }
    // the 32x32 bytes of the block being pushed, 8 words per lane (conflict-free), rows past the buffer ignored
    __device__ __forceinline__ bool block_quiet(int L) const {
        const uint32_t* wds = reinterpret_cast<const uint32_t*>(gpush - lane);
        const uint32_t lim = (uint32_t)L * 0x01010101u;
        const int rows_here = nrows - cur_blk * 32;  // rows of this block that exist (>= 1)
        uint32_t m = 0xFFFFFFFFu;
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            const uint32_t v = wds[lane + 32 * i];            // row (lane >> 3) + 4*i, columns 4*(lane & 7)..+3
            const bool live = (lane >> 3) + 4 * i < rows_here;
            m = __vminu4(m, live ? v : 0xFFFFFFFFu);
        }
        return __all_sync(0xffffffffu, __vcmpgeu4(m, lim) == 0xFFFFFFFFu);
    }
    __device__ __forceinline__ bool cls_uniform(uint32_t w, int* bit) const {
        const bool z = __all_sync(0xffffffffu, w == 0u || !store_ok);
        const bool o = __all_sync(0xffffffffu, w == 0xFFFFFFFFu || !store_ok);
        *bit = z ? 0 : 1;
        return z || o;
    }
    __device__ __forceinline__ uint32_t cls_fetch(int) { return warp_transpose32(craw, lane); }
    __device__ __forceinline__ uint32_t g_push(int j) const { return gpush[j * 32]; }
    __device__ __forceinline__ uint32_t g_at(int blk, int slot, int j) const {
        if (GMODE) return __ldg(rec_lane + (unsigned)(blk * 32 + j) * rs32);
        return gring[slot * 1024 + j * 32];
    }
    __device__ __forceinline__ uint32_t hull_ld(int s) const { return hring[s * 32]; }
    __device__ __forceinline__ void hull_st(int s, uint32_t v) { hring[s * 32] = v; }
    __device__ __forceinline__ uint32_t cls_ld(int s) const { return cring[s * 32]; }
    __device__ __forceinline__ void cls_st(int s, uint32_t v) { cring[s * 32] = v; }
    // the byte is a step function of d2c (tables.h); consecutive rows mostly stay on the same step or move to a
    // neighbouring one, so keep the current step [qlo, qhi) and move one step without branching; anything
    // else (class change, a jump of several steps) takes the slow path
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
    __device__ __forceinline__ void out(int cls, int d2c) {
        const int step = (d2c >= qhi ? 1 : 0) - (d2c < qlo ? 1 : 0);
        if (step != 0) {  // one step up or down: two table reads, no loop
            qm += step;
            qbyte += step * qsign;
            qlo = qthr[qm];
            qhi = qthr[qm + 1];
        }
        if (CQ_UNLIKELY(d2c >= qhi || d2c < qlo || cls != qcls)) requant(cls, d2c);
        #ifdef CQ_STCS
        if (store_ok) __stcs(optr, (uint8_t)qbyte);
#else
        if (store_ok) *optr = (uint8_t)qbyte;
#endif
        optr += ostride;
    }
    __device__ __forceinline__ void out_run(int cls, int d2c, int n) {
        out(cls, d2c);
        const uint8_t b = (uint8_t)qbyte;
        for (int i = 1; i < n; ++i) {
            if (store_ok) *optr = b;
            optr += ostride;
        }
    }
};

template <bool GMODE>
__global__ void __launch_bounds__(kSweepWarps * 32, GMODE ? 7 : 2) sweep_kernel(SweepParams p) {
    extern __shared__ __align__(16) uint8_t s_sweep[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    int32_t* s_thr = reinterpret_cast<int32_t*>(s_sweep);
    for (int i = threadIdx.x; i < 2 * 260; i += kSweepWarps * 32) s_thr[i] = (&p.table->thr[0][0])[i];
    __syncthreads();
    uint8_t* base = s_sweep + kSweepTableBytes + (size_t)warp * sweep_warp_smem(p.nb, GMODE);
    const int gblocks = GMODE ? kSweepStageBlocks : p.nb;

    SweepCtx<GMODE> c;
    c.gslot = 0;
    c.gring = base + lane;
    c.gpush = c.gring;
    c.hring = reinterpret_cast<uint32_t*>(base + (size_t)gblocks * 1024) + lane;
    c.cring = c.hring + p.nb * 32;
    c.gring_s = smem_u32(base);
    c.nb = p.nb; c.lane = lane; c.rs = p.rs;
    c.thr = s_thr;
    c.ostride = p.w;

    const int out_rows = p.out_end - p.out_begin;
    {
    long long lin, lin_end;
    if (p.cuts) {
        lin = p.cuts[blockIdx.x];
        lin_end = p.cuts[blockIdx.x + 1];
    } else {
        lin = (long long)blockIdx.x * p.rows_per_cta;
        lin_end = lin + p.rows_per_cta;
    }
    if (lin_end > p.total_rows) lin_end = p.total_rows;
    while (lin < lin_end) {
        const long long piece = lin / out_rows;         // (image, strip)
        const int ya = (int)(lin - piece * out_rows);   // first output row (relative to out_begin)
        const long long rest = lin_end - lin;
        const int yb = (int)(rest < (long long)(out_rows - ya) ? ya + rest : out_rows);
        const int img = (int)(piece / p.nstrips), strip = (int)(piece - (long long)img * p.nstrips);
        const int cg = strip * kSweepWarps + warp;
        if (cg < p.ncg) {
            const int y0 = p.out_begin + ya, y1 = p.out_begin + yb;  // buffer rows
            const int row0 = (y0 - p.L > 0 ? y0 - p.L : 0) & ~31;
            const uint8_t* recbase = p.inter + ((long long)img * p.buf_rows + row0) * p.rs;
            c.rec0 = recbase + cg * 32;
            c.rec_lane = c.rec0 + lane;
            c.rs32 = (unsigned)p.rs;
            c.cls0 = recbase + p.wp + cg * 4;
            c.nrows = p.buf_rows - row0;
            const int tend = (c.nrows < (y1 - row0) + p.L) ? c.nrows : (y1 - row0) + p.L;
            c.nblk = (tend + 31) >> 5;
            const int x = cg * 32 + lane;
            c.store_ok = x < p.w;
            c.optr = p.out + ((long long)img * out_rows + ya) * (long long)p.w + (x < p.w ? x : 0);
            c.qcls = -1; c.qm = 1; c.qlo = 0; c.qhi = 0x7fffffff; c.qbyte = 0; c.qsign = 1; c.qthr = s_thr;
            int lag = p.L;
            asm volatile("" : "+r"(lag));  // keep it in a register: it is used several times per row
            ColumnSweep<SweepCtx<GMODE>> sw(c, p.nb, lag, p.c2);
            sw.run(c.nrows, y0 - row0, y1 - row0);
            __syncwarp();
        }
        lin += yb - ya;
    }
    }
}

}  // namespace chaq
