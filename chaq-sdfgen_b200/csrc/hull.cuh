// hull.cuh -- pass 2, banded fast path (spread <= 256): build and merge of the hull bitmask in global memory
// (hull_core.h); the evaluation is paint.cuh. Replaces dist_transform_axis #2 (reference openmp/df.c:133 -> df.c:29-97, incl.
// the sqrtf at df.c:95) for the inside and the outside field of openmp/sdfgen.c:277-289, and fuses
// transform_float_sub (sdfgen.c:98-106) and transform_float_to_byte (sdfgen.c:75-96) as an exact integer
// threshold look-up (tables.h).
//
//   hull_build_kernel  CTA = kHullWarps (2) warps = one strip of 32-column groups x one run of <= 512 rows; lane = column. g8 rows
//                      arrive by cp.async (16-byte copies, one 32-row block per group, one block ahead); the
//                      run's hull words live in shared memory until the run is done, then go out as 128-byte
//                      lines. No context rows: a run only pushes its own rows. Also leaves one flag per
//                      (block, 32-column group): 0/1 = no site and all class 0/1, 2 = anything else.
//   hull_build4_kernel the same for small grids: 16 warps = 4 groups x 4 sub-runs, inner junctions repaired in
//                      shared memory by the CTA itself (shorter dependent chain per warp).
//   hull_merge_kernel  thread = column; repairs the junctions between runs top to bottom.
//   hull_merge_par_kernel  thread = (junction, column); junctions of one launch are far enough apart to be independent.
// Grids are one CTA per run (thousands), so the hardware scheduler balances dense and empty areas.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "hull_core.h"
#include "common.cuh"
#include "tables.h"

namespace chaq {

#ifndef CQ_HULL_WARPS
#define CQ_HULL_WARPS 2
#endif
#ifndef CQ_BUILD_MINB
#define CQ_BUILD_MINB 12  // CTAs per SM the build kernel is compiled for (10: 0.533 ms, 12: 0.511, 14: 0.519, 16: 0.515 at 16384^2)
#endif
constexpr int kHullWarps = CQ_HULL_WARPS;  // column groups (warps) per build CTA

// Row distances in pass 1's record: one byte (spread <= 256) or, WIDE, one uint16 (spread <= 1024). Same kernels.
template <bool WIDE> struct GVal { typedef uint8_t T; };
template <> struct GVal<true> { typedef uint16_t T; };
template <bool WIDE>
__device__ __forceinline__ uint32_t g_load(const uint8_t* p) {
    return (uint32_t)__ldg(reinterpret_cast<const typename GVal<WIDE>::T*>(p));
}
#ifndef CQ_EMIT_WARPS
#define CQ_EMIT_WARPS 4
#endif
constexpr int kEmitWarps = CQ_EMIT_WARPS;  // warps (32-column groups) per CTA of the emit kernel

struct HullParams {
    const uint8_t* inter;    // banded intermediate, [n_img][buf_rows] records
    uint8_t* out;            // [n_img][out_end-out_begin][w]
    const LevelTable* table;
    uint32_t* hull;          // [n_img][nblk][wc] hull words
    uint8_t* flags;          // [n_img][nblk][ncg] block flags
    uint32_t* clsT;          // [n_img][nblk][wc] class bits, transposed like the hull words (bit j = row 32*blk+j)
    long long rs;            // record stride, bytes
    int cls_off;             // byte offset of the class bit-plane inside a record (after the row distances)
    int wp, w, buf_rows, nblk;
    int wc;                  // ncg * 32
    int ncg, nstrips, n_img;
    int cg0, cg1;            // column groups [cg0, cg1) of this launch (the image may be cut into column parts that
                             // run on different streams); nstrips / nstrips_e count the strips of this range
    int nstrips_e;           // strips of kEmitWarps column groups (emit kernel)
    int out_begin, out_end;
    int pb;                  // blocks per build run
    int nsub, sb;            // a build CTA splits its run into nsub sub-runs of sb blocks (one warp each per group)
    int nruns_b;             // build runs per strip
    int eb;                  // blocks per emit run
    int nruns_e;             // emit runs per strip
    int L, c2;
    int merge_phase, merge_nphase;  // parallel merge: this launch repairs junctions k = 1 + phase + i*nphase
    const uint8_t* lut;      // [2][lut_stride] byte for (class, clamped D2 in [0, c2]); spreads <= 256 only, else null
    int lut_stride, lut_nbk; // bytes per class; buckets per class of the bucketed form (0: one byte per D2)
#ifdef CQ_BUILD_TRACE
    const uint32_t* order;   // experiment (tools/build_order_exp.py): build CTA i works on item order[i]
#endif
    unsigned* work;          // work-item counter of the persistent paint kernel of this launch (zeroed by the build kernel)
};

// Staged blocks of row distances per warp (a ring, filled one block ahead by cp.async). Two: the block being pushed and
// the next. A longer ring, so that pops find heights in shared memory, measured neutral (8 blocks: 0.52 -> 0.54 ms at
// 16384^2): the slow warps of this kernel are bound by instruction issue, not by those loads (tools/build_trace.py).
constexpr int kHullRing = 2;
__host__ __device__ inline size_t hull_build_smem(int pb, int nsub, bool wide) {
    return (size_t)kHullWarps * ((size_t)nsub * kHullRing * (wide ? 2048 : 1024) + (size_t)pb * 128);
}

template <bool WIDE>
struct HullBuildCtx {
    typedef typename GVal<WIDE>::T G;
    static constexpr int kSlot = 1024 * (int)sizeof(G);  // one staged block: 32 rows x 32 columns
    const uint8_t* gst;      // staging ring (kHullRing blocks, block b in slot b % kHullRing), generic pointer at lane 0
    uint32_t gst_s;          // the same, shared-space address
    uint32_t* bits;          // run's hull words, already offset by lane
    const uint8_t* rec0;     // record of buffer row 0 at this warp's first column
    const uint8_t* rec_lane; // the same at this lane's column
    const uint8_t* cls0;     // class word of buffer row 0 for this column group
    uint8_t* flag0;          // flags of block 0 for this column group (stride ncg per block)
    uint32_t* clst0;         // transposed class word of block 0 for this lane's column (stride wc per block)
    int wc;
    long long rs;
    unsigned rs32;
    uint32_t vm;             // columns of the group that exist
    int nrows, lane, cur_blk, slot, blk_end, blk_first, ncg;
    uint32_t craw;           // class word of row cur_blk*32 + lane

    __device__ __forceinline__ void issue(int blk, int s) {
        if (blk < blk_end) {
            constexpr int kPieces = 2 * (int)sizeof(G);  // 16-byte pieces per row of 32 columns
#pragma unroll
            for (int i = 0; i < kPieces; ++i) {
                const int q = lane + 32 * i, row = q / kPieces, piece = q % kPieces;
                const int r = blk * 32 + row;
                if (r < nrows) cp_async16(gst_s + s * kSlot + row * (32 * (int)sizeof(G)) + piece * 16, rec0 + (long long)r * rs + piece * 16);
            }
        }
        cp_async_commit();
    }
    __device__ __forceinline__ void block_begin(int blk) {
        __syncwarp();
        slot = blk & (kHullRing - 1);
        if (blk == blk_first) issue(blk, slot);
        issue(blk + 1, (blk + 1) & (kHullRing - 1));  // takes the slot of block blk + 1 - kHullRing
        int r = blk * 32 + lane;
        r = r < nrows ? r : nrows - 1;
        craw = __ldg(reinterpret_cast<const uint32_t*>(cls0 + (long long)r * rs));
        clst0[(long long)blk * wc] = warp_transpose32(craw, lane);  // for the paint kernel: lane = column again
        cp_async_wait1();
        __syncwarp();
        cur_blk = blk;
    }
    __device__ __forceinline__ bool block_quiet(int L) const {
        const uint32_t* wds = reinterpret_cast<const uint32_t*>(gst + slot * kSlot);
        const int rows_here = nrows - cur_blk * 32;
        uint32_t m = 0xFFFFFFFFu;
        if (WIDE) {
            const uint32_t lim = (uint32_t)L * 0x00010001u;
#pragma unroll
            for (int i = 0; i < 16; ++i) {  // 16 words per row
                const uint32_t v = wds[lane + 32 * i];
                const bool live = (lane >> 4) + 2 * i < rows_here;
                m = __vminu2(m, live ? v : 0xFFFFFFFFu);
            }
            return __all_sync(0xffffffffu, __vcmpgeu2(m, lim) == 0xFFFFFFFFu);
        }
        // byte-wise SIMD minima are emulated on sm_100a (~7 instructions); 16-bit ones are native: even and odd bytes apart
        uint32_t me = 0x00FF00FFu, mo = 0x00FF00FFu;
#pragma unroll
        for (int i = 0; i < 8; ++i) {  // 8 words per row
            uint32_t v = wds[lane + 32 * i];
            if (rows_here < 32) v = (lane >> 3) + 4 * i < rows_here ? v : 0xFFFFFFFFu;
            me = __vminu2(me, __byte_perm(v, 0u, 0x4240));  // bytes 0 and 2 as halfwords
            mo = __vminu2(mo, __byte_perm(v, 0u, 0x4341));  // bytes 1 and 3
        }
        const uint32_t m2 = __vminu2(me, mo);
        const uint32_t mn = min(m2 & 0xFFFFu, m2 >> 16);
        return __all_sync(0xffffffffu, mn >= (uint32_t)L);
    }
    __device__ __forceinline__ uint32_t g_push(int j) const { return reinterpret_cast<const G*>(gst + slot * kSlot)[j * 32 + lane]; }
    __device__ __forceinline__ uint32_t g_row(int r) const {
        const int b = r >> 5;
        if (b > cur_blk - (kHullRing - 1) && b >= blk_first)  // blocks cur_blk - (kHullRing - 2) .. cur_blk are in the ring
            return reinterpret_cast<const G*>(gst + (b & (kHullRing - 1)) * kSlot)[(r & 31) * 32 + lane];
        return g_load<WIDE>(rec_lane + (unsigned)r * rs32);
    }
    __device__ __forceinline__ uint32_t bits_ld(int i) const { return bits[i * 32]; }
    __device__ __forceinline__ void bits_st(int i, uint32_t v) { bits[i * 32] = v; }
    __device__ __forceinline__ void block_end(int blk, uint32_t word) {
        const bool any = __any_sync(0xffffffffu, word != 0u);
        const bool all0 = __all_sync(0xffffffffu, (craw & vm) == 0u);
        const bool all1 = __all_sync(0xffffffffu, (craw & vm) == vm);
        if (lane == 0) flag0[(long long)blk * ncg] = any ? 2 : (all0 ? 0 : (all1 ? 1 : 2));
    }
};

// junction repair inside a build CTA: the run's hull words are still in shared memory
template <bool WIDE>
struct HullRunMergeCtx {
    uint32_t* bits;          // run's hull words of this column (stride 32 words per block), block blk0 first
    const uint8_t* rec_x;
    unsigned rs32;
    int blk0;
    __device__ __forceinline__ uint32_t h_ld(int b) const { return bits[(b - blk0) * 32]; }
    __device__ __forceinline__ void h_st(int b, uint32_t v) { bits[(b - blk0) * 32] = v; }
    __device__ __forceinline__ uint32_t g_row(int r) const { return g_load<WIDE>(rec_x + (unsigned)r * rs32); }
};

// CTA = kHullWarps column groups x nsub sub-runs. Each warp builds the hull of its sub-run alone; then one warp per
// group repairs the nsub-1 junctions inside the run, looking only at the run's own rows (triples that reach across
// a run boundary are the merge kernel's), and the words go out. Splitting the run shortens the chain of dependent
// pushes per warp, which is what a small band's pass 2 waits for.
template <int NSUB, bool WIDE>
__device__ __forceinline__ void hull_build_body(const HullParams p) {
    constexpr int kG = (int)sizeof(typename GVal<WIDE>::T);
    extern __shared__ __align__(16) uint8_t s_hull[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int cgi = warp & (kHullWarps - 1), sub = NSUB > 1 ? warp / kHullWarps : 0;
    long long id = blockIdx.x;
#ifdef CQ_BUILD_TRACE
    if (p.order) id = (long long)__ldg(p.order + blockIdx.x);
#endif
    const int run = (int)(id % p.nruns_b); id /= p.nruns_b;
    const int strip = (int)(id % p.nstrips);
    const int img = (int)(id / p.nstrips);
    const int cg = p.cg0 + strip * kHullWarps + cgi;
    const int b0 = run * p.pb, b1 = b0 + p.pb < p.nblk ? b0 + p.pb : p.nblk;
    const int s0 = b0 + sub * p.sb, s1 = s0 + p.sb < b1 ? s0 + p.sb : b1;
    const bool active = cg < p.cg1;
    if (blockIdx.x == 0 && threadIdx.x == 0 && p.work) *p.work = 0u;  // nobody reads it before the paint kernel of this stream
    if (NSUB == 1 && !active) return;  // no barrier below in this shape
    uint32_t* gbits = reinterpret_cast<uint32_t*>(s_hull + (size_t)cgi * p.pb * 128) + lane;
    const uint8_t* recbase = p.inter + (long long)img * p.buf_rows * p.rs;
    int lag = p.L;
    asm volatile("" : "+r"(lag));
    if (active && s0 < s1) {
        uint8_t* stage = s_hull + (size_t)kHullWarps * p.pb * 128 + (size_t)warp * kHullRing * HullBuildCtx<WIDE>::kSlot;
        HullBuildCtx<WIDE> c;
        c.gst = stage; c.gst_s = smem_u32(stage);
        c.bits = gbits + (s0 - b0) * 32;
        c.rec0 = recbase + cg * 32 * kG; c.rec_lane = c.rec0 + lane * kG; c.cls0 = recbase + p.cls_off + cg * 4;
        c.flag0 = p.flags + (long long)img * p.nblk * p.ncg + cg;
        c.clst0 = p.clsT + (long long)img * p.nblk * p.wc + cg * 32 + lane; c.wc = p.wc;
        c.rs = p.rs; c.rs32 = (unsigned)p.rs;
        const int ncol = p.w - cg * 32;
        c.vm = ncol >= 32 ? 0xFFFFFFFFu : ((1u << ncol) - 1u);
        c.nrows = p.buf_rows; c.lane = lane; c.ncg = p.ncg; c.cur_blk = 0; c.slot = 0;
        c.blk_first = s0; c.blk_end = s1;
        HullBuild<HullBuildCtx<WIDE>, WIDE> hb(c, lag);
        hb.run(s0, s1, p.buf_rows);
    }
    if (NSUB > 1) {
        __syncthreads();
        if (active && sub == 0 && cg * 32 + lane < p.w) {
            HullRunMergeCtx<WIDE> m;
            m.bits = gbits; m.rec_x = recbase + (cg * 32 + lane) * kG; m.rs32 = (unsigned)p.rs; m.blk0 = b0;
            const int hi = (b1 * 32 < p.buf_rows ? b1 * 32 : p.buf_rows) - 1;
            HullMerge<HullRunMergeCtx<WIDE>, WIDE> hm(m, lag, b0 * 32, hi);
            hm.run_rows = p.sb * 32;
            for (int j = b0 + p.sb; j < b1; j += p.sb) hm.junction(j * 32);
        }
        __syncthreads();
    }
    if (active) {
        uint32_t* hw = p.hull + ((long long)img * p.nblk + s0) * p.wc + cg * 32 + lane;
        for (int i = 0; i < s1 - s0; ++i) hw[(long long)i * p.wc] = gbits[(s0 - b0 + i) * 32];
    }
}

#ifdef CQ_BUILD_TRACE
__device__ unsigned long long g_build_trace[3 * 16384];
__device__ __forceinline__ unsigned long long gtimer() { unsigned long long t; asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t)); return t; }
#endif
template <bool WIDE>
__global__ void __launch_bounds__(kHullWarps * 32, WIDE ? 8 : CQ_BUILD_MINB) hull_build_kernel(HullParams p) {
#ifdef CQ_BUILD_TRACE
    unsigned long long t0 = gtimer();
#endif
    hull_build_body<1, WIDE>(p);
#ifdef CQ_BUILD_TRACE
    __syncthreads();
    if (threadIdx.x == 0 && blockIdx.x < 16384) {
        unsigned sm; asm volatile("mov.u32 %0, %smid;" : "=r"(sm));
        const unsigned item = p.order ? p.order[blockIdx.x] : blockIdx.x;
        g_build_trace[3 * item] = t0; g_build_trace[3 * item + 1] = gtimer(); g_build_trace[3 * item + 2] = sm;
    }
#endif
}
template <bool WIDE>
__global__ void __launch_bounds__(kHullWarps * 32 * 4, 2) hull_build4_kernel(HullParams p) { hull_build_body<4, WIDE>(p); }
template <bool WIDE>
__global__ void __launch_bounds__(kHullWarps * 32 * 2, 5) hull_build2_kernel(HullParams p) { hull_build_body<2, WIDE>(p); }

template <bool WIDE>
struct HullMergeCtx {
    uint32_t* hcol;          // hull word of block 0 of this column (stride wc per block)
    const uint8_t* rec_x;    // row distance of buffer row 0 of this column
    unsigned rs32;
    int wc;
    __device__ __forceinline__ uint32_t h_ld(int b) const { return hcol[(long long)b * wc]; }
    __device__ __forceinline__ void h_st(int b, uint32_t v) { hcol[(long long)b * wc] = v; }
    __device__ __forceinline__ uint32_t g_row(int r) const { return g_load<WIDE>(rec_x + (unsigned)r * rs32); }
};

template <bool WIDE>
__global__ void __launch_bounds__(256) hull_merge_kernel(HullParams p) {
    constexpr int kG = (int)sizeof(typename GVal<WIDE>::T);
    const long long t = (long long)blockIdx.x * 256 + threadIdx.x;
    const int x0 = p.cg0 * 32, wpart = min(p.cg1 * 32, p.w) - x0;
    if (t >= (long long)p.n_img * wpart) return;
    const int img = (int)(t / wpart), x = x0 + (int)(t - (long long)img * wpart);
    HullMergeCtx<WIDE> c;
    c.hcol = p.hull + (long long)img * p.nblk * p.wc + x;
    c.rec_x = p.inter + (long long)img * p.buf_rows * p.rs + x * kG;
    c.rs32 = (unsigned)p.rs; c.wc = p.wc;
    HullMerge<HullMergeCtx<WIDE>, WIDE> hm(c, p.L, p.buf_rows);
    hm.run_rows = WIDE ? 0 : p.pb * 32;
    for (int b = p.pb; b < p.nblk; b += p.pb) hm.junction(b * 32);
}

// A junction only ever touches sites within 2L rows of itself, so junctions far enough apart can be repaired at
// the same time: hull_merge_phases(L, run blocks) launches, launch i takes every nphase-th junction. Any order of
// repairs ends in the same kind of hull (each pop re-tests both triples around the adjacency it creates).
// thread = (junction, column).
template <bool WIDE>
__global__ void __launch_bounds__(256) hull_merge_par_kernel(HullParams p) {
    constexpr int kG = (int)sizeof(typename GVal<WIDE>::T);
    const long long t = (long long)blockIdx.x * 256 + threadIdx.x;
    const int njun = p.nruns_b - 1;                       // junctions 1 .. njun (junction k is at block k*pb)
    const int nmine = (njun - p.merge_phase + p.merge_nphase - 1) / p.merge_nphase;  // k = 1 + phase + i*nphase
    const int x0 = p.cg0 * 32, wpart = min(p.cg1 * 32, p.w) - x0;
    const long long per_img = (long long)nmine * wpart;
    if (t >= per_img * p.n_img) return;
    const int img = (int)(t / per_img);
    const long long r = t - (long long)img * per_img;
    const int k = 1 + p.merge_phase + p.merge_nphase * (int)(r / wpart), x = x0 + (int)(r % wpart);
    HullMergeCtx<WIDE> c;
    c.hcol = p.hull + (long long)img * p.nblk * p.wc + x;
    c.rec_x = p.inter + (long long)img * p.buf_rows * p.rs + x * kG;
    c.rs32 = (unsigned)p.rs; c.wc = p.wc;
    HullMerge<HullMergeCtx<WIDE>, WIDE> hm(c, p.L, p.buf_rows);
    hm.run_rows = WIDE ? 0 : p.pb * 32;  // hull words in global memory: galloping's re-scans of long, sparse windows cost more than they save
    hm.junction(k * p.pb * 32);
}

// The same with the hull words around the junction in shared memory: the repair finds sites by scanning hull words
// (galloping re-scans them), and from global memory every word is a dependent L2 round trip. Every thread first copies
// the words of its column for the blocks the repair can touch (kMergeWin each way, asynchronous copies all in flight at
// once) into a private column of shared memory, scans there, and writes changed words through to global memory.
// Blocks beyond the window (spreads above 256) are read from global memory as before.
constexpr int kMergeWin = 16;      // blocks each way: 2L <= 512 rows
constexpr int kMergeThreads = 128;

template <bool WIDE>
struct HullMergeWinCtx {
    uint32_t* hcol;
    const uint8_t* rec_x;
    uint32_t* win;           // this thread's words: slot i at win[i * kMergeThreads], block b0 + i
    unsigned rs32;
    int wc, b0;
    __device__ __forceinline__ uint32_t h_ld(int b) const {
        const unsigned i = (unsigned)(b - b0);
        return i < 2u * kMergeWin ? win[i * kMergeThreads] : hcol[(long long)b * wc];
    }
    __device__ __forceinline__ void h_st(int b, uint32_t v) {
        const unsigned i = (unsigned)(b - b0);
        if (i < 2u * kMergeWin) win[i * kMergeThreads] = v;
        hcol[(long long)b * wc] = v;
    }
    __device__ __forceinline__ uint32_t g_row(int r) const { return g_load<WIDE>(rec_x + (unsigned)r * rs32); }
};

template <bool WIDE>
__global__ void __launch_bounds__(kMergeThreads) hull_merge_win_kernel(HullParams p) {
    __shared__ uint32_t s_win[2 * kMergeWin * kMergeThreads];
    constexpr int kG = (int)sizeof(typename GVal<WIDE>::T);
    const long long t = (long long)blockIdx.x * kMergeThreads + threadIdx.x;
    const int njun = p.nruns_b - 1;
    const int nmine = (njun - p.merge_phase + p.merge_nphase - 1) / p.merge_nphase;
    const int x0 = p.cg0 * 32, wpart = min(p.cg1 * 32, p.w) - x0;
    const long long per_img = (long long)nmine * wpart;
    if (t >= per_img * p.n_img) return;
    const int img = (int)(t / per_img);
    const long long r = t - (long long)img * per_img;
    const int k = 1 + p.merge_phase + p.merge_nphase * (int)(r / wpart), x = x0 + (int)(r % wpart);
    HullMergeWinCtx<WIDE> c;
    c.hcol = p.hull + (long long)img * p.nblk * p.wc + x;
    c.rec_x = p.inter + (long long)img * p.buf_rows * p.rs + x * kG;
    c.rs32 = (unsigned)p.rs; c.wc = p.wc;
    c.win = s_win + threadIdx.x;
    const int kb = k * p.pb;
    c.b0 = kb - kMergeWin;
    {
        const uint32_t dst = smem_u32(c.win);
#pragma unroll
        for (int i = 0; i < 2 * kMergeWin; ++i) {
            const int b = c.b0 + i;
            if (b >= 0 && b < p.nblk) cp_async4(dst + i * kMergeThreads * 4, c.hcol + (long long)b * p.wc);
            else c.win[i * kMergeThreads] = 0u;
        }
        cp_async_commit();
        cp_async_wait0();
    }
    HullMerge<HullMergeWinCtx<WIDE>, WIDE> hm(c, p.L, p.buf_rows);
    hm.run_rows = p.pb * 32;
    hm.junction(kb * 32);
}

}  // namespace chaq
