// sweep_plan.cuh -- work split of the banded pass 2 (sweep.cuh): equal COST per CTA instead of equal rows.
//
// Pass 2 costs little where nothing is within reach (sweep_core.h skips whole 32-row blocks there) and a lot
// near class boundaries, so equal row counts leave most SMs idle while a few finish the dense runs. Pass 1
// (rowscan8.cuh) leaves one "busy" byte per (row, 128-column strip); two tiny kernels turn them into cut
// points of the linearised (image, strip, row) space:
//   sweep_busy_kernel   busy bytes of 32 rows -> one byte per (piece, 32-row block)    [many CTAs]
//   sweep_cuts_kernel   cost per block = rows * (c0 + cp*[block busy] + ce*[a busy block within reach]),
//                       prefix sum, cuts[k] = first row whose prefix cost reaches k/ncta of the total [1 CTA]
// The reference has no counterpart (its OpenMP loop is `schedule(static)` over lines, df.c:116); this is
// launch geometry only and never changes a result.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace chaq {

constexpr int kPlanThreads = 1024;
constexpr int kSweepMaxCtas = 4096;

struct SweepPlanParams {
    const uint8_t* inter;  // records
    long long rs;          // record stride
    int flag_off;          // offset of the busy bytes in a record
    int buf_rows, nstrips, n_img;
    int out_begin, out_end;
    int nblk;              // 32-row blocks of the buffer = ceil(buf_rows/32)
    int b0, nbo;           // first block holding output rows, number of such blocks
    int reach;             // blocks above/below whose sites can reach a row = ceil(L/32)
    uint8_t* pb;           // [n_img*nstrips][nblk] block-busy bytes
    long long* cuts;       // [ncta+1]
    int ncta;
    int c0, cp, ce;        // cost model (per row)
};

__global__ void __launch_bounds__(256) sweep_busy_kernel(SweepPlanParams p) {
    const long long idx = (long long)blockIdx.x * 256 + threadIdx.x;  // ((img*nblk + blk)*nstrips + strip)
    const long long n = (long long)p.n_img * p.nblk * p.nstrips;
    if (idx >= n) return;
    const int strip = (int)(idx % p.nstrips);
    const long long t = idx / p.nstrips;
    const int blk = (int)(t % p.nblk), img = (int)(t / p.nblk);
    const int r0 = blk * 32, r1 = r0 + 32 < p.buf_rows ? r0 + 32 : p.buf_rows;
    const uint8_t* f = p.inter + ((long long)img * p.buf_rows + r0) * p.rs + p.flag_off + strip;
    uint32_t any = 0;
    for (int r = r0; r < r1; ++r, f += p.rs) any |= __ldg(f);
    p.pb[((long long)img * p.nstrips + strip) * p.nblk + blk] = any ? 1 : 0;
}

__global__ void __launch_bounds__(kPlanThreads) sweep_cuts_kernel(SweepPlanParams p) {
    __shared__ unsigned long long s_warp[32];
    const int tid = threadIdx.x;
    const long long n_items = (long long)p.n_img * p.nstrips * p.nbo;
    const long long chunk = (n_items + kPlanThreads - 1) / kPlanThreads;
    const long long i0 = (long long)tid * chunk, i1 = i0 + chunk < n_items ? i0 + chunk : n_items;
    const int out_rows = p.out_end - p.out_begin;

    // cost of item i = (piece, output block ob); the window count of busy blocks within reach slides along a piece
    auto walk = [&](auto&& visit) {
        long long piece = -1;
        int ob = 0, win = 0;
        const uint8_t* pbp = nullptr;
        for (long long i = i0; i < i1; ++i) {
            const long long pc = i / p.nbo;
            if (pc != piece) {
                piece = pc; ob = (int)(i - pc * p.nbo);
                pbp = p.pb + pc * p.nblk;
                const int b = p.b0 + ob;
                win = 0;
                for (int q = b - p.reach; q <= b + p.reach; ++q) if (q >= 0 && q < p.nblk) win += pbp[q];
            }
            const int b = p.b0 + ob;
            const int r0 = b * 32 > p.out_begin ? b * 32 : p.out_begin;
            const int r1 = b * 32 + 32 < p.out_end ? b * 32 + 32 : p.out_end;
            const int cpr = p.c0 + (pbp[b] ? p.cp : 0) + (win ? p.ce : 0);
            visit(piece, r0, r1 - r0, cpr);
            // slide to b+1
            const int add = b + 1 + p.reach, sub = b - p.reach;
            if (add < p.nblk) win += pbp[add];
            if (sub >= 0) win -= pbp[sub];
            ++ob;
        }
    };

    unsigned long long mine = 0;
    walk([&](long long, int, int rows, int cpr) { mine += (unsigned long long)rows * cpr; });
    // exclusive scan of the 1024 partial sums
    const int lane = tid & 31, wid = tid >> 5;
    unsigned long long v = mine;
    for (int d = 1; d < 32; d <<= 1) {
        const unsigned long long o = __shfl_up_sync(0xffffffffu, v, d);
        if (lane >= d) v += o;
    }
    if (lane == 31) s_warp[wid] = v;
    __syncthreads();
    if (wid == 0) {
        unsigned long long w = s_warp[lane];
        for (int d = 1; d < 32; d <<= 1) {
            const unsigned long long o = __shfl_up_sync(0xffffffffu, w, d);
            if (lane >= d) w += o;
        }
        s_warp[lane] = w;
    }
    __syncthreads();
    const unsigned long long total = s_warp[31];
    unsigned long long pre = v - mine + (wid ? s_warp[wid - 1] : 0ull);
    if (tid == 0) { p.cuts[0] = 0; p.cuts[p.ncta] = (long long)p.n_img * p.nstrips * out_rows; }
    if (total == 0 || i0 >= i1) return;
    // target k sits at prefix cost T_k = floor(k*total/ncta) < total; it falls into exactly one item (every item
    // costs > 0). This thread owns the targets inside [pre, pre + mine).
    const unsigned long long ncta = (unsigned long long)p.ncta;
    unsigned long long k = (pre * ncta + total - 1) / total;  // smallest k with T_k >= pre
    if (k == 0) k = 1;
    walk([&](long long piece, int r0, int rows, int cpr) {
        const unsigned long long cost = (unsigned long long)rows * cpr;
        while (k < ncta) {
            const unsigned long long T = k * total / ncta;
            if (T >= pre + cost) break;
            const int off = (int)((T - pre) / cpr);
            p.cuts[k] = piece * out_rows + (r0 - p.out_begin) + off;
            ++k;
        }
        pre += cost;
    });
}

}  // namespace chaq
