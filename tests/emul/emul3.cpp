// Host emulation of the build / merge / paint organisation of the banded column pass (hull_core.h, paint_core.h)
// and the level tables (tables.h): the SAME headers the CUDA kernels compile, driven column by column on the CPU, so
// the logic (run boundaries, junction repair, owner lists, saturated blocks) can be checked against the oracle
// without a GPU.
// Test infrastructure only; not part of the product library.
#include <cstdint>
#include <cstdlib>
#include <algorithm>
#include <cstring>
#include <vector>
#include "../../chaq-sdfgen_b200/csrc/tables.h"
#include "../../chaq-sdfgen_b200/csrc/hull_core.h"
#include "../../chaq-sdfgen_b200/csrc/paint_core.h"

using namespace chaq;

// wide = 1: the same organisation with a 16-bit row distance and the 64-bit domination test, for L up to 1024
typedef uint16_t gval_t;
static int g_max_cap = 256;

static void rowscan_naive(const uint8_t* mask, int w, int h, int C, std::vector<gval_t>& g8) {
    g8.resize((size_t)w * h);
    for (int y = 0; y < h; ++y) {
        const uint8_t* m = mask + (size_t)y * w;
        for (int x = 0; x < w; ++x) {
            int best = C;
            for (int d = 1; d < C; ++d) {
                bool hit = (x - d >= 0 && m[x - d] != m[x]) || (x + d < w && m[x + d] != m[x]);
                if (hit) { best = d; break; }
            }
            g8[(size_t)y * w + x] = (gval_t)(best - 1);
        }
    }
}

struct Col {
    const gval_t* g; const uint8_t* m; size_t stride; int nrows, nblk;
    std::vector<uint32_t>* H;   // hull words of this column
    std::vector<uint8_t>* Q;    // block flags of this column (the "group" is this one column)
    long oob = 0;
    uint32_t g_row(int r) { if (r < 0 || r >= nrows) { ++oob; return g_max_cap - 1; } return g[(size_t)r * stride]; }
};

struct BuildCtx : Col {
    int cur = 0, b0 = 0;
    std::vector<uint32_t> bits;
    void block_begin(int blk) { cur = blk; }
    bool block_quiet(int L) {
        for (int j = 0; j < 32; ++j) { int r = cur * 32 + j; if (r < nrows && g[(size_t)r * stride] < L) return false; }
        return true;
    }
    uint32_t g_push(int j) { return g_row(cur * 32 + j); }
    uint32_t bits_ld(int i) { if (i < 0 || i >= (int)bits.size()) { ++oob; return 0; } return bits[i]; }
    void bits_st(int i, uint32_t v) { if (i < 0 || i >= (int)bits.size()) { ++oob; return; } bits[i] = v; }
    void block_end(int blk, uint32_t word) {
        int c0 = 1, c1 = 1;
        for (int j = 0; j < 32; ++j) { int r = blk * 32 + j; if (r >= nrows) r = nrows - 1; if (m[(size_t)r * stride] & 1) c0 = 0; else c1 = 0; }
        (*Q)[blk] = word ? 2 : (c0 ? 0 : (c1 ? 1 : 2));
    }
};

struct MergeCtx : Col {
    // junctions of one round run concurrently on the GPU (plain loads and stores): no two of them may store the
    // same word. writer[b] = junction that stored word b in the current round (-1: none).
    std::vector<int>* writer = nullptr;
    int junction_id = -1;
    long conflicts = 0;
    uint32_t h_ld(int b) { if (b < 0 || b >= nblk) { ++oob; return 0; } return (*H)[b]; }
    void h_st(int b, uint32_t v) {
        if (b < 0 || b >= nblk) { ++oob; return; }
        if (writer) {
            if ((*writer)[b] >= 0 && (*writer)[b] != junction_id) ++conflicts;
            (*writer)[b] = junction_id;
        }
        (*H)[b] = v;
    }
};

struct EmitCtx : Col {
    int* d2; uint8_t* ocls; int nout = 0;
    uint32_t h_ld(int b) { if (b < 0) { ++oob; return 0; } return b < nblk ? (*H)[b] : 0u; }
    uint32_t cls_fetch(int blk) {
        uint32_t wd = 0;
        if (blk < 0) { ++oob; return 0; }
        for (int j = 0; j < 32; ++j) { int r = blk * 32 + j; if (r >= nrows) r = nrows - 1; wd |= (uint32_t)(m[(size_t)r * stride] & 1) << j; }
        return wd;
    }
    uint32_t cls_t(int blk) { return cls_fetch(blk); }
    void prefetch_block(int) {}
    void prefetch_g(int) {}
    long n_live = 0, n_dead = 0;
    bool any(bool b) { if (b) ++n_live; else ++n_dead; return b; }
    int q_flag(int blk) { return (blk < 0 || blk >= nblk) ? 3 : (*Q)[blk]; }
    void out(int c, int v) { d2[nout] = v; ocls[nout] = (uint8_t)c; ++nout; }
    void out_run(int c, int v, int n) { for (int i = 0; i < n; ++i) out(c, v); }
    // owner list of the block being painted (paint_core.h); slots past 32 would be a bug
    uint32_t seg[33];
    void seg_st(int k, uint32_t v) { if (k < 0 || k > 32) { ++oob; return; } seg[k] = v; }
    uint32_t seg_ld(int k) { if (k < 0 || k > 32) { ++oob; return 0; } return seg[k]; }
};

static long g_live = 0, g_dead = 0;  // (column, block) pairs painted / skipped as saturated by HullPaint
extern "C" void emul3_paint_stats(long* live, long* dead) { *live = g_live; *dead = g_dead; g_live = g_dead = 0; }

// mask: h*w bytes of 0/1. build_blocks: blocks per build run; emit_rows: rows per emit run (0 = whole column);
// y_begin/y_end: output row range (context rows outside it still exist, as in a row-band buffer).
// nsub > 1: every build run is built as nsub sub-runs whose junctions are repaired first, looking only at the
// run's own rows (what one CTA of the build kernel does in shared memory), before the junctions between runs.
// merge_order 0: junctions top to bottom; 1: in hull_merge_phases(L, run blocks) rounds, round i taking every
// nphase-th junction (what the parallel merge kernel does; same-round junctions must never store the same hull
// word -- rc 2500 if they do).
template <bool WIDE>
static int emul3_run(const uint8_t* mask, int w, int h, size_t spread, int asym, int build_blocks, int emit_rows,
                     int y_begin, int y_end, uint8_t* out, int* cap_eff, int merge_order, int nsub) {
    g_max_cap = WIDE ? 1025 : 256;
    QuantTable q; build_quant_table(spread, asym, &q);
    LevelTable lt; build_level_table(q, &lt);
    const int C = lt.cap_eff, L = C - 1;
    if (cap_eff) *cap_eff = C;
    if (C > g_max_cap) return 1;
    if (build_blocks < 1 || build_blocks > hull_max_run_blocks(WIDE)) return 3;
    std::vector<gval_t> g8;
    rowscan_naive(mask, w, h, C, g8);
    const int nblk = (h + 31) / 32;
    if (emit_rows <= 0) emit_rows = h;
    std::vector<int> cd2(h); std::vector<uint8_t> ccls(h);
    for (int x = 0; x < w; ++x) {
        std::vector<uint32_t> H(nblk, 0xdeadbeefu);
        std::vector<uint8_t> Q(nblk, 2);
        if (nsub < 1) nsub = 1;
        const int sb = (build_blocks + nsub - 1) / nsub;  // blocks per sub-run
        for (int b0 = 0; b0 < nblk; b0 += build_blocks) {
            const int b1 = b0 + build_blocks < nblk ? b0 + build_blocks : nblk;
            for (int s0 = b0; s0 < b1; s0 += sb) {
                BuildCtx c;
                c.g = g8.data() + x; c.m = mask + x; c.stride = (size_t)w; c.nrows = h; c.nblk = nblk; c.H = &H; c.Q = &Q;
                const int s1 = s0 + sb < b1 ? s0 + sb : b1;
                c.b0 = s0; c.bits.assign(s1 - s0, 0xdeadbeefu);
                HullBuild<BuildCtx, WIDE> hb(c, L);
                hb.run(s0, s1, h);
                for (int b = s0; b < s1; ++b) H[b] = c.bits[b - s0];
                if (c.oob) return 1000;
            }
            if (nsub > 1) {  // junctions inside the run, restricted to the run's rows
                MergeCtx c;
                c.g = g8.data() + x; c.m = mask + x; c.stride = (size_t)w; c.nrows = h; c.nblk = nblk; c.H = &H; c.Q = &Q;
                const int hi = (b1 * 32 < h ? b1 * 32 : h) - 1;
                HullMerge<MergeCtx, WIDE> hm(c, L, b0 * 32, hi);
                hm.run_rows = sb * 32;
                for (int s0 = b0 + sb; s0 < b1; s0 += sb) hm.junction(s0 * 32);
                if (c.oob) return 1500;
            }
        }
        {
            MergeCtx c;
            c.g = g8.data() + x; c.m = mask + x; c.stride = (size_t)w; c.nrows = h; c.nblk = nblk; c.H = &H; c.Q = &Q;
            HullMerge<MergeCtx, WIDE> hm(c, L, h);
            hm.run_rows = build_blocks * 32;
            if (merge_order == 1) {
                const int nphase = hull_merge_phases(L, build_blocks);
                std::vector<int> writer(nblk);
                c.writer = &writer;
                for (int phase = 0; phase < nphase; ++phase) {
                    std::fill(writer.begin(), writer.end(), -1);
                    for (int k = 1 + phase; k * build_blocks < nblk; k += nphase) {
                        c.junction_id = k;
                        hm.junction(k * build_blocks * 32);
                    }
                }
                if (c.conflicts) return 2500;
            } else {
                for (int b0 = build_blocks; b0 < nblk; b0 += build_blocks) hm.junction(b0 * 32);
            }
            if (c.oob) return 2000;
        }
        for (int y0 = y_begin; y0 < y_end; y0 += emit_rows) {
            const int y1 = y0 + emit_rows < y_end ? y0 + emit_rows : y_end;
            EmitCtx c;
            c.g = g8.data() + x; c.m = mask + x; c.stride = (size_t)w; c.nrows = h; c.nblk = nblk; c.H = &H; c.Q = &Q;
            c.d2 = cd2.data() + y0; c.ocls = ccls.data() + y0;
            HullPaint<EmitCtx> hp(c, L, lt.c2);
            hp.run(y0, y1);
            g_live += c.n_live; g_dead += c.n_dead;
            if (c.nout != y1 - y0) return 2;
            if (c.oob) return 3000;
        }
        for (int y = y_begin; y < y_end; ++y) {
            const int cls = ccls[y];
            const int mlev = level_of(lt.thr[cls], cd2[y]);
            out[(size_t)y * w + x] = (uint8_t)(cls ? mlev : 255 - mlev);
        }
    }
    return 0;
}

extern "C" int emul3_sdf(const uint8_t* mask, int w, int h, size_t spread, int asym, int build_blocks, int emit_rows,
                         int y_begin, int y_end, uint8_t* out, int* cap_eff, int merge_order, int nsub) {
    return emul3_run<false>(mask, w, h, spread, asym, build_blocks, emit_rows, y_begin, y_end, out, cap_eff, merge_order, nsub);
}
extern "C" int emul3_sdf_wide(const uint8_t* mask, int w, int h, size_t spread, int asym, int build_blocks, int emit_rows,
                              int y_begin, int y_end, uint8_t* out, int* cap_eff, int merge_order, int nsub) {
    return emul3_run<true>(mask, w, h, spread, asym, build_blocks, emit_rows, y_begin, y_end, out, cap_eff, merge_order, nsub);
}
