// Host emulation of the banded column pass (sweep_core.h) and the level tables (tables.h): the SAME headers
// the CUDA kernels compile, driven column by column on the CPU with physically emulated rings (a stale slot
// returns whatever block overwrote it), so the logic can be checked against the oracle without a GPU.
// Test infrastructure only; not part of the product library.
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <vector>
#include "../../chaq-sdfgen_b200/csrc/tables.h"
#include "../../chaq-sdfgen_b200/csrc/sweep_core.h"

using namespace chaq;

// pass 1 restated naively: g8 = min(G, C) - 1 and the class, per pixel
static void rowscan_naive(const uint8_t* mask, int w, int h, int C, std::vector<uint8_t>& g8) {
    g8.resize((size_t)w * h);
    for (int y = 0; y < h; ++y) {
        const uint8_t* m = mask + (size_t)y * w;
        for (int x = 0; x < w; ++x) {
            int best = C;
            for (int d = 1; d < C; ++d) {
                bool hit = (x - d >= 0 && m[x - d] != m[x]) || (x + d < w && m[x + d] != m[x]);
                if (hit) { best = d; break; }
            }
            g8[(size_t)y * w + x] = (uint8_t)(best - 1);
        }
    }
}

struct Ctx2 {
    const uint8_t* gcol; const uint8_t* mcol; size_t stride;  // column base at relative row 0
    int nrows;                                                  // relative rows that exist
    int NB;
    std::vector<int> slot_blk;
    std::vector<uint32_t> hull, cls;
    int* d2; uint8_t* ocls; int y0;
    long oob = 0;
    int cur_slot = 0, nout = 0;
    int Lg = 0;
    void block_begin(int blk, int slot) {
        cur_blk = blk;
        cur_slot = slot;
        slot_blk[slot] = blk;
        slot_blk[(slot + 1) % NB] = blk + 1;  // the prefetch of the next block lands there now
    }
    uint32_t cls_fetch(int blk) {
        uint32_t wd = 0;
        for (int j = 0; j < 32; ++j) {
            int r = blk * 32 + j;
            if (r >= nrows) r = nrows - 1;
            wd |= (uint32_t)(mcol[(size_t)r * stride] & 1) << j;
        }
        return wd;
    }
    bool gmode = false;  // g8 re-read from "global memory" by block number; only hull/class live in rings
    int cur_blk = 0;
    uint32_t g_push(int j) { return g_at(gmode ? cur_blk : slot_blk[cur_slot], cur_slot, j); }
    uint32_t g_at(int blk, int slot, int j) {
        if (gmode) {
            const int r = blk * 32 + j;
            if (r < 0 || r >= nrows || blk > cur_blk || blk * 32 + 31 < cur_blk * 32 - 2 * Lg) { ++oob; return 0; }
            return gcol[(size_t)r * stride];
        }
        if (slot < 0 || slot >= NB) { ++oob; return 0; }
        if (slot_blk[slot] != blk) ++oob;  // the ring no longer holds that block
        int r = slot_blk[slot] * 32 + j;
        if (r < 0 || r >= nrows) { ++oob; return 0; }
        return gcol[(size_t)r * stride];
    }
    uint32_t hull_ld(int s) { if (s < 0 || s >= NB) { ++oob; return 0; } return hull[s]; }
    void hull_st(int s, uint32_t v) { if (s < 0 || s >= NB) { ++oob; return; } hull[s] = v; }
    uint32_t cls_ld(int s) { if (s < 0 || s >= NB) { ++oob; return 0; } return cls[s]; }
    void cls_st(int s, uint32_t v) { if (s < 0 || s >= NB) { ++oob; return; } cls[s] = v; }
    void out(int c, int v) { d2[nout] = v; ocls[nout] = (uint8_t)c; ++nout; }
    void out_run(int c, int v, int n) { for (int i = 0; i < n; ++i) out(c, v); }
    // the group that shares control flow is this one column
    bool block_quiet(int L) {
        const int blk = cur_blk;
        for (int j = 0; j < 32; ++j) {
            const int r = blk * 32 + j;
            if (r < nrows && gcol[(size_t)r * stride] < L) return false;
        }
        return true;
    }
    bool cls_uniform(uint32_t w, int* bit) { *bit = w ? 1 : 0; return w == 0u || w == 0xFFFFFFFFu; }
};

// mask: h*w bytes of 0/1. band_rows: split each column into bands finished independently (0 = one band).
// Outputs: the byte image; optionally the clamped squared distances and classes. Returns 0, or 1000+n on
// out-of-range ring/row accesses.
extern "C" int emul2_sdf(const uint8_t* mask, int w, int h, size_t spread, int asym, int band_rows, uint8_t* out,
                         int32_t* d2_out, int* cap_eff, int gmode) {
    QuantTable q; build_quant_table(spread, asym, &q);
    LevelTable lt; build_level_table(q, &lt);
    const int C = lt.cap_eff, L = C - 1;
    if (cap_eff) *cap_eff = C;
    if (C > 256) return 1;
    std::vector<uint8_t> g8;
    rowscan_naive(mask, w, h, C, g8);
    if (band_rows <= 0) band_rows = h;
    const int NB = gmode ? sweep_ring_blocks_g(L) : sweep_ring_blocks(L);
    std::vector<int> cd2(h); std::vector<uint8_t> ccls(h);
    for (int x = 0; x < w; ++x) {
        for (int y0 = 0; y0 < h; y0 += band_rows) {
            const int y1 = y0 + band_rows < h ? y0 + band_rows : h;
            const int row0 = (y0 - L > 0 ? y0 - L : 0) & ~31;
            Ctx2 c;
            c.gcol = g8.data() + (size_t)row0 * w + x; c.mcol = mask + (size_t)row0 * w + x; c.stride = (size_t)w;
            c.nrows = h - row0; c.NB = NB; c.gmode = gmode != 0; c.Lg = L;
            c.slot_blk.assign(NB, -1000000);
            c.hull.assign(NB, 0xdeadbeefu); c.cls.assign(NB, 0xdeadbeefu);
            c.d2 = cd2.data() + y0; c.ocls = ccls.data() + y0; c.y0 = y0 - row0;
            ColumnSweep<Ctx2> sw(c, NB, L, lt.c2);
            sw.run(h - row0, y0 - row0, y1 - row0);
            if (c.nout != y1 - y0) return 2;
            if (c.oob) return 1000 + (int)(c.oob > 1000000 ? 1000000 : c.oob);
        }
        for (int y = 0; y < h; ++y) {
            const int cls = ccls[y];
            const int m = level_of(lt.thr[cls], cd2[y]);
            out[(size_t)y * w + x] = (uint8_t)(cls ? m : 255 - m);
            if (d2_out) d2_out[(size_t)y * w + x] = cd2[y];
        }
    }
    return 0;
}
