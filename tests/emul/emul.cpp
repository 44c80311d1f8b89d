// Host emulation of the pass-2 column algorithm (envelope_core.h) and the quantiser (tables.h): the SAME
// headers the CUDA kernels compile, driven column by column on the CPU so that the logic can be checked
// against the oracle without a GPU. Test infrastructure only; not part of the product library.
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <vector>
#include "../../chaq-sdfgen_b200/csrc/tables.h"

using namespace chaq;

// pass 1 restated naively: per row, distance to nearest pixel of the other class, saturated at cap
static void rowscan_naive(const uint8_t* mask, int w, int h, int cap, std::vector<uint16_t>& inter) {
    inter.resize((size_t)w * h);
    for (int y = 0; y < h; ++y) {
        const uint8_t* m = mask + (size_t)y * w;
        for (int x = 0; x < w; ++x) {
            int best = cap;
            for (int d = 1; d < cap; ++d) {
                bool hit = (x - d >= 0 && m[x - d] != m[x]) || (x + d < w && m[x + d] != m[x]);
                if (hit) { best = d; break; }
            }
            inter[(size_t)y * w + x] = (uint16_t)((m[x] ? 0x8000 : 0) | best);
        }
    }
}

struct Ctx {
    const uint16_t* col; size_t stride;
    std::vector<uint32_t> hull, chg;
    int* d2; uint8_t* cls; int row0;
    long nld = 0;
    uint32_t ld(int r) { ++nld; if (r < 0 || r >= nrows) { ++oob; return 0; } return col[(size_t)r * stride]; }
    int nrows = 0; long oob = 0;
    bool ok(int i) { if (i < 0 || i >= (int)hull.size()) { ++oob; return false; } return true; }
    uint32_t hull_get(int i) { return ok(i) ? hull[i] : 0; }
    void hull_set(int i, uint32_t v) { if (ok(i)) hull[i] = v; }
    uint32_t chg_get(int i) { return ok(i) ? chg[i] : 0; }
    void chg_set(int i, uint32_t v) { if (ok(i)) chg[i] = v; }
    void out(int r, int c, int v) { d2[r - row0] = v; cls[r - row0] = (uint8_t)c; }
};

// cap = spread+1 (windowed) or anything >= 0x7fff for the exact unwindowed field.
// band_rows: split the column into bands emitted independently (0 = one band).
extern "C" int emul_column_pass(const uint8_t* mask, int w, int h, int cap, int band_rows, int wide, int32_t* d2,
                                uint8_t* cls, long* loads) {
    std::vector<uint16_t> inter;
    int scap = cap > 0x7fff ? 0x7fff : cap;
    rowscan_naive(mask, w, h, scap, inter);
    if (band_rows <= 0) band_rows = h;
    int nw = ring_words(scap, h);
    long total = 0;
    std::vector<int> cd2(h); std::vector<uint8_t> ccls(h);
    for (int x = 0; x < w; ++x) {
        for (int y0 = 0; y0 < h; y0 += band_rows) {
            int y1 = y0 + band_rows < h ? y0 + band_rows : h;
            Ctx c; c.col = inter.data() + x; c.stride = (size_t)w;
            c.hull.assign(nw, 0xdeadbeefu); c.chg.assign(nw, 0xdeadbeefu);  // poison: must be zeroed by the algorithm
            c.d2 = cd2.data() + y0; c.cls = ccls.data() + y0; c.row0 = y0; c.nrows = h;
            if (wide) { ColumnEnvelope<Ctx, true> ce(c, nw, scap); ce.run(h, y0, y1); }
            else { ColumnEnvelope<Ctx, false> ce(c, nw, scap); ce.run(h, y0, y1); }
            total += c.nld;
            if (c.oob) return 1000 + (int)(c.oob > 1000000 ? 1000000 : c.oob);  // out-of-bounds ring/row access
        }
        for (int y = 0; y < h; ++y) { d2[(size_t)y * w + x] = cd2[y]; cls[(size_t)y * w + x] = ccls[y]; }
    }
    if (loads) *loads = total;
    return 0;
}

extern "C" void emul_quantise(const int32_t* d2, const uint8_t* cls, size_t n, size_t spread, int asym, uint8_t* out) {
    QuantTable q; build_quant_table(spread, asym, &q);
    for (size_t i = 0; i < n; ++i)
        out[i] = (uint8_t)quantise(q.tin, q.tout, q.cap2, q.smin, q.smax, q.scale, cls[i], d2[i]);
}

extern "C" uint8_t emul_ref_byte(int cls, long long d2, size_t spread, int asym) { return ref_byte(cls, d2, spread, asym); }
