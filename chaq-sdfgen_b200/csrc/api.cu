// api.cu -- the C ABI of include/chaqsdf.h over the sm_100a kernels. No CPU compute path exists here:
// every entry point either launches the CUDA kernels or fails with an error code.
#include <cuda_runtime.h>

#include <atomic>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <mutex>
#include <string>
#include <tuple>

#include "../../include/chaqsdf.h"
#include "envelope.cuh"
#include "rowscan.cuh"
#include "rowscan8.cuh"
#include "hull.cuh"
#include "paint.cuh"
#include "paint2.cuh"
#include "glyph.cuh"
#include "transpose.cuh"
#include "tuning.h"

using namespace chaq;

namespace {

thread_local std::string g_err;
std::atomic<unsigned long long> g_launches{0};

int fail(int code, const std::string& msg) {
    g_err = msg;
    return code;
}
#define CQ_CUDA(expr)                                                                                   \
    do {                                                                                                \
        cudaError_t e_ = (expr);                                                                        \
        if (e_ != cudaSuccess)                                                                          \
            return fail(CHAQSDF_ERR_CUDA, std::string(#expr) + ": " + cudaGetErrorString(e_));          \
    } while (0)

#define CQ_LAUNCHED(what, grid, smem)                                                                   \
    do {                                                                                                \
        cudaError_t e_ = cudaGetLastError();                                                            \
        if (e_ != cudaSuccess)                                                                          \
            return fail(CHAQSDF_ERR_CUDA, std::string(what) + " launch (grid " + std::to_string(grid) + ", smem " +  \
                                              std::to_string(smem) + "): " + cudaGetErrorString(e_));   \
        g_launches.fetch_add(1, std::memory_order_relaxed);                                             \
    } while (0)

size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

Tuning read_tuning() {
    Tuning t;
    auto num = [](const char* k) { const char* e = getenv(k); return e ? atoll(e) : 0LL; };
    t.band_rows = num("CHAQSDF_BAND_ROWS");
    t.hull_pb = num("CHAQSDF_HULL_PB");
    t.hull_sub = num("CHAQSDF_HULL_SUB");
    t.hull_eb = num("CHAQSDF_HULL_EB");
    t.hull_seq = getenv("CHAQSDF_HULL_SEQ") != nullptr;
    t.merge_plain = getenv("CHAQSDF_MERGE_PLAIN") != nullptr;
    t.hull_parts = num("CHAQSDF_HULL_PARTS");
    t.row8_wide = getenv("CHAQSDF_ROW8_WIDE") != nullptr;
    t.pipe_rows = num("CHAQSDF_PIPE_ROWS");
    t.pipe_first = num("CHAQSDF_PIPE_FIRST");
    if (const char* m = getenv("CHAQSDF_DT2D")) t.dt2d_mode = m[0] == 's' ? 1 : (m[0] == 'e' ? 2 : 0);
    if (const char* m = getenv("CHAQSDF_PAINT")) t.paint = m[0] == 'l' ? 1 : (m[0] == 'b' ? 3 : 0);  // "level": the level-tracking kernel; "bucket": never the direct table
    if (const char* m = getenv("CHAQSDF_GLYPH")) t.glyph = m[0] == '0' || m[0] == 'n' ? 1 : 2;
    return t;
}
Tuning g_tuning = read_tuning();

struct DevInfo {
    int sms = 0;
    int max_smem_optin = 0;
};
std::mutex g_mu;
std::map<int, DevInfo> g_dev;
struct Tables {
    QuantTable* dq = nullptr;   // device copies
    LevelTable* dl = nullptr;
    LevelTable hl;              // host copy (cap_eff decides the geometry of the banded path)
    uint8_t* dlut = nullptr;    // cap_eff <= kGlyphMaxC: byte for (class, clamped D2), for the fused small-image kernel
    uint8_t* dlutb = nullptr;   // cap_eff <= 256: the same table at its own stride, for the banded paint kernel
    int lutb_stride = 0;        // bytes per class (c2 + 1 rounded up to 16)
    uint8_t* dlutk = nullptr;   // bucketed form: per class nbk 32-bit step masks, then nbk first bytes (paint.cuh)
    int lutk_stride = 0, lutk_nbk = 0;
};
std::map<std::tuple<int, size_t, int>, Tables> g_tables;

int device_info(DevInfo* out) {
    int dev = 0;
    CQ_CUDA(cudaGetDevice(&dev));
    std::lock_guard<std::mutex> lk(g_mu);
    auto it = g_dev.find(dev);
    if (it == g_dev.end()) {
        DevInfo d;
        CQ_CUDA(cudaDeviceGetAttribute(&d.sms, cudaDevAttrMultiProcessorCount, dev));
        CQ_CUDA(cudaDeviceGetAttribute(&d.max_smem_optin, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev));
        it = g_dev.emplace(dev, d).first;
    }
    *out = it->second;
    return CHAQSDF_OK;
}

// per-(device, spread, asymmetric) quantisation thresholds, built on the host from the reference's float
// formula (tables.h) and kept on the device for the life of the process.
int get_table(size_t spread, int asymmetric, const Tables** out) {
    int dev = 0;
    CQ_CUDA(cudaGetDevice(&dev));
    std::lock_guard<std::mutex> lk(g_mu);
    auto key = std::make_tuple(dev, spread, asymmetric ? 1 : 0);
    auto it = g_tables.find(key);
    if (it == g_tables.end()) {
        QuantTable h;
        build_quant_table(spread, asymmetric, &h);
        Tables t;
        build_level_table(h, &t.hl);
        CQ_CUDA(cudaMalloc(&t.dq, sizeof(QuantTable)));
        CQ_CUDA(cudaMemcpy(t.dq, &h, sizeof(QuantTable), cudaMemcpyHostToDevice));
        CQ_CUDA(cudaMalloc(&t.dl, sizeof(LevelTable)));
        CQ_CUDA(cudaMemcpy(t.dl, &t.hl, sizeof(LevelTable), cudaMemcpyHostToDevice));
        if (t.hl.cap_eff <= kGlyphMaxC) {
            std::string lut((size_t)2 * kGlyphLutStride, '\0');
            for (int cls = 0; cls < 2; ++cls)
                for (int d2 = 0; d2 <= t.hl.c2; ++d2) {
                    const int m = level_of(t.hl.thr[cls], d2);
                    lut[(size_t)cls * kGlyphLutStride + d2] = (char)(uint8_t)(cls ? m : 255 - m);
                }
            CQ_CUDA(cudaMalloc(&t.dlut, lut.size()));
            CQ_CUDA(cudaMemcpy(t.dlut, lut.data(), lut.size(), cudaMemcpyHostToDevice));
        }
        if (t.hl.cap_eff <= 256) {
            t.lutb_stride = (int)align_up((size_t)t.hl.c2 + 1, 16);
            std::string lut((size_t)2 * t.lutb_stride, '\0');
            for (int cls = 0; cls < 2; ++cls) {
                int m = 0;  // thresholds rise with the level: walk them once
                for (int d2 = 0; d2 <= t.hl.c2; ++d2) {
                    while (m < 255 && t.hl.thr[cls][m + 1] <= d2) ++m;
                    lut[(size_t)cls * t.lutb_stride + d2] = (char)(uint8_t)(cls ? m : 255 - m);
                }
            }
            CQ_CUDA(cudaMalloc(&t.dlutb, lut.size()));
            CQ_CUDA(cudaMemcpy(t.dlutb, lut.data(), lut.size(), cudaMemcpyHostToDevice));
            // bucketed form, possible when the byte never moves by more than one step (inside up, outside down):
            // per class nbk entries of {32-bit step mask, first byte} (paint2.cuh)
            const int nbk = (t.hl.c2 >> 5) + 1;
            const int kstride = (int)align_up((size_t)nbk * 8, 16);
            std::string kb((size_t)2 * kstride, '\0');
            bool ok = true;
            for (int cls = 0; cls < 2 && ok; ++cls) {
                const uint8_t* bytes = (const uint8_t*)lut.data() + (size_t)cls * t.lutb_stride;
                uint32_t* ent = (uint32_t*)(&kb[(size_t)cls * kstride]);
                for (int i = 0; i < nbk; ++i) {
                    uint32_t m = 0;
                    for (int j = 1; j < 32 && i * 32 + j <= t.hl.c2; ++j) {
                        const int d = (int)bytes[i * 32 + j] - (int)bytes[i * 32 + j - 1];
                        if (d == (cls ? 1 : -1)) m |= 1u << j;
                        else if (d != 0) ok = false;
                    }
                    ent[2 * i] = m;
                    ent[2 * i + 1] = bytes[i * 32];
                }
            }
            if (ok) {
                t.lutk_stride = kstride; t.lutk_nbk = nbk;
                CQ_CUDA(cudaMalloc(&t.dlutk, kb.size()));
                CQ_CUDA(cudaMemcpy(t.dlutk, kb.data(), kb.size(), cudaMemcpyHostToDevice));
            }
        }
        it = g_tables.emplace(key, t).first;
    }
    *out = &it->second;
    return CHAQSDF_OK;
}

int check_dims(size_t w, size_t h, size_t spread) {
    if (w == 0 || h == 0) return fail(CHAQSDF_ERR_ARG, "empty image");
    if (w > CHAQSDF_MAX_DIM || h > CHAQSDF_MAX_DIM) return fail(CHAQSDF_ERR_ARG, "image side exceeds 32768");
    if (spread == 0) return fail(CHAQSDF_ERR_ARG, "spread must be a positive integer");  // sdfgen.c:233-236
    if (spread >= 32767 && (w > 32767 || h > 32767))
        return fail(CHAQSDF_ERR_ARG, "spread >= 32767 on a 32768-pixel side is not representable");
    return CHAQSDF_OK;
}

int cap_for(size_t spread) { return spread + 1 > 0x7fff ? 0x7fff : (int)(spread + 1); }

// ---- intermediate formats --------------------------------------------------------------------------------
// banded (spread <= 256): per row [Wp bytes g8 | Wp/8 bytes class bits | one busy byte per 128-column strip,
//                        padded to 16], Wp = w rounded up to 128
// banded, 257..1024    : per row [2*Wp bytes g16 | Wp/8 bytes class bits]
// general               : per row w uint16 (bit 15 class, bits 0..14 row distance saturated at spread+1)
// banded path: hull build / merge / paint over records. Row distances are bytes up to spread 256, uint16 up to 1024.
bool banded(size_t spread) { return spread <= 1024; }
bool banded_wide(size_t spread) { return spread > 256 && spread <= 1024; }
int banded_wp(size_t w) { return (int)align_up(w, 128); }
size_t banded_rs(size_t w, size_t spread) {
    const size_t wp = (size_t)banded_wp(w);
    if (banded_wide(spread)) return 2 * wp + wp / 8;
    return wp + wp / 8 + align_up(wp / 128, 16);
}
// scratch of the banded pass 2 (hull.cuh, paint.cuh): hull bitmask + transposed class bits (one word per
// (32-row block, column) each) + one flag byte per (block, 32-column group)
constexpr size_t kWorkBytes = 256;  // work-item counters of the persistent paint kernels (one per column part)
size_t hull_words_bytes(size_t n_img, size_t w, size_t buf_rows) {
    const size_t ncg = (w + 31) / 32, nblk = (buf_rows + 31) / 32;
    return align_up(n_img * nblk * ncg * 32 * sizeof(uint32_t), 256);
}
size_t hull_ws_bytes(size_t n_img, size_t w, size_t buf_rows) {
    const size_t nblk = (buf_rows + 31) / 32, ncg = (w + 31) / 32;
    return 2 * hull_words_bytes(n_img, w, buf_rows) + align_up(n_img * nblk * ncg, 256) + kWorkBytes;
}
size_t inter_row_bytes(size_t w, size_t spread) { return banded(spread) ? banded_rs(w, spread) : w * sizeof(uint16_t); }

// launch geometry of the general pass 2
struct EnvPlan {
    int cap, nw, band_rows, nbands;
    bool wide, global_ring;
    size_t smem_bytes, ring_bytes;
    dim3 grid;
};

int plan_envelope(size_t w, size_t buf_rows, size_t out_rows, size_t n_img, int cap, EnvPlan* pl) {
    DevInfo di;
    if (int rc = device_info(&di)) return rc;
    pl->cap = cap;
    pl->nw = ring_words(cap, (int)buf_rows);
    pl->wide = cap > 1023;
    const int lag = cap - 1 < (int)buf_rows ? cap - 1 : (int)buf_rows;
    const size_t colblocks = (w + kEnvThreads - 1) / kEnvThreads;
    // aim for ~1536 resident threads per SM; keep the halo re-read (2*lag rows per band) under ~50 %
    long long target_threads = (long long)di.sms * 1536;
    long long nb = target_threads / (long long)(colblocks * kEnvThreads * n_img);
    if (nb < 1) nb = 1;
    long long band = ((long long)out_rows + nb - 1) / nb;
    long long min_band = 4LL * lag;
    if (g_tuning.band_rows > 0) band = g_tuning.band_rows, min_band = 1;
    if (band < min_band) band = min_band;
    if (band > (long long)out_rows) band = (long long)out_rows;
    if (band < 1) band = 1;
    pl->band_rows = (int)band;
    pl->nbands = (int)(((long long)out_rows + band - 1) / band);
    const size_t ring_smem = (size_t)2 * pl->nw * kEnvThreads * sizeof(uint32_t);
    pl->global_ring = ring_smem > 96 * 1024;
    pl->smem_bytes = 516 * sizeof(uint32_t) + (pl->global_ring ? 0 : ring_smem);
    pl->grid = dim3((unsigned)colblocks, (unsigned)pl->nbands, (unsigned)n_img);
    pl->ring_bytes = pl->global_ring
                         ? (size_t)2 * pl->nw * colblocks * kEnvThreads * pl->nbands * n_img * sizeof(uint32_t)
                         : 0;
    if (n_img > 65535 || pl->nbands > 65535) return fail(CHAQSDF_ERR_ARG, "batch or band count exceeds grid limits");
    return CHAQSDF_OK;
}

template <bool WIDE, bool GR, bool FIELDS>
int launch_env_t(const EnvPlan& pl, const EnvParams& p, cudaStream_t st) {
    auto k = envelope_kernel<WIDE, GR, FIELDS>;
    if (pl.smem_bytes > 48 * 1024)
        CQ_CUDA(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)pl.smem_bytes));
    k<<<pl.grid, kEnvThreads, pl.smem_bytes, st>>>(p);
    CQ_LAUNCHED("envelope_kernel", pl.grid.x * pl.grid.y * pl.grid.z, pl.smem_bytes);
    return CHAQSDF_OK;
}

int launch_env(const EnvPlan& pl, const EnvParams& p, bool fields, cudaStream_t st) {
    const int sel = (pl.wide ? 4 : 0) | (pl.global_ring ? 2 : 0) | (fields ? 1 : 0);
    switch (sel) {
        case 0: return launch_env_t<false, false, false>(pl, p, st);
        case 1: return launch_env_t<false, false, true>(pl, p, st);
        case 2: return launch_env_t<false, true, false>(pl, p, st);
        case 3: return launch_env_t<false, true, true>(pl, p, st);
        case 4: return launch_env_t<true, false, false>(pl, p, st);
        case 5: return launch_env_t<true, false, true>(pl, p, st);
        case 6: return launch_env_t<true, true, false>(pl, p, st);
        default: return launch_env_t<true, true, true>(pl, p, st);
    }
}

int launch_rowscan(const uint8_t* d_img, size_t w, size_t rows, size_t stride, size_t offset, int test_above, int cap,
                   uint16_t* d_inter, cudaStream_t st) {
    RowScanParams rp{};
    rp.img = d_img; rp.inter = d_inter; rp.rows = (long long)rows; rp.w = (int)w;
    rp.stride = (int)stride; rp.offset = (int)offset; rp.above = test_above ? 1 : 0; rp.cap = cap;
    rp.nwords = (int)((w + 31) / 32);
    const size_t smem = (size_t)kRowWarps * rp.nwords * sizeof(uint32_t);
    const unsigned blocks = (unsigned)((rows + kRowWarps - 1) / kRowWarps);
    rowscan_kernel<<<blocks, kRowWarps * 32, smem, st>>>(rp);
    CQ_LAUNCHED("rowscan_kernel", blocks, smem);
    return CHAQSDF_OK;
}

// pass 1 of the banded path for spreads 257..1024: the general scan, writing records with 16-bit row distances
int launch_rowscan_rec(const uint8_t* d_img, size_t w, size_t rows, size_t stride, size_t offset, int test_above,
                       size_t spread, uint8_t* d_inter, cudaStream_t st) {
    RowScanParams rp{};
    rp.img = d_img; rp.inter = (uint16_t*)d_inter; rp.rows = (long long)rows; rp.w = (int)w;
    rp.stride = (int)stride; rp.offset = (int)offset; rp.above = test_above ? 1 : 0; rp.cap = (int)spread + 1;
    rp.nwords = (int)((w + 31) / 32);
    rp.record = 1; rp.wp = banded_wp(w); rp.rs_bytes = (long long)banded_rs(w, spread);
    const size_t smem = (size_t)kRowWarps * rp.nwords * sizeof(uint32_t);
    const unsigned blocks = (unsigned)((rows + kRowWarps - 1) / kRowWarps);
    rowscan_kernel<<<blocks, kRowWarps * 32, smem, st>>>(rp);
    CQ_LAUNCHED("rowscan_kernel<records>", blocks, smem);
    return CHAQSDF_OK;
}

// ---- banded path -------------------------------------------------------------------------------------------
int launch_rowscan8(const uint8_t* d_img, size_t w, size_t rows, size_t stride, size_t offset, int test_above,
                    size_t spread, uint8_t* d_inter, cudaStream_t st) {
    RowScan8Params rp;
    rp.img = d_img; rp.inter = d_inter; rp.rows = (long long)rows; rp.rs = (long long)banded_rs(w, spread);
    rp.wp = banded_wp(w); rp.w = (int)w; rp.stride = (int)stride; rp.offset = (int)offset;
    rp.above = test_above ? 1 : 0; rp.cap = spread + 1 > 256 ? 256 : (int)(spread + 1);
    rp.nwords = (int)((w + 31) / 32);
    rp.nstrips = rp.wp / 128;
    if (rp.nwords <= 16 && !g_tuning.row8_wide) {  // several rows per warp
        int shift = 0;
        while ((1 << shift) < rp.nwords) ++shift;
        const size_t rows_per_cta = (size_t)kRow8Warps * (32u >> shift);
        const unsigned blocks = (unsigned)((rows + rows_per_cta - 1) / rows_per_cta);
        rowscan8_narrow_kernel<<<blocks, kRow8Warps * 32, 0, st>>>(rp, shift);
        CQ_LAUNCHED("rowscan8_narrow_kernel", blocks, 0);
        return CHAQSDF_OK;
    }
    const size_t smem = (size_t)kRow8Warps * (rp.nwords + 1) * sizeof(uint32_t);
    const unsigned blocks = (unsigned)((rows + kRow8Warps - 1) / kRow8Warps);
    rowscan8_kernel<<<blocks, kRow8Warps * 32, smem, st>>>(rp);
    CQ_LAUNCHED("rowscan8_kernel", blocks, smem);
    return CHAQSDF_OK;
}

// ---- fused small-image path (glyph.cuh) -----------------------------------------------------------------------
// Does the batch go through one CTA per image? Needs: byte-contiguous pixels, width a multiple of 32 up to 256,
// C <= 33, and the image's planes in shared memory.
int glyph_lmax(int L) { return L <= 8 ? 8 : (L <= 16 ? 16 : 32); }
bool glyph_fits(size_t w, size_t h, size_t stride, const Tables* tab, const DevInfo& di) {
    if (g_tuning.glyph == 1) return false;
    if (stride != 1 || w % 32 != 0 || w > 256 || w < 32 || h < 1 || h > 1024 || !tab->dlut) return false;
    if (tab->hl.cap_eff > kGlyphMaxC) return false;
    return glyph_smem((int)w, (int)h, glyph_lmax(tab->hl.cap_eff - 1)) <= (size_t)di.max_smem_optin;
}
template <int LMAX>
int launch_glyph_t(const GlyphParams& p, size_t smem, cudaStream_t st) {
    static int raised[64] = {0};
    int dev = 0;
    CQ_CUDA(cudaGetDevice(&dev));
    {
        std::lock_guard<std::mutex> lk(g_mu);
        if (dev < 64 && (int)smem > raised[dev]) {
            CQ_CUDA(cudaFuncSetAttribute(glyph_kernel<LMAX>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
            raised[dev] = (int)smem;
        }
    }
    glyph_kernel<LMAX><<<(unsigned)p.n, kGlyphThreads, smem, st>>>(p);
    CQ_LAUNCHED("glyph_kernel", p.n, smem);
    return CHAQSDF_OK;
}
int launch_glyph(const uint8_t* d_imgs, size_t n, size_t w, size_t h, int test_above, const Tables* tab, uint8_t* d_out,
                 cudaStream_t st) {
    GlyphParams p{};
    p.img = d_imgs; p.out = d_out; p.lut = tab->dlut;
    p.w = (int)w; p.h = (int)h; p.n = (int)n; p.wq = (int)(w / 32); p.hb = (int)((h + 31) / 32);
    p.C = tab->hl.cap_eff; p.C2 = tab->hl.c2; p.above = test_above ? 1 : 0;
    const int lmax = glyph_lmax(p.C - 1);
    const size_t smem = glyph_smem(p.w, p.h, lmax);
    if (lmax == 8) return launch_glyph_t<8>(p, smem, st);
    if (lmax == 16) return launch_glyph_t<16>(p, smem, st);
    return launch_glyph_t<32>(p, smem, st);
}

// optional per-kernel timing (chaqsdf_profile_enable): events around the launches of the banded pass 2
constexpr int kProfSets = 64;
bool g_prof_on = false;
int g_prof_n = 0;  // profiled calls since it was enabled (the first kProfSets are kept)
cudaEvent_t g_prof_ev[kProfSets][4] = {};
void prof_mark(int i, cudaStream_t st) {
    if (!g_prof_on || g_prof_n >= kProfSets) return;
    cudaEvent_t& e = g_prof_ev[g_prof_n][i];
    if (!e && cudaEventCreate(&e) != cudaSuccess) return;
    cudaEventRecord(e, st);
    if (i == 3) ++g_prof_n;
}

// helper streams for the column parts of one pass-2 call (per device, created on first use, never destroyed)
constexpr int kMaxParts = 16;
struct PartStreams {
    std::mutex mu;  // held while one call enqueues its parts (the enqueue is short; the work itself overlaps)
    cudaStream_t s[kMaxParts - 1] = {};
    cudaEvent_t fork = nullptr, join[kMaxParts - 1] = {};
};
std::map<int, PartStreams*> g_part_streams;
int part_streams(PartStreams** out) {
    int dev = 0;
    CQ_CUDA(cudaGetDevice(&dev));
    std::lock_guard<std::mutex> lk(g_mu);
    PartStreams*& ps = g_part_streams[dev];
    if (!ps) {
        PartStreams* n = new PartStreams();
        CQ_CUDA(cudaEventCreateWithFlags(&n->fork, cudaEventDisableTiming));
        for (int i = 0; i < kMaxParts - 1; ++i) {
            CQ_CUDA(cudaStreamCreateWithFlags(&n->s[i], cudaStreamNonBlocking));
            CQ_CUDA(cudaEventCreateWithFlags(&n->join[i], cudaEventDisableTiming));
        }
        ps = n;
    }
    *out = ps;
    return CHAQSDF_OK;
}
#define CQ_CUDA_UNLOCK(expr, ps)                                                                        \
    do {                                                                                                \
        cudaError_t e_ = (expr);                                                                        \
        if (e_ != cudaSuccess) {                                                                        \
            (ps)->mu.unlock();                                                                          \
            return fail(CHAQSDF_ERR_CUDA, std::string(#expr) + ": " + cudaGetErrorString(e_));          \
        }                                                                                               \
    } while (0)

#ifdef CQ_BUILD_TRACE
const uint32_t* g_dbg_order = nullptr;  // experiment: an externally supplied order of the build grid
unsigned g_dbg_order_n = 0;
#endif
// banded pass 2 as build / merge / paint (hull.cuh, paint.cuh)
template <bool WIDE>
int launch_hull_t(const uint8_t* d_inter, size_t w, size_t buf_rows, size_t out_begin, size_t out_end, size_t n_img,
                const Tables* tab, uint8_t* d_out, void* d_ws, size_t ws_bytes, cudaStream_t st, bool one_stream = false,
                int col_begin = -1, int col_end = -1) {
    DevInfo di;
    if (int rc = device_info(&di)) return rc;
    if (!d_ws || ws_bytes < hull_ws_bytes(n_img, w, buf_rows)) return fail(CHAQSDF_ERR_WORKSPACE, "workspace too small");
    HullParams p{};
    p.inter = d_inter; p.out = d_out; p.table = tab->dl;
    p.hull = (uint32_t*)d_ws;
    p.clsT = (uint32_t*)((uint8_t*)d_ws + hull_words_bytes(n_img, w, buf_rows));
    p.flags = (uint8_t*)d_ws + 2 * hull_words_bytes(n_img, w, buf_rows);
    unsigned* const work0 = (unsigned*)((uint8_t*)d_ws + hull_ws_bytes(n_img, w, buf_rows) - kWorkBytes);
    // staged paint kernel: the direct table when it fits beside the warps' staging, else the bucketed one
    const bool lut_direct = !WIDE && tab->dlutb && hull_paint2_smem(2 * tab->lutb_stride) <= (size_t)di.max_smem_optin &&
                            g_tuning.paint != 3;
    const bool lut_bucket = !WIDE && !lut_direct && tab->dlutk && hull_paint2_smem(2 * tab->lutk_stride) <= (size_t)di.max_smem_optin;
    const bool use_lut = (lut_direct || lut_bucket) && g_tuning.paint != 1;
    p.lut = lut_direct ? tab->dlutb : tab->dlutk; p.lut_stride = lut_direct ? tab->lutb_stride : tab->lutk_stride;
    p.lut_nbk = lut_direct ? 0 : tab->lutk_nbk; p.work = nullptr;
    if (!use_lut) p.lut = nullptr;
    p.wp = banded_wp(w); p.w = (int)w; p.buf_rows = (int)buf_rows;
    p.rs = (long long)banded_rs(w, WIDE ? 1024 : 256); p.cls_off = WIDE ? 2 * p.wp : p.wp;
    p.nblk = (int)((buf_rows + 31) / 32);
    p.ncg = (int)((w + 31) / 32); p.wc = p.ncg * 32;
    // column groups [cgb, cge) are processed: the whole width, or the groups covering columns [col_begin, col_end)
    const int cgb = col_begin >= 0 ? col_begin / 32 : 0, cge = col_end >= 0 ? (col_end + 31) / 32 : p.ncg;
    const int ncgr = cge - cgb;
    p.nstrips = (ncgr + kHullWarps - 1) / kHullWarps; p.n_img = (int)n_img;
    p.nstrips_e = (ncgr + kEmitWarps - 1) / kEmitWarps;
    p.cg0 = cgb; p.cg1 = cge;
    p.out_begin = (int)out_begin; p.out_end = (int)out_end;
    p.L = tab->hl.cap_eff - 1; p.c2 = tab->hl.c2;
    // run lengths: enough CTAs to fill the machine a few times over, runs no longer than 512 rows (build) / 1024 (emit)
    const long long target = (long long)di.sms * 24;
    const long long strips = (long long)n_img * p.nstrips;
    long long pb = (strips * p.nblk + target - 1) / target;
    pb = pb < 2 ? 2 : (pb > 16 ? 16 : pb);
    // runs of >= 2L rows need only two rounds of junction repair; shorter runs cost more in rounds than they gain
    const long long pb_par = (2 * p.L + 31) / 32;
    if (pb < pb_par) pb = pb_par;
    if (g_tuning.hull_pb > 0) pb = g_tuning.hull_pb;
    if (pb < 1) pb = 1;
    if (pb > hull_max_run_blocks(WIDE)) pb = hull_max_run_blocks(WIDE);
    if (pb > p.nblk) pb = p.nblk;
    p.pb = (int)pb; p.nruns_b = (p.nblk + p.pb - 1) / p.pb;
    // Few build CTAs (a band of a pipelined or sharded image): what matters is the chain of pushes down one run, so
    // four warps share each run and the CTA repairs its own inner junctions. Many CTAs: one warp per run is cheaper.
    long long nsub = strips * p.nruns_b <= (long long)di.sms * 16 ? 4 : 1;
    if (g_tuning.hull_sub > 0) nsub = g_tuning.hull_sub;
    nsub = nsub >= 4 && p.nblk >= 8 ? 4 : (nsub == 2 && p.nblk >= 4 ? 2 : 1);
    if (nsub == 4 && p.pb < 8 && g_tuning.hull_pb <= 0) { p.pb = 8; p.nruns_b = (p.nblk + p.pb - 1) / p.pb; }  // sub-runs of >= 2 blocks
    p.sb = (p.pb + (int)nsub - 1) / (int)nsub; p.nsub = (int)nsub;
    const int bfirst = p.out_begin >> 5, nbo = ((p.out_end - 1) >> 5) - bfirst + 1;
    // paint runs: starting one is cheap (the cursor is found by a local search), so short runs -- more CTAs than the
    // build, evening out dense and empty areas -- win: ~8 blocks at 16384^2
    const long long target_e = (long long)di.sms * 56;
    const long long strips_e = (long long)n_img * p.nstrips_e;
    long long eb = (strips_e * nbo + target_e - 1) / target_e;
    eb = eb < 2 ? 2 : (eb > 32 ? 32 : eb);
    if (use_lut) {  // persistent kernel, one item per (column group, run): ~10 items per resident warp (measured: 8-12 blocks per run at 16384^2)
        eb = ((long long)n_img * ncgr * nbo) / ((long long)di.sms * kP2Warps * 10);
        eb = eb < 4 ? 4 : (eb > 32 ? 32 : eb);
    }
    if (g_tuning.hull_eb > 0) eb = g_tuning.hull_eb;
    if (eb < 1) eb = 1;
    if (eb > nbo) eb = nbo;
    p.eb = (int)eb; p.nruns_e = (nbo + p.eb - 1) / p.eb;
    const long long gb = strips * p.nruns_b, ge = (long long)n_img * p.nstrips_e * p.nruns_e;
    if (gb > 0x7fffffffLL || ge > 0x7fffffffLL) return fail(CHAQSDF_ERR_ARG, "too many runs for one grid");
    const size_t smem_b = hull_build_smem(p.pb, p.nsub, WIDE), smem_e = hull_paint_smem();
    const size_t smem_l = use_lut ? hull_paint2_smem(2 * p.lut_stride) : 0;
    if (use_lut) {
        if (lut_direct) CQ_CUDA(cudaFuncSetAttribute(hull_paint2_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_l));
        else CQ_CUDA(cudaFuncSetAttribute(hull_paint2_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_l));
    }
    if (smem_b > 48 * 1024) {
        CQ_CUDA(cudaFuncSetAttribute(hull_build_kernel<WIDE>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_b));
        CQ_CUDA(cudaFuncSetAttribute(hull_build4_kernel<WIDE>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_b));
        CQ_CUDA(cudaFuncSetAttribute(hull_build2_kernel<WIDE>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_b));
    }
    // Column parts on separate streams: the chains build -> junction rounds -> paint of different parts overlap, so
    // the tail of one part's build (SMs idle while the last heavy CTAs finish) and its latency-bound junction rounds
    // are filled with other parts' work. One part when the grid is small, when profiling per kernel, or when asked.
    int nparts = 1;
    if (n_img == 1 && !g_prof_on && p.nblk >= 4) nparts = ncgr >= 256 ? 8 : (ncgr >= 64 ? 4 : 1);
    // the persistent paint kernel fills the machine by itself: two parts are enough to overlap one part's junction rounds
    // with the other's build (measured at 16384^2: 1 part 1.19 ms, 2 parts 1.18, 4 parts 1.19, 8 parts 1.29)
    if (use_lut && nparts > 1) nparts = ncgr >= 256 ? 2 : 1;
    if (g_tuning.hull_parts > 0) nparts = (int)g_tuning.hull_parts;
    if (nparts > kMaxParts) nparts = kMaxParts;
    if (nparts > ncgr / 4) nparts = ncgr / 4 > 0 ? ncgr / 4 : 1;
    if (n_img != 1 || g_prof_on || one_stream) nparts = 1;
    PartStreams* ps = nullptr;
    if (nparts > 1) {
        if (int rc = part_streams(&ps)) return rc;
        ps->mu.lock();
        CQ_CUDA_UNLOCK(cudaEventRecord(ps->fork, st), ps);
    }
    const int per = ((ncgr + nparts - 1) / nparts + 3) & ~3;  // parts are multiples of 4 column groups
    int rc_all = CHAQSDF_OK;
    for (int k = 0; k < nparts && rc_all == CHAQSDF_OK; ++k) {
        HullParams q = p;
        q.cg0 = cgb + k * per; q.cg1 = q.cg0 + per < cge ? q.cg0 + per : cge;
        if (q.cg0 >= q.cg1) break;
        q.work = use_lut ? work0 + k : nullptr;
        cudaStream_t sk = k == 0 ? st : ps->s[k - 1];
        if (k > 0) CQ_CUDA_UNLOCK(cudaStreamWaitEvent(sk, ps->fork, 0), ps);
        q.nstrips = (q.cg1 - q.cg0 + kHullWarps - 1) / kHullWarps;
        q.nstrips_e = (q.cg1 - q.cg0 + kEmitWarps - 1) / kEmitWarps;
        const long long gbk = (long long)n_img * q.nstrips * q.nruns_b, gek = (long long)n_img * q.nstrips_e * q.nruns_e;
        const long long wpart = (long long)(q.cg1 * 32 < q.w ? q.cg1 * 32 : q.w) - q.cg0 * 32;
        rc_all = [&]() -> int {
            if (k == 0) prof_mark(0, st);
#ifdef CQ_BUILD_TRACE
            q.order = (g_dbg_order && gbk == (long long)g_dbg_order_n && q.nsub == 1) ? g_dbg_order : nullptr;
#endif
            if (q.nsub == 4) hull_build4_kernel<WIDE><<<(unsigned)gbk, kHullWarps * 32 * 4, smem_b, sk>>>(q);
            else if (q.nsub == 2) hull_build2_kernel<WIDE><<<(unsigned)gbk, kHullWarps * 32 * 2, smem_b, sk>>>(q);
            else hull_build_kernel<WIDE><<<(unsigned)gbk, kHullWarps * 32, smem_b, sk>>>(q);
            CQ_LAUNCHED(q.nsub == 4 ? "hull_build4_kernel" : "hull_build_kernel", gbk, smem_b);
            if (k == 0) prof_mark(1, st);
            if (q.nruns_b > 1 && !g_tuning.hull_seq) {
                const int njun = q.nruns_b - 1;
                int nphase = hull_merge_phases(q.L, q.pb);
                nphase = nphase < 1 ? 1 : (nphase > njun ? njun : nphase);
                q.merge_nphase = nphase;
                for (int phase = 0; phase < nphase; ++phase) {
                    const int nmine = (njun - phase + nphase - 1) / nphase;
                    if (nmine <= 0) continue;
                    q.merge_phase = phase;
                    const long long thr = (long long)n_img * nmine * wpart;
                    if (!WIDE && !g_tuning.merge_plain) {  // the repair's window of hull words fits the shared-memory copy
                        const unsigned gm = (unsigned)((thr + kMergeThreads - 1) / kMergeThreads);
                        hull_merge_win_kernel<WIDE><<<gm, kMergeThreads, 0, sk>>>(q);
                        CQ_LAUNCHED("hull_merge_win_kernel", gm, 0);
                    } else {
                        const unsigned gm = (unsigned)((thr + 255) / 256);
                        hull_merge_par_kernel<WIDE><<<gm, 256, 0, sk>>>(q);
                        CQ_LAUNCHED("hull_merge_par_kernel", gm, 0);
                    }
                }
            } else if (q.nruns_b > 1) {
                const long long cols = (long long)n_img * wpart;
                const unsigned gm = (unsigned)((cols + 255) / 256);
                hull_merge_kernel<WIDE><<<gm, 256, 0, sk>>>(q);
                CQ_LAUNCHED("hull_merge_kernel", gm, 0);
            }
            if (k == 0) prof_mark(2, st);
            if (use_lut) {
                const long long items = (long long)n_img * (q.cg1 - q.cg0) * q.nruns_e;
                long long gl = (items + kP2Warps - 1) / kP2Warps;
                const long long cap = (di.sms + nparts - 1) / nparts;
                gl = gl > cap ? cap : gl;
                if (lut_direct) hull_paint2_kernel<true><<<(unsigned)gl, kP2Warps * 32, smem_l, sk>>>(q);
                else hull_paint2_kernel<false><<<(unsigned)gl, kP2Warps * 32, smem_l, sk>>>(q);
                CQ_LAUNCHED("hull_paint2_kernel", gl, smem_l);
            } else {
                hull_paint_kernel<WIDE><<<(unsigned)gek, kEmitWarps * 32, smem_e, sk>>>(q);
                CQ_LAUNCHED("hull_paint_kernel", gek, smem_e);
            }
            if (k == 0) prof_mark(3, st);
            return CHAQSDF_OK;
        }();
        if (rc_all == CHAQSDF_OK && k > 0) {
            if (cudaEventRecord(ps->join[k - 1], sk) != cudaSuccess || cudaStreamWaitEvent(st, ps->join[k - 1], 0) != cudaSuccess)
                rc_all = fail(CHAQSDF_ERR_CUDA, "joining the column parts failed");
        }
    }
    if (ps) ps->mu.unlock();
    return rc_all;
}

int launch_hull(size_t spread, const uint8_t* d_inter, size_t w, size_t buf_rows, size_t out_begin, size_t out_end, size_t n_img,
                const Tables* tab, uint8_t* d_out, void* d_ws, size_t ws_bytes, cudaStream_t st, bool one_stream = false,
                int col_begin = -1, int col_end = -1) {
    if (tab->hl.cap_eff > (banded_wide(spread) ? 1025 : 256)) return fail(CHAQSDF_ERR_ARG, "internal: banded path beyond its reach");
    if (banded_wide(spread)) return launch_hull_t<true>(d_inter, w, buf_rows, out_begin, out_end, n_img, tab, d_out, d_ws, ws_bytes, st, one_stream, col_begin, col_end);
    return launch_hull_t<false>(d_inter, w, buf_rows, out_begin, out_end, n_img, tab, d_out, d_ws, ws_bytes, st, one_stream, col_begin, col_end);
}
int launch_pass1_banded(const uint8_t* d_img, size_t w, size_t rows, size_t stride, size_t offset, int test_above,
                        size_t spread, uint8_t* d_inter, cudaStream_t st) {
    if (banded_wide(spread)) return launch_rowscan_rec(d_img, w, rows, stride, offset, test_above, spread, d_inter, st);
    return launch_rowscan8(d_img, w, rows, stride, offset, test_above, spread, d_inter, st);
}

}  // namespace

namespace chaq {
const Tuning& tuning() { return g_tuning; }
int set_error(int code, const char* msg) { return fail(code, msg); }
void note_launch() { g_launches.fetch_add(1, std::memory_order_relaxed); }
}  // namespace chaq

// ------------------------------------------------------------------------------------------------
//                                          C ABI
// ------------------------------------------------------------------------------------------------
extern "C" {

int chaqsdf_version(void) { return 110; }
const char* chaqsdf_last_error_string(void) { return g_err.c_str(); }
unsigned long long chaqsdf_launch_count(void) { return g_launches.load(); }
void chaqsdf_reload_tuning(void) { g_tuning = read_tuning(); }
void chaqsdf_profile_enable(int on) { g_prof_on = on != 0; if (on) g_prof_n = 0; }
// returns the number of calls averaged; 0 (and an error string) if there is none or a CUDA call fails
int chaqsdf_profile_mean_ms(float ms[3]) {
    if (!ms) { fail(CHAQSDF_ERR_ARG, "null pointer"); return 0; }
    ms[0] = ms[1] = ms[2] = 0.f;
    const int n = g_prof_n < kProfSets ? g_prof_n : kProfSets;
    if (n == 0) { fail(CHAQSDF_ERR_ARG, "no profiled call of the banded pass 2 yet"); return 0; }
    float acc[3] = {0.f, 0.f, 0.f};
    for (int k = 0; k < n; ++k) {
        cudaError_t e = cudaEventSynchronize(g_prof_ev[k][3]);
        for (int i = 0; i < 3 && e == cudaSuccess; ++i) {
            float t = 0.f;
            e = cudaEventElapsedTime(&t, g_prof_ev[k][i], g_prof_ev[k][i + 1]);
            acc[i] += t / (float)n;
        }
        if (e != cudaSuccess) { fail(CHAQSDF_ERR_CUDA, std::string("profile events: ") + cudaGetErrorString(e)); return 0; }
    }
    for (int i = 0; i < 3; ++i) ms[i] = acc[i];
    return n;
}

size_t chaqsdf_inter_row_bytes(size_t w, size_t spread) {
    if (w == 0 || w > CHAQSDF_MAX_DIM || spread == 0) return 0;
    return inter_row_bytes(w, spread);
}

size_t chaqsdf_batched_workspace_bytes(size_t n, size_t w, size_t h, size_t spread) {
    if (check_dims(w, h, spread) || n == 0) return 0;
    if (banded(spread)) return align_up(n * h * banded_rs(w, spread), 256) + hull_ws_bytes(n, w, h);
    EnvPlan pl;
    if (plan_envelope(w, h, h, n, cap_for(spread), &pl)) return 0;
    return align_up(n * w * h * sizeof(uint16_t), 256) + pl.ring_bytes;
}
size_t chaqsdf_workspace_bytes(size_t w, size_t h, size_t spread) {
    return chaqsdf_batched_workspace_bytes(1, w, h, spread);
}
size_t chaqsdf_envelope_workspace_bytes(size_t w, size_t buf_rows, size_t spread) {
    if (check_dims(w, buf_rows, spread)) return 0;
    if (banded(spread)) return hull_ws_bytes(1, w, buf_rows);
    EnvPlan pl;
    if (plan_envelope(w, buf_rows, buf_rows, 1, cap_for(spread), &pl)) return 0;
    return pl.ring_bytes;
}
size_t chaqsdf_fields_workspace_bytes(size_t w, size_t h) {
    if (check_dims(w, h, 1)) return 0;
    EnvPlan pl;
    if (plan_envelope(w, h, h, 1, 0x7fff, &pl)) return 0;
    return align_up(w * h * sizeof(uint16_t), 256) + pl.ring_bytes;
}

int chaqsdf_rowscan_device(const uint8_t* d_img, size_t w, size_t rows, size_t stride, size_t offset, int test_above,
                           size_t spread, void* d_inter, void* stream) {
    if (!d_img || !d_inter) return fail(CHAQSDF_ERR_ARG, "null pointer");
    if (w == 0 || rows == 0) return fail(CHAQSDF_ERR_ARG, "empty image");
    if (w > CHAQSDF_MAX_DIM) return fail(CHAQSDF_ERR_ARG, "image side exceeds 32768");
    if (spread == 0) return fail(CHAQSDF_ERR_ARG, "spread must be a positive integer");
    if (stride == 0 || offset >= stride) return fail(CHAQSDF_ERR_ARG, "bad stride/offset");
    if (banded(spread))
        return launch_pass1_banded(d_img, w, rows, stride, offset, test_above, spread, (uint8_t*)d_inter, (cudaStream_t)stream);
    return launch_rowscan(d_img, w, rows, stride, offset, test_above, cap_for(spread), (uint16_t*)d_inter,
                          (cudaStream_t)stream);
}

int chaqsdf_envelope_device(const void* d_inter, size_t w, size_t buf_rows, size_t out_begin, size_t out_end,
                            size_t spread, int asymmetric, uint8_t* d_out, void* d_workspace, size_t workspace_bytes,
                            void* stream) {
    if (!d_inter || !d_out) return fail(CHAQSDF_ERR_ARG, "null pointer");
    if (int rc = check_dims(w, buf_rows, spread)) return rc;
    if (out_begin >= out_end || out_end > buf_rows) return fail(CHAQSDF_ERR_ARG, "bad output row range");
    const Tables* tab;
    if (int rc = get_table(spread, asymmetric, &tab)) return rc;
    if (banded(spread)) {
        return launch_hull(spread, (const uint8_t*)d_inter, w, buf_rows, out_begin, out_end, 1, tab, d_out, d_workspace,
                           workspace_bytes, (cudaStream_t)stream);
    }
    EnvPlan pl;
    if (int rc = plan_envelope(w, buf_rows, out_end - out_begin, 1, cap_for(spread), &pl)) return rc;
    if (pl.ring_bytes > workspace_bytes || (pl.ring_bytes && !d_workspace))
        return fail(CHAQSDF_ERR_WORKSPACE, "workspace too small");
    EnvParams p{};
    p.inter = (const uint16_t*)d_inter; p.w = (int)w; p.buf_rows = (int)buf_rows;
    p.out_begin = (int)out_begin; p.out_end = (int)out_end; p.band_rows = pl.band_rows;
    p.cap = pl.cap; p.nw = pl.nw; p.table = tab->dq; p.out = d_out; p.gring = (uint32_t*)d_workspace;
    return launch_env(pl, p, false, (cudaStream_t)stream);
}

int chaqsdf_envelope_cols_device(const void* d_inter, size_t w, size_t buf_rows, size_t out_begin, size_t out_end,
                                 size_t col_begin, size_t col_end, size_t spread, int asymmetric, uint8_t* d_out,
                                 void* d_workspace, size_t workspace_bytes, void* stream) {
    if (!d_inter || !d_out) return fail(CHAQSDF_ERR_ARG, "null pointer");
    if (int rc = check_dims(w, buf_rows, spread)) return rc;
    if (out_begin >= out_end || out_end > buf_rows) return fail(CHAQSDF_ERR_ARG, "bad output row range");
    if (col_begin >= col_end || col_end > w) return fail(CHAQSDF_ERR_ARG, "bad column range");
    if (!banded(spread)) return fail(CHAQSDF_ERR_ARG, "column ranges need spread <= 1024");
    const Tables* tab;
    if (int rc = get_table(spread, asymmetric, &tab)) return rc;
    return launch_hull(spread, (const uint8_t*)d_inter, w, buf_rows, out_begin, out_end, 1, tab, d_out, d_workspace,
                       workspace_bytes, (cudaStream_t)stream, false, (int)col_begin, (int)col_end);
}

// dst[c * dst_pitch + r] = src[r * src_pitch + c] (transpose.cuh)
int chaqsdf_transpose_u8_device(const uint8_t* d_src, size_t src_pitch, size_t rows, size_t cols, uint8_t* d_dst,
                                size_t dst_pitch, void* stream) {
    if (!d_src || !d_dst) return fail(CHAQSDF_ERR_ARG, "null pointer");
    if (rows == 0 || cols == 0 || src_pitch < cols || dst_pitch < rows) return fail(CHAQSDF_ERR_ARG, "bad transpose shape");
    const size_t tr = (rows + kTposeTile - 1) / kTposeTile, tc = (cols + kTposeTile - 1) / kTposeTile;
    if (tr > 65535 || tc > 0x7fffffff) return fail(CHAQSDF_ERR_ARG, "transpose shape too large");
    TposeParams tp;
    tp.src = d_src; tp.dst = d_dst; tp.src_pitch = (long long)src_pitch; tp.dst_pitch = (long long)dst_pitch;
    tp.rows = (int)rows; tp.cols = (int)cols;
    tp.vec_in = src_pitch % 16 == 0 && reinterpret_cast<uintptr_t>(d_src) % 16 == 0;
    tp.vec_out = dst_pitch % 16 == 0 && reinterpret_cast<uintptr_t>(d_dst) % 16 == 0;
    transpose_u8_kernel<<<dim3((unsigned)tc, (unsigned)tr), kTposeThreads, 0, (cudaStream_t)stream>>>(tp);
    CQ_LAUNCHED("transpose_u8_kernel", tc * tr, 0);
    return CHAQSDF_OK;
}

int chaqsdf_u8_batched_device(const uint8_t* d_imgs, size_t n, size_t w, size_t h, size_t stride, size_t offset,
                              int test_above, size_t spread, int asymmetric, uint8_t* d_out, void* d_workspace,
                              size_t workspace_bytes, void* stream) {
    if (!d_imgs || !d_out || !d_workspace) return fail(CHAQSDF_ERR_ARG, "null pointer");
    if (n == 0) return fail(CHAQSDF_ERR_ARG, "empty batch");
    if (int rc = check_dims(w, h, spread)) return rc;
    if (stride == 0 || offset >= stride) return fail(CHAQSDF_ERR_ARG, "bad stride/offset");
    const Tables* tab;
    if (int rc = get_table(spread, asymmetric, &tab)) return rc;
    cudaStream_t st = (cudaStream_t)stream;
    if (banded(spread)) {
        DevInfo di;
        if (int rc = device_info(&di)) return rc;
        // batches of small images with a small spread: one CTA per image, nothing but the bytes in and out touches HBM
        if (n <= 0x7fffffff && (n >= 2 || g_tuning.glyph == 2) && glyph_fits(w, h, stride, tab, di))
            return launch_glyph(d_imgs, n, w, h, test_above, tab, d_out, st);
        const size_t inter_bytes = align_up(n * h * banded_rs(w, spread), 256);
        if (workspace_bytes < inter_bytes + hull_ws_bytes(n, w, h)) return fail(CHAQSDF_ERR_WORKSPACE, "workspace too small");
        if (int rc = launch_pass1_banded(d_imgs, w, n * h, stride, offset, test_above, spread, (uint8_t*)d_workspace, st)) return rc;
        return launch_hull(spread, (const uint8_t*)d_workspace, w, h, 0, h, n, tab, d_out, (char*)d_workspace + inter_bytes,
                           workspace_bytes - inter_bytes, st);
    }
    const int cap = cap_for(spread);
    EnvPlan pl;
    if (int rc = plan_envelope(w, h, h, n, cap, &pl)) return rc;
    const size_t inter_bytes = align_up(n * w * h * sizeof(uint16_t), 256);
    if (workspace_bytes < inter_bytes + pl.ring_bytes) return fail(CHAQSDF_ERR_WORKSPACE, "workspace too small");
    uint16_t* inter = (uint16_t*)d_workspace;
    if (int rc = launch_rowscan(d_imgs, w, n * h, stride, offset, test_above, cap, inter, st)) return rc;
    EnvParams p{};
    p.inter = inter; p.w = (int)w; p.buf_rows = (int)h; p.out_begin = 0; p.out_end = (int)h;
    p.band_rows = pl.band_rows; p.cap = cap; p.nw = pl.nw; p.table = tab->dq; p.out = d_out;
    p.gring = (uint32_t*)((char*)d_workspace + inter_bytes);
    return launch_env(pl, p, false, st);
}

int chaqsdf_u8_device(const uint8_t* d_img, size_t w, size_t h, size_t stride, size_t offset, int test_above,
                      size_t spread, int asymmetric, uint8_t* d_out, void* d_workspace, size_t workspace_bytes,
                      void* stream) {
    return chaqsdf_u8_batched_device(d_img, 1, w, h, stride, offset, test_above, spread, asymmetric, d_out,
                                     d_workspace, workspace_bytes, stream);
}

// ---- multi-GPU exact path (variant E1: all-to-all between the passes) ---------------------------------------------
size_t chaqsdf_col_part(size_t w, size_t nparts, size_t q, size_t* c0, size_t* c1) {
    if (nparts == 0 || w == 0) return 0;
    const size_t pw = align_up((w + nparts - 1) / nparts, 32);
    const size_t a = q * pw < w ? q * pw : w, b = (q + 1) * pw < w ? (q + 1) * pw : w;
    if (c0) *c0 = a;
    if (c1) *c1 = b;
    return pw;
}

int chaqsdf_rowscan_scatter_device(const uint8_t* d_img, size_t w, size_t rows, size_t stride, size_t offset, int test_above,
                                   size_t row0, size_t nparts, void* const* d_parts, void* stream) {
    if (!d_img || !d_parts) return fail(CHAQSDF_ERR_ARG, "null pointer");
    if (w == 0 || rows == 0) return fail(CHAQSDF_ERR_ARG, "empty image");
    if (w > 32767) return fail(CHAQSDF_ERR_ARG, "exact fields need sides <= 32767");
    if (stride == 0 || offset >= stride) return fail(CHAQSDF_ERR_ARG, "bad stride/offset");
    if (nparts == 0 || nparts > (size_t)kRowMaxParts) return fail(CHAQSDF_ERR_ARG, "1..16 column parts");
    RowScanParams rp{};
    rp.img = d_img; rp.inter = nullptr; rp.rows = (long long)rows; rp.w = (int)w;
    rp.stride = (int)stride; rp.offset = (int)offset; rp.above = test_above ? 1 : 0; rp.cap = 0x7fff;
    rp.nwords = (int)((w + 31) / 32);
    rp.nparts = (int)nparts; rp.part_w = (int)chaqsdf_col_part(w, nparts, 0, nullptr, nullptr); rp.row0 = (long long)row0;
    for (size_t q = 0; q < nparts; ++q) {
        size_t c0 = 0, c1 = 0;
        chaqsdf_col_part(w, nparts, q, &c0, &c1);
        if (c1 > c0 && !d_parts[q]) return fail(CHAQSDF_ERR_ARG, "null part pointer");
        rp.part[q] = (uint16_t*)d_parts[q];
    }
    const size_t smem = (size_t)kRowWarps * rp.nwords * sizeof(uint32_t);
    const unsigned blocks = (unsigned)((rows + kRowWarps - 1) / kRowWarps);
    rowscan_kernel<<<blocks, kRowWarps * 32, smem, (cudaStream_t)stream>>>(rp);
    CQ_LAUNCHED("rowscan_kernel<scatter>", blocks, smem);
    return CHAQSDF_OK;
}

int chaqsdf_fields_from_inter_device(const void* d_inter, size_t w, size_t h, size_t spread, int asymmetric,
                                     int32_t* d_d2_in, int32_t* d_d2_out, float* d_f_in, float* d_f_out, float* d_sdf,
                                     uint8_t* d_out, void* d_workspace, size_t workspace_bytes, void* stream) {
    if (!d_inter) return fail(CHAQSDF_ERR_ARG, "null pointer");
    if (int rc = check_dims(w, h, spread)) return rc;
    if (w > 32767 || h > 32767) return fail(CHAQSDF_ERR_ARG, "exact fields need sides <= 32767");
    const int cap = 0x7fff;
    EnvPlan pl;
    if (int rc = plan_envelope(w, h, h, 1, cap, &pl)) return rc;
    if (pl.ring_bytes > workspace_bytes || (pl.ring_bytes && !d_workspace)) return fail(CHAQSDF_ERR_WORKSPACE, "workspace too small");
    const Tables* tab;
    if (int rc = get_table(spread, asymmetric, &tab)) return rc;
    EnvParams p{};
    p.inter = (const uint16_t*)d_inter; p.w = (int)w; p.buf_rows = (int)h; p.out_begin = 0; p.out_end = (int)h;
    p.band_rows = pl.band_rows; p.cap = cap; p.nw = pl.nw; p.table = tab->dq; p.out = d_out;
    p.d2_in = d_d2_in; p.d2_out = d_d2_out; p.f_in = d_f_in; p.f_out = d_f_out; p.sdf = d_sdf;
    p.gring = (uint32_t*)d_workspace;
    return launch_env(pl, p, true, (cudaStream_t)stream);
}
size_t chaqsdf_fields_from_inter_workspace_bytes(size_t w, size_t h) {
    if (check_dims(w, h, 1)) return 0;
    EnvPlan pl;
    if (plan_envelope(w, h, h, 1, 0x7fff, &pl)) return 0;
    return pl.ring_bytes;
}

int chaqsdf_fields_device(const uint8_t* d_img, size_t w, size_t h, size_t stride, size_t offset, int test_above,
                          size_t spread, int asymmetric, int32_t* d_d2_in, int32_t* d_d2_out, float* d_f_in,
                          float* d_f_out, float* d_sdf, uint8_t* d_out, void* d_workspace, size_t workspace_bytes,
                          void* stream) {
    if (!d_img || !d_workspace) return fail(CHAQSDF_ERR_ARG, "null pointer");
    if (int rc = check_dims(w, h, spread)) return rc;
    if (stride == 0 || offset >= stride) return fail(CHAQSDF_ERR_ARG, "bad stride/offset");
    if (w > 32767 || h > 32767) return fail(CHAQSDF_ERR_ARG, "exact fields need sides <= 32767");
    const int cap = 0x7fff;  // nothing saturates: every line distance is < 32767
    EnvPlan pl;
    if (int rc = plan_envelope(w, h, h, 1, cap, &pl)) return rc;
    const size_t inter_bytes = align_up(w * h * sizeof(uint16_t), 256);
    if (workspace_bytes < inter_bytes + pl.ring_bytes) return fail(CHAQSDF_ERR_WORKSPACE, "workspace too small");
    const Tables* tab;
    if (int rc = get_table(spread, asymmetric, &tab)) return rc;
    uint16_t* inter = (uint16_t*)d_workspace;
    cudaStream_t st = (cudaStream_t)stream;
    if (int rc = launch_rowscan(d_img, w, h, stride, offset, test_above, cap, inter, st)) return rc;
    EnvParams p{};
    p.inter = inter; p.w = (int)w; p.buf_rows = (int)h; p.out_begin = 0; p.out_end = (int)h;
    p.band_rows = pl.band_rows; p.cap = cap; p.nw = pl.nw; p.table = tab->dq; p.out = d_out;
    p.d2_in = d_d2_in; p.d2_out = d_d2_out; p.f_in = d_f_in; p.f_out = d_f_out; p.sdf = d_sdf;
    p.gring = (uint32_t*)((char*)d_workspace + inter_bytes);
    return launch_env(pl, p, true, st);
}

}  // extern "C"

// ---- no-allocation variant (no counterpart in the reference) --------
// chaqsdf_u8_device with its workspace AND its quantisation tables in a file-scope device array, for callers that
// must not allocate (CUDA graph capture, serving loops). Every GPU that loads the library holds the array until exit.
namespace {

constexpr size_t kStaticTabBytes = 180224;  // LevelTable @0 (4 KB), bucketed table @4096 (36 KB), byte table @40960
static_assert(CHAQSDF_STATIC_SCRATCH_BYTES > kStaticTabBytes && kStaticTabBytes % 256 == 0, "static scratch layout");
static __device__ __align__(256) unsigned char g_static_scratch[CHAQSDF_STATIC_SCRATCH_BYTES];

struct StaticDev {
    uint8_t* base = nullptr;      // device address of g_static_scratch on this device (cudaGetSymbolAddress)
    bool have = false;            // tables of (spread, asym) were enqueued into the array by the last call
    size_t spread = 0;
    int asym = 0;
    Tables tab;                   // pointers into the array
    std::string lutb, lutk;       // host copies (kept alive: they are the source of asynchronous copies)
};
std::mutex g_static_mu;
std::map<int, StaticDev> g_static;

// the tables of get_table() for the banded path of spreads <= 256, built on the host, nothing uploaded
void static_build_tables(size_t spread, int asymmetric, StaticDev* d) {
    QuantTable h;
    build_quant_table(spread, asymmetric, &h);
    Tables& t = d->tab;
    build_level_table(h, &t.hl);
    t.dq = nullptr; t.dlut = nullptr;
    t.lutb_stride = (int)align_up((size_t)t.hl.c2 + 1, 16);
    d->lutb.assign((size_t)2 * t.lutb_stride, '\0');
    for (int cls = 0; cls < 2; ++cls) {
        int m = 0;
        for (int d2 = 0; d2 <= t.hl.c2; ++d2) {
            while (m < 255 && t.hl.thr[cls][m + 1] <= d2) ++m;
            d->lutb[(size_t)cls * t.lutb_stride + d2] = (char)(uint8_t)(cls ? m : 255 - m);
        }
    }
    const int nbk = (t.hl.c2 >> 5) + 1;
    const int kstride = (int)align_up((size_t)nbk * 8, 16);
    d->lutk.assign((size_t)2 * kstride, '\0');
    bool ok = true;
    for (int cls = 0; cls < 2 && ok; ++cls) {
        const uint8_t* bytes = (const uint8_t*)d->lutb.data() + (size_t)cls * t.lutb_stride;
        uint32_t* ent = (uint32_t*)(&d->lutk[(size_t)cls * kstride]);
        for (int i = 0; i < nbk; ++i) {
            uint32_t m = 0;
            for (int j = 1; j < 32 && i * 32 + j <= t.hl.c2; ++j) {
                const int dd = (int)bytes[i * 32 + j] - (int)bytes[i * 32 + j - 1];
                if (dd == (cls ? 1 : -1)) m |= 1u << j;
                else if (dd != 0) ok = false;
            }
            ent[2 * i] = m;
            ent[2 * i + 1] = bytes[i * 32];
        }
    }
    t.lutk_stride = ok ? kstride : 0; t.lutk_nbk = ok ? nbk : 0;
    t.dl = (LevelTable*)d->base;
    t.dlutk = ok ? d->base + 4096 : nullptr;
    t.dlutb = d->base + 40960;
}

}  // namespace

extern "C" int chaqsdf_u8_device_static(const uint8_t* d_img, size_t w, size_t h, size_t stride, size_t offset,
                                        int test_above, size_t spread, int asymmetric, uint8_t* d_out, void* stream) {
    if (!d_img || !d_out) return fail(CHAQSDF_ERR_ARG, "null pointer");
    if (int rc = check_dims(w, h, spread)) return rc;
    if (stride == 0 || offset >= stride) return fail(CHAQSDF_ERR_ARG, "bad stride/offset");
    if (spread > 256) return fail(CHAQSDF_ERR_ARG, "chaqsdf_u8_device_static: spread > 256 is not supported by the static-scratch variant");
    const size_t inter_bytes = align_up(h * banded_rs(w, spread), 256);
    const size_t need = inter_bytes + hull_ws_bytes(1, w, h);
    if (need > CHAQSDF_STATIC_SCRATCH_BYTES - kStaticTabBytes)
        return fail(CHAQSDF_ERR_WORKSPACE, "chaqsdf_u8_device_static: image needs " + std::to_string(need) + " bytes of scratch, the static array holds " +
                                               std::to_string(CHAQSDF_STATIC_SCRATCH_BYTES - kStaticTabBytes) + " (use chaqsdf_u8_device with a caller workspace)");
    int dev = 0;
    CQ_CUDA(cudaGetDevice(&dev));
    cudaStream_t st = (cudaStream_t)stream;
    std::lock_guard<std::mutex> lk(g_static_mu);
    StaticDev& d = g_static[dev];
    if (!d.base) {  // once per device, with that device current: the array's name is not a device address in host code
        void* ptr = nullptr;
        CQ_CUDA(cudaGetSymbolAddress(&ptr, g_static_scratch));
        if (!ptr) return fail(CHAQSDF_ERR_CUDA, "cudaGetSymbolAddress returned a null address");
        d.base = (uint8_t*)ptr;
    }
    if (!d.have || d.spread != spread || d.asym != (asymmetric ? 1 : 0)) {
        d.have = false;
        static_build_tables(spread, asymmetric, &d);
        if (d.tab.hl.cap_eff > 256) return fail(CHAQSDF_ERR_ARG, "internal: banded path beyond its reach");
        CQ_CUDA(cudaMemcpyAsync(d.tab.dl, &d.tab.hl, sizeof(LevelTable), cudaMemcpyHostToDevice, st));
        CQ_CUDA(cudaMemcpyAsync(d.tab.dlutb, d.lutb.data(), d.lutb.size(), cudaMemcpyHostToDevice, st));
        if (d.tab.dlutk) CQ_CUDA(cudaMemcpyAsync(d.tab.dlutk, d.lutk.data(), d.lutk.size(), cudaMemcpyHostToDevice, st));
        d.have = true; d.spread = spread; d.asym = asymmetric ? 1 : 0;
    }
    uint8_t* ws = d.base + kStaticTabBytes;
    // nothing in the workspace has to start at zero: pass 1 writes every record, the build kernel writes every hull /
    // class word and flag and clears the paint kernel's work counter before the paint kernel of the same stream reads it
    if (int rc = launch_pass1_banded(d_img, w, h, stride, offset, test_above, spread, ws, st)) return rc;
    return launch_hull(spread, ws, w, h, 0, h, 1, &d.tab, d_out, ws + inter_bytes, CHAQSDF_STATIC_SCRATCH_BYTES - kStaticTabBytes - inter_bytes, st,
                       /*one_stream=*/true);
}

#ifdef CQ_BUILD_TRACE
extern "C" void chaqsdf_debug_set_build_order(const uint32_t* d_order, unsigned n) { g_dbg_order = d_order; g_dbg_order_n = n; }
extern "C" int chaqsdf_debug_build_trace(unsigned long long* out, int n) {
    return cudaMemcpyFromSymbol(out, g_build_trace, (size_t)n * 3 * sizeof(unsigned long long)) == cudaSuccess ? 0 : 2;
}
#endif
