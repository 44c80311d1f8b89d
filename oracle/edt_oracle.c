/*
 * oracle/edt_oracle.c -- TEST INFRASTRUCTURE ONLY. Never linked into, imported by, or called from the
 * product path (chaq-sdfgen_b200/). Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * --impl reference legs may load this.
 *
 * A CPU restatement, in exact integer arithmetic, of the hot path of chaquator/chaq-sdfgen's OpenMP
 * program (citations are relative to /root/reference/):
 *
 *   a1  transform_img_to_bool      openmp/sdfgen.c:52-62    -> oracle_img_to_mask
 *   a2  transform_bool_to_float    openmp/sdfgen.c:65-72    -> (implicit: seeds of the EDT)
 *   a3-a6 dist_transform_2d        openmp/df.c:9-136        -> oracle_edt_d2 (exact squared EDT)
 *   a8  transform_float_sub        openmp/sdfgen.c:98-106   -> oracle_signed_from_d2
 *   a9  transform_float_to_byte    openmp/sdfgen.c:75-96    -> oracle_byte_from_signed
 *   a7  the two-field orchestration openmp/sdfgen.c:277-300 -> oracle_sdf_u8 / oracle_sdf_fields
 *
 * The reference computes the EDT with the Felzenszwalb-Huttenlocher lower envelope in float32; that is
 * exact only for sides <= 4098. This oracle computes the same mathematical
 * object -- D2(p) = min over seeds q of |p-q|^2 -- in integers (row scan + Meijster-style column envelope
 * with 64-bit separators), so it is the oracle of record at every size and for non-square images, and is
 * pinned against the stock build (oracle/_ref, compiled from the reference tree) where the stock build is
 * itself exact, against the reference's single golden pair, and against a brute-force search on small
 * images (tests/test_oracle.py).
 *
 * Float semantics of a8/a9 follow the reference build flags (openmp/CMakeLists.txt:7-12: C11, no
 * extensions => no FMA contraction, one IEEE rounding per operation). Compile with -ffp-contract=off.
 */
#include <math.h>
#include <stdbool.h>
#include <stddef.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define ORACLE_INF_D2 INT32_MAX /* "no seed anywhere" (reference: +INFINITY, sdfgen.c:70) */

/* a1: sdfgen.c:52-62. value 127 is "outside" for both directions of the test. */
void oracle_img_to_mask(const uint8_t* img, uint8_t* mask, size_t w, size_t h, size_t stride, size_t offset,
                        int test_above) {
    size_t n = w * h;
#pragma omp parallel for schedule(static)
    for (ptrdiff_t i = 0; i < (ptrdiff_t)n; ++i) {
        uint8_t px = img[(size_t)i * stride + offset];
        mask[i] = test_above ? (px > 127) : (px < 127);
    }
}

/* 1-D distance (not squared) along each row to the nearest pixel whose mask value == seed_value.
 * INT32_MAX/2 if the row has none. This is what the first F-H pass (df.c:130) yields for {0,INF} input:
 * the lower envelope of unit parabolas rooted at the seeds is (x - nearest seed)^2. */
static void row_scan(const uint8_t* mask, int32_t* g, size_t w, size_t h, uint8_t seed_value) {
    const int32_t none = INT32_MAX / 2;
#pragma omp parallel for schedule(static)
    for (ptrdiff_t y = 0; y < (ptrdiff_t)h; ++y) {
        const uint8_t* m = mask + (size_t)y * w;
        int32_t* o = g + (size_t)y * w;
        int32_t d = none;
        for (size_t x = 0; x < w; ++x) {
            d = (m[x] == seed_value) ? 0 : (d >= none ? none : d + 1);
            o[x] = d;
        }
        d = none;
        for (size_t x = w; x-- > 0;) {
            d = (m[x] == seed_value) ? 0 : (d >= none ? none : d + 1);
            if (d < o[x]) o[x] = d;
        }
    }
}

/* Column pass: D2(x,y) = min_{y'} (y-y')^2 + g(x,y')^2, by the integer lower envelope
 * (Meijster, Roerdink, Hesselink 2000 -- the integer sibling of df.c:57-96). Output int32 squared
 * distance, ORACLE_INF_D2 when the image has no seed at all. */
static void column_envelope(const int32_t* g, int32_t* d2, size_t w, size_t h) {
    const int32_t none = INT32_MAX / 2;
#pragma omp parallel
    {
        int64_t* s = (int64_t*)malloc(sizeof(int64_t) * h); /* site rows on the envelope   */
        int64_t* t = (int64_t*)malloc(sizeof(int64_t) * h); /* first row each site owns    */
        int64_t* f = (int64_t*)malloc(sizeof(int64_t) * h); /* g^2 of the column, or -1    */
#pragma omp for schedule(static)
        for (ptrdiff_t x = 0; x < (ptrdiff_t)w; ++x) {
            ptrdiff_t q = -1;
            for (size_t y = 0; y < h; ++y) {
                int32_t gv = g[y * w + (size_t)x];
                f[y] = gv >= none ? -1 : (int64_t)gv * gv;
            }
            for (size_t u = 0; u < h; ++u) {
                if (f[u] < 0) continue; /* no parabola here (df.c:60) */
                while (q >= 0) {
                    /* value of top site vs new site at the row where the top starts to own */
                    int64_t r = t[q];
                    int64_t a = (r - s[q]) * (r - s[q]) + f[s[q]];
                    int64_t b = (r - (int64_t)u) * (r - (int64_t)u) + f[u];
                    if (a > b) --q; else break;
                }
                if (q < 0) {
                    q = 0; s[0] = (int64_t)u; t[0] = 0;
                } else {
                    /* first integer row where u beats s[q]: floor(sep)+1 */
                    int64_t i = s[q], uu = (int64_t)u;
                    int64_t num = uu * uu - i * i + f[u] - f[i];
                    int64_t den = 2 * (uu - i);
                    int64_t sep = num >= 0 ? num / den : -((-num + den - 1) / den);
                    int64_t wrow = sep + 1;
                    if (wrow < (int64_t)h) { ++q; s[q] = uu; t[q] = wrow < 0 ? 0 : wrow; }
                }
            }
            if (q < 0) {
                for (size_t y = 0; y < h; ++y) d2[y * w + (size_t)x] = ORACLE_INF_D2;
                continue;
            }
            for (size_t y = h; y-- > 0;) {
                int64_t dy = (int64_t)y - s[q];
                d2[y * w + (size_t)x] = (int32_t)(dy * dy + f[s[q]]);
                if ((int64_t)y == t[q] && q > 0) --q;
            }
        }
        free(s); free(t); free(f);
    }
}

/* Exact squared EDT to the nearest pixel with mask == seed_value. d2: int32[w*h]. */
void oracle_edt_d2(const uint8_t* mask, size_t w, size_t h, int seed_value, int32_t* d2) {
    int32_t* g = (int32_t*)malloc(sizeof(int32_t) * w * h);
    row_scan(mask, g, w, h, (uint8_t)(seed_value != 0));
    column_envelope(g, d2, w, h);
    free(g);
}

/* O(n^2) per pixel brute force, for pinning the envelope code on small images. */
void oracle_edt_d2_bruteforce(const uint8_t* mask, size_t w, size_t h, int seed_value, int32_t* d2) {
    uint8_t sv = (uint8_t)(seed_value != 0);
#pragma omp parallel for schedule(static)
    for (ptrdiff_t p = 0; p < (ptrdiff_t)(w * h); ++p) {
        int64_t px = p % (ptrdiff_t)w, py = p / (ptrdiff_t)w, best = ORACLE_INF_D2;
        for (size_t qy = 0; qy < h; ++qy)
            for (size_t qx = 0; qx < w; ++qx)
                if (mask[qy * w + qx] == sv) {
                    int64_t dx = px - (int64_t)qx, dy = py - (int64_t)qy, d = dx * dx + dy * dy;
                    if (d < best) best = d;
                }
        d2[p] = (int32_t)best;
    }
}

/* The value the reference's second pass stores: sqrtf(displacement^2 + h) (df.c:89-95). For D2 < 2^24 the
 * float sum is exact, so this is sqrtf((float)D2) bit for bit. */
static inline float dist_from_d2(int32_t d2) { return d2 == ORACLE_INF_D2 ? INFINITY : sqrtf((float)d2); }

/* a8: sdfgen.c:98-106. dst = outside field (distance to nearest outside px), by = inside field. */
static inline float signed_from_fields(float d_out, float d_in) {
    float bias = -1.f;
    float val = d_in > 0.f ? d_in + bias : d_in;
    return d_out - val;
}

/* a9: sdfgen.c:75-96, operation for operation. */
static inline uint8_t byte_from_signed(float v, size_t spread, int asymmetric) {
    float s_min = asymmetric ? 0 : -(float)spread;
    float s_max = (float)spread;
    float d_min = 0.f;
    float d_max = 255.f;
    float sn = s_max - s_min;
    float nd = d_max - d_min;
    v = v > s_max ? s_max : v;
    v = v < s_min ? s_min : v;
    float remap = (((v - s_min) * nd) / sn) + d_min;
    return (uint8_t)remap;
}

float oracle_signed_from_d2(int32_t d2_out, int32_t d2_in) {
    return signed_from_fields(dist_from_d2(d2_out), dist_from_d2(d2_in));
}
uint8_t oracle_byte_from_signed(float v, size_t spread, int asymmetric) {
    return byte_from_signed(v, spread, asymmetric);
}

/* Everything between stbi_load and stbi_write (sdfgen.c:261-300). Any of the optional outputs may be NULL.
 *   d2_in  : squared distance to the nearest INSIDE pixel  (the reference's img_float_inside, squared)
 *   d2_out : squared distance to the nearest OUTSIDE pixel (img_float_outside before the subtraction)
 *   sdf    : the float field handed to transform_float_to_byte
 *   out    : the 8-bit bitmap */
int oracle_sdf_fields(const uint8_t* img, size_t w, size_t h, size_t stride, size_t offset, int test_above,
                      size_t spread, int asymmetric, uint8_t* out, float* sdf, int32_t* d2_in, int32_t* d2_out) {
    size_t n = w * h;
    if (n == 0) return 0;
    uint8_t* mask = (uint8_t*)malloc(n);
    int32_t* din = d2_in ? d2_in : (int32_t*)malloc(sizeof(int32_t) * n);
    int32_t* dout = d2_out ? d2_out : (int32_t*)malloc(sizeof(int32_t) * n);
    if (!mask || !din || !dout) return -1;
    oracle_img_to_mask(img, mask, w, h, stride, offset, test_above);
    oracle_edt_d2(mask, w, h, 1, din);  /* inside field: seeds are the inside pixels (sdfgen.c:281) */
    oracle_edt_d2(mask, w, h, 0, dout); /* outside field (sdfgen.c:286) */
#pragma omp parallel for schedule(static)
    for (ptrdiff_t i = 0; i < (ptrdiff_t)n; ++i) {
        float v = signed_from_fields(dist_from_d2(dout[i]), dist_from_d2(din[i]));
        if (sdf) sdf[i] = v;
        if (out) out[i] = byte_from_signed(v, spread, asymmetric);
    }
    free(mask);
    if (!d2_in) free(din);
    if (!d2_out) free(dout);
    return 0;
}

int oracle_sdf_u8(const uint8_t* img, size_t w, size_t h, size_t stride, size_t offset, int test_above,
                  size_t spread, int asymmetric, uint8_t* out) {
    return oracle_sdf_fields(img, w, h, stride, offset, test_above, spread, asymmetric, out, NULL, NULL, NULL);
}

/* df.h:6 semantics for the only input domain the CLI produces ({0,+INF}): in place, sqrt of the exact EDT.
 * Returns -1 if the buffer holds anything else (general heights are not restated here). */
int oracle_dist_transform_2d_binary(float* img, size_t w, size_t h) {
    size_t n = w * h;
    uint8_t* mask = (uint8_t*)malloc(n);
    int32_t* d2 = (int32_t*)malloc(sizeof(int32_t) * n);
    for (size_t i = 0; i < n; ++i) {
        if (img[i] == 0.f) mask[i] = 1;
        else if (isinf(img[i]) && img[i] > 0) mask[i] = 0;
        else { free(mask); free(d2); return -1; }
    }
    oracle_edt_d2(mask, w, h, 1, d2);
    for (size_t i = 0; i < n; ++i) img[i] = dist_from_d2(d2[i]);
    free(mask); free(d2);
    return 0;
}

/* ---- synthetic inputs of SURVEY.md section 8(d) (xorshift64, seed 88172645463325252) ---- */
static inline uint64_t xs64(uint64_t* s) { *s ^= *s << 13; *s ^= *s >> 7; *s ^= *s << 17; return *s; }

void oracle_syn_discs(uint8_t* img, size_t n_w, size_t n_h, uint64_t seed, int n_discs) {
    uint64_t s = seed ? seed : 88172645463325252ULL;
    size_t n = n_w < n_h ? n_w : n_h;
    memset(img, 0, n_w * n_h);
    for (int k = 0; k < n_discs; ++k) {
        int64_t cx = (int64_t)(xs64(&s) % n_w), cy = (int64_t)(xs64(&s) % n_h);
        int64_t r = 8 + (int64_t)(xs64(&s) % (n / 16 ? n / 16 : 1));
        int64_t y0 = cy - r < 0 ? 0 : cy - r, y1 = cy + r >= (int64_t)n_h ? (int64_t)n_h - 1 : cy + r;
#pragma omp parallel for schedule(static)
        for (int64_t y = y0; y <= y1; ++y) {
            int64_t dy = y - cy, rem = r * r - dy * dy;
            int64_t hw = (int64_t)floor(sqrt((double)rem));
            while ((hw + 1) * (hw + 1) <= rem) ++hw;
            while (hw * hw > rem) --hw;
            int64_t x0 = cx - hw < 0 ? 0 : cx - hw, x1 = cx + hw >= (int64_t)n_w ? (int64_t)n_w - 1 : cx + hw;
            if (x1 >= x0) memset(img + (size_t)y * n_w + (size_t)x0, 255, (size_t)(x1 - x0 + 1));
        }
    }
}

void oracle_syn_noise(uint8_t* img, size_t w, size_t h, uint64_t seed) {
    uint64_t s = seed ? seed : 88172645463325252ULL;
    for (size_t i = 0; i < w * h; ++i) img[i] = ((xs64(&s) & 63) == 0) ? 255 : 0;
}

/* SYN-GLYPH(i) of SURVEY.md section 8(d): an n x n canvas (n >= 96), xorshift64 seeded with base + i, union of 2-5
 * strokes -- annuli and thick segments, 6-20 px wide -- kept >= 16 px from the border. Integer arithmetic only, so
 * bench.py's numpy restatement reproduces it bit for bit. */
void oracle_syn_glyph(uint8_t* img, size_t n, uint64_t index) {
    uint64_t s = 88172645463325252ULL + index;
    const int64_t N = (int64_t)n;
    memset(img, 0, n * n);
    int k = 2 + (int)(xs64(&s) % 4);
    for (int st = 0; st < k; ++st) {
        int type = (int)(xs64(&s) % 2);
        int64_t wd = 6 + (int64_t)(xs64(&s) % 15);
        if (type == 0) {
            int64_t R = 24 + (int64_t)(xs64(&s) % (uint64_t)(N / 2 - 40));
            int64_t span = N - 32 - 2 * R;
            int64_t cx = 16 + R + (int64_t)(xs64(&s) % (uint64_t)span);
            int64_t cy = 16 + R + (int64_t)(xs64(&s) % (uint64_t)span);
            int64_t inner = R - wd > 0 ? R - wd : 0;
            for (int64_t y = cy - R; y <= cy + R; ++y)
                for (int64_t x = cx - R; x <= cx + R; ++x) {
                    int64_t d2 = (x - cx) * (x - cx) + (y - cy) * (y - cy);
                    if (d2 <= R * R && d2 >= inner * inner) img[(size_t)y * n + (size_t)x] = 255;
                }
        } else {
            int64_t hw = wd / 2;
            int64_t span = N - 32 - 2 * hw;
            int64_t x0 = 16 + hw + (int64_t)(xs64(&s) % (uint64_t)span), y0 = 16 + hw + (int64_t)(xs64(&s) % (uint64_t)span);
            int64_t x1 = 16 + hw + (int64_t)(xs64(&s) % (uint64_t)span), y1 = 16 + hw + (int64_t)(xs64(&s) % (uint64_t)span);
            int64_t dx = x1 - x0, dy = y1 - y0, L2 = dx * dx + dy * dy;
            int64_t xa = (x0 < x1 ? x0 : x1) - hw, xb = (x0 > x1 ? x0 : x1) + hw;
            int64_t ya = (y0 < y1 ? y0 : y1) - hw, yb = (y0 > y1 ? y0 : y1) + hw;
            for (int64_t y = ya; y <= yb; ++y)
                for (int64_t x = xa; x <= xb; ++x) {
                    int64_t px = x - x0, py = y - y0, t = px * dx + py * dy;
                    int hit;
                    if (t <= 0) hit = px * px + py * py <= hw * hw;
                    else if (t >= L2) hit = (x - x1) * (x - x1) + (y - y1) * (y - y1) <= hw * hw;
                    else { int64_t c = px * dy - py * dx; hit = c * c <= hw * hw * L2; }
                    if (hit) img[(size_t)y * n + (size_t)x] = 255;
                }
        }
    }
}
