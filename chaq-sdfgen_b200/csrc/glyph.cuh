// glyph.cuh -- both passes fused for batches of small images (glyph atlases, BASELINE config 3): one CTA per image,
// the image's class bits and row distances never leave shared memory, so HBM sees 1 byte in and 1 byte out per
// pixel. Replaces, per image, everything between stbi_load and stbi_write of reference openmp/sdfgen.c:261-300
// (transform_img_to_bool :52-62, both dist_transform_2d calls :277-289 -> df.c:125-136, transform_float_sub
// :98-106, transform_float_to_byte :75-96).
//
// Used when the spread is small (C = cap_eff <= 33, i.e. spread <= 32): every result is clamped at C^2, so only rows
// and columns within L = C-1 <= 32 pixels matter and the column pass can simply take the minimum over that window --
// no hull, no data-dependent control flow:
//   phase 1 (warp per row)   threshold (px > 127 / px < 127) -> class words; per pixel the distance along the row
//                            to the nearest class change from two funnel shifts + count-leading/trailing-zeros on
//                            the 96-bit window around it; stored as a byte min(G, C) in shared memory.
//   phase 2 (thread = column, 32 rows at a time)  D2(y) = min_r (y-r)^2 + G(r)^2 over |y-r| <= LMAX (the lower
//                            envelope of df.c:57-96 by exhaustion), two output rows per DPX instruction
//                            (VIADDMNMX.U16x2: min(a + b, c) on 16-bit halves; all values < 2^15), then the
//                            in-column distance to the nearest row of the other class from the transposed class
//                            words, min, and the byte from a look-up table indexed by the clamped D2 (built on
//                            the host from tables.h: exact by construction).
// One envelope serves both fields (hull_core.h has the argument). Images are independent.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "common.cuh"
#include "rowscan.cuh"  // class_nibble

namespace chaq {

constexpr int kGlyphThreads = 512;
constexpr int kGlyphMaxC = 33;                      // C <= 33: a 32-bit window on each side sees every distance < C
constexpr int kGlyphLutStride = kGlyphMaxC * kGlyphMaxC + 7;  // bytes per class in the D2 -> byte table (1096)

struct GlyphParams {
    const uint8_t* img;   // [n][h][w] bytes (stride 1)
    uint8_t* out;         // [n][h][w]
    const uint8_t* lut;   // [2][kGlyphLutStride]: byte for (class, clamped D2)
    int w, h, n;
    int wq;               // w / 32
    int hb;               // ceil(h / 32)
    int C, C2;            // cap_eff and its square
    int above;
};

// shared memory: g plane [hb*32 + 2*LMAX][w] | transposed class words [hb + 2][w] | class words, row-major [h][wq] | lut
__host__ __device__ inline size_t glyph_smem(int w, int h, int lmax) {
    const int hb = (h + 31) / 32;
    return (size_t)(hb * 32 + 2 * lmax) * w + (size_t)(hb + 2) * w * 4 + (size_t)h * (w / 32) * 4 +
           (size_t)2 * kGlyphLutStride + 8;
}

template <int LMAX>
__global__ void __launch_bounds__(kGlyphThreads, 2) glyph_kernel(GlyphParams p) {
    extern __shared__ __align__(16) uint8_t s_g[];
    const int w = p.w, h = p.h, wq = p.wq, hb = p.hb;
    uint8_t* s_gp = s_g;
    uint32_t* s_clsT = reinterpret_cast<uint32_t*>(s_g + (size_t)(hb * 32 + 2 * LMAX) * w);
    uint32_t* s_cls = s_clsT + (size_t)(hb + 2) * w;
    uint8_t* s_lut = reinterpret_cast<uint8_t*>(s_cls + (size_t)h * wq);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int C = p.C;
    const uint8_t* img = p.img + (size_t)blockIdx.x * w * h;
    uint8_t* out = p.out + (size_t)blockIdx.x * w * h;

    for (int i = tid; i < 2 * kGlyphLutStride; i += kGlyphThreads) s_lut[i] = __ldg(p.lut + i);
    // rows of the g plane outside the image hold "nothing within reach"
    {
        const uint32_t fill = (uint32_t)C * 0x01010101u;
        uint32_t* top = reinterpret_cast<uint32_t*>(s_gp);
        for (int i = tid; i < LMAX * w / 4; i += kGlyphThreads) top[i] = fill;
        uint32_t* bot = reinterpret_cast<uint32_t*>(s_gp + (size_t)(LMAX + h) * w);
        const int nbot = (hb * 32 - h + LMAX) * w / 4;
        for (int i = tid; i < nbot; i += kGlyphThreads) bot[i] = fill;
    }

    // ---- phase 1: class words and row distances, one warp per row, 8 pixels per lane ----
    const bool above = p.above != 0;
    const int nl = w >> 3;  // lanes that hold pixels
    for (int y = warp; y < h; y += kGlyphThreads / 32) {
        uint32_t m8 = 0u;
        if (lane < nl) {
            const uint2 q = __ldg(reinterpret_cast<const uint2*>(img + (size_t)y * w) + lane);
            m8 = class_nibble(q.x, above) | (class_nibble(q.y, above) << 4);
        }
        uint32_t word = m8 << (8 * (lane & 3));
        word |= __shfl_xor_sync(0xffffffffu, word, 1);
        word |= __shfl_xor_sync(0xffffffffu, word, 2);  // lanes 4k..4k+3: class word k of the row
        const int k = lane >> 2;
        if ((lane & 3) == 0 && k < wq) s_cls[y * wq + k] = word;
        // neighbours; beyond the image the edge pixel's class goes on for ever (no class change out there)
        uint32_t wprev = __shfl_up_sync(0xffffffffu, word, 4);
        uint32_t wnext = __shfl_down_sync(0xffffffffu, word, 4);
        if (k == 0) wprev = (word & 1u) ? 0xFFFFFFFFu : 0u;
        if (k == wq - 1) wnext = (word >> 31) ? 0xFFFFFFFFu : 0u;
        if (lane < nl) {
            uint32_t packed[2];
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                const int b = 8 * (lane & 3) + i;
                const uint32_t flip = ((word >> b) & 1u) ? 0xFFFFFFFFu : 0u;
                const uint32_t R = __funnelshift_rc(word, wnext, b + 1) ^ flip;    // bit 0 = pixel x+1
                const uint32_t Lw = __funnelshift_lc(wprev, word, 32 - b) ^ flip; // bit 31 = pixel x-1
                // 33 (> any reach) when the window holds no change: clz(0) = 32
                const int dr = __clz((int)__brev(R)) + 1;
                const int dl = __clz((int)Lw) + 1;
                const int g = min(min(dl, dr), C);
                if (i == 0 || i == 4) packed[i >> 2] = 0u;
                packed[i >> 2] |= (uint32_t)g << (8 * (i & 3));
            }
            *reinterpret_cast<uint2*>(s_gp + (size_t)(LMAX + y) * w + lane * 8) = make_uint2(packed[0], packed[1]);
        }
    }
    __syncthreads();

    // ---- phase 1b: class words per column (bit j of word m+1 = row 32*m + j), one padding word above and below ----
    for (int t = warp; t < hb * wq; t += kGlyphThreads / 32) {
        const int m = t / wq, k = t - m * wq;
        int r = 32 * m + lane;
        r = r < h ? r : h - 1;
        const uint32_t colw = warp_transpose32(s_cls[r * wq + k], lane);
        s_clsT[(size_t)(m + 1) * w + 32 * k + lane] = colw;
        if (m == 0) s_clsT[32 * k + lane] = (colw & 1u) ? 0xFFFFFFFFu : 0u;
        if (m == hb - 1) s_clsT[(size_t)(hb + 1) * w + 32 * k + lane] = (colw >> 31) ? 0xFFFFFFFFu : 0u;
    }
    __syncthreads();

    // ---- phase 2: columns ----
    const int nslices = kGlyphThreads / w;
    const int x = tid % w, slice = tid / w;
    if (slice >= nslices) return;
    const uint32_t c2 = (uint32_t)p.C2;
    for (int m = slice; m < hb; m += nslices) {
        uint32_t acc[16];
#pragma unroll
        for (int i = 0; i < 16; ++i) acc[i] = c2 | (c2 << 16);
        const uint8_t* gcol = s_gp + (size_t)(32 * m) * w + x;  // plane row (32m + r + LMAX) holds image row 32m + r
#pragma unroll
        for (int r = -LMAX; r < 32 + LMAX; ++r) {
            const uint32_t g = gcol[(size_t)(r + LMAX) * w];
            const uint32_t hp = (g * g) * 0x10001u;  // G^2 in both halves (<= 33^2)
#pragma unroll
            for (int i = 0; i < 16; ++i) {
                const int d0 = 2 * i - r, d1 = 2 * i + 1 - r;
                const bool in0 = d0 >= -LMAX && d0 <= LMAX, in1 = d1 >= -LMAX && d1 <= LMAX;
                if (in0 || in1) {
                    const uint32_t k0 = in0 ? (uint32_t)(d0 * d0) : 0x7000u, k1 = in1 ? (uint32_t)(d1 * d1) : 0x7000u;
                    acc[i] = __viaddmin_u16x2(hp, k0 | (k1 << 16), acc[i]);
                }
            }
        }
        const uint32_t cw = s_clsT[(size_t)(m + 1) * w + x];
        const uint32_t cup = s_clsT[(size_t)m * w + x], cdn = s_clsT[(size_t)(m + 2) * w + x];
        uint8_t* o = out + (size_t)(32 * m) * w + x;
        const int jn = min(32, h - 32 * m);
        // no class change in this column within 32 rows of the block, for every column of the warp: V is out of reach
        const bool vflat = cw == cup && cw == cdn && (cw == 0u || cw == 0xFFFFFFFFu);
        if (__all_sync(0xffffffffu, vflat)) {
            const uint8_t* lut = s_lut + (cw & 1u) * kGlyphLutStride;
            uint32_t any = 0u;
#pragma unroll
            for (int i = 0; i < 16; ++i) any |= acc[i] ^ (c2 | (c2 << 16));
            if (__all_sync(0xffffffffu, any == 0u)) {
                // nothing within reach of any pixel of the 32 x 32 block: one saturated byte per column
                const uint8_t b = lut[c2];
                for (int j = 0; j < jn; ++j, o += w) *o = b;
            } else {
#pragma unroll
                for (int j = 0; j < 32; ++j) {
                    if (j < jn) {
                        *o = lut[(acc[j >> 1] >> (16 * (j & 1))) & 0xFFFFu];
                        o += w;
                    }
                }
            }
        } else {
#pragma unroll
            for (int j = 0; j < 32; ++j) {
                if (j < jn) {
                    const uint32_t cls = (cw >> j) & 1u;
                    const uint32_t flip = 0u - cls;
                    const uint32_t U = __funnelshift_lc(cup, cw, 32 - j) ^ flip;  // bit 31 = row y-1
                    const uint32_t D = __funnelshift_rc(cw, cdn, j + 1) ^ flip;   // bit 0 = row y+1
                    const uint32_t v = (uint32_t)min(__clz((int)U), __clz((int)__brev(D))) + 1u;  // 33 if none in the window
                    const uint32_t d2 = min((acc[j >> 1] >> (16 * (j & 1))) & 0xFFFFu, v * v);  // acc never exceeds C^2
                    *o = (cls ? s_lut + kGlyphLutStride : s_lut)[d2];
                    o += w;
                }
            }
        }
    }
}

}  // namespace chaq
