// tables.h -- exact 8-bit quantisation of integer squared distances.
//
// The reference turns a squared distance into a byte through float32 arithmetic:
//   d = sqrtf(D2)                                  (openmp/df.c:95)
//   inside pixel : sdf = d_out - 0                 (openmp/sdfgen.c:102-104, d_in == 0)
//   outside pixel: sdf = 0 - (d_in + -1.f)         (same lines, d_out == 0)
//   byte = (uchar)(((clamp(sdf) - s_min) * 255.f) / (s_max - s_min) + 0.f)   (openmp/sdfgen.c:81-94)
// Every step is monotone in D2, so for each class the byte is a monotone step function of the integer D2
// with at most 256 steps. The host evaluates the float formula (the very same operation sequence, compiled
// with -ffp-contract=off) to find, by bisection, the D2 at which each step happens; the device then needs
// only integer compares against those 2x257 thresholds, and is bit-exact by construction.
//
//   tin[k]  = smallest D2 with byte_inside(D2)  >= k   (non-decreasing in k, tin[0] = 0,  kBig if never)
//   tout[k] = smallest D2 with byte_outside(D2) <= k   (non-increasing in k,             kBig if never)
// All D2 >= cap2 = (spread+1)^2 (and "no pixel of the other class at all") give the same byte as cap2.
#pragma once
#include <stdint.h>
#include "envelope_core.h"

namespace chaq {

struct QuantTable {
    int32_t tin[257];
    int32_t tout[257];
    int32_t cap2;       // (spread+1)^2 clamped to int32
    float smin, scale;  // for the device-side first guess only: byte ~ (v - smin) * scale
    float smax;
    int32_t pad;
};

// first guess + integer correction. d2 >= 0 (kBig allowed).
CQ_HD int quantise(const int32_t* tin, const int32_t* tout, int cap2, float smin, float smax, float scale, int cls,
                   int d2) {
    if (d2 > cap2) d2 = cap2;
#if defined(__CUDA_ARCH__)
    float d = __fsqrt_rn((float)d2);
#else
    float d = __builtin_sqrtf((float)d2);
#endif
    float v = cls ? d : 1.0f - d;
    v = v > smax ? smax : v;
    v = v < smin ? smin : v;
    int k = (int)((v - smin) * scale);
    k = k < 0 ? 0 : (k > 255 ? 255 : k);
    if (cls) {
        while (k < 255 && d2 >= tin[k + 1]) ++k;
        while (k > 0 && d2 < tin[k]) --k;
    } else {
        while (k > 0 && d2 >= tout[k - 1]) --k;
        while (k < 255 && d2 < tout[k]) ++k;
    }
    return k;
}

// ---- host side: the reference's float formula and the threshold search ----
inline uint8_t ref_byte_from_signed(float v, float s_min, float s_max) {
    float d_min = 0.f, d_max = 255.f;
    float sn = s_max - s_min;
    float nd = d_max - d_min;
    v = v > s_max ? s_max : v;
    v = v < s_min ? s_min : v;
    float remap = (((v - s_min) * nd) / sn) + d_min;
    return (uint8_t)remap;
}
inline uint8_t ref_byte(int cls, int64_t d2, size_t spread, int asymmetric) {
    float s_min = asymmetric ? 0 : -(float)spread;
    float s_max = (float)spread;
    float d = __builtin_sqrtf((float)d2);
    float sdf;
    if (cls) {
        sdf = d - 0.f;
    } else {
        float val = d > 0.f ? d + -1.f : d;
        sdf = 0.f - val;
    }
    return ref_byte_from_signed(sdf, s_min, s_max);
}

inline void build_quant_table(size_t spread, int asymmetric, QuantTable* q) {
    int64_t cap = (int64_t)spread + 1;
    int64_t cap2 = cap * cap;
    if (cap2 > 0x7ffffffe) cap2 = 0x7ffffffe;
    q->cap2 = (int32_t)cap2;
    q->smin = asymmetric ? 0.f : -(float)spread;
    q->smax = (float)spread;
    q->scale = 255.f / (q->smax - q->smin);
    q->pad = 0;
    for (int k = 0; k <= 256; ++k) {
        // inside: smallest d2 in [0,cap2] with byte >= k
        if (k <= 255 && ref_byte(1, cap2, spread, asymmetric) >= k) {
            int64_t lo = 0, hi = cap2;  // invariant: byte(hi) >= k
            while (lo < hi) {
                int64_t mid = (lo + hi) / 2;
                if (ref_byte(1, mid, spread, asymmetric) >= k) hi = mid; else lo = mid + 1;
            }
            q->tin[k] = (int32_t)lo;
        } else {
            q->tin[k] = kBig;
        }
        // outside: smallest d2 in [0,cap2] with byte <= k
        if (k <= 255 && ref_byte(0, cap2, spread, asymmetric) <= k) {
            int64_t lo = 0, hi = cap2;
            while (lo < hi) {
                int64_t mid = (lo + hi) / 2;
                if (ref_byte(0, mid, spread, asymmetric) <= k) hi = mid; else lo = mid + 1;
            }
            q->tout[k] = (int32_t)lo;
        } else {
            q->tout[k] = k == 256 ? 0 : kBig;
        }
    }
}

}  // namespace chaq

// ---- unified level tables for the banded (fast) path ------------------------------------------------------
// For class c (1 = inside) the byte is a monotone step function of the clamped squared distance; written as
// a level m in [0,255] that never decreases with D2:
//     m(c, D2) = max { m : thr[c][m] <= D2 },   byte = c ? m : 255 - m
//   thr[1][m] = tin[m]          (inside : byte = m)
//   thr[0][m] = tout[255 - m]   (outside: byte = 255 - m)
// thr[c][0] = 0 and thr[c][256] = kBig (sentinel). cap_eff is the smallest integer C such that every D2 >= C*C
// gives the saturated byte of both classes: pixels (or row distances) that far away need not be told apart,
// so the banded path only looks C-1 rows up and down and only keeps row distances below C.
namespace chaq {

struct LevelTable {
    int32_t thr[2][260];
    int32_t cap_eff;   // C
    int32_t c2;        // C*C: squared distances are clamped here before the look-up
    int32_t pad[2];
};

inline void build_level_table(const QuantTable& q, LevelTable* t) {
    int64_t last = 1;  // largest finite threshold of either class
    for (int m = 0; m <= 255; ++m) {
        t->thr[1][m] = q.tin[m];
        t->thr[0][m] = q.tout[255 - m];
        if (q.tin[m] != kBig && q.tin[m] > last) last = q.tin[m];
        if (q.tout[255 - m] != kBig && q.tout[255 - m] > last) last = q.tout[255 - m];
    }
    for (int m = 256; m < 260; ++m) t->thr[0][m] = t->thr[1][m] = kBig;
    t->thr[0][0] = t->thr[1][0] = 0;
    int64_t c = 1;
    while (c * c < last) ++c;
    t->cap_eff = (int32_t)c;
    t->c2 = (int32_t)(c * c);
    t->pad[0] = t->pad[1] = 0;
}

// plain search, for the host emulation and for re-synchronising after a class change
CQ_HD int level_of(const int32_t* thr, int d2c) {
    int lo = 0, hi = 255;  // invariant: thr[lo] <= d2c
    while (lo < hi) {
        int mid = (lo + hi + 1) >> 1;
        if (thr[mid] <= d2c) lo = mid; else hi = mid - 1;
    }
    return lo;
}

}  // namespace chaq
