// envelope_core.h -- the per-column (pass 2) algorithm, written once for host and device.
//
// Replaces the second dist_transform_axis call (reference openmp/df.c:133 -> dist_transform_1d, df.c:29-97)
// for BOTH fields of openmp/sdfgen.c:277-289 in one sweep, in exact integer arithmetic.
//
// Input per column: one uint16 per row from pass 1 (rowscan): bit 15 = class c of the pixel, bits 0..14 =
// g = distance along the row to the nearest pixel of the OTHER class (>= cap means "none within reach").
// A pixel of class c needs D2 = min over rows r of (y-r)^2 + h_c(r), with h_c(r) = 0 when class(r) != c,
// else g(r)^2. A zero-height site hides everything behind it, so for a pixel in the vertical run [a,b] of
// its class only three kinds of candidate exist: the pixel above the run (a-1, height 0), the pixel below
// it (b+1, height 0), and the rows of the run itself with height g^2. Only the last needs a lower envelope
// of parabolas (Felzenszwalb-Huttenlocher / Meijster); it is kept as a bitmask of "rows on the hull" in a
// ring of 32-bit words (shared memory on the GPU), with g re-read from the intermediate when a site is
// needed again, so the state per column is O(lag/32) words however deep the hull gets.
//
// One sweep: row t is pushed, then row e = t - lag is finished (all sites within lag rows of e are known),
// where lag = spread (cap-1): sites further than that cannot change an 8-bit result, and every value
// >= cap^2 quantises identically (see tables.h). With cap >= number of rows the sweep degenerates to
// "push everything, then emit everything" and the result is the exact unwindowed EDT.
//
// Domination test (site b between a and c is never the minimum), all integers:
//   w*(h_b-h_a) - u*(h_c-h_b) >= u*w*(u+w),  u = b-a, w = c-b      (cross-multiplied form of df.c:67)
//
// The same source is compiled by g++ in tests/emul (host emulation, checked against the oracle) and by nvcc.
#pragma once
#include <stdint.h>

#if defined(__CUDACC__)
#define CQ_HD __host__ __device__ __forceinline__
#else
#define CQ_HD inline
#endif

namespace chaq {

constexpr int kBig = 0x7fffffff;

CQ_HD int cq_lowbit(uint32_t v) {  // index of lowest set bit, v != 0
#if defined(__CUDA_ARCH__)
    return __ffs((int)v) - 1;
#else
    return __builtin_ctz(v);
#endif
}
CQ_HD int cq_highbit(uint32_t v) {  // index of highest set bit, v != 0
#if defined(__CUDA_ARCH__)
    return 31 - __clz((int)v);
#else
    return 31 - __builtin_clz(v);
#endif
}
CQ_HD int cq_max(int a, int b) { return a > b ? a : b; }
CQ_HD int cq_min(int a, int b) { return a < b ? a : b; }

// Ctx must provide:
//   uint32_t ld(int row)                       intermediate value of this column at buffer row `row`
//   uint32_t hull_get(int i) / hull_set(i,v)   ring word i in [0,NW) of the hull bitmask
//   uint32_t chg_get(int i)  / chg_set(i,v)    ring word i of the class-change bitmask
//   void out(int row, int cls, int d2)         result for buffer row `row` (d2 == kBig: nothing in reach)
template <class Ctx, bool WIDE>
struct ColumnEnvelope {
    Ctx& c;
    int NW, cap, lag;
    int tw, twi;  // word number of the newest pushed row and its ring index
    int tbegin, tlast;
    // push side: the two newest hull sites of the current run
    int top_p, top_g, sec_p, sec_g, run0, cprev;
    // emission side
    bool einit;
    int ecls, lb, nb;  // class of the run being emitted, row above it (-1: none in reach), first row below it
    int cur_p, cur_g;  // hull site currently best (-1: run has no site yet)
    int nxt_p, nxt_g;  // its successor: >=0 cached, -1 unknown, -2 none as of tlast

    CQ_HD ColumnEnvelope(Ctx& ctx, int nw, int cap_) : c(ctx), NW(nw), cap(cap_) {}

    CQ_HD int ridx(int w) const {
        int i = twi - (tw - w);
        return i < 0 ? i + NW : i;
    }
    template <bool HULL>
    CQ_HD uint32_t rget(int w) {
        int i = ridx(w);
        return HULL ? c.hull_get(i) : c.chg_get(i);
    }
    // lowest set bit at a row in [from, to], or -1
    template <bool HULL>
    CQ_HD int scan_up(int from, int to) {
        if (from > to) return -1;
        int w = from >> 5, wend = to >> 5;
        uint32_t bits = rget<HULL>(w) & (0xFFFFFFFFu << (from & 31));
        for (;;) {
            if (w == wend) bits &= (0xFFFFFFFFu >> (31 - (to & 31)));
            if (bits) return (w << 5) + cq_lowbit(bits);
            if (w == wend) return -1;
            ++w;
            bits = rget<HULL>(w);
        }
    }
    // highest set bit at a row in [lo, from], or -1
    template <bool HULL>
    CQ_HD int scan_down(int from, int lo) {
        if (from < lo) return -1;
        int w = from >> 5, wend = lo >> 5;
        uint32_t bits = rget<HULL>(w) & (0xFFFFFFFFu >> (31 - (from & 31)));
        for (;;) {
            if (w == wend) bits &= (0xFFFFFFFFu << (lo & 31));
            if (bits) return (w << 5) + cq_highbit(bits);
            if (w == wend) return -1;
            --w;
            bits = rget<HULL>(w);
        }
    }

    CQ_HD bool dominated(int pa, int ga, int pb, int gb, int pc, int gc) const {
        if (WIDE) {
            long long u = pb - pa, w = pc - pb;
            long long ha = (long long)ga * ga, hb = (long long)gb * gb, hc = (long long)gc * gc;
            return w * (hb - ha) - u * (hc - hb) >= u * w * (u + w);
        } else {
            int u = pb - pa, w = pc - pb;
            int ha = ga * ga, hb = gb * gb, hc = gc * gc;
            return w * (hb - ha) - u * (hc - hb) >= u * w * (u + w);
        }
    }

    CQ_HD void push(int t) {
        int w = t >> 5;
        if (t == tbegin) {
            tw = w; twi = 0;
            c.hull_set(0, 0u); c.chg_set(0, 0u);
        } else if (w != tw) {
            tw = w; twi = (twi + 1 == NW) ? 0 : twi + 1;
            c.hull_set(twi, 0u); c.chg_set(twi, 0u);
        }
        tlast = t;
        uint32_t v = c.ld(t);
        int cls = (int)(v >> 15) & 1, g = (int)(v & 0x7fffu);
        uint32_t bit = 1u << (t & 31);
        if (t != tbegin && cls != cprev) {
            c.chg_set(twi, c.chg_get(twi) | bit);
            run0 = t; top_p = -1; sec_p = -1;
            if (einit && nb == kBig) nb = t;
        }
        cprev = cls;
        if (g >= cap) return;
        int popfloor = kBig;
        int lowb = cq_max(run0, t - 2 * lag);
        // sites older than 2*lag rows are out of the ring and out of reach of every row still to be
        // finished: forget them (they can neither be popped nor serve as the predecessor in a test)
        if (top_p < lowb) { top_p = -1; sec_p = -1; }
        else if (sec_p < lowb) sec_p = -1;
        while (top_p >= 0 && sec_p >= 0 && dominated(sec_p, sec_g, top_p, top_g, t, g)) {
            int i = ridx(top_p >> 5);
            c.hull_set(i, c.hull_get(i) & ~(1u << (top_p & 31)));
            popfloor = top_p;
            top_p = sec_p; top_g = sec_g;
            sec_p = scan_down<true>(top_p - 1, lowb);
            sec_g = sec_p >= 0 ? (int)(c.ld(sec_p) & 0x7fffu) : 0;
        }
        c.hull_set(twi, c.hull_get(twi) | bit);
        sec_p = top_p; sec_g = top_g;
        top_p = t; top_g = g;
        if (einit && nb == kBig) {  // the run being emitted is still open: its hull just changed
            if (cur_p < 0 || popfloor <= cur_p) { cur_p = t; cur_g = g; nxt_p = -2; }
            else if (nxt_p == -2 || (nxt_p >= 0 && popfloor <= nxt_p)) { nxt_p = t; nxt_g = g; }
        }
    }

    CQ_HD void setup_cur(int from) {
        int hi = nb == kBig ? tlast : cq_min(nb - 1, tlast);
        int p = scan_up<true>(from, hi);
        cur_p = p;
        cur_g = p >= 0 ? (int)(c.ld(p) & 0x7fffu) : 0;
        nxt_p = p >= 0 ? -1 : -2;
    }

    CQ_HD void emit(int e) {
        if (!einit) {
            einit = true;
            ecls = (int)(c.ld(e) >> 15) & 1;
            int lo = cq_max(tbegin, e - lag);
            int a = scan_down<false>(e, lo + 1);
            lb = a >= 0 ? a - 1 : -1;
            nb = scan_up<false>(e + 1, tlast);
            if (nb < 0) nb = kBig;
            setup_cur(a >= 0 ? a : lo);
        } else if (e == nb) {
            ecls ^= 1;
            lb = e - 1;
            nb = scan_up<false>(e + 1, tlast);
            if (nb < 0) nb = kBig;
            setup_cur(e);
        }
        // A site more than lag rows above e can no longer matter, and its neighbours in the bitmask may be
        // sites that only survived because their own predecessor had already been forgotten (pops look back
        // 2*lag rows at most), so the walk below must not start from it: restart at the first site in reach.
        if (cur_p >= 0 && cur_p < e - lag) setup_cur(e - lag);
        int cand = kBig;
        if (cur_p >= 0) {
            int d = e - cur_p;
            int fc = d * d + cur_g * cur_g;
            for (;;) {
                if (nxt_p == -1) {
                    int hi = nb == kBig ? tlast : cq_min(nb - 1, tlast);
                    int p = scan_up<true>(cq_max(cur_p + 1, e - lag), hi);
                    if (p < 0) nxt_p = -2;
                    else { nxt_p = p; nxt_g = (int)(c.ld(p) & 0x7fffu); }
                }
                if (nxt_p < 0) break;
                int dn = e - nxt_p;
                int fn = dn * dn + nxt_g * nxt_g;
                if (fn > fc) break;
                cur_p = nxt_p; cur_g = nxt_g; fc = fn; nxt_p = -1;
            }
            cand = fc;
        }
        if (lb >= 0) { int d = e - lb; cand = cq_min(cand, d * d); }
        if (nb != kBig) { int d = nb - e; cand = cq_min(cand, d * d); }
        c.out(e, ecls, cand);
    }

    // buffer rows [0,nrows); results for rows [y0,y1)
    CQ_HD void run(int nrows, int y0, int y1) {
        lag = cq_min(cap - 1, nrows);
        tbegin = cq_max(0, y0 - lag);
        int tend = cq_min(nrows, y1 + lag);
        tw = 0; twi = 0; tlast = tbegin;
        top_p = sec_p = -1; top_g = sec_g = 0; run0 = tbegin; cprev = 0;
        einit = false; ecls = 0; lb = -1; nb = kBig; cur_p = -1; cur_g = 0; nxt_p = -2; nxt_g = 0;
        int stop = y1 + lag;
        for (int step = tbegin; step < stop; ++step) {
            if (step < tend) push(step);
            int e = step - lag;
            if (e >= y0) emit(e);
        }
    }
};

// ring words needed so that rows [t-2*lag, t] stay addressable
CQ_HD int ring_words(int cap, int nrows) {
    int lag = cq_min(cap - 1, nrows);
    int need = (2 * lag + 31) / 32 + 1;
    int have = (nrows + 31) / 32 + 1;
    return cq_min(need, have);
}

}  // namespace chaq
