// hull_core.h -- the banded column pass (pass 2) as build / merge over a hull bitmask in global memory, written once
// for host and device; the evaluation of the finished hull is paint_core.h. After the band structure of PBA (Cao et al.).
//
// Replaces the second dist_transform_axis call (reference openmp/df.c:133 -> dist_transform_1d, df.c:29-97) for BOTH
// fields of openmp/sdfgen.c:277-289, in exact integer arithmetic.
//
// Input per column and row r (from pass 1): the pixel's class c(r) and G(r) = distance along the row to the nearest
// pixel of the OTHER class, stored as g8 = min(G, C) - 1 (C = cap_eff of tables.h, g8 == C-1: none within reach).
// A pixel of class c at row y needs D2(y) = min over opposite-class pixels q of |p - q|^2. One lower envelope serves
// both classes:
//     D2(y) = min( V(y)^2 , min_r (y - r)^2 + G(r)^2 )             (clamped at C^2)
// where V(y) is the distance along the column to the nearest row of the other class and r runs over ALL rows
// regardless of class. Proof sketch: for r of y's class the term is the squared distance to a real opposite pixel
// (the one G(r) points at); for r of the other class the term is >= (y-r)^2 >= V(y)^2; and the true nearest opposite
// pixel is either in y's column (V) or is the row-nearest opposite pixel of a same-class row. So the class bit
// matters only for V and for the sign of the result, and the envelope is the plain Felzenszwalb-Huttenlocher one
// (df.c:57-79) over heights G^2, kept as a bitmask of "rows on the hull".
//
//   build   each (32-column group, run of <= 1024 rows) builds the Felzenszwalb-Huttenlocher hull of ITS rows only
//           (reference openmp/df.c:57-79, integer cross-multiplied domination test), no context rows;
//   merge   per column, the junctions between consecutive runs are repaired: sites next to a junction that are
//           dominated across it are removed (alternating pops on both sides);
//   paint   (paint_core.h) each (group, run of rows) evaluates the finished, immutable hull (df.c:83-96), adds the
//           in-column distance V to the nearest row of the other class, clamps at C^2 and quantises.
//
// Because every height is < C^2 and every result is clamped at C^2 = (L+1)^2, only triples of consecutive
// sites spanning <= 2L rows matter: a row's walk is restarted at the first site within L rows above it, a site
// more than L rows below it yields >= C^2, so a dominated site can only mislead the walk if it and both its
// dominators lie within one window of 2L+1 rows. Build and merge therefore guarantee: no site is dominated by
// its two neighbours on the hull when those are <= 2L rows apart. Sites that are removed are never the strict
// minimum anywhere (removal is by two sites that stay or are themselves removed later by the same rule).
//
// The hull is one bit per pixel, word (32-row block b, column x); g8 is re-read from pass 1's record whenever
// a site's height is needed again.
// The same source is compiled by g++ in tests/emul (checked against the oracle) and by nvcc.
#pragma once
#include <stdint.h>
#include "envelope_core.h"  // CQ_HD, cq_min/cq_max, cq_lowbit/cq_highbit

namespace chaq {

constexpr int kFarRow = 1 << 28;  // "no such row"; survives +- image-size arithmetic without overflow
#define CQ_UNLIKELY(x) __builtin_expect(!!(x), 0)


// site b is never the minimum between its neighbours a < b < c (cross-multiplied df.c:67).
// WIDE = false: int32 (|terms| < 2^29 while spans stay <= 1024 rows and heights <= 2^16: the spreads <= 256 of the
// byte-wide intermediate). WIDE = true: 64-bit, for row distances up to 1024 (16-bit intermediate) and longer runs.
template <bool WIDE>
CQ_HD bool cq_dominated(int ap, int ah, int bp, int bh, int cp, int ch) {
    if (WIDE) {
        const long long u = bp - ap, w = cp - bp;
        return w * (bh - ah) - u * (ch - bh) >= u * w * (u + w);
    }
    const int u = bp - ap, w = cp - bp;
    return w * (bh - ah) - u * (ch - bh) >= u * w * (u + w);
}
constexpr int hull_max_run_blocks(bool wide) { return wide ? 128 : 32; }  // rows per build run / 32

// ---------------------------------------------------------------------------------------------------------
// build. Ctx provides:
//   void     block_begin(int blk)         rows of buffer block blk are readable through g_push from now on
//   bool     block_quiet(int L)           no g8 < L in that block, for every column sharing control flow
//   uint32_t g_push(int j)                g8 of row j of that block
//   uint32_t g_row(int r)                 g8 of buffer row r (any row of this run)
//   uint32_t bits_ld(int i) / void bits_st(int i, uint32_t)   hull word of block b0 + i of this run
//   void     block_end(int blk, uint32_t word)                 the block's pushes are done (word = its hull bits now)
template <class Ctx, bool WIDE = false>
struct HullBuild {
    Ctx& c;
    int L, b0, pblk;
    uint32_t hwp;
    // the three newest sites; p == -1: none, thr_p == -2: not loaded (look it up in the bitmask when needed)
    int top_p, top_h, sec_p, sec_h, thr_p, thr_h;

    CQ_HD HullBuild(Ctx& ctx, int lag) : c(ctx), L(lag) {}

    CQ_HD uint32_t word(int blk) { return blk == pblk ? hwp : c.bits_ld(blk - b0); }
    CQ_HD int height(int r) {
        const int g = (int)c.g_row(r) + 1;
        return g * g;
    }
    CQ_HD void clear(int r) {
        const uint32_t bit = 1u << (r & 31);
        const int b = r >> 5;
        if (b == pblk) hwp &= ~bit;
        else c.bits_st(b - b0, c.bits_ld(b - b0) & ~bit);
    }
    CQ_HD int down(int from) {  // highest site at a row <= from in this run, or -1
        if (from < b0 * 32) return -1;
        int b = from >> 5;
        uint32_t bits = word(b) & (0xFFFFFFFFu >> (31 - (from & 31)));
        for (;;) {
            if (bits) return (b << 5) + cq_highbit(bits);
            if (b == b0) return -1;
            --b;
            bits = word(b);
        }
    }
    CQ_HD void pop() {
        clear(top_p);
        top_p = sec_p; top_h = sec_h;
        if (thr_p == -2) {
            thr_p = down(top_p - 1);
            if (thr_p >= 0) thr_h = height(thr_p);
        }
        sec_p = thr_p; sec_h = thr_h;
        thr_p = -2;
    }
    CQ_HD void push(int rr, int h) {
        {
            while (sec_p >= 0 && cq_dominated<WIDE>(sec_p, sec_h, top_p, top_h, rr, h)) pop();
            hwp |= 1u << (rr & 31);
            thr_p = sec_p; thr_h = sec_h;
            sec_p = top_p; sec_h = top_h;
            top_p = rr; top_h = h;
        }
    }
    // blocks [blk0, blk1) of a buffer of nrows rows
    CQ_HD void run(int blk0, int blk1, int nrows) {
        b0 = blk0;
        top_p = sec_p = thr_p = -1; top_h = sec_h = thr_h = 0;
        for (int blk = blk0; blk < blk1; ++blk) {
            pblk = blk;
            hwp = 0u;
            c.block_begin(blk);
            if (!c.block_quiet(L)) {
                // A site that lies on or above the chord of its two row neighbours (both sites) is never the strict
                // minimum anywhere: 2 h_j >= h_{j-1} + h_{j+1} + 2 with h = G^2 (the domination test for u = w = 1).
                // It would be popped by the very next push; not pushing it saves the push and the pop. Whether the
                // neighbours themselves end up on the hull does not matter (a point above a chord of input points is
                // not a vertex of their lower hull). Rows at the ends of a block are always pushed.
                const int jn = cq_min(32, nrows - blk * 32);
                const int big = 1 << 24;  // "no site": never a dominator, never pushed
                int hp = big, hc, hn;
                {
                    const int g0 = (int)c.g_push(0);
                    hc = g0 < L ? (g0 + 1) * (g0 + 1) : big;
                }
                for (int j = 0; j < jn; ++j) {
                    hn = big;
                    if (j + 1 < jn) {
                        const int g1 = (int)c.g_push(j + 1);
                        hn = g1 < L ? (g1 + 1) * (g1 + 1) : big;
                    }
                    // Likewise a site that is never the minimum between the newest site and its successor row (both input
                    // points) is not a vertex: after a step up in G the next ~sqrt(2G) rows are of this kind, and each of
                    // them would be pushed and then popped by the row after it.
                    if (hc < big && !(2 * hc >= hp + hn + 2) &&
                        !(top_p >= 0 && hn < big && cq_dominated<WIDE>(top_p, top_h, blk * 32 + j, hc, blk * 32 + j + 1, hn)))
                        push(blk * 32 + j, hc);
                    hp = hc; hc = hn;
                }
            }
            c.bits_st(blk - b0, hwp);
            c.block_end(blk, hwp);
        }
    }
};

// Junctions repaired at the same time (different threads, plain loads and stores on the hull words) must not touch
// the same WORD. The repair of the junction at row J (a multiple of 32) clears sites in rows [J-2L+1, J+2L-2], i.e.
// words J/32 - ceil((2L-1)/32) .. J/32 + floor((2L-2)/32); junctions k and k + nphase are pb*nphase words apart.
// Returns the number of rounds: round i repairs every nphase-th junction starting with the i-th.
CQ_HD int hull_merge_phases(int L, int pb) {
    if (L < 1) return 1;
    const int below = (2 * L - 2) >> 5, above = (2 * L - 1 + 31) >> 5;
    return (below + above) / pb + 1;
}

// ---------------------------------------------------------------------------------------------------------
// merge. Ctx provides: uint32_t h_ld(int blk) / void h_st(int blk, uint32_t) (this column's hull words, all
// blocks of the buffer) and uint32_t g_row(int r).
template <class Ctx, bool WIDE = false>
struct HullMerge {
    Ctx& c;
    int L;
    int lo_row, hi_row;  // the repair looks at sites of rows [lo_row, hi_row] only (a whole buffer, or one build run)
    int run_rows = 0;    // rows per build run if the hulls on either side of a junction are whole runs (0: unknown, walk site by site)
    CQ_HD HullMerge(Ctx& ctx, int lag, int rows) : c(ctx), L(lag), lo_row(0), hi_row(rows - 1) {}
    CQ_HD HullMerge(Ctx& ctx, int lag, int lo, int hi) : c(ctx), L(lag), lo_row(lo), hi_row(hi) {}
    CQ_HD int height(int r) {
        const int g = (int)c.g_row(r) + 1;
        return g * g;
    }
    CQ_HD int up(int from, int to) {  // lowest site in [from, to], or -1
        if (from > to) return -1;
        int b = from >> 5;
        const int bend = to >> 5;
        uint32_t bits = c.h_ld(b) & (0xFFFFFFFFu << (from & 31));
        for (;;) {
            if (b == bend) bits &= (0xFFFFFFFFu >> (31 - (to & 31)));
            if (bits) return (b << 5) + cq_lowbit(bits);
            if (b == bend) return -1;
            ++b;
            bits = c.h_ld(b);
        }
    }
    CQ_HD int down(int from, int lo) {  // highest site in [lo, from], or -1
        if (from < lo) return -1;
        int b = from >> 5;
        const int bend = lo >> 5;
        uint32_t bits = c.h_ld(b) & (0xFFFFFFFFu >> (31 - (from & 31)));
        for (;;) {
            if (b == bend) bits &= (0xFFFFFFFFu << (lo & 31));
            if (bits) return (b << 5) + cq_highbit(bits);
            if (b == bend) return -1;
            --b;
            bits = c.h_ld(b);
        }
    }
    CQ_HD void clear(int r) { c.h_st(r >> 5, c.h_ld(r >> 5) & ~(1u << (r & 31))); }

    static CQ_HD int popc32(uint32_t v) {
#if defined(__CUDA_ARCH__)
        return __popc(v);
#else
        return __builtin_popcount(v);
#endif
    }
    // the n-th site (n >= 1) below row `from`, no lower than row lo; -1 if there are fewer
    CQ_HD int nth_down(int from, int n, int lo) {
        int r = from - 1;
        if (r < lo) return -1;
        int b = r >> 5;
        const int bend = lo >> 5;
        uint32_t bits = c.h_ld(b) & (0xFFFFFFFFu >> (31 - (r & 31)));
        for (;;) {
            if (b == bend) bits &= (0xFFFFFFFFu << (lo & 31));
            const int cnt = popc32(bits);
            if (cnt >= n) {
                for (int k = 1; k < n; ++k) bits &= ~(1u << cq_highbit(bits));
                return (b << 5) + cq_highbit(bits);
            }
            n -= cnt;
            if (b == bend) return -1;
            --b;
            bits = c.h_ld(b);
        }
    }
    // the n-th site (n >= 1) above row `from`, no higher than row hi; -1 if there are fewer
    CQ_HD int nth_up(int from, int n, int hi) {
        int r = from + 1;
        if (r > hi) return -1;
        int b = r >> 5;
        const int bend = hi >> 5;
        uint32_t bits = c.h_ld(b) & (0xFFFFFFFFu << (r & 31));
        for (;;) {
            if (b == bend) bits &= (0xFFFFFFFFu >> (31 - (hi & 31)));
            const int cnt = popc32(bits);
            if (cnt >= n) {
                for (int k = 1; k < n; ++k) bits &= bits - 1u;
                return (b << 5) + cq_lowbit(bits);
            }
            n -= cnt;
            if (b == bend) return -1;
            ++b;
            bits = c.h_ld(b);
        }
    }
    // sites of rows [from, to] are removed
    CQ_HD void clear_rows(int from, int to) {
        for (int b = from >> 5; b <= (to >> 5); ++b) {
            uint32_t m = 0xFFFFFFFFu;
            if (b == (from >> 5)) m &= 0xFFFFFFFFu << (from & 31);
            if (b == (to >> 5)) m &= 0xFFFFFFFFu >> (31 - (to & 31));
            const uint32_t w = c.h_ld(b);
            if (w & m) c.h_st(b, w & ~m);
        }
    }

    // Along one side of a junction the sites form a convex chain (proper inside the 2L window), so "site i is dominated by
    // its chain neighbour i+1 and the fixed site d across the junction" holds for a prefix of the chain and for no site
    // after it: the supporting lines of a convex chain, evaluated at a point beyond its end, fall monotonically along the
    // chain. The prefix is found by galloping (i = 0, 1, 3, 7, ...) and bisection instead of site by site: a repair that
    // removes n sites takes O(log n) dependent look-ups of row distances, and the common repairs (0..2 sites) take the same
    // tests as the one-by-one walk. These look-ups were what the merge kernel's duration consisted of (up to ~100 in one
    // column at 16384^2).
    // The chain is only known to be convex inside ONE build run (the junction at the run's other end may not have been
    // repaired yet), so galloping stays inside the run next to the junction (rows >= lo_g); if it runs out of sites
    // there, the walk goes on site by site down to lo.
    // Up side: s_0 = u (height uh), s_i = the i-th site below it. Returns the number of sites to remove;
    // nu / nuh = s_k, the first that stays.
    CQ_HD int gallop_up(int u, int uh, int d, int dh, int lo, int lo_g, int& nu, int& nuh) {
        int k = 0, kp = u, kh = uh;        // P(i) holds for all i < k; s_k = kp
        int step = 1;
        int fail = -1;                     // smallest index known to fail (-1: none yet)
        bool fail_absent = false;          // ... because the chain ran out of sites inside the run
        while (kp >= lo_g) {
            const int j = fail < 0 ? k + step - 1 : k + (fail - k) / 2;   // gallop, then bisect in [k, fail)
            const int sj = j == k ? kp : nth_down(kp, j - k, lo_g);
            const int sj1 = sj >= 0 ? nth_down(sj, 1, lo_g) : -1;
            bool ok = false;
            int sjh = 0, sj1h = 0;
            if (sj1 >= 0) {
                sjh = j == k ? kh : height(sj);
                sj1h = height(sj1);
                ok = cq_dominated<WIDE>(sj1, sj1h, sj, sjh, d, dh);
            }
            if (ok) {
                k = j + 1; kp = sj1; kh = sj1h;
                if (fail < 0) step *= 2;
            } else {
                fail = j;
                fail_absent = sj1 < 0;
            }
            if (fail >= 0 && fail == k) break;
        }
        if (fail < 0 || fail_absent) {     // site by site from s_k on (all of it when run_rows is unknown)
            for (;;) {
                const int n = down(kp - 1, lo);
                if (n < 0) break;
                const int nh = height(n);
                if (!cq_dominated<WIDE>(n, nh, kp, kh, d, dh)) break;
                ++k; kp = n; kh = nh;
            }
        }
        nu = kp; nuh = kh;
        return k;
    }
    // Down side: t_0 = d, t_i = the i-th site above it; u fixed; galloping inside rows <= hi_g.
    CQ_HD int gallop_down(int u, int uh, int d, int dh, int hi, int hi_g, int& nd, int& ndh) {
        int k = 0, kp = d, kh = dh;
        int step = 1;
        int fail = -1;
        bool fail_absent = false;
        while (kp <= hi_g) {
            const int j = fail < 0 ? k + step - 1 : k + (fail - k) / 2;
            const int tj = j == k ? kp : nth_up(kp, j - k, hi_g);
            const int tj1 = tj >= 0 ? nth_up(tj, 1, hi_g) : -1;
            bool ok = false;
            int tjh = 0, tj1h = 0;
            if (tj1 >= 0) {
                tjh = j == k ? kh : height(tj);
                tj1h = height(tj1);
                ok = cq_dominated<WIDE>(u, uh, tj, tjh, tj1, tj1h);
            }
            if (ok) {
                k = j + 1; kp = tj1; kh = tj1h;
                if (fail < 0) step *= 2;
            } else {
                fail = j;
                fail_absent = tj1 < 0;
            }
            if (fail >= 0 && fail == k) break;
        }
        if (fail < 0 || fail_absent) {
            for (;;) {
                const int n = up(kp + 1, hi);
                if (n < 0) break;
                const int nh = height(n);
                if (!cq_dominated<WIDE>(u, uh, kp, kh, n, nh)) break;
                ++k; kp = n; kh = nh;
            }
        }
        nd = kp; ndh = kh;
        return k;
    }

    // rows < J were built (and merged) as one hull, rows >= J as another: make the union proper within 2L
    CQ_HD void junction(int J) {
        const int span = 2 * L;
        int d1 = up(J, cq_min(hi_row, J + span - 1));
        if (d1 < 0) return;
        int u1 = down(J - 1, cq_max(lo_row, d1 - span));
        if (u1 < 0) return;
        int d1h = height(d1), u1h = height(u1);
        for (;;) {
            bool changed = false;
            int nu, nuh;
            const int lo = cq_max(lo_row, d1 - span);
            // run_rows unknown: lo_g above every site, so the galloping loop is not entered
            if (gallop_up(u1, u1h, d1, d1h, lo, run_rows > 0 ? cq_max(lo, J - run_rows) : J, nu, nuh) > 0) {
                clear_rows(nu + 1, u1);
                u1 = nu; u1h = nuh;
                changed = true;
            }
            int nd, ndh;
            const int hi = cq_min(hi_row, u1 + span);
            if (gallop_down(u1, u1h, d1, d1h, hi, run_rows > 0 ? cq_min(hi, J + run_rows - 1) : J - 1, nd, ndh) > 0) {
                clear_rows(d1, nd - 1);
                d1 = nd; d1h = ndh;
                changed = true;
            }
            if (!changed || d1 - u1 > span) break;
        }
    }
};


}  // namespace chaq
