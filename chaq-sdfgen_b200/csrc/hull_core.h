// hull_core.h -- the banded column pass (pass 2) as three steps over a hull bitmask in global memory, written
// once for host and device. Same mathematics as sweep_core.h (see there for the proof that ONE lower envelope
// over heights G^2 serves both classes); different organisation, after the band structure of PBA (Cao et al.):
//
//   build   each (32-column group, run of <= 1024 rows) builds the Felzenszwalb-Huttenlocher hull of ITS rows only
//           (reference openmp/df.c:57-79, integer cross-multiplied domination test), no context rows;
//   merge   per column, the junctions between consecutive runs are repaired top to bottom: sites next to a
//           junction that are dominated across it are removed (alternating pops on both sides);
//   emit    each (group, run of rows) walks the finished, immutable hull (df.c:83-96) with a cursor, adds the
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
// a site's height is needed again. All int32: spans <= 1024 rows, heights <= 2^16 -> |terms| < 2^29.
#pragma once
#include <stdint.h>
#include "envelope_core.h"  // CQ_HD, cq_min/cq_max, cq_lowbit/cq_highbit
#include "sweep_core.h"     // kFarRow, kNoCurH, kNoNxtH, CQ_UNLIKELY

namespace chaq {

#ifndef CQ_HULL_WIDE
constexpr int kHullMaxPieceBlocks = 32;  // build runs are at most 1024 rows (int32 range of the domination test)

// site b is never the minimum between its neighbours a < b < c (cross-multiplied df.c:67)
CQ_HD bool cq_dominated(int ap, int ah, int bp, int bh, int cp, int ch) {
    const int u = bp - ap, w = cp - bp;
    return w * (bh - ah) - u * (ch - bh) >= u * w * (u + w);
}
#else
// Wide variant (not compiled into the library yet; tests/emul builds it to check the organisation for L up to 1024,
// i.e. spreads beyond the banded path): u*w*(u+w) passes 2^31 once spans reach ~1300 rows, so the test is 64-bit,
// and runs may be longer.
constexpr int kHullMaxPieceBlocks = 128;
CQ_HD bool cq_dominated(int ap, int ah, int bp, int bh, int cp, int ch) {
    const long long u = bp - ap, w = cp - bp;
    return w * (bh - ah) - u * (ch - bh) >= u * w * (u + w);
}
#endif

// ---------------------------------------------------------------------------------------------------------
// build. Ctx provides:
//   void     block_begin(int blk)         rows of buffer block blk are readable through g_push from now on
//   bool     block_quiet(int L)           no g8 < L in that block, for every column sharing control flow
//   uint32_t g_push(int j)                g8 of row j of that block
//   uint32_t g_row(int r)                 g8 of buffer row r (any row of this run)
//   uint32_t bits_ld(int i) / void bits_st(int i, uint32_t)   hull word of block b0 + i of this run
//   void     block_end(int blk, uint32_t word)                 the block's pushes are done (word = its hull bits now)
template <class Ctx>
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
    CQ_HD void push(int rr, int j) {
        const int g8 = (int)c.g_push(j);
        if (g8 < L) {
            const int h = (g8 + 1) * (g8 + 1);
            while (sec_p >= 0 && cq_dominated(sec_p, sec_h, top_p, top_h, rr, h)) pop();
            hwp |= 1u << j;
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
                const int jn = cq_min(32, nrows - blk * 32);
                for (int j = 0; j < jn; ++j) push(blk * 32 + j, j);
            }
            c.bits_st(blk - b0, hwp);
            c.block_end(blk, hwp);
        }
    }
};

// ---------------------------------------------------------------------------------------------------------
// merge. Ctx provides: uint32_t h_ld(int blk) / void h_st(int blk, uint32_t) (this column's hull words, all
// blocks of the buffer) and uint32_t g_row(int r).
template <class Ctx>
struct HullMerge {
    Ctx& c;
    int L;
    int lo_row, hi_row;  // the repair looks at sites of rows [lo_row, hi_row] only (a whole buffer, or one build run)
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

    // rows < J were built (and merged) as one hull, rows >= J as another: make the union proper within 2L
    CQ_HD void junction(int J) {
        const int span = 2 * L;
        int d1 = up(J, cq_min(hi_row, J + span - 1));
        if (d1 < 0) return;
        int u1 = down(J - 1, cq_max(lo_row, d1 - span));
        if (u1 < 0) return;
        int d1h = height(d1), u1h = height(u1);
        int u2 = -2, u2h = 0, d2 = -2, d2h = 0;  // -2: look it up
        for (;;) {
            bool changed = false;
            if (u2 == -2) {
                u2 = down(u1 - 1, cq_max(lo_row, d1 - span));
                if (u2 >= 0) u2h = height(u2);
            }
            if (u2 >= 0 && cq_dominated(u2, u2h, u1, u1h, d1, d1h)) {
                clear(u1);
                u1 = u2; u1h = u2h; u2 = -2;
                if (d2 >= 0 && d2 - u1 > span) d2 = -1;
                changed = true;
            }
            if (d2 == -2) {
                d2 = up(d1 + 1, cq_min(hi_row, u1 + span));
                if (d2 >= 0) d2h = height(d2);
            }
            if (d2 >= 0 && cq_dominated(u1, u1h, d1, d1h, d2, d2h)) {
                clear(d1);
                d1 = d2; d1h = d2h; d2 = -2;
                if (u2 >= 0 && d1 - u2 > span) u2 = -1;
                changed = true;
            }
            if (!changed || d1 - u1 > span) break;
        }
    }
};


// ---------------------------------------------------------------------------------------------------------
// emit. Ctx provides (blk = 32-row block of the buffer; blocks past the buffer read as "no site", class of the
// last real row):
//   uint32_t h_ld(int blk)                 this column's finished hull word
//   uint32_t g_row(int r)
//   uint32_t cls_fetch(int blk)            class bits of this column for the block's rows (bit j = row 32*blk+j)
//   uint32_t cls_ld(int slot) / void cls_st(int slot, uint32_t)   ring of kClsRing class words (slot = blk & mask)
//   int      q_flag(int blk)               0/1: block holds no site in any column of the group and is all class
//                                          0/1 there; 2: anything else; 3: block outside the buffer
//   void     out(int cls, int d2c) / void out_run(int cls, int d2c, int n)
#ifndef CQ_HULL_WIDE
constexpr int kClsRing = 16;  // class words kept ahead of the row being finished: ceil(L/32) + 2 <= 10
#else
constexpr int kClsRing = 64;  // ceil(1024/32) + 2 = 34
#endif

template <class Ctx>
struct HullEmit {
    Ctx& c;
    int L, C2, reach;
    int ecls, upd, nb, lastknown;
    int cur_p, cur_h, nxt_p, nxt_h;
    int eblk;       // cursor: the sites after nxt (after cur if there is no nxt) are the bits of ehw in block eblk,
    uint32_t ehw;   // then the blocks after eblk
    int n0, n1, n2; // blocks of the quiet window with flag 0 / 1 / 2
    int ev;         // see set_ev

    CQ_HD HullEmit(Ctx& ctx, int lag, int c2) : c(ctx), L(lag), C2(c2), reach((lag + 31) / 32) {}

    CQ_HD uint32_t cls_word(int blk) { return c.cls_ld(blk & (kClsRing - 1)); }
    // first row in [from, to] whose class differs from cls, or -1 (rows must be in the ring)
    CQ_HD int cls_up(int from, int to, int cls) {
        if (from > to) return -1;
        const uint32_t flip = cls ? 0xFFFFFFFFu : 0u;
        int b = from >> 5;
        const int bend = to >> 5;
        uint32_t bits = (cls_word(b) ^ flip) & (0xFFFFFFFFu << (from & 31));
        for (;;) {
            if (b == bend) bits &= (0xFFFFFFFFu >> (31 - (to & 31)));
            if (bits) return (b << 5) + cq_lowbit(bits);
            if (b == bend) return -1;
            ++b;
            bits = cls_word(b) ^ flip;
        }
    }
    CQ_HD void class_flip(int er) {
        ecls ^= 1;
        upd = 1;
        const int n = cls_up(er + 1, lastknown, ecls);
        nb = n >= 0 ? n : kFarRow;
    }
    // class words of block blk enter the ring; an open-ended run may end there
    CQ_HD void cls_push(int blk) {
        const uint32_t w = c.cls_fetch(blk);
        c.cls_st(blk & (kClsRing - 1), w);
        lastknown = blk * 32 + 31;
        if (nb == kFarRow) {
            const uint32_t x = w ^ (ecls ? 0xFFFFFFFFu : 0u);
            if (x) { nb = blk * 32 + cq_lowbit(x); ev = cq_min(ev, nb); }
        }
    }
    CQ_HD int height(int r) {
        const int g = (int)c.g_row(r) + 1;
        return g * g;
    }
    // nxt := the next site of the hull if one is known at a row <= limit or in the block being scanned
    CQ_HD void next_site(int limit) {
        for (;;) {
            if (ehw) {
                const int j = cq_lowbit(ehw);
                ehw &= ehw - 1u;
                nxt_p = eblk * 32 + j;
                nxt_h = height(nxt_p);
                return;
            }
            if ((eblk + 1) * 32 > limit) { nxt_p = -1; nxt_h = kNoNxtH; return; }
            ++eblk;
            ehw = c.h_ld(eblk);
        }
    }
    // ev = the first row at which something other than the plain walk has to happen: the class run ends (nb), the
    // current site drops out of reach (cur_p + L + 1), or the successor has to be looked up again because a new
    // block of the hull has come within reach. One compare per row replaces three.
    CQ_HD void set_ev(int er) {
        int t = nb;
        if (cur_p >= 0) t = cq_min(t, cur_p + L + 1);
        if (nxt_p < 0) t = cq_min(t, ehw != 0u ? er : (eblk + 1) * 32 - L);
        ev = t;
    }
    CQ_HD int advance(int er, int fn) {
        int fc;
        do {
            cur_p = nxt_p; cur_h = nxt_h; fc = fn;
            next_site(er + L);
            const int dn = er - nxt_p;
            fn = dn * dn + nxt_h;
        } while (fn <= fc);
        set_ev(er + 1);
        return fc;
    }
    CQ_HD void housekeeping(int er) {
        if (er == nb) class_flip(er);
        if (nxt_p < 0 && (ehw != 0u || (eblk + 1) * 32 <= er + L)) next_site(er + L);
        // a site more than L rows above er no longer matters: move on to the first one within reach
        const int lowrow = cq_max(er - L, 0);
        while ((unsigned)cur_p < (unsigned)lowrow) {
            cur_p = nxt_p; cur_h = nxt_p < 0 ? kNoCurH : nxt_h;
            if (nxt_p >= 0) next_site(er + L);
        }
        set_ev(er + 1);
    }
    CQ_HD void emit(int er) {
        ++upd;
        if (CQ_UNLIKELY(er >= ev)) housekeeping(er);
        const int d = er - cur_p;
        int fc = d * d + cur_h;
        const int dn = er - nxt_p;
        const int fn = dn * dn + nxt_h;
        if (fn <= fc) fc = advance(er, fn);
        const int v = cq_min(cq_min(upd, nb - er), L + 1);
        c.out(ecls, cq_min(fc, v * v));
    }
    CQ_HD void q_add(int blk, int s) {
        const int f = c.q_flag(blk);
        if (f == 0) n0 += s; else if (f == 1) n1 += s; else if (f == 2) n2 += s;
    }

    // results for buffer rows [y0, y1)
    CQ_HD void run(int y0, int y1) {
        const int B0 = y0 >> 5;
        // class state at y0: the ring gets blocks B0 .. B0+reach, the run's start is looked up above y0
        nb = 0;  // no open-ended check while the ring fills
        for (int b = B0; b <= B0 + reach; ++b) cls_push(b);
        ecls = (int)(cls_word(B0) >> (y0 & 31)) & 1;
        {
            const uint32_t flip = ecls ? 0xFFFFFFFFu : 0u;
            upd = kFarRow;
            const int blo = cq_max(0, (y0 - L - 1) >> 5);
            for (int b = B0; b >= blo; --b) {
                uint32_t x = (b == B0 ? cls_word(b) : c.cls_fetch(b)) ^ flip;
                if (b == B0) x &= (y0 & 31) ? (0xFFFFFFFFu >> (32 - (y0 & 31))) : 0u;
                if (x && upd == kFarRow) upd = y0 - (b * 32 + cq_highbit(x)) - 1;  // emit() adds the 1
            }
            const int n = cls_up(y0 + 1, lastknown, ecls);
            nb = n >= 0 ? n : kFarRow;
        }
        // hull cursor: nothing taken yet, scanning starts L rows above y0
        cur_p = -1; cur_h = kNoCurH; nxt_p = -1; nxt_h = kNoNxtH;
        {
            const int s = cq_max(y0 - L, 0);
            eblk = s >> 5;
            ehw = c.h_ld(eblk) & (0xFFFFFFFFu << (s & 31));
        }
        set_ev(y0);
        // quiet window of block B: blocks B-reach .. B+reach
        n0 = n1 = n2 = 0;
        for (int b = B0 - reach; b <= B0 + reach; ++b) q_add(b, 1);
        for (int B = B0; B * 32 < y1; ++B) {
            if (B > B0) {
                cls_push(B + reach);
                q_add(B + reach, 1);
                q_add(B - reach - 1, -1);
            }
            const int r0 = cq_max(B * 32, y0), r1 = cq_min(B * 32 + 32, y1);
            if (r1 - r0 == 32 && n2 == 0 && (n0 == 0 || n1 == 0)) {
                // no site and no class change within reach of any of the 32 rows, in any column of the group
                ++upd;
                if (CQ_UNLIKELY(r0 == nb)) { class_flip(r0); set_ev(r0 + 1); }  // only possible for L == 0
                upd += 31;
                c.out_run(ecls, C2, 32);
                continue;
            }
            for (int er = r0; er < r1; ++er) emit(er);
        }
    }
};

}  // namespace chaq
