// sweep_core.h -- the banded column pass (pass 2, fast path), written once for host and device.
//
// Replaces the second dist_transform_axis call (reference openmp/df.c:133 -> dist_transform_1d, df.c:29-97)
// for BOTH fields of openmp/sdfgen.c:277-289 in one sweep over a column, in exact integer arithmetic.
//
// Input per column and row r (from pass 1): the pixel's class c(r) and G(r) = distance along the row to the
// nearest pixel of the OTHER class, stored as g8 = min(G, C) - 1 (C = cap_eff of tables.h, g8 == C-1 means
// "none within reach"). A pixel of class c at row y needs
//     D2(y) = min over opposite-class pixels q of |p - q|^2.
// One lower envelope serves both classes:
//     D2(y) = min( V(y)^2 , min_r (y - r)^2 + G(r)^2 )             (clamped at C^2)
// where V(y) is the distance along the column to the nearest row of the other class and r runs over ALL rows
// regardless of class. Proof sketch: for r of y's class the term is the squared distance to a real opposite
// pixel (the one G(r) points at); for r of the other class the term is >= (y-r)^2 >= V(y)^2; and the true
// nearest opposite pixel is either in y's column (V) or is the row-nearest opposite pixel of a same-class
// row. So the class bit matters only for V and for the sign of the result, and the envelope is the plain
// Felzenszwalb-Huttenlocher one (df.c:57-79) over heights G^2, kept as a bitmask of "rows on the hull".
//
// One sweep: rows are pushed in blocks of 32; row e = t - L (L = C-1) is finished right after row t was
// pushed, when every site within L rows of e is known. Sites more than 2L rows behind the newest are
// forgotten. All state that is re-read (g8, hull bits, class bits) lives in rings of NB blocks addressed by
// slot = block mod NB; the device keeps them in shared memory.
//
// Domination test (site b between a and c never the minimum), all int32 (|terms| < 2^29 for C <= 256):
//   w*(h_b-h_a) - u*(h_c-h_b) >= u*w*(u+w),  u = b-a, w = c-b      (cross-multiplied form of df.c:67)
//
// The same source is compiled by g++ in tests/emul (checked against the oracle) and by nvcc.
#pragma once
#include <stdint.h>
#include "envelope_core.h"  // CQ_HD, cq_min/cq_max, cq_lowbit/cq_highbit, kBig

namespace chaq {

constexpr int kFarRow = 1 << 28;  // "no such row"; survives +- image-size arithmetic without overflow
constexpr int kNoCurH = 1 << 28;  // height of the absent current site: every value it yields is >= C^2
constexpr int kNoNxtH = 1 << 29;  // height of the absent successor: never better than any current site

#if defined(__CUDA_ARCH__)
#define CQ_UNLIKELY(x) __builtin_expect(!!(x), 0)
#else
#define CQ_UNLIKELY(x) __builtin_expect(!!(x), 0)
#endif

// Ctx must provide (slot = ring slot of a 32-row block, j = row within the block):
//   void     block_begin(int blk, int slot)   block `blk` is about to be pushed into `slot`: its g8 rows must be
//                                             readable from now on
//   uint32_t cls_fetch(int blk)               class bits of this column for the 32 rows of block blk
//                                             (rows past the end of the buffer repeat the last real row)
//   uint32_t g_push(int j)                    g8 of row j of the block being pushed
//   uint32_t g_at(int blk, int slot, int j)   g8 of row j of block blk (ring slot `slot`), blk within 2L rows
//   uint32_t hull_ld(int slot) / void hull_st(int slot, uint32_t)
//   uint32_t cls_ld(int slot)  / void cls_st(int slot, uint32_t)
//   void     out(int cls, int d2c)            next result (rows arrive in order from y0; d2c clamped to C^2)
//   void     out_run(int cls, int d2c, int n) the next n results are all (cls, d2c)
//   bool     block_quiet(int L)               no row of the block being pushed holds a g8 < L -- for EVERY column
//                                             that shares this control flow (the warp on the device), so that
//                                             skipping the block's pushes is a uniform decision
//   bool     cls_uniform(uint32_t w, int* b)  the class word w is all-0 or all-1, the same for every such column
template <class Ctx>
struct ColumnSweep {
    Ctx& c;
    int NB, L, C2;
    // push side
    int pblk, pslot;        // block being pushed and its ring slot
    uint32_t hwp, cwp;      // hull / class word of that block (registers until the block is complete)
    int lowb;               // sites above this row are forgotten (checked once per block): pblk*32 - 2L
    int lastknown;          // last row whose class bit is known (end of the push block)
    int top_p, top_h, sec_p, sec_h;  // the two newest hull sites (-1: none)
    // emission side
    bool einit;
    int ecls, upd, nb;      // class of the run being emitted, rows since it began, first row of the next run
    int cur_p, cur_h;       // best site so far (none: -1, kNoCurH)
    int nxt_p, nxt_h;       // its successor on the hull (none: -1, kNoNxtH)
    // successor cursor: the hull sites after nxt are the set bits of (hull word of block eblk) & emask, then of
    // the blocks after eblk; no block between nxt's and eblk holds a hull bit. ehw caches that word while
    // eblk < pblk (for eblk == pblk the live register hwp is used).
    int eblk, eslot;
    uint32_t ehw, emask;
    // quiet tracking: the last qrun pushed blocks (ending with pblk) held no site in any column of the group and
    // one class (qruncls) throughout
    int qrun, qruncls, qneed;

    CQ_HD ColumnSweep(Ctx& ctx, int nb_, int lag, int c2) : c(ctx), NB(nb_), L(lag), C2(c2) {}

    CQ_HD int slot_of(int blk) const {
        int s = pslot - (pblk - blk);
        return s < 0 ? s + NB : s;
    }
    CQ_HD uint32_t hull_word(int blk) { return blk == pblk ? hwp : c.hull_ld(slot_of(blk)); }
    CQ_HD uint32_t cls_word(int blk) { return blk == pblk ? cwp : c.cls_ld(slot_of(blk)); }
    CQ_HD int height_at(int rr) {
        int g = (int)c.g_at(rr >> 5, slot_of(rr >> 5), rr & 31) + 1;
        return g * g;
    }
    // lowest set hull bit at a row in [from, end of the push block], or -1
    CQ_HD int hull_up(int from) {
        int b = from >> 5;
        uint32_t bits = hull_word(b) & (0xFFFFFFFFu << (from & 31));
        for (;;) {
            if (bits) return (b << 5) + cq_lowbit(bits);
            if (b == pblk) return -1;
            ++b;
            bits = hull_word(b);
        }
    }
    // highest set hull bit at a row in [lo, from], or -1
    CQ_HD int hull_down(int from, int lo) {
        if (from < lo) return -1;
        int b = from >> 5, bend = lo >> 5;
        uint32_t bits = hull_word(b) & (0xFFFFFFFFu >> (31 - (from & 31)));
        for (;;) {
            if (b == bend) bits &= (0xFFFFFFFFu << (lo & 31));
            if (bits) return (b << 5) + cq_highbit(bits);
            if (b == bend) return -1;
            --b;
            bits = hull_word(b);
        }
    }
    // first row in [from, to] whose class differs from cls, or -1
    CQ_HD int cls_up(int from, int to, int cls) {
        if (from > to) return -1;
        const uint32_t flip = cls ? 0xFFFFFFFFu : 0u;
        int b = from >> 5, bend = to >> 5;
        uint32_t bits = (cls_word(b) ^ flip) & (0xFFFFFFFFu << (from & 31));
        for (;;) {
            if (b == bend) bits &= (0xFFFFFFFFu >> (31 - (to & 31)));
            if (bits) return (b << 5) + cq_lowbit(bits);
            if (b == bend) return -1;
            ++b;
            bits = cls_word(b) ^ flip;
        }
    }
    // last row in [lo, from] whose class differs from cls, or -1
    CQ_HD int cls_down(int from, int lo, int cls) {
        if (from < lo) return -1;
        const uint32_t flip = cls ? 0xFFFFFFFFu : 0u;
        int b = from >> 5, bend = lo >> 5;
        uint32_t bits = (cls_word(b) ^ flip) & (0xFFFFFFFFu >> (31 - (from & 31)));
        for (;;) {
            if (b == bend) bits &= (0xFFFFFFFFu << (lo & 31));
            if (bits) return (b << 5) + cq_highbit(bits);
            if (b == bend) return -1;
            --b;
            bits = cls_word(b) ^ flip;
        }
    }
    CQ_HD void hull_clear(int rr) {
        const uint32_t bit = 1u << (rr & 31);
        const int b = rr >> 5;
        if (b == pblk) hwp &= ~bit;
        else {
            int s = slot_of(b);
            c.hull_st(s, c.hull_ld(s) & ~bit);
            if (b == eblk) ehw &= ~bit;
        }
    }

    CQ_HD void reset() {
        pblk = 0; pslot = 0; hwp = 0u; cwp = 0u; lowb = -2 * L; lastknown = -1;
        top_p = sec_p = -1; top_h = sec_h = 0;
        einit = false; ecls = 0; upd = kFarRow; nb = kFarRow;
        cur_p = -1; cur_h = kNoCurH; nxt_p = -1; nxt_h = kNoNxtH;
        eblk = 0; eslot = 0; ehw = 0u; emask = 0xFFFFFFFFu;
        qrun = 0; qruncls = 0; qneed = (2 * L + 31) / 32 + 1;
    }

    // after begin_block: is the block free of sites for the whole group, and does it extend a quiet run?
    CQ_HD bool note_quiet() {
        const bool pq = c.block_quiet(L);
        int bit = 0;
        if (pq && c.cls_uniform(cwp, &bit)) {
            qrun = (qrun > 0 && bit == qruncls) ? qrun + 1 : 1;
            qruncls = bit;
        } else {
            qrun = 0;
        }
        return pq;
    }
    // Rows [base-L, base-L+32) when the blocks covering [base-2L, base+31] (or the buffer top) are one quiet run:
    // no site and no class change within reach of any of them, in any column of the group -> D2 = C^2.
    CQ_HD bool emit_quiet_ok(int blk) const { return einit && qrun >= cq_min(qneed, blk + 1); }
    CQ_HD void emit_quiet_block(int e0) {
        ++upd;
        if (CQ_UNLIKELY(e0 == nb)) class_flip(e0);  // only possible for L == 0 (row e0-1 is within reach otherwise)
        upd += 31;
        c.out_run(ecls, C2, 32);
    }

    // a new block of 32 rows starts: retire the previous one to the rings, fetch the class bits of the new one
    CQ_HD void begin_block(int blk, bool first) {
        if (!first) {
            c.hull_st(pslot, hwp);
            c.cls_st(pslot, cwp);
            if (eblk == pblk) ehw = hwp;  // the cursor's block is no longer live: cache its word
            pslot = (pslot + 1 == NB) ? 0 : pslot + 1;
        }
        pblk = blk;
        hwp = 0u;
        c.block_begin(blk, pslot);
        cwp = c.cls_fetch(blk);
        lastknown = blk * 32 + 31;
        // sites more than 2L rows above this block are out of the rings and out of reach of every row still to
        // be finished: forget them
        lowb = blk * 32 - 2 * L;
        if (top_p < lowb) { top_p = -1; sec_p = -1; }
        else if (sec_p < lowb) sec_p = -1;
        if (einit && nb == kFarRow) {  // the run being emitted was open-ended: does it end in this block?
            const uint32_t x = cwp ^ (ecls ? 0xFFFFFFFFu : 0u);
            if (x) nb = blk * 32 + cq_lowbit(x);
        }
    }

    // the top site is dominated by (sec, new site at rr with height h): pop until it is not. Returns the lowest
    // popped row.
    CQ_HD int pop_sites(int rr, int h) {
        int popfloor;
        for (;;) {
            hull_clear(top_p);
            popfloor = top_p;
            top_p = sec_p; top_h = sec_h;
            sec_p = hull_down(top_p - 1, cq_max(lowb, 0));
            if (sec_p < 0) break;
            sec_h = height_at(sec_p);
            const int u = top_p - sec_p, w = rr - top_p;
            if (w * (top_h - sec_h) - u * (h - top_h) < u * w * (u + w)) break;
        }
        return popfloor;
    }

    CQ_HD void push(int rr, int j) {
        const int g8 = (int)c.g_push(j);
        if (g8 < L) {  // else: no opposite pixel within reach along this row
            const int h = (g8 + 1) * (g8 + 1);
            int popfloor = kFarRow;
            if (sec_p >= 0) {
                const int u = top_p - sec_p, w = rr - top_p;
                if (CQ_UNLIKELY(w * (top_h - sec_h) - u * (h - top_h) >= u * w * (u + w))) popfloor = pop_sites(rr, h);
            }
            hwp |= 1u << j;
            sec_p = top_p; sec_h = top_h;
            top_p = rr; top_h = h;
            // emission side: every site from popfloor up to rr-1 has just left the hull (unsigned: -1 = none)
            if ((unsigned)popfloor <= (unsigned)cur_p) {
                cur_p = rr; cur_h = h; nxt_p = -1; nxt_h = kNoNxtH;
                eblk = pblk; eslot = pslot; emask = 0xFFFFFFFEu << j;
            } else if ((unsigned)popfloor <= (unsigned)nxt_p) {
                nxt_p = rr; nxt_h = h;
                eblk = pblk; eslot = pslot; emask = 0xFFFFFFFEu << j;
            }
        }
    }

    // nxt := the hull site after nxt
    CQ_HD void next_site() {
        for (;;) {
            const uint32_t wd = (eblk == pblk ? hwp : ehw) & emask;
            if (wd) {
                const int j = cq_lowbit(wd);
                emask = 0xFFFFFFFEu << j;
                nxt_p = eblk * 32 + j;
                const int g = (int)c.g_at(eblk, eslot, j) + 1;
                nxt_h = g * g;
                return;
            }
            if (eblk == pblk) { nxt_p = -1; nxt_h = kNoNxtH; return; }  // later pushes into this block stay visible
            ++eblk;
            eslot = (eslot + 1 == NB) ? 0 : eslot + 1;
            emask = 0xFFFFFFFFu;
            if (eblk != pblk) ehw = c.hull_ld(eslot);
        }
    }

    CQ_HD void seek(int from) {  // cur := first hull site at or after `from`, nxt := its successor
        const int p = hull_up(from);
        cur_p = p;
        if (p >= 0) {
            cur_h = height_at(p);
            eblk = p >> 5; eslot = slot_of(eblk);
            emask = 0xFFFFFFFEu << (p & 31);
            if (eblk != pblk) ehw = c.hull_ld(eslot);
            next_site();
        } else {
            cur_h = kNoCurH;
            nxt_p = -1; nxt_h = kNoNxtH;
            eblk = pblk; eslot = pslot; emask = 0xFFFFFFFFu;
        }
    }

    CQ_HD void init_emit(int er) {
        einit = true;
        ecls = (int)(cls_word(er >> 5) >> (er & 31)) & 1;
        const int a = cls_down(er - 1, 0, ecls);
        upd = a >= 0 ? er - a - 1 : kFarRow;  // emit() adds the 1
        const int n = cls_up(er + 1, lastknown, ecls);
        nb = n >= 0 ? n : kFarRow;
    }

    CQ_HD void class_flip(int er) {
        ecls ^= 1;
        upd = 1;
        const int n = cls_up(er + 1, lastknown, ecls);
        nb = n >= 0 ? n : kFarRow;
    }

    CQ_HD int advance(int er, int fn) {
        int fc;
        do {
            cur_p = nxt_p; cur_h = nxt_h; fc = fn;
            next_site();
            const int dn = er - nxt_p;
            fn = dn * dn + nxt_h;
        } while (fn <= fc);
        return fc;
    }

    CQ_HD void emit(int er) {
        ++upd;
        if (CQ_UNLIKELY(er == nb)) class_flip(er);
        // A site more than L rows above er no longer matters, and what follows it in the bitmask may have
        // survived only because its own predecessor had been forgotten: restart at the first site in reach.
        const int lowrow = cq_max(er - L, 0);
        if (CQ_UNLIKELY((unsigned)cur_p < (unsigned)lowrow)) seek(lowrow);
        const int d = er - cur_p;
        int fc = d * d + cur_h;
        const int dn = er - nxt_p;
        const int fn = dn * dn + nxt_h;
        if (fn <= fc) fc = advance(er, fn);
        const int v = cq_min(cq_min(upd, nb - er), L + 1);  // (L+1)^2 = C^2: the clamp comes for free
        c.out(ecls, cq_min(fc, v * v));
    }

    template <bool PUSH, bool EMIT>
    CQ_HD void rows(int base, int j0, int j1) {
        for (int j = j0; j < j1; ++j) {
            if (PUSH) push(base + j, j);
            if (EMIT) emit(base + j - L);
        }
    }

    // relative rows [0, nrows) exist; results for rows [y0, y1). Row 0 of the piece is a multiple of 32 rows
    // into the buffer and at least L rows above y0 unless it is the top of the buffer.
    CQ_HD void run(int nrows, int y0, int y1) {
        reset();
        const int tend = cq_min(nrows, y1 + L);  // rows [0, tend) are pushed
        const int stop = y1 + L;                 // row er = rr - L is emitted for rr in [y0 + L, stop)
        for (int base = 0; base < stop; base += 32) {
            bool pq = false;  // nothing to push in this block (for the whole group)
            if (base < tend) {
                begin_block(base >> 5, base == 0);
                pq = note_quiet();
            } else if (qrun > 0) {
                ++qrun;  // rows past the end of the buffer: no sites, no class change
            }
            const int jstop = cq_min(32, stop - base);
            const int jpush = cq_max(0, cq_min(jstop, tend - base));
            const int jemit = cq_max(0, cq_min(jstop, y0 + L - base));
            if (jemit == 0 && jstop == 32 && (jpush == 32 || jpush == 0) && emit_quiet_ok(base >> 5)) {
                emit_quiet_block(base - L);
                continue;
            }
            const int lo = cq_min(jpush, jemit), hi = cq_max(jpush, jemit);
            if (pq) {  // same as pushing 32 rows that hold no site
                if (!einit && jemit < jstop) init_emit(y0);
                rows<false, true>(base, jemit, jstop);
                continue;
            }
            rows<true, false>(base, 0, lo);
            if (!einit && jemit < jstop) init_emit(y0);
            if (jemit < jpush) rows<true, true>(base, lo, hi);
            rows<false, true>(base, hi, jstop);
        }
    }
};

// ring blocks needed: rows [t-2L-31, t] stay addressable while the next block (32 rows) is being prefetched
CQ_HD int sweep_ring_blocks(int L) { return (2 * L + 64 + 31) / 32; }
// when g8 is not kept in the ring (re-read from global memory), the hull/class rings only span the blocks
// that can still hold a site, floor((32*pblk - 2L)/32) .. pblk
CQ_HD int sweep_ring_blocks_g(int L) { return (2 * L + 31) / 32 + 1; }

}  // namespace chaq
