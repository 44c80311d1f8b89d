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
template <class Ctx>
struct ColumnSweep {
    Ctx& c;
    int NB, L, C2;
    // push side
    int pblk, pslot;        // block being pushed and its ring slot
    uint32_t hwp, cwp;      // hull / class word of that block (registers until the block is complete)
    int tlast;              // newest pushed row (relative), -1 before the first
    int lastknown;          // last row whose class bit is known (end of the push block)
    int top_p, top_h, sec_p, sec_h;
    // emission side
    bool einit;
    int ecls, upd, nb;      // class of the run being emitted, rows since it began (+1), first row of the next run
    int cur_p, cur_h, nxt_p, nxt_h;  // best site and its successor on the hull (-1: none)
    // successor cursor: every hull bit after nxt_p up to the end of block eblk is in ehw, and no block between
    // nxt's and eblk holds a hull bit. nxt_p < 0 (cur has no successor yet): eblk == pblk at that time, ehw == 0.
    int eblk, eslot;
    uint32_t ehw;

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
    // lowest set hull bit at a row in [from, to], or -1
    CQ_HD int hull_up(int from, int to) {
        if (from > to) return -1;
        int b = from >> 5, bend = to >> 5;
        uint32_t bits = hull_word(b) & (0xFFFFFFFFu << (from & 31));
        for (;;) {
            if (b == bend) bits &= (0xFFFFFFFFu >> (31 - (to & 31)));
            if (bits) return (b << 5) + cq_lowbit(bits);
            if (b == bend) return -1;
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
        }
        if (b == eblk) ehw &= ~bit;
    }

    CQ_HD void reset() {
        pblk = 0; pslot = 0; hwp = 0u; cwp = 0u; tlast = -1; lastknown = -1;
        top_p = sec_p = -1; top_h = sec_h = 0;
        einit = false; ecls = 0; upd = kFarRow; nb = kFarRow;
        cur_p = nxt_p = -1; cur_h = nxt_h = 0;
        eblk = 0; eslot = 0; ehw = 0u;
    }

    // a new block of 32 rows starts: retire the previous one to the rings, fetch the class bits of the new one
    CQ_HD void begin_block(int blk, bool first) {
        if (!first) {
            c.hull_st(pslot, hwp);
            c.cls_st(pslot, cwp);
            pslot = (pslot + 1 == NB) ? 0 : pslot + 1;
        }
        pblk = blk;
        hwp = 0u;
        c.block_begin(blk, pslot);
        cwp = c.cls_fetch(blk);
        lastknown = blk * 32 + 31;
        if (einit && nb == kFarRow) {  // the run being emitted was open-ended: does it end in this block?
            const uint32_t x = cwp ^ (ecls ? 0xFFFFFFFFu : 0u);
            if (x) nb = blk * 32 + cq_lowbit(x);
        }
    }

    CQ_HD void push(int rr, int j) {
        tlast = rr;
        const int g8 = (int)c.g_push(j);
        if (g8 >= L) return;  // no opposite pixel within reach along this row
        const int h = (g8 + 1) * (g8 + 1);
        const int lowb = rr - 2 * L;
        // sites more than 2L rows back are out of the rings and out of reach of every row still to be finished
        if (top_p < lowb) { top_p = -1; sec_p = -1; }
        else if (sec_p < lowb) sec_p = -1;
        int popfloor = kFarRow;
        while (sec_p >= 0) {
            const int u = top_p - sec_p, w = rr - top_p;
            if (w * (top_h - sec_h) - u * (h - top_h) < u * w * (u + w)) break;  // top stays on the hull
            hull_clear(top_p);
            popfloor = top_p;
            top_p = sec_p; top_h = sec_h;
            sec_p = hull_down(top_p - 1, cq_max(lowb, 0));
            sec_h = sec_p >= 0 ? height_at(sec_p) : 0;
        }
        const uint32_t bit = 1u << j;
        hwp |= bit;
        sec_p = top_p; sec_h = top_h;
        top_p = rr; top_h = h;
        // emission side: every site from popfloor up to rr-1 has just left the hull
        if (cur_p < 0 || popfloor <= cur_p) { cur_p = rr; cur_h = h; nxt_p = -1; eblk = pblk; eslot = pslot; ehw = 0u; }
        else if (nxt_p < 0 || popfloor <= nxt_p) { nxt_p = rr; nxt_h = h; eblk = pblk; eslot = pslot; ehw = 0u; }
        else if (eblk == pblk) ehw |= bit;
    }

    // nxt := the hull site after nxt (nxt_p >= 0 on entry)
    CQ_HD void next_site() {
        if (ehw == 0u) {
            int b = eblk;
            while (b < pblk) {
                ++b;
                const uint32_t wd = hull_word(b);
                if (wd) { ehw = wd; eblk = b; eslot = slot_of(b); break; }
            }
            if (ehw == 0u) { nxt_p = -1; eblk = pblk; eslot = pslot; return; }
        }
        const int j = cq_lowbit(ehw);
        ehw &= ehw - 1u;
        nxt_p = eblk * 32 + j;
        const int g = (int)c.g_at(eblk, eslot, j) + 1;
        nxt_h = g * g;
    }

    CQ_HD void seek(int from) {  // cur := first hull site at or after `from`, nxt := its successor
        const int p = hull_up(from, tlast);
        cur_p = p;
        if (p >= 0) {
            cur_h = height_at(p);
            eblk = p >> 5; eslot = slot_of(eblk);
            ehw = hull_word(eblk) & ~((2u << (p & 31)) - 1u);
            nxt_p = p;
            next_site();
        } else {
            nxt_p = -1; eblk = pblk; eslot = pslot; ehw = 0u;
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

    CQ_HD void emit(int er) {
        if (er == nb) {
            ecls ^= 1;
            upd = 1;
            const int n = cls_up(er + 1, lastknown, ecls);
            nb = n >= 0 ? n : kFarRow;
        } else {
            ++upd;
        }
        // A site more than L rows above er no longer matters, and what follows it in the bitmask may have
        // survived only because its own predecessor had been forgotten: restart at the first site in reach.
        if (cur_p >= 0 && cur_p < er - L) seek(er - L);
        int cand = C2;
        if (cur_p >= 0) {
            const int d = er - cur_p;
            int fc = d * d + cur_h;
            while (nxt_p >= 0) {
                const int dn = er - nxt_p;
                const int fn = dn * dn + nxt_h;
                if (fn > fc) break;
                cur_p = nxt_p; cur_h = nxt_h; fc = fn;
                next_site();
            }
            cand = cq_min(cand, fc);
        }
        const int v = cq_min(upd, nb - er);
        if (v <= L) cand = cq_min(cand, v * v);
        c.out(ecls, cand);
    }

    template <bool PUSH, bool EMIT>
    CQ_HD void rows(int base, int j0, int j1) {
#if defined(__CUDA_ARCH__)
#pragma unroll 2
#endif
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
            if (base < tend) begin_block(base >> 5, base == 0);
            const int jstop = cq_min(32, stop - base);
            const int jpush = cq_max(0, cq_min(jstop, tend - base));
            const int jemit = cq_max(0, cq_min(jstop, y0 + L - base));
            const int lo = cq_min(jpush, jemit), hi = cq_max(jpush, jemit);
            rows<true, false>(base, 0, lo);
            if (!einit && jemit < jstop) init_emit(y0);
            if (jemit < jpush) rows<true, true>(base, lo, hi);
            rows<false, true>(base, hi, jstop);
        }
    }
};

// ring blocks needed: rows [t-2L, t] stay addressable while the next block (32 rows) is being prefetched
CQ_HD int sweep_ring_blocks(int L) { return (2 * L + 64 + 31) / 32; }

}  // namespace chaq
