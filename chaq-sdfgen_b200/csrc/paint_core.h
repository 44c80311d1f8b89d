// paint_core.h -- evaluation of the finished column hull (pass 2c), written once for host and device.
//
// Same mathematics and the same inputs as HullEmit (hull_core.h): the hull bitmask left by build + merge, pass 1's
// record (g8 + class bits) and the block flags; same results, row for row. What differs is the shape of the work.
// HullEmit walks the hull row by row: every row compares the current site with its successor and, when the
// successor wins, looks up the next one -- in a warp (lane = column) that rare, long path runs on nearly every
// row because one of 32 lanes needs it. Here a block of 32 rows is done in two phases:
//
//   plan   (per lane, data dependent)  walk the hull sites that own rows of this block and note, per site, the FIRST
//          row it owns: one integer division per site (the row where two parabolas cross, openmp/df.c:9-15 in exact
//          arithmetic) instead of one comparison per row. Result: a 32-bit mask T (bit j: a new owner starts at row j
//          of the block) and the owners in order, in a small list.
//   paint  (all lanes in step, no data-dependent control flow but the rare class change) for each row: take the
//          next owner from the list if T says so, evaluate its parabola (df.c:89-93), the in-column distance V to the
//          nearest row of the other class, min, quantise, store.
//
// Owner of a row, exactly as HullEmit defines it: the walk position after (1) dropping the current site once it is
// more than L rows above the row and (2) moving on while the successor's value is <= the current one's. Values of
// two sites differ by a function that is linear in the row, so (2) happens at one row: the smallest y with
//   2*y*(n - c) >= (n^2 - c^2) + (h_n - h_c)          (c < n the two site rows, h their heights)
// and never un-happens. Sites that would start at or before the start of their predecessor own no row and are
// skipped. Rows relative to the block keep every product below 2^20.
#pragma once
#include <stdint.h>
#include "hull_core.h"

namespace chaq {

constexpr int kPaintNoSite = 4095;  // "height" g+1 of the absent owner: 4095^2 >= any C^2 (C <= 1025)
constexpr int kPaintRelBias = 2048; // list entries hold (row - block start) + bias in 16 bits: rows within ~L+64 of the block

// smallest integer q with q * den >= num (den > 0, |num| < 2^22, den < 2^12)
CQ_HD int cq_ceil_div(int num, int den) {
#if defined(__CUDA_ARCH__)
    int q = __float2int_ru(__fdividef((float)num, (float)den));
    const int t = q * den - num;          // want 0 <= t < den
    q += (t < 0) ? 1 : 0;
    q -= (t >= den) ? 1 : 0;
    return q;
#else
    int q = num / den;                    // truncates toward zero
    if (num > 0 && q * den != num) ++q;
    return q;
#endif
}

// Ctx provides (blk = 32-row block of the buffer; blocks past the buffer read as "no site", class of the last row):
//   uint32_t h_ld(int blk)      this column's finished hull word
//   uint32_t g_row(int r)       g8 of buffer row r
//   uint32_t cls_t(int blk)     class bits of this column for the block's rows (bit j = row 32*blk+j); the build
//                               kernel leaves them transposed like the hull words
//   int      q_flag(int blk)    0/1: no site in any column of the group and all class 0/1; 2: anything else; 3: outside
//   void     seg_st(int k, uint32_t v) / uint32_t seg_ld(int k)   owner list of the block being painted, k in [0, 32]
//   bool     any(bool)          true if it is for any column that shares control flow with this one (warp vote)
//   void     prefetch_block(int blk) / prefetch_g(int r)   hints: the words of block blk / g8 of row r are needed soon
//   void     out(int cls, int d2c) / void out_run(int cls, int d2c, int n)
template <class Ctx>
struct HullPaint {
    Ctx& c;
    int L, C2, reach;
    int ecls, upd, nb, lastknown;  // class run of the row being finished (as HullEmit)
    int cur_p, cur_g;              // owner so far: row, g8 + 1 (none: -1, kPaintNoSite)
    int nxt_p, nxt_g;              // its successor if already looked up (else -1)
    int eblk;                      // the sites after nxt (after cur if there is no nxt) are the bits of ehw in block
    uint32_t ehw;                  // eblk, then the blocks after eblk
    int n0, n1, n2;                // blocks of the quiet window with flag 0 / 1 / 2

    CQ_HD HullPaint(Ctx& ctx, int lag, int c2) : c(ctx), L(lag), C2(c2), reach((lag + 31) / 32) {}

    CQ_HD uint32_t cls_word(int blk) { return c.cls_t(blk); }
    CQ_HD int cls_up(int from, int to, int cls) {  // first row in [from, to] whose class differs from cls, or -1
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
    // block blk comes within reach: an open-ended class run may end there
    CQ_HD void cls_push(int blk) {
        lastknown = blk * 32 + 31;
        if (nb == kFarRow) {
            const uint32_t x = c.cls_t(blk) ^ (ecls ? 0xFFFFFFFFu : 0u);
            if (x) nb = blk * 32 + cq_lowbit(x);
        }
    }
    CQ_HD void q_add(int blk, int s) {
        const int f = c.q_flag(blk);
        if (f == 0) n0 += s; else if (f == 1) n1 += s; else if (f == 2) n2 += s;
    }

    // nxt := the next site of the hull if one is known in the word being scanned or in a block that starts at a
    // row <= limit
    CQ_HD void fetch_next(int limit) {
        for (;;) {
            if (ehw) {
                const int j = cq_lowbit(ehw);
                ehw &= ehw - 1u;
                nxt_p = eblk * 32 + j;
                if (ehw) c.prefetch_g(eblk * 32 + cq_lowbit(ehw));  // the site after this one
                nxt_g = (int)c.g_row(nxt_p) + 1;
                return;
            }
            if ((eblk + 1) * 32 > limit) { nxt_p = -1; return; }
            ++eblk;
            ehw = c.h_ld(eblk);
        }
    }
    // highest site in [lo, from], or -1
    CQ_HD int prev_site(int from, int lo) {
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
    // Where the walk stands when it reaches row y0. Walking down from y0 - L would visit every site in between (about
    // L/6 of them on shapes, each a dependent load); the hull is proper inside any window of 2L+1 rows, so the values
    // its sites take at row y0 fall and then rise along the hull: start at the site nearest above y0 and climb only
    // while the site before is strictly better. Sites below y0 are plan()'s business (it moves on while the successor
    // is at least as good).
    CQ_HD void start_cursor(int y0) {
        const int lo = cq_max(y0 - L, 0);
        cur_p = -1; cur_g = kPaintNoSite; nxt_p = -1; nxt_g = 0;
        int s = prev_site(y0, lo);
        if (s < 0) {  // nothing within reach above: the cursor stands before the first site below y0
            const int t = y0 + 1;
            eblk = t >> 5;
            ehw = c.h_ld(eblk) & (0xFFFFFFFFu << (t & 31));
            return;
        }
        int sg = (int)c.g_row(s) + 1;
        for (;;) {
            const int q = prev_site(s - 1, lo);
            if (q < 0) break;
            const int qg = (int)c.g_row(q) + 1;
            const int ds = y0 - s, dq = y0 - q;
            if (!(dq * dq + qg * qg < ds * ds + sg * sg)) break;
            s = q; sg = qg;
        }
        cur_p = s; cur_g = sg;
        eblk = s >> 5;
        ehw = (s & 31) == 31 ? 0u : (c.h_ld(eblk) & (0xFFFFFFFEu << (s & 31)));
    }
    static CQ_HD uint32_t pack(int rel, int g) { return ((uint32_t)(rel + kPaintRelBias) << 16) | (uint32_t)g; }

    // owners of rows [r0, r1) of the block that starts at row Y. Entry 0 of the list is the owner of r0.
    // live: some row of the block gets a value below C^2 from its owner.
    CQ_HD uint32_t plan(int Y, int r0, int r1, bool& live) {
        uint32_t T = 0u;
        int n = 0, seg_start = r0;
        const int limit = r1 - 1 + L;
        live = false;
        for (;;) {
            if (nxt_p < 0) fetch_next(limit);
            int s = kFarRow;
            if (cur_p >= 0) s = cur_p + L + 1;  // out of reach from this row on
            if (nxt_p >= 0) {
                if (cur_p < 0) {
                    s = seg_start;
                } else {
                    const int dn = nxt_p - cur_p;
                    const int num = dn * ((nxt_p - Y) + (cur_p - Y)) + (nxt_g * nxt_g - cur_g * cur_g);
                    s = cq_min(s, Y + cq_ceil_div(num, 2 * dn));
                }
            }
            const int e = cq_min(s, r1);  // cur owns rows [seg_start, e)
            if (e > seg_start && cur_p >= 0) {
                const int d = cq_max(cq_max(seg_start - cur_p, cur_p - (e - 1)), 0);
                live = live || d * d + cur_g * cur_g < C2;
            }
            if (s >= r1) break;
            if (s > seg_start) {
                ++n;
                T |= 1u << (s - Y);
                seg_start = s;
            }
            cur_p = nxt_p;
            cur_g = nxt_p < 0 ? kPaintNoSite : nxt_g;
            nxt_p = -1;
            c.seg_st(n, pack(cur_p < 0 ? 0 : cur_p - Y, cur_g));
        }
        return T;
    }

    CQ_HD void paint(int Y, int r0, int r1, uint32_t T) {
        int k = 0;
        uint32_t e = c.seg_ld(0);
        int cg = (int)(e & 0xFFFFu);
        int ch = cg * cg;
        int d = (r0 - Y) - ((int)(e >> 16) - kPaintRelBias);  // row - owner's row
        uint32_t t = T >> (r0 - Y);                           // bit 0: a new owner starts at this row
        int nbd = nb - r0;                                    // rows until the class changes
        const int lp1 = L + 1;
#if defined(__CUDA_ARCH__)
#pragma unroll 4
#endif
        for (int y = r0; y < r1; ++y) {
            if (t & 1u) {
                e = c.seg_ld(++k);
                cg = (int)(e & 0xFFFFu);
                ch = cg * cg;
                d = (y - Y) - ((int)(e >> 16) - kPaintRelBias);
            }
            ++upd;
            if (CQ_UNLIKELY(nbd == 0)) {
                class_flip(y);
                nbd = nb - y;
            }
            const int fc = d * d + ch;
            const int v = cq_min(cq_min(upd, nbd), lp1);
            c.out(ecls, cq_min(fc, v * v));
            ++d;
            --nbd;
            t >>= 1;
        }
    }

    // results for buffer rows [y0, y1)
    CQ_HD void run(int y0, int y1) {
        const int B0 = y0 >> 5;
        lastknown = (B0 + reach) * 32 + 31;
        ecls = (int)(cls_word(B0) >> (y0 & 31)) & 1;
        {
            const uint32_t flip = ecls ? 0xFFFFFFFFu : 0u;
            upd = kFarRow;
            const int blo = cq_max(0, (y0 - L - 1) >> 5);
            for (int b = B0; b >= blo && upd == kFarRow; --b) {
                uint32_t x = cls_word(b) ^ flip;
                if (b == B0) x &= (y0 & 31) ? (0xFFFFFFFFu >> (32 - (y0 & 31))) : 0u;
                if (x) upd = y0 - (b * 32 + cq_highbit(x)) - 1;  // paint() adds the 1
            }
            const int n = cls_up(y0 + 1, lastknown, ecls);
            nb = n >= 0 ? n : kFarRow;
        }
        start_cursor(y0);
        c.seg_st(0, pack(cur_p < 0 ? 0 : cur_p - B0 * 32, cur_g));
        n0 = n1 = n2 = 0;
        for (int b = B0 - reach; b <= B0 + reach; ++b) q_add(b, 1);
        for (int B = B0; B * 32 < y1; ++B) {
            c.prefetch_block(B + reach + 2);
            if (B > B0) {
                cls_push(B + reach);
                q_add(B + reach, 1);
                q_add(B - reach - 1, -1);
            }
            const int r0 = cq_max(B * 32, y0), r1 = cq_min(B * 32 + 32, y1);
            if (r1 - r0 == 32 && n2 == 0 && (n0 == 0 || n1 == 0)) {
                // no site and no class change within reach of any of the 32 rows, in any column of the group
                ++upd;
                if (CQ_UNLIKELY(r0 == nb)) class_flip(r0);  // only possible for L == 0
                upd += 31;
                c.out_run(ecls, C2, 32);
                // whatever the cursor holds is out of reach by the end of this block
                cur_p = -1; cur_g = kPaintNoSite;
                c.seg_st(0, pack(0, kPaintNoSite));
                continue;
            }
            bool live;
            const uint32_t T = plan(B * 32, r0, r1, live);
            // the in-column class distance matters if a class change lies within L rows of some row of the block
            live = live || upd + 1 <= L || nb - (r1 - 1) <= L;
            if (c.any(live)) {
                paint(B * 32, r0, r1, T);
            } else {
                upd += r1 - r0;
                c.out_run(ecls, C2, r1 - r0);
            }
            // the list's last entry is the owner carried into the next block: re-base it
            c.seg_st(0, pack(cur_p < 0 ? 0 : cur_p - (B + 1) * 32, cur_g));
        }
    }
};

}  // namespace chaq
