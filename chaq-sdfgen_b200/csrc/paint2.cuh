// paint2.cuh -- pass 2c of the banded path for spreads <= 256: evaluation of the finished hull, staged through shared
// memory. Replaces the fill loop of dist_transform_1d (reference openmp/df.c:83-96, incl. the sqrtf at df.c:95) for the
// inside and the outside field of openmp/sdfgen.c:277-289, and fuses transform_float_sub (sdfgen.c:98-106) and
// transform_float_to_byte (sdfgen.c:75-96) as an exact table look-up (tables.h).
//
// Same mathematics as paint_core.h (plan + paint, lane = column): per half block of 16 rows each lane lists the sites
// that own its rows (one exact integer division per site), then all lanes paint the rows in step. What differs is
// that nothing on the per-site or per-row path waits for global memory, and the per-row work is branch free:
//   * persistent kernel: kP2Warps (19) warps per CTA, one CTA per SM; each warp takes (column group, run of blocks) items
//     from a counter in the workspace (zeroed by the build kernel of the same stream);
//   * per warp, rings in shared memory filled by cp.async ahead of use: the hull words of the blocks B-reach-1 ..
//     B+reach+1 around the block B being painted (20 slots), the transposed class words of B .. B+reach+1 (12 slots),
//     and pass 1's row distances of the blocks B-1 .. B+2 (128 rows x 32 bytes). Sites outside that ring -- few, in
//     sparse areas -- are read from global memory;
//   * the block flags of the window are three ballot masks in registers;
//   * the byte for (class, clamped D2) comes from a table in shared memory. DIRECT: one byte per D2 (small spreads).
//     Otherwise bucketed: per 32 consecutive D2 the byte of the first and a 32-bit mask of the D2 at which the byte
//     steps by one (2 x 16 KB for C = 256): byte = b0 +- popc(mask & low bits);
//   * the row loop has no data-dependent branch: the owner list is indexed by a running count of the plan's "new owner
//     starts here" bits; a class change in any column of the group is a warp-uniform test of one bit per row.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "hull.cuh"
#include "paint_core.h"

namespace chaq {

#ifndef CQ_P2_WARPS
#define CQ_P2_WARPS 19
#endif
constexpr int kP2Warps = CQ_P2_WARPS;
#ifndef CQ_P2_HALF
#define CQ_P2_HALF 16
#endif
constexpr int kP2Half = CQ_P2_HALF;                      // rows per plan + paint step
constexpr int kP2ListWords = kP2Half + 1;        // owners of one half block per lane: one per row plus the one carried in
// Ring sizes in blocks. The hull ring must hold blocks B-reach-1 .. B+reach+1 (19 for reach 8), the class ring blocks
// B .. B+reach+1 (10). Powers of two (32 / 16) make the slot a mask; 20 / 12 slots save 2 KB per warp, which is three
// more warps per SM (the kernel's time falls with its warps), for a multiply-shift per slot computation.
#ifndef CQ_P2_HSLOTS
#define CQ_P2_HSLOTS 20
#endif
#ifndef CQ_P2_CSLOTS
#define CQ_P2_CSLOTS 12
#endif
constexpr int kP2HSlots = CQ_P2_HSLOTS, kP2CSlots = CQ_P2_CSLOTS;
constexpr int kP2HullOff = 0, kP2ClsOff = kP2HSlots * 128, kP2GOff = kP2ClsOff + kP2CSlots * 128, kP2ListOff = kP2GOff + 4096;
// slot of block b in a ring of N slots; b >= -64 (blocks above the image near its top edge)
template <int N>
__device__ __forceinline__ uint32_t p2_slot(int b) {
    if ((N & (N - 1)) == 0) return (uint32_t)b & (uint32_t)(N - 1);
    return (uint32_t)(b + 64 * N) % (uint32_t)N;
}
constexpr int kP2WarpBytes = kP2ListOff + kP2ListWords * 128;
constexpr int kP2RelBias = 1024;                 // list entry = (site row - block start + bias) << 17 | G^2
constexpr int kP2SinkOff = 128;                  // bytes after the work counters: where lanes past the image's width store

__host__ __device__ inline size_t hull_paint2_smem(int lut_bytes) {
    return (size_t)((lut_bytes + 127) & ~127) + (size_t)kP2Warps * kP2WarpBytes;
}

__device__ __forceinline__ uint32_t lds_u32(uint32_t a) {
    uint32_t v;
    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(a));
    return v;
}
__device__ __forceinline__ uint32_t lds_u8(uint32_t a) {
    uint32_t v;
    asm volatile("ld.shared.u8 %0, [%1];" : "=r"(v) : "r"(a));
    return v;
}
__device__ __forceinline__ uint2 lds_u64(uint32_t a) {
    uint2 v;
    asm volatile("ld.shared.v2.u32 {%0, %1}, [%2];" : "=r"(v.x), "=r"(v.y) : "r"(a));
    return v;
}
__device__ __forceinline__ void sts_u32(uint32_t a, uint32_t v) { asm volatile("st.shared.u32 [%0], %1;" ::"r"(a), "r"(v) : "memory"); }
__device__ __forceinline__ float rcp_approx(float x) {
    float r;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
    return r;
}
// smallest integer q with q * den >= num (den > 0, |num| < 2^22, den < 2^12): float estimate, integer correction
__device__ __forceinline__ int p2_ceil_div(int num, int den) {
    int q = __float2int_ru(__fmul_rn((float)num, rcp_approx((float)den)));
    const int t = q * den - num;  // want 0 <= t < den
    q += (t < 0) ? 1 : 0;
    q -= (t >= den) ? 1 : 0;
    return q;
}

// first row in [from, lastknown] whose class differs from cls, or kFarRow; lastknown is the last row of a block whose
// class word is in the ring (slot p2_slot<kP2CSlots>(block), stride 32 words, ring already offset by lane)
__device__ __noinline__ int p2_find_change(uint32_t cring_s, int cls, int from, int lastknown) {
    if (from > lastknown) return kFarRow;
    const uint32_t flip = cls ? 0xFFFFFFFFu : 0u;
    int b = from >> 5;
    const int bend = lastknown >> 5;
    uint32_t bits = (lds_u32(cring_s + p2_slot<kP2CSlots>(b) * 128) ^ flip) & (0xFFFFFFFFu << (from & 31));
    for (;;) {
        if (bits) return (b << 5) + (__ffs(bits) - 1);
        if (b == bend) return kFarRow;
        ++b;
        bits = lds_u32(cring_s + p2_slot<kP2CSlots>(b) * 128) ^ flip;
    }
}

// a row distance that is not in the ring (a site more than a block away from the block being painted)
__device__ __noinline__ int p2_g_global(const uint8_t* rec_grp, long long rs, int r, int lane) {
    return (int)__ldg(rec_grp + (long long)r * rs + lane);
}

template <bool DIRECT>
struct Paint2 {
    // ---- geometry of this lane's column
    const uint32_t* hcol;    // hull word of block 0 (stride wc per block)
    const uint32_t* clst;    // transposed class word of block 0
    const uint8_t* rec_grp;  // g8 of buffer row 0 at the group's first column
    const uint8_t* flag0;    // flag of block 0 of this column group (stride ncg per block)
    long long rs;
    int wc, ncg, nblk, nrows, lane;
    int L, C2, reach;
    // ---- shared memory of this warp (shared-space addresses, rings already offset by lane)
    uint32_t hring_s, cring_s, gring_s, seg_s;
    uint32_t lut_s;          // the table
    int stride;              // bytes per class
    // ---- output
    uint8_t* optr;
    long long ostride;
    bool wide_ok;
    // ---- walk state
    int curB;                // block whose neighbourhood the g ring holds
    uint32_t F0, F1, F2;     // bit k: flag of block fb + k is 0 / 1 / 2
    int fb;
    int ecls, up, nb;        // class of the current row; distance to the nearest other-class row above; next change row
    int cur_p, cur_h;        // owner so far (row, G^2); none: -1, C2
    int nxt_p, nxt_h;        // its successor if already fetched (else -1)
    uint32_t nzw;            // bit (b & 31): the hull word of block b is not empty (blocks of the window)
    int eblk;                // the sites after nxt (after cur) are the bits of ehw in block eblk, then the blocks after
    uint32_t ehw;

    // ------------------------------------------------------------------------------------------------ staging
    __device__ __forceinline__ void stage_hull(int b) {
        const uint32_t hd = hring_s + p2_slot<kP2HSlots>(b) * 128;
        if (b >= 0 && b < nblk) cp_async4(hd, hcol + (long long)b * wc);
        else sts_u32(hd, 0u);
    }
    __device__ __forceinline__ void stage_cls(int b) {
        const uint32_t cd = cring_s + p2_slot<kP2CSlots>(b) * 128;
        if (b >= 0 && b < nblk) cp_async4(cd, clst + (long long)b * wc);
        // past the buffer: the class of the last row (bit 31 of the last word; the build kernel replicates it)
        else sts_u32(cd, b < 0 ? 0u : (uint32_t)((int32_t)__ldg(clst + (long long)(nblk - 1) * wc) >> 31));
    }
    __device__ __forceinline__ void stage_g(int b) {
        const int k = b - fb;
        if (b >= 0 && b < nblk && k >= 0 && k < 32 && ((F2 >> k) & 1u)) {
#pragma unroll
            for (int i = 0; i < 2; ++i) {
                const int q = lane + 32 * i, row = q >> 1, piece = q & 1;
                const int r = b * 32 + row;
                if (r < nrows) cp_async16(gring_s + (r & 127) * 32 + piece * 16, rec_grp + (long long)r * rs + piece * 16);
            }
        }
    }
    __device__ __forceinline__ void load_flags(int first) {
        fb = first;
        const int b = first + lane;
        const int f = (b < 0 || b >= nblk) ? 3 : (int)__ldg(flag0 + (long long)b * ncg);
        F0 = __ballot_sync(0xffffffffu, f == 0);
        F1 = __ballot_sync(0xffffffffu, f == 1);
        F2 = __ballot_sync(0xffffffffu, f == 2);
    }
    __device__ __forceinline__ uint32_t hword(int b) const { return lds_u32(hring_s + p2_slot<kP2HSlots>(b) * 128); }
    __device__ __forceinline__ uint32_t cword(int b) const { return lds_u32(cring_s + p2_slot<kP2CSlots>(b) * 128); }
    __device__ __forceinline__ int height(int r) const {
        const int b = r >> 5;
        int g;
        if (CQ_UNLIKELY((unsigned)(b - curB + 1) >= 3u)) g = p2_g_global(rec_grp, rs, r, lane);
        else g = (int)lds_u8(gring_s + (r & 127) * 32 + lane);
        return (g + 1) * (g + 1);
    }

    // ------------------------------------------------------------------------------------------------ sites
    // highest site in [lo, from], or -1
    __device__ __forceinline__ int prev_site(int from, int lo) const {
        if (from < lo) return -1;
        int b = from >> 5;
        const int bend = lo >> 5;
        uint32_t bits = hword(b) & (0xFFFFFFFFu >> (31 - (from & 31)));
        for (;;) {
            if (b == bend) bits &= (0xFFFFFFFFu << (lo & 31));
            if (bits) return (b << 5) + (31 - __clz(bits));
            if (b == bend) return -1;
            --b;
            bits = hword(b);
        }
    }
    // Where the walk stands when it reaches row y0 (see paint_core.h start_cursor): the hull is proper inside any
    // window of 2L+1 rows, so the values its sites take at y0 fall and then rise along the hull: start at the site
    // nearest above y0 and climb only while the site before is strictly better.
    __device__ __forceinline__ void start_cursor(int y0) {
        const int lo = y0 - L > 0 ? y0 - L : 0;
        cur_p = -1; cur_h = C2; nxt_p = -1; nxt_h = 0;
        int s = prev_site(y0, lo);
        if (s < 0) {
            const int t = y0 + 1;
            eblk = t >> 5;
            ehw = hword(eblk) & (0xFFFFFFFFu << (t & 31));
            return;
        }
        int sh = height(s);
        for (;;) {
            const int q = prev_site(s - 1, lo);
            if (q < 0) break;
            const int qh = height(q);
            const int ds = y0 - s, dq = y0 - q;
            if (!(dq * dq + qh < ds * ds + sh)) break;
            s = q; sh = qh;
        }
        cur_p = s; cur_h = sh;
        eblk = s >> 5;
        ehw = (s & 31) == 31 ? 0u : (hword(eblk) & (0xFFFFFFFEu << (s & 31)));
    }
    // nxt := the next site of the hull if there is one in the word being scanned or in a block that starts at a row <= limit
    __device__ __forceinline__ void note_word(int b) {
        const uint32_t bit = 1u << (b & 31);
        nzw = hword(b) ? (nzw | bit) : (nzw & ~bit);
    }
    __device__ __forceinline__ void fetch_next(int limit) {
        if (ehw == 0u) {
            // the next block with a site, no further than the block of row `limit`
            const int n = (limit >> 5) - eblk;   // blocks eblk+1 .. eblk+n may be looked at
            if (n <= 0) { nxt_p = -1; return; }
            uint32_t m = __funnelshift_r(nzw, nzw, (eblk + 1) & 31);  // bit k: block eblk + 1 + k
            if (n < 32) m &= (1u << n) - 1u;
            if (m == 0u) { eblk += n; nxt_p = -1; return; }
            eblk += __ffs(m);
            ehw = hword(eblk);
        }
        const int j = __ffs(ehw) - 1;
        ehw &= ehw - 1u;
        nxt_p = eblk * 32 + j;
        nxt_h = height(nxt_p);
    }
    static __device__ __forceinline__ uint32_t pack(int rel, int h) { return ((uint32_t)(rel + kP2RelBias) << 17) | (uint32_t)h; }

    // owners of rows [r0, r1) of the block that starts at row Y: entry 0 of the list is the owner of r0, bit j of the
    // result says that the next entry takes over at row Y + j (paint_core.h plan())
    __device__ __forceinline__ uint32_t plan(int Y, int r0, int r1) {
        uint32_t T = 0u;
        uint32_t sp = seg_s;
        int seg_start = r0;
        const int limit = r1 - 1 + L;
        sts_u32(sp, pack(cur_p < 0 ? 0 : cur_p - Y, cur_h));
        for (;;) {
            if (nxt_p < 0) fetch_next(limit);
            int s;
            if (nxt_p >= 0) {
                if (cur_p >= 0) {
                    const int dn = nxt_p - cur_p;
                    const int num = dn * ((nxt_p - Y) + (cur_p - Y)) + (nxt_h - cur_h);
                    const int q = Y + p2_ceil_div(num, 2 * dn);
                    s = cur_p + L + 1;  // out of reach from this row on
                    s = s < q ? s : q;
                } else {
                    s = seg_start;
                }
            } else {
                s = cur_p >= 0 ? cur_p + L + 1 : kFarRow;
            }
            if (s >= r1) break;
            if (s > seg_start) {
                sp += 128;
                T |= 1u << (s - Y);
                seg_start = s;
            }
            uint32_t e = pack(nxt_p - Y, nxt_h);
            if (nxt_p < 0) e = pack(0, C2);
            cur_p = nxt_p;
            cur_h = nxt_p < 0 ? C2 : nxt_h;
            nxt_p = -1;
            sts_u32(sp, e);
        }
        return T;
    }

    // ------------------------------------------------------------------------------------------------ output
    __device__ __forceinline__ uint32_t byte_of(uint32_t tb, int sgn, int d2c) const {
        if (DIRECT) return lds_u8(tb + d2c);
        const uint2 e = lds_u64(tb + ((d2c >> 5) << 3));
        const int pc = __popc(e.x & (0xFFFFFFFFu >> ((~d2c) & 31)));
        return (uint32_t)(pc * sgn + (int)e.y);
    }
    // 32 rows of one saturated value per column: whole blocks go out as two 16-byte stores per lane
    __device__ __forceinline__ void out_block(uint32_t b) {
        if (wide_ok) {
            const uint32_t b01 = b | (__shfl_down_sync(0xffffffffu, b, 1) << 8);
            const uint32_t wq = b01 | (__shfl_down_sync(0xffffffffu, b01, 2) << 16);  // valid in lanes 4q
            const int h4 = (lane & 1) * 16;
            uint4 v;
            v.x = __shfl_sync(0xffffffffu, wq, h4);
            v.y = __shfl_sync(0xffffffffu, wq, h4 + 4);
            v.z = __shfl_sync(0xffffffffu, wq, h4 + 8);
            v.w = __shfl_sync(0xffffffffu, wq, h4 + 12);
            uint8_t* dst = optr - lane + (long long)(lane >> 1) * ostride + h4;
            *reinterpret_cast<uint4*>(dst) = v;
            *reinterpret_cast<uint4*>(dst + 16 * ostride) = v;
            optr += 32 * ostride;
            return;
        }
        for (int i = 0; i < 32; ++i) {
            *optr = (uint8_t)b;
            optr += ostride;
        }
    }

    // one row, all lanes in step. y = the row, jb = y - Y + bias; sp walks the owner list; bit U of t / xs / as belongs
    // to this row: a new owner starts here / this column's class changes here / some column's class changes here
    template <int U, bool FLIPS>
    __device__ __forceinline__ void row(int y, int jb, uint32_t& sp, uint32_t t, uint32_t xs, uint32_t as, int& nbd,
                                        uint32_t& tb, int& sgn, int blk_end, int lastknown) {
        sp += ((t >> U) & 1u) << 7;
        const uint32_t e = lds_u32(sp);
        const int h = (int)(e & 0x1FFFFu);
        const int d = (jb + U) - (int)(e >> 17);  // row - owner's row
        if (FLIPS && ((as >> U) & 1u)) {  // warp-uniform: the class changes at this row in some column of the group
            if ((xs >> U) & 1u) {
                ecls ^= 1;
                up = 1;
                sgn = -sgn;
                tb = lut_s + (uint32_t)(ecls * stride);
                const uint32_t rest = (xs >> U) >> 1;
                nbd = rest ? __ffs(rest) : p2_find_change(cring_s, ecls, blk_end, lastknown) - (y + U);
            }
        }
        const int fc = d * d + h;
        const int v = min(min(up, nbd), L + 1);
        const int d2 = min(fc, v * v);
        *optr = (uint8_t)byte_of(tb, sgn, d2);
        optr += ostride;
        ++up;
        --nbd;
    }
    // rows [r0, r1) of the block that starts at row Y (at most kP2Half rows, one plan's worth)
    __device__ __forceinline__ void paint(int Y, int r0, int r1, uint32_t T, uint32_t xw, uint32_t anyx, int lastknown) {
        uint32_t sp = seg_s;
        uint32_t t = T >> (r0 - Y), xs = xw >> (r0 - Y), as = anyx >> (r0 - Y);  // bit 0: row r0
        int nbd = nb - r0;
        uint32_t tb = lut_s + (uint32_t)(ecls * stride);
        int sgn = ecls ? 1 : -1;
        int y = r0, jb = (r0 - Y) + kP2RelBias;
#pragma unroll 1
        for (; y + 4 <= r1; y += 4, jb += 4, t >>= 4, xs >>= 4, as >>= 4) {
            if (as & 0xFu) {   // a class change in one of the four rows, in some column: rows one after the other
                row<0, true>(y, jb, sp, t, xs, as, nbd, tb, sgn, Y + 32, lastknown);
                row<1, true>(y, jb, sp, t, xs, as, nbd, tb, sgn, Y + 32, lastknown);
                row<2, true>(y, jb, sp, t, xs, as, nbd, tb, sgn, Y + 32, lastknown);
                row<3, true>(y, jb, sp, t, xs, as, nbd, tb, sgn, Y + 32, lastknown);
            } else {           // straight-line code: the four rows' loads overlap
                row<0, false>(y, jb, sp, t, xs, as, nbd, tb, sgn, Y + 32, lastknown);
                row<1, false>(y, jb, sp, t, xs, as, nbd, tb, sgn, Y + 32, lastknown);
                row<2, false>(y, jb, sp, t, xs, as, nbd, tb, sgn, Y + 32, lastknown);
                row<3, false>(y, jb, sp, t, xs, as, nbd, tb, sgn, Y + 32, lastknown);
            }
        }
#pragma unroll 1
        for (; y < r1; ++y, ++jb, t >>= 1, xs >>= 1, as >>= 1) row<0, true>(y, jb, sp, t, xs, as, nbd, tb, sgn, Y + 32, lastknown);
        nb = nbd + r1;
    }

    static __device__ __forceinline__ bool is_far(int row) { return row >= (kFarRow >> 1); }

    // results for buffer rows [y0, y1); the rings are this warp's, all lanes converged
    __device__ __forceinline__ void run(int y0, int y1) {
        const int B0 = y0 >> 5;
        __syncwarp();
        load_flags(B0 - reach);
        for (int b = B0 - reach - 1; b <= B0 + reach; ++b) stage_hull(b);
        for (int b = B0; b <= B0 + reach; ++b) stage_cls(b);
        stage_g(B0 - 1);
        stage_g(B0);
        stage_g(B0 + 1);
        cp_async_commit();
        // distance to the nearest row of the other class above y0 (within L + 1 rows): words straight from global memory,
        // all loads in flight at once
        uint32_t back[9];
        {
            const int blo = (y0 - L - 1) >> 5 > 0 ? (y0 - L - 1) >> 5 : 0;
#pragma unroll
            for (int i = 0; i < 9; ++i) {
                const int b = B0 - 1 - i;
                back[i] = b >= blo ? __ldg(clst + (long long)b * wc) : 0u;
            }
        }
        cp_async_wait0();
        __syncwarp();
        curB = B0;
        nzw = 0u;
        for (int b = B0 - reach - 1; b <= B0 + reach; ++b) note_word(b);
        const uint32_t cw0 = cword(B0);
        ecls = (int)(cw0 >> (y0 & 31)) & 1;
        {
            const uint32_t flip = ecls ? 0xFFFFFFFFu : 0u;
            up = kFarRow;
            uint32_t x = (cw0 ^ flip) & ((y0 & 31) ? (0xFFFFFFFFu >> (32 - (y0 & 31))) : 0u);
            if (x) up = y0 - (B0 * 32 + (31 - __clz(x)));
            const int blo = (y0 - L - 1) >> 5 > 0 ? (y0 - L - 1) >> 5 : 0;
#pragma unroll
            for (int i = 0; i < 9; ++i) {
                const int b = B0 - 1 - i;
                x = back[i] ^ flip;
                if (b >= blo && up == kFarRow && x) up = y0 - (b * 32 + (31 - __clz(x)));
            }
            nb = p2_find_change(cring_s, ecls, y0 + 1, (B0 + reach) * 32 + 31);
        }
        start_cursor(y0);
        const uint32_t wbits = (2u << (2 * reach)) - 1u;  // 2*reach+1 blocks (reach <= 8)
        for (int B = B0; B * 32 < y1; ++B) {
            __syncwarp();                    // every lane is done with block B - 2's rows (the slot block B + 2 takes)
            if (B + (reach > 2 ? reach : 2) - fb > 31) load_flags(B - reach);
            stage_hull(B + reach + 1);
            stage_cls(B + reach + 1);
            stage_g(B + 2);
            cp_async_commit();
            cp_async_wait1();                // everything but the newest group: words <= B + reach, rows <= block B + 1
            __syncwarp();
            curB = B;
            note_word(B + reach);
            const int lastknown = (B + reach) * 32 + 31;
            if (B > B0 && is_far(nb)) {      // block B + reach has come within reach: an open-ended class run may end there
                const uint32_t x = cword(B + reach) ^ (ecls ? 0xFFFFFFFFu : 0u);
                nb = x ? (B + reach) * 32 + (__ffs(x) - 1) : kFarRow;
            }
            const int Y = B * 32;
            const int r0 = Y > y0 ? Y : y0, r1 = Y + 32 < y1 ? Y + 32 : y1;
            const uint32_t wm = wbits << (B - reach - fb);
            if (r1 - r0 == 32 && !(F2 & wm) && !((F0 & wm) && (F1 & wm))) {
                // no site and no class change within reach of any of the 32 rows, in any column of the group
                if (CQ_UNLIKELY(r0 == nb)) {  // only possible for L == 0
                    ecls ^= 1;
                    up = 1;
                    nb = p2_find_change(cring_s, ecls, r0 + 1, lastknown);
                }
                up += 32;
                out_block(byte_of(lut_s + (uint32_t)(ecls * stride), ecls ? 1 : -1, C2) & 0xFFu);
                cur_p = -1; cur_h = C2; nxt_p = -1;
                if (eblk < B + reach) { eblk = B + reach; ehw = 0u; }
                continue;
            }
            // rows at which this column's class changes: bit j = row Y + j differs from the row before
            const uint32_t cw = cword(B);
            uint32_t xw = cw ^ ((cw << 1) | (uint32_t)ecls);
            if (r0 > Y) xw &= ~((2u << (r0 - Y)) - 1u);   // a run's first row has no row before it
            const uint32_t anyx = __reduce_or_sync(0xffffffffu, xw);
            for (int h0 = r0; h0 < r1;) {
                int h1 = (h0 & ~(kP2Half - 1)) + kP2Half;
                h1 = h1 < r1 ? h1 : r1;
                const uint32_t T = plan(Y, h0, h1);
                paint(Y, h0, h1, T, xw, anyx, lastknown);
                h0 = h1;
            }
        }
    }
};

template <bool DIRECT>
__global__ void __launch_bounds__(kP2Warps * 32, 1) hull_paint2_kernel(HullParams p) {
    extern __shared__ __align__(16) uint8_t s_hull[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int lut_bytes = 2 * p.lut_stride;
    {
        const uint4* src = reinterpret_cast<const uint4*>(p.lut);
        uint4* dst = reinterpret_cast<uint4*>(s_hull);
        for (int i = threadIdx.x; i < lut_bytes / 16; i += kP2Warps * 32) dst[i] = __ldg(src + i);
    }
    __syncthreads();
    const uint32_t mine = smem_u32(s_hull) + (uint32_t)((lut_bytes + 127) & ~127) + (uint32_t)warp * kP2WarpBytes;
    const int ncgp = p.cg1 - p.cg0;
    const long long n_items = (long long)p.n_img * ncgp * p.nruns_e;
    int lag = p.L;
    asm volatile("" : "+r"(lag));
    const int bfirst = p.out_begin >> 5;
    const int out_rows = p.out_end - p.out_begin;
    Paint2<DIRECT> c;
    c.hring_s = mine + kP2HullOff + lane * 4;
    c.cring_s = mine + kP2ClsOff + lane * 4;
    c.gring_s = mine + kP2GOff;
    c.seg_s = mine + kP2ListOff + lane * 4;
    c.rs = p.rs; c.wc = p.wc; c.ncg = p.ncg; c.nblk = p.nblk; c.nrows = p.buf_rows; c.lane = lane;
    c.L = lag; c.C2 = p.c2; c.reach = (lag + 31) / 32;
    c.lut_s = smem_u32(s_hull); c.stride = p.lut_stride;
    uint8_t* const sink = reinterpret_cast<uint8_t*>(p.work) + kP2SinkOff + lane;
    for (;;) {
        unsigned it = 0;
        if (lane == 0) it = atomicAdd(p.work, 1u);
        it = __shfl_sync(0xffffffffu, it, 0);
        if ((long long)it >= n_items) break;
        const int cg = p.cg0 + (int)(it % (unsigned)ncgp);
        const unsigned rest = it / (unsigned)ncgp;
        const int run = (int)(rest % (unsigned)p.nruns_e), img = (int)(rest / (unsigned)p.nruns_e);
        int y0 = (bfirst + run * p.eb) * 32, y1 = y0 + p.eb * 32;
        y0 = y0 > p.out_begin ? y0 : p.out_begin;
        y1 = y1 < p.out_end ? y1 : p.out_end;
        if (y0 >= y1) continue;
        const int x = cg * 32 + lane;
        c.hcol = p.hull + (long long)img * p.nblk * p.wc + x;
        c.clst = p.clsT + (long long)img * p.nblk * p.wc + x;
        c.rec_grp = p.inter + (long long)img * p.buf_rows * p.rs + cg * 32;
        c.flag0 = p.flags + (long long)img * p.nblk * p.ncg + cg;
        // lanes past the image's width compute like the others and store into a sink
        c.ostride = x < p.w ? p.w : 0;
        c.optr = x < p.w ? p.out + ((long long)img * out_rows + (y0 - p.out_begin)) * (long long)p.w + x : sink;
        c.wide_ok = cg * 32 + 32 <= p.w && (p.w & 15) == 0 && (reinterpret_cast<uintptr_t>(p.out) & 15) == 0;
        c.run(y0, y1);
    }
}

}  // namespace chaq
