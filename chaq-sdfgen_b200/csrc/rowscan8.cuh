// rowscan8.cuh -- pass 1, banded fast path (spread <= 256): threshold + per-row distance to the nearest
// pixel of the other class, 8 bits per pixel plus a class bit-plane.
//
// Replaces, fused into one kernel and for both fields at once:
//   transform_img_to_bool    reference openmp/sdfgen.c:52-62   (px > 127 / px < 127 on a strided byte image)
//   transform_bool_to_float  reference openmp/sdfgen.c:65-72   ({0,INF} seeds for each field)
//   dist_transform_axis #1   reference openmp/df.c:130 -> df.c:29-97 (for {0,INF} input the lower envelope
//                            along a row is (x - nearest seed)^2, i.e. a run-length scan)
// Output record per row (see sweep.cuh): Wp bytes g8 = min(G, C) - 1, then Wp/32 words of class bits
// (bit = 1: inside, the reference's `true`), then one "busy" byte per 128-column strip (1: some pixel of the
// strip has an opposite-class pixel within reach along this row) -- a hint for pass 2's work split only. G = distance along the row to the nearest pixel of the OTHER
// class. The distance is class-agnostic: it is the distance to the nearest class change, so one pair of
// ramps serves both fields.
//
// Mapping: one warp per row, 8 rows per CTA. Phase A turns the row into a bit string in shared memory
// (16 pixels per lane per 128-bit load on the fast path, ballot on the generic path). Phase B gives each lane
// one 32-pixel word: the nearest change outside the word comes from warp ballots plus a carry from earlier
// 1024-pixel chunks and a one-chunk look-ahead; inside the word two 32-step ramps finish the job in registers,
// and the 32 result bytes leave as two 128-bit stores.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "rowscan.cuh"  // class_nibble, kFar

namespace chaq {

#ifndef CQ_ROW8_WARPS
#define CQ_ROW8_WARPS 8
#endif
#ifndef CQ_ROW8_MINB
#define CQ_ROW8_MINB 4
#endif
constexpr int kRow8Warps = CQ_ROW8_WARPS;

struct RowScan8Params {
    const uint8_t* img;
    uint8_t* inter;      // records
    long long rows;
    long long rs;        // record stride
    int wp;              // padded width
    int w;
    int stride, offset;
    int above;
    int cap;             // C_row = min(spread+1, 256): g8 = min(G, cap) - 1
    int nwords;          // ceil(w/32)
    int nstrips;         // 128-column strips = wp/128; busy bytes live at record offset wp + wp/8
};

// Ramp tables for words WITH class changes inside, in shared memory (2 x 4 KB per CTA). For a byte m of the change
// mask, entry m holds eight 16-bit values, one per pixel i of the byte:
//   left  table: i - j, j the highest set bit of m at or below i   (distance - 1 to the change on the left inside the byte)
//   right table: j - i, j the lowest set bit of m at or above i
// and 0xFFFF where the byte has no such bit. A pixel's ramp is then min(table, incoming + offset), which is ONE DPX
// instruction for two pixels (VIADDMNMX.U16x2), instead of the bit-by-bit ramp's ~9 instructions per pixel.
struct Row8Tables {
    uint4 tl[256], tr[256];
};
__device__ __forceinline__ void row8_build_tables(Row8Tables* t, int tid, int nthreads) {
    for (int m = tid; m < 256; m += nthreads) {
        uint32_t l[8], r[8];
        int last = -1;
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            if ((m >> i) & 1) last = i;
            l[i] = last >= 0 ? (uint32_t)(i - last) : 0xFFFFu;
        }
        int next = -1;
#pragma unroll
        for (int i = 7; i >= 0; --i) {
            if ((m >> i) & 1) next = i;
            r[i] = next >= 0 ? (uint32_t)(next - i) : 0xFFFFu;
        }
        t->tl[m] = make_uint4(l[0] | (l[1] << 16), l[2] | (l[3] << 16), l[4] | (l[5] << 16), l[6] | (l[7] << 16));
        t->tr[m] = make_uint4(r[0] | (r[1] << 16), r[2] | (r[3] << 16), r[4] | (r[5] << 16), r[6] | (r[7] << 16));
    }
}

// One lane's 32-pixel word: the two ramps (distance to the nearest class change on the left / on the right), their
// minimum clamped at cap, minus one, as 32 bytes in two 128-bit stores; then the class word.
//   PL: last pixel left of the word with a class change immediately to its left (or -kFar)
//   PR: first pixel right of the word with a class change immediately to its right (or >= kFar)
//   flat: no word of the warp has a class change inside it (warp-uniform on purpose: a per-lane choice would make
//         a mixed warp run both paths). Both ramps are then straight lines, and two pixels go per DPX instruction
//         (VIADDMNMX.S16x2: min(a + b, c) on 16-bit halves), about six times fewer instructions than the bit-by-bit ramps.
__device__ __forceinline__ void row8_store(uint8_t* rec, uint32_t* crec, int wi, int x0, int nvalid, int PL, int PR,
                                           uint32_t Lw, uint32_t R, uint32_t v, int cap, bool flat, const Row8Tables* tab) {
    uint32_t packed[8];
    if (flat) {
        // stored byte of pixel b = min(z0 + b, y31 + 31 - b, cap) - 1; both starts are clamped at cap first, so that
        // every halfword stays small (<= cap + 31)
        const uint32_t zs = (uint32_t)(min(x0 - PL + 1, cap) - 1);          // pixel 0, left ramp, minus one
        const uint32_t ys = (uint32_t)(min(PR - (x0 + 31) + 1, cap) - 1);   // pixel 31, right ramp, minus one
        const uint32_t z2 = zs | ((zs + 1u) << 16);                         // pixels (0, 1)
        const uint32_t y2 = (ys + 31u) | ((ys + 30u) << 16);
        const uint32_t c2 = (uint32_t)(cap - 1) * 0x00010001u;
#pragma unroll
        for (int i = 0; i < 8; ++i) {  // four pixels per step: halfword pairs (4i, 4i+1) and (4i+2, 4i+3)
            const uint32_t p0 = 4u * i, p1 = 4u * i + 2u;
            const uint32_t up0 = p0 * 0x00010001u, up1 = p1 * 0x00010001u;                       // +p in both halves
            const uint32_t dn0 = ((0u - p0) & 0xFFFFu) * 0x00010001u, dn1 = ((0u - p1) & 0xFFFFu) * 0x00010001u;  // -p
            const uint32_t a = __viaddmin_s16x2(y2, dn0, __viaddmin_s16x2(z2, up0, c2));
            const uint32_t b = __viaddmin_s16x2(y2, dn1, __viaddmin_s16x2(z2, up1, c2));
            packed[i] = __byte_perm(a, b, 0x6420);
        }
    } else {
        // all values are "distance - 1"; 16 bits per pixel, pairs (2q, 2q+1) per register
        uint32_t lr[16];
        uint32_t zin = (uint32_t)min(x0 - PL, 1023);                // pixel 0 of the word, no change inside so far
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const uint4 e = tab->tl[(Lw >> (8 * k)) & 0xFFu];
            const uint32_t z2 = zin * 0x00010001u;
            lr[4 * k + 0] = __viaddmin_u16x2(z2, 0x00010000u, e.x);  // min(incoming + i, table)
            lr[4 * k + 1] = __viaddmin_u16x2(z2, 0x00030002u, e.y);
            lr[4 * k + 2] = __viaddmin_u16x2(z2, 0x00050004u, e.z);
            lr[4 * k + 3] = __viaddmin_u16x2(z2, 0x00070006u, e.w);
            zin = (lr[4 * k + 3] >> 16) + 1u;                        // pixel 0 of the next byte
        }
        uint32_t yin = (uint32_t)min(PR - (x0 + 31), 1023);          // pixel 31 of the word
        const uint32_t c2 = (uint32_t)(cap - 1) * 0x00010001u;
#pragma unroll
        for (int k = 3; k >= 0; --k) {
            const uint4 e = tab->tr[(R >> (8 * k)) & 0xFFu];
            const uint32_t y2 = yin * 0x00010001u;
            const uint32_t r0 = __viaddmin_u16x2(y2, 0x00060007u, e.x);  // pixel i: incoming + (7 - i)
            const uint32_t r1 = __viaddmin_u16x2(y2, 0x00040005u, e.y);
            const uint32_t r2 = __viaddmin_u16x2(y2, 0x00020003u, e.z);
            const uint32_t r3 = __viaddmin_u16x2(y2, 0x00000001u, e.w);
            yin = (r0 & 0xFFFFu) + 1u;                                   // pixel 7 of the byte below
            const uint32_t a0 = __vimin3_u16x2(lr[4 * k + 0], r0, c2), a1 = __vimin3_u16x2(lr[4 * k + 1], r1, c2);
            const uint32_t a2 = __vimin3_u16x2(lr[4 * k + 2], r2, c2), a3 = __vimin3_u16x2(lr[4 * k + 3], r3, c2);
            packed[2 * k] = __byte_perm(a0, a1, 0x6420);
            packed[2 * k + 1] = __byte_perm(a2, a3, 0x6420);
        }
    }
    if (nvalid < 32) {  // pixels past the end of the row: "none"
#pragma unroll
        for (int b = 0; b < 32; ++b)
            if (b >= nvalid) packed[b >> 2] = (packed[b >> 2] & ~(0xFFu << (8 * (b & 3)))) | ((uint32_t)(cap - 1) << (8 * (b & 3)));
    }
    uint4* d4 = reinterpret_cast<uint4*>(rec + x0);
    d4[0] = make_uint4(packed[0], packed[1], packed[2], packed[3]);
    d4[1] = make_uint4(packed[4], packed[5], packed[6], packed[7]);
    crec[wi] = v;
}

__global__ void __launch_bounds__(kRow8Warps * 32, CQ_ROW8_MINB) rowscan8_kernel(RowScan8Params p) {
    extern __shared__ uint32_t s_bits8[];
    __shared__ Row8Tables s_tab;
    row8_build_tables(&s_tab, threadIdx.x, kRow8Warps * 32);
    __syncthreads();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const long long row = (long long)blockIdx.x * kRow8Warps + warp;
    if (row >= p.rows) return;  // whole warp leaves together; no block-level barrier is used below
    uint32_t* bits = s_bits8 + warp * (p.nwords + 1);
    const int w = p.w;
    const bool above = p.above != 0;
    const uint8_t* src = p.img + (size_t)row * (size_t)w * (size_t)p.stride + (size_t)p.offset;

    // ---------------- phase A: class bits of the row -> shared memory ----------------
    int word0 = 0;
    if (p.stride == 1 && ((reinterpret_cast<uintptr_t>(src) & 15) == 0)) {
        const int nfull = w >> 9;  // chunks of 512 pixels: 16 per lane
        const uint4* s4 = reinterpret_cast<const uint4*>(src);
#pragma unroll 8
        for (int i = 0; i < nfull; ++i) {  // unrolled: eight 512-byte loads of the warp in flight; the pass is load-latency bound otherwise
            uint4 q = __ldg(s4 + (size_t)i * 32 + lane);
            uint32_t m = class_nibble(q.x, above) | (class_nibble(q.y, above) << 4) |
                         (class_nibble(q.z, above) << 8) | (class_nibble(q.w, above) << 12);
            uint32_t o = __shfl_xor_sync(0xffffffffu, m, 1);
            if (!(lane & 1)) bits[i * 16 + (lane >> 1)] = m | (o << 16);
        }
        word0 = nfull * 16;
        if ((w & 15) == 0 && (w & 511) != 0) {
            // the rest of the row (or the whole of a narrow one, e.g. a 256-pixel glyph) as one partial 512-pixel
            // chunk: lanes whose 16 pixels lie inside the row load them, the others contribute no bits
            const int x = nfull * 512 + lane * 16;
            uint32_t m = 0u;
            if (x < w) {
                const uint4 q = __ldg(s4 + (size_t)nfull * 32 + lane);
                m = class_nibble(q.x, above) | (class_nibble(q.y, above) << 4) | (class_nibble(q.z, above) << 8) |
                    (class_nibble(q.w, above) << 12);
            }
            const uint32_t o = __shfl_xor_sync(0xffffffffu, m, 1);
            const int wi = nfull * 16 + (lane >> 1);
            if (!(lane & 1) && wi < p.nwords) bits[wi] = m | (o << 16);
            word0 = p.nwords;
        }
    }
    for (int wi = word0; wi < p.nwords; ++wi) {
        int x = wi * 32 + lane;
        bool c = false;
        if (x < w) {
            uint8_t px = __ldg(src + (size_t)x * (size_t)p.stride);
            c = above ? px > 127 : px < 127;
        }
        uint32_t b = __ballot_sync(0xffffffffu, c);
        if (lane == 0) bits[wi] = b;
    }
    __syncwarp();

    // ---------------- phase B: distances ----------------
    const int cap = p.cap;
    uint8_t* rec = p.inter + (size_t)row * (size_t)p.rs;
    uint32_t* crec = reinterpret_cast<uint32_t*>(rec + p.wp);
    const uint32_t lt = (1u << lane) - 1u, gt = lane == 31 ? 0u : (0xFFFFFFFEu << lane);
    int carryL = -kFar;  // last pixel of earlier chunks that has a class change immediately to its left
    const int nchunks = (p.nwords + 31) >> 5;
    for (int ch = 0; ch < nchunks; ++ch) {
        const int wi = ch * 32 + lane;
        const int x0 = wi * 32;
        int nvalid = w - x0;
        nvalid = nvalid < 0 ? 0 : (nvalid > 32 ? 32 : nvalid);
        int evalid = w - 1 - x0;  // positions b whose right neighbour b+1 exists
        evalid = evalid < 0 ? 0 : (evalid > 32 ? 32 : evalid);
        const uint32_t v = wi < p.nwords ? bits[wi] : 0u;
        const uint32_t nbit = wi + 1 < p.nwords ? (bits[wi + 1] & 1u) : 0u;
        const uint32_t pbit = (wi > 0 && wi < p.nwords) ? (bits[wi - 1] >> 31) : (v & 1u);
        const uint32_t em = evalid == 32 ? 0xFFFFFFFFu : ((1u << evalid) - 1u);
        const uint32_t R = (v ^ ((v >> 1) | (nbit << 31))) & em;     // bit b: classes of b and b+1 differ
        const uint32_t Lw = nvalid ? ((R << 1) | ((pbit ^ v) & 1u)) : 0u;  // bit b: classes of b-1 and b differ
        const uint32_t BL = __ballot_sync(0xffffffffu, Lw != 0), BR = __ballot_sync(0xffffffffu, R != 0);
        const int lastL = Lw ? x0 + (31 - __clz(Lw)) : -kFar;
        const int firstR = R ? x0 + __ffs(R) - 1 : kFar;

        // look-ahead: the first pixel of the next chunk whose class differs from this chunk's last pixel
        int look = 2 * kFar;
        if (ch + 1 < nchunks) {  // uniform; one chunk ahead is always enough (1024 >= cap)
            const uint32_t lastv = __shfl_sync(0xffffffffu, v, 31);
            const uint32_t flip = (lastv >> 31) ? 0xFFFFFFFFu : 0u;
            const int wj = wi + 32;
            int nv = w - wj * 32;
            nv = nv < 0 ? 0 : (nv > 32 ? 32 : nv);
            const uint32_t m2 = nv == 32 ? 0xFFFFFFFFu : ((1u << nv) - 1u);
            const uint32_t d2 = wj < p.nwords ? ((bits[wj] ^ flip) & m2) : 0u;
            const uint32_t C = __ballot_sync(0xffffffffu, d2 != 0);
            const int f = d2 ? wj * 32 + __ffs(d2) - 2 : kFar;  // the pixel left of the first differing one
            const int g = __shfl_sync(0xffffffffu, f, C ? __ffs(C) - 1 : 0);
            if (C) look = g;
        }
        const uint32_t pl = BL & lt, nr = BR & gt;
        const int sl = __shfl_sync(0xffffffffu, lastL, pl ? 31 - __clz(pl) : 0);
        const int sr = __shfl_sync(0xffffffffu, firstR, nr ? __ffs(nr) - 1 : 0);
        const int PL = pl ? sl : carryL;
        const int PR = nr ? sr : look;
        const int tl = __shfl_sync(0xffffffffu, lastL, BL ? 31 - __clz(BL) : 0);
        if (BL) carryL = tl;
        {   // busy hint: does any of this lane's pixels see a class change within reach? (4 lanes = one strip)
            const bool busy = nvalid > 0 && ((Lw | R) != 0u || x0 - PL + 1 < cap || PR - (x0 + nvalid - 1) + 1 < cap);
            const uint32_t B = __ballot_sync(0xffffffffu, busy);
            const int strip = ch * 8 + lane;
            if (lane < 8 && strip < p.nstrips) rec[p.wp + (p.wp >> 3) + strip] = ((B >> (4 * lane)) & 0xFu) ? 1 : 0;
        }

        if (nvalid > 0) row8_store(rec, crec, wi, x0, nvalid, PL, PR, Lw, R, v, cap, (BL | BR) == 0u, &s_tab);
    }
}

// Rows of at most 512 pixels (glyph atlases: thousands of 256-pixel rows): a warp per row would leave most lanes
// without a word, so a warp takes 32/S rows, S = lanes per row (a power of two >= nwords). Each lane loads its own
// 32 pixels (two 128-bit loads on the fast path), so no shared memory is needed; ballots are masked to the lane's
// row. The arithmetic per word is the wide kernel's.
__global__ void __launch_bounds__(kRow8Warps * 32, CQ_ROW8_MINB) rowscan8_narrow_kernel(RowScan8Params p, int shift_s) {
    __shared__ Row8Tables s_tab;
    row8_build_tables(&s_tab, threadIdx.x, kRow8Warps * 32);
    __syncthreads();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int S = 1 << shift_s, rpw = 32 >> shift_s;
    const int seg = lane >> shift_s, wi = lane & (S - 1);
    const long long row_w = ((long long)blockIdx.x * kRow8Warps + warp) * rpw;
    if (row_w >= p.rows) return;  // whole warp
    const long long row = row_w + seg;
    const int w = p.w, x0 = wi * 32;
    const bool above = p.above != 0;
    const bool live = row < p.rows && wi < p.nwords;
    int nvalid = w - x0, evalid = w - 1 - x0;
    nvalid = !live || nvalid < 0 ? 0 : (nvalid > 32 ? 32 : nvalid);
    evalid = !live || evalid < 0 ? 0 : (evalid > 32 ? 32 : evalid);

    uint32_t v = 0u;
    const bool fast = p.stride == 1 && (w & 15) == 0 && ((reinterpret_cast<uintptr_t>(p.img + p.offset) & 15) == 0);
    if (nvalid > 0) {
        const uint8_t* src = p.img + ((size_t)row * (size_t)w + (size_t)x0) * (size_t)p.stride + (size_t)p.offset;
        if (fast) {  // nvalid is 16 or 32
            const uint4 q = __ldg(reinterpret_cast<const uint4*>(src));
            v = class_nibble(q.x, above) | (class_nibble(q.y, above) << 4) | (class_nibble(q.z, above) << 8) |
                (class_nibble(q.w, above) << 12);
            if (nvalid > 16) {
                const uint4 r = __ldg(reinterpret_cast<const uint4*>(src) + 1);
                v |= (class_nibble(r.x, above) | (class_nibble(r.y, above) << 4) | (class_nibble(r.z, above) << 8) |
                      (class_nibble(r.w, above) << 12)) << 16;
            }
        } else {
            for (int b = 0; b < nvalid; ++b) {
                const uint8_t px = __ldg(src + (size_t)b * (size_t)p.stride);
                v |= (uint32_t)(above ? px > 127 : px < 127) << b;
            }
        }
    }
    const uint32_t vn = __shfl_down_sync(0xffffffffu, v, 1), vp = __shfl_up_sync(0xffffffffu, v, 1);
    const uint32_t nbit = wi + 1 < p.nwords ? (vn & 1u) : 0u;
    const uint32_t pbit = (wi > 0 && wi < p.nwords) ? (vp >> 31) : (v & 1u);
    const uint32_t em = evalid == 32 ? 0xFFFFFFFFu : ((1u << evalid) - 1u);
    const uint32_t R = (v ^ ((v >> 1) | (nbit << 31))) & em;
    const uint32_t Lw = nvalid ? ((R << 1) | ((pbit ^ v) & 1u)) : 0u;
    const uint32_t segmask = (S == 32 ? 0xFFFFFFFFu : ((1u << S) - 1u)) << (seg * S);
    const uint32_t BLw = __ballot_sync(0xffffffffu, Lw != 0), BRw = __ballot_sync(0xffffffffu, R != 0);
    const uint32_t BL = BLw & segmask, BR = BRw & segmask;
    const int lastL = Lw ? x0 + (31 - __clz(Lw)) : -kFar;
    const int firstR = R ? x0 + __ffs(R) - 1 : kFar;
    const uint32_t lt = (1u << lane) - 1u, gt = lane == 31 ? 0u : (0xFFFFFFFEu << lane);
    const uint32_t pl = BL & lt, nr = BR & gt;
    const int sl = __shfl_sync(0xffffffffu, lastL, pl ? 31 - __clz(pl) : 0);
    const int sr = __shfl_sync(0xffffffffu, firstR, nr ? __ffs(nr) - 1 : 0);
    const int PL = pl ? sl : -kFar;
    const int PR = nr ? sr : 2 * kFar;
    const int cap = p.cap;
    const bool busy = nvalid > 0 && ((Lw | R) != 0u || x0 - PL + 1 < cap || PR - (x0 + nvalid - 1) + 1 < cap);
    const uint32_t B = __ballot_sync(0xffffffffu, busy);
    if (nvalid > 0) {
        uint8_t* rec = p.inter + (size_t)row * (size_t)p.rs;
        if ((wi & 3) == 0) rec[p.wp + (p.wp >> 3) + (wi >> 2)] = (B & (0xFu << lane) & segmask) ? 1 : 0;  // 4 lanes = one strip
        row8_store(rec, reinterpret_cast<uint32_t*>(rec + p.wp), wi, x0, nvalid, PL, PR, Lw, R, v, cap, (BLw | BRw) == 0u, &s_tab);
    }
}

}  // namespace chaq
