// rowscan.cuh -- pass 1: threshold + per-row distance to the nearest pixel of the other class.
//
// Replaces, fused into one kernel and for both fields at once:
//   transform_img_to_bool    reference openmp/sdfgen.c:52-62   (px > 127 / px < 127 on a strided byte image)
//   transform_bool_to_float  reference openmp/sdfgen.c:65-72   ({0,INF} seeds for each field)
//   dist_transform_axis #1   reference openmp/df.c:130 -> df.c:29-97 (for {0,INF} input the lower envelope
//                            along a row is (x - nearest seed)^2, i.e. a run-length scan)
// Output: uint16 per pixel, bit 15 = class (1 = inside, the reference's `true`), bits 0..14 = min(distance
// along the row to the nearest pixel of the other class, cap). No transposed store (df.c:92): the second
// pass walks columns with adjacent threads on adjacent columns instead.
// Record mode (RowScanParams::record, spreads 257..1024 of the banded path): per row [wp x uint16 min(G, cap) - 1 |
// wp/8 bytes of class bits], the 16-bit sibling of rowscan8.cuh's record, read by the hull kernels (hull.cuh, WIDE).
//
// Mapping: one warp per row. Phase A turns the row into a bit string in shared memory (16 pixels per lane
// per 128-bit load on the fast path, ballot on the generic path). Phase B gives each lane one 32-pixel
// word; the nearest other-class pixel outside the word comes from warp ballots (clz/ffs on the ballot of
// "word has a pixel of class k") plus a carry from earlier chunks and a bounded look-ahead into later
// ones; inside the word two 32-step run-length ramps (left, right) finish the job in registers.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace chaq {

constexpr int kRowWarps = 4;  // warps (= rows) per CTA
constexpr int kRowMaxParts = 16;  // column parts of the scatter mode (one per GPU)
constexpr int kFar = 1 << 20; // "no such pixel" position offset; larger than any cap, small enough not to overflow

__device__ __forceinline__ uint32_t class_nibble(uint32_t wd, bool above) {
    // 4 pixels in a 32-bit word -> 4 class bits. above: px > 127 <=> bit 7 set; else px < 127.
    // One formula for both tests, plain 32-bit operations (the byte-wise SIMD compare is emulated on this architecture):
    // px >= 127 <=> bit 7 of px, or its low seven bits are all ones (adding 1 to them carries into bit 7).
    const uint32_t add = above ? 0u : 0x01010101u;      // above: bit 7 of px itself; else: px >= 127 ...
    const uint32_t flip = above ? 0u : 0x01010101u;     // ... inverted: px < 127
    const uint32_t t = ((wd & 0x7f7f7f7fu) + add) | wd;
    const uint32_t x = ((t >> 7) & 0x01010101u) ^ flip;
    return (x * 0x01020408u) >> 24;
}

struct RowScanParams {
    const uint8_t* img;
    uint16_t* inter;
    long long rows;
    int w;
    int stride, offset;
    int above;
    int cap;
    int nwords;  // ceil(w/32)
    int record;          // 0: [rows][w] uint16 with the class in bit 15; 1: records (below)
    int wp;              // record mode: w rounded up to 128
    long long rs_bytes;  // record mode: bytes per record = 2*wp + wp/8
    // scatter mode (nparts > 0; the all-to-all of the multi-GPU exact path fused into the store): the row goes out in
    // column parts of part_w columns (a multiple of 32; the last part takes what is left), part q to part[q], which
    // holds ALL image rows of that part's columns: element (image row r, column x) at part[q][r * width(q) + x - q*part_w]
    int nparts, part_w;
    long long row0;      // image row of row 0 of this launch
    uint16_t* part[kRowMaxParts];
};

__global__ void __launch_bounds__(kRowWarps * 32) rowscan_kernel(RowScanParams p) {
    extern __shared__ uint32_t s_bits[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const long long row = (long long)blockIdx.x * kRowWarps + warp;
    if (row >= p.rows) return;  // whole warp leaves together; no block-level barrier is used below
    uint32_t* bits = s_bits + warp * p.nwords;
    const int w = p.w;
    const bool above = p.above != 0;
    const uint8_t* src = p.img + (size_t)row * (size_t)w * (size_t)p.stride + (size_t)p.offset;

    // ---------------- phase A: class bits of the row -> shared memory ----------------
    int word0 = 0;
    if (p.stride == 1 && ((reinterpret_cast<uintptr_t>(src) & 15) == 0)) {
        const int nfull = w >> 9;  // chunks of 512 pixels: 16 per lane
        const uint4* s4 = reinterpret_cast<const uint4*>(src);
#pragma unroll 8
        for (int i = 0; i < nfull; ++i) {  // eight loads in flight per warp: the loop is load-latency bound otherwise
            uint4 q = __ldg(s4 + (size_t)i * 32 + lane);
            uint32_t m = class_nibble(q.x, above) | (class_nibble(q.y, above) << 4) |
                         (class_nibble(q.z, above) << 8) | (class_nibble(q.w, above) << 12);
            uint32_t o = __shfl_xor_sync(0xffffffffu, m, 1);
            if (!(lane & 1)) bits[i * 16 + (lane >> 1)] = m | (o << 16);
        }
        word0 = nfull * 16;
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
    const bool rec = p.record != 0;
    uint8_t* recbase = reinterpret_cast<uint8_t*>(p.inter) + (size_t)row * (size_t)p.rs_bytes;
    uint16_t* dst = rec ? reinterpret_cast<uint16_t*>(recbase) : p.inter + (size_t)row * (size_t)w;
    if (rec) {  // class words of the row, as rowscan8 leaves them
        uint32_t* crec = reinterpret_cast<uint32_t*>(recbase + 2 * (size_t)p.wp);
        for (int wi = lane; wi < p.nwords; wi += 32) crec[wi] = bits[wi];
    }
    const bool vec_ok = (reinterpret_cast<uintptr_t>(dst) & 15) == 0;
    const uint32_t lt = (1u << lane) - 1u, gt = lane == 31 ? 0u : (0xFFFFFFFEu << lane);
    int carry0 = -kFar, carry1 = -kFar;  // last pixel of class 0 / 1 in earlier chunks
    const int nchunks = (p.nwords + 31) >> 5;
    for (int ch = 0; ch < nchunks; ++ch) {
        const int wi = ch * 32 + lane;
        const int x0 = wi * 32;
        int nvalid = w - x0;
        nvalid = nvalid < 0 ? 0 : (nvalid > 32 ? 32 : nvalid);
        const uint32_t vm = nvalid == 32 ? 0xFFFFFFFFu : ((1u << nvalid) - 1u);
        const uint32_t v = wi < p.nwords ? bits[wi] : 0u;
        const uint32_t ones = v & vm, zeros = ~v & vm;
        const uint32_t B0 = __ballot_sync(0xffffffffu, zeros != 0), B1 = __ballot_sync(0xffffffffu, ones != 0);
        const int last0 = zeros ? x0 + (31 - __clz(zeros)) : -kFar, first0 = zeros ? x0 + __ffs(zeros) - 1 : kFar;
        const int last1 = ones ? x0 + (31 - __clz(ones)) : -kFar, first1 = ones ? x0 + __ffs(ones) - 1 : kFar;

        // look-ahead: first pixel of each class in later chunks, as far as cap can reach
        int look0 = kFar * 2, look1 = kFar * 2;
        {
            const int chunk_end = ch * 1024 + 1023;
            for (int cc = ch + 1; cc < nchunks; ++cc) {
                if (cc * 1024 - chunk_end >= cap) break;
                const int wj = cc * 32 + lane;
                int nv = w - wj * 32;
                nv = nv < 0 ? 0 : (nv > 32 ? 32 : nv);
                const uint32_t m2 = nv == 32 ? 0xFFFFFFFFu : ((1u << nv) - 1u);
                const uint32_t v2 = wj < p.nwords ? bits[wj] : 0u;
                const uint32_t o2 = v2 & m2, z2 = ~v2 & m2;
                const uint32_t C0 = __ballot_sync(0xffffffffu, z2 != 0), C1 = __ballot_sync(0xffffffffu, o2 != 0);
                const int f0 = z2 ? wj * 32 + __ffs(z2) - 1 : kFar, f1 = o2 ? wj * 32 + __ffs(o2) - 1 : kFar;
                const int g0 = __shfl_sync(0xffffffffu, f0, C0 ? __ffs(C0) - 1 : 0);
                const int g1 = __shfl_sync(0xffffffffu, f1, C1 ? __ffs(C1) - 1 : 0);
                if (look0 >= kFar && C0) look0 = g0;
                if (look1 >= kFar && C1) look1 = g1;
                if (look0 < kFar && look1 < kFar) break;
            }
        }
        // nearest pixel of class k strictly outside this lane's word
        const uint32_t p0 = B0 & lt, p1 = B1 & lt, n0 = B0 & gt, n1 = B1 & gt;
        const int sl0 = __shfl_sync(0xffffffffu, last0, p0 ? 31 - __clz(p0) : 0);
        const int sl1 = __shfl_sync(0xffffffffu, last1, p1 ? 31 - __clz(p1) : 0);
        const int sf0 = __shfl_sync(0xffffffffu, first0, n0 ? __ffs(n0) - 1 : 0);
        const int sf1 = __shfl_sync(0xffffffffu, first1, n1 ? __ffs(n1) - 1 : 0);
        const int L0 = p0 ? sl0 : carry0, L1 = p1 ? sl1 : carry1;
        const int N0 = n0 ? sf0 : look0, N1 = n1 ? sf1 : look1;
        const int t0 = __shfl_sync(0xffffffffu, last0, B0 ? 31 - __clz(B0) : 0);
        const int t1 = __shfl_sync(0xffffffffu, last1, B1 ? 31 - __clz(B1) : 0);
        if (B0) carry0 = t0;
        if (B1) carry1 = t1;

        if (nvalid > 0) {
            // left ramp: z_k = distance from the current pixel to the last class-k pixel before it
            int dl[32];
            int z0 = x0 - L0, z1 = x0 - L1;
#pragma unroll
            for (int b = 0; b < 32; ++b) {
                const bool c = (v >> b) & 1u;
                dl[b] = c ? z0 : z1;
                z0 = c ? z0 + 1 : 1;
                z1 = c ? 1 : z1 + 1;
            }
            // right ramp + combine; pixels past the row end must not act as neighbours
            const int xl = x0 + nvalid - 1;
            int y0 = N0 - xl, y1 = N1 - xl;  // distances from the last valid pixel to the first class-k pixel after it
            uint32_t packed[16];
#pragma unroll
            for (int b = 31; b >= 0; --b) {
                const bool c = (v >> b) & 1u;
                const bool valid = b < nvalid;
                int dr = c ? y0 : y1;
                int m = min(min(dl[b], dr), cap);
                uint32_t o = rec ? (uint32_t)((valid ? m : cap) - 1) : ((c ? 0x8000u : 0u) | (uint32_t)m);
                if (b & 1) packed[b >> 1] = o << 16; else packed[b >> 1] |= o;
                if (valid) {
                    y0 = c ? y0 + 1 : 1;
                    y1 = c ? 1 : y1 + 1;
                }
            }
            uint16_t* wdst = dst + x0;
            bool v_ok = vec_ok;
            if (p.nparts > 0) {  // this word belongs to one column part (parts start at multiples of 32 columns)
                const int q = x0 / p.part_w;
                const int pw = min(p.part_w, w - q * p.part_w);
                wdst = p.part[q] + (size_t)(p.row0 + row) * (size_t)pw + (size_t)(x0 - q * p.part_w);
                v_ok = (reinterpret_cast<uintptr_t>(wdst) & 15) == 0;
            }
            if ((nvalid == 32 || rec) && v_ok) {  // a record has room for the whole last word (padding: nothing in reach)
                uint4* d4 = reinterpret_cast<uint4*>(wdst);
#pragma unroll
                for (int q = 0; q < 4; ++q)
                    d4[q] = make_uint4(packed[4 * q], packed[4 * q + 1], packed[4 * q + 2], packed[4 * q + 3]);
            } else {
#pragma unroll
                for (int b = 0; b < 32; ++b)  // static indices keep packed[] in registers
                    if (b < nvalid) wdst[b] = (uint16_t)((b & 1) ? packed[b >> 1] >> 16 : packed[b >> 1] & 0xffffu);
            }
        }
    }
}

}  // namespace chaq
