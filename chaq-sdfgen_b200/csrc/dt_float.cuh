// dt_float.cuh -- dist_transform_2d for GENERAL heights (any finite value or +INF), in the reference's own float32
// arithmetic: a restatement of dist_transform_1d (reference openmp/df.c:29-97) with every operation rounded once
// (__f*_rn, no contraction), so the result is bit-identical to the stock OpenMP build wherever that build is
// defined (square images; it also reproduces the stock build's inexactness beyond side 4098, where q*q is no
// longer representable in float32 -- df.c:9-15). The binary {0,+INF} case normally takes the exact integer
// kernels instead; this path serves sampled functions and, on request, bug-compatible output.
//
// Mapping: one thread per line (rows in pass 1, columns of the transposed image in pass 2). The vertex and
// break-point stacks (df.c's v[] and z[]; h[] is re-read from the line) live in global scratch, interleaved by
// line so that lines at the same depth touch neighbouring words. Output is written transposed (df.c:92), which
// makes the stores of a warp contiguous. Throughput is modest (16K-32K threads); this is the general path, not
// the hot one.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace chaq {

struct DtFloatParams {
    const float* in;   // [nlines][len]
    float* out;        // [len][nlines] (transposed)
    int* v;            // [len][nlines] vertex stack
    float* z;          // [len][nlines] break points
    int len, nlines;
    int do_sqrt;
};

// df.c:9-15, one rounding per operation
__device__ __forceinline__ float dt_intersect(const float* f, int p, int q) {
    const float p1x = (float)p, p2x = (float)q;
    const float p1y = __ldg(f + p), p2y = __ldg(f + q);
    const float num = __fadd_rn(__fsub_rn(p2y, p1y), __fsub_rn(__fmul_rn(p2x, p2x), __fmul_rn(p1x, p1x)));
    return __fdiv_rn(num, __fmul_rn(2.f, __fsub_rn(p2x, p1x)));
}

__global__ void __launch_bounds__(128) dt_float_kernel(DtFloatParams p) {
    const int line = blockIdx.x * 128 + threadIdx.x;
    if (line >= p.nlines) return;
    const float* f = p.in + (size_t)line * p.len;
    float* out = p.out + line;
    const size_t ns = (size_t)p.nlines;
    const int w = p.len;
    const float inf = __int_as_float(0x7f800000);
    if (w <= 1) {  // a line of one sample is its own transform (df.c:32-36 writes it back unchanged)
        float r = __ldg(f);
        if (p.do_sqrt) r = __fsqrt_rn(r);
        out[0] = r;
        return;
    }
    int offset = 0;
    while (offset < w && fabsf(__ldg(f + offset)) == inf) ++offset;  // isinf, df.c:41
    if (offset == w) {  // df.c:44-51
        for (int i = 0; i < w; ++i) out[ns * i] = inf;
        return;
    }
    int* v = p.v + line;
    float* z = p.z + line;
    v[0] = offset;
    int k = 0, vk = offset;  // vk = v[k]
    for (int q = offset + 1; q < w; ++q) {
        if (fabsf(__ldg(f + q)) == inf) continue;     // df.c:60
        float s = dt_intersect(f, vk, q);
        while (k > 0 && s <= z[ns * (k - 1)]) {        // df.c:67-70
            --k;
            vk = v[ns * k];
            s = dt_intersect(f, vk, q);
        }
        z[ns * k] = s;
        ++k;
        v[ns * k] = q;
        vk = q;
    }
    int j = 0;
    int vj = v[0];
    float hj = __ldg(f + vj);
    float zj = k > 0 ? z[0] : 0.f;
    for (int q = 0; q < w; ++q) {
        const float fq = (float)q;
        while (j < k && zj < fq) {                      // df.c:85
            ++j;
            vj = v[ns * j];
            hj = __ldg(f + vj);
            if (j < k) zj = z[ns * j];
        }
        const float d = __fsub_rn(fq, (float)vj);
        float r = __fadd_rn(__fmul_rn(d, d), hj);
        if (p.do_sqrt) r = __fsqrt_rn(r);
        out[ns * q] = r;
    }
}

}  // namespace chaq
