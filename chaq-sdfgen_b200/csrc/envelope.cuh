// envelope.cuh -- pass 2 kernel: column lower envelope for both fields + fused epilogue.
//
// Replaces dist_transform_axis #2 (reference openmp/df.c:133 -> df.c:29-97, incl. the sqrtf at df.c:95) for
// the inside and the outside field of openmp/sdfgen.c:277-289 in ONE launch, and fuses
// transform_float_sub (sdfgen.c:98-106) and transform_float_to_byte (sdfgen.c:75-96) as an exact integer
// threshold look-up (tables.h). The algorithm per column is envelope_core.h.
//
// Mapping: thread = (column, row band); adjacent threads own adjacent columns, so every row access of a
// warp is one contiguous 64-byte piece of the uint16 intermediate and one 32-byte piece of the output;
// nothing is transposed. The hull / class-change bitmask rings live in shared memory laid out
// [word][thread] (bank = lane, conflict-free whatever each lane's depth); for the unwindowed exact mode the
// rings are too long for shared memory and sit in the caller's workspace with the same layout.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "envelope_core.h"
#include "tables.h"

namespace chaq {

constexpr int kEnvThreads = 128;

struct EnvParams {
    const uint16_t* inter;  // [n_img][buf_rows][w]
    int w, buf_rows;
    int out_begin, out_end;  // rows of the buffer to produce
    int band_rows;
    int cap, nw;
    const QuantTable* table;
    uint8_t* out;        // [n_img][out_end-out_begin][w] or null
    int32_t* d2_in;      // fields mode, [n_img][out rows][w], any may be null
    int32_t* d2_out;
    float* f_in;
    float* f_out;
    float* sdf;
    uint32_t* gring;     // global ring scratch (GLOBAL_RING): [2][nw][total threads]
};

template <bool FIELDS>
struct DevCtx {
    const uint16_t* col;  // inter + image offset + x
    size_t w;
    uint32_t* hull;       // this thread's slot of ring word 0
    uint32_t* chg;
    int rstride;          // distance between ring words
    // epilogue
    const int32_t* tin; const int32_t* tout;
    int cap2; float smin, smax, scale;
    uint8_t* o8; int32_t* d2_in; int32_t* d2_out; float* f_in; float* f_out; float* sdf;
    int out_begin;

    __device__ __forceinline__ uint32_t ld(int r) const { return __ldg(col + (size_t)r * w); }
    __device__ __forceinline__ uint32_t hull_get(int i) const { return hull[i * rstride]; }
    __device__ __forceinline__ void hull_set(int i, uint32_t v) { hull[i * rstride] = v; }
    __device__ __forceinline__ uint32_t chg_get(int i) const { return chg[i * rstride]; }
    __device__ __forceinline__ void chg_set(int i, uint32_t v) { chg[i * rstride] = v; }
    __device__ __forceinline__ void out_px(int r, int cls, int d2) {
        const size_t o = (size_t)(r - out_begin) * w;
        if (o8) o8[o] = (uint8_t)quantise(tin, tout, cap2, smin, smax, scale, cls, d2);
        if (FIELDS) {
            const bool none = d2 == kBig;
            const float d = none ? __int_as_float(0x7f800000) : __fsqrt_rn((float)d2);
            if (d2_in) d2_in[o] = cls ? 0 : d2;
            if (d2_out) d2_out[o] = cls ? d2 : 0;
            if (f_in) f_in[o] = cls ? 0.f : d;
            if (f_out) f_out[o] = cls ? d : 0.f;
            if (sdf) sdf[o] = cls ? d : __fsub_rn(0.f, __fadd_rn(d, -1.f));  // sdfgen.c:102-104
        }
    }
    __device__ __forceinline__ void out(int r, int cls, int d2) { out_px(r, cls, d2); }
};

template <bool WIDE, bool GLOBAL_RING, bool FIELDS>
__global__ void __launch_bounds__(kEnvThreads) envelope_kernel(EnvParams p) {
    extern __shared__ uint32_t s_mem[];
    // quantisation thresholds -> shared memory
    int32_t* s_tin = reinterpret_cast<int32_t*>(s_mem);
    int32_t* s_tout = s_tin + 257;
    uint32_t* s_ring = s_mem + 516;
    for (int i = threadIdx.x; i < 257; i += kEnvThreads) {
        s_tin[i] = p.table->tin[i];
        s_tout[i] = p.table->tout[i];
    }
    __syncthreads();
    const int x = blockIdx.x * kEnvThreads + threadIdx.x;
    if (x >= p.w) return;
    const int img = blockIdx.z;
    const int y0 = p.out_begin + blockIdx.y * p.band_rows;
    if (y0 >= p.out_end) return;
    const int y1 = min(y0 + p.band_rows, p.out_end);

    DevCtx<FIELDS> c;
    const size_t n_out_rows = (size_t)(p.out_end - p.out_begin);
    c.w = (size_t)p.w;
    c.col = p.inter + (size_t)img * (size_t)p.buf_rows * c.w + x;
    if (GLOBAL_RING) {
        const size_t nthreads = (size_t)gridDim.x * kEnvThreads * gridDim.y * gridDim.z;
        const size_t tid = ((size_t)(blockIdx.z * gridDim.y + blockIdx.y) * gridDim.x + blockIdx.x) * kEnvThreads + threadIdx.x;
        c.hull = p.gring + tid;
        c.chg = p.gring + (size_t)p.nw * nthreads + tid;
        c.rstride = (int)nthreads;
    } else {
        c.hull = s_ring + threadIdx.x;
        c.chg = s_ring + p.nw * kEnvThreads + threadIdx.x;
        c.rstride = kEnvThreads;
    }
    c.tin = s_tin; c.tout = s_tout;
    c.cap2 = p.table->cap2; c.smin = p.table->smin; c.smax = p.table->smax; c.scale = p.table->scale;
    const size_t obase = (size_t)img * n_out_rows * c.w + x;
    c.o8 = p.out ? p.out + obase : nullptr;
    c.d2_in = FIELDS && p.d2_in ? p.d2_in + obase : nullptr;
    c.d2_out = FIELDS && p.d2_out ? p.d2_out + obase : nullptr;
    c.f_in = FIELDS && p.f_in ? p.f_in + obase : nullptr;
    c.f_out = FIELDS && p.f_out ? p.f_out + obase : nullptr;
    c.sdf = FIELDS && p.sdf ? p.sdf + obase : nullptr;
    c.out_begin = p.out_begin;

    ColumnEnvelope<DevCtx<FIELDS>, WIDE> ce(c, p.nw, p.cap);
    ce.run(p.buf_rows, y0, y1);
}

}  // namespace chaq
