// host_api.cu -- host-pointer entry points of include/chaqsdf.h: the df.h:6 drop-in and the fused
// replacement of reference openmp/sdfgen.c:261-300. They stage through a grow-only per-process device pool
// (one lock: concurrent callers, e.g. the two OpenMP section threads of sdfgen.c:277-289, serialise) and
// run the same device entry points as everything else. No CPU compute path.
#include <cuda_runtime.h>

#include <cmath>
#include <cstring>
#include <mutex>
#include <string>

#include "../../include/chaqsdf.h"

namespace chaq {
int set_error(int code, const char* msg);
void note_launch();
}  // namespace chaq
using chaq::set_error;

namespace {

struct Pool {
    std::mutex mu;
    int dev = -1;
    void* buf[4] = {nullptr, nullptr, nullptr, nullptr};
    size_t cap[4] = {0, 0, 0, 0};
    cudaStream_t stream = nullptr;
    int* flag = nullptr;
};
Pool g_pool;

cudaError_t pool_get(int slot, size_t bytes, void** out) {
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    if (g_pool.dev != dev) {  // device changed: drop everything (buffers belong to the old device)
        for (int i = 0; i < 4; ++i) { g_pool.buf[i] = nullptr; g_pool.cap[i] = 0; }
        g_pool.stream = nullptr; g_pool.flag = nullptr;
        g_pool.dev = dev;
    }
    if (!g_pool.stream) {
        e = cudaStreamCreateWithFlags(&g_pool.stream, cudaStreamNonBlocking);
        if (e != cudaSuccess) return e;
        e = cudaMalloc(&g_pool.flag, sizeof(int));
        if (e != cudaSuccess) return e;
    }
    if (g_pool.cap[slot] < bytes) {
        if (g_pool.buf[slot]) cudaFree(g_pool.buf[slot]);
        g_pool.buf[slot] = nullptr; g_pool.cap[slot] = 0;
        e = cudaMalloc(&g_pool.buf[slot], bytes);
        if (e != cudaSuccess) return e;
        g_pool.cap[slot] = bytes;
    }
    *out = g_pool.buf[slot];
    return cudaSuccess;
}

// {0,+INF} float heights -> 255/0 mask bytes; anything else raises the flag (general heights: not this path)
__global__ void heights_to_mask_kernel(const float* __restrict__ f, uint8_t* __restrict__ m, size_t n, int* flag) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    bool bad = false;
    for (; i < n; i += stride) {
        const float v = f[i];
        const bool seed = v == 0.f;
        const bool none = v == __int_as_float(0x7f800000);
        bad |= !(seed || none);
        m[i] = seed ? 255 : 0;
    }
    if (bad) *flag = 1;
}

}  // namespace

extern "C" {

int chaqsdf_u8(const uint8_t* img, size_t w, size_t h, size_t stride, size_t offset, int test_above, size_t spread,
               int asymmetric, uint8_t* out) {
    if (!img || !out) return set_error(CHAQSDF_ERR_ARG, "null pointer");
    if (stride == 0 || offset >= stride) return set_error(CHAQSDF_ERR_ARG, "bad stride/offset");
    if (w == 0 || h == 0) return set_error(CHAQSDF_ERR_ARG, "empty image");
    if (w > CHAQSDF_MAX_DIM || h > CHAQSDF_MAX_DIM) return set_error(CHAQSDF_ERR_ARG, "image side exceeds 32768");
    if (spread == 0) return set_error(CHAQSDF_ERR_ARG, "spread must be a positive integer");
    const size_t ws_bytes = chaqsdf_workspace_bytes(w, h, spread);
    if (ws_bytes == 0) return CHAQSDF_ERR_CUDA;  // error string already set by the planner
    std::lock_guard<std::mutex> lk(g_pool.mu);
    void *d_img, *d_out, *d_ws;
    const size_t in_bytes = w * h * stride, out_bytes = w * h;
    if (pool_get(0, in_bytes, &d_img) != cudaSuccess || pool_get(1, out_bytes, &d_out) != cudaSuccess ||
        pool_get(2, ws_bytes, &d_ws) != cudaSuccess)
        return set_error(CHAQSDF_ERR_CUDA, "device allocation failed");
    cudaStream_t st = g_pool.stream;
    if (cudaMemcpyAsync(d_img, img, in_bytes, cudaMemcpyHostToDevice, st) != cudaSuccess) return CHAQSDF_ERR_CUDA;
    int rc = chaqsdf_u8_device((const uint8_t*)d_img, w, h, stride, offset, test_above, spread, asymmetric,
                               (uint8_t*)d_out, d_ws, ws_bytes, st);
    if (rc) return rc;
    if (cudaMemcpyAsync(out, d_out, out_bytes, cudaMemcpyDeviceToHost, st) != cudaSuccess) return CHAQSDF_ERR_CUDA;
    if (cudaStreamSynchronize(st) != cudaSuccess) return CHAQSDF_ERR_CUDA;
    return CHAQSDF_OK;
}

// df.h:6. Status of the last call on this thread is readable through chaqsdf_dt2d_last_status().
static thread_local int g_dt_status = 0;
int chaqsdf_dt2d_last_status(void) { return g_dt_status; }

void dist_transform_2d(float* img, size_t w, size_t h) {
    g_dt_status = CHAQSDF_ERR_ARG;
    if (!img || w == 0 || h == 0) return;
    const size_t ws_bytes = chaqsdf_fields_workspace_bytes(w, h);
    if (ws_bytes == 0) { g_dt_status = CHAQSDF_ERR_CUDA; return; }
    std::lock_guard<std::mutex> lk(g_pool.mu);
    void *d_f, *d_mask, *d_ws;
    const size_t n = w * h;
    g_dt_status = CHAQSDF_ERR_CUDA;
    if (pool_get(0, n * sizeof(float), &d_f) != cudaSuccess || pool_get(1, n, &d_mask) != cudaSuccess ||
        pool_get(2, ws_bytes, &d_ws) != cudaSuccess)
        return;
    cudaStream_t st = g_pool.stream;
    if (cudaMemsetAsync(g_pool.flag, 0, sizeof(int), st) != cudaSuccess) return;
    if (cudaMemcpyAsync(d_f, img, n * sizeof(float), cudaMemcpyHostToDevice, st) != cudaSuccess) return;
    const unsigned blocks = (unsigned)((n + 1023) / 1024 < 65535 ? (n + 1023) / 1024 : 65535);
    heights_to_mask_kernel<<<blocks, 256, 0, st>>>((const float*)d_f, (uint8_t*)d_mask, n, g_pool.flag);
    if (cudaGetLastError() != cudaSuccess) return;
    chaq::note_launch();
    int flag = 0;
    if (cudaMemcpyAsync(&flag, g_pool.flag, sizeof(int), cudaMemcpyDeviceToHost, st) != cudaSuccess) return;
    if (cudaStreamSynchronize(st) != cudaSuccess) return;
    if (flag) { g_dt_status = set_error(CHAQSDF_ERR_ARG, "dist_transform_2d: heights other than 0/+INF"); return; }  // heights other than {0,+INF}: untouched
    // seeds are class 1; f_in = distance to the nearest class-1 pixel (0 on the seeds themselves)
    int rc = chaqsdf_fields_device((const uint8_t*)d_mask, w, h, 1, 0, 1, 1, 0, nullptr, nullptr, (float*)d_f, nullptr,
                                   nullptr, nullptr, d_ws, ws_bytes, st);
    if (rc) { g_dt_status = rc; return; }
    if (cudaMemcpyAsync(img, d_f, n * sizeof(float), cudaMemcpyDeviceToHost, st) != cudaSuccess) return;
    if (cudaStreamSynchronize(st) != cudaSuccess) return;
    g_dt_status = CHAQSDF_OK;
}

}  // extern "C"
