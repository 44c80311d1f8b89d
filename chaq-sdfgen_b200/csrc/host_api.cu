// host_api.cu -- host-pointer entry points of include/chaqsdf.h: the df.h:6 drop-in and the fused
// replacement of reference openmp/sdfgen.c:261-300. They stage through one grow-only pool per device (one lock per
// device: concurrent callers on the same device, e.g. the two OpenMP section threads of sdfgen.c:277-289,
// serialise; callers on different devices do not) and run the same device entry points as everything else.
// No CPU compute path.
#include <cuda_runtime.h>

#include <cmath>
#include <cstdlib>
#include <cstring>
#include <map>
#include <mutex>
#include <string>
#include <vector>

#include "../../include/chaqsdf.h"
#include "dt_float.cuh"
#include "tuning.h"

namespace chaq {
int set_error(int code, const char* msg);
void note_launch();
}  // namespace chaq
using chaq::set_error;

namespace {

struct Pool {
    std::mutex mu;
    void* buf[4] = {nullptr, nullptr, nullptr, nullptr};
    size_t cap[4] = {0, 0, 0, 0};
    cudaStream_t stream = nullptr;
    cudaStream_t s_in = nullptr, s_out = nullptr;  // copy streams of the banded pipeline
    std::vector<cudaEvent_t> ev;                   // grow-only event cache
    int* flag = nullptr;
};
std::mutex g_pools_mu;
std::map<int, Pool*> g_pools;  // one per device, never freed (process lifetime)

// the pool of the calling thread's current device
Pool* current_pool() {
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) return nullptr;
    std::lock_guard<std::mutex> lk(g_pools_mu);
    Pool*& p = g_pools[dev];
    if (!p) p = new Pool();
    return p;
}

cudaError_t pool_get(Pool& g_pool, int slot, size_t bytes, void** out) {
    cudaError_t e;
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

// after a failure inside the pipeline: nothing of this call may still be running when the caller gets its buffers back
int drained(Pool& g_pool, int code, const char* msg) {
    if (g_pool.s_in) cudaStreamSynchronize(g_pool.s_in);
    if (g_pool.stream) cudaStreamSynchronize(g_pool.stream);
    if (g_pool.s_out) cudaStreamSynchronize(g_pool.s_out);
    return msg ? set_error(code, msg) : code;
}

cudaError_t pool_events(Pool& g_pool, size_t n) {
    while (g_pool.ev.size() < n) {
        cudaEvent_t e;
        cudaError_t r = cudaEventCreateWithFlags(&e, cudaEventDisableTiming);
        if (r != cudaSuccess) return r;
        g_pool.ev.push_back(e);
    }
    return cudaSuccess;
}

// Host buffers, banded path: the image goes through in row bands so that the upload of band k+1, the kernels of
// band k and the download of band k-1 overlap (three streams). Pass 1 is per row; pass 2 of output rows
// [done, b - spread) only needs pass 1 up to row b, and a sub-buffer starting >= spread rows above `done` (both
// are the exact-context rule of chaqsdf_envelope_device). Results are identical to the one-shot path.
// Band heights: short bands first (the download can start early) and last (little is left to do once the upload
// ends), `cap` rows in between (large launches of pass 2 waste less on the context rows of each piece).
std::vector<size_t> pipeline_bands(size_t h_all, size_t cap, size_t first, size_t min_rows) {
    std::vector<size_t> up, down, v;
    const size_t tail = h_all % 32, h = h_all - tail;  // every band a multiple of 32 rows; the odd rows join the last band
    size_t rem = h;
    for (size_t s = first; s < cap && rem > 4 * s; s *= 2) { up.push_back(s); rem -= s; }
    size_t dsum = 0;
    for (size_t t = first; t < cap && dsum + t < rem / 2; t *= 2) { down.push_back(t); dsum += t; }
    const size_t mid = rem - dsum;
    const size_t nmid = (mid + cap - 1) / cap;
    v = up;
    size_t left = mid;
    const size_t nup = v.size();
    for (size_t i = 0; i < nmid; ++i) {
        size_t b = ((mid / nmid) + 31) & ~(size_t)31;
        if (i + 1 == nmid || b > left) b = left;
        if (b) v.push_back(b);
        left -= b;
    }
    // a leftover too short to finish any output rows (rounding the middle bands up to 32 rows) joins the band before it
    if (v.size() >= nup + 2 && v.back() < min_rows) { const size_t b = v.back(); v.pop_back(); v.back() += b; }
    for (size_t i = down.size(); i-- > 0;) v.push_back(down[i]);
    if (!v.empty()) v.back() += tail; else if (tail) v.push_back(tail);
    return v;
}

// Does chaqsdf_u8 pipeline this shape, and with which band heights (middle bands / first and last band)?
// Bands are multiples of 32 rows, a middle band is at least 2*spread + 64 rows and the first at least spread + 32,
// so that every band finishes some output rows (pass 2 of rows [done, b - spread) runs once row b is there).
bool pipeline_shape(size_t w, size_t h, size_t stride, size_t spread, size_t* band_rows, size_t* first_rows) {
    if (!(spread <= 256 && w * h >= ((size_t)1 << 22))) return false;
    size_t band = ((size_t)16 << 20) / (w * stride);  // ~16 MB of input per band (measured at 16384^2: 16 MB 6.44 ms, 28 MB 6.72, 64 MB 7.14)
    if (chaq::tuning().pipe_rows > 0) band = (size_t)chaq::tuning().pipe_rows;
    band = (band + 31) & ~(size_t)31;
    if (band < 2 * spread + 64) band = (2 * spread + 64 + 31) & ~(size_t)31;
    size_t first = ((size_t)4 << 20) / (w * stride);   // ~4 MB first/last band (ramps up by doubling)
    if (chaq::tuning().pipe_first > 0) first = (size_t)chaq::tuning().pipe_first;
    first = (first + 31) & ~(size_t)31;
    if (first < spread + 32) first = (spread + 32 + 31) & ~(size_t)31;
    if (first > band) first = band;
    if (!(band > 0 && band * 2 <= h)) return false;
    *band_rows = band; *first_rows = first;
    return true;
}

int u8_pipelined(Pool& g_pool, const uint8_t* img, size_t w, size_t h, size_t stride, size_t offset, int test_above, size_t spread,
                 int asymmetric, uint8_t* out, size_t band_rows, size_t first_rows) {
    const size_t rs = chaqsdf_inter_row_bytes(w, spread);
    const size_t plan_bytes = chaqsdf_envelope_workspace_bytes(w, h, spread);
    void *d_img, *d_out, *d_inter, *d_plan;
    if (pool_get(g_pool, 0, w * h * stride, &d_img) != cudaSuccess || pool_get(g_pool, 1, w * h, &d_out) != cudaSuccess ||
        pool_get(g_pool, 2, rs * h, &d_inter) != cudaSuccess || pool_get(g_pool, 3, plan_bytes ? plan_bytes : 256, &d_plan) != cudaSuccess)
        return set_error(CHAQSDF_ERR_CUDA, "device allocation failed");
    if (!g_pool.s_in) {
        if (cudaStreamCreateWithFlags(&g_pool.s_in, cudaStreamNonBlocking) != cudaSuccess ||
            cudaStreamCreateWithFlags(&g_pool.s_out, cudaStreamNonBlocking) != cudaSuccess)
            return set_error(CHAQSDF_ERR_CUDA, "stream creation failed");
    }
    const std::vector<size_t> bands = pipeline_bands(h, band_rows, first_rows, spread + 32);
    const size_t nbands = bands.size();
    if (pool_events(g_pool, 2 * nbands) != cudaSuccess) return set_error(CHAQSDF_ERR_CUDA, "event creation failed");
    cudaStream_t s_in = g_pool.s_in, s_cmp = g_pool.stream, s_out = g_pool.s_out;
    const uint8_t* dimg = (const uint8_t*)d_img;
    uint8_t* dinter = (uint8_t*)d_inter;
    uint8_t* dout = (uint8_t*)d_out;
    size_t done = 0;  // output rows [0, done) are computed
    size_t a = 0, b = 0;
    for (size_t k = 0; k < nbands; ++k) {
        a = b; b = a + bands[k];
        cudaEvent_t ev_in = g_pool.ev[2 * k], ev_cmp = g_pool.ev[2 * k + 1];
        const size_t row_bytes = w * stride;
        if (cudaMemcpyAsync((void*)(dimg + a * row_bytes), img + a * row_bytes, (b - a) * row_bytes, cudaMemcpyHostToDevice,
                            s_in) != cudaSuccess || cudaEventRecord(ev_in, s_in) != cudaSuccess ||
            cudaStreamWaitEvent(s_cmp, ev_in, 0) != cudaSuccess)
            return drained(g_pool, CHAQSDF_ERR_CUDA, "upload failed");
        int rc = chaqsdf_rowscan_device(dimg + a * row_bytes, w, b - a, stride, offset, test_above, spread, dinter + a * rs, s_cmp);
        if (rc) return drained(g_pool, rc, nullptr);
        const size_t upto = b == h ? h : (b > spread ? b - spread : 0);
        if (upto > done) {
            size_t top = done > spread ? done - spread : 0;
            top &= ~(size_t)31;
            rc = chaqsdf_envelope_device(dinter + top * rs, w, b - top, done - top, upto - top, spread, asymmetric,
                                         dout + done * w, d_plan, plan_bytes, s_cmp);
            if (rc) return drained(g_pool, rc, nullptr);
            if (cudaEventRecord(ev_cmp, s_cmp) != cudaSuccess || cudaStreamWaitEvent(s_out, ev_cmp, 0) != cudaSuccess ||
                cudaMemcpyAsync(out + done * w, dout + done * w, (upto - done) * w, cudaMemcpyDeviceToHost, s_out) != cudaSuccess)
                return drained(g_pool, CHAQSDF_ERR_CUDA, "download failed");
            done = upto;
        }
    }
    if (cudaStreamSynchronize(s_out) != cudaSuccess || cudaStreamSynchronize(s_cmp) != cudaSuccess)
        return drained(g_pool, CHAQSDF_ERR_CUDA, "pipeline failed");
    return CHAQSDF_OK;
}

// ---- several GPUs of one box, one process (chaqsdf_u8_multi) -------------------------------------------------
// Pools of the multi-GPU entry are keyed by (device, position in the caller's device list), so that a device may
// appear more than once (every position is a band with its own buffers; this is also how a one-GPU box tests it).
std::map<std::pair<int, int>, Pool*> g_multi_pools;
std::mutex g_multi_mu;  // one multi-GPU call at a time

struct Shard {
    int dev = 0;
    size_t r0 = 0, r1 = 0, top = 0, bot = 0;  // image rows [r0, r1); halo rows held above / below them
    Pool* pool = nullptr;
    uint8_t *d_img = nullptr, *d_out = nullptr, *d_inter = nullptr;
    void* d_ws = nullptr;
    size_t ws_bytes = 0;
    cudaEvent_t ev_p1 = nullptr;
};

int multi_fail(std::vector<Shard>& sh, int prev_dev, int code, const char* msg) {
    for (auto& s : sh)
        if (s.pool && s.pool->stream) { cudaSetDevice(s.dev); cudaStreamSynchronize(s.pool->stream); }
    cudaSetDevice(prev_dev);
    return msg ? set_error(code, msg) : code;
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

size_t chaqsdf_pipeline_plan(size_t w, size_t h, size_t stride, size_t spread, size_t* bands, size_t max_bands) {
    size_t band = 0, first = 0;
    if (w == 0 || h == 0 || stride == 0 || !pipeline_shape(w, h, stride, spread, &band, &first)) return 0;
    const std::vector<size_t> v = pipeline_bands(h, band, first, spread + 32);
    for (size_t i = 0; i < v.size() && i < max_bands; ++i) bands[i] = v[i];
    return v.size();
}

int chaqsdf_u8(const uint8_t* img, size_t w, size_t h, size_t stride, size_t offset, int test_above, size_t spread,
               int asymmetric, uint8_t* out) {
    if (!img || !out) return set_error(CHAQSDF_ERR_ARG, "null pointer");
    if (stride == 0 || offset >= stride) return set_error(CHAQSDF_ERR_ARG, "bad stride/offset");
    if (w == 0 || h == 0) return set_error(CHAQSDF_ERR_ARG, "empty image");
    if (w > CHAQSDF_MAX_DIM || h > CHAQSDF_MAX_DIM) return set_error(CHAQSDF_ERR_ARG, "image side exceeds 32768");
    if (spread == 0) return set_error(CHAQSDF_ERR_ARG, "spread must be a positive integer");
    const size_t ws_bytes = chaqsdf_workspace_bytes(w, h, spread);
    if (ws_bytes == 0) return CHAQSDF_ERR_CUDA;  // error string already set by the planner
    Pool* pool = current_pool();
    if (!pool) return set_error(CHAQSDF_ERR_CUDA, "no current CUDA device");
    Pool& g_pool = *pool;
    std::lock_guard<std::mutex> lk(g_pool.mu);
    // large images on the banded path: overlap upload, kernels and download band by band
    size_t band = 0, first = 0;
    if (pipeline_shape(w, h, stride, spread, &band, &first))
        return u8_pipelined(g_pool, img, w, h, stride, offset, test_above, spread, asymmetric, out, band, first);
    void *d_img, *d_out, *d_ws;
    const size_t in_bytes = w * h * stride, out_bytes = w * h;
    if (pool_get(g_pool, 0, in_bytes, &d_img) != cudaSuccess || pool_get(g_pool, 1, out_bytes, &d_out) != cudaSuccess ||
        pool_get(g_pool, 2, ws_bytes, &d_ws) != cudaSuccess)
        return set_error(CHAQSDF_ERR_CUDA, "device allocation failed");
    cudaStream_t st = g_pool.stream;
    if (cudaMemcpyAsync(d_img, img, in_bytes, cudaMemcpyHostToDevice, st) != cudaSuccess)
        return drained(g_pool, CHAQSDF_ERR_CUDA, "upload failed");
    int rc = chaqsdf_u8_device((const uint8_t*)d_img, w, h, stride, offset, test_above, spread, asymmetric,
                               (uint8_t*)d_out, d_ws, ws_bytes, st);
    if (rc) return drained(g_pool, rc, nullptr);
    if (cudaMemcpyAsync(out, d_out, out_bytes, cudaMemcpyDeviceToHost, st) != cudaSuccess)
        return drained(g_pool, CHAQSDF_ERR_CUDA, "download failed");
    if (cudaStreamSynchronize(st) != cudaSuccess) return set_error(CHAQSDF_ERR_CUDA, "stream synchronisation failed");
    return CHAQSDF_OK;
}

// Row bands over several devices of this process (reference path: the same sdfgen.c:261-300, sharded). Pass 1 is per
// row, so each device scans its own rows; pass 2 needs min(spread, distance to the image edge) rows of the
// INTERMEDIATE from the bands above and below, which arrive by cudaMemcpyPeerAsync (NVLink when peer access exists,
// staged by the driver otherwise) as soon as the owner's pass 1 has finished (cross-device event wait, no host
// synchronisation). No torch, no NCCL: the exchange is a few MB between neighbours.
int chaqsdf_u8_multi(const uint8_t* img, size_t w, size_t h, size_t stride, size_t offset, int test_above, size_t spread,
                     int asymmetric, const int* devices, int n_devices, uint8_t* out) {
    if (!img || !out || !devices) return set_error(CHAQSDF_ERR_ARG, "null pointer");
    if (n_devices < 1 || n_devices > 64) return set_error(CHAQSDF_ERR_ARG, "device count must be 1..64");
    if (stride == 0 || offset >= stride) return set_error(CHAQSDF_ERR_ARG, "bad stride/offset");
    if (w == 0 || h == 0) return set_error(CHAQSDF_ERR_ARG, "empty image");
    if (w > CHAQSDF_MAX_DIM || h > CHAQSDF_MAX_DIM) return set_error(CHAQSDF_ERR_ARG, "image side exceeds 32768");
    if (spread == 0) return set_error(CHAQSDF_ERR_ARG, "spread must be a positive integer");
    int ndev = 0, prev_dev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || cudaGetDevice(&prev_dev) != cudaSuccess)
        return set_error(CHAQSDF_ERR_CUDA, "no CUDA device");
    for (int k = 0; k < n_devices; ++k)
        if (devices[k] < 0 || devices[k] >= ndev) return set_error(CHAQSDF_ERR_ARG, "device index out of range");
    std::lock_guard<std::mutex> lk(g_multi_mu);
    const size_t rs = chaqsdf_inter_row_bytes(w, spread);
    const size_t halo = spread < h ? spread : h;
    std::vector<Shard> sh;
    const size_t nb = (size_t)n_devices < h ? (size_t)n_devices : h;  // no empty bands
    for (size_t k = 0; k < nb; ++k) {
        Shard s;
        s.dev = devices[k];
        s.r0 = h * k / nb; s.r1 = h * (k + 1) / nb;
        s.top = halo < s.r0 ? halo : s.r0;
        s.bot = halo < h - s.r1 ? halo : h - s.r1;
        sh.push_back(s);
    }
    // buffers, streams, events
    for (size_t k = 0; k < nb; ++k) {
        Shard& s = sh[k];
        if (cudaSetDevice(s.dev) != cudaSuccess) return multi_fail(sh, prev_dev, CHAQSDF_ERR_CUDA, "cudaSetDevice failed");
        Pool*& p = g_multi_pools[std::make_pair(s.dev, (int)k)];
        if (!p) p = new Pool();
        s.pool = p;
        const size_t rows = s.r1 - s.r0, buf_rows = s.top + rows + s.bot;
        s.ws_bytes = chaqsdf_envelope_workspace_bytes(w, buf_rows, spread);
        void *a, *b, *c, *d;
        if (pool_get(*p, 0, rows * w * stride, &a) != cudaSuccess || pool_get(*p, 1, rows * w, &b) != cudaSuccess ||
            pool_get(*p, 2, buf_rows * rs, &c) != cudaSuccess || pool_get(*p, 3, s.ws_bytes ? s.ws_bytes : 256, &d) != cudaSuccess ||
            pool_events(*p, 1) != cudaSuccess)
            return multi_fail(sh, prev_dev, CHAQSDF_ERR_CUDA, "device allocation failed");
        s.d_img = (uint8_t*)a; s.d_out = (uint8_t*)b; s.d_inter = (uint8_t*)c; s.d_ws = d;
        s.ev_p1 = p->ev[0];
        // direct NVLink copies between the bands' devices where the hardware allows (failure is fine: staged copies)
        for (size_t q = 0; q < k; ++q)
            if (sh[q].dev != s.dev) {
                int can = 0;
                if (cudaDeviceCanAccessPeer(&can, s.dev, sh[q].dev) == cudaSuccess && can) {
                    cudaDeviceEnablePeerAccess(sh[q].dev, 0);
                    cudaSetDevice(sh[q].dev);
                    cudaDeviceEnablePeerAccess(s.dev, 0);
                    cudaSetDevice(s.dev);
                    cudaGetLastError();  // "already enabled" is not an error
                }
            }
    }
    // pass 1 on every band
    for (Shard& s : sh) {
        cudaSetDevice(s.dev);
        cudaStream_t st = s.pool->stream;
        const size_t rows = s.r1 - s.r0, row_bytes = w * stride;
        if (cudaMemcpyAsync(s.d_img, img + s.r0 * row_bytes, rows * row_bytes, cudaMemcpyHostToDevice, st) != cudaSuccess)
            return multi_fail(sh, prev_dev, CHAQSDF_ERR_CUDA, "upload failed");
        int rc = chaqsdf_rowscan_device(s.d_img, w, rows, stride, offset, test_above, spread, s.d_inter + s.top * rs, st);
        if (rc) return multi_fail(sh, prev_dev, rc, nullptr);
        if (cudaEventRecord(s.ev_p1, st) != cudaSuccess) return multi_fail(sh, prev_dev, CHAQSDF_ERR_CUDA, "event record failed");
    }
    // halo rows from the owners, pass 2, download
    for (size_t k = 0; k < nb; ++k) {
        Shard& s = sh[k];
        cudaSetDevice(s.dev);
        cudaStream_t st = s.pool->stream;
        const size_t lo = s.r0 - s.top, hi = s.r1 + s.bot;  // image rows this band's buffer holds
        for (size_t q = 0; q < nb; ++q) {
            if (q == k) continue;
            const Shard& o = sh[q];
            const size_t a = o.r0 > lo ? o.r0 : lo, b = o.r1 < hi ? o.r1 : hi;  // rows owned by q that k needs
            if (a >= b) continue;
            if (cudaStreamWaitEvent(st, o.ev_p1, 0) != cudaSuccess ||
                cudaMemcpyPeerAsync(s.d_inter + (a - lo) * rs, s.dev, o.d_inter + (o.top + a - o.r0) * rs, o.dev, (b - a) * rs,
                                    st) != cudaSuccess)
                return multi_fail(sh, prev_dev, CHAQSDF_ERR_CUDA, "halo copy failed");
        }
        const size_t rows = s.r1 - s.r0;
        int rc = chaqsdf_envelope_device(s.d_inter, w, s.top + rows + s.bot, s.top, s.top + rows, spread, asymmetric, s.d_out,
                                         s.d_ws, s.ws_bytes, st);
        if (rc) return multi_fail(sh, prev_dev, rc, nullptr);
        if (cudaMemcpyAsync(out + s.r0 * w, s.d_out, rows * w, cudaMemcpyDeviceToHost, st) != cudaSuccess)
            return multi_fail(sh, prev_dev, CHAQSDF_ERR_CUDA, "download failed");
    }
    int rc = CHAQSDF_OK;
    for (Shard& s : sh) {
        cudaSetDevice(s.dev);
        if (cudaStreamSynchronize(s.pool->stream) != cudaSuccess) rc = set_error(CHAQSDF_ERR_CUDA, "multi-GPU pipeline failed");
    }
    cudaSetDevice(prev_dev);
    return rc;
}

// General heights (or CHAQSDF_DT2D=stock): the reference's float32 envelope, one thread per line (dt_float.cuh).
// d_f holds the image on the device; on success it holds the result.
static int dt2d_float_device(Pool& g_pool, float* d_f, size_t w, size_t h, cudaStream_t st) {
    const size_t n = w * h;
    void *d_tmp, *d_v, *d_z;
    if (pool_get(g_pool, 1, n * sizeof(float), &d_tmp) != cudaSuccess || pool_get(g_pool, 2, n * sizeof(int), &d_v) != cudaSuccess ||
        pool_get(g_pool, 3, n * sizeof(float), &d_z) != cudaSuccess)
        return set_error(CHAQSDF_ERR_CUDA, "device allocation failed");
    chaq::DtFloatParams p;
    p.v = (int*)d_v; p.z = (float*)d_z;
    p.in = d_f; p.out = (float*)d_tmp; p.len = (int)w; p.nlines = (int)h; p.do_sqrt = 0;   // df.c:130
    chaq::dt_float_kernel<<<(unsigned)((h + 127) / 128), 128, 0, st>>>(p);
    if (cudaGetLastError() != cudaSuccess) return set_error(CHAQSDF_ERR_CUDA, "dt_float_kernel launch failed");
    chaq::note_launch();
    p.in = (const float*)d_tmp; p.out = d_f; p.len = (int)h; p.nlines = (int)w; p.do_sqrt = 1;  // df.c:133
    chaq::dt_float_kernel<<<(unsigned)((w + 127) / 128), 128, 0, st>>>(p);
    if (cudaGetLastError() != cudaSuccess) return set_error(CHAQSDF_ERR_CUDA, "dt_float_kernel launch failed");
    chaq::note_launch();
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
    Pool* pool = current_pool();
    if (!pool) { g_dt_status = CHAQSDF_ERR_CUDA; return; }
    Pool& g_pool = *pool;
    std::lock_guard<std::mutex> lk(g_pool.mu);
    void *d_f, *d_mask, *d_ws;
    const size_t n = w * h;
    g_dt_status = CHAQSDF_ERR_CUDA;
    if (pool_get(g_pool, 0, n * sizeof(float), &d_f) != cudaSuccess || pool_get(g_pool, 1, n, &d_mask) != cudaSuccess ||
        pool_get(g_pool, 2, ws_bytes, &d_ws) != cudaSuccess)
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
    if (flag || chaq::tuning().dt2d_mode == 1) {  // sampled function, or stock-compatible float32 output asked for
        if (w > CHAQSDF_MAX_DIM || h > CHAQSDF_MAX_DIM) { g_dt_status = set_error(CHAQSDF_ERR_ARG, "image side exceeds 32768"); return; }
        if (n == 1) { g_dt_status = CHAQSDF_OK; return; }  // df.c:32-36: a single cell is returned as it is
        int rc = dt2d_float_device(g_pool, (float*)d_f, w, h, st);
        if (rc) { g_dt_status = rc; return; }
        if (cudaMemcpyAsync(img, d_f, n * sizeof(float), cudaMemcpyDeviceToHost, st) != cudaSuccess) return;
        if (cudaStreamSynchronize(st) != cudaSuccess) return;
        g_dt_status = CHAQSDF_OK;
        return;
    }
    // seeds are class 1; f_in = distance to the nearest class-1 pixel (0 on the seeds themselves)
    int rc = chaqsdf_fields_device((const uint8_t*)d_mask, w, h, 1, 0, 1, 1, 0, nullptr, nullptr, (float*)d_f, nullptr,
                                   nullptr, nullptr, d_ws, ws_bytes, st);
    if (rc) { g_dt_status = rc; return; }
    if (cudaMemcpyAsync(img, d_f, n * sizeof(float), cudaMemcpyDeviceToHost, st) != cudaSuccess) return;
    if (cudaStreamSynchronize(st) != cudaSuccess) return;
    g_dt_status = CHAQSDF_OK;
}

}  // extern "C"
