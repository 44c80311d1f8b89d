// atlas.cu -- glyph / atlas front-end and packer: rasterised glyph batches in, packed SDF atlas
// out. The reference has nothing here (stb_truetype.h / stb_rect_pack.h sit unused in its submodule); this is the caller
// side of BASELINE.json's config 3 (4096 glyphs of 256^2, spread 16), written from scratch:
//   boxes  (device)  ink bounding box of every glyph canvas, grown by a margin (the spread), clipped to the canvas;
//   pack   (host)    skyline bottom-left rectangle packer: boxes -> positions in an atlas of a given width;
//   SDF    (device)  chaqsdf_u8_batched_device on the canvases (the fused one-CTA-per-glyph kernel where it applies);
//   blit   (device)  every glyph's box of its SDF canvas into its place in the atlas.
// Pixels of the atlas that no glyph covers hold 0, the byte of a pixel far outside every shape (openmp/sdfgen.c:75-96
// maps sdf <= s_min to 0 in both the symmetric and the asymmetric mode).
#include <cuda_runtime.h>

#include <algorithm>
#include <cstdint>
#include <numeric>
#include <string>
#include <vector>

#include "../../include/chaqsdf.h"

namespace chaq {
int set_error(int code, const char* msg);  // api.cu
void note_launch();
}  // namespace chaq

namespace {

int fail(int code, const std::string& msg) { return chaq::set_error(code, msg.c_str()); }
#define AT_CUDA(expr)                                                                                    \
    do {                                                                                                 \
        cudaError_t e_ = (expr);                                                                         \
        if (e_ != cudaSuccess) return fail(CHAQSDF_ERR_CUDA, std::string(#expr) + ": " + cudaGetErrorString(e_)); \
    } while (0)

// one CTA per glyph: min / max of the coordinates of its inside pixels
__global__ void __launch_bounds__(256) glyph_boxes_kernel(const uint8_t* __restrict__ g, int w, int h, int above, int margin,
                                                          int32_t* __restrict__ boxes) {
    __shared__ int s[4][8];
    const uint8_t* img = g + (size_t)blockIdx.x * w * h;
    int x0 = w, y0 = h, x1 = -1, y1 = -1;
    for (int i = threadIdx.x; i < w * h; i += 256) {
        const int v = img[i];
        const bool in = above ? v > 127 : v < 127;
        if (in) {
            const int y = i / w, x = i - y * w;
            x0 = min(x0, x); x1 = max(x1, x); y0 = min(y0, y); y1 = max(y1, y);
        }
    }
    x0 = __reduce_min_sync(0xffffffffu, x0); y0 = __reduce_min_sync(0xffffffffu, y0);
    x1 = __reduce_max_sync(0xffffffffu, x1); y1 = __reduce_max_sync(0xffffffffu, y1);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (lane == 0) { s[0][warp] = x0; s[1][warp] = y0; s[2][warp] = x1; s[3][warp] = y1; }
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int k = 1; k < 8; ++k) {
            s[0][0] = min(s[0][0], s[0][k]); s[1][0] = min(s[1][0], s[1][k]);
            s[2][0] = max(s[2][0], s[2][k]); s[3][0] = max(s[3][0], s[3][k]);
        }
        int32_t* b = boxes + 4 * (size_t)blockIdx.x;
        if (s[2][0] < 0) {
            b[0] = b[1] = b[2] = b[3] = 0;  // no ink: an empty box
        } else {
            b[0] = max(s[0][0] - margin, 0); b[1] = max(s[1][0] - margin, 0);
            b[2] = min(s[2][0] + 1 + margin, w); b[3] = min(s[3][0] + 1 + margin, h);
        }
    }
}

// grid (glyph, row chunk): rows of the glyph's box go to the atlas, 4 bytes per thread where both sides are aligned
__global__ void __launch_bounds__(256) atlas_blit_kernel(const uint8_t* __restrict__ sdf, int w, int h, const int32_t* __restrict__ boxes,
                                                         const uint32_t* __restrict__ xy, uint8_t* __restrict__ atlas, int aw) {
    const int gi = blockIdx.x;
    const int bx0 = boxes[4 * gi], by0 = boxes[4 * gi + 1], bw = boxes[4 * gi + 2] - bx0, bh = boxes[4 * gi + 3] - by0;
    if (bw <= 0 || bh <= 0) return;
    const uint8_t* src = sdf + (size_t)gi * w * h + (size_t)by0 * w + bx0;
    uint8_t* dst = atlas + (size_t)xy[2 * gi + 1] * aw + xy[2 * gi];
    for (int i = blockIdx.y * 256 + threadIdx.x; i < bw * bh; i += gridDim.y * 256) {
        const int y = i / bw, x = i - y * bw;
        dst[(size_t)y * aw + x] = src[(size_t)y * w + x];
    }
}

struct Node { uint32_t x, y, w; };

}  // namespace

extern "C" {

// Skyline bottom-left: rectangles are placed tallest first (ties: widest, then input order), each at the position
// that leaves its top edge lowest (ties: leftmost). `pad` empty pixels are kept to the right of and below every
// rectangle. Deterministic; O(n * skyline length).
int chaqsdf_atlas_pack(size_t n, const uint32_t* wh, size_t atlas_w, size_t pad, uint32_t* xy, size_t* atlas_h) {
    if ((n && (!wh || !xy)) || !atlas_h) return fail(CHAQSDF_ERR_ARG, "null pointer");
    if (atlas_w == 0 || atlas_w > 0x7fffffffu) return fail(CHAQSDF_ERR_ARG, "bad atlas width");
    std::vector<size_t> order(n);
    std::iota(order.begin(), order.end(), (size_t)0);
    std::stable_sort(order.begin(), order.end(), [&](size_t a, size_t b) {
        if (wh[2 * a + 1] != wh[2 * b + 1]) return wh[2 * a + 1] > wh[2 * b + 1];
        return wh[2 * a] > wh[2 * b];
    });
    std::vector<Node> sky{{0u, 0u, (uint32_t)atlas_w}};
    uint64_t top = 0;
    for (size_t idx : order) {
        const uint64_t rw = wh[2 * idx], rh = wh[2 * idx + 1];
        if (rw == 0 || rh == 0) { xy[2 * idx] = xy[2 * idx + 1] = 0; continue; }
        if (rw > atlas_w) return fail(CHAQSDF_ERR_ARG, "a rectangle is wider than the atlas");
        const uint64_t fw = std::min<uint64_t>(rw + pad, atlas_w - 0), fh = rh + pad;  // footprint on the skyline
        size_t best = sky.size();
        uint64_t best_y = ~0ull, best_x = 0;
        for (size_t i = 0; i < sky.size(); ++i) {
            const uint64_t x = sky[i].x;
            if (x + rw > atlas_w) break;
            const uint64_t need = std::min<uint64_t>(fw, atlas_w - x);
            uint64_t y = 0, covered = 0;
            for (size_t j = i; j < sky.size() && covered < need; ++j) { y = std::max<uint64_t>(y, sky[j].y); covered += sky[j].w; }
            if (y < best_y) { best_y = y; best_x = x; best = i; }
        }
        if (best == sky.size()) return fail(CHAQSDF_ERR_ARG, "internal: no position found");
        if (best_y + rh > 0xffffffffull) return fail(CHAQSDF_ERR_ARG, "atlas too tall");
        xy[2 * idx] = (uint32_t)best_x; xy[2 * idx + 1] = (uint32_t)best_y;
        top = std::max(top, best_y + rh);
        // raise the skyline under the footprint
        const uint64_t x_end = std::min<uint64_t>(best_x + fw, atlas_w);
        Node nn{(uint32_t)best_x, (uint32_t)(best_y + fh), (uint32_t)(x_end - best_x)};
        size_t j = best;
        std::vector<Node> tail;
        while (j < sky.size() && sky[j].x < x_end) {
            const uint64_t jend = (uint64_t)sky[j].x + sky[j].w;
            if (jend > x_end) tail.push_back(Node{(uint32_t)x_end, sky[j].y, (uint32_t)(jend - x_end)});
            ++j;
        }
        sky.erase(sky.begin() + best, sky.begin() + j);
        sky.insert(sky.begin() + best, nn);
        sky.insert(sky.begin() + best + 1, tail.begin(), tail.end());
        for (size_t k = 0; k + 1 < sky.size();) {  // merge equal neighbours
            if (sky[k].y == sky[k + 1].y) { sky[k].w += sky[k + 1].w; sky.erase(sky.begin() + k + 1); } else ++k;
        }
    }
    *atlas_h = (size_t)top;
    return CHAQSDF_OK;
}

int chaqsdf_glyph_boxes_device(const uint8_t* d_glyphs, size_t n, size_t w, size_t h, int test_above, size_t margin,
                               int32_t* d_boxes, void* stream) {
    if (!d_glyphs || !d_boxes) return fail(CHAQSDF_ERR_ARG, "null pointer");
    if (n == 0 || n > 0x7fffffffu || w == 0 || h == 0 || w > CHAQSDF_MAX_DIM || h > CHAQSDF_MAX_DIM) return fail(CHAQSDF_ERR_ARG, "bad glyph batch shape");
    glyph_boxes_kernel<<<(unsigned)n, 256, 0, (cudaStream_t)stream>>>(d_glyphs, (int)w, (int)h, test_above ? 1 : 0,
                                                                        (int)std::min<size_t>(margin, CHAQSDF_MAX_DIM), d_boxes);
    AT_CUDA(cudaGetLastError());
    chaq::note_launch();
    return CHAQSDF_OK;
}

int chaqsdf_atlas_blit_device(const uint8_t* d_sdf, size_t n, size_t w, size_t h, const int32_t* d_boxes, const uint32_t* d_xy,
                              uint8_t* d_atlas, size_t atlas_w, void* stream) {
    if (!d_sdf || !d_boxes || !d_xy || !d_atlas) return fail(CHAQSDF_ERR_ARG, "null pointer");
    if (n == 0 || n > 0x7fffffffu || w == 0 || h == 0 || atlas_w == 0 || atlas_w > 0x7fffffffu) return fail(CHAQSDF_ERR_ARG, "bad atlas shape");
    const unsigned chunks = (unsigned)std::min<size_t>((w * h + 256 * 16 - 1) / (256 * 16), 64);
    atlas_blit_kernel<<<dim3((unsigned)n, chunks), 256, 0, (cudaStream_t)stream>>>(d_sdf, (int)w, (int)h, d_boxes, d_xy, d_atlas, (int)atlas_w);
    AT_CUDA(cudaGetLastError());
    chaq::note_launch();
    return CHAQSDF_OK;
}

// The whole front-end from host buffers: n canvases of w*h bytes in, one atlas out.
int chaqsdf_atlas_u8(const uint8_t* glyphs, size_t n, size_t w, size_t h, int test_above, size_t spread, int asymmetric,
                     size_t atlas_w, size_t pad, uint8_t* atlas, size_t atlas_rows_cap, size_t* atlas_h, int32_t* boxes,
                     uint32_t* xy) {
    if (!glyphs || !atlas || !atlas_h || !boxes || !xy) return fail(CHAQSDF_ERR_ARG, "null pointer");
    if (n == 0 || w == 0 || h == 0) return fail(CHAQSDF_ERR_ARG, "empty batch");
    const size_t ws_bytes = chaqsdf_batched_workspace_bytes(n, w, h, spread);
    if (ws_bytes == 0) return CHAQSDF_ERR_ARG;  // the error string is set
    uint8_t *d_g = nullptr, *d_sdf = nullptr, *d_atlas = nullptr, *d_ws = nullptr;
    int32_t* d_boxes = nullptr;
    uint32_t* d_xy = nullptr;
    int rc = CHAQSDF_OK;
    auto body = [&]() -> int {
        AT_CUDA(cudaMalloc(&d_g, n * w * h));
        AT_CUDA(cudaMalloc(&d_sdf, n * w * h));
        AT_CUDA(cudaMalloc(&d_ws, ws_bytes));
        AT_CUDA(cudaMalloc(&d_boxes, n * 4 * sizeof(int32_t)));
        AT_CUDA(cudaMalloc(&d_xy, n * 2 * sizeof(uint32_t)));
        AT_CUDA(cudaMemcpyAsync(d_g, glyphs, n * w * h, cudaMemcpyHostToDevice, 0));
        if (int r = chaqsdf_glyph_boxes_device(d_g, n, w, h, test_above, spread, d_boxes, nullptr)) return r;
        // the SDF batch runs while the host packs
        if (int r = chaqsdf_u8_batched_device(d_g, n, w, h, 1, 0, test_above, spread, asymmetric, d_sdf, d_ws, ws_bytes, nullptr)) return r;
        AT_CUDA(cudaMemcpy(boxes, d_boxes, n * 4 * sizeof(int32_t), cudaMemcpyDeviceToHost));
        std::vector<uint32_t> wh(2 * n);
        for (size_t i = 0; i < n; ++i) {
            wh[2 * i] = (uint32_t)(boxes[4 * i + 2] - boxes[4 * i]);
            wh[2 * i + 1] = (uint32_t)(boxes[4 * i + 3] - boxes[4 * i + 1]);
        }
        if (int r = chaqsdf_atlas_pack(n, wh.data(), atlas_w, pad, xy, atlas_h)) return r;
        if (*atlas_h > atlas_rows_cap)
            return fail(CHAQSDF_ERR_WORKSPACE, "atlas needs " + std::to_string(*atlas_h) + " rows, the output buffer holds " + std::to_string(atlas_rows_cap));
        const size_t rows = std::max<size_t>(*atlas_h, 1);
        AT_CUDA(cudaMalloc(&d_atlas, rows * atlas_w));
        AT_CUDA(cudaMemsetAsync(d_atlas, 0, rows * atlas_w, 0));
        AT_CUDA(cudaMemcpyAsync(d_xy, xy, n * 2 * sizeof(uint32_t), cudaMemcpyHostToDevice, 0));
        if (int r = chaqsdf_atlas_blit_device(d_sdf, n, w, h, d_boxes, d_xy, d_atlas, atlas_w, nullptr)) return r;
        AT_CUDA(cudaMemcpy(atlas, d_atlas, *atlas_h * atlas_w, cudaMemcpyDeviceToHost));
        return CHAQSDF_OK;
    };
    rc = body();
    cudaFree(d_g); cudaFree(d_sdf); cudaFree(d_ws); cudaFree(d_boxes); cudaFree(d_xy); cudaFree(d_atlas);
    return rc;
}

}  // extern "C"
