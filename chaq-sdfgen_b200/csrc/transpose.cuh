// transpose.cuh -- byte-matrix transposition, for the transposed row-band sharding of big images over several GPUs
// (DESIGN.md section 6): a rank that owns a band of ROWS runs the two passes on the band's transpose, so that the
// expensive envelope pass (reference openmp/df.c:57-96) runs along the direction that is whole inside the band and
// only the cheap scan (the first dist_transform_axis, df.c:130) looks across the band's edges.
// One CTA moves a 64 x 64 tile through shared memory: 16-byte loads along the source rows, 16-byte stores along the
// destination rows; tiles at the ragged edges and unaligned operands go byte by byte.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace chaq {

constexpr int kTposeTile = 64;
constexpr int kTposeThreads = 256;

struct TposeParams {
    const uint8_t* src;
    uint8_t* dst;
    long long src_pitch, dst_pitch;
    int rows, cols;          // of the source
    bool vec_in, vec_out;    // 16-byte accesses are aligned on that side
};

__global__ void __launch_bounds__(kTposeThreads) transpose_u8_kernel(TposeParams p) {
    __shared__ __align__(16) uint8_t tile[kTposeTile][kTposeTile + 16];  // [source row][source column]
    const int r0 = blockIdx.y * kTposeTile, c0 = blockIdx.x * kTposeTile;
    const int t = threadIdx.x;
    const bool full = r0 + kTposeTile <= p.rows && c0 + kTposeTile <= p.cols;
    {
        const int r = t >> 2, q = (t & 3) * 16;  // source row of the tile, 16-byte piece
        if (full && p.vec_in) {
            const uint4 v = *reinterpret_cast<const uint4*>(p.src + (long long)(r0 + r) * p.src_pitch + c0 + q);
            *reinterpret_cast<uint4*>(&tile[r][q]) = v;
        } else if (r0 + r < p.rows) {
            for (int k = 0; k < 16; ++k)
                if (c0 + q + k < p.cols) tile[r][q + k] = p.src[(long long)(r0 + r) * p.src_pitch + c0 + q + k];
        }
    }
    __syncthreads();
    {
        const int c = t & 63, q = (t >> 6) * 16;  // destination row of the tile (= source column), 16-byte piece
        if (full && p.vec_out) {
            uint32_t w[4];
#pragma unroll
            for (int i = 0; i < 4; ++i)
                w[i] = (uint32_t)tile[q + 4 * i][c] | ((uint32_t)tile[q + 4 * i + 1][c] << 8) |
                       ((uint32_t)tile[q + 4 * i + 2][c] << 16) | ((uint32_t)tile[q + 4 * i + 3][c] << 24);
            *reinterpret_cast<uint4*>(p.dst + (long long)(c0 + c) * p.dst_pitch + r0 + q) = make_uint4(w[0], w[1], w[2], w[3]);
        } else if (c0 + c < p.cols) {
            for (int k = 0; k < 16; ++k)
                if (r0 + q + k < p.rows) p.dst[(long long)(c0 + c) * p.dst_pitch + r0 + q + k] = tile[q + k][c];
        }
    }
}

}  // namespace chaq
