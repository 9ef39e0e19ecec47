// stitch_exact.cuh -- the exact per-pixel kernel: every canvas pixel evaluates the cameras from
// the highest index down in fp64 and gathers straight from global memory.  It is the path for
// seam / edge / hole tiles and the fallback when buffers do not meet the staged path's
// alignment rules.  Replaces the loop body of /root/reference/tools/stitch/main.cpp:145-152.
#pragma once

#include "stitch_common.cuh"

namespace bs {

constexpr int kExactTileW = 128;   // 32 threads x 4 pixels
constexpr int kExactTileH = 8;

// Computes and stores 4 horizontally consecutive canvas pixels starting at (x0, y).
__device__ __forceinline__ void exact_quad(const FrameArgs &A, int x0, int y)
{
    if (y >= A.H || x0 >= A.W) return;
    uint32_t px[4];
    bool cov[4];
    const int n = min(4, A.W - x0);
    bool all = (n == 4);
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        px[k] = 0;
        cov[k] = false;
        if (k < n) cov[k] = composite_exact(A, x0 + k, y, px[k]);
        if (A.keep_canvas && !cov[k]) all = false;
    }
    uint8_t *row = A.canvas + (long long)y * A.cpitch + (long long)x0 * 3;
    if (all && ((reinterpret_cast<uintptr_t>(row) & 3) == 0)) {
        uint32_t *o = reinterpret_cast<uint32_t *>(row);
        o[0] = px[0] | (px[1] << 24);
        o[1] = (px[1] >> 8) | (px[2] << 16);
        o[2] = (px[2] >> 16) | (px[3] << 8);
    } else {
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            if (k < n && (cov[k] || !A.keep_canvas)) {
                row[3 * k + 0] = (uint8_t)(px[k]);
                row[3 * k + 1] = (uint8_t)(px[k] >> 8);
                row[3 * k + 2] = (uint8_t)(px[k] >> 16);
            }
        }
    }
}

// Whole-canvas launch: grid = (ceil(W/128), ceil(H/8)), block = (32, 8).
__global__ void __launch_bounds__(256) stitch_exact_kernel(const __grid_constant__ FrameArgs A)
{
    const int x0 = (blockIdx.x * 32 + threadIdx.x) * 4;
    const int y = blockIdx.y * kExactTileH + threadIdx.y;
    exact_quad(A, x0, y);
}

}  // namespace bs
