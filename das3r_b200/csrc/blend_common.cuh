// Pieces shared by the forward and backward blend kernels (warp-autonomous streaming over a tile's depth-sorted list).
#pragma once
#include "common.cuh"

namespace das3r {

constexpr int BLEND_WARPS = 2;   // warps per block; every warp owns one 8x4 pixel rectangle of one tile
constexpr int RING = 64;         // per-warp shared-memory ring of survivor records
constexpr int QCAP = 8;          // per-lane queue of ring slots waiting for the blend step (8 bits each)

struct RingRec {
    float4 A;  // {x, y, list position (bits), -tau'}
    float4 B;  // {conic a, b, c, opacity}
    float4 C;  // {r, g, b, gaussian index (bits)}
};

// power = -0.5f*(c.x*d.x*d.x + c.z*d.y*d.y) - c.y*d.x*d.y with the contraction ptxas chose for the reference's
// renderCUDA (sm_100 build, CUDA 12.9): the x term is fused, the z term and the cross term are rounded products.
// Pinned with intrinsics so forward, backward and the reference take the same `alpha < 1/255` decisions.
__device__ __forceinline__ float splat_power(const float4& con_o, const float2& d) {
    const float zz = __fmul_rn(__fmul_rn(con_o.z, d.y), d.y);
    const float s = __fmaf_rn(__fmul_rn(con_o.x, d.x), d.x, zz);
    const float u = __fmul_rn(__fmul_rn(con_o.y, d.x), d.y);
    return __fmaf_rn(s, -0.5f, -u);
}

// Which tile / pixel rectangle a warp works on: 4 blocks of 2 warps per tile.
struct WarpRect {
    int tile, wx0, wy0, px, py;
    bool inside;
};
__device__ __forceinline__ WarpRect warp_rect(int W, int H, unsigned gx) {
    WarpRect r;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    r.tile = blockIdx.x / (8 / BLEND_WARPS);
    const int rect = (blockIdx.x % (8 / BLEND_WARPS)) * BLEND_WARPS + warp;  // 0..7
    const int tx = r.tile % gx, ty = r.tile / gx;
    r.wx0 = tx * TILE_X + (rect & 1) * 8;
    r.wy0 = ty * TILE_Y + (rect >> 1) * 4;
    r.px = r.wx0 + (lane & 7);
    r.py = r.wy0 + (lane >> 3);
    r.inside = r.px < W && r.py < H;
    return r;
}

}  // namespace das3r
