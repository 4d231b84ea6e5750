// Pieces shared by the forward and backward blend kernels.
//
// Work decomposition: every 16x16 tile is cut into eight 8x4 pixel rectangles; one warp owns one rectangle (one lane
// per pixel) and walks, on its own (no block barrier), the RECT LIST of that rectangle: the positions, in the tile's
// depth-sorted list, of the Gaussians whose conservative radius h reaches the rectangle. The lists are built once per
// forward by rect_lists_kernel and used by both the forward and the backward blend (before, each of the 8 warps of a
// tile re-derived its list by scanning the whole tile list in both passes).
//
// Walking a list:
//   gather  - 32 entries at a time, every lane fetches one entry's record into a per-warp shared-memory ring
//             ({x, y, position, -tau'} through registers, the other 32 bytes with cp.async);
//   scan    - for each ring entry every lane evaluates only the quadratic form `power` (bit-identical to the reference
//             expression) and appends the entry to ITS OWN list iff power >= -tau', i.e. iff alpha >= 1/255 is possible
//             for its pixel (tau' = ln(255*opacity) with a safety margin): a pair rejected here is one the reference
//             skips at forward.cu:340 / backward.cu:499;
//   blend   - listed (slot, power) pairs sit in a per-lane queue of 8 entries in shared memory; when a lane's queue is
//             full, or the ring is about to be overwritten, all lanes blend their queued entries in lockstep, in list
//             order. (Depth order correlates with screen position, so at any moment only a few pixels of the rectangle
//             receive entries and the lockstep blend runs at ~20 % lane utilisation. A variant with 64-entry per-lane
//             lists blended once at the end reaches ~78 % but has to re-fetch every record per lane from L1/L2 and
//             measured 40 % slower; see DESIGN.md.)
#pragma once
#include "common.cuh"

namespace das3r {

constexpr int BLEND_WARPS = 2;   // warps per block
constexpr int CHUNK = 32;        // list entries gathered and scanned per round (one per lane)
constexpr int RING = 2 * CHUNK;  // per-warp ring of gathered records: the current chunk and the previous one
constexpr int QCAP = 8;          // per-lane queue capacity (ring slots waiting for the blend step)

struct RingRec {
    float4 A;  // {x, y, tile-list position (bits), -tau'}
    float4 B;  // {conic a, b, c, opacity}
    float4 C;  // {r, g, b, gaussian index (bits)}
};

struct BlendSmem {
    RingRec ring[BLEND_WARPS][RING];
    float qpower[BLEND_WARPS][QCAP * 32];  // [entry][lane]: the pair's `power`, evaluated once in the scan
    uint8_t qslot[BLEND_WARPS][QCAP * 32];  // [entry][lane]: ring slot of the queued entry
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

// Which tile / pixel rectangle a warp works on: 8 / BLEND_WARPS blocks per tile.
struct WarpRect {
    int tile, rect, wx0, wy0, px, py;
    bool inside;
};
__device__ __forceinline__ WarpRect warp_rect(int W, int H, unsigned gx) {
    WarpRect r;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    r.tile = blockIdx.x / (8 / BLEND_WARPS);
    r.rect = (blockIdx.x % (8 / BLEND_WARPS)) * BLEND_WARPS + warp;  // 0..7
    const int tx = r.tile % gx, ty = r.tile / gx;
    r.wx0 = tx * TILE_X + (r.rect & 1) * 8;
    r.wy0 = ty * TILE_Y + (r.rect >> 1) * 4;
    r.px = r.wx0 + (lane & 7);
    r.py = r.wy0 + (lane >> 3);
    r.inside = r.px < W && r.py < H;
    return r;
}

// Does the Gaussian with tile-ordered record head A = {x, y, h, -tau'} reach rectangle `rect` of tile (tx, ty)?
__device__ __forceinline__ bool reaches_rect(const float4& A, int tx, int ty, int rect) {
    const float rx0 = (float)(tx * TILE_X + (rect & 1) * 8), ry0 = (float)(ty * TILE_Y + (rect >> 1) * 4);
    return (A.z >= 0.f) && (A.x + A.z >= rx0) && (A.x - A.z <= rx0 + 7.f) && (A.y + A.z >= ry0) &&
           (A.y - A.z <= ry0 + 3.f);
}

// Walks list[0..cnt) (REVERSE: from the end) as described at the top of this file.
//   keep(pos)           per-lane extra condition for queueing tile-list position `pos` (forward: pixel not finished;
//                       backward: pos < the pixel's last contributor)
//   blend(rec, power)   per-lane blend step for one queued record; `power` is the value the scan computed
//   finished()          warp-uniform early exit test, evaluated after every chunk
//   pos_limit           no lane needs list positions >= pos_limit: chunks entirely beyond it are skipped (backward)
template <bool REVERSE, typename Keep, typename Blend, typename Finished>
__device__ __forceinline__ void walk_rect_list(RingRec* __restrict__ ring, float* __restrict__ qpower,
                                               uint8_t* __restrict__ qslot, const uint32_t* __restrict__ list, int cnt,
                                               const float4* __restrict__ gA, const float4* __restrict__ gBC,
                                               const float2 pixf, const uint32_t pos_limit, Keep keep, Blend blend,
                                               Finished finished) {
    const int lane = threadIdx.x & 31;
    uint32_t qn = 0;          // this lane's queue length * 32 (offset of its next entry row)
    bool prev_live = false;   // may the previous chunk's half of the ring still be referenced by a queue?

    auto drain = [&]() {
        const uint32_t qmax = __reduce_max_sync(0xffffffffu, qn);
        for (uint32_t i = 0; i < qmax; i += 32) {
            if (i < qn) blend(ring[qslot[i + lane]], qpower[i + lane]);
        }
        qn = 0;
        prev_live = false;
    };

    const int nchunks = (cnt + CHUNK - 1) / CHUNK;
    if (nchunks == 0) return;
    int c = REVERSE ? nchunks - 1 : 0;
    const int step = REVERSE ? -1 : 1;
    // software pipeline: list entries two chunks ahead, record heads one chunk ahead
    uint32_t li = (c * CHUNK + lane < cnt) ? list[c * CHUNK + lane] : 0u;
    float4 a = gA[li];
    int cn = c + step;
    uint32_t li_n = (cn >= 0 && cn < nchunks && cn * CHUNK + lane < cnt) ? list[cn * CHUNK + lane] : 0u;
    uint32_t half = 0;  // which half of the ring the current chunk uses
    for (int it = 0; it < nchunks; it++, c += step) {
        const uint32_t li_c = li;
        const float4 a_c = a;
        li = li_n;
        if (it + 1 < nchunks) a = gA[li];
        cn = c + 2 * step;
        li_n = (cn >= 0 && cn < nchunks && cn * CHUNK + lane < cnt) ? list[cn * CHUNK + lane] : 0u;

        const int ns = min(CHUNK, cnt - c * CHUNK);
        if (__all_sync(0xffffffffu, lane >= ns || li_c >= pos_limit)) continue;  // nothing in this chunk is needed
        RingRec* cur = ring + half * CHUNK;
        if (lane < ns) {
            // scan order = list order (REVERSE: descending positions)
            RingRec& r = cur[REVERSE ? ns - 1 - lane : lane];
            r.A = make_float4(a_c.x, a_c.y, __uint_as_float(li_c), a_c.w);
            cp_async_16(&r.B, gBC + 2 * (size_t)li_c);  // LDGSTS: global -> ring, no register round trip
            cp_async_16(&r.C, gBC + 2 * (size_t)li_c + 1);
        }
        cp_async_commit();
        cp_async_wait_all();
        __syncwarp();
        const uint32_t slot0 = half * CHUNK;
        for (int s = 0; s < ns; s += 2) {
            // up to two entries are queued per lane and round: keep two free slots in every lane's queue
            if (__reduce_max_sync(0xffffffffu, qn) > (uint32_t)(QCAP - 2) * 32u) drain();
#pragma unroll
            for (int u = 0; u < 2; u++) {
                if (s + u < ns) {
                    const float4 A = cur[s + u].A;
                    const float4 con_o = cur[s + u].B;
                    const float2 d = {A.x - pixf.x, A.y - pixf.y};
                    const float power = splat_power(con_o, d);
                    if (!(power > 0.0f) && power >= A.w && keep(__float_as_uint(A.z))) {
                        qslot[qn + lane] = (uint8_t)(slot0 + s + u);
                        qpower[qn + lane] = power;
                        qn += 32;
                    }
                }
            }
        }
        if (finished()) break;
        // The next chunk overwrites the other half of the ring, i.e. the previous chunk's records: entries of that
        // chunk that are still queued have to be blended first. Entries queued during this chunk may wait one more
        // round (drain() empties every queue, so after it nothing of either chunk is pending).
        if (prev_live) {
            drain();
        } else {
            prev_live = true;
        }
        half ^= 1u;
        __syncwarp();
    }
    drain();
}

}  // namespace das3r
