// Pieces shared by the forward and backward blend kernels.
//
// Work decomposition: every 16x16 tile is cut into eight 8x4 pixel rectangles; one warp owns one rectangle (one lane
// per pixel) and walks, on its own (no block barrier), the RECT LIST of that rectangle. rect_lists_kernel builds the
// lists once per forward; both blend passes use them. A list entry is {position in the tile's depth-sorted list,
// 32-bit PIXEL MASK}: bit p is set iff the pair (entry, pixel p of the rectangle) can reach alpha >= 1/255, decided with
// the blend's own fp32 `power` expression (power <= 0 and power >= -tau', tau' = ln(255*opacity) plus a margin). A pair
// whose bit is clear is one the reference skips at forward.cu:340 / backward.cu:499; entries with an empty mask are not
// listed at all.
//
// Walking a list (walk_rect_list):
//   gather   - 32 entries at a time, every lane starts the copy of one entry's 48-byte record into a per-warp
//              shared-memory ring of K chunks with cp.async (LDGSTS, no register round trip) and the warp transposes
//              the 32 pixel masks (5 shuffle stages), so that every lane ends up with the bitmask of the chunk's
//              entries that touch ITS pixel; the masks go to a small per-lane table in shared memory;
//   blend    - every lane pops the lowest set bit of its current mask and blends that ring record; lanes move on to
//              their next chunk on their own, so a pixel with many contributions in one chunk does not hold the others
//              back as long as the ring (a sliding window of K chunks) still has records for them. One chunk is always
//              in flight while the others are blended.
// History: per-pixel evaluation of `power` for every ring entry inside the blend kernels plus an
// 8-entry per-lane queue spent ~half of the instructions on the test and ran the blend at ~20 % lane utilisation,
// because depth order correlates with screen position.
#pragma once
#include "common.cuh"

namespace das3r {

constexpr int BLEND_WARPS = 2;   // warps per block
constexpr int CHUNK = 32;        // list entries gathered per round (one per lane)
// Tunables (overridable at compile time for experiments: tools/variants.sh)
#ifndef DAS3R_FWD_K
#define DAS3R_FWD_K 4
#endif
#ifndef DAS3R_BWD_K
#define DAS3R_BWD_K 4
#endif
#ifndef DAS3R_FWD_NPAIR
#define DAS3R_FWD_NPAIR 3
#endif
#ifndef DAS3R_BWD_NPAIR
#define DAS3R_BWD_NPAIR 2
#endif
#ifndef DAS3R_LIST_AHEAD
#define DAS3R_LIST_AHEAD 0
#endif
#ifndef DAS3R_BLEND_INFLIGHT
#define DAS3R_BLEND_INFLIGHT 1
#endif
#ifndef DAS3R_BWD_MIN_BLOCKS
#define DAS3R_BWD_MIN_BLOCKS 16
#endif
#ifndef DAS3R_BLEND_BURST
#define DAS3R_BLEND_BURST 3
#endif
constexpr int FWD_K = DAS3R_FWD_K;  // ring depth (chunks) of the forward / backward blend
constexpr int BWD_K = DAS3R_BWD_K;
constexpr int FWD_NPAIR = DAS3R_FWD_NPAIR;  // records prepared side by side per blend round
constexpr int BWD_NPAIR = DAS3R_BWD_NPAIR;
constexpr int LIST_AHEAD = DAS3R_LIST_AHEAD;  // chunks of list entries fetched ahead of the record gather
constexpr int BLEND_INFLIGHT = DAS3R_BLEND_INFLIGHT;  // chunks whose record copies may be in flight (1..3)
static_assert(BLEND_INFLIGHT >= 1 && BLEND_INFLIGHT <= 3, "wait_group immediates below cover 1..3");
constexpr int BLEND_BURST = DAS3R_BLEND_BURST;  // blend rounds between two checks of the gather side

struct RingRec {
    float4 A;  // {x, y, tile-list position (bits), -tau'}
    float4 B;  // {conic a, b, c, opacity}
    float4 C;  // {r, g, b, gaussian index (bits)}
};

template <int K>
struct BlendSmem {
    RingRec ring[BLEND_WARPS][K * CHUNK];
    uint32_t qmask[BLEND_WARPS][K * 32];  // [chunk slot][lane]: which of the chunk's entries touch the lane's pixel
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

// Transposes the 32x32 bit matrix held one row per lane: on return, bit e of lane p's word is bit p of lane e's input.
__device__ __forceinline__ uint32_t transpose_bits(uint32_t x, int lane) {
#pragma unroll
    for (int j = 16; j >= 1; j >>= 1) {
        const uint32_t msk = (j == 16) ? 0x0000ffffu : (j == 8) ? 0x00ff00ffu : (j == 4) ? 0x0f0f0f0fu
                           : (j == 2) ? 0x33333333u : 0x55555555u;
        const uint32_t y = __shfl_xor_sync(0xffffffffu, x, j);
        x = (lane & j) ? ((x & ~msk) | ((y >> j) & msk)) : ((x & msk) | ((y & msk) << j));
    }
    return x;
}

// Walks list[0..cnt) (REVERSE: from the end) as described at the top of this file.
//   keep()        per-lane: does this lane still want records? (forward: pixel not finished)
//   prep(rec)     per-lane, order-independent part of the blend step for one record whose mask bit is set for this
//                 lane's pixel (power, exp, alpha); returns what apply() needs
//   apply(rec, x) per-lane, the part that carries the pixel's recurrence (T, colour sums); called in list order.
//                 Two records are prepared per round so that their independent chains overlap.
//   pos_limit     no lane needs tile-list positions >= pos_limit (backward: the warp's highest n_contrib)
// Chunks are numbered in walk order; `iss` chunks have been started, `vis` of them have landed (records readable, masks
// stored), up to BLEND_INFLIGHT are in flight. A lane holds the unblended rest `m` of its mask of chunk nx-1. Chunk c lives in
// ring slot c % K, so chunk `iss` may be started as soon as no lane needs chunk iss-K any more.
#ifdef DAS3R_BLEND_TIMING  // developer instrumentation: cycles per phase of the walk, per warp
__device__ unsigned long long g_walk_prof[8 * 65536];
#define WALK_PROF_DECL long long wp_t = clock64(); unsigned long long wp[8] = {0, 0, 0, 0, 0, 0, 0, 0}; long long wq_t = 0; unsigned long long wq[6] = {0, 0, 0, 0, 0, 0};
#define WALK_Q0 wq_t = clock64();
#define WALK_Q(i) { const long long wq_n = clock64(); wq[i] += (unsigned long long)(wq_n - wq_t); wq_t = wq_n; }
#define WALK_PROF(i) { const long long wp_n = clock64(); wp[i] += (unsigned long long)(wp_n - wp_t); if ((i) >= 2) wp[(i) + 4]++; wp_t = wp_n; }
#define WALK_COUNT_ROUND { wp[4]++; wp[5] += __popc(__ballot_sync(0xffffffffu, m != 0)); }
#define WALK_PROF_STORE if (lane == 0) { unsigned long long* o = g_walk_prof + 8 * (size_t)(blockIdx.x * BLEND_WARPS + (threadIdx.x >> 5)); for (int q = 0; q < 8; q++) o[q] = wp[q]; unsigned long long* o2 = g_walk_prof + 8 * (size_t)32768 + 8 * (size_t)(blockIdx.x * BLEND_WARPS + (threadIdx.x >> 5)); for (int q = 0; q < 6; q++) o2[q] = wq[q]; }
#else
#define WALK_PROF_DECL
#define WALK_PROF(i)
#define WALK_Q0
#define WALK_Q(i)
#define WALK_COUNT_ROUND
#define WALK_PROF_STORE
#endif

template <bool REVERSE, int K, int NPAIR, typename Keep, typename Prep, typename Apply>
__device__ __forceinline__ void walk_rect_list(RingRec* __restrict__ ring, uint32_t* __restrict__ qmask,
                                               const uint2* __restrict__ list, const uint32_t* __restrict__ gid_list,
                                               int cnt, const GaussRec* __restrict__ rec, const uint32_t pos_limit,
                                               Keep keep, Prep prep, Apply apply) {
    static_assert((K & (K - 1)) == 0, "K must be a power of two");
    const int lane = threadIdx.x & 31;
    const int nchunks = (cnt + CHUNK - 1) / CHUNK;
    if (nchunks == 0) return;
    constexpr int step = REVERSE ? -1 : 1;
    // this lane's entry of list chunk c, in walk order: {position, mask, Gaussian index}; {~0, 0, 0} where there is none
    auto load_entry = [&](int c) -> uint3 {
        if (c < 0 || c >= nchunks) return make_uint3(0xffffffffu, 0u, 0u);
        const int ns = min(CHUNK, cnt - c * CHUNK);
        if (lane >= ns) return make_uint3(0xffffffffu, 0u, 0u);
        const int i = c * CHUNK + (REVERSE ? ns - 1 - lane : lane);
        const uint2 e = list[i];
        return make_uint3(e.x, e.y, gid_list[i]);
    };
    int lc = REVERSE ? nchunks - 1 : 0;  // list chunk to start next
    int left = nchunks;                  // list chunks not started yet
    // list entries are fetched LIST_AHEAD chunks before they are needed: a chunk start must never wait for them
    uint3 e_q[LIST_AHEAD + 1];
#pragma unroll
    for (int j = 0; j <= LIST_AHEAD; j++) e_q[j] = load_entry(lc + j * step);

    uint32_t m = 0;
    int nx = 0, vis = 0, iss = 0;
    // chunks in flight, oldest first: their transposed masks and this lane's entry position wait in registers
    uint32_t pend_mask[BLEND_INFLIGHT], pend_pos[BLEND_INFLIGHT], pend_gid[BLEND_INFLIGHT];
#pragma unroll
    for (int f = 0; f < BLEND_INFLIGHT; f++) {
        pend_mask[f] = 0;
        pend_pos[f] = 0xffffffffu;
        pend_gid[f] = 0;
    }

    WALK_PROF_DECL
    for (;;) {
        if (!keep()) {
            m = 0;
            nx = vis;
        }
        while (m == 0 && nx < vis) {
            m = qmask[(nx & (K - 1)) * 32 + lane];
            nx++;
        }
        const bool idle = (m == 0);  // nothing landed is left for this lane
        if (__all_sync(0xffffffffu, !keep())) break;
        const int oldest = __reduce_min_sync(0xffffffffu, idle ? vis : nx - 1);  // lowest chunk still needed
        int nfly = iss - vis;
        bool acted = false;
        WALK_PROF(0)
        // start list chunks while the window and the pipeline have room (chunks nobody needs are skipped)
        while (left > 0 && nfly < BLEND_INFLIGHT && iss - oldest < K) {
            const uint3 e_cur = e_q[0];
            const bool valid = e_cur.x < pos_limit;
            if (__any_sync(0xffffffffu, valid)) {
                RingRec& r = ring[(iss & (K - 1)) * 32 + lane];
                if (valid) {  // the entry's 48-byte record, straight from the per-Gaussian array
                    const float4* src = reinterpret_cast<const float4*>(rec + e_cur.z);
                    cp_async_16(&r.A, src);
                    cp_async_16(&r.B, src + 1);
                    cp_async_16(&r.C, src + 2);
                }
                cp_async_commit();
                const uint32_t tm = transpose_bits(valid ? e_cur.y : 0u, lane);
#pragma unroll
                for (int f = 0; f < BLEND_INFLIGHT; f++)
                    if (f == nfly) {
                        pend_mask[f] = tm;
                        pend_pos[f] = valid ? e_cur.x : 0xffffffffu;
                        pend_gid[f] = e_cur.z;
                    }
                iss++;
                nfly++;
            }
            lc += step;
            left--;
#pragma unroll
            for (int j = 0; j < LIST_AHEAD; j++) e_q[j] = e_q[j + 1];
            e_q[LIST_AHEAD] = load_entry(lc + LIST_AHEAD * step);
            acted = true;
        }
        WALK_PROF(1)
        if (nfly > 0 && (oldest == vis || nfly == BLEND_INFLIGHT || left == 0)) {
            // land the oldest chunk in flight: the pipeline is full, or everybody is out of work, or nothing is left
            // to start
            if (nfly == 1) cp_async_wait_group<0>();
            else if (nfly == 2) cp_async_wait_group<1>();
            else cp_async_wait_group<2>();
            const int slot = (vis & (K - 1)) * 32 + lane;
            if (pend_pos[0] != 0xffffffffu) {  // the blends want the list position and the Gaussian index with the record
                ring[slot].A.z = __uint_as_float(pend_pos[0]);
                ring[slot].C.w = __uint_as_float(pend_gid[0]);
            }
            qmask[slot] = pend_mask[0];
#pragma unroll
            for (int f = 0; f + 1 < BLEND_INFLIGHT; f++) {
                pend_mask[f] = pend_mask[f + 1];
                pend_pos[f] = pend_pos[f + 1];
                pend_gid[f] = pend_gid[f + 1];
            }
            __syncwarp();
            vis++;
            acted = true;
            WALK_PROF(2)
        }
        if (!acted && oldest == vis) break;  // nothing landed, nothing in flight, nothing left to start
        // blend: a few rounds between two looks at the gather side; lanes step to their next landed chunk on their own
#pragma unroll 1
        for (int round = 0; round < BLEND_BURST; round++) {
            if (m == 0 && nx < vis) {
                m = qmask[(nx & (K - 1)) * 32 + lane];
                nx++;
            }
            WALK_COUNT_ROUND
            WALK_Q0
            if (m) {
                // up to NPAIR records per round: their order-independent parts are evaluated side by side (a missing
                // one repeats the first record and is not applied)
                const RingRec* base = ring + ((nx - 1) & (K - 1)) * 32;
                const RingRec* r[NPAIR];
                bool have[NPAIR];
#pragma unroll
                for (int j = 0; j < NPAIR; j++) {
                    have[j] = (m != 0);
                    const int bit = have[j] ? __ffs(m) - 1 : 0;
                    r[j] = have[j] ? base + bit : r[0];
                    m &= m - 1;
                }
                WALK_Q(0)
                decltype(prep(*r[0])) x[NPAIR];
#pragma unroll
                for (int j = 0; j < NPAIR; j++) x[j] = prep(*r[j]);
                WALK_Q(1)
                apply(*r[0], x[0]);
                WALK_Q(2)
#pragma unroll
                for (int j = 1; j < NPAIR; j++)
                    if (have[j] && keep()) apply(*r[j], x[j]);
                WALK_Q(3)
                if (!keep()) {
                    m = 0;
                    nx = vis;
                }
            }
            WALK_Q(4)
            if (!__any_sync(0xffffffffu, m != 0 || nx < vis)) break;  // nothing landed is left for anybody
            WALK_Q(5)
        }
        WALK_PROF(3)
    }
    WALK_PROF_STORE
    cp_async_wait_all();  // never leave with a copy into this block's shared memory in flight
}

}  // namespace das3r
