// distCUDA2: mean squared distance to the 3 nearest OTHER points (exact), replacing simple-knn
// (reference submodules/simple-knn/simple_knn.cu:186-222).
//
// Same idea as the reference (Morton order -> fixed-size boxes with AABBs -> exact scan of the boxes that can still
// hold a closer point), re-laid for coalesced/broadcast access: the points are physically gathered into Morton order
// once (float4 {x,y,z,original index}); a warp owns 32 consecutive sorted points, prunes 32 boxes per step (one per
// lane) against the warp's own bounding box, and scans surviving boxes with warp-uniform (broadcast) loads.
// No host round trip: the bounding box stays on the device (the reference copies it back twice, simple_knn.cu:197-201).
// The result is the exact 3-NN set, so it equals the reference's up to its own fp32 pruning edge cases; the distance
// and mean expressions are the reference's (updateKBest :132-146, final mean :183).
#include "kernels.cuh"
#include <cub/device/device_radix_sort.cuh>
#include <cfloat>

namespace das3r {

constexpr int KNN_BOX = 512;

static size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

__device__ __forceinline__ uint32_t enc_f(float f) {
    uint32_t b = __float_as_uint(f);
    return (b & 0x80000000u) ? ~b : (b | 0x80000000u);
}
__device__ __forceinline__ float dec_f(uint32_t e) {
    uint32_t b = (e & 0x80000000u) ? (e & 0x7fffffffu) : ~e;
    return __uint_as_float(b);
}

__global__ void knn_init_bbox(uint32_t* bbox) {
    if (threadIdx.x < 3) bbox[threadIdx.x] = 0xffffffffu;      // running min (encoded)
    else if (threadIdx.x < 6) bbox[threadIdx.x] = 0u;          // running max (encoded)
}

__global__ void __launch_bounds__(256) knn_bbox(int P, const float* __restrict__ pts, uint32_t* bbox) {
    float mn[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, mx[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < P; i += gridDim.x * blockDim.x) {
#pragma unroll
        for (int k = 0; k < 3; k++) {
            float v = pts[3 * i + k];
            mn[k] = fminf(mn[k], v);
            mx[k] = fmaxf(mx[k], v);
        }
    }
#pragma unroll
    for (int k = 0; k < 3; k++)
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            mn[k] = fminf(mn[k], __shfl_xor_sync(0xffffffffu, mn[k], o));
            mx[k] = fmaxf(mx[k], __shfl_xor_sync(0xffffffffu, mx[k], o));
        }
    if ((threadIdx.x & 31) == 0) {
#pragma unroll
        for (int k = 0; k < 3; k++) {
            atomicMin(&bbox[k], enc_f(mn[k]));
            atomicMax(&bbox[3 + k], enc_f(mx[k]));
        }
    }
}

__device__ __forceinline__ uint32_t prep_morton(uint32_t x) {  // simple_knn.cu:46-53
    x = (x | (x << 16)) & 0x030000FF;
    x = (x | (x << 8)) & 0x0300F00F;
    x = (x | (x << 4)) & 0x030C30C3;
    x = (x | (x << 2)) & 0x09249249;
    return x;
}

__global__ void __launch_bounds__(256) knn_morton(int P, const float* __restrict__ pts, const uint32_t* __restrict__ bbox,
                                                  uint32_t* __restrict__ codes, uint32_t* __restrict__ idx) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= P) return;
    uint32_t c = 0;
#pragma unroll
    for (int k = 0; k < 3; k++) {
        float lo = dec_f(bbox[k]), hi = dec_f(bbox[3 + k]);
        float ext = hi - lo;
        float u = ext > 0.f ? (pts[3 * i + k] - lo) / ext : 0.f;
        u = fminf(fmaxf(u, 0.f), 1.f);
        c |= prep_morton((uint32_t)(u * ((1 << 10) - 1))) << k;
    }
    codes[i] = c;
    idx[i] = (uint32_t)i;
}

__global__ void __launch_bounds__(256) knn_gather(int P, const float* __restrict__ pts,
                                                  const uint32_t* __restrict__ idx_sorted, float4* __restrict__ sp) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= P) return;
    uint32_t g = idx_sorted[i];
    sp[i] = make_float4(pts[3 * g], pts[3 * g + 1], pts[3 * g + 2], __uint_as_float(g));
}

// One block per box: AABB of KNN_BOX consecutive sorted points. boxes[2b] = min, boxes[2b+1] = max.
__global__ void __launch_bounds__(KNN_BOX) knn_boxes(int P, const float4* __restrict__ sp, float4* __restrict__ boxes) {
    __shared__ float smn[3][KNN_BOX / 32], smx[3][KNN_BOX / 32];
    int i = blockIdx.x * KNN_BOX + threadIdx.x;
    float mn[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, mx[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
    if (i < P) {
        float4 p = sp[i];
        mn[0] = mx[0] = p.x;
        mn[1] = mx[1] = p.y;
        mn[2] = mx[2] = p.z;
    }
#pragma unroll
    for (int k = 0; k < 3; k++)
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            mn[k] = fminf(mn[k], __shfl_xor_sync(0xffffffffu, mn[k], o));
            mx[k] = fmaxf(mx[k], __shfl_xor_sync(0xffffffffu, mx[k], o));
        }
    int w = threadIdx.x >> 5, l = threadIdx.x & 31;
    if (l == 0)
#pragma unroll
        for (int k = 0; k < 3; k++) {
            smn[k][w] = mn[k];
            smx[k][w] = mx[k];
        }
    __syncthreads();
    if (w == 0) {
#pragma unroll
        for (int k = 0; k < 3; k++) {
            float a = l < KNN_BOX / 32 ? smn[k][l] : FLT_MAX, b = l < KNN_BOX / 32 ? smx[k][l] : -FLT_MAX;
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
                a = fminf(a, __shfl_xor_sync(0xffffffffu, a, o));
                b = fmaxf(b, __shfl_xor_sync(0xffffffffu, b, o));
            }
            mn[k] = a;
            mx[k] = b;
        }
        if (l == 0) {
            boxes[2 * blockIdx.x] = make_float4(mn[0], mn[1], mn[2], 0.f);
            boxes[2 * blockIdx.x + 1] = make_float4(mx[0], mx[1], mx[2], 0.f);
        }
    }
}

__device__ __forceinline__ void update3(const float4& ref, const float4& point, float* knn) {  // simple_knn.cu:132-146
    float3 d = {point.x - ref.x, point.y - ref.y, point.z - ref.z};
    float dist = d.x * d.x + d.y * d.y + d.z * d.z;
#pragma unroll
    for (int j = 0; j < 3; j++) {
        if (knn[j] > dist) {
            float t = knn[j];
            knn[j] = dist;
            dist = t;
        }
    }
}

__device__ __forceinline__ float dist_box_point(const float4& bmin, const float4& bmax, const float4& p) {  // :120-130
    float3 diff = {0, 0, 0};
    if (p.x < bmin.x || p.x > bmax.x) diff.x = fminf(fabsf(p.x - bmin.x), fabsf(p.x - bmax.x));
    if (p.y < bmin.y || p.y > bmax.y) diff.y = fminf(fabsf(p.y - bmin.y), fabsf(p.y - bmax.y));
    if (p.z < bmin.z || p.z > bmax.z) diff.z = fminf(fabsf(p.z - bmin.z), fabsf(p.z - bmax.z));
    return diff.x * diff.x + diff.y * diff.y + diff.z * diff.z;
}

__global__ void __launch_bounds__(256) knn_search(int P, const float4* __restrict__ sp, const float4* __restrict__ boxes,
                                                  float* __restrict__ out) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const int lane = threadIdx.x & 31;
    const bool valid = i < P;
    const int ic = valid ? i : P - 1;
    const float4 me = sp[ic];
    float best[3] = {FLT_MAX, FLT_MAX, FLT_MAX};
    if (valid)
        for (int k = max(0, i - 3); k <= min(P - 1, i + 3); k++) {  // simple_knn.cu:157-162
            if (k == i) continue;
            update3(me, sp[k], best);
        }
    const float reject = best[2];
    best[0] = best[1] = best[2] = FLT_MAX;

    // warp bounding box and the largest rejection radius in the warp (conservative warp-level prune)
    float wmn[3] = {me.x, me.y, me.z}, wmx[3] = {me.x, me.y, me.z};
    float wrej = valid ? reject : 0.f;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
#pragma unroll
        for (int k = 0; k < 3; k++) {
            wmn[k] = fminf(wmn[k], __shfl_xor_sync(0xffffffffu, wmn[k], o));
            wmx[k] = fmaxf(wmx[k], __shfl_xor_sync(0xffffffffu, wmx[k], o));
        }
        wrej = fmaxf(wrej, __shfl_xor_sync(0xffffffffu, wrej, o));
    }
    const int nb = (P + KNN_BOX - 1) / KNN_BOX;
    for (int b0 = 0; b0 < nb; b0 += 32) {
        const int bl = b0 + lane;
        bool cand = false;
        if (bl < nb) {
            float4 bmin = boxes[2 * bl], bmax = boxes[2 * bl + 1];
            float gx = fmaxf(0.f, fmaxf(bmin.x - wmx[0], wmn[0] - bmax.x));
            float gy = fmaxf(0.f, fmaxf(bmin.y - wmx[1], wmn[1] - bmax.y));
            float gz = fmaxf(0.f, fmaxf(bmin.z - wmx[2], wmn[2] - bmax.z));
            // slightly deflated so fp32 rounding can never prune a box the per-point test below would accept
            float g2 = (gx * gx + gy * gy + gz * gz) * 0.9999f;
            cand = !(g2 > wrej);
        }
        unsigned m = __ballot_sync(0xffffffffu, cand);
        while (m) {
            const int j = __ffs(m) - 1;
            m &= m - 1;
            const int b = b0 + j;
            const float4 bmin = boxes[2 * b], bmax = boxes[2 * b + 1];
            const float dist = dist_box_point(bmin, bmax, me);  // simple_knn.cu:169-172
            const bool need = valid && !(dist > reject || dist > best[2]);
            if (__any_sync(0xffffffffu, need)) {
                const int kend = min(P, (b + 1) * KNN_BOX);
                for (int k = b * KNN_BOX; k < kend; k++) {
                    const float4 q = sp[k];  // warp-uniform address: one broadcast transaction
                    if (need && k != i) update3(me, q, best);
                }
            }
        }
    }
    if (valid) out[__float_as_uint(me.w)] = (best[0] + best[1] + best[2]) / 3.0f;  // simple_knn.cu:183
}

struct KnnLayout {
    size_t bbox, codes, codes_sorted, idx, idx_sorted, sp, boxes, cub, cub_bytes, total;
};

static KnnLayout knn_layout(int P) {
    KnnLayout L;
    size_t off = 0;
    auto take = [&](size_t bytes) {
        size_t o = off;
        off += align_up(bytes, 256);
        return o;
    };
    const size_t n = (size_t)(P > 0 ? P : 1);
    L.bbox = take(8 * 4);
    L.codes = take(n * 4);
    L.codes_sorted = take(n * 4);
    L.idx = take(n * 4);
    L.idx_sorted = take(n * 4);
    L.sp = take(n * 16);
    L.boxes = take(((n + KNN_BOX - 1) / KNN_BOX) * 32);
    L.cub_bytes = 0;
    cub::DeviceRadixSort::SortPairs((void*)nullptr, L.cub_bytes, (const uint32_t*)nullptr, (uint32_t*)nullptr,
                                    (const uint32_t*)nullptr, (uint32_t*)nullptr, (int)n, 0, 30);
    L.cub = take(L.cub_bytes);
    L.total = off;
    return L;
}

size_t knn_scratch_bytes(int P) { return knn_layout(P).total; }

cudaError_t knn_dist2(int P, const float* points, float* out, void* scratch, size_t scratch_bytes, cudaStream_t s) {
    if (P <= 0) return cudaSuccess;
    KnnLayout L = knn_layout(P);
    if (scratch_bytes < L.total) return cudaErrorInvalidValue;
    char* base = static_cast<char*>(scratch);
    uint32_t* bbox = reinterpret_cast<uint32_t*>(base + L.bbox);
    uint32_t* codes = reinterpret_cast<uint32_t*>(base + L.codes);
    uint32_t* codes_sorted = reinterpret_cast<uint32_t*>(base + L.codes_sorted);
    uint32_t* idx = reinterpret_cast<uint32_t*>(base + L.idx);
    uint32_t* idx_sorted = reinterpret_cast<uint32_t*>(base + L.idx_sorted);
    float4* sp = reinterpret_cast<float4*>(base + L.sp);
    float4* boxes = reinterpret_cast<float4*>(base + L.boxes);
    const int blocks = (P + 255) / 256;
    knn_init_bbox<<<1, 32, 0, s>>>(bbox);
    knn_bbox<<<min(blocks, 148 * 8), 256, 0, s>>>(P, points, bbox);
    knn_morton<<<blocks, 256, 0, s>>>(P, points, bbox, codes, idx);
    size_t cub_bytes = L.cub_bytes;
    cudaError_t e = cub::DeviceRadixSort::SortPairs(base + L.cub, cub_bytes, codes, codes_sorted, idx, idx_sorted, P, 0,
                                                    30, s);
    if (e != cudaSuccess) return e;
    knn_gather<<<blocks, 256, 0, s>>>(P, points, idx_sorted, sp);
    knn_boxes<<<(P + KNN_BOX - 1) / KNN_BOX, KNN_BOX, 0, s>>>(P, sp, boxes);
    knn_search<<<blocks, 256, 0, s>>>(P, sp, boxes, out);
    return cudaGetLastError();
}

}  // namespace das3r
