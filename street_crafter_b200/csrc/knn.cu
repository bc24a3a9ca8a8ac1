// distCUDA2 (A10): mean squared distance to the 3 nearest other points — exact.
//
// Replaces /root/reference/submodules/simple-knn (spatial.cu:16-25 -> simple_knn.cu:186-222).
// The reference Morton-sorts, cuts 1024-point boxes and lets EVERY point test EVERY box
// (O(P * P/1024) box tests with uncoalesced gathers).  This version keeps the Morton order but
//   * gathers the points once into sorted float4 order (all later reads are coalesced),
//   * builds a 3-level 32-ary box hierarchy (leaf = 32 Morton-consecutive points = one warp),
//   * answers the 32 queries of a leaf COOPERATIVELY: one lane tests one child box against the
//     leaf's own AABB and the warp's largest 3rd-best distance (conservative => exact result),
//     surviving leaves are loaded coalesced and broadcast by shuffle, no divergence.
//   * never synchronises the host (the reference does two blocking cudaMemcpy for the bounds).
// Distances use the same mul/fma/fma structure for points and boxes, so box pruning is
// conservative under fp32 rounding (each op is monotone).
#include <cfloat>
#include "common.cuh"

namespace {

constexpr int LEAF = 32;

struct Box { float lo[3], hi[3]; };

__device__ __forceinline__ int float_to_ordered(float f) {
    const int b = __float_as_int(f);
    return b >= 0 ? b : b ^ 0x7fffffff;
}
__device__ __forceinline__ float ordered_to_float(int k) { return __int_as_float(k >= 0 ? k : k ^ 0x7fffffff); }

__global__ void knn_init_kernel(int* bounds) {
    if (threadIdx.x < 3) bounds[threadIdx.x] = float_to_ordered(FLT_MAX);
    else if (threadIdx.x < 6) bounds[threadIdx.x] = float_to_ordered(-FLT_MAX);
}

__global__ void __launch_bounds__(256)
knn_bounds_kernel(int64_t P, const float* __restrict__ pts, int* __restrict__ bounds) {
    float lo[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, hi[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < P; i += stride) {
#pragma unroll
        for (int a = 0; a < 3; ++a) {
            const float v = __ldg(pts + i * 3 + a);
            lo[a] = fminf(lo[a], v);
            hi[a] = fmaxf(hi[a], v);
        }
    }
#pragma unroll
    for (int a = 0; a < 3; ++a) {
#pragma unroll
        for (int o = 16; o >= 1; o >>= 1) {
            lo[a] = fminf(lo[a], __shfl_xor_sync(0xffffffffu, lo[a], o));
            hi[a] = fmaxf(hi[a], __shfl_xor_sync(0xffffffffu, hi[a], o));
        }
    }
    if ((threadIdx.x & 31) == 0) {
#pragma unroll
        for (int a = 0; a < 3; ++a) {
            atomicMin(bounds + a, float_to_ordered(lo[a]));
            atomicMax(bounds + 3 + a, float_to_ordered(hi[a]));
        }
    }
}

__device__ __forceinline__ uint32_t spread10(uint32_t x) {
    x = (x | (x << 16)) & 0x030000FFu;
    x = (x | (x << 8)) & 0x0300F00Fu;
    x = (x | (x << 4)) & 0x030C30C3u;
    x = (x | (x << 2)) & 0x09249249u;
    return x;
}

__global__ void __launch_bounds__(256)
knn_morton_kernel(int64_t P, const float* __restrict__ pts, const int* __restrict__ bounds,
                  uint64_t* __restrict__ keys, uint32_t* __restrict__ vals) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= P) return;
    uint32_t code = 0;
#pragma unroll
    for (int a = 0; a < 3; ++a) {
        const float lo = ordered_to_float(bounds[a]), hi = ordered_to_float(bounds[3 + a]);
        const float ext = hi - lo;
        const float v = __ldg(pts + i * 3 + a);
        float q = ext > 0.f ? (v - lo) / ext * 1023.0f : 0.f;
        q = fminf(fmaxf(q, 0.f), 1023.f);
        code |= spread10((uint32_t)q) << a;
    }
    keys[i] = code;
    vals[i] = (uint32_t)i;
}

__global__ void __launch_bounds__(256)
knn_gather_kernel(int64_t P, const float* __restrict__ pts, const uint32_t* __restrict__ order,
                  float4* __restrict__ sorted) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= P) return;
    const uint32_t src = __ldg(order + i);
    sorted[i] = make_float4(__ldg(pts + (int64_t)src * 3 + 0), __ldg(pts + (int64_t)src * 3 + 1),
                            __ldg(pts + (int64_t)src * 3 + 2), __uint_as_float(src));
}

__device__ __forceinline__ void warp_minmax(float (&lo)[3], float (&hi)[3]) {
#pragma unroll
    for (int a = 0; a < 3; ++a) {
#pragma unroll
        for (int o = 16; o >= 1; o >>= 1) {
            lo[a] = fminf(lo[a], __shfl_xor_sync(0xffffffffu, lo[a], o));
            hi[a] = fmaxf(hi[a], __shfl_xor_sync(0xffffffffu, hi[a], o));
        }
    }
}

// one warp per leaf: AABB of its (up to) 32 points
__global__ void __launch_bounds__(256)
knn_leaf_box_kernel(int64_t P, int64_t n_leaf, const float4* __restrict__ sorted, Box* __restrict__ boxes) {
    const int lane = threadIdx.x & 31;
    const int64_t leaf = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (leaf >= n_leaf) return;
    const int64_t i = leaf * LEAF + lane;
    float lo[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, hi[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
    if (i < P) {
        const float4 p = sorted[i];
        lo[0] = hi[0] = p.x; lo[1] = hi[1] = p.y; lo[2] = hi[2] = p.z;
    }
    warp_minmax(lo, hi);
    if (lane == 0) {
        Box b;
#pragma unroll
        for (int a = 0; a < 3; ++a) { b.lo[a] = lo[a]; b.hi[a] = hi[a]; }
        boxes[leaf] = b;
    }
}

// one warp per parent: AABB of its (up to) 32 child boxes
__global__ void __launch_bounds__(256)
knn_parent_box_kernel(int64_t n_child, int64_t n_parent, const Box* __restrict__ child, Box* __restrict__ parent) {
    const int lane = threadIdx.x & 31;
    const int64_t pidx = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (pidx >= n_parent) return;
    const int64_t c = pidx * 32 + lane;
    float lo[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, hi[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
    if (c < n_child) {
        const Box b = child[c];
#pragma unroll
        for (int a = 0; a < 3; ++a) { lo[a] = b.lo[a]; hi[a] = b.hi[a]; }
    }
    warp_minmax(lo, hi);
    if (lane == 0) {
        Box b;
#pragma unroll
        for (int a = 0; a < 3; ++a) { b.lo[a] = lo[a]; b.hi[a] = hi[a]; }
        parent[pidx] = b;
    }
}

// Same contraction nvcc picks for the reference's `d.x*d.x + d.y*d.y + d.z*d.z`
// (simple_knn.cu:135-136; SASS: FMUL dy², FFMA dx², FFMA dz²) so results are bit-identical.
__device__ __forceinline__ float sqdist3(float dx, float dy, float dz) {
    return fmaf(dz, dz, fmaf(dx, dx, dy * dy));
}

__device__ __forceinline__ float box_box_dist2(const float (&alo)[3], const float (&ahi)[3], const Box& b) {
    const float gx = fmaxf(0.f, fmaxf(alo[0] - b.hi[0], b.lo[0] - ahi[0]));
    const float gy = fmaxf(0.f, fmaxf(alo[1] - b.hi[1], b.lo[1] - ahi[1]));
    const float gz = fmaxf(0.f, fmaxf(alo[2] - b.hi[2], b.lo[2] - ahi[2]));
    return sqdist3(gx, gy, gz);
}

__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
    for (int o = 16; o >= 1; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}

__global__ void __launch_bounds__(128)
knn_query_kernel(int64_t P, int64_t n0, int64_t n1, int64_t n2, const float4* __restrict__ sorted,
                 const Box* __restrict__ box0, const Box* __restrict__ box1, const Box* __restrict__ box2,
                 float* __restrict__ out) {
    const unsigned full = 0xffffffffu;
    const int lane = threadIdx.x & 31;
    const int64_t own = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (own >= n0) return;
    const int64_t qi = own * LEAF + lane;
    const bool qvalid = qi < P;
    float4 q = make_float4(0.f, 0.f, 0.f, 0.f);
    if (qvalid) q = sorted[qi];
    float b0 = FLT_MAX, b1 = FLT_MAX, b2 = FLT_MAX;

    auto visit_leaf = [&](int64_t leaf) {
        const int64_t pi = leaf * LEAF + lane;
        float4 p = make_float4(0.f, 0.f, 0.f, 0.f);
        if (pi < P) p = sorted[pi];
        const int cnt = (int)min((int64_t)LEAF, P - leaf * LEAF);
        const bool is_own = leaf == own;
        for (int jj = 0; jj < cnt; ++jj) {
            const float x = __shfl_sync(full, p.x, jj), y = __shfl_sync(full, p.y, jj),
                        z = __shfl_sync(full, p.z, jj);
            float d = sqdist3(x - q.x, y - q.y, z - q.z);
            if (is_own && jj == lane) d = FLT_MAX;  // self excluded by index (simple_knn.cu:174-176)
            const float x0 = fmaxf(b0, d);
            b0 = fminf(b0, d);
            const float x1 = fmaxf(b1, x0);
            b1 = fminf(b1, x0);
            b2 = fminf(b2, x1);
        }
    };

    // own leaf AABB (query box)
    float qlo[3], qhi[3];
    {
        const Box ob = box0[own];
#pragma unroll
        for (int a = 0; a < 3; ++a) { qlo[a] = ob.lo[a]; qhi[a] = ob.hi[a]; }
    }
    visit_leaf(own);
    float rmax = warp_max(qvalid ? b2 : 0.f);

    for (int64_t top = 0; top < n2; top += 32) {
        const int64_t c = top + lane;
        bool hit2 = false;
        if (c < n2) hit2 = box_box_dist2(qlo, qhi, box2[c]) <= rmax;
        unsigned m2 = __ballot_sync(full, hit2);
        while (m2) {
            const int64_t c2 = top + (__ffs(m2) - 1);
            m2 &= m2 - 1;
            const int64_t id1 = c2 * 32 + lane;
            bool hit1 = false;
            if (id1 < n1) hit1 = box_box_dist2(qlo, qhi, box1[id1]) <= rmax;
            unsigned m1 = __ballot_sync(full, hit1);
            while (m1) {
                const int64_t c1 = c2 * 32 + (__ffs(m1) - 1);
                m1 &= m1 - 1;
                const int64_t id0 = c1 * 32 + lane;
                bool hit0 = false;
                if (id0 < n0 && id0 != own) hit0 = box_box_dist2(qlo, qhi, box0[id0]) <= rmax;
                unsigned m0 = __ballot_sync(full, hit0);
                while (m0) {
                    const int b = __ffs(m0) - 1;
                    m0 &= m0 - 1;
                    const int64_t c0 = c1 * 32 + b;
                    // rmax may have shrunk since the ballot: cheap uniform re-test
                    if (box_box_dist2(qlo, qhi, box0[c0]) > rmax) continue;
                    visit_leaf(c0);
                    rmax = warp_max(qvalid ? b2 : 0.f);
                }
            }
        }
    }
    if (qvalid) out[__float_as_uint(q.w)] = (b0 + b1 + b2) / 3.0f;
}

struct KnnPlan {
    int64_t n0, n1, n2;
    int64_t off_bounds, off_keys_a, off_keys_b, off_vals_a, off_vals_b, off_sorted, off_box0, off_box1,
        off_box2, off_sort_ws, sort_ws_bytes, total;
};

KnnPlan knn_plan(int64_t P) {
    KnnPlan p;
    p.n0 = sc_div_up(P, LEAF);
    p.n1 = sc_div_up(p.n0, 32);
    p.n2 = sc_div_up(p.n1, 32);
    int64_t o = 0;
    auto take = [&](int64_t bytes) { const int64_t r = o; o += sc_align_up(bytes, 256); return r; };
    p.off_bounds = take(32);
    p.off_keys_a = take(P * 8);
    p.off_keys_b = take(P * 8);
    p.off_vals_a = take(P * 4);
    p.off_vals_b = take(P * 4);
    p.off_sorted = take(P * 16);
    p.off_box0 = take(p.n0 * (int64_t)sizeof(Box));
    p.off_box1 = take(p.n1 * (int64_t)sizeof(Box));
    p.off_box2 = take(p.n2 * (int64_t)sizeof(Box));
    p.sort_ws_bytes = sc_sort_workspace_bytes(P);
    p.off_sort_ws = take(p.sort_ws_bytes);
    p.total = o;
    return p;
}

}  // namespace

extern "C" int64_t sc_knn_workspace_bytes(int64_t P) { return P <= 0 ? 256 : knn_plan(P).total; }

extern "C" int sc_knn3_mean_dist2(int64_t P, const float* points, float* mean_dist2, void* workspace,
                                  int64_t workspace_bytes, void* stream) {
    SC_CHECK_ARG(P >= 0 && P < (1ll << 30), "P");
    if (P == 0) return SC_OK;
    SC_CHECK_ARG(points && mean_dist2 && workspace, "null pointer");
    const KnnPlan pl = knn_plan(P);
    if (workspace_bytes < pl.total) {
        sc_set_error("knn workspace too small: %lld < %lld", (long long)workspace_bytes, (long long)pl.total);
        return SC_ERR_WORKSPACE;
    }
    cudaStream_t st = (cudaStream_t)stream;
    unsigned char* ws = (unsigned char*)workspace;
    int* bounds = (int*)(ws + pl.off_bounds);
    uint64_t* keys_a = (uint64_t*)(ws + pl.off_keys_a);
    uint64_t* keys_b = (uint64_t*)(ws + pl.off_keys_b);
    uint32_t* vals_a = (uint32_t*)(ws + pl.off_vals_a);
    uint32_t* vals_b = (uint32_t*)(ws + pl.off_vals_b);
    float4* sorted = (float4*)(ws + pl.off_sorted);
    Box* box0 = (Box*)(ws + pl.off_box0);
    Box* box1 = (Box*)(ws + pl.off_box1);
    Box* box2 = (Box*)(ws + pl.off_box2);

    knn_init_kernel<<<1, 32, 0, st>>>(bounds);
    SC_CHECK_LAUNCH("knn_init_kernel");
    knn_bounds_kernel<<<(unsigned)min(sc_div_up(P, 256), (int64_t)SC_NUM_SMS * 8), 256, 0, st>>>(P, points, bounds);
    SC_CHECK_LAUNCH("knn_bounds_kernel");
    knn_morton_kernel<<<(unsigned)sc_div_up(P, 256), 256, 0, st>>>(P, points, bounds, keys_a, vals_a);
    SC_CHECK_LAUNCH("knn_morton_kernel");
    const int where = sc_radix_sort_pairs(P, keys_a, vals_a, keys_b, vals_b, 0, 30, ws + pl.off_sort_ws,
                                          pl.sort_ws_bytes, stream);
    if (where < 0) return where;
    const uint32_t* order = where == 0 ? vals_b : vals_a;
    knn_gather_kernel<<<(unsigned)sc_div_up(P, 256), 256, 0, st>>>(P, points, order, sorted);
    SC_CHECK_LAUNCH("knn_gather_kernel");
    knn_leaf_box_kernel<<<(unsigned)sc_div_up(pl.n0 * 32, 256), 256, 0, st>>>(P, pl.n0, sorted, box0);
    SC_CHECK_LAUNCH("knn_leaf_box_kernel");
    knn_parent_box_kernel<<<(unsigned)sc_div_up(pl.n1 * 32, 256), 256, 0, st>>>(pl.n0, pl.n1, box0, box1);
    SC_CHECK_LAUNCH("knn_parent_box_kernel");
    knn_parent_box_kernel<<<(unsigned)sc_div_up(pl.n2 * 32, 256), 256, 0, st>>>(pl.n1, pl.n2, box1, box2);
    SC_CHECK_LAUNCH("knn_parent_box_kernel");
    knn_query_kernel<<<(unsigned)sc_div_up(pl.n0 * 32, 128), 128, 0, st>>>(P, pl.n0, pl.n1, pl.n2, sorted, box0,
                                                                         box1, box2, mean_dist2);
    SC_CHECK_LAUNCH("knn_query_kernel");
    return SC_OK;
}
