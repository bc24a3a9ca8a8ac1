// Fused projection (A1) and tile binning (A3/A4) — the integer-feeding chain.
//
// THIS FILE IS COMPILED WITH -fmad=false (see build.py): every fp32 operation below is a
// separately rounded IEEE op in exactly the association order of oracle/gsplat_torch.py, so
// radii, tile bounds and the depth bits of the sort key are bit-identical to the CPU oracle.
// The kernels are HBM-bound streaming passes (80 B/Gaussian), so giving up FMA costs nothing.
//
// Replaces gsplat.fully_fused_projection / isect_tiles / isect_offset_encode as called at
// /root/reference/street_gaussian/models/street_gaussian_renderer.py:219-253.
#include <algorithm>
#include <cooperative_groups.h>
#include "common.cuh"
#include "sort_internal.cuh"

namespace {

#ifndef SC_PROJ_THREADS
#define SC_PROJ_THREADS 128
#endif
constexpr int PROJ_THREADS = SC_PROJ_THREADS;

struct Cam {
    float W[3][3];
    float t[3];
    float fx, fy, cx, cy;
};

__device__ __forceinline__ Cam load_cam(const float* __restrict__ viewmat, const float* __restrict__ K) {
    Cam c;
#pragma unroll
    for (int i = 0; i < 3; ++i) {
#pragma unroll
        for (int j = 0; j < 3; ++j) c.W[i][j] = __ldg(viewmat + i * 4 + j);
        c.t[i] = __ldg(viewmat + i * 4 + 3);
    }
    c.fx = __ldg(K + 0);
    c.cx = __ldg(K + 2);
    c.fy = __ldg(K + 4);
    c.cy = __ldg(K + 5);
    return c;
}

// quat (wxyz, un-normalised) + scale -> 6 unique entries of the world covariance
__device__ __forceinline__ void quat_scale_to_covar(const float4 q, const float s0, const float s1,
                                                    const float s2, float S[6]) {
    float w = q.x, x = q.y, y = q.z, z = q.w;
    float n2 = ((w * w + x * x) + y * y) + z * z;
    float nrm = fmaxf(sqrtf(n2), 1e-12f);
    w = w / nrm; x = x / nrm; y = y / nrm; z = z / nrm;
    float x2 = x * x, y2 = y * y, z2 = z * z;
    float xy = x * y, xz = x * z, yz = y * z;
    float wx = w * x, wy = w * y, wz = w * z;
    float R[3][3] = {{1.0f - 2.0f * (y2 + z2), 2.0f * (xy - wz), 2.0f * (xz + wy)},
                     {2.0f * (xy + wz), 1.0f - 2.0f * (x2 + z2), 2.0f * (yz - wx)},
                     {2.0f * (xz - wy), 2.0f * (yz + wx), 1.0f - 2.0f * (x2 + y2)}};
    float M[3][3];
#pragma unroll
    for (int i = 0; i < 3; ++i) {
        M[i][0] = R[i][0] * s0;
        M[i][1] = R[i][1] * s1;
        M[i][2] = R[i][2] * s2;
    }
    S[0] = (M[0][0] * M[0][0] + M[0][1] * M[0][1]) + M[0][2] * M[0][2];
    S[1] = (M[0][0] * M[1][0] + M[0][1] * M[1][1]) + M[0][2] * M[1][2];
    S[2] = (M[0][0] * M[2][0] + M[0][1] * M[2][1]) + M[0][2] * M[2][2];
    S[3] = (M[1][0] * M[1][0] + M[1][1] * M[1][1]) + M[1][2] * M[1][2];
    S[4] = (M[1][0] * M[2][0] + M[1][1] * M[2][1]) + M[1][2] * M[2][2];
    S[5] = (M[2][0] * M[2][0] + M[2][1] * M[2][1]) + M[2][2] * M[2][2];
}

#ifndef SC_PF_SPAN
#define SC_PF_SPAN 2  // Gaussians per thread in the pre-cull scan (a CTA owns PROJ_THREADS * SPAN of them)
#endif
constexpr int PF_SPAN = SC_PF_SPAN;
constexpr int PF_CHUNK = PROJ_THREADS * PF_SPAN;

struct ProjOut {
    int32_t* __restrict__ radii;
    float* __restrict__ means2d;
    float* __restrict__ depths;
    float* __restrict__ conics;
    float* __restrict__ comps;
    const float* __restrict__ opacities;
    float* __restrict__ op_out;
    uint8_t* __restrict__ vis_out;
    float* __restrict__ radii_norm;
    float inv_max_dim;
};

__device__ __forceinline__ void proj_store(const ProjOut& O, int64_t o, int64_t g, int32_t radius_i, float mx,
                                           float my, float z, float a, float b, float c, float comp) {
    O.radii[o] = radius_i;
    reinterpret_cast<float2*>(O.means2d)[o] = make_float2(mx, my);
    O.depths[o] = z;
    O.conics[o * 3 + 0] = a; O.conics[o * 3 + 1] = b; O.conics[o * 3 + 2] = c;
    if (O.comps) O.comps[o] = comp;
    // A2 fused: opacity * compensation (plain opacity in classic mode)
    if (O.op_out) O.op_out[o] = O.comps ? __ldg(O.opacities + g) * comp : __ldg(O.opacities + g);
    // A11 fused: the renderer's `visibility_filter` (radii > 0) and `radii` (radii / max(H, W),
    // which torch evaluates as float(radii) * (1 / max)) (street_gaussian_renderer.py:293-294)
    if (O.vis_out) O.vis_out[o] = radius_i > 0 ? 1 : 0;
    if (O.radii_norm) O.radii_norm[o] = (float)radius_i * O.inv_max_dim;
}

// the full chain for one Gaussian that survived the pre-cull (canonical IEEE op order)
__device__ __forceinline__ void
proj_one(int64_t g, int64_t o, const Cam& c, const float* __restrict__ means, const float* __restrict__ quats,
         const float* __restrict__ scales, int width, int height, float eps2d, float near_plane, float far_plane,
         float radius_clip, const ProjOut& O) {
    int32_t radius_i = 0;
    float o_mx = 0.f, o_my = 0.f, o_z = 0.f, o_a = 0.f, o_b = 0.f, o_c = 0.f, o_comp = 0.f;
    {
        const float px = __ldg(means + g * 3 + 0), py = __ldg(means + g * 3 + 1), pz = __ldg(means + g * 3 + 2);
        const float4 q = __ldg(reinterpret_cast<const float4*>(quats) + g);
        float S[6];
        quat_scale_to_covar(q, __ldg(scales + g * 3 + 0), __ldg(scales + g * 3 + 1), __ldg(scales + g * 3 + 2), S);
        const float Sm[3][3] = {{S[0], S[1], S[2]}, {S[1], S[3], S[4]}, {S[2], S[4], S[5]}};
        const float xc = ((c.W[0][0] * px + c.W[0][1] * py) + c.W[0][2] * pz) + c.t[0];
        const float yc = ((c.W[1][0] * px + c.W[1][1] * py) + c.W[1][2] * pz) + c.t[1];
        const float zc = ((c.W[2][0] * px + c.W[2][1] * py) + c.W[2][2] * pz) + c.t[2];
        float A[3][3];
#pragma unroll
        for (int i = 0; i < 3; ++i)
#pragma unroll
            for (int j = 0; j < 3; ++j)
                A[i][j] = (c.W[i][0] * Sm[0][j] + c.W[i][1] * Sm[1][j]) + c.W[i][2] * Sm[2][j];
#define CC(i, j) ((A[i][0] * c.W[j][0] + A[i][1] * c.W[j][1]) + A[i][2] * c.W[j][2])
        const float c00 = CC(0, 0), c01 = CC(0, 1), c02 = CC(0, 2), c11 = CC(1, 1), c12 = CC(1, 2),
                    c22 = CC(2, 2);
#undef CC
        const float Wf = (float)width, Hf = (float)height;
        const float tan_fovx = 0.5f * Wf / c.fx;
        const float tan_fovy = 0.5f * Hf / c.fy;
        const float lim_x_pos = (Wf - c.cx) / c.fx + SC_JAC_PAD * tan_fovx;
        const float lim_x_neg = c.cx / c.fx + SC_JAC_PAD * tan_fovx;
        const float lim_y_pos = (Hf - c.cy) / c.fy + SC_JAC_PAD * tan_fovy;
        const float lim_y_neg = c.cy / c.fy + SC_JAC_PAD * tan_fovy;
        const float zs = (zc == 0.0f) ? 1.0f : zc;
        const float tx = zs * fminf(fmaxf(xc / zs, -lim_x_neg), lim_x_pos);
        const float ty = zs * fminf(fmaxf(yc / zs, -lim_y_neg), lim_y_pos);
        const float z2 = zs * zs;
        const float J00 = c.fx / zs;
        const float J02 = -(c.fx * tx) / z2;
        const float J11 = c.fy / zs;
        const float J12 = -(c.fy * ty) / z2;
        const float T00 = J00 * c00 + J02 * c02;
        const float T01 = J00 * c01 + J02 * c12;
        const float T02 = J00 * c02 + J02 * c22;
        const float T11 = J11 * c11 + J12 * c12;
        const float T12 = J11 * c12 + J12 * c22;
        const float v00 = T00 * J00 + T02 * J02;
        const float v01 = T01 * J11 + T02 * J12;
        const float v11 = T11 * J11 + T12 * J12;
        const float mx = (c.fx * xc) / zs + c.cx;
        const float my = (c.fy * yc) / zs + c.cy;
        const float det_orig = v00 * v11 - v01 * v01;
        const float b00 = v00 + eps2d;
        const float b11 = v11 + eps2d;
        const float det = b00 * b11 - v01 * v01;
        bool valid = (det > 0.0f) && (zc > near_plane) && (zc < far_plane);
        const float dets = (det > 0.0f) ? det : 1.0f;
        const float comp = sqrtf(fmaxf(det_orig / dets, 0.0f));
        const float b = 0.5f * (b00 + b11);
        const float v1 = b + sqrtf(fmaxf(b * b - det, SC_EIGEN_FLOOR));
        const float radius = ceilf(3.0f * sqrtf(v1));
        valid = valid && (radius > radius_clip);
        const bool inside = (mx + radius > 0.0f) && (mx - radius < Wf) && (my + radius > 0.0f) &&
                            (my - radius < Hf);
        valid = valid && inside;
        if (valid) {
            radius_i = (int32_t)radius;
            o_mx = mx; o_my = my; o_z = zc;
            o_a = b11 / dets; o_b = -v01 / dets; o_c = b00 / dets;
            o_comp = comp;
        }
    }
    proj_store(O, o, g, radius_i, o_mx, o_my, o_z, o_a, o_b, o_c, o_comp);
}

// A CTA owns PF_CHUNK consecutive Gaussians of one camera.  On a street scene two thirds of them
// are culled, most because they lie outside the frustum: each thread first runs a cheap
// CONSERVATIVE test on PF_SPAN Gaussians (exact depth test; lateral test against an upper bound
// of the radius: lambda_max(cov2d) <= max(scale)^2 |J|^2 + eps), writes the all-zero rows of the
// ones it rejects (coalesced) and compacts the ids of the rest, which then run the full chain
// with full warps.  A Gaussian rejected here is rejected by the full chain too, so the outputs
// are bit-identical to evaluating every Gaussian.
__global__ void __launch_bounds__(PROJ_THREADS)
projection_fwd_kernel(int64_t N, const float* __restrict__ means, const float* __restrict__ quats,
                      const float* __restrict__ scales, const float* __restrict__ viewmats,
                      const float* __restrict__ Ks, int width, int height, float eps2d, float near_plane,
                      float far_plane, float radius_clip, ProjOut O) {
    __shared__ uint16_t s_list[PF_CHUNK];
    __shared__ int s_wsum[PROJ_THREADS / 32];
    const int cam = blockIdx.y;
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const int64_t base = (int64_t)blockIdx.x * PF_CHUNK;
    const Cam c = load_cam(viewmats + cam * 16, Ks + cam * 9);
    const float Wf = (float)width, Hf = (float)height;
    // |tx / z| and |ty / z| never exceed these (the Jacobian clamp of the full chain)
    const float Lx = fmaxf((Wf - c.cx) / c.fx, c.cx / c.fx) + SC_JAC_PAD * (0.5f * Wf / c.fx);
    const float Ly = fmaxf((Hf - c.cy) / c.fy, c.cy / c.fy) + SC_JAC_PAD * (0.5f * Hf / c.fy);
    // |W|_2^2 <= |W|_1 |W|_inf (1 for a rotation up to this bound's slack; holds for any view matrix)
    float w1 = 0.f, winf = 0.f;
#pragma unroll
    for (int i = 0; i < 3; ++i) {
        w1 = fmaxf(w1, fabsf(c.W[0][i]) + fabsf(c.W[1][i]) + fabsf(c.W[2][i]));
        winf = fmaxf(winf, fabsf(c.W[i][0]) + fabsf(c.W[i][1]) + fabsf(c.W[i][2]));
    }
    const float jac2 = 1.02f * (w1 * winf) * (c.fx * c.fx * (1.0f + Lx * Lx) + c.fy * c.fy * (1.0f + Ly * Ly));
    bool keep[PF_SPAN];
#pragma unroll
    for (int k = 0; k < PF_SPAN; ++k) {
        const int64_t g = base + k * PROJ_THREADS + tid;
        keep[k] = false;
        if (g < N) {
            const float px = __ldg(means + g * 3 + 0), py = __ldg(means + g * 3 + 1), pz = __ldg(means + g * 3 + 2);
            const float smax = fmaxf(fmaxf(fabsf(__ldg(scales + g * 3 + 0)), fabsf(__ldg(scales + g * 3 + 1))),
                                     fabsf(__ldg(scales + g * 3 + 2)));
            const float xc = ((c.W[0][0] * px + c.W[0][1] * py) + c.W[0][2] * pz) + c.t[0];
            const float yc = ((c.W[1][0] * px + c.W[1][1] * py) + c.W[1][2] * pz) + c.t[1];
            const float zc = ((c.W[2][0] * px + c.W[2][1] * py) + c.W[2][2] * pz) + c.t[2];
            bool dead = !((zc > near_plane) && (zc < far_plane));  // exactly the full chain's depth test
            if (!dead && zc > 0.0f) {
                const float mx = (c.fx * xc) / zc + c.cx;  // same expressions as the full chain
                const float my = (c.fy * yc) / zc + c.cy;
                // lambda_max(cov2d + eps I) <= smax^2 (|J0|^2 + |J1|^2) + 2 eps + 0.1, radius = ceil(3 sqrt(.))
                const float lam = (smax * smax) * (jac2 / (zc * zc)) + (2.0f * eps2d + 0.1f);
                const float rub = 3.0f * sqrtf(lam) * 1.001f + 2.0f;
                dead = (mx + rub <= 0.0f) || (mx - rub >= Wf) || (my + rub <= 0.0f) || (my - rub >= Hf);
            }
            keep[k] = !dead;
            if (dead) proj_store(O, (int64_t)cam * N + g, g, 0, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f);
        }
    }
    int round_base = 0;
#pragma unroll
    for (int k = 0; k < PF_SPAN; ++k) {
        const unsigned b = __ballot_sync(0xffffffffu, keep[k]);
        if (lane == 0) s_wsum[wid] = __popc(b);
        __syncthreads();
        int off = round_base;
#pragma unroll
        for (int w = 0; w < PROJ_THREADS / 32; ++w) {
            const int cnt = s_wsum[w];
            if (w < wid) off += cnt;
            round_base += cnt;
        }
        if (keep[k]) s_list[off + __popc(b & ((1u << lane) - 1u))] = (uint16_t)(k * PROJ_THREADS + tid);
        __syncthreads();
    }
    const int total = round_base;
    for (int i = tid; i < total; i += PROJ_THREADS) {
        const int64_t g = base + s_list[i];
        proj_one(g, (int64_t)cam * N + g, c, means, quats, scales, width, height, eps2d, near_plane, far_plane,
                 radius_clip, O);
    }
}

// ------------------------------------------------------------------ tile bounds (A3 pass 1)
__device__ __forceinline__ int64_t sc_dev_round_up(int64_t a, int64_t b) { return (a + b - 1) / b * b; }

struct TileBox { int x0, x1, y0, y1; };

__device__ __forceinline__ TileBox tile_box(float mx, float my, int32_t radius, float ts, int tw, int th) {
    const float r = (float)radius / ts;
    const float tx = mx / ts;
    const float ty = my / ts;
    TileBox b;
    b.x0 = (int)fminf(fmaxf(floorf(tx - r), 0.0f), (float)tw);
    b.x1 = (int)fminf(fmaxf(ceilf(tx + r), 0.0f), (float)tw);
    b.y0 = (int)fminf(fmaxf(floorf(ty - r), 0.0f), (float)th);
    b.y1 = (int)fminf(fmaxf(ceilf(ty + r), 0.0f), (float)th);
    return b;
}

__global__ void __launch_bounds__(256)
isect_count_kernel(int64_t CN, const float2* __restrict__ means2d, const int32_t* __restrict__ radii,
                   const float* __restrict__ depths, float ts, int tw, int th,
                   int32_t* __restrict__ tiles_per_gauss, uint32_t* __restrict__ depth_keys,
                   uint32_t* __restrict__ order_init) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= CN) return;
    const int32_t r = __ldg(radii + i);
    int32_t n = 0;
    if (r > 0) {
        const float2 m = __ldg(means2d + i);
        const TileBox b = tile_box(m.x, m.y, r, ts, tw, th);
        n = (b.x1 - b.x0) * (b.y1 - b.y0);
    }
    tiles_per_gauss[i] = n;
    if (depth_keys) {
        // entries that emit nothing sort last; the others by the raw depth bits (the low 32 key bits)
        depth_keys[i] = n > 0 ? __float_as_uint(__ldg(depths + i)) : 0xffffffffu;
        order_init[i] = (uint32_t)i;
    }
}

// counts in depth order: out[j] = tiles_per_gauss[order[j]]
__global__ void __launch_bounds__(256)
isect_gather_counts_kernel(int64_t CN, const uint32_t* __restrict__ order,
                           const int32_t* __restrict__ tiles_per_gauss, int32_t* __restrict__ out) {
    const int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= CN) return;
    out[j] = __ldg(tiles_per_gauss + __ldg(order + j));
}

// ------------------------------------------------------------------ emit (A3 pass 2)
// Output-centric and load-balanced: a CTA owns EMIT_OUT CONSECUTIVE output slots, whatever the
// per-entry tile counts are (in depth order the near, screen-filling Gaussians are neighbours:
// an entry-centric split would leave a few warps with all the work).  The CTA finds the first
// entry that owns one of its slots by binary search in the exclusive scan, stages entries 256
// at a time in shared memory (tile box, depth bits, id), and every thread resolves its slots'
// owners with a shared-memory binary search.  All stores are consecutive 8 B + 4 B per lane.
constexpr int EMIT_THREADS = 256;
constexpr int EMIT_OUT = 2048;

__global__ void __launch_bounds__(EMIT_THREADS)
isect_emit_kernel(int64_t N, int64_t CN, const float2* __restrict__ means2d,
                  const int32_t* __restrict__ radii, const float* __restrict__ depths,
                  const int64_t* __restrict__ cum_tiles /*[CN+1]*/, const uint32_t* __restrict__ order,
                  float ts, int tw, int th, int tile_n_bits, int64_t n_isects,
                  int64_t* __restrict__ isect_ids, int32_t* __restrict__ flatten_ids) {
    __shared__ int s_start[EMIT_THREADS];   // first output slot of the entry, relative to p0
    __shared__ int s_x0[EMIT_THREADS], s_y0[EMIT_THREADS], s_w[EMIT_THREADS];
    __shared__ uint32_t s_dbits[EMIT_THREADS];
    __shared__ int s_id[EMIT_THREADS], s_camhi[EMIT_THREADS];
    __shared__ int64_t s_j0;
    const int tid = threadIdx.x;
    const int64_t p0 = (int64_t)blockIdx.x * EMIT_OUT;
    const int64_t p1 = min(p0 + EMIT_OUT, n_isects);
    if (tid == 0) {
        // largest j in [0, CN) with cum[j] <= p0 (the last of equal starts owns the slot)
        int64_t lo = 0, hi = CN;  // invariant: cum[lo] <= p0, cum[hi] > p0 (cum[CN] = n_isects > p0)
        while (hi - lo > 1) {
            const int64_t mid = (lo + hi) >> 1;
            if (__ldg(cum_tiles + mid) <= p0) lo = mid; else hi = mid;
        }
        s_j0 = lo;
    }
    __syncthreads();
    for (int64_t je = s_j0; je < CN; je += EMIT_THREADS) {
        const int64_t chunk_first = __ldg(cum_tiles + je);
        if (chunk_first >= p1) break;
        const int64_t jn = min(je + EMIT_THREADS, CN);
        const int64_t chunk_end = __ldg(cum_tiles + jn);  // first slot owned by a later chunk
        __syncthreads();                                   // previous chunk's readers are done
        {
            const int64_t j = je + tid;
            int start = 0x7fffffff, x0 = 0, y0 = 0, w = 1, id = 0, camhi = 0;
            uint32_t dbits = 0;
            if (j < jn) {
                const int64_t st = __ldg(cum_tiles + j);
                start = (int)max(min(st - p0, (int64_t)EMIT_OUT), (int64_t)-0x40000000);
                if (st < p1 && __ldg(cum_tiles + j + 1) > st) {
                    const int64_t i = order ? (int64_t)__ldg(order + j) : j;
                    const float2 m = __ldg(means2d + i);
                    const TileBox b = tile_box(m.x, m.y, __ldg(radii + i), ts, tw, th);
                    x0 = b.x0; y0 = b.y0;
                    w = max(b.x1 - b.x0, 1);
                    dbits = __float_as_uint(__ldg(depths + i));
                    id = (int)i;
                    camhi = (int)(i / N) << tile_n_bits;
                }
            }
            s_start[tid] = start; s_x0[tid] = x0; s_y0[tid] = y0; s_w[tid] = w;
            s_dbits[tid] = dbits; s_id[tid] = id; s_camhi[tid] = camhi;
        }
        __syncthreads();
        const int qlo = (int)(max(p0, chunk_first) - p0);
        const int qhi = (int)(min(p1, chunk_end) - p0);
        for (int q = qlo + tid; q < qhi; q += EMIT_THREADS) {
            // last t with s_start[t] <= q (padding entries hold INT_MAX)
            int t = 0;
#pragma unroll
            for (int step = EMIT_THREADS / 2; step >= 1; step >>= 1)
                if (s_start[t + step] <= q) t += step;
            const int k = q - s_start[t];
            const int w = s_w[t];
            const int ty = s_y0[t] + k / w;
            const int tx = s_x0[t] + k % w;
            const int64_t hi32 = (int64_t)(s_camhi[t] | (ty * tw + tx));
            isect_ids[p0 + q] = (hi32 << 32) | (int64_t)s_dbits[t];
            flatten_ids[p0 + q] = s_id[t];
        }
    }
}

// ------------------------------------------------------------------ fused binning pipeline
// sc_isect_prepare / sc_isect_emit_sorted: the same result as count -> depth pre-sort ->
// gather -> scan -> emit -> tile sort, in fewer launches and over fewer bytes:
//   * isect_count_compact_kernel writes tiles_per_gauss AND the ORDER-PRESERVING compaction of
//     the entries that emit something (depth bits, entry index): the depth pre-sort then moves
//     V (visible) pairs instead of C*N; the visible count stays on the device (every consumer
//     reads it from memory, grids are sized for C*N and surplus CTAs exit);
//   * isect_scan_kernel gathers the tile counts in depth order and scans them in ONE pass
//     (decoupled look-back), and records for every 2048-slot output block the entry that owns
//     its first slot, so the emit CTAs start without a 21-step binary search;
//   * the emit kernel accumulates the digit histograms of the tile sort while it writes the
//     keys (the keys are not re-read for a histogram pass).
// Look-back words pack {flag (2 bits), value (62 bits)} in one 64-bit word: a single store
// publishes both, no fences.
#define LB_AGG (1ull << 62)
#define LB_INC (2ull << 62)
#define LB_VAL ((1ull << 62) - 1ull)

// exclusive prefix of tile `tile` over the published states of tiles [0, tile): called by ONE
// full warp, every lane returns the result
__device__ __forceinline__ unsigned long long lookback_exclusive(volatile unsigned long long* state,
                                                                 int64_t tile, int lane) {
    unsigned long long excl = 0;
    int64_t idx = tile - 1;
    while (idx >= 0) {
        const int64_t i = idx - lane;
        unsigned long long s;
        do {
            s = (i >= 0) ? state[i] : LB_INC;
        } while (__any_sync(0xffffffffu, (s >> 62) == 0));
        const unsigned inc = __ballot_sync(0xffffffffu, (s >> 62) == 2);
        const int first = inc ? __ffs(inc) - 1 : 31;  // nearest tile with an inclusive prefix
        unsigned long long v = (lane <= first) ? (s & LB_VAL) : 0ull;
#pragma unroll
        for (int o = 16; o >= 1; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
        excl += v;
        if (inc) break;
        idx -= 32;
    }
    return excl;
}

constexpr int CC_THREADS = 256;
constexpr int CC_ITEMS = 4;
constexpr int CC_TILE = CC_THREADS * CC_ITEMS;
constexpr int CC_CTAS_PER_SM = 4;  // persistent grid: the digit histograms are flushed once per CTA

// Persistent: a CTA takes tiles of CC_TILE entries from an atomic ticket until none is left.
//   hist     (nullable): RAW 8-bit digit counts [4][256] of the compacted depth keys, i.e. the
//            histogram pass of the depth sort, accumulated here while the keys are in registers.
__global__ void __launch_bounds__(CC_THREADS)
isect_count_compact_kernel(int64_t CN, int64_t n_tiles, const float2* __restrict__ means2d,
                           const int32_t* __restrict__ radii, const float* __restrict__ depths, float ts, int tw,
                           int th, int32_t* __restrict__ tiles_per_gauss, uint32_t* __restrict__ depth_keys,
                           uint32_t* __restrict__ entry_ids,
                           volatile unsigned long long* __restrict__ state, uint32_t* __restrict__ ticket,
                           int64_t* __restrict__ totals, uint32_t* __restrict__ hist) {
    __shared__ uint32_t s_cnt[CC_ITEMS][CC_THREADS / 32];
    __shared__ uint32_t s_tile;
    __shared__ unsigned long long s_excl;
    __shared__ uint32_t s_hist[4][SC_SORT_RADIX];
    const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
    if (hist)
        for (int i = tid; i < 4 * SC_SORT_RADIX; i += CC_THREADS) (&s_hist[0][0])[i] = 0;
    for (;;) {
        if (tid == 0) s_tile = atomicAdd(ticket, 1u);
        __syncthreads();
        const int64_t tile = s_tile;
        if (tile >= n_tiles) break;
        const int64_t base = tile * CC_TILE;
        // striped: item k of thread t is entry base + k*256 + t (coalesced); the compaction order is
        // entry order = (k, t) lexicographic
        uint32_t key[CC_ITEMS];
        bool vis[CC_ITEMS];
        uint32_t pre[CC_ITEMS];  // visible entries of the same item in lower lanes of this warp
#pragma unroll
        for (int k = 0; k < CC_ITEMS; ++k) {
            const int64_t i = base + k * CC_THREADS + tid;
            int32_t n = 0;
            key[k] = 0;
            if (i < CN) {
                const int32_t r = __ldg(radii + i);
                if (r > 0) {
                    const float2 m = __ldg(means2d + i);
                    const TileBox b = tile_box(m.x, m.y, r, ts, tw, th);
                    n = (b.x1 - b.x0) * (b.y1 - b.y0);
                    key[k] = __float_as_uint(__ldg(depths + i));
                }
                tiles_per_gauss[i] = n;
            }
            vis[k] = n > 0;
            const unsigned bal = __ballot_sync(0xffffffffu, vis[k]);
            pre[k] = __popc(bal & ((1u << lane) - 1u));
            if (lane == 0) s_cnt[k][w] = __popc(bal);
            if (hist) {
                // low digits are random (plain shared atomics); the top digit (sign + exponent) takes a
                // handful of values over a whole scene: one atomic per distinct value of the warp
                if (vis[k]) {
                    atomicAdd(&s_hist[0][key[k] & 255u], 1u);
                    atomicAdd(&s_hist[1][(key[k] >> 8) & 255u], 1u);
                    atomicAdd(&s_hist[2][(key[k] >> 16) & 255u], 1u);
                }
                const uint32_t d3 = key[k] >> 24;
                const unsigned peers = __match_any_sync(0xffffffffu, vis[k] ? d3 : 0x100u);
                if (vis[k] && (peers & ((1u << lane) - 1u)) == 0) atomicAdd(&s_hist[3][d3], (uint32_t)__popc(peers));
            }
        }
        __syncthreads();
        // block-level exclusive offsets of (item, warp) groups, in (k, w) order
        uint32_t my_off[CC_ITEMS];
        uint32_t run = 0;
#pragma unroll
        for (int k = 0; k < CC_ITEMS; ++k) {
#pragma unroll
            for (int ww = 0; ww < CC_THREADS / 32; ++ww) {
                if (ww == w) my_off[k] = run;
                run += s_cnt[k][ww];
            }
        }
        const uint32_t block_total = run;
        if (w == 0) {
            if (lane == 0) state[tile] = (tile == 0 ? LB_INC : LB_AGG) | (unsigned long long)block_total;
            unsigned long long excl = 0;
            if (tile > 0) {
                excl = lookback_exclusive(state, tile, lane);
                if (lane == 0) state[tile] = LB_INC | (excl + block_total);
            }
            if (lane == 0) {
                s_excl = excl;
                if (base + CC_TILE >= CN) totals[0] = (int64_t)(excl + block_total);  // the last tile
            }
        }
        __syncthreads();
        const int64_t out0 = (int64_t)s_excl;
#pragma unroll
        for (int k = 0; k < CC_ITEMS; ++k) {
            if (vis[k]) {
                const int64_t o = out0 + my_off[k] + pre[k];
                depth_keys[o] = key[k];
                entry_ids[o] = (uint32_t)(base + k * CC_THREADS + tid);
            }
        }
    }
    if (hist) {
        __syncthreads();
        for (int i = tid; i < 4 * SC_SORT_RADIX; i += CC_THREADS) {
            const uint32_t v = (&s_hist[0][0])[i];
            if (v) atomicAdd(&hist[i], v);
        }
    }
}

// cum[j] = sum_{j' < j} tiles_per_gauss[order[j']] for j in [0, V], V read from the device;
// first_entry[b] = the entry that owns output slot b * EMIT_OUT.
constexpr int SG_THREADS = 256;
constexpr int SG_ITEMS = 8;
constexpr int SG_TILE = SG_THREADS * SG_ITEMS;
constexpr int EMIT_OUT_LOG2 = 11;

__global__ void __launch_bounds__(SG_THREADS)
isect_scan_kernel(const int64_t* __restrict__ totals, const uint32_t* __restrict__ order,
                  const int32_t* __restrict__ tiles_per_gauss, int64_t* __restrict__ cum,
                  uint32_t* __restrict__ first_entry, int64_t first_cap,
                  volatile unsigned long long* __restrict__ state, uint32_t* __restrict__ ticket,
                  int64_t* __restrict__ total_out) {
    __shared__ unsigned long long s_warp[SG_THREADS / 32];
    __shared__ uint32_t s_tile;
    __shared__ unsigned long long s_excl;
    const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
    const int64_t V = totals[0];
    if (tid == 0) s_tile = atomicAdd(ticket, 1u);
    __syncthreads();
    const int64_t tile = s_tile;
    const int64_t base = tile * SG_TILE;
    if (base >= V) {
        if (tile == 0 && tid == 0) { cum[0] = 0; *total_out = 0; }
        return;
    }
    // blocked: thread t owns entries base + t*8 .. +7
    const int64_t j0 = base + (int64_t)tid * SG_ITEMS;
    int32_t v[SG_ITEMS];
    unsigned long long sum = 0;
#pragma unroll
    for (int k = 0; k < SG_ITEMS; ++k) {
        v[k] = (j0 + k < V) ? __ldg(tiles_per_gauss + __ldg(order + j0 + k)) : 0;
        sum += (unsigned long long)v[k];
    }
    unsigned long long incl = sum;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const unsigned long long t = __shfl_up_sync(0xffffffffu, incl, d);
        if (lane >= d) incl += t;
    }
    if (lane == 31) s_warp[w] = incl;
    __syncthreads();
    unsigned long long wpre = 0, block_total = 0;
#pragma unroll
    for (int ww = 0; ww < SG_THREADS / 32; ++ww) {
        const unsigned long long t = s_warp[ww];
        if (ww < w) wpre += t;
        block_total += t;
    }
    if (w == 0) {
        if (lane == 0) state[tile] = (tile == 0 ? LB_INC : LB_AGG) | block_total;
        unsigned long long excl = 0;
        if (tile > 0) {
            excl = lookback_exclusive(state, tile, lane);
            if (lane == 0) state[tile] = LB_INC | (excl + block_total);
        }
        if (lane == 0) s_excl = excl;
    }
    __syncthreads();
    int64_t ex = (int64_t)(s_excl + wpre + incl - sum);
#pragma unroll
    for (int k = 0; k < SG_ITEMS; ++k) {
        if (j0 + k < V) {
            cum[j0 + k] = ex;
            const int64_t end = ex + v[k];
            // output blocks whose first slot b*EMIT_OUT lies in [ex, end)
            for (int64_t b = (ex + (1 << EMIT_OUT_LOG2) - 1) >> EMIT_OUT_LOG2; (b << EMIT_OUT_LOG2) < end; ++b)
                if (b < first_cap) first_entry[b] = (uint32_t)(j0 + k);
            ex = end;
            if (j0 + k == V - 1) { cum[V] = end; *total_out = end; }
        }
    }
}

// emit with the entry table and the fused sort histogram (see the pipeline header above)
#ifndef SC_EMIT_ILP
#define SC_EMIT_ILP 2
#endif
constexpr int EMIT_ILP = SC_EMIT_ILP;  // output slots per lane and iteration (independent search chains)

__global__ void __launch_bounds__(EMIT_THREADS)
isect_emit_hist_kernel(int64_t N, int64_t V, const float2* __restrict__ means2d,
                       const int32_t* __restrict__ radii, const float* __restrict__ depths,
                       const int64_t* __restrict__ cum_tiles /*[V+1]*/, const uint32_t* __restrict__ order,
                       const uint32_t* __restrict__ first_entry, float ts, int tw, int th, int tile_n_bits,
                       int64_t n_isects, int64_t* __restrict__ isect_ids, int32_t* __restrict__ flatten_ids,
                       ScPassPlan pp, uint32_t* __restrict__ hist /*[passes][256]*/) {
    static_assert(EMIT_OUT == (1 << EMIT_OUT_LOG2), "first_entry granularity");
    __shared__ int s_start[EMIT_THREADS];
    __shared__ int s_x0[EMIT_THREADS], s_y0[EMIT_THREADS], s_w[EMIT_THREADS];
    __shared__ uint32_t s_dbits[EMIT_THREADS];
    __shared__ int s_id[EMIT_THREADS], s_camhi[EMIT_THREADS];
    __shared__ uint32_t s_hist[4][SC_SORT_RADIX];
    const int tid = threadIdx.x, lane = tid & 31;
    const int64_t p0 = (int64_t)blockIdx.x * EMIT_OUT;
    const int64_t p1 = min(p0 + EMIT_OUT, n_isects);
    for (int i = tid; i < pp.passes * SC_SORT_RADIX; i += EMIT_THREADS) (&s_hist[0][0])[i] = 0;
    const uint32_t lt_mask = (1u << lane) - 1u;
    for (int64_t je = __ldg(first_entry + blockIdx.x); je < V; je += EMIT_THREADS) {
        const int64_t chunk_first = __ldg(cum_tiles + je);
        if (chunk_first >= p1) break;
        const int64_t jn = min(je + EMIT_THREADS, V);
        const int64_t chunk_end = __ldg(cum_tiles + jn);
        __syncthreads();
        {
            const int64_t j = je + tid;
            int start = 0x7fffffff, x0 = 0, y0 = 0, w = 1, id = 0, camhi = 0;
            uint32_t dbits = 0;
            if (j < jn) {
                const int64_t st = __ldg(cum_tiles + j);
                start = (int)max(min(st - p0, (int64_t)EMIT_OUT), (int64_t)-0x40000000);
                if (st < p1) {
                    const int64_t i = (int64_t)__ldg(order + j);
                    const float2 m = __ldg(means2d + i);
                    const TileBox b = tile_box(m.x, m.y, __ldg(radii + i), ts, tw, th);
                    x0 = b.x0; y0 = b.y0;
                    w = max(b.x1 - b.x0, 1);
                    dbits = __float_as_uint(__ldg(depths + i));
                    id = (int)i;
                    camhi = (int)(i / N) << tile_n_bits;
                }
            }
            s_start[tid] = start; s_x0[tid] = x0; s_y0[tid] = y0; s_w[tid] = w;
            s_dbits[tid] = dbits; s_id[tid] = id; s_camhi[tid] = camhi;
        }
        __syncthreads();
        const int qlo = (int)(max(p0, chunk_first) - p0);
        const int qhi = (int)(min(p1, chunk_end) - p0);
        // all lanes of a warp iterate together (the histogram uses warp votes); EMIT_ILP slots per
        // lane and iteration, so that the two dependent chains (8-step search over the staged entries,
        // integer divide) interleave
        for (int qb = qlo + (tid & ~31); qb < qhi; qb += EMIT_ILP * EMIT_THREADS) {
            int q[EMIT_ILP], t[EMIT_ILP];
            bool in[EMIT_ILP];
            uint32_t hi32[EMIT_ILP];
#pragma unroll
            for (int u = 0; u < EMIT_ILP; ++u) {
                q[u] = qb + u * EMIT_THREADS + lane;
                in[u] = q[u] < qhi;
                t[u] = 0;
                hi32[u] = 0u;
            }
#pragma unroll
            for (int step = EMIT_THREADS / 2; step >= 1; step >>= 1) {
#pragma unroll
                for (int u = 0; u < EMIT_ILP; ++u)
                    if (s_start[t[u] + step] <= q[u]) t[u] += step;
            }
#pragma unroll
            for (int u = 0; u < EMIT_ILP; ++u) {
                if (in[u]) {
                    const int k = q[u] - s_start[t[u]];
                    const int w = s_w[t[u]];
                    const int ty = s_y0[t[u]] + k / w;
                    const int tx = s_x0[t[u]] + k % w;
                    hi32[u] = (uint32_t)(s_camhi[t[u]] | (ty * tw + tx));
                    isect_ids[p0 + q[u]] = (int64_t)(((uint64_t)hi32[u] << 32) | (uint64_t)s_dbits[t[u]]);
                    flatten_ids[p0 + q[u]] = s_id[t[u]];
                }
            }
#pragma unroll
            for (int u = 0; u < EMIT_ILP; ++u) {
                if (u > 0 && qb + u * EMIT_THREADS >= qhi) break;  // warp-uniform
                // pass 0 digits of neighbouring slots are consecutive tiles (distinct): plain atomics;
                // higher digits are shared by most of the warp: one atomic per distinct digit
                if (in[u]) atomicAdd(&s_hist[0][(hi32[u] >> (pp.shift[0] - 32)) & ((1u << pp.nbits[0]) - 1u)], 1u);
                for (int p = 1; p < pp.passes; ++p) {
                    const uint32_t d = (hi32[u] >> (pp.shift[p] - 32)) & ((1u << pp.nbits[p]) - 1u);
                    // neighbouring slots share the high digits: match.any resolves in one or two rounds here
                    const unsigned peers = __match_any_sync(0xffffffffu, in[u] ? d : 0xffffffffu);
                    if (in[u] && (peers & lt_mask) == 0) atomicAdd(&s_hist[p][d], (uint32_t)__popc(peers));
                }
            }
        }
    }
    __syncthreads();
    for (int i = tid; i < pp.passes * SC_SORT_RADIX; i += EMIT_THREADS) {
        const uint32_t v = (&s_hist[0][0])[i];
        if (v) atomicAdd(&hist[i], v);
    }
}

// ------------------------------------------------------------------ fast binning path
// C == 1 and a tile grid of at most 128 x 128 tiles (images up to 2048 x 2048 at tile 16 — every
// configuration of the reference).  The sorted list is ordered by (tile row ty, tile column tx,
// depth, Gaussian index).  With the visible Gaussians already in stable depth order, that is a
// stable 2-digit LSD sort by tx, then by ty — and because tx / ty are TILE-BOX coordinates:
//   * both digit histograms follow from the boxes alone (h added to columns x0..x1-1, w to rows
//     y0..y1-1), so the scan kernel builds them and no pass over the intersections is needed;
//   * the FIRST sorting pass is fused into the emit kernel: a CTA expands 4096 consecutive slots
//     of the depth-ordered emission in shared memory, ranks them by tx and scatters 8-byte
//     (ty | tx | Gaussian) pairs — the unsorted (isect_id, flatten_id) list is never written;
//   * the SECOND pass writes the final arrays: flatten_ids, (optionally) the 64-bit isect_ids
//     rebuilt from depths[gaussian], and the tile offsets through one atomicMin per (tile, CTA).
// Traffic per intersection: 8 B written + 8 B read + 4 B (12 B more with isect_ids) written,
// against 12 + 2 x 24 + 8 B of the generic emit -> 2 x u64 onesweep -> offset-encode sequence.
constexpr int FB_THREADS = 256;
constexpr int FB_WARPS = FB_THREADS / 32;
constexpr int FB_KPT = 16;
constexpr int FB_TILE = FB_THREADS * FB_KPT;  // 4096 slots per CTA
constexpr int FB_TILE_LOG2 = 12;
constexpr int FB_RADIX = 128;
#ifndef SC_FB_MIN_CTAS
#define SC_FB_MIN_CTAS 4   // measured 2 / 3 / 4 CTAs per SM: finish half 0.701 / 0.683 / 0.659 ms per cfg2 step
#endif
constexpr int FB_MIN_CTAS = SC_FB_MIN_CTAS;   // resident CTAs per SM the two sorting kernels are compiled for
#ifndef SC_FB_LB_W
#define SC_FB_LB_W 2   // measured on cfg2: look-back window 16 / 8 / 4 -> finish half 0.724 / 0.733 / 0.683 ms per step; at 4 CTAs per SM: 4 / 2 -> 0.660 / 0.630
#endif
constexpr int FB_LB_W = SC_FB_LB_W;
static_assert(FB_TILE == (1 << FB_TILE_LOG2), "tile size");
// item blocks of the emit kernel: an item is one (Gaussian, tile column) pair, i.e. a run of h
// consecutive tile rows that lands contiguously in the column-sorted list
constexpr int FI_KPT = 8;
constexpr int FI_TILE = FB_THREADS * FI_KPT;  // 2048 items per CTA
constexpr int FI_TILE_LOG2 = 11;
static_assert(FI_TILE == (1 << FI_TILE_LOG2), "item tile size");
#define FB_FLAG_AGG (1u << 30)
#define FB_FLAG_INC (2u << 30)
#define FB_FLAG_MASK (3u << 30)

// lanes of the warp whose 7-bit digit equals this lane's (7 independent ballots)
__device__ __forceinline__ unsigned fb_match_digit7(unsigned d) {
    unsigned peers = 0xffffffffu;
#pragma unroll
    for (int b = 0; b < 7; ++b) {
        const bool bit = (d >> b) & 1u;
        const unsigned bal = __ballot_sync(0xffffffffu, bit);
        peers &= bit ? bal : ~bal;
    }
    return peers;
}

// ------------------------------------------------------------------ cooperative prepare (fast path)
// The prepare half of the binning works on V ~ 0.3 N visible Gaussians (0.6 M of 2 M on the bench
// scene): count + compaction, a 4-pass depth sort of V (key, index) pairs, a scan.  As separate
// launches each of those is latency, not bandwidth: ~300 tiles start at once and every one walks
// back through the others' look-back words (10-25 us per launch for 5 MB of data).  Here the whole
// half is ONE cooperative kernel, one CTA of 1024 threads per SM, with grid-wide barriers between
// the phases — the classic three-step radix pass (per-CTA digit counts -> barrier -> every CTA sums
// the counts of its predecessors -> stable scatter), no look-back chains, no launch gaps:
//   phase 0  tile boxes + tiles_per_gauss for the CTA's contiguous slice of the N Gaussians; the
//            visible ones are staged IN ORDER in the slice's own region (no global compaction:
//            pass 0 of the sort reads the slices in CTA order, which is entry order);
//   pass p   (p = 0..3, 8-bit digits of the depth bits) count -> barrier -> bases -> scatter -> barrier;
//   scan     tile counts of the sorted order, exclusive scan (CTA totals exchanged through one
//            more barrier), boxes / ids in depth order, the block table of the emit kernel, and
//            the tile column / row histograms.
namespace cg = cooperative_groups;
#ifndef SC_CP_THREADS
#define SC_CP_THREADS 512
#endif
constexpr int CP_THREADS = SC_CP_THREADS;          // 512 x 2 CTAs per SM: one CTA's barrier waits overlap the other's work
constexpr int CP_CTAS_PER_SM = 1024 / CP_THREADS;
constexpr int CP_WARPS = CP_THREADS / 32;
constexpr int CP_KPT = 8;   // 4096 keys per CTA and tile: V up to 1.2 M sorts with one tile per CTA and pass
constexpr int CP_TILE = CP_THREADS * CP_KPT;
constexpr int CP_MAX_CTAS = CP_THREADS;            // one thread per CTA sums the per-CTA totals

struct CoopArgs {
    int64_t N;
    const float2* means2d;
    const int32_t* radii;
    const float* depths;
    float ts;
    int tw, th;
    int32_t* tiles_per_gauss;
    uint32_t *stage_key, *stage_gid, *stage_rect;  // [N] slices, visible entries of a slice packed at its start
    uint32_t *keyA, *valA, *keyB, *valB;            // [N] ping-pong buffers of the depth sort
    uint32_t* cnt_g;                                // [G] visible per CTA
    unsigned long long* tot_g;                      // [G][2] intersections / items per CTA chunk (scan phase)
    uint32_t* hist_g;                               // [G][256] digit counts per CTA
    int64_t* totals;                                // {V, I, M}: visible, intersections, (entry, tile column) items
    int64_t* cum;                                   // [V + 1] exclusive scan of the tile counts w * h
    uint32_t* cumw;                                 // [V + 1] exclusive scan of the box widths w (items)
    uint32_t* first_item;                           // entry that owns the first item of every FI_TILE item block
    uint32_t *rect_s, *gid_s;                       // [V] depth order
    uint32_t* first_entry;
    int64_t first_cap;
    uint32_t* hist_xy;                              // [2][128] raw tile column / row counts (zeroed by the host)
    unsigned long long* phase_ns;                   // [16] globaltimer stamps of CTA 0 at the phase boundaries
};

__device__ __forceinline__ void coop_stamp(const CoopArgs& a, int i) {
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        unsigned long long t;
        asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
        a.phase_ns[i] = t;
    }
}

struct CoopSmem {
    uint32_t whist[CP_WARPS][256];   // per-warp digit counters -> absolute positions
    uint32_t base[256];              // next free global position of every digit for this CTA
    uint32_t hist[256];
    uint32_t cnt[CP_KPT][CP_WARPS];
    unsigned long long wsum[CP_WARPS];
    unsigned long long carry;
    int32_t dx[FB_RADIX + 1], dy[FB_RADIX + 1];
    uint32_t run;
};

// stable scatter of in_key[begin, end) (value = in_val[i], or the index i itself when in_val is null)
// by digit (key >> shift) & 255 to the positions that sm.base hands out; sm.base advances.
// (every buffer here is produced earlier in the SAME kernel by other SMs: L2-coherent __ldcg loads only)
__device__ __forceinline__ void coop_scatter(CoopSmem& sm, const uint32_t* in_key, const uint32_t* in_val,
                                             int64_t begin, int64_t end, int shift, uint32_t* out_key,
                                             uint32_t* out_val) {
    const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
    const uint32_t lt_mask = (1u << lane) - 1u;
    for (int64_t tb = begin; tb < end; tb += CP_TILE) {
        for (int i = tid; i < CP_WARPS * 256; i += CP_THREADS) (&sm.whist[0][0])[i] = 0;
        __syncthreads();
        uint32_t key[CP_KPT], val[CP_KPT], peers[CP_KPT];
        uint16_t rank[CP_KPT];
        bool ok[CP_KPT];
#pragma unroll
        for (int k = 0; k < CP_KPT; ++k) {
            const int64_t i = tb + w * (32 * CP_KPT) + k * 32 + lane;  // warp-striped: order = (w, k, lane)
            ok[k] = i < end;
            key[k] = ok[k] ? __ldcg(in_key + i) : 0u;
            val[k] = ok[k] ? (in_val ? __ldcg(in_val + i) : (uint32_t)i) : 0u;
        }
#pragma unroll
        for (int k = 0; k < CP_KPT; ++k)
            peers[k] = sc_match_digit8((key[k] >> shift) & 255u) & __ballot_sync(0xffffffffu, ok[k]);
#pragma unroll
        for (int k = 0; k < CP_KPT; ++k) {
            const uint32_t d = (key[k] >> shift) & 255u;
            const uint32_t pre = sm.whist[w][d];
            rank[k] = (uint16_t)(pre + __popc(peers[k] & lt_mask));
            __syncwarp();
            if (ok[k] && (peers[k] & lt_mask) == 0) sm.whist[w][d] = pre + __popc(peers[k]);
            __syncwarp();
        }
        __syncthreads();
        if (tid < 256) {
            uint32_t acc = sm.base[tid];
#pragma unroll 8
            for (int ww = 0; ww < CP_WARPS; ++ww) {
                const uint32_t c = sm.whist[ww][tid];
                sm.whist[ww][tid] = acc;
                acc += c;
            }
            sm.base[tid] = acc;
        }
        __syncthreads();
#pragma unroll
        for (int k = 0; k < CP_KPT; ++k) {
            if (ok[k]) {
                const uint32_t pos = sm.whist[w][(key[k] >> shift) & 255u] + rank[k];
                out_key[pos] = key[k];
                out_val[pos] = val[k];
            }
        }
        __syncthreads();
    }
}

// sm.base[d] = (sum over digits d' < d of all CTAs' counts) + (counts of digit d in CTAs before this one)
__device__ __forceinline__ void coop_bases(CoopSmem& sm, const uint32_t* hist_g, int c, int G) {
    const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
    // thread (g, q): the 4 digits 4g .. 4g+3 (one 16-byte load per CTA row) over the CTAs cc = q (mod CP_THREADS / 64),
    // eight L2 loads in flight; the partial sums meet in shared memory
    constexpr int NQ = CP_THREADS / 64;
    const int g4 = tid & 63, q = tid >> 6;
    uint4 tot4 = make_uint4(0, 0, 0, 0), mine4 = make_uint4(0, 0, 0, 0);
    for (int cc0 = q; cc0 < G; cc0 += 8 * NQ) {
        uint4 v[8];
#pragma unroll
        for (int u = 0; u < 8; ++u)
            v[u] = (cc0 + NQ * u < G) ? __ldcg(reinterpret_cast<const uint4*>(hist_g + (size_t)(cc0 + NQ * u) * 256) + g4)
                                      : make_uint4(0, 0, 0, 0);
#pragma unroll
        for (int u = 0; u < 8; ++u) {
            tot4.x += v[u].x; tot4.y += v[u].y; tot4.z += v[u].z; tot4.w += v[u].w;
            if (cc0 + NQ * u < c) { mine4.x += v[u].x; mine4.y += v[u].y; mine4.z += v[u].z; mine4.w += v[u].w; }
        }
    }
    uint32_t* scratch = &sm.whist[0][0];   // [q][256] totals, then [q][256] mine
    reinterpret_cast<uint4*>(scratch + q * 256)[g4] = tot4;
    reinterpret_cast<uint4*>(scratch + (NQ + q) * 256)[g4] = mine4;
    __syncthreads();
    uint32_t total = 0, mine = 0;
    if (tid < 256) {
#pragma unroll
        for (int qq = 0; qq < NQ; ++qq) {
            total += scratch[qq * 256 + tid];
            mine += scratch[(NQ + qq) * 256 + tid];
        }
    }
    uint32_t incl = total;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t t = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += t;
    }
    if (lane == 31) sm.cnt[0][w] = incl;
    __syncthreads();
    if (tid < 256) {
        uint32_t wsum = 0;
#pragma unroll
        for (int ww = 0; ww < 8; ++ww)
            if (ww < w) wsum += sm.cnt[0][ww];
        sm.base[tid] = wsum + incl - total + mine;
    }
    __syncthreads();
}

__global__ void __launch_bounds__(CP_THREADS, CP_CTAS_PER_SM)
bin_prepare_coop_kernel(const CoopArgs a) {
    cg::grid_group grid = cg::this_grid();
    __shared__ CoopSmem sm;
    const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
    const int c = blockIdx.x, G = gridDim.x;
    const int64_t N = a.N;
    coop_stamp(a, 0);
    // ------------------------------------------------ phase 0: boxes, counts, in-order staging
    const int64_t sliceN = sc_dev_round_up((N + G - 1) / G, (int64_t)CP_TILE);
    const int64_t n0 = min(N, (int64_t)c * sliceN), n1 = min(N, n0 + sliceN);
    if (tid < 256) sm.hist[tid] = 0;
    if (tid == 0) sm.run = 0;
    __syncthreads();
    for (int64_t base = n0; base < n1; base += CP_TILE) {
        uint32_t key[CP_KPT], box[CP_KPT], pre[CP_KPT];
        bool vis[CP_KPT];
#pragma unroll
        for (int k = 0; k < CP_KPT; ++k) {
            const int64_t i = base + k * CP_THREADS + tid;  // striped: entry order = (k, tid)
            int32_t n = 0;
            key[k] = 0;
            box[k] = 0;
            if (i < n1) {
                const int32_t r = __ldg(a.radii + i);
                if (r > 0) {
                    const float2 m = __ldg(a.means2d + i);
                    const TileBox b = tile_box(m.x, m.y, r, a.ts, a.tw, a.th);
                    n = (b.x1 - b.x0) * (b.y1 - b.y0);
                    key[k] = __float_as_uint(__ldg(a.depths + i));
                    box[k] = (uint32_t)b.x0 | ((uint32_t)b.y0 << 8) | ((uint32_t)(b.x1 - b.x0) << 16) |
                             ((uint32_t)(b.y1 - b.y0) << 24);
                }
                a.tiles_per_gauss[i] = n;
            }
            vis[k] = n > 0;
            const unsigned bal = __ballot_sync(0xffffffffu, vis[k]);
            pre[k] = __popc(bal & ((1u << lane) - 1u));
            if (lane == 0) sm.cnt[k][w] = __popc(bal);
            if (vis[k]) atomicAdd(&sm.hist[key[k] & 255u], 1u);
        }
        __syncthreads();
        // exclusive offsets of the (k, w) groups in (k, w) order: CP_KPT * CP_WARPS counts, two-level scan
        constexpr int NG = CP_KPT * CP_WARPS, NGW = (NG + 31) / 32;
        uint32_t gv = 0, gincl = 0;
        if (tid < NGW * 32) {
            gv = tid < NG ? (&sm.cnt[0][0])[tid] : 0u;
            gincl = gv;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const uint32_t t = __shfl_up_sync(0xffffffffu, gincl, o);
                if (lane >= o) gincl += t;
            }
            if (lane == 31) sm.wsum[w] = gincl;
        }
        __syncthreads();
        uint32_t tile_total = 0;
#pragma unroll
        for (int ww = 0; ww < NGW; ++ww) tile_total += (uint32_t)sm.wsum[ww];
        if (tid < NG) {
            uint32_t gpre = 0;
#pragma unroll
            for (int ww = 0; ww < NGW; ++ww)
                if (ww < w) gpre += (uint32_t)sm.wsum[ww];
            (&sm.cnt[0][0])[tid] = gpre + gincl - gv;
        }
        __syncthreads();
        const uint32_t run = sm.run;
#pragma unroll
        for (int k = 0; k < CP_KPT; ++k) {
            if (vis[k]) {
                const int64_t o = n0 + run + sm.cnt[k][w] + pre[k];
                a.stage_key[o] = key[k];
                a.stage_gid[o] = (uint32_t)(base + k * CP_THREADS + tid);
                a.stage_rect[o] = box[k];
            }
        }
        __syncthreads();
        if (tid == 0) sm.run = run + tile_total;
        __syncthreads();
    }
    const uint32_t my_visible = sm.run;
    if (tid == 0) a.cnt_g[c] = my_visible;
    if (tid < 256) a.hist_g[(size_t)c * 256 + tid] = sm.hist[tid];
    coop_stamp(a, 1);
    grid.sync();
    coop_stamp(a, 2);
    // V and the chunk of the sorted order this CTA owns from here on
    int64_t V;
    {
        unsigned long long v = (tid < G) ? (unsigned long long)__ldcg(a.cnt_g + tid) : 0ull;
#pragma unroll
        for (int o = 16; o >= 1; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
        if (lane == 0) sm.wsum[w] = v;
        __syncthreads();
        unsigned long long t = 0;
        for (int ww = 0; ww < CP_WARPS; ++ww) t += sm.wsum[ww];
        V = (int64_t)t;
        __syncthreads();
    }
    const int64_t chunkV = (V + G - 1) / G;
    const int64_t v0 = min(V, (int64_t)c * chunkV), v1 = min(V, v0 + chunkV);
    // ------------------------------------------------ depth sort: 4 passes of 8 bits
    coop_bases(sm, a.hist_g, c, G);
    coop_scatter(sm, a.stage_key, nullptr, n0, n0 + my_visible, 0, a.keyA, a.valA);
    coop_stamp(a, 3);
    grid.sync();
    coop_stamp(a, 4);
    const uint32_t* kin = a.keyA;
    const uint32_t* vin = a.valA;
    uint32_t* kout = a.keyB;
    uint32_t* vout = a.valB;
    for (int p = 1; p < 4; ++p) {
        const int shift = 8 * p;
        if (tid < 256) sm.hist[tid] = 0;
        __syncthreads();
        for (int64_t i0 = v0; i0 < v1; i0 += CP_TILE) {   // warp-uniform trip count
            uint32_t kk[CP_KPT];
#pragma unroll
            for (int k = 0; k < CP_KPT; ++k) {
                const int64_t i = i0 + k * CP_THREADS + tid;
                kk[k] = (i < v1) ? __ldcg(kin + i) : 0u;
            }
#pragma unroll
            for (int k = 0; k < CP_KPT; ++k) {
                const bool ok = i0 + k * CP_THREADS + tid < v1;
                const uint32_t d = ok ? ((kk[k] >> shift) & 255u) : 0x100u;
                if (p == 3) {
                    // the top digit (sign + exponent) takes a handful of values: one atomic per distinct value
                    const unsigned peers = __match_any_sync(0xffffffffu, d);
                    if (ok && (peers & ((1u << lane) - 1u)) == 0) atomicAdd(&sm.hist[d], (uint32_t)__popc(peers));
                } else if (ok) {
                    atomicAdd(&sm.hist[d], 1u);
                }
            }
        }
        __syncthreads();
        if (tid < 256) a.hist_g[(size_t)c * 256 + tid] = sm.hist[tid];
        if (p == 1) coop_stamp(a, 5);
        grid.sync();
        if (p == 1) coop_stamp(a, 6);
        coop_bases(sm, a.hist_g, c, G);
        if (p == 1) coop_stamp(a, 7);
        coop_scatter(sm, kin, vin, v0, v1, shift, kout, vout);
        if (p == 1) coop_stamp(a, 8);
        grid.sync();
        if (p == 1) coop_stamp(a, 9);
        if (p == 3) coop_stamp(a, 10);
        const uint32_t* tk = kin; kin = kout; kout = const_cast<uint32_t*>(tk);
        const uint32_t* tv = vin; vin = vout; vout = const_cast<uint32_t*>(tv);
    }
    // `vin` holds the staging index of every visible entry in (depth, index) order
    // ------------------------------------------------ scan of the tile counts (slots) and box widths (items)
    for (int i = tid; i <= FB_RADIX; i += CP_THREADS) { sm.dx[i] = 0; sm.dy[i] = 0; }
    for (int pass = 0; pass < 2; ++pass) {
        unsigned long long carry = 0, carry_w = 0;
        if (pass == 1) {
            // exclusive prefixes of the chunk totals over the CTAs before this one, and the grand totals
            unsigned long long v = (tid < G) ? __ldcg(a.tot_g + 2 * tid) : 0ull;
            unsigned long long vw = (tid < G) ? __ldcg(a.tot_g + 2 * tid + 1) : 0ull;
            unsigned long long mine = (tid < c) ? v : 0ull, mine_w = (tid < c) ? vw : 0ull;
#pragma unroll
            for (int o = 16; o >= 1; o >>= 1) {
                v += __shfl_xor_sync(0xffffffffu, v, o);
                vw += __shfl_xor_sync(0xffffffffu, vw, o);
                mine += __shfl_xor_sync(0xffffffffu, mine, o);
                mine_w += __shfl_xor_sync(0xffffffffu, mine_w, o);
            }
            __syncthreads();
            unsigned long long* scr = reinterpret_cast<unsigned long long*>(&sm.whist[0][0]);
            if (lane == 0) { scr[4 * w] = v; scr[4 * w + 1] = vw; scr[4 * w + 2] = mine; scr[4 * w + 3] = mine_w; }
            __syncthreads();
            unsigned long long tot = 0, tot_w = 0;
            for (int ww = 0; ww < CP_WARPS; ++ww) {
                tot += scr[4 * ww]; tot_w += scr[4 * ww + 1]; carry += scr[4 * ww + 2]; carry_w += scr[4 * ww + 3];
            }
            __syncthreads();
            if (c == 0 && tid == 0) {
                a.totals[0] = V;
                a.totals[1] = (int64_t)tot;
                a.totals[2] = (int64_t)tot_w;
                if (V == 0) { a.cum[0] = 0; a.cumw[0] = 0; }
            }
        }
        for (int64_t tb = v0; tb < v1; tb += CP_TILE) {
            // blocked: thread t owns entries tb + t*KPT .. +KPT-1
            const int64_t j0 = tb + (int64_t)tid * CP_KPT;
            uint32_t box[CP_KPT], o[CP_KPT];
            unsigned long long sum = 0;   // slots in the low 40 bits, items above: one scan for both
#pragma unroll
            for (int k = 0; k < CP_KPT; ++k) {
                box[k] = 0;
                o[k] = 0;
                if (j0 + k < v1) {
                    o[k] = __ldcg(vin + j0 + k);
                    box[k] = __ldcg(a.stage_rect + o[k]);
                }
                const unsigned long long bw = (box[k] >> 16) & 255u, bh = box[k] >> 24;
                sum += bw * bh + (bw << 40);
            }
            unsigned long long incl = sum;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                const unsigned long long t = __shfl_up_sync(0xffffffffu, incl, d);
                if (lane >= d) incl += t;
            }
            if (lane == 31) sm.wsum[w] = incl;
            __syncthreads();
            unsigned long long wpre = 0, tile_total = 0;
            for (int ww = 0; ww < CP_WARPS; ++ww) {
                const unsigned long long t = sm.wsum[ww];
                if (ww < w) wpre += t;
                tile_total += t;
            }
            __syncthreads();
            if (pass == 1) {
                const unsigned long long exb = wpre + incl - sum;
                int64_t ex = (int64_t)(carry + (exb & ((1ull << 40) - 1ull)));
                uint32_t exw = (uint32_t)(carry_w + (exb >> 40));
#pragma unroll
                for (int k = 0; k < CP_KPT; ++k) {
                    if (j0 + k < v1) {
                        a.cum[j0 + k] = ex;
                        a.cumw[j0 + k] = exw;
                        a.rect_s[j0 + k] = box[k];
                        a.gid_s[j0 + k] = __ldcg(a.stage_gid + o[k]);
                        const int x0 = box[k] & 255u, y0 = (box[k] >> 8) & 255u, bw = (box[k] >> 16) & 255u, bh = box[k] >> 24;
                        atomicAdd(&sm.dx[x0], bh); atomicAdd(&sm.dx[x0 + bw], -bh);
                        atomicAdd(&sm.dy[y0], bw); atomicAdd(&sm.dy[y0 + bh], -bw);
                        const int64_t end = ex + (int64_t)(bw * bh);
                        for (int64_t b = (ex + FB_TILE - 1) >> FB_TILE_LOG2; (b << FB_TILE_LOG2) < end; ++b)
                            if (b < a.first_cap) a.first_entry[b] = (uint32_t)(j0 + k);
                        const uint32_t endw = exw + (uint32_t)bw;
                        for (uint32_t b = (exw + FI_TILE - 1) >> FI_TILE_LOG2; (b << FI_TILE_LOG2) < endw; ++b)
                            if ((int64_t)b < a.first_cap) a.first_item[b] = (uint32_t)(j0 + k);
                        ex = end;
                        exw = endw;
                        if (j0 + k == V - 1) { a.cum[V] = end; a.cumw[V] = endw; }
                    }
                }
            }
            carry += tile_total & ((1ull << 40) - 1ull);
            carry_w += tile_total >> 40;
        }
        if (pass == 0) {
            if (tid == 0) { a.tot_g[2 * c] = carry; a.tot_g[2 * c + 1] = carry_w; }
            coop_stamp(a, 11);
            grid.sync();
            coop_stamp(a, 12);
        }
    }
    coop_stamp(a, 13);
    __syncthreads();
    if (tid < 2 * 64) {
        // tile column / row histograms of this CTA's chunk: prefix sums of the difference arrays
        const int32_t* dsrc = tid < 64 ? sm.dx : sm.dy;
        uint32_t* hdst = a.hist_xy + (tid < 64 ? 0 : FB_RADIX);
        const int d0 = (tid & 63) * 2;
        int32_t acc = 0;
        for (int i = 0; i <= d0; ++i) acc += dsrc[i];
        if (acc) atomicAdd(hdst + d0, (uint32_t)acc);
        acc += dsrc[d0 + 1];
        if (acc) atomicAdd(hdst + d0 + 1, (uint32_t)acc);
    }
}

struct FbSmem {
    uint64_t pairs[FB_TILE];           // the block in sorted order
    uint32_t warp_hist[FB_WARPS][FB_RADIX];
    uint32_t digit_start[FB_RADIX];
    int64_t gbase[FB_RADIX];
    uint32_t scan_tmp[FB_WARPS];
    uint32_t tile_id;
};

// Shared tail of the two sorting kernels: the per-warp digit counts are in sm.warp_hist; resolves
// the tile's global bases (chained look-back with an FB_LB_W window, digit totals scanned from the
// RAW histogram) into sm.gbase / sm.digit_start and turns sm.warp_hist into warp-exclusive offsets.
template <typename SmemT>
__device__ __forceinline__ void fb_resolve_bases(SmemT& sm, uint32_t tile, const uint32_t* __restrict__ hist_raw,
                                                 volatile uint32_t* __restrict__ lookback) {
    const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
    uint32_t count = 0, excl = 0, hin = 0, hraw = 0;
    if (tid < FB_RADIX) {
        const int d = tid;
        uint32_t acc = 0;
#pragma unroll
        for (int ww = 0; ww < FB_WARPS; ++ww) {
            const uint32_t c = sm.warp_hist[ww][d];
            sm.warp_hist[ww][d] = acc;
            acc += c;
        }
        count = acc;
        volatile uint32_t* my = lookback + (size_t)tile * FB_RADIX + d;
        if (tile == 0) {
            *my = count | FB_FLAG_INC;
        } else {
            *my = count | FB_FLAG_AGG;
            int64_t t = (int64_t)tile - 1;
            bool done = false;
            while (!done) {
                uint32_t v[FB_LB_W];
#pragma unroll
                for (int u = 0; u < FB_LB_W; ++u)
                    v[u] = (t - u >= 0) ? lookback[(size_t)(t - u) * FB_RADIX + d] : FB_FLAG_INC;
                int used = 0;
#pragma unroll
                for (int u = 0; u < FB_LB_W; ++u) {
                    const uint32_t f = v[u] & FB_FLAG_MASK;
                    if (!done && used == u && f != 0) {
                        excl += v[u] & ~FB_FLAG_MASK;
                        used = u + 1;
                        done = f == FB_FLAG_INC;
                    }
                }
                t -= used;
            }
            *my = (excl + count) | FB_FLAG_INC;
        }
        hraw = hist_raw[d];
    }
    // block exclusive scans over the digits: tile-local counts and global digit totals
    uint32_t incl = count;
    hin = hraw;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t t2 = __shfl_up_sync(0xffffffffu, incl, o);
        const uint32_t t3 = __shfl_up_sync(0xffffffffu, hin, o);
        if (lane >= o) { incl += t2; hin += t3; }
    }
    __shared__ uint32_t s_scan2[FB_WARPS];
    if (lane == 31) { sm.scan_tmp[w] = incl; s_scan2[w] = hin; }
    __syncthreads();
    if (tid < FB_RADIX) {
        uint32_t wsum = 0, hsum = 0;
#pragma unroll
        for (int ww = 0; ww < FB_RADIX / 32; ++ww)
            if (ww < w) { wsum += sm.scan_tmp[ww]; hsum += s_scan2[ww]; }
        const uint32_t dstart = wsum + incl - count;
        sm.digit_start[tid] = dstart;
        sm.gbase[tid] = (int64_t)(hsum + hin - hraw) + (int64_t)excl - (int64_t)dstart;
    }
    __syncthreads();
}

struct FeSmem {
    union {
        struct {
            int32_t start[FB_TILE + 4];    // first slot of staged entry t, relative to the block (<= FB_TILE + 1
            uint32_t rect[FB_TILE + 4];    // entries; + 4 keeps `owner` 16-byte aligned)
            uint16_t owner[FB_TILE];       // staged entry that owns slot q
        };
        uint64_t pairs[FB_TILE];           // the block in sorted order
    };
    uint32_t warp_hist[FB_WARPS][FB_RADIX];
    uint32_t digit_start[FB_RADIX];
    int64_t gbase[FB_RADIX];
    uint32_t scan_tmp[FB_WARPS];
    uint32_t tile_id;
};

// emit + first sorting pass (digit = tile column tx)
__global__ void __launch_bounds__(FB_THREADS, FB_MIN_CTAS)
isect_emit_sort_tx_kernel(int64_t V, int64_t n_isects, const int64_t* __restrict__ totals_dev,
                          const int64_t* __restrict__ cum /*[V+1]*/,
                          const uint32_t* __restrict__ rect_s, const uint32_t* __restrict__ gid_s,
                          const uint32_t* __restrict__ first_entry, const uint32_t* __restrict__ hist_tx,
                          volatile uint32_t* __restrict__ lookback, uint32_t* __restrict__ ticket,
                          uint64_t* __restrict__ pairs_out) {
    extern __shared__ __align__(16) unsigned char fb_smem_raw[];
    FeSmem& sm = *reinterpret_cast<FeSmem*>(fb_smem_raw);
    const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
    if (totals_dev) {  // sync-free mode: the counts stay on the device, `n_isects` is the capacity of the outputs
        V = totals_dev[0];
        n_isects = totals_dev[1] <= n_isects ? totals_dev[1] : 0;  // a list that does not fit is DROPPED (empty view)
    }
    if (tid == 0) sm.tile_id = atomicAdd(ticket, 1u);
    for (int i = tid; i < FB_WARPS * FB_RADIX; i += FB_THREADS) (&sm.warp_hist[0][0])[i] = 0;
    for (int q = tid; q < FB_TILE / 2; q += FB_THREADS) reinterpret_cast<uint32_t*>(sm.owner)[q] = 0u;
    __syncthreads();
    const uint32_t tile = sm.tile_id;
    const int64_t p0 = (int64_t)tile * FB_TILE;
    if (p0 >= n_isects) return;
    const int64_t p1 = min(p0 + FB_TILE, n_isects);
    const int n_valid = (int)(p1 - p0);
    // ---- stage every entry that owns one of the block's slots
    const int64_t j0 = __ldg(first_entry + tile);
    const int64_t j1 = (p1 < n_isects) ? (int64_t)__ldg(first_entry + tile + 1) : V - 1;
    const int ne = (int)(j1 - j0 + 1);  // <= FB_TILE + 1: every visible entry owns at least one slot
    for (int t = tid; t < ne; t += FB_THREADS) {
        const int64_t c = __ldg(cum + j0 + t) - p0;
        const int st = (int)max(c, (int64_t)-0x40000000);
        sm.start[t] = st;
        sm.rect[t] = __ldg(rect_s + j0 + t);
    }
    __syncthreads();
    for (int t = tid + 1; t < ne; t += FB_THREADS) {
        const int st = sm.start[t];   // > 0: entry j0 is the LAST one that starts at or before p0
        if (st < FB_TILE) sm.owner[st] = (uint16_t)t;
    }
    __syncthreads();
    {   // inclusive max-scan of the head marks: thread t owns slots t*16 .. +15
        uint32_t o[FB_KPT];
        const uint4* src = reinterpret_cast<const uint4*>(sm.owner + tid * FB_KPT);
        const uint4 v0 = src[0], v1 = src[1];
        const uint32_t raw[8] = {v0.x, v0.y, v0.z, v0.w, v1.x, v1.y, v1.z, v1.w};
        uint32_t run = 0;
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            run = max(run, raw[i] & 0xffffu); o[2 * i] = run;
            run = max(run, raw[i] >> 16); o[2 * i + 1] = run;
        }
        uint32_t incl = run;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const uint32_t t2 = __shfl_up_sync(0xffffffffu, incl, d);
            if (lane >= d) incl = max(incl, t2);
        }
        if (lane == 31) sm.scan_tmp[w] = incl;
        __syncthreads();
        uint32_t pre = __shfl_up_sync(0xffffffffu, incl, 1);
        if (lane == 0) pre = 0;
#pragma unroll
        for (int ww = 0; ww < FB_WARPS; ++ww)
            if (ww < w) pre = max(pre, sm.scan_tmp[ww]);
        uint32_t outw[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) outw[i] = max(o[2 * i], pre) | (max(o[2 * i + 1], pre) << 16);
        uint4* dst = reinterpret_cast<uint4*>(sm.owner + tid * FB_KPT);
        dst[0] = make_uint4(outw[0], outw[1], outw[2], outw[3]);
        dst[1] = make_uint4(outw[4], outw[5], outw[6], outw[7]);
    }
    __syncthreads();
    // ---- expand: slot -> (tx, ty, Gaussian); warp-striped (item i of lane l = warp_base + i*32 + l)
    uint32_t key[FB_KPT], gid[FB_KPT];
    const int warp_base = w * (32 * FB_KPT);
#pragma unroll
    for (int i = 0; i < FB_KPT; ++i) {
        const int local = warp_base + i * 32 + lane;
        key[i] = 0x7f7fu;  // padding sorts last
        gid[i] = 0;
        if (local < n_valid) {
            const int t = sm.owner[local];
            const uint32_t box = sm.rect[t];
            const int k = local - sm.start[t];
            const int bw = (box >> 16) & 255u;
            int qy = (int)((float)k * __fdividef(1.0f, (float)bw));
            int rx = k - qy * bw;
            if (rx < 0) { --qy; rx += bw; }
            if (rx >= bw) { ++qy; rx -= bw; }
            key[i] = (((box >> 8) & 255u) + qy) << 8 | ((box & 255u) + rx);
            gid[i] = __ldg(gid_s + j0 + t);
        }
    }
    // ---- stable warp-level ranking by tx
    const uint32_t lt_mask = (1u << lane) - 1u;
    uint16_t rank[FB_KPT];
    uint32_t peers_of[FB_KPT];
#pragma unroll
    for (int i = 0; i < FB_KPT; ++i) peers_of[i] = fb_match_digit7(key[i] & 127u);
#pragma unroll
    for (int i = 0; i < FB_KPT; ++i) {
        const uint32_t d = key[i] & 127u;
        const uint32_t pre = sm.warp_hist[w][d];
        rank[i] = (uint16_t)(pre + __popc(peers_of[i] & lt_mask));
        __syncwarp();
        if ((peers_of[i] & lt_mask) == 0) sm.warp_hist[w][d] = pre + __popc(peers_of[i]);
        __syncwarp();
    }
    __syncthreads();   // staging arrays are dead from here on (pairs aliases them)
    fb_resolve_bases(sm, tile, hist_tx, lookback);
#pragma unroll
    for (int i = 0; i < FB_KPT; ++i) {
        const uint32_t d = key[i] & 127u;
        const uint32_t pos = sm.digit_start[d] + sm.warp_hist[w][d] + rank[i];
        sm.pairs[pos] = ((uint64_t)key[i] << 32) | gid[i];
    }
    __syncthreads();
#pragma unroll
    for (int i = 0; i < FB_KPT; ++i) {
        const int local = i * FB_THREADS + tid;
        if (local < n_valid) {
            const uint64_t pr = sm.pairs[local];
            pairs_out[sm.gbase[(uint32_t)(pr >> 32) & 127u] + local] = pr;
        }
    }
}

// second (final) sorting pass (digit = tile row ty): flatten_ids, isect_ids (nullable), offsets
__global__ void __launch_bounds__(FB_THREADS, FB_MIN_CTAS)
isect_sort_ty_final_kernel(int64_t n_isects, const int64_t* __restrict__ totals_dev,
                           const uint64_t* __restrict__ pairs_in,
                           const uint32_t* __restrict__ hist_ty, volatile uint32_t* __restrict__ lookback,
                           uint32_t* __restrict__ ticket, const float* __restrict__ depths, int tw,
                           int32_t* __restrict__ flatten_ids, int64_t* __restrict__ isect_ids,
                           int32_t* __restrict__ offsets) {
    extern __shared__ __align__(16) unsigned char fb_smem_raw[];
    FbSmem& sm = *reinterpret_cast<FbSmem*>(fb_smem_raw);
    const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
    if (totals_dev) n_isects = totals_dev[1] <= n_isects ? totals_dev[1] : 0;
    if (tid == 0) sm.tile_id = atomicAdd(ticket, 1u);
    for (int i = tid; i < FB_WARPS * FB_RADIX; i += FB_THREADS) (&sm.warp_hist[0][0])[i] = 0;
    __syncthreads();
    const uint32_t tile = sm.tile_id;
    const int64_t p0 = (int64_t)tile * FB_TILE;
    if (p0 >= n_isects) return;
    const int n_valid = (int)min((int64_t)FB_TILE, n_isects - p0);
    uint64_t pr[FB_KPT];
    const int warp_base = w * (32 * FB_KPT);
#pragma unroll
    for (int i = 0; i < FB_KPT; ++i) {
        const int local = warp_base + i * 32 + lane;
        pr[i] = (local < n_valid) ? __ldg(pairs_in + p0 + local) : ((uint64_t)0x7f7fu << 32);
    }
    const uint32_t lt_mask = (1u << lane) - 1u;
    uint16_t rank[FB_KPT];
    uint32_t peers_of[FB_KPT];
#pragma unroll
    for (int i = 0; i < FB_KPT; ++i) peers_of[i] = fb_match_digit7((uint32_t)(pr[i] >> 40) & 127u);
#pragma unroll
    for (int i = 0; i < FB_KPT; ++i) {
        const uint32_t d = (uint32_t)(pr[i] >> 40) & 127u;
        const uint32_t pre = sm.warp_hist[w][d];
        rank[i] = (uint16_t)(pre + __popc(peers_of[i] & lt_mask));
        __syncwarp();
        if ((peers_of[i] & lt_mask) == 0) sm.warp_hist[w][d] = pre + __popc(peers_of[i]);
        __syncwarp();
    }
    __syncthreads();
    fb_resolve_bases(sm, tile, hist_ty, lookback);
#pragma unroll
    for (int i = 0; i < FB_KPT; ++i) {
        const uint32_t d = (uint32_t)(pr[i] >> 40) & 127u;
        sm.pairs[sm.digit_start[d] + sm.warp_hist[w][d] + rank[i]] = pr[i];
    }
    __syncthreads();
#pragma unroll
    for (int i = 0; i < FB_KPT; ++i) {
        const int local = i * FB_THREADS + tid;
        if (local < n_valid) {
            const uint64_t q = sm.pairs[local];
            const uint32_t k16 = (uint32_t)(q >> 32), g = (uint32_t)q;
            const uint32_t ty = k16 >> 8, tx = k16 & 255u;
            const int64_t pos = sm.gbase[ty] + local;
            const uint32_t tile_id = ty * (uint32_t)tw + tx;
            flatten_ids[pos] = (int32_t)g;
            if (isect_ids)
                isect_ids[pos] = (int64_t)(((uint64_t)tile_id << 32) | (uint64_t)__float_as_uint(__ldg(depths + g)));
            // the first entry of a tile inside this block: candidate for the tile's offset
            if (local == 0 || (uint32_t)(sm.pairs[local - 1] >> 32) != k16) atomicMin(offsets + tile_id, (int32_t)pos);
        }
    }
}

// offsets[t] = min(offsets[t..]) with offsets[n_tiles] = n_isects: empty tiles take the next start
__global__ void __launch_bounds__(1024)
isect_offsets_fixup_kernel(int n_tiles, const int64_t* __restrict__ n_isects_dev, int64_t n_isects,
                           int32_t* __restrict__ offsets) {
    __shared__ int32_t s_warp[32];
    const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
    // with the device count, `n_isects` is the capacity of the list (a list that does not fit is dropped: all
    // offsets 0, the view renders empty)
    const int32_t total = (int32_t)(n_isects_dev ? (*n_isects_dev <= n_isects ? *n_isects_dev : 0) : n_isects);
    const int per = (n_tiles + 1023) / 1024;
    const int lo = tid * per, hi = min(lo + per, n_tiles);
    int32_t m = total;
    for (int i = hi - 1; i >= lo; --i) m = min(m, offsets[i]);
    // suffix minimum over the per-thread minima: inside the warp by shuffles, across the 32 warps through one
    // shared-memory exchange (two barriers instead of twenty)
    int32_t suf = m;   // min over threads tid .. end of the warp
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const int32_t o = __shfl_down_sync(0xffffffffu, suf, d);
        if (lane + d < 32) suf = min(suf, o);
    }
    if (lane == 0) s_warp[w] = suf;
    __syncthreads();
    int32_t later = total;   // min over the warps behind this one
    for (int ww = w + 1; ww < 32; ++ww) later = min(later, s_warp[ww]);
    // what follows this thread: the rest of its warp, then the later warps
    int32_t next_in_warp = __shfl_down_sync(0xffffffffu, suf, 1);
    if (lane == 31) next_in_warp = total;
    int32_t run = min(next_in_warp, later);
    for (int i = hi - 1; i >= lo; --i) {
        run = min(run, offsets[i]);
        offsets[i] = run;
    }
}

// ------------------------------------------------------------------ offsets (A4)
// 4 consecutive keys per thread (two 16-byte loads); only the high words (camera | tile) matter
__global__ void __launch_bounds__(256)
isect_offset_encode_kernel(int64_t n_isects, const int64_t* __restrict__ isect_ids, int n_tiles_total,
                           int n_tiles, int tile_n_bits, int32_t* __restrict__ offsets) {
    const int64_t i0 = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) * 4;
    if (i0 >= n_isects) return;
    const uint32_t mask = (1u << tile_n_bits) - 1u;
    uint32_t hi[4];
    if (i0 + 4 <= n_isects) {
        const int4 a = __ldg(reinterpret_cast<const int4*>(isect_ids + i0));
        const int4 b = __ldg(reinterpret_cast<const int4*>(isect_ids + i0 + 2));
        hi[0] = (uint32_t)a.y; hi[1] = (uint32_t)a.w; hi[2] = (uint32_t)b.y; hi[3] = (uint32_t)b.w;
    } else {
#pragma unroll
        for (int k = 0; k < 4; ++k)
            hi[k] = (i0 + k < n_isects) ? (uint32_t)(__ldg(isect_ids + i0 + k) >> 32) : 0u;
    }
    int64_t slot_p = -1;  // slot of the previous key (-1 before the first)
    if (i0 > 0) {
        const uint32_t hp = (uint32_t)(__ldg(isect_ids + i0 - 1) >> 32);
        slot_p = (int64_t)(hp >> tile_n_bits) * n_tiles + (hp & mask);
    }
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        const int64_t i = i0 + k;
        if (i < n_isects) {
            const int64_t slot = (int64_t)(hi[k] >> tile_n_bits) * n_tiles + (hi[k] & mask);
            for (int64_t s = slot_p + 1; s <= slot; ++s) offsets[s] = (int32_t)i;  // i == 0: zeros up to slot
            slot_p = slot;
            if (i == n_isects - 1)
                for (int64_t s = slot + 1; s < n_tiles_total; ++s) offsets[s] = (int32_t)n_isects;
        }
    }
}

}  // namespace

static int projection_fwd_launch(int64_t N, int C, const float* means, const float* quats, const float* scales,
                                 const float* viewmats, const float* Ks, int width, int height, float eps2d,
                                 float near_plane, float far_plane, float radius_clip, int32_t* radii,
                                 float* means2d, float* depths, float* conics, float* compensations,
                                 const float* opacities, float* op_out, uint8_t* vis_out, float* radii_norm,
                                 void* stream) {
    SC_CHECK_ARG(N >= 0 && C >= 1, "N, C");
    SC_CHECK_ARG(width > 0 && height > 0, "image size");
    if (N == 0) return SC_OK;
    SC_CHECK_ARG(means && quats && scales && viewmats && Ks && radii && means2d && depths && conics,
                 "null pointer");
    SC_CHECK_ARG((opacities == nullptr) == (op_out == nullptr), "opacities and op_out go together");
    SC_CHECK_ARG(((uintptr_t)means & 15) == 0 && ((uintptr_t)scales & 15) == 0 &&
                     ((uintptr_t)quats & 15) == 0 && ((uintptr_t)means2d & 7) == 0 &&
                     ((uintptr_t)conics & 15) == 0,
                 "16-byte alignment of means/quats/scales/conics");
    dim3 grid((unsigned)sc_div_up(N, PF_CHUNK), (unsigned)C);
    const ProjOut O{radii, means2d, depths, conics, compensations, opacities, op_out, vis_out, radii_norm,
                    1.0f / (float)(width > height ? width : height)};
    projection_fwd_kernel<<<grid, PROJ_THREADS, 0, (cudaStream_t)stream>>>(
        N, means, quats, scales, viewmats, Ks, width, height, eps2d, near_plane, far_plane, radius_clip, O);
    SC_CHECK_LAUNCH("projection_fwd_kernel");
    return SC_OK;
}

extern "C" int sc_projection_fwd(int64_t N, int C, const float* means, const float* quats,
                                 const float* scales, const float* viewmats, const float* Ks, int width,
                                 int height, float eps2d, float near_plane, float far_plane,
                                 float radius_clip, int32_t* radii, float* means2d, float* depths,
                                 float* conics, float* compensations, void* stream) {
    return projection_fwd_launch(N, C, means, quats, scales, viewmats, Ks, width, height, eps2d, near_plane,
                                 far_plane, radius_clip, radii, means2d, depths, conics, compensations, nullptr,
                                 nullptr, nullptr, nullptr, stream);
}

extern "C" int sc_projection_fwd_opacity(int64_t N, int C, const float* means, const float* quats,
                                         const float* scales, const float* opacities, const float* viewmats,
                                         const float* Ks, int width, int height, float eps2d, float near_plane,
                                         float far_plane, float radius_clip, int32_t* radii, float* means2d,
                                         float* depths, float* conics, float* compensations,
                                         float* opacities_out, uint8_t* visibility, float* radii_norm,
                                         void* stream) {
    SC_CHECK_ARG(opacities && opacities_out, "null pointer");
    return projection_fwd_launch(N, C, means, quats, scales, viewmats, Ks, width, height, eps2d, near_plane,
                                 far_plane, radius_clip, radii, means2d, depths, conics, compensations,
                                 opacities, opacities_out, visibility, radii_norm, stream);
}

extern "C" int sc_isect_count(int64_t CN, const float* means2d, const int32_t* radii, const float* depths,
                              int tile_size, int tile_width, int tile_height, int32_t* tiles_per_gauss,
                              uint32_t* depth_keys, uint32_t* order_init, void* stream) {
    SC_CHECK_ARG(CN >= 0 && CN < (1ll << 31) && tile_size > 0 && tile_width > 0 && tile_height > 0, "sizes");
    if (CN == 0) return SC_OK;
    SC_CHECK_ARG(means2d && radii && tiles_per_gauss, "null pointer");
    SC_CHECK_ARG((depth_keys == nullptr) == (order_init == nullptr), "depth_keys and order_init go together");
    SC_CHECK_ARG(depth_keys == nullptr || depths != nullptr, "depths needed for depth_keys");
    isect_count_kernel<<<(unsigned)sc_div_up(CN, 256), 256, 0, (cudaStream_t)stream>>>(
        CN, reinterpret_cast<const float2*>(means2d), radii, depths, (float)tile_size, tile_width, tile_height,
        tiles_per_gauss, depth_keys, order_init);
    SC_CHECK_LAUNCH("isect_count_kernel");
    return SC_OK;
}

extern "C" int sc_isect_gather_counts(int64_t CN, const uint32_t* order, const int32_t* tiles_per_gauss,
                                      int32_t* counts_in_order, void* stream) {
    SC_CHECK_ARG(CN >= 0, "CN");
    if (CN == 0) return SC_OK;
    SC_CHECK_ARG(order && tiles_per_gauss && counts_in_order, "null pointer");
    isect_gather_counts_kernel<<<(unsigned)sc_div_up(CN, 256), 256, 0, (cudaStream_t)stream>>>(
        CN, order, tiles_per_gauss, counts_in_order);
    SC_CHECK_LAUNCH("isect_gather_counts_kernel");
    return SC_OK;
}

extern "C" int sc_isect_emit(int64_t N, int C, const float* means2d, const int32_t* radii,
                             const float* depths, const int64_t* cum_tiles, const uint32_t* order,
                             int tile_size, int tile_width, int tile_height, int tile_n_bits,
                             int64_t n_isects, int64_t* isect_ids, int32_t* flatten_ids, void* stream) {
    SC_CHECK_ARG(N >= 0 && C >= 1 && n_isects >= 0, "sizes");
    SC_CHECK_ARG(tile_n_bits > 0 && tile_n_bits < 24, "tile_n_bits");
    if (N == 0 || n_isects == 0) return SC_OK;
    SC_CHECK_ARG(means2d && radii && depths && cum_tiles && isect_ids && flatten_ids, "null pointer");
    const int64_t CN = N * C;
    SC_CHECK_ARG(CN < (1ll << 31), "C*N must fit int32 (flatten_ids)");
    isect_emit_kernel<<<(unsigned)sc_div_up(n_isects, EMIT_OUT), EMIT_THREADS, 0, (cudaStream_t)stream>>>(
        N, CN, reinterpret_cast<const float2*>(means2d), radii, depths, cum_tiles, order, (float)tile_size,
        tile_width, tile_height, tile_n_bits, n_isects, isect_ids, flatten_ids);
    SC_CHECK_LAUNCH("isect_emit_kernel");
    return SC_OK;
}

// ---- fused binning pipeline (see the kernels' header comment) ----------------------------
namespace {
struct IsectWs {
    ScSortPlan plan32;
    int64_t cc_tiles, sg_tiles, first_cap;
    int64_t totals_off, ticket_off, cc_state_off, sg_state_off, sort32_off, ctrl_bytes;
    int64_t dk_off, od_off, cum_off, first_off, total;
    // fast path (C == 1, tile grid <= 128 x 128)
    int64_t hist_xy_off, rects_off, gids_off, rect_s_off, gid_s_off;
    int64_t kb_off, coop_cnt_off, coop_tot_off, coop_hist_off, cumw_off, first_item_off, coop_dbg_off;   // cooperative prepare
};
IsectWs isect_ws_layout(int64_t CN, int64_t n_tiles_total) {
    IsectWs L;
    L.plan32 = sc_sort_make_plan(CN, 0, 32, 4);
    L.cc_tiles = sc_div_up(CN, CC_TILE);
    L.sg_tiles = sc_div_up(CN, SG_TILE);
    // n_isects < 2^31 (int32 flatten positions) bounds the number of emit blocks
    const int64_t max_isects = (CN > 0 && n_tiles_total > (1ll << 31) / CN) ? (1ll << 31) : CN * n_tiles_total;
    L.first_cap = sc_div_up(max_isects, EMIT_OUT) + 1;
    int64_t o = 0;
    L.totals_off = o; o += 256;
    L.ticket_off = o; o += 256;
    L.cc_state_off = o; o = sc_align_up(o + L.cc_tiles * 8, 256);
    L.sg_state_off = o; o = sc_align_up(o + L.sg_tiles * 8, 256);
    L.hist_xy_off = o; o += 2 * FB_RADIX * 4;
    L.sort32_off = o;
    L.ctrl_bytes = o + L.plan32.ctrl_bytes;
    o += L.plan32.total;
    L.dk_off = o; o = sc_align_up(o + 2 * CN * 4, 256);
    L.od_off = o; o = sc_align_up(o + 2 * CN * 4, 256);
    L.cum_off = o; o = sc_align_up(o + (CN + 1) * 8, 256);
    L.first_off = o; o = sc_align_up(o + L.first_cap * 4, 256);
    L.rects_off = o; o = sc_align_up(o + CN * 4, 256);
    L.gids_off = o; o = sc_align_up(o + CN * 4, 256);
    L.rect_s_off = o; o = sc_align_up(o + CN * 4, 256);
    L.gid_s_off = o; o = sc_align_up(o + CN * 4, 256);
    L.kb_off = o; o = sc_align_up(o + CN * 4, 256);
    L.coop_cnt_off = o; o += CP_MAX_CTAS * 4;
    L.coop_tot_off = o; o += CP_MAX_CTAS * 16;
    L.coop_dbg_off = o; o += 256;
    L.cumw_off = o; o = sc_align_up(o + (CN + 1) * 4, 256);
    L.first_item_off = o; o = sc_align_up(o + L.first_cap * 4, 256);
    L.coop_hist_off = o; o += (int64_t)CP_MAX_CTAS * 256 * 4;
    L.total = o;
    return L;
}
}  // namespace

extern "C" int64_t sc_isect_workspace_bytes(int64_t CN, int tile_width, int tile_height) {
    return isect_ws_layout(CN, (int64_t)tile_width * tile_height).total;
}

extern "C" int sc_isect_prepare(int64_t CN, const float* means2d, const int32_t* radii, const float* depths,
                                int tile_size, int tile_width, int tile_height, int32_t* tiles_per_gauss,
                                void* workspace, int64_t workspace_bytes, void* stream) {
    SC_CHECK_ARG(CN >= 0 && CN < (1ll << 30) && tile_size > 0 && tile_width > 0 && tile_height > 0, "sizes");
    SC_CHECK_ARG(workspace && ((uintptr_t)workspace & 255) == 0, "workspace (256-byte aligned)");
    const IsectWs L = isect_ws_layout(CN, (int64_t)tile_width * tile_height);
    if (workspace_bytes < L.total) {
        sc_set_error("isect workspace too small: %lld < %lld", (long long)workspace_bytes, (long long)L.total);
        return SC_ERR_WORKSPACE;
    }
    cudaStream_t st = (cudaStream_t)stream;
    unsigned char* ws = (unsigned char*)workspace;
    SC_CUDA(cudaMemsetAsync(ws, 0, (size_t)L.ctrl_bytes, st));
    if (CN == 0) return SC_OK;
    SC_CHECK_ARG(means2d && radii && depths && tiles_per_gauss, "null pointer");
    int64_t* totals = (int64_t*)(ws + L.totals_off);
    uint32_t* tickets = (uint32_t*)(ws + L.ticket_off);
    uint32_t* dk0 = (uint32_t*)(ws + L.dk_off);
    uint32_t* dk1 = dk0 + CN;
    uint32_t* od0 = (uint32_t*)(ws + L.od_off);
    uint32_t* od1 = od0 + CN;
    void* sws = ws + L.sort32_off;
    const unsigned cc_grid = (unsigned)std::min<int64_t>(L.cc_tiles, (int64_t)SC_NUM_SMS * CC_CTAS_PER_SM);
    isect_count_compact_kernel<<<cc_grid, CC_THREADS, 0, st>>>(
        CN, L.cc_tiles, reinterpret_cast<const float2*>(means2d), radii, depths, (float)tile_size, tile_width,
        tile_height, tiles_per_gauss, dk0, od0, (volatile unsigned long long*)(ws + L.cc_state_off),
        tickets + 0, totals, (uint32_t*)((unsigned char*)sws + L.plan32.hist_off));
    SC_CHECK_LAUNCH("isect_count_compact_kernel");
    int rc = sc_sort_run_u32(CN, totals, dk0, od0, dk1, od1, L.plan32, sws, st, /*hist_raw=*/true);
    if (rc < 0) return rc;
    const uint32_t* order = rc == 0 ? od1 : od0;
    isect_scan_kernel<<<(unsigned)L.sg_tiles, SG_THREADS, 0, st>>>(
        totals, order, tiles_per_gauss, (int64_t*)(ws + L.cum_off), (uint32_t*)(ws + L.first_off), L.first_cap,
        (volatile unsigned long long*)(ws + L.sg_state_off), tickets + 1, totals + 1);
    SC_CHECK_LAUNCH("isect_scan_kernel");
    return SC_OK;
}

extern "C" int sc_isect_emit_sorted(int64_t N, int C, const float* means2d, const int32_t* radii,
                                    const float* depths, int tile_size, int tile_width, int tile_height,
                                    int tile_n_bits, int cam_bits, int64_t n_visible, int64_t n_isects,
                                    const void* prepare_workspace, int64_t* isect_ids, int32_t* flatten_ids,
                                    int64_t* tmp_ids, int32_t* tmp_flatten, void* sort_workspace,
                                    int64_t sort_workspace_bytes, void* stream) {
    SC_CHECK_ARG(N >= 0 && C >= 1 && n_isects >= 0 && n_visible >= 0 && n_visible <= N * C, "sizes");
    SC_CHECK_ARG(tile_n_bits > 0 && cam_bits >= 0 && tile_n_bits + cam_bits <= 32, "key bits");
    SC_CHECK_ARG(n_isects < (1ll << 30), "n_isects (must be < 2^30)");
    if (n_isects == 0) return SC_OK;
    SC_CHECK_ARG(means2d && radii && depths && prepare_workspace && isect_ids && flatten_ids && tmp_ids &&
                     tmp_flatten && sort_workspace, "null pointer");
    const int64_t CN = N * C;
    const IsectWs L = isect_ws_layout(CN, (int64_t)tile_width * tile_height);
    SC_CHECK_ARG(sc_div_up(n_isects, EMIT_OUT) <= L.first_cap, "n_isects exceeds the entry table");
    const ScSortPlan plan = sc_sort_make_plan(n_isects, 32, 32 + tile_n_bits + cam_bits, 8);
    if (sort_workspace_bytes < plan.total) {
        sc_set_error("sort workspace too small: %lld < %lld", (long long)sort_workspace_bytes,
                     (long long)plan.total);
        return SC_ERR_WORKSPACE;
    }
    cudaStream_t st = (cudaStream_t)stream;
    const unsigned char* pws = (const unsigned char*)prepare_workspace;
    const uint32_t* od0 = (const uint32_t*)(pws + L.od_off);
    const uint32_t* order = (L.plan32.pp.passes % 2 == 0) ? od0 : od0 + CN;
    SC_CUDA(cudaMemsetAsync(sort_workspace, 0, (size_t)plan.ctrl_bytes, st));
    // emit into the buffer from which `passes` ping-pong steps end in (isect_ids, flatten_ids)
    const bool into_tmp = (plan.pp.passes % 2) != 0;
    uint64_t* k_a = (uint64_t*)(into_tmp ? tmp_ids : isect_ids);
    uint32_t* v_a = (uint32_t*)(into_tmp ? tmp_flatten : flatten_ids);
    uint64_t* k_b = (uint64_t*)(into_tmp ? isect_ids : tmp_ids);
    uint32_t* v_b = (uint32_t*)(into_tmp ? flatten_ids : tmp_flatten);
    isect_emit_hist_kernel<<<(unsigned)sc_div_up(n_isects, EMIT_OUT), EMIT_THREADS, 0, st>>>(
        N, n_visible, reinterpret_cast<const float2*>(means2d), radii, depths,
        (const int64_t*)(pws + L.cum_off), order, (const uint32_t*)(pws + L.first_off), (float)tile_size,
        tile_width, tile_height, tile_n_bits, n_isects, (int64_t*)k_a, (int32_t*)v_a, plan.pp,
        (uint32_t*)((unsigned char*)sort_workspace + plan.hist_off));
    SC_CHECK_LAUNCH("isect_emit_hist_kernel");
    const int rc = sc_sort_run_u64(n_isects, nullptr, k_a, v_a, k_b, v_b, plan, sort_workspace, st);
    if (rc < 0) return rc;
    const uint64_t* res = rc == 0 ? k_b : k_a;
    if (res != (const uint64_t*)isect_ids) {
        sc_set_error("internal: sorted list ended in the wrong buffer");
        return SC_ERR_CUDA;
    }
    return SC_OK;
}

// ---- fast binning path (C == 1, tile grid <= 128 x 128): see the kernels' header comment ----
namespace {
struct BinSortWs {
    int64_t tiles, item_tiles_max, ticket_off, lookback_off, lookback2_off, ctrl_bytes, pairs_off, total;
};
BinSortWs bin_sort_layout(int64_t n_isects) {
    BinSortWs L;
    L.tiles = sc_div_up(n_isects, FB_TILE);
    L.item_tiles_max = sc_div_up(n_isects, FI_TILE);   // items <= intersections
    L.ticket_off = 0;
    L.lookback_off = 256;
    L.lookback2_off = L.lookback_off + L.item_tiles_max * FB_RADIX * 4;
    L.ctrl_bytes = L.lookback2_off + L.tiles * FB_RADIX * 4;
    L.pairs_off = sc_align_up(L.ctrl_bytes, 256);
    L.total = sc_align_up(L.pairs_off + n_isects * 8, 256);
    return L;
}
}  // namespace

namespace {
int g_reserved_sms = 0;   // sc_bin_reserve_sms

// CTAs of the cooperative prepare kernel on the current device (CP_CTAS_PER_SM per SM, minus the SMs left to
// concurrent kernels), 0 when cooperative launch is not available; device properties cached per device
int coop_grid_size() {
    static int sms_of[64], per_sm_of[64];
    static bool known[64];
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 0;
    if (!known[dev]) {
        int coop = 0, sms = 0, per_sm = 0;
        cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, dev);
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, bin_prepare_coop_kernel, CP_THREADS, 0);
        sms_of[dev] = (coop && per_sm >= 1) ? sms : 0;
        per_sm_of[dev] = std::min(per_sm, CP_CTAS_PER_SM);
        known[dev] = true;
    }
    const int sms = std::max(sms_of[dev] - g_reserved_sms, sms_of[dev] > 0 ? 8 : 0);
    return std::min(sms * per_sm_of[dev], CP_MAX_CTAS);
}
}  // namespace

// A cooperative launch needs all its CTAs resident at once: while another kernel (an NCCL collective on a side
// stream) holds SMs, a grid sized for the whole device cannot start until that kernel ends — which would
// serialise exactly the transfer it is meant to overlap.  Leaves `n_sms` SMs out of the cooperative grid.
extern "C" int sc_bin_reserve_sms(int n_sms) {
    SC_CHECK_ARG(n_sms >= 0 && n_sms < 1024, "n_sms");
    g_reserved_sms = n_sms;
    return SC_OK;
}

extern "C" int sc_bin_supported(int C, int tile_width, int tile_height) {
    return (C == 1 && tile_width <= FB_RADIX && tile_height <= FB_RADIX && coop_grid_size() > 0) ? 1 : 0;
}

extern "C" int sc_bin_prepare(int64_t N, const float* means2d, const int32_t* radii, const float* depths,
                              int tile_size, int tile_width, int tile_height, int32_t* tiles_per_gauss,
                              void* workspace, int64_t workspace_bytes, void* stream) {
    SC_CHECK_ARG(N >= 0 && N < (1ll << 30) && tile_size > 0 && tile_width > 0 && tile_height > 0, "sizes");
    SC_CHECK_ARG(tile_width <= FB_RADIX && tile_height <= FB_RADIX, "tile grid (at most 128 x 128 on the fast path)");
    SC_CHECK_ARG(workspace && ((uintptr_t)workspace & 255) == 0, "workspace (256-byte aligned)");
    const IsectWs L = isect_ws_layout(N, (int64_t)tile_width * tile_height);
    if (workspace_bytes < L.total) {
        sc_set_error("isect workspace too small: %lld < %lld", (long long)workspace_bytes, (long long)L.total);
        return SC_ERR_WORKSPACE;
    }
    cudaStream_t st = (cudaStream_t)stream;
    unsigned char* ws = (unsigned char*)workspace;
    SC_CUDA(cudaMemsetAsync(ws, 0, (size_t)L.ctrl_bytes, st));
    if (N == 0) return SC_OK;
    SC_CHECK_ARG(means2d && radii && depths && tiles_per_gauss, "null pointer");
    int64_t* totals = (int64_t*)(ws + L.totals_off);
    uint32_t* tickets = (uint32_t*)(ws + L.ticket_off);
    uint32_t* dk0 = (uint32_t*)(ws + L.dk_off);
    uint32_t* dk1 = dk0 + N;
    uint32_t* od0 = (uint32_t*)(ws + L.od_off);
    uint32_t* od1 = od0 + N;
    const int coop_grid = coop_grid_size();
    if (coop_grid > 0) {
        CoopArgs a;
        a.N = N;
        a.means2d = reinterpret_cast<const float2*>(means2d);
        a.radii = radii;
        a.depths = depths;
        a.ts = (float)tile_size;
        a.tw = tile_width;
        a.th = tile_height;
        a.tiles_per_gauss = tiles_per_gauss;
        a.stage_key = dk0;
        a.stage_gid = (uint32_t*)(ws + L.gids_off);
        a.stage_rect = (uint32_t*)(ws + L.rects_off);
        a.keyA = dk1;
        a.valA = od0;
        a.keyB = (uint32_t*)(ws + L.kb_off);
        a.valB = od1;
        a.cnt_g = (uint32_t*)(ws + L.coop_cnt_off);
        a.tot_g = (unsigned long long*)(ws + L.coop_tot_off);
        a.hist_g = (uint32_t*)(ws + L.coop_hist_off);
        a.totals = totals;
        a.cum = (int64_t*)(ws + L.cum_off);
        a.cumw = (uint32_t*)(ws + L.cumw_off);
        a.first_item = (uint32_t*)(ws + L.first_item_off);
        a.rect_s = (uint32_t*)(ws + L.rect_s_off);
        a.gid_s = (uint32_t*)(ws + L.gid_s_off);
        a.first_entry = (uint32_t*)(ws + L.first_off);
        a.first_cap = L.first_cap;
        a.hist_xy = (uint32_t*)(ws + L.hist_xy_off);
        a.phase_ns = (unsigned long long*)(ws + L.coop_dbg_off);
        void* args[] = {&a};
        SC_CUDA(cudaLaunchCooperativeKernel((const void*)bin_prepare_coop_kernel, dim3((unsigned)coop_grid),
                                            dim3(CP_THREADS), args, 0, st));
        SC_CHECK_LAUNCH("bin_prepare_coop_kernel");
        return SC_OK;
    }
    sc_set_error("cooperative launch is not available on this device (sc_bin_supported() == 0)");
    return SC_ERR_CUDA;
}

extern "C" int64_t sc_bin_phase_stamps_offset(int64_t N, int tile_width, int tile_height) {
    return isect_ws_layout(N, (int64_t)tile_width * tile_height).coop_dbg_off;
}

extern "C" int64_t sc_bin_sort_workspace_bytes(int64_t n_isects) { return bin_sort_layout(n_isects).total; }

// shared body of sc_bin_finish (counts known on the host) and sc_bin_finish_dev (counts on the device, outputs and
// workspace sized for `n_isects` = the caller's capacity)
static int bin_finish_impl(int64_t N, const float* depths, int tile_width, int tile_height, int64_t n_visible,
                           int64_t n_isects, bool dev_counts, const void* prepare_workspace, int32_t* flatten_ids,
                           int64_t* isect_ids, int32_t* offsets, void* sort_workspace, int64_t sort_workspace_bytes,
                           cudaStream_t st) {
    const int n_tiles = tile_width * tile_height;
    const IsectWs L = isect_ws_layout(N, (int64_t)n_tiles);
    const BinSortWs B = bin_sort_layout(n_isects);
    SC_CHECK_ARG(B.tiles + 1 <= L.first_cap, "n_isects exceeds the entry table");
    if (sort_workspace_bytes < B.total) {
        sc_set_error("bin sort workspace too small: %lld < %lld", (long long)sort_workspace_bytes, (long long)B.total);
        return SC_ERR_WORKSPACE;
    }
    const unsigned char* pws = (const unsigned char*)prepare_workspace;
    unsigned char* sws = (unsigned char*)sort_workspace;
    SC_CUDA(cudaMemsetAsync(sws, 0, (size_t)B.ctrl_bytes, st));
    SC_CUDA(cudaMemsetAsync(offsets, 0x7f, sizeof(int32_t) * (size_t)n_tiles, st));
    static bool attr_set = false;
    if (!attr_set) {
        SC_CUDA(cudaFuncSetAttribute(isect_emit_sort_tx_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                     (int)sizeof(FeSmem)));
        SC_CUDA(cudaFuncSetAttribute(isect_sort_ty_final_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                     (int)sizeof(FbSmem)));
        attr_set = true;
    }
    uint32_t* tickets = (uint32_t*)(sws + B.ticket_off);
    uint64_t* pairs = (uint64_t*)(sws + B.pairs_off);
    const uint32_t* hist_xy = (const uint32_t*)(pws + L.hist_xy_off);
    const int64_t* totals_dev = dev_counts ? (const int64_t*)(pws + L.totals_off) : nullptr;
    // (an ITEM-wise variant of this kernel — one weighted rank per (Gaussian, tile column) run instead of one
    // per slot, every lane writing its run — measured 2x SLOWER: 254 us vs 120 us on 13.6 M intersections; its
    // scattered 8-byte stores and the divergent run loops cost more than the ranking work it saves)
    isect_emit_sort_tx_kernel<<<(unsigned)B.tiles, FB_THREADS, sizeof(FeSmem), st>>>(
        n_visible, n_isects, totals_dev, (const int64_t*)(pws + L.cum_off), (const uint32_t*)(pws + L.rect_s_off),
        (const uint32_t*)(pws + L.gid_s_off), (const uint32_t*)(pws + L.first_off), hist_xy,
        (uint32_t*)(sws + B.lookback_off), tickets + 0, pairs);
    SC_CHECK_LAUNCH("isect_emit_sort_tx_kernel");
    isect_sort_ty_final_kernel<<<(unsigned)B.tiles, FB_THREADS, sizeof(FbSmem), st>>>(
        n_isects, totals_dev, pairs, hist_xy + FB_RADIX, (uint32_t*)(sws + B.lookback2_off), tickets + 1, depths,
        tile_width, flatten_ids, isect_ids, offsets);
    SC_CHECK_LAUNCH("isect_sort_ty_final_kernel");
    isect_offsets_fixup_kernel<<<1, 1024, 0, st>>>(n_tiles, totals_dev ? totals_dev + 1 : nullptr, n_isects, offsets);
    SC_CHECK_LAUNCH("isect_offsets_fixup_kernel");
    return SC_OK;
}

extern "C" int sc_bin_finish(int64_t N, const float* depths, int tile_width, int tile_height, int64_t n_visible,
                             int64_t n_isects, int64_t n_items, const void* prepare_workspace, int32_t* flatten_ids,
                             int64_t* isect_ids, int32_t* offsets, void* sort_workspace,
                             int64_t sort_workspace_bytes, void* stream) {
    SC_CHECK_ARG(N >= 0 && n_isects >= 0 && n_visible >= 0 && n_visible <= N, "sizes");
    SC_CHECK_ARG(n_items >= n_visible && n_items <= n_isects, "n_items");
    SC_CHECK_ARG(n_isects < (1ll << 30), "n_isects (must be < 2^30)");
    SC_CHECK_ARG(tile_width > 0 && tile_height > 0 && tile_width <= FB_RADIX && tile_height <= FB_RADIX, "tile grid");
    SC_CHECK_ARG(offsets, "null pointer");
    cudaStream_t st = (cudaStream_t)stream;
    if (n_isects == 0) {
        SC_CUDA(cudaMemsetAsync(offsets, 0, sizeof(int32_t) * (size_t)(tile_width * tile_height), st));
        return SC_OK;
    }
    SC_CHECK_ARG(depths && prepare_workspace && flatten_ids && sort_workspace, "null pointer");
    return bin_finish_impl(N, depths, tile_width, tile_height, n_visible, n_isects, false, prepare_workspace, flatten_ids,
                           isect_ids, offsets, sort_workspace, sort_workspace_bytes, st);
}

// Sync-free form: no host read-back between the two halves of the binning.  The counts {n_visible, n_isects,
// n_items} stay in the first 24 bytes of the prepare workspace; `capacity` bounds the list the outputs (and the sort
// workspace, sc_bin_sort_workspace_bytes(capacity)) can hold.  A list longer than the capacity is DROPPED (all tile
// offsets 0: the view renders empty, nothing is accessed out of bounds): the caller compares the counts with its
// capacity once the stream has passed — e.g. one step later — and re-runs with a larger one.  Nothing here depends on a value the
// host has not got yet, so a camera pass built on it can be captured in a CUDA graph.
extern "C" int sc_bin_finish_dev(int64_t N, const float* depths, int tile_width, int tile_height, int64_t capacity,
                                 const void* prepare_workspace, int32_t* flatten_ids, int64_t* isect_ids,
                                 int32_t* offsets, void* sort_workspace, int64_t sort_workspace_bytes, void* stream) {
    SC_CHECK_ARG(N >= 0 && capacity >= 1 && capacity < (1ll << 30), "sizes (1 <= capacity < 2^30)");
    SC_CHECK_ARG(tile_width > 0 && tile_height > 0 && tile_width <= FB_RADIX && tile_height <= FB_RADIX, "tile grid");
    SC_CHECK_ARG(depths && prepare_workspace && flatten_ids && sort_workspace && offsets, "null pointer");
    return bin_finish_impl(N, depths, tile_width, tile_height, 0, capacity, true, prepare_workspace, flatten_ids,
                           isect_ids, offsets, sort_workspace, sort_workspace_bytes, (cudaStream_t)stream);
}

extern "C" int sc_isect_offset_encode(int64_t n_isects, const int64_t* isect_ids, int C, int tile_width,
                                      int tile_height, int tile_n_bits, int32_t* offsets, void* stream) {
    SC_CHECK_ARG(n_isects >= 0 && C >= 1 && tile_width > 0 && tile_height > 0, "sizes");
    SC_CHECK_ARG(offsets, "null pointer");
    const int n_tiles = tile_width * tile_height;
    if (n_isects == 0) {
        SC_CUDA(cudaMemsetAsync(offsets, 0, sizeof(int32_t) * (size_t)n_tiles * C, (cudaStream_t)stream));
        return SC_OK;
    }
    SC_CHECK_ARG(isect_ids, "null pointer");
    SC_CHECK_ARG(((uintptr_t)isect_ids & 15) == 0, "16-byte alignment of isect_ids");
    isect_offset_encode_kernel<<<(unsigned)sc_div_up(n_isects, 1024), 256, 0, (cudaStream_t)stream>>>(
        n_isects, isect_ids, n_tiles * C, n_tiles, tile_n_bits, offsets);
    SC_CHECK_LAUNCH("isect_offset_encode_kernel");
    return SC_OK;
}
