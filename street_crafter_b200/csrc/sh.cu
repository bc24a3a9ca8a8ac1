// Spherical-harmonics colour evaluation (A5) and its backward (A8b).
//
// HBM-bound: the [M,K,3] coefficient block is 192 B per Gaussian at K=16.  Lane-per-coefficient
// layout (see ShLayout below): a group of lanes owns one Gaussian, lane k its k-th coefficient,
// so the group's accesses cover the Gaussian's contiguous span with fully used sectors, without
// shared memory; masked-out Gaussians cost a predicate and no traffic.
//
// Replaces gsplat.spherical_harmonics as called at
// /root/reference/street_gaussian/models/street_gaussian_renderer.py:256-260; basis signs and
// constants follow /root/reference/street_gaussian/utils/sh_utils.py:24-110.
#include "common.cuh"

namespace {

constexpr int SH_WARPS = 4;
constexpr int SH_THREADS = SH_WARPS * 32;

#define C0 0.28209479177387814f
#define C1 0.4886025119029199f
#define C20 1.0925484305920792f
#define C21 -1.0925484305920792f
#define C22 0.31539156525252005f
#define C23 -1.0925484305920792f
#define C24 0.5462742152960396f
#define C30 -0.5900435899266435f
#define C31 2.890611442640554f
#define C32 -0.4570457994644658f
#define C33 0.3731763325901154f
#define C34 -0.4570457994644658f
#define C35 1.445305721320277f
#define C36 -0.5900435899266435f
#define C40 2.5033429417967046f
#define C41 -1.7701307697799304f
#define C42 0.9461746957575601f
#define C43 -0.6690465435572892f
#define C44 0.10578554691520431f
#define C45 -0.6690465435572892f
#define C46 0.47308734787878004f
#define C47 -1.7701307697799304f
#define C48 0.6258357354491761f

template <int DEG>
__device__ __forceinline__ void sh_basis(float x, float y, float z, float* Y) {
    Y[0] = C0;
    if (DEG > 0) {
        Y[1] = -C1 * y; Y[2] = C1 * z; Y[3] = -C1 * x;
    }
    if (DEG > 1) {
        const float xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
        Y[4] = C20 * xy; Y[5] = C21 * yz; Y[6] = C22 * (2.f * zz - xx - yy); Y[7] = C23 * xz;
        Y[8] = C24 * (xx - yy);
        if (DEG > 2) {
            Y[9] = C30 * y * (3.f * xx - yy); Y[10] = C31 * xy * z; Y[11] = C32 * y * (4.f * zz - xx - yy);
            Y[12] = C33 * z * (2.f * zz - 3.f * xx - 3.f * yy); Y[13] = C34 * x * (4.f * zz - xx - yy);
            Y[14] = C35 * z * (xx - yy); Y[15] = C36 * x * (xx - 3.f * yy);
        }
        if (DEG > 3) {
            Y[16] = C40 * xy * (xx - yy); Y[17] = C41 * yz * (3.f * xx - yy); Y[18] = C42 * xy * (7.f * zz - 1.f);
            Y[19] = C43 * yz * (7.f * zz - 3.f); Y[20] = C44 * (zz * (35.f * zz - 30.f) + 3.f);
            Y[21] = C45 * xz * (7.f * zz - 3.f); Y[22] = C46 * (xx - yy) * (7.f * zz - 1.f);
            Y[23] = C47 * xz * (xx - 3.f * yy); Y[24] = C48 * (xx * (xx - 3.f * yy) - yy * (3.f * xx - yy));
        }
    }
}

// gradient of sum_k s[k]*Y_k wrt (x,y,z) treating them as independent variables
template <int DEG>
__device__ __forceinline__ void sh_basis_vjp(float x, float y, float z, const float* s, float& gx,
                                             float& gy, float& gz) {
    gx = gy = gz = 0.f;
    if (DEG > 0) {
        gy += -C1 * s[1]; gz += C1 * s[2]; gx += -C1 * s[3];
    }
    if (DEG > 1) {
        const float xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
        gx += C20 * y * s[4] + C22 * (-2.f * x) * s[6] + C23 * z * s[7] + C24 * 2.f * x * s[8];
        gy += C20 * x * s[4] + C21 * z * s[5] + C22 * (-2.f * y) * s[6] + C24 * (-2.f * y) * s[8];
        gz += C21 * y * s[5] + C22 * 4.f * z * s[6] + C23 * x * s[7];
        if (DEG > 2) {
            gx += C30 * 6.f * xy * s[9] + C31 * yz * s[10] + C32 * (-2.f * xy) * s[11] + C33 * (-6.f * xz) * s[12] +
                  C34 * (4.f * zz - 3.f * xx - yy) * s[13] + C35 * 2.f * xz * s[14] + C36 * (3.f * xx - 3.f * yy) * s[15];
            gy += C30 * (3.f * xx - 3.f * yy) * s[9] + C31 * xz * s[10] + C32 * (4.f * zz - xx - 3.f * yy) * s[11] +
                  C33 * (-6.f * yz) * s[12] + C34 * (-2.f * xy) * s[13] + C35 * (-2.f * yz) * s[14] +
                  C36 * (-6.f * xy) * s[15];
            gz += C31 * xy * s[10] + C32 * 8.f * yz * s[11] + C33 * (6.f * zz - 3.f * xx - 3.f * yy) * s[12] +
                  C34 * 8.f * xz * s[13] + C35 * (xx - yy) * s[14];
        }
        if (DEG > 3) {
            gx += C40 * y * (3.f * xx - yy) * s[16] + C41 * 6.f * xy * z * s[17] + C42 * y * (7.f * zz - 1.f) * s[18] +
                  C45 * z * (7.f * zz - 3.f) * s[21] + C46 * 2.f * x * (7.f * zz - 1.f) * s[22] +
                  C47 * z * (3.f * xx - 3.f * yy) * s[23] + C48 * (4.f * xx * x - 12.f * x * yy) * s[24];
            gy += C40 * x * (xx - 3.f * yy) * s[16] + C41 * z * (3.f * xx - 3.f * yy) * s[17] +
                  C42 * x * (7.f * zz - 1.f) * s[18] + C43 * z * (7.f * zz - 3.f) * s[19] +
                  C46 * (-2.f * y) * (7.f * zz - 1.f) * s[22] + C47 * (-6.f * xy * z) * s[23] +
                  C48 * (-12.f * xx * y + 4.f * yy * y) * s[24];
            gz += C41 * y * (3.f * xx - yy) * s[17] + C42 * 14.f * xy * z * s[18] + C43 * y * (21.f * zz - 3.f) * s[19] +
                  C44 * (140.f * zz * z - 60.f * z) * s[20] + C45 * x * (21.f * zz - 3.f) * s[21] +
                  C46 * (xx - yy) * 14.f * z * s[22] + C47 * x * (xx - 3.f * yy) * s[23];
        }
    }
}

// ---------------------------------------------------------------------------------------
// Lane-per-coefficient layout.  A group of LPG = pow2 >= K lanes owns ONE Gaussian, lane k of
// the group owns coefficient k (its 3 channels, 12 contiguous bytes): the group's accesses to
// coeffs / v_coeffs cover the Gaussian's contiguous K*12-byte span (K = 16: 192 B per half
// warp, fully used sectors), no shared memory, no transposes, ~40 registers of state, and a
// masked-out Gaussian costs one predicate.  Every lane evaluates the whole basis (a few dozen
// FMULs, redundant but free next to the HBM traffic) and keeps Y_k; colours are a 3-value
// shuffle reduction over the group.  Two Gaussians per group are in flight per iteration.
template <int K>
struct ShLayout {
    static constexpr int LPG = K <= 1 ? 1 : K <= 4 ? 4 : K <= 16 ? 16 : 32;  // lanes per Gaussian
    static constexpr int GPW = 32 / LPG;                                      // Gaussians per warp
};
constexpr int SH_UNROLL = 2;

template <int NB>
__device__ __forceinline__ float select_k(const float* Y, int k) {
    float y = 0.f;
#pragma unroll
    for (int i = 0; i < NB; ++i) y = (i == k) ? Y[i] : y;
    return y;
}

template <int K, int DEG>
__global__ void __launch_bounds__(SH_THREADS)
sh_fwd_kernel(int64_t M, const float* __restrict__ dirs, const float* __restrict__ coeffs,
              const uint8_t* __restrict__ masks, float* __restrict__ colors) {
    constexpr int NB = (DEG + 1) * (DEG + 1);
    constexpr int LPG = ShLayout<K>::LPG, GPW = ShLayout<K>::GPW;
    const int lane = threadIdx.x & 31;
    const int k = lane % LPG, grp = lane / LPG;
    const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int64_t n_warps = ((int64_t)gridDim.x * blockDim.x) >> 5;
    const int64_t n_iter = (M + GPW * SH_UNROLL - 1) / (GPW * SH_UNROLL);
    for (int64_t it = warp; it < n_iter; it += n_warps) {
        int64_t g[SH_UNROLL];
        bool live[SH_UNROLL];
        float dx[SH_UNROLL], dy[SH_UNROLL], dz[SH_UNROLL], c0[SH_UNROLL], c1[SH_UNROLL], c2[SH_UNROLL];
#pragma unroll
        for (int u = 0; u < SH_UNROLL; ++u) {
            g[u] = (it * SH_UNROLL + u) * GPW + grp;
            live[u] = (g[u] < M) && (masks == nullptr || __ldg(masks + g[u]) != 0);
        }
#pragma unroll
        for (int u = 0; u < SH_UNROLL; ++u) {
            dx[u] = dy[u] = dz[u] = c0[u] = c1[u] = c2[u] = 0.f;
            if (live[u]) {
                dx[u] = __ldg(dirs + g[u] * 3 + 0); dy[u] = __ldg(dirs + g[u] * 3 + 1); dz[u] = __ldg(dirs + g[u] * 3 + 2);
                if (k < NB) {
                    const float* c = coeffs + (g[u] * K + k) * 3;
                    c0[u] = __ldg(c + 0); c1[u] = __ldg(c + 1); c2[u] = __ldg(c + 2);
                }
            }
        }
#pragma unroll
        for (int u = 0; u < SH_UNROLL; ++u) {
            float r = 0.f, gg = 0.f, b = 0.f;
            if (__any_sync(0xffffffffu, live[u])) {
                const float inv = 1.0f / fmaxf(sqrtf(dx[u] * dx[u] + dy[u] * dy[u] + dz[u] * dz[u]), 1e-12f);
                float Y[25];
                sh_basis<DEG>(dx[u] * inv, dy[u] * inv, dz[u] * inv, Y);
                const float yk = select_k<NB>(Y, k);
                r = yk * c0[u]; gg = yk * c1[u]; b = yk * c2[u];
#pragma unroll
                for (int o = LPG / 2; o >= 1; o >>= 1) {
                    r += __shfl_xor_sync(0xffffffffu, r, o);
                    gg += __shfl_xor_sync(0xffffffffu, gg, o);
                    b += __shfl_xor_sync(0xffffffffu, b, o);
                }
            }
            if (g[u] < M) {
                if (LPG >= 4) {
                    if (k < 3) colors[g[u] * 3 + k] = live[u] ? (k == 0 ? r : k == 1 ? gg : b) : 0.f;
                } else {
                    colors[g[u] * 3 + 0] = live[u] ? r : 0.f;
                    colors[g[u] * 3 + 1] = live[u] ? gg : 0.f;
                    colors[g[u] * 3 + 2] = live[u] ? b : 0.f;
                }
            }
        }
    }
}

template <int K, int DEG>
__global__ void __launch_bounds__(SH_THREADS)
sh_bwd_kernel(int64_t M, const float* __restrict__ dirs, const float* __restrict__ coeffs,
              const uint8_t* __restrict__ masks, const float* __restrict__ v_colors,
              float* __restrict__ v_coeffs, float* __restrict__ v_dirs, int accumulate) {
    constexpr int NB = (DEG + 1) * (DEG + 1);
    constexpr int LPG = ShLayout<K>::LPG, GPW = ShLayout<K>::GPW;
    const int lane = threadIdx.x & 31;
    const int k = lane % LPG, grp = lane / LPG;
    const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int64_t n_warps = ((int64_t)gridDim.x * blockDim.x) >> 5;
    const int64_t n_iter = (M + GPW * SH_UNROLL - 1) / (GPW * SH_UNROLL);
    const bool want_dirs = v_dirs != nullptr;
    for (int64_t it = warp; it < n_iter; it += n_warps) {
        int64_t g[SH_UNROLL];
        bool live[SH_UNROLL];
        float dx[SH_UNROLL], dy[SH_UNROLL], dz[SH_UNROLL], vr[SH_UNROLL], vg[SH_UNROLL], vb[SH_UNROLL];
        float c0[SH_UNROLL], c1[SH_UNROLL], c2[SH_UNROLL], o0[SH_UNROLL], o1[SH_UNROLL], o2[SH_UNROLL];
#pragma unroll
        for (int u = 0; u < SH_UNROLL; ++u) {
            g[u] = (it * SH_UNROLL + u) * GPW + grp;
            live[u] = (g[u] < M) && (masks == nullptr || __ldg(masks + g[u]) != 0);
        }
#pragma unroll
        for (int u = 0; u < SH_UNROLL; ++u) {
            dx[u] = dy[u] = dz[u] = vr[u] = vg[u] = vb[u] = 0.f;
            c0[u] = c1[u] = c2[u] = o0[u] = o1[u] = o2[u] = 0.f;
            if (live[u]) {
                dx[u] = __ldg(dirs + g[u] * 3 + 0); dy[u] = __ldg(dirs + g[u] * 3 + 1); dz[u] = __ldg(dirs + g[u] * 3 + 2);
                vr[u] = __ldg(v_colors + g[u] * 3 + 0); vg[u] = __ldg(v_colors + g[u] * 3 + 1);
                vb[u] = __ldg(v_colors + g[u] * 3 + 2);
                if (k < K) {
                    const int64_t e = (g[u] * K + k) * 3;
                    if (want_dirs && k < NB) { c0[u] = __ldg(coeffs + e); c1[u] = __ldg(coeffs + e + 1); c2[u] = __ldg(coeffs + e + 2); }
                    if (accumulate) { o0[u] = v_coeffs[e]; o1[u] = v_coeffs[e + 1]; o2[u] = v_coeffs[e + 2]; }
                }
            }
        }
#pragma unroll
        for (int u = 0; u < SH_UNROLL; ++u) {
            if (!__any_sync(0xffffffffu, live[u])) {
                // nothing visible in this warp slice: zeros (or untouched in accumulate mode)
                if (!accumulate && g[u] < M) {
                    if (k < K) { const int64_t e = (g[u] * K + k) * 3; v_coeffs[e] = 0.f; v_coeffs[e + 1] = 0.f; v_coeffs[e + 2] = 0.f; }
                    if (want_dirs && k < 3 && (LPG >= 4 || k == 0)) {
                        if (LPG >= 4) v_dirs[g[u] * 3 + k] = 0.f;
                        else { v_dirs[g[u] * 3] = 0.f; v_dirs[g[u] * 3 + 1] = 0.f; v_dirs[g[u] * 3 + 2] = 0.f; }
                    }
                }
                continue;
            }
            const float inv = 1.0f / fmaxf(sqrtf(dx[u] * dx[u] + dy[u] * dy[u] + dz[u] * dz[u]), 1e-12f);
            const float x = dx[u] * inv, y = dy[u] * inv, z = dz[u] * inv;
            float Y[25];
            sh_basis<DEG>(x, y, z, Y);
            const float yk = select_k<NB>(Y, k);  // 0 for k >= NB
            if (g[u] < M && k < K && (live[u] || !accumulate)) {
                const int64_t e = (g[u] * K + k) * 3;
                const float m = live[u] ? yk : 0.f;
                v_coeffs[e] = o0[u] + m * vr[u]; v_coeffs[e + 1] = o1[u] + m * vg[u]; v_coeffs[e + 2] = o2[u] + m * vb[u];
            }
            if (want_dirs) {
                // s_k = <coeff[k,:], v_colour>, gathered to every lane of the group
                const float sk = c0[u] * vr[u] + c1[u] * vg[u] + c2[u] * vb[u];
                float s[25];
#pragma unroll
                for (int i = 0; i < NB; ++i) s[i] = __shfl_sync(0xffffffffu, sk, grp * LPG + (i % LPG));
                float ax, ay, az;
                sh_basis_vjp<DEG>(x, y, z, s, ax, ay, az);
                const float dp = ax * x + ay * y + az * z;
                const float gx = (ax - x * dp) * inv, gy = (ay - y * dp) * inv, gz = (az - z * dp) * inv;
                if (g[u] < M && (live[u] || !accumulate)) {
                    const float m = live[u] ? 1.f : 0.f;
                    if (LPG >= 4) {
                        if (k < 3) {
                            const float val = m * (k == 0 ? gx : k == 1 ? gy : gz);
                            if (accumulate) v_dirs[g[u] * 3 + k] += val; else v_dirs[g[u] * 3 + k] = val;
                        }
                    } else {
                        if (accumulate) { v_dirs[g[u] * 3] += gx; v_dirs[g[u] * 3 + 1] += gy; v_dirs[g[u] * 3 + 2] += gz; }
                        else { v_dirs[g[u] * 3] = m * gx; v_dirs[g[u] * 3 + 1] = m * gy; v_dirs[g[u] * 3 + 2] = m * gz; }
                    }
                }
            }
        }
    }
}

// ---------------------------------------------------------------------------------------
// Transposing kernels for K*3 % 4 == 0 (K = 4, 16 — the reference's sh_degree 1 and 3).
// Compute is lane-per-Gaussian (one warp instruction serves 32 Gaussians); global traffic is
// float4 over the warp's contiguous 32*K*12-byte span; the transpose goes through a padded
// shared-memory tile (row stride K3+4 floats: float4 accesses are conflict-free per quarter
// warp both row-wise and lane-per-row).  Everything is streamed float4 by float4 with
// compile-time indices, so the kernels stay near 64 registers and 32 warps fit on an SM —
// these kernels are latency-bound otherwise (each chunk is a mask -> coefficients -> result
// chain of dependent DRAM round trips).
// "View mode" of the transposing kernels (the fused renderer path): directions are
// means - camera centre, the mask is radii > 0, colours carry +0.5 / clamp_min 0 and the depth
// as a 4th channel (rows A5 + A6 of SURVEY.md section 8 in one pass); the backward reads the
// 4-channel upstream gradient and applies the clamp mask from the saved forward colours.
struct ShView {
    const float* campos;        // DEVICE pointer to the camera centre (nullptr: dirs are given)
    const int32_t* radii;       // mask source (nullptr: use `masks`)
    const float* depths;        // forward: 4th output channel (nullptr: 3-channel output)
    const float* colors_fwd;    // backward: saved [M,4] forward colours (clamp mask), nullable
    int vc_stride;              // backward: floats per row of v_colors (3, or 4 in view mode)
};

template <int K>
struct ShT {
    static constexpr int K3 = K * 3;
    static constexpr int Q = K3 / 4;        // float4 per Gaussian
    static constexpr int STRIDE = K3 + 4;   // floats per padded row
    static constexpr int TILE_FLOATS = 32 * STRIDE;
};
#ifndef SC_SH_WARPS
#define SC_SH_WARPS 8
#endif
constexpr int SHT_WARPS = SC_SH_WARPS;
#ifndef SC_SH_MINB_FWD
#define SC_SH_MINB_FWD 3  // resident CTAs per SM the register budget is capped for
#endif
#ifndef SC_SH_MINB_BWD
#define SC_SH_MINB_BWD 3
#endif

// coalesced: the chunk's n*Q float4 -> padded rows; only Gaussians in `vis`.  The copies are
// asynchronous (cp.async, 16 B each, no staging registers): all Q of a lane are in flight at
// once; sht_wait() is called right before the rows are read.
template <int K>
__device__ __forceinline__ void sht_issue(const float4* __restrict__ src4, unsigned vis, float* tile, int lane) {
    constexpr int Q = ShT<K>::Q, STRIDE = ShT<K>::STRIDE;
    if (vis) {
#pragma unroll
        for (int u = 0; u < Q; ++u) {
            const int i = lane + 32 * u;
            const int gl = i / Q;
            if ((vis >> gl) & 1u) {
                const unsigned dst = (unsigned)__cvta_generic_to_shared(tile + gl * STRIDE + 4 * (i - gl * Q));
                asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst), "l"(src4 + i) : "memory");
            }
        }
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
}

__device__ __forceinline__ void sht_wait() { asm volatile("cp.async.wait_group 0;" ::: "memory"); }

// Per warp: chunks of 32 consecutive Gaussians.  The per-lane inputs (mask, position, upstream
// gradient) of the NEXT chunk are requested before the current chunk's coefficient copies are
// waited for, so a chunk exposes one DRAM round trip (its coefficient span, all Q float4 of a
// lane in flight at once); 24 resident warps per SM overlap those.  (A two-tile software
// pipeline per warp was measured slower: its shared memory halves the resident warps.)
template <int K, int DEG>
__global__ void __launch_bounds__(SHT_WARPS * 32, SC_SH_MINB_FWD)
sh_fwd_t_kernel(int64_t M, const float* __restrict__ dirs, const float4* __restrict__ coeffs4,
                const uint8_t* __restrict__ masks, float* __restrict__ colors, ShView vw) {
    extern __shared__ __align__(16) float smem[];
    constexpr int NB = (DEG + 1) * (DEG + 1);
    constexpr int Q = ShT<K>::Q, STRIDE = ShT<K>::STRIDE, TILE = ShT<K>::TILE_FLOATS;
    const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
    float* tile = smem + wib * TILE;
    const int64_t n_chunks = (M + 31) >> 5;
    const float cx = vw.campos ? __ldg(vw.campos + 0) : 0.f, cy = vw.campos ? __ldg(vw.campos + 1) : 0.f,
                cz = vw.campos ? __ldg(vw.campos + 2) : 0.f;
    struct LaneIn { bool m; float x, y, z, d; };
    auto fetch = [&](int64_t chunk) {
        LaneIn in{false, 0.f, 0.f, 0.f, 0.f};
        const int64_t g = (chunk << 5) + lane;
        if (chunk < n_chunks && g < M) {
            in.m = vw.radii ? __ldg(vw.radii + g) > 0 : (masks == nullptr || __ldg(masks + g) != 0);
            in.x = __ldg(dirs + g * 3 + 0) - cx; in.y = __ldg(dirs + g * 3 + 1) - cy;
            in.z = __ldg(dirs + g * 3 + 2) - cz;
            if (vw.depths) in.d = __ldg(vw.depths + g);
        }
        return in;
    };
    const int64_t first = (int64_t)blockIdx.x * SHT_WARPS + wib, step = (int64_t)gridDim.x * SHT_WARPS;
    if (first >= n_chunks) return;
    LaneIn cur = fetch(first);
    for (int64_t chunk = first; chunk < n_chunks; chunk += step) {
        sht_issue<K>(coeffs4 + (chunk << 5) * Q, __ballot_sync(0xffffffffu, cur.m), tile, lane);
        const LaneIn nxt = fetch(chunk + step);
        const int64_t g = (chunk << 5) + lane;
        float r = 0.f, gg = 0.f, b = 0.f;
        float Y[25];
        if (cur.m) {
            const float inv = 1.0f / fmaxf(sqrtf(cur.x * cur.x + cur.y * cur.y + cur.z * cur.z), 1e-12f);
            sh_basis<DEG>(cur.x * inv, cur.y * inv, cur.z * inv, Y);
        }
        sht_wait();
        __syncwarp();
        if (cur.m) {
            const float4* row = reinterpret_cast<const float4*>(tile + lane * STRIDE);
            float acc[3] = {0.f, 0.f, 0.f};
#pragma unroll
            for (int q = 0; q < Q; ++q) {
                if (4 * q < NB * 3) {  // float4s wholly beyond the used degree are skipped
                    const float4 c = row[q];
                    const float cv[4] = {c.x, c.y, c.z, c.w};
#pragma unroll
                    for (int j = 0; j < 4; ++j) {
                        const int e = 4 * q + j;
                        if (e < NB * 3) acc[e % 3] = fmaf(Y[e / 3], cv[j], acc[e % 3]);
                    }
                }
            }
            r = acc[0]; gg = acc[1]; b = acc[2];
        }
        __syncwarp();  // the tile is overwritten by the next chunk's copies
        if (g < M) {
            if (vw.depths) {
                // view mode (A5 + A6 fused): clamp_min(colour + 0.5, 0) and the depth channel
                reinterpret_cast<float4*>(colors)[g] = make_float4(fmaxf(r + 0.5f, 0.f), fmaxf(gg + 0.5f, 0.f),
                                                                   fmaxf(b + 0.5f, 0.f), cur.d);
            } else {
                colors[g * 3 + 0] = r; colors[g * 3 + 1] = gg; colors[g * 3 + 2] = b;
            }
        }
        cur = nxt;
    }
}

template <int K, int DEG>
__global__ void __launch_bounds__(SHT_WARPS * 32, SC_SH_MINB_BWD)
sh_bwd_t_kernel(int64_t M, const float* __restrict__ dirs, const float4* __restrict__ coeffs4,
                const uint8_t* __restrict__ masks, const float* __restrict__ v_colors,
                float4* __restrict__ v_coeffs4, float* __restrict__ v_dirs, int accumulate, ShView vw) {
    extern __shared__ __align__(16) float smem[];
    constexpr int NB = (DEG + 1) * (DEG + 1);
    constexpr int Q = ShT<K>::Q, STRIDE = ShT<K>::STRIDE, TILE = ShT<K>::TILE_FLOATS;
    const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
    float* tile = smem + wib * TILE;
    const int64_t n_chunks = (M + 31) >> 5;
    const bool want_dirs = v_dirs != nullptr;
    const int vs = vw.vc_stride;
    const float cx = vw.campos ? __ldg(vw.campos + 0) : 0.f, cy = vw.campos ? __ldg(vw.campos + 1) : 0.f,
                cz = vw.campos ? __ldg(vw.campos + 2) : 0.f;
    struct LaneIn { bool m; float x, y, z, r, g, b; };
    auto fetch = [&](int64_t chunk) {
        LaneIn in{false, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
        const int64_t g = (chunk << 5) + lane;
        if (chunk < n_chunks && g < M) {
            in.m = vw.radii ? __ldg(vw.radii + g) > 0 : (masks == nullptr || __ldg(masks + g) != 0);
            in.x = __ldg(dirs + g * 3 + 0) - cx; in.y = __ldg(dirs + g * 3 + 1) - cy;
            in.z = __ldg(dirs + g * 3 + 2) - cz;
            if (vs == 4) {
                // view mode: one 16-byte load per row; unconditional (a load gated on the mask
                // would wait for the mask's DRAM round trip)
                const float4 v = __ldg(reinterpret_cast<const float4*>(v_colors) + g);
                in.r = v.x; in.g = v.y; in.b = v.z;
                if (vw.colors_fwd) {  // backward of clamp_min(colour + 0.5, 0)
                    const float4 c = __ldg(reinterpret_cast<const float4*>(vw.colors_fwd) + g);
                    in.r = c.x > 0.f ? in.r : 0.f; in.g = c.y > 0.f ? in.g : 0.f; in.b = c.z > 0.f ? in.b : 0.f;
                }
            } else {
                in.r = __ldg(v_colors + g * 3 + 0); in.g = __ldg(v_colors + g * 3 + 1); in.b = __ldg(v_colors + g * 3 + 2);
            }
        }
        return in;
    };
    const int64_t first = (int64_t)blockIdx.x * SHT_WARPS + wib, step = (int64_t)gridDim.x * SHT_WARPS;
    if (first >= n_chunks) return;
    LaneIn cur = fetch(first);
    for (int64_t chunk = first; chunk < n_chunks; chunk += step) {
        const unsigned vis = __ballot_sync(0xffffffffu, cur.m);
        sht_issue<K>(coeffs4 + (chunk << 5) * Q, want_dirs ? vis : 0u, tile, lane);
        const LaneIn nxt = fetch(chunk + step);
        const int64_t g0 = chunk << 5;
        const int n = (int)min((int64_t)32, M - g0);
        const int64_t g = g0 + lane;
        const bool live = cur.m;
        float4* dst4 = v_coeffs4 + g0 * Q;
        sht_wait();
        __syncwarp();
        if (!vis) {
            // nothing visible in the chunk: zeros, or untouched in accumulate mode
            if (!accumulate) {
#pragma unroll
                for (int u = 0; u < Q; ++u) {
                    const int i = lane + 32 * u;
                    if (i < n * Q) dst4[i] = make_float4(0.f, 0.f, 0.f, 0.f);
                }
                if (want_dirs && lane < n) { v_dirs[g * 3] = 0.f; v_dirs[g * 3 + 1] = 0.f; v_dirs[g * 3 + 2] = 0.f; }
            }
        } else {
            float x = cur.x, y = cur.y, z = cur.z, inv = 0.f;
            const float vc[3] = {live ? cur.r : 0.f, live ? cur.g : 0.f, live ? cur.b : 0.f};
            if (live) {
                inv = 1.0f / fmaxf(sqrtf(x * x + y * y + z * z), 1e-12f);
                x *= inv; y *= inv; z *= inv;
            }
            float gx = 0.f, gy = 0.f, gz = 0.f;
            float4* row = reinterpret_cast<float4*>(tile + lane * STRIDE);
            if (want_dirs) {
                if (live) {
                    float s[25];
#pragma unroll
                    for (int k = 0; k < NB; ++k) s[k] = 0.f;
#pragma unroll
                    for (int q = 0; q < Q; ++q) {
                        if (4 * q < NB * 3) {
                            const float4 c = row[q];
                            const float cv[4] = {c.x, c.y, c.z, c.w};
#pragma unroll
                            for (int j = 0; j < 4; ++j) {
                                const int e = 4 * q + j;
                                if (e < NB * 3) s[e / 3] = fmaf(cv[j], vc[e % 3], s[e / 3]);
                            }
                        }
                    }
                    float ax, ay, az;
                    sh_basis_vjp<DEG>(x, y, z, s, ax, ay, az);
                    const float dp = ax * x + ay * y + az * z;
                    gx = (ax - x * dp) * inv; gy = (ay - y * dp) * inv; gz = (az - z * dp) * inv;
                }
                __syncwarp();
            }
            // v_coeffs rows: Y_k * v_colour, zeros beyond the used degree and for masked Gaussians
            // (the basis is evaluated here, after the direction gradient, so that Y[] and s[] are
            // never live together)
            if (lane < n) {
                float Y[25];
                sh_basis<DEG>(x, y, z, Y);
#pragma unroll
                for (int q = 0; q < Q; ++q) {
                    float o[4];
#pragma unroll
                    for (int j = 0; j < 4; ++j) {
                        const int e = 4 * q + j;
                        o[j] = (live && e < NB * 3) ? Y[e / 3 < NB ? e / 3 : 0] * vc[e % 3] : 0.f;
                    }
                    row[q] = make_float4(o[0], o[1], o[2], o[3]);
                }
            }
            __syncwarp();
            if (accumulate) {
                // red.global.add.v4.f32: fire-and-forget, no read of the running sums
#pragma unroll
                for (int u = 0; u < Q; ++u) {
                    const int i = lane + 32 * u;
                    const int gl = i / Q;
                    if ((gl < n) && ((vis >> gl) & 1u))
                        atomicAdd(dst4 + i, *reinterpret_cast<const float4*>(tile + gl * STRIDE + 4 * (i - gl * Q)));
                }
            } else {
#pragma unroll
                for (int u = 0; u < Q; ++u) {
                    const int i = lane + 32 * u;
                    const int gl = i / Q;
                    if (gl < n) dst4[i] = *reinterpret_cast<const float4*>(tile + gl * STRIDE + 4 * (i - gl * Q));
                }
            }
            if (want_dirs && lane < n) {
                if (accumulate) {
                    if (live) { atomicAdd(v_dirs + g * 3 + 0, gx); atomicAdd(v_dirs + g * 3 + 1, gy); atomicAdd(v_dirs + g * 3 + 2, gz); }
                } else {
                    v_dirs[g * 3 + 0] = gx; v_dirs[g * 3 + 1] = gy; v_dirs[g * 3 + 2] = gz;
                }
            }
        }
        __syncwarp();  // the tile is overwritten by the next chunk's copies
        cur = nxt;
    }
}

template <int K, int DEG>
int launch_fwd(int64_t M, const float* dirs, const float* coeffs, const uint8_t* masks, float* colors,
               ShView vw, cudaStream_t st) {
    if constexpr ((K * 3) % 4 == 0) {
        const int smem = SHT_WARPS * ShT<K>::TILE_FLOATS * (int)sizeof(float);
        auto kern = sh_fwd_t_kernel<K, DEG>;
        if (smem > 48 * 1024) SC_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
        const int grid = (int)min(sc_div_up((M + 31) >> 5, SHT_WARPS), (int64_t)SC_NUM_SMS * 16);
        kern<<<grid, SHT_WARPS * 32, smem, st>>>(M, dirs, reinterpret_cast<const float4*>(coeffs), masks, colors, vw);
        SC_CHECK_LAUNCH("sh_fwd_t_kernel");
        return SC_OK;
    }
    if (vw.radii || vw.depths) {
        sc_set_error("SH view mode needs K*3 %% 4 == 0 (K = 4, 16), got K=%d", K);
        return SC_ERR_UNSUPPORTED;
    }
    const int64_t iters = sc_div_up(M, ShLayout<K>::GPW * SH_UNROLL);
    const int grid = (int)min(sc_div_up(iters, SH_WARPS), (int64_t)SC_NUM_SMS * 64);
    sh_fwd_kernel<K, DEG><<<grid, SH_THREADS, 0, st>>>(M, dirs, coeffs, masks, colors);
    SC_CHECK_LAUNCH("sh_fwd_kernel");
    return SC_OK;
}

template <int K, int DEG>
int launch_bwd(int64_t M, const float* dirs, const float* coeffs, const uint8_t* masks, const float* v_colors,
               float* v_coeffs, float* v_dirs, int accumulate, ShView vw, cudaStream_t st) {
    if constexpr ((K * 3) % 4 == 0) {
        const int smem = SHT_WARPS * ShT<K>::TILE_FLOATS * (int)sizeof(float);
        auto kern = sh_bwd_t_kernel<K, DEG>;
        if (smem > 48 * 1024) SC_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
        const int grid = (int)min(sc_div_up((M + 31) >> 5, SHT_WARPS), (int64_t)SC_NUM_SMS * 16);
        kern<<<grid, SHT_WARPS * 32, smem, st>>>(M, dirs, reinterpret_cast<const float4*>(coeffs), masks, v_colors,
                                                 reinterpret_cast<float4*>(v_coeffs), v_dirs, accumulate, vw);
        SC_CHECK_LAUNCH("sh_bwd_t_kernel");
        return SC_OK;
    }
    if (vw.radii || vw.colors_fwd || vw.vc_stride != 3) {
        sc_set_error("SH view mode needs K*3 %% 4 == 0 (K = 4, 16), got K=%d", K);
        return SC_ERR_UNSUPPORTED;
    }
    const int64_t iters = sc_div_up(M, ShLayout<K>::GPW * SH_UNROLL);
    const int grid = (int)min(sc_div_up(iters, SH_WARPS), (int64_t)SC_NUM_SMS * 64);
    sh_bwd_kernel<K, DEG><<<grid, SH_THREADS, 0, st>>>(M, dirs, coeffs, masks, v_colors, v_coeffs, v_dirs,
                                                       accumulate);
    SC_CHECK_LAUNCH("sh_bwd_kernel");
    return SC_OK;
}

}  // namespace

#define SH_DISPATCH(FN, ...)                                                        \
    switch (K) {                                                                    \
        case 1: return FN<1, 0>(__VA_ARGS__);                                       \
        case 4: return degree == 0 ? FN<4, 0>(__VA_ARGS__) : FN<4, 1>(__VA_ARGS__); \
        case 9:                                                                     \
            return degree == 0 ? FN<9, 0>(__VA_ARGS__)                              \
                 : degree == 1 ? FN<9, 1>(__VA_ARGS__) : FN<9, 2>(__VA_ARGS__);     \
        case 16:                                                                    \
            return degree == 0 ? FN<16, 0>(__VA_ARGS__)                             \
                 : degree == 1 ? FN<16, 1>(__VA_ARGS__)                             \
                 : degree == 2 ? FN<16, 2>(__VA_ARGS__) : FN<16, 3>(__VA_ARGS__);   \
        case 25:                                                                    \
            return degree == 0 ? FN<25, 0>(__VA_ARGS__)                             \
                 : degree == 1 ? FN<25, 1>(__VA_ARGS__)                             \
                 : degree == 2 ? FN<25, 2>(__VA_ARGS__)                             \
                 : degree == 3 ? FN<25, 3>(__VA_ARGS__) : FN<25, 4>(__VA_ARGS__);   \
        default: break;                                                             \
    }

extern "C" int sc_sh_fwd(int64_t M, int K, int degree, const float* dirs, const float* coeffs,
                         const uint8_t* masks, float* colors, void* stream) {
    SC_CHECK_ARG(M >= 0, "M");
    SC_CHECK_ARG(degree >= 0 && degree <= 4 && (degree + 1) * (degree + 1) <= K, "degree vs K");
    if (M == 0) return SC_OK;
    SC_CHECK_ARG(dirs && coeffs && colors, "null pointer");
    SC_CHECK_ARG(((uintptr_t)coeffs & 15) == 0, "16-byte alignment of coeffs");
    cudaStream_t st = (cudaStream_t)stream;
    const ShView vw{nullptr, nullptr, nullptr, nullptr, 3};
    SH_DISPATCH(launch_fwd, M, dirs, coeffs, masks, colors, vw, st)
    sc_set_error("unsupported SH coefficient count K=%d (supported 1,4,9,16,25)", K);
    return SC_ERR_UNSUPPORTED;
}

extern "C" int sc_sh_fwd_view(int64_t M, int K, int degree, const float* means, const float* campos,
                              const float* coeffs, const int32_t* radii, const float* depths,
                              float* colors4, void* stream) {
    SC_CHECK_ARG(M >= 0, "M");
    SC_CHECK_ARG(degree >= 0 && degree <= 4 && (degree + 1) * (degree + 1) <= K, "degree vs K");
    if (M == 0) return SC_OK;
    SC_CHECK_ARG(means && campos && coeffs && radii && depths && colors4, "null pointer");
    SC_CHECK_ARG(((uintptr_t)coeffs & 15) == 0 && ((uintptr_t)colors4 & 15) == 0, "16-byte alignment");
    cudaStream_t st = (cudaStream_t)stream;
    const ShView vw{campos, radii, depths, nullptr, 3};
    const float* dirs = means;
    const uint8_t* masks = nullptr;
    float* colors = colors4;
    SH_DISPATCH(launch_fwd, M, dirs, coeffs, masks, colors, vw, st)
    sc_set_error("unsupported SH coefficient count K=%d (view mode: 4, 16)", K);
    return SC_ERR_UNSUPPORTED;
}

extern "C" int sc_sh_bwd(int64_t M, int K, int degree, const float* dirs, const float* coeffs,
                         const uint8_t* masks, const float* v_colors, float* v_coeffs, float* v_dirs,
                         int accumulate, void* stream) {
    SC_CHECK_ARG(M >= 0, "M");
    SC_CHECK_ARG(degree >= 0 && degree <= 4 && (degree + 1) * (degree + 1) <= K, "degree vs K");
    if (M == 0) return SC_OK;
    SC_CHECK_ARG(dirs && coeffs && v_colors && v_coeffs, "null pointer");
    SC_CHECK_ARG(((uintptr_t)coeffs & 15) == 0 && ((uintptr_t)v_coeffs & 15) == 0,
                 "16-byte alignment of coeffs");
    cudaStream_t st = (cudaStream_t)stream;
    const ShView vw{nullptr, nullptr, nullptr, nullptr, 3};
    SH_DISPATCH(launch_bwd, M, dirs, coeffs, masks, v_colors, v_coeffs, v_dirs, accumulate, vw, st)
    sc_set_error("unsupported SH coefficient count K=%d (supported 1,4,9,16,25)", K);
    return SC_ERR_UNSUPPORTED;
}

extern "C" int sc_sh_bwd_view(int64_t M, int K, int degree, const float* means, const float* campos,
                              const float* coeffs, const int32_t* radii, const float* colors4_fwd,
                              const float* v_colors4, float* v_coeffs, float* v_means, int accumulate,
                              void* stream) {
    SC_CHECK_ARG(M >= 0, "M");
    SC_CHECK_ARG(degree >= 0 && degree <= 4 && (degree + 1) * (degree + 1) <= K, "degree vs K");
    if (M == 0) return SC_OK;
    SC_CHECK_ARG(means && campos && coeffs && radii && colors4_fwd && v_colors4 && v_coeffs, "null pointer");
    SC_CHECK_ARG(((uintptr_t)coeffs & 15) == 0 && ((uintptr_t)v_coeffs & 15) == 0, "16-byte alignment of coeffs");
    SC_CHECK_ARG(((uintptr_t)colors4_fwd & 15) == 0 && ((uintptr_t)v_colors4 & 15) == 0,
                 "16-byte alignment of the 4-channel colour rows");
    cudaStream_t st = (cudaStream_t)stream;
    const ShView vw{campos, radii, nullptr, colors4_fwd, 4};
    const float* dirs = means;
    const uint8_t* masks = nullptr;
    const float* v_colors = v_colors4;
    float* v_dirs = v_means;
    SH_DISPATCH(launch_bwd, M, dirs, coeffs, masks, v_colors, v_coeffs, v_dirs, accumulate, vw, st)
    sc_set_error("unsupported SH coefficient count K=%d (view mode: 4, 16)", K);
    return SC_ERR_UNSUPPORTED;
}
