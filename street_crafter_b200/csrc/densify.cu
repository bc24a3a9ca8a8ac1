// Densification statistics (second half): what the reference does after
// every backward pass with ~10 boolean-index / cat / norm / max launches
// (/root/reference/street_gaussian/models/street_gaussian_model.py:486-497 set_max_radii2D, :503-519
// add_densification_stats, called at /root/reference/train.py:283-284) as ONE pass over the Gaussians:
//
//   visible i:  accum[i,0] += |(absgrad_i) * 0.5 * (W,H)|     accum[i,1] += |(grad_i) * 0.5 * (W,H)|
//               denom[i]   += 1                               max_radii2D[i] = max(max_radii2D[i], radii_i)
//
// HBM-bound: 21 B read + 16 B read-modify-write per visible Gaussian, 1 B per invisible one.
#include "common.cuh"

namespace {

__global__ void __launch_bounds__(256)
densify_stats_kernel(int64_t N, const float2* __restrict__ absgrad, const float2* __restrict__ grad,
                     const uint8_t* __restrict__ vis, const float* __restrict__ radii, float half_w, float half_h,
                     float2* __restrict__ accum, float* __restrict__ denom, float* __restrict__ max_radii) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= N || !vis[i]) return;
    const float2 a = __ldg(absgrad + i), g = __ldg(grad + i);
    const float ax = a.x * half_w, ay = a.y * half_h, gx = g.x * half_w, gy = g.y * half_h;
    float2 acc = accum[i];
    acc.x += sqrtf(ax * ax + ay * ay);
    acc.y += sqrtf(gx * gx + gy * gy);
    accum[i] = acc;
    denom[i] += 1.0f;
    max_radii[i] = fmaxf(max_radii[i], __ldg(radii + i));
}

}  // namespace

extern "C" int sc_densification_stats(int64_t N, const float* means2d_absgrad, const float* means2d_grad,
                                      const uint8_t* visibility, const float* radii, int width, int height,
                                      float* xyz_gradient_accum, float* denom, float* max_radii2D, void* stream) {
    SC_CHECK_ARG(N >= 0 && width > 0 && height > 0, "sizes");
    if (N == 0) return SC_OK;
    SC_CHECK_ARG(means2d_absgrad && means2d_grad && visibility && radii && xyz_gradient_accum && denom &&
                     max_radii2D, "null pointer");
    SC_CHECK_ARG((((uintptr_t)means2d_absgrad | (uintptr_t)means2d_grad | (uintptr_t)xyz_gradient_accum) & 7) == 0,
                 "8-byte alignment of the [N,2] arrays");
    densify_stats_kernel<<<(unsigned)sc_div_up(N, 256), 256, 0, (cudaStream_t)stream>>>(
        N, reinterpret_cast<const float2*>(means2d_absgrad), reinterpret_cast<const float2*>(means2d_grad),
        visibility, radii, 0.5f * (float)width, 0.5f * (float)height, reinterpret_cast<float2*>(xyz_gradient_accum),
        denom, max_radii2D);
    SC_CHECK_LAUNCH("densify_stats_kernel");
    return SC_OK;
}
