// Fused photometric loss: masked L1 + 11x11 Gaussian SSIM,
// forward and backward, straight on the compositor's [H,W,CH] output — the gradient comes out
// in the layout the rasterizer's backward consumes, so no permute / slice / cat nodes and no
// six depthwise convolutions (and their backward) per view.
//
// Replaces /root/reference/street_gaussian/utils/loss_utils.py:21-38 (l1_loss), :84-131 (ssim)
// as combined at /root/reference/train.py:183-186:
//     loss = (1 - lambda_dssim) * lambda_l1 * L1 + lambda_dssim * (1 - SSIM)
// (the LPIPS term of the same line is a network evaluation: out of scope.)
//
// HBM-bound.  One CTA = one 16x16 pixel tile, all 3 channels: the 26x26 halo of both images is
// staged once in shared memory (zero outside the image and outside the mask, which is exactly
// the reference's zero padding + torch.where), the 11-tap window is applied separably.
//   forward : writes the three partial-derivative maps d ssim / d{mu1, E[x^2], E[xy]} and
//             accumulates sum(ssim), sum|x-y| over the mask, and the mask count in doubles;
//             the last CTA turns them into {loss, L1, SSIM} (floats, on the device).
//   backward: d loss / d image = window (*) dmu1 + 2 x window (*) dE11 + y window (*) dE12
//             (the window is symmetric), plus the L1 sign term; scaled by the upstream
//             gradient read from device memory (no host sync).
#include "common.cuh"

namespace {

constexpr int LT = 16;            // tile edge
constexpr int LR = 5;             // window radius
constexpr int LH = LT + 2 * LR;   // 26: halo edge
constexpr float SSIM_C1 = 0.0001f, SSIM_C2 = 0.0009f;

// exp(-(k-5)^2 / (2 * 1.5^2)) / sum, k = 0..10 (loss_utils.py:84-86)
__constant__ float c_win[11] = {0.0010283802f, 0.0075987582f, 0.0360007721f, 0.1093606895f, 0.2130055377f,
                                0.2660117249f, 0.2130055377f, 0.1093606895f, 0.0360007721f, 0.0075987582f,
                                0.0010283802f};

struct ImgView {        // element (c, y, x) of an image lives at base[c * cs + (y * W + x) * ps]
    const float* base;
    int64_t cs, ps;
};

template <int WARPS>
__device__ __forceinline__ float block_sum(float v, float* s_red) {
#pragma unroll
    for (int o = 16; o >= 1; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    __syncthreads();
    if (lane == 0) s_red[w] = v;
    __syncthreads();
    float t = 0.f;
#pragma unroll
    for (int i = 0; i < WARPS; ++i) t += s_red[i];
    return t;
}

// Both passes of the separable window keep a sliding window in REGISTERS: a horizontal task is
// (channel, halo row, half row) and produces 8 outputs from 18 staged values; a vertical task
// is (channel, column, group of 4 rows) and produces 4 outputs from 14 values — 3-4 shared
// loads per output instead of 11 per tap set.
constexpr int HSEG = 8;   // outputs per horizontal task
constexpr int VSEG = 4;   // outputs per vertical task
constexpr int LPAD = LT + 1;  // odd row stride: the horizontal tasks of a warp walk down the rows
// forward CTA: 192 threads = the 3 * 4 * 16 vertical tasks exactly (the horizontal pass has 156)
constexpr int LF_THREADS = 3 * (LT / VSEG) * LT;

// sums[0] = sum ssim, sums[1] = sum |x - y| over mask (all channels), sums[2] = masked pixels,
// counter: CTAs done; out[0..2] = loss, L1, SSIM
__global__ void __launch_bounds__(LF_THREADS)
photometric_fwd_kernel(int H, int W, ImgView img, const float* __restrict__ gt,
                       const uint8_t* __restrict__ mask, float* __restrict__ dmaps, double* __restrict__ sums,
                       unsigned* __restrict__ counter, float lambda_dssim, float lambda_l1,
                       float* __restrict__ out) {
    __shared__ float s_x[3][LH][LH + 1], s_y[3][LH][LH + 1];
    __shared__ float s_h[3][5][LH][LPAD];
    __shared__ float s_red[8];
    const int tid = threadIdx.x;
    const int x0 = blockIdx.x * LT, y0 = blockIdx.y * LT;
    const int64_t HW = (int64_t)H * W;
    const bool hwc4 = img.cs == 1 && img.ps == 4 && (((uintptr_t)img.base) & 15) == 0;
    float l1_part = 0.f, cnt_part = 0.f;
    for (int i = tid; i < LH * LH; i += LF_THREADS) {
        const int hy = i / LH, hx = i - hy * LH;
        const int y = y0 + hy - LR, x = x0 + hx - LR;
        float a[3] = {0.f, 0.f, 0.f}, b[3] = {0.f, 0.f, 0.f};
        if (y >= 0 && y < H && x >= 0 && x < W) {
            const int64_t p = (int64_t)y * W + x;
            if (mask == nullptr || mask[p]) {
                if (hwc4) {
                    const float4 v = __ldg(reinterpret_cast<const float4*>(img.base) + p);
                    a[0] = v.x; a[1] = v.y; a[2] = v.z;
                } else {
#pragma unroll
                    for (int c = 0; c < 3; ++c) a[c] = __ldg(img.base + c * img.cs + p * img.ps);
                }
#pragma unroll
                for (int c = 0; c < 3; ++c) b[c] = __ldg(gt + c * HW + p);
                // the tile's own pixels (not the halo) carry the L1 term
                if (hy >= LR && hy < LR + LT && hx >= LR && hx < LR + LT) {
                    l1_part += fabsf(a[0] - b[0]) + fabsf(a[1] - b[1]) + fabsf(a[2] - b[2]);
                    cnt_part += 1.f;
                }
            }
        }
#pragma unroll
        for (int c = 0; c < 3; ++c) { s_x[c][hy][hx] = a[c]; s_y[c][hy][hx] = b[c]; }
    }
    __syncthreads();
    // horizontal pass, all channels: 3 * 26 * 2 tasks.  The products x^2, y^2, xy are formed once
    // per staged value, so every tap is one FMA per moment map.
    for (int t = tid; t < 3 * LH * (LT / HSEG); t += LF_THREADS) {
        // consecutive lanes take consecutive halo rows (odd strides 27 / 17: no bank conflicts)
        const int ch = t / LH, r = t - ch * LH, c = ch >> 1, c0 = (ch & 1) * HSEG;
        float xa[HSEG + 10], ya[HSEG + 10], xx[HSEG + 10], yy[HSEG + 10], xy[HSEG + 10];
#pragma unroll
        for (int j = 0; j < HSEG + 10; ++j) {
            xa[j] = s_x[c][r][c0 + j]; ya[j] = s_y[c][r][c0 + j];
            xx[j] = xa[j] * xa[j]; yy[j] = ya[j] * ya[j]; xy[j] = xa[j] * ya[j];
        }
#pragma unroll
        for (int o = 0; o < HSEG; ++o) {
            float hx_ = 0.f, hy_ = 0.f, hxx = 0.f, hyy = 0.f, hxy = 0.f;
#pragma unroll
            for (int k = 0; k < 11; ++k) {
                const float w = c_win[k];
                hx_ = fmaf(w, xa[o + k], hx_); hy_ = fmaf(w, ya[o + k], hy_);
                hxx = fmaf(w, xx[o + k], hxx); hyy = fmaf(w, yy[o + k], hyy); hxy = fmaf(w, xy[o + k], hxy);
            }
            s_h[c][0][r][c0 + o] = hx_; s_h[c][1][r][c0 + o] = hy_; s_h[c][2][r][c0 + o] = hxx;
            s_h[c][3][r][c0 + o] = hyy; s_h[c][4][r][c0 + o] = hxy;
        }
    }
    __syncthreads();
    // vertical pass + SSIM: 3 * 4 * 16 tasks, 4 rows each
    float ssim_part = 0.f;
    {
        const int c = tid / ((LT / VSEG) * LT), rem = tid - c * (LT / VSEG) * LT;
        const int r0 = (rem >> 4) * VSEG, col = rem & 15;
        float acc[VSEG][5];
#pragma unroll
        for (int o = 0; o < VSEG; ++o)
#pragma unroll
            for (int q = 0; q < 5; ++q) acc[o][q] = 0.f;
#pragma unroll
        for (int q = 0; q < 5; ++q) {
            float v[VSEG + 10];
#pragma unroll
            for (int j = 0; j < VSEG + 10; ++j) v[j] = s_h[c][q][r0 + j][col];
#pragma unroll
            for (int o = 0; o < VSEG; ++o)
#pragma unroll
                for (int k = 0; k < 11; ++k) acc[o][q] = fmaf(c_win[k], v[o + k], acc[o][q]);
        }
        const int x = x0 + col;
#pragma unroll
        for (int o = 0; o < VSEG; ++o) {
            const int y = y0 + r0 + o;
            if (x < W && y < H) {
                const float mu1 = acc[o][0], mu2 = acc[o][1];
                const float s1 = acc[o][2] - mu1 * mu1, s2 = acc[o][3] - mu2 * mu2, s12 = acc[o][4] - mu1 * mu2;
                const float A1 = 2.f * mu1 * mu2 + SSIM_C1, A2 = 2.f * s12 + SSIM_C2;
                const float B1 = mu1 * mu1 + mu2 * mu2 + SSIM_C1, B2 = s1 + s2 + SSIM_C2;
                const float inv = 1.0f / (B1 * B2);
                const float m = A1 * A2 * inv;
                ssim_part += m;
                const int64_t p = (int64_t)y * W + x;
                // partial derivatives with E[x^2], E[xy] held fixed (sigma's depend on mu1 through them)
                dmaps[(0 * 3 + c) * HW + p] = 2.f * mu2 * (A2 - A1) * inv - 2.f * mu1 * m * (B2 - B1) * inv;
                dmaps[(1 * 3 + c) * HW + p] = -m / B2;
                dmaps[(2 * 3 + c) * HW + p] = 2.f * A1 * inv;
            }
        }
    }
    const float bs = block_sum<LF_THREADS / 32>(ssim_part, s_red);
    const float bl = block_sum<LF_THREADS / 32>(l1_part, s_red);
    const float bc = block_sum<LF_THREADS / 32>(cnt_part, s_red);
    if (tid == 0) {
        atomicAdd(sums + 0, (double)bs);
        atomicAdd(sums + 1, (double)bl);
        atomicAdd(sums + 2, (double)bc);
        __threadfence();
        const unsigned done = atomicAdd(counter, 1u);
        if (done == gridDim.x * gridDim.y - 1) {  // last CTA: every partial sum is visible
            __threadfence();
            const double ssum = atomicAdd(sums + 0, 0.0), lsum = atomicAdd(sums + 1, 0.0),
                         cnt = atomicAdd(sums + 2, 0.0);
            const double l1 = cnt > 0.0 ? lsum / (cnt * 3.0) : 0.0;
            const double ss = ssum / (3.0 * (double)HW);
            out[0] = (float)((1.0 - lambda_dssim) * lambda_l1 * l1 + lambda_dssim * (1.0 - ss));
            out[1] = (float)l1;
            out[2] = (float)ss;
        }
    }
}

__global__ void __launch_bounds__(256)
photometric_bwd_kernel(int H, int W, ImgView img, const float* __restrict__ gt, const uint8_t* __restrict__ mask,
                       const float* __restrict__ dmaps, const double* __restrict__ sums,
                       const float* __restrict__ upstream, float lambda_dssim, float lambda_l1,
                       float* __restrict__ v_img, int64_t v_cs, int64_t v_ps, int zero_extra) {
    __shared__ float s_d[9][LH][LH + 1];   // [map * 3 + channel], with halo
    __shared__ float s_h[9][LH][LPAD];
    const int tid = threadIdx.x;
    const int x0 = blockIdx.x * LT, y0 = blockIdx.y * LT;
    const int64_t HW = (int64_t)H * W;
    const float up = upstream ? __ldg(upstream) : 1.0f;
    const double cnt = sums[2];
    const float w_ssim = -lambda_dssim * up / (3.0f * (float)HW);
    const float w_l1 = cnt > 0.0 ? (1.0f - lambda_dssim) * lambda_l1 * up / (float)(cnt * 3.0) : 0.f;
    for (int i = tid; i < LH * LH; i += 256) {
        const int hy = i / LH, hx = i - hy * LH;
        const int yy = y0 + hy - LR, xx = x0 + hx - LR;
        const bool ok = yy >= 0 && yy < H && xx >= 0 && xx < W;
        const int64_t q = (int64_t)yy * W + xx;
#pragma unroll
        for (int m = 0; m < 9; ++m) s_d[m][hy][hx] = ok ? __ldg(dmaps + m * HW + q) : 0.f;
    }
    __syncthreads();
    for (int t = tid; t < 9 * LH * (LT / HSEG); t += 256) {
        const int mh = t / LH, r = t - mh * LH, m = mh >> 1, c0 = (mh & 1) * HSEG;  // lanes walk down the rows
        float v[HSEG + 10];
#pragma unroll
        for (int j = 0; j < HSEG + 10; ++j) v[j] = s_d[m][r][c0 + j];
#pragma unroll
        for (int o = 0; o < HSEG; ++o) {
            float h = 0.f;
#pragma unroll
            for (int k = 0; k < 11; ++k) h = fmaf(c_win[k], v[o + k], h);
            s_h[m][r][c0 + o] = h;
        }
    }
    __syncthreads();
    if (tid < 3 * (LT / VSEG) * LT) {
        const int c = tid / ((LT / VSEG) * LT), rem = tid - c * (LT / VSEG) * LT;
        const int r0 = (rem >> 4) * VSEG, col = rem & 15;
        float g[VSEG][3];
#pragma unroll
        for (int m = 0; m < 3; ++m) {
            float v[VSEG + 10];
#pragma unroll
            for (int j = 0; j < VSEG + 10; ++j) v[j] = s_h[m * 3 + c][r0 + j][col];
#pragma unroll
            for (int o = 0; o < VSEG; ++o) {
                float a = 0.f;
#pragma unroll
                for (int k = 0; k < 11; ++k) a = fmaf(c_win[k], v[o + k], a);
                g[o][m] = a;
            }
        }
        const int x = x0 + col;
#pragma unroll
        for (int o = 0; o < VSEG; ++o) {
            const int y = y0 + r0 + o;
            if (x < W && y < H) {
                const int64_t p = (int64_t)y * W + x;
                float gr = 0.f;
                if (mask == nullptr || mask[p]) {
                    const float a = __ldg(img.base + c * img.cs + p * img.ps), b = __ldg(gt + c * HW + p);
                    const float d = a - b;
                    gr = w_ssim * (g[o][0] + 2.f * a * g[o][1] + b * g[o][2]) +
                         w_l1 * (d > 0.f ? 1.f : d < 0.f ? -1.f : 0.f);
                }
                v_img[c * v_cs + p * v_ps] = gr;
                // channels beyond RGB of an interleaved gradient (the compositor's depth channel) get
                // their zeros here instead of a separate fill of the whole buffer
                if (c == 0)
                    for (int e = 0; e < zero_extra; ++e) v_img[(3 + e) * v_cs + p * v_ps] = 0.f;
            }
        }
    }
}

}  // namespace

extern "C" int64_t sc_photometric_workspace_bytes(int H, int W) {
    // 9 derivative maps (padded to 256 B) + {3 double sums, counter}
    return sc_align_up((int64_t)9 * H * W * 4, 256) + 256;
}

extern "C" int sc_photometric_fwd(int H, int W, const float* image, int64_t channel_stride,
                                  int64_t pixel_stride, const float* gt, const uint8_t* mask,
                                  float lambda_dssim, float lambda_l1, void* workspace,
                                  int64_t workspace_bytes, float* out3, void* stream) {
    SC_CHECK_ARG(H > 0 && W > 0 && channel_stride > 0 && pixel_stride > 0, "sizes / strides");
    SC_CHECK_ARG(image && gt && workspace && out3, "null pointer");
    if (workspace_bytes < sc_photometric_workspace_bytes(H, W)) {
        sc_set_error("photometric workspace too small: %lld < %lld", (long long)workspace_bytes,
                     (long long)sc_photometric_workspace_bytes(H, W));
        return SC_ERR_WORKSPACE;
    }
    cudaStream_t st = (cudaStream_t)stream;
    unsigned char* ws = (unsigned char*)workspace;
    const int64_t maps_bytes = sc_align_up((int64_t)9 * H * W * 4, 256);
    SC_CUDA(cudaMemsetAsync(ws + maps_bytes, 0, 256, st));
    dim3 grid((unsigned)sc_div_up(W, LT), (unsigned)sc_div_up(H, LT));
    photometric_fwd_kernel<<<grid, LF_THREADS, 0, st>>>(H, W, ImgView{image, channel_stride, pixel_stride}, gt, mask,
                                                 (float*)ws, (double*)(ws + maps_bytes),
                                                 (unsigned*)(ws + maps_bytes + 32), lambda_dssim, lambda_l1, out3);
    SC_CHECK_LAUNCH("photometric_fwd_kernel");
    return SC_OK;
}

extern "C" int sc_photometric_bwd(int H, int W, const float* image, int64_t channel_stride,
                                  int64_t pixel_stride, const float* gt, const uint8_t* mask,
                                  float lambda_dssim, float lambda_l1, const void* workspace,
                                  const float* upstream, float* v_image, int64_t v_channel_stride,
                                  int64_t v_pixel_stride, int zero_extra_channels, void* stream) {
    SC_CHECK_ARG(H > 0 && W > 0 && channel_stride > 0 && pixel_stride > 0 && v_channel_stride > 0 &&
                     v_pixel_stride > 0, "sizes / strides");
    SC_CHECK_ARG(image && gt && workspace && v_image, "null pointer");
    SC_CHECK_ARG(zero_extra_channels >= 0 && zero_extra_channels <= 8, "zero_extra_channels");
    const unsigned char* ws = (const unsigned char*)workspace;
    const int64_t maps_bytes = sc_align_up((int64_t)9 * H * W * 4, 256);
    dim3 grid((unsigned)sc_div_up(W, LT), (unsigned)sc_div_up(H, LT));
    photometric_bwd_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(
        H, W, ImgView{image, channel_stride, pixel_stride}, gt, mask, (const float*)ws,
        (const double*)(ws + maps_bytes), upstream, lambda_dssim, lambda_l1, v_image, v_channel_stride,
        v_pixel_stride, zero_extra_channels);
    SC_CHECK_LAUNCH("photometric_bwd_kernel");
    return SC_OK;
}
