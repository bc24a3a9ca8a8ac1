// Spherical-harmonics colour evaluation (A5) and its backward (A8b).
//
// HBM-bound: the [M,K,3] coefficient block is 192 B per Gaussian at K=16.  A warp owns 32
// consecutive Gaussians whose coefficients form ONE contiguous 32*K*3-float span: it is
// loaded with fully coalesced float4 accesses (skipping spans of masked-out Gaussians),
// transposed through odd-stride shared memory (conflict-free), and each lane then evaluates
// its own Gaussian.  The backward writes v_coeffs the same way in reverse.
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

template <int K>
struct ShCfg {
    static constexpr int K3 = K * 3;
    static constexpr int STRIDE = K3 | 1;  // odd -> conflict-free column access
    static constexpr int SMEM_FLOATS = 32 * STRIDE;
};

// coalesced span load: the warp's 32 Gaussians occupy floats [g0*K3, g0*K3 + n*K3)
template <int K>
__device__ __forceinline__ void warp_load_coeffs(const float* __restrict__ coeffs, int64_t g0, int n,
                                                 unsigned vis, float* sm, int lane) {
    constexpr int K3 = ShCfg<K>::K3;
    constexpr int STRIDE = ShCfg<K>::STRIDE;
    const float* src = coeffs + g0 * K3;  // 128*K3 bytes from an aligned base: 16B aligned
    const int nf = n * K3;
    const int nf4 = nf >> 2;
    for (int i = lane; i < nf4; i += 32) {
        const int e = i << 2;
        const int ga = e / K3, gb = (e + 3) / K3;
        if (((vis >> ga) | (vis >> gb)) & 1u) {
            const float4 v = __ldg(reinterpret_cast<const float4*>(src) + i);
            const float vv[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const int ee = e + j;
                const int gl = ee / K3;
                sm[gl * STRIDE + (ee - gl * K3)] = vv[j];
            }
        }
    }
    for (int ee = (nf4 << 2) + lane; ee < nf; ee += 32) {
        const int gl = ee / K3;
        sm[gl * STRIDE + (ee - gl * K3)] = __ldg(src + ee);
    }
}

template <int K>
__device__ __forceinline__ void warp_store_coeffs(float* __restrict__ dst_base, int64_t g0, int n,
                                                  const float* sm, int lane) {
    constexpr int K3 = ShCfg<K>::K3;
    constexpr int STRIDE = ShCfg<K>::STRIDE;
    float* dst = dst_base + g0 * K3;
    const int nf = n * K3;
    const int nf4 = nf >> 2;
    for (int i = lane; i < nf4; i += 32) {
        const int e = i << 2;
        float vv[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int ee = e + j;
            const int gl = ee / K3;
            vv[j] = sm[gl * STRIDE + (ee - gl * K3)];
        }
        reinterpret_cast<float4*>(dst)[i] = make_float4(vv[0], vv[1], vv[2], vv[3]);
    }
    for (int ee = (nf4 << 2) + lane; ee < nf; ee += 32) {
        const int gl = ee / K3;
        dst[ee] = sm[gl * STRIDE + (ee - gl * K3)];
    }
}

template <int K, int DEG>
__global__ void __launch_bounds__(SH_THREADS)
sh_fwd_kernel(int64_t M, const float* __restrict__ dirs, const float* __restrict__ coeffs,
              const uint8_t* __restrict__ masks, float* __restrict__ colors) {
    extern __shared__ float smem[];
    constexpr int NB = (DEG + 1) * (DEG + 1);
    constexpr int STRIDE = ShCfg<K>::STRIDE;
    const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
    float* sm = smem + wib * ShCfg<K>::SMEM_FLOATS;
    const int64_t n_chunks = (M + 31) >> 5;
    for (int64_t chunk = (int64_t)blockIdx.x * SH_WARPS + wib; chunk < n_chunks;
         chunk += (int64_t)gridDim.x * SH_WARPS) {
        const int64_t g0 = chunk << 5;
        const int n = (int)min((int64_t)32, M - g0);
        const int64_t g = g0 + lane;
        const bool live = (lane < n) && (masks == nullptr || __ldg(masks + g) != 0);
        const unsigned vis = __ballot_sync(0xffffffffu, live);
        float r = 0.f, gg = 0.f, b = 0.f;
        if (vis) {
            warp_load_coeffs<K>(coeffs, g0, n, vis, sm, lane);
            __syncwarp();
            if (live) {
                float x = __ldg(dirs + g * 3 + 0), y = __ldg(dirs + g * 3 + 1), z = __ldg(dirs + g * 3 + 2);
                const float inv = 1.0f / fmaxf(sqrtf(x * x + y * y + z * z), 1e-12f);
                x *= inv; y *= inv; z *= inv;
                float Y[25];
                sh_basis<DEG>(x, y, z, Y);
                const float* c = sm + lane * STRIDE;
#pragma unroll
                for (int k = 0; k < NB; ++k) {
                    r += Y[k] * c[k * 3 + 0];
                    gg += Y[k] * c[k * 3 + 1];
                    b += Y[k] * c[k * 3 + 2];
                }
            }
            __syncwarp();
        }
        if (lane < n) {
            colors[g * 3 + 0] = r; colors[g * 3 + 1] = gg; colors[g * 3 + 2] = b;
        }
    }
}

template <int K, int DEG>
__global__ void __launch_bounds__(SH_THREADS)
sh_bwd_kernel(int64_t M, const float* __restrict__ dirs, const float* __restrict__ coeffs,
              const uint8_t* __restrict__ masks, const float* __restrict__ v_colors,
              float* __restrict__ v_coeffs, float* __restrict__ v_dirs) {
    extern __shared__ float smem[];
    constexpr int NB = (DEG + 1) * (DEG + 1);
    constexpr int K3 = ShCfg<K>::K3;
    constexpr int STRIDE = ShCfg<K>::STRIDE;
    const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
    float* sm = smem + wib * ShCfg<K>::SMEM_FLOATS;
    const int64_t n_chunks = (M + 31) >> 5;
    for (int64_t chunk = (int64_t)blockIdx.x * SH_WARPS + wib; chunk < n_chunks;
         chunk += (int64_t)gridDim.x * SH_WARPS) {
        const int64_t g0 = chunk << 5;
        const int n = (int)min((int64_t)32, M - g0);
        const int64_t g = g0 + lane;
        const bool live = (lane < n) && (masks == nullptr || __ldg(masks + g) != 0);
        const unsigned vis = __ballot_sync(0xffffffffu, live);
        float x = 0.f, y = 0.f, z = 0.f, inv = 0.f, vr = 0.f, vg = 0.f, vb = 0.f;
        float Y[25];
        if (live) {
            x = __ldg(dirs + g * 3 + 0); y = __ldg(dirs + g * 3 + 1); z = __ldg(dirs + g * 3 + 2);
            inv = 1.0f / fmaxf(sqrtf(x * x + y * y + z * z), 1e-12f);
            x *= inv; y *= inv; z *= inv;
            vr = __ldg(v_colors + g * 3 + 0); vg = __ldg(v_colors + g * 3 + 1); vb = __ldg(v_colors + g * 3 + 2);
            sh_basis<DEG>(x, y, z, Y);
        }
        float gx = 0.f, gy = 0.f, gz = 0.f;
        if (v_dirs != nullptr && vis) {
            // s[k] = <coeff[k,:], v_color> needs the coefficients
            warp_load_coeffs<K>(coeffs, g0, n, vis, sm, lane);
            __syncwarp();
            if (live) {
                const float* c = sm + lane * STRIDE;
                float s[25];
#pragma unroll
                for (int k = 0; k < NB; ++k) s[k] = c[k * 3 + 0] * vr + c[k * 3 + 1] * vg + c[k * 3 + 2] * vb;
                float ax, ay, az;
                sh_basis_vjp<DEG>(x, y, z, s, ax, ay, az);
                const float dp = ax * x + ay * y + az * z;
                gx = (ax - x * dp) * inv; gy = (ay - y * dp) * inv; gz = (az - z * dp) * inv;
            }
            __syncwarp();
        }
        // v_coeffs: Y_k * v_color, zeros for k >= NB and for masked Gaussians
        if (lane < n) {
            float* c = sm + lane * STRIDE;
            if (live) {
#pragma unroll
                for (int k = 0; k < NB; ++k) {
                    c[k * 3 + 0] = Y[k] * vr; c[k * 3 + 1] = Y[k] * vg; c[k * 3 + 2] = Y[k] * vb;
                }
                for (int e = NB * 3; e < K3; ++e) c[e] = 0.f;
            } else {
                for (int e = 0; e < K3; ++e) c[e] = 0.f;
            }
        }
        __syncwarp();
        warp_store_coeffs<K>(v_coeffs, g0, n, sm, lane);
        __syncwarp();
        if (v_dirs != nullptr && lane < n) {
            v_dirs[g * 3 + 0] = gx; v_dirs[g * 3 + 1] = gy; v_dirs[g * 3 + 2] = gz;
        }
    }
}

template <int K, int DEG>
int launch_fwd(int64_t M, const float* dirs, const float* coeffs, const uint8_t* masks, float* colors,
               cudaStream_t st) {
    const int smem = SH_WARPS * ShCfg<K>::SMEM_FLOATS * (int)sizeof(float);
    auto kern = sh_fwd_kernel<K, DEG>;
    if (smem > 48 * 1024) SC_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    const int64_t chunks = (M + 31) >> 5;
    const int grid = (int)min(sc_div_up(chunks, SH_WARPS), (int64_t)SC_NUM_SMS * 16);
    kern<<<grid, SH_THREADS, smem, st>>>(M, dirs, coeffs, masks, colors);
    SC_CHECK_LAUNCH("sh_fwd_kernel");
    return SC_OK;
}

template <int K, int DEG>
int launch_bwd(int64_t M, const float* dirs, const float* coeffs, const uint8_t* masks, const float* v_colors,
               float* v_coeffs, float* v_dirs, cudaStream_t st) {
    const int smem = SH_WARPS * ShCfg<K>::SMEM_FLOATS * (int)sizeof(float);
    auto kern = sh_bwd_kernel<K, DEG>;
    if (smem > 48 * 1024) SC_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    const int64_t chunks = (M + 31) >> 5;
    const int grid = (int)min(sc_div_up(chunks, SH_WARPS), (int64_t)SC_NUM_SMS * 16);
    kern<<<grid, SH_THREADS, smem, st>>>(M, dirs, coeffs, masks, v_colors, v_coeffs, v_dirs);
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
    SH_DISPATCH(launch_fwd, M, dirs, coeffs, masks, colors, st)
    sc_set_error("unsupported SH coefficient count K=%d (supported 1,4,9,16,25)", K);
    return SC_ERR_UNSUPPORTED;
}

extern "C" int sc_sh_bwd(int64_t M, int K, int degree, const float* dirs, const float* coeffs,
                         const uint8_t* masks, const float* v_colors, float* v_coeffs, float* v_dirs,
                         void* stream) {
    SC_CHECK_ARG(M >= 0, "M");
    SC_CHECK_ARG(degree >= 0 && degree <= 4 && (degree + 1) * (degree + 1) <= K, "degree vs K");
    if (M == 0) return SC_OK;
    SC_CHECK_ARG(dirs && coeffs && v_colors && v_coeffs, "null pointer");
    SC_CHECK_ARG(((uintptr_t)coeffs & 15) == 0 && ((uintptr_t)v_coeffs & 15) == 0,
                 "16-byte alignment of coeffs");
    cudaStream_t st = (cudaStream_t)stream;
    SH_DISPATCH(launch_bwd, M, dirs, coeffs, masks, v_colors, v_coeffs, v_dirs, st)
    sc_set_error("unsupported SH coefficient count K=%d (supported 1,4,9,16,25)", K);
    return SC_ERR_UNSUPPORTED;
}
