// FP32-FMA and SFU(ex2) issue-rate probes: the measured denominators of the compositing
// kernels' roofline (MEASURED_PEAKS.json only carries HBM and bf16-tensor peaks).
#include "common.cuh"

namespace {
// Blackwell packed FP32 (PTX fma.rn.f32x2 -> SASS FFMA2): two FMAs per lane and ISSUE SLOT
__device__ __forceinline__ unsigned long long probe_pack(float a, float b) {
    unsigned long long r;
    asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(a), "f"(b));
    return r;
}
__global__ void probe_ffma2_kernel(int iters, float* __restrict__ sink) {
    unsigned long long a[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) a[k] = probe_pack(1.0f + 1e-3f * (float)(threadIdx.x + k), 1.0f - 1e-3f * (float)k);
    const unsigned long long m = probe_pack(0.999f + 1e-9f * (float)blockIdx.x, 0.998f), c = probe_pack(1e-4f, 2e-4f);
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int k = 0; k < 8; ++k) asm("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(a[k]) : "l"(m), "l"(c));
    }
    float s = 0.f;
#pragma unroll
    for (int k = 0; k < 8; ++k) {
        float lo, hi;
        asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(a[k]));
        s += lo + hi;
    }
    sink[(size_t)blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <int KIND>
__global__ void probe_kernel(int iters, float* __restrict__ sink) {
    float a[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) a[k] = 1.0f + 1e-3f * (float)(threadIdx.x + k);
    const float m = 0.999f + 1e-9f * (float)blockIdx.x, c = 1e-4f;
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            if (KIND == 0) a[k] = fmaf(a[k], m, c);
            else a[k] = exp2f(a[k] * 0.5f - 1.0f);
        }
    }
    float s = 0.f;
#pragma unroll
    for (int k = 0; k < 8; ++k) s += a[k];
    sink[(size_t)blockIdx.x * blockDim.x + threadIdx.x] = s;
}
}  // namespace

extern "C" int sc_probe_fp32(int kind, int grid, int block, int iters, float* sink, void* stream) {
    SC_CHECK_ARG(grid > 0 && block > 0 && block <= 1024 && iters > 0 && sink, "probe args");
    if (kind == 2) probe_ffma2_kernel<<<grid, block, 0, (cudaStream_t)stream>>>(iters, sink);   // 32 FLOP / iteration and lane
    else if (kind == 0) probe_kernel<0><<<grid, block, 0, (cudaStream_t)stream>>>(iters, sink);
    else probe_kernel<1><<<grid, block, 0, (cudaStream_t)stream>>>(iters, sink);
    SC_CHECK_LAUNCH("probe_kernel");
    return SC_OK;
}
