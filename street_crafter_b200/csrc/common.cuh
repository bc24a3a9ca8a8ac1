// Shared helpers for libscb200 (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include "../../include/scb200.h"

#ifndef __CUDA_ARCH__
#define SC_HOST_ONLY 1
#endif

// thread-local error message + launch counter live in api.cu
void sc_set_error(const char* fmt, ...);
void sc_count_launch(int n = 1);

#define SC_CHECK_ARG(cond, msg)                                  \
    do {                                                         \
        if (!(cond)) {                                           \
            sc_set_error("bad argument: %s (%s)", msg, #cond);   \
            return SC_ERR_BAD_ARG;                               \
        }                                                        \
    } while (0)

#define SC_CHECK_LAUNCH(name)                                                        \
    do {                                                                             \
        cudaError_t e__ = cudaGetLastError();                                        \
        if (e__ != cudaSuccess) {                                                    \
            sc_set_error("CUDA error after %s: %s", name, cudaGetErrorString(e__));  \
            return SC_ERR_CUDA;                                                      \
        }                                                                            \
        sc_count_launch();                                                           \
    } while (0)

#define SC_CUDA(call)                                                                \
    do {                                                                             \
        cudaError_t e__ = (call);                                                    \
        if (e__ != cudaSuccess) {                                                    \
            sc_set_error("CUDA error in %s: %s", #call, cudaGetErrorString(e__));    \
            return SC_ERR_CUDA;                                                      \
        }                                                                            \
    } while (0)

static inline int64_t sc_div_up(int64_t a, int64_t b) { return (a + b - 1) / b; }
static inline int64_t sc_align_up(int64_t a, int64_t b) { return sc_div_up(a, b) * b; }

constexpr int SC_NUM_SMS = 148;  // B200

// Frozen constants of the compositing spec (see oracle/gsplat_torch.py).
#define SC_ALPHA_CAP 0.999f
#define SC_ALPHA_FLOOR (1.0f / 255.0f)
#define SC_T_STOP 1e-4f
#define SC_EIGEN_FLOOR 0.01f
#define SC_JAC_PAD 0.3f

// Lanes of the (full) warp whose 8-bit digit equals this lane's, built from 8 ballots.  The
// hardware match.any iterates over the distinct values in the warp (up to 32 rounds on random
// digits); the ballots are fixed-latency and independent across keys, so they pipeline.
#ifdef __CUDACC__
__device__ __forceinline__ unsigned sc_match_digit8(unsigned d) {
    unsigned peers = 0xffffffffu;
#pragma unroll
    for (int b = 0; b < 8; ++b) {
        const bool bit = (d >> b) & 1u;
        const unsigned bal = __ballot_sync(0xffffffffu, bit);
        peers &= bit ? bal : ~bal;
    }
    return peers;
}
#endif
