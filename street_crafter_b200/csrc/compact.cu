// Order-preserving row compaction of many arrays at once (the
// prune / clone / split half): what the reference does with one boolean-index kernel chain per
// tensor — parameter, Adam exp_avg and exp_avg_sq of every attribute, plus the densification
// statistics (/root/reference/street_gaussian/models/gaussian_model.py:363-382 prune_optimizer,
// :416-431 prune_points) — as ONE launch over a table of (src, dst, row width) entries that share
// the keep mask and its exclusive prefix sum.  Rows keep their relative order, so parameters and
// optimizer moments stay aligned.  HBM-bound: every kept float is read once and written once.
#include "common.cuh"

#include "../../include/scb200.h"

namespace {

struct RowTable {
    const float* src[SC_COMPACT_MAX_ARRAYS];
    float* dst[SC_COMPACT_MAX_ARRAYS];
    int width[SC_COMPACT_MAX_ARRAYS];
    unsigned block_end[SC_COMPACT_MAX_ARRAYS];  // inclusive prefix of the CTAs each array owns
    int n;
};

// One flat grid over all arrays: a CTA finds its array in the (<= 32 entry) prefix table; a thread moves one
// 4-float piece of one row (a float4 when the row width and both bases allow it).
__global__ void __launch_bounds__(256)
compact_rows_kernel(int64_t N, const uint8_t* __restrict__ keep, const int64_t* __restrict__ dest, RowTable T) {
    int a = 0;
    while (a < T.n - 1 && blockIdx.x >= T.block_end[a]) ++a;
    const unsigned b0 = a ? T.block_end[a - 1] : 0u;
    const int w = T.width[a];
    const int pieces = (w + 3) >> 2;
    const int64_t idx = (int64_t)(blockIdx.x - b0) * 256 + threadIdx.x;
    const int64_t row = idx / pieces;
    if (row >= N || !keep[row]) return;
    const int c = (int)(idx - row * pieces) * 4;
    const float* s = T.src[a] + row * w + c;
    float* d = T.dst[a] + __ldg(dest + row) * w + c;
    if (((w & 3) == 0) && ((((uintptr_t)T.src[a]) | ((uintptr_t)T.dst[a])) & 15) == 0) {
        *reinterpret_cast<float4*>(d) = __ldg(reinterpret_cast<const float4*>(s));
    } else {
#pragma unroll
        for (int k = 0; k < 4; ++k)
            if (c + k < w) d[k] = __ldg(s + k);
    }
}

}  // namespace

extern "C" int sc_compact_rows(int64_t N, const uint8_t* keep, const int64_t* dest_index, int n_arrays,
                               const float* const* src, float* const* dst, const int* row_floats, void* stream) {
    SC_CHECK_ARG(N >= 0 && n_arrays >= 0 && n_arrays <= SC_COMPACT_MAX_ARRAYS, "sizes");
    if (N == 0 || n_arrays == 0) return SC_OK;
    SC_CHECK_ARG(keep && dest_index && src && dst && row_floats, "null pointer");
    RowTable T;
    T.n = n_arrays;
    int64_t blocks = 0;
    for (int a = 0; a < SC_COMPACT_MAX_ARRAYS; ++a) {
        if (a < n_arrays) {
            SC_CHECK_ARG(src[a] && dst[a] && row_floats[a] > 0, "array table entry");
            T.src[a] = src[a]; T.dst[a] = dst[a]; T.width[a] = row_floats[a];
            blocks += sc_div_up(N * ((row_floats[a] + 3) / 4), 256);
            SC_CHECK_ARG(blocks < (1ll << 31), "N * row width");
        } else {
            T.src[a] = nullptr; T.dst[a] = nullptr; T.width[a] = 1;
        }
        T.block_end[a] = (unsigned)blocks;
    }
    compact_rows_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(N, keep, dest_index, T);
    SC_CHECK_LAUNCH("compact_rows_kernel");
    return SC_OK;
}
