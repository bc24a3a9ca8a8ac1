// Order-preserving row compaction of many arrays at once (the
// prune / clone / split half): what the reference does with one boolean-index kernel chain per
// tensor — parameter, Adam exp_avg and exp_avg_sq of every attribute, plus the densification
// statistics (/root/reference/street_gaussian/models/gaussian_model.py:363-382 prune_optimizer,
// :416-431 prune_points) — as ONE launch over a table of (src, dst, row width) entries that share
// the keep mask and its exclusive prefix sum.  Rows keep their relative order, so parameters and
// optimizer moments stay aligned.  HBM-bound: every kept float is read once and written once.
#include "common.cuh"
#include "row_move.cuh"

#include "../../include/scb200.h"

namespace {

struct RowTable {
    const float* src[SC_COMPACT_MAX_ARRAYS];
    float* dst[SC_COMPACT_MAX_ARRAYS];
    int width[SC_COMPACT_MAX_ARRAYS];
    unsigned block_end[SC_COMPACT_MAX_ARRAYS];  // inclusive prefix of the CTAs each array owns
    int rows_per_cta[SC_COMPACT_MAX_ARRAYS];    // > 0: ragged width staged through shared memory
    int n;
};

// One flat grid over all arrays: a CTA finds its array in the (<= 32 entry) prefix table.  A thread moves one float4
// of one row when the row width and both bases allow it (width < 0 in the table marks the other case), else one
// float per access with consecutive lanes on consecutive floats of the flattened array — ragged widths (xyz 3,
// f_rest 45) still move whole sectors, where 4-float strided pieces per thread made the L2 fill every partially
// written sector from DRAM.
constexpr int COMPACT_EPT = 4;   // element mode: floats per thread (a CTA covers 1024 consecutive floats)

__global__ void __launch_bounds__(256)
compact_rows_kernel(int64_t N, const uint8_t* __restrict__ keep, const int64_t* __restrict__ dest, RowTable T) {
    int a = 0;
    while (a < T.n - 1 && blockIdx.x >= T.block_end[a]) ++a;
    const unsigned b0 = a ? T.block_end[a - 1] : 0u;
    const float* __restrict__ src = T.src[a];
    float* __restrict__ dst = T.dst[a];
    if (T.width[a] > 0) {
        const int w = T.width[a], pieces = w >> 2;
        const int64_t idx = (int64_t)(blockIdx.x - b0) * 256 + threadIdx.x;
        const int64_t row = idx / pieces;
        if (row >= N || !keep[row]) return;
        const int c = (int)(idx - row * pieces) * 4;
        *reinterpret_cast<float4*>(dst + __ldg(dest + row) * w + c) =
            __ldg(reinterpret_cast<const float4*>(src + row * w + c));
        return;
    }
    const int w = -T.width[a];
    if (T.rows_per_cta[a] > 0) {    // ragged width, staged through shared memory (row_move.cuh)
        rowmove_staged_cta(N, dest, src, dst, w, T.rows_per_cta[a], blockIdx.x - b0);
        return;
    }
    const int64_t base = (int64_t)(blockIdx.x - b0) * (256 * COMPACT_EPT);
    const int64_t row0 = base / w, total = N * w;
    const int rem0 = (int)(base - row0 * w);
    // all loads of the four elements are independent of each other (the source is read whether or not the row is
    // kept: ~all rows are) and issue before the first store
    float v[COMPACT_EPT];
    int64_t shift[COMPACT_EPT];
    bool k[COMPACT_EPT];
#pragma unroll
    for (int j = 0; j < COMPACT_EPT; ++j) {
        const int64_t e = base + j * 256 + threadIdx.x;
        const bool in = e < total;
        const int64_t row = row0 + (rem0 + j * 256 + (int)threadIdx.x) / w;
        v[j] = in ? __ldg(src + e) : 0.f;
        k[j] = in && keep[row];
        shift[j] = in ? (__ldg(dest + row) - row) * w : 0;
    }
#pragma unroll
    for (int j = 0; j < COMPACT_EPT; ++j)
        if (k[j]) dst[base + j * 256 + threadIdx.x + shift[j]] = v[j];
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
            const int w = row_floats[a];
            const bool vec = (w & 3) == 0 && ((((uintptr_t)src[a]) | ((uintptr_t)dst[a])) & 15) == 0;
            T.src[a] = src[a]; T.dst[a] = dst[a]; T.width[a] = vec ? w : -w;
            const int R = vec ? 0 : rowmove_rows_per_cta(w);
            T.rows_per_cta[a] = R;
            blocks += vec ? sc_div_up(N * (w / 4), 256) : R > 0 ? sc_div_up(N, R) : sc_div_up(N * w, 256 * COMPACT_EPT);
            SC_CHECK_ARG(blocks < (1ll << 31), "N * row width");
        } else {
            T.src[a] = nullptr; T.dst[a] = nullptr; T.width[a] = 1; T.rows_per_cta[a] = 0;
        }
        T.block_end[a] = (unsigned)blocks;
    }
    compact_rows_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(N, keep, dest_index, T);
    SC_CHECK_LAUNCH("compact_rows_kernel");
    return SC_OK;
}
