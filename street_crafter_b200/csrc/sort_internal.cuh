// Internal (C++ linkage) interface of the radix sorter in scan_sort.cu, used by the fused tile
// binning pipeline in projection.cu.  Not part of the C ABI.
#pragma once
#include "common.cuh"

constexpr int SC_SORT_RADIX = 256;
constexpr int SC_SORT_MAX_PASSES = 8;

// pass p sorts key bits [shift[p], shift[p] + nbits[p])
struct ScPassPlan {
    int passes;
    int shift[SC_SORT_MAX_PASSES];
    int nbits[SC_SORT_MAX_PASSES];
};

// layout of the sort workspace: [digit histograms | tickets | look-back words]
struct ScSortPlan {
    ScPassPlan pp;
    int64_t tiles;         // onesweep tiles at the maximum element count
    int64_t hist_off;      // u32 [MAX_PASSES][RADIX] raw digit counts, scanned in place by run()
    int64_t ticket_off;    // u32 [MAX_PASSES]
    int64_t lookback_off;  // u32 [passes][tiles][RADIX]
    int64_t ctrl_bytes;    // prefix of the workspace that must be ZERO before the histograms are built
    int64_t total;
};

// key_bytes: 4 or 8 (tile size differs per key width); 0 = workspace upper bound for either
ScSortPlan sc_sort_make_plan(int64_t n_max, int begin_bit, int end_bit, int key_bytes);

// accumulate the digit histograms of `keys` into ws + hist_off (workspace control region zeroed
// by the caller).  n_dev (nullable) overrides n_max with a device-resident count <= n_max.
int sc_sort_histogram_u32(int64_t n_max, const int64_t* n_dev, const uint32_t* keys, const ScSortPlan& plan,
                          void* ws, cudaStream_t st);
int sc_sort_histogram_u64(int64_t n_max, const int64_t* n_dev, const uint64_t* keys, const ScSortPlan& plan,
                          void* ws, cudaStream_t st);

// scan the histograms and run the passes.  Returns 0 if the result is in keys_out/vals_out,
// 1 if it is in keys_in/vals_in, < 0 on error.
// hist_raw: the histograms in ws are raw counts accumulated by the caller's own kernel (the passes
// scan them in their prologue; no scan launch).
int sc_sort_run_u32(int64_t n_max, const int64_t* n_dev, uint32_t* keys_in, uint32_t* vals_in,
                    uint32_t* keys_out, uint32_t* vals_out, const ScSortPlan& plan, void* ws, cudaStream_t st,
                    bool hist_raw = false);
int sc_sort_run_u64(int64_t n_max, const int64_t* n_dev, uint64_t* keys_in, uint32_t* vals_in,
                    uint64_t* keys_out, uint32_t* vals_out, const ScSortPlan& plan, void* ws, cudaStream_t st);
