// The densification cycle (clone / split half): the reference's
// `densify_and_prune` (/root/reference/street_gaussian/models/gaussian_model.py:521-551) runs
// densify_and_clone (:493-519), densify_and_split (:452-491), two cat_optimizer rebuilds (:384-408) and two
// prune_points rebuilds (:416-431) — about 150 boolean-index / cat / repeat launches that copy every parameter
// and both Adam moments four times.  Every row of the result descends from exactly ONE input row:
//
//   stay   : the row itself            (not split, not pruned)            parameters + moments copied
//   clone  : a copy of the row         (clone-selected, not pruned)       parameters copied, moments 0
//   child c: c-th sample of a split row (split-selected, child not pruned) xyz = R(q)(z*std) + xyz,
//            c in [0, n_copies)                                           scaling = log(exp(s)/(0.8 n_copies)),
//                                                                         the rest copied, moments 0
//
// laid out as [stay rows | clones | children 0 | children 1 ...], each block in input order — the order the
// reference's cat / repeat / boolean-index chain produces.  So the cycle is: ONE classification pass (28 B read,
// 16 B written per row), ONE exclusive scan over the four flag segments, and ONE move pass that reads every
// input float once and writes every output float once.  Clone and split are mutually exclusive (max scale <= / >
// percent_dense * extent), and a clone is never split (same scale as its source).  HBM-bound.
//
// Compiled WITHOUT fast math: expf / logf / the divisions are the accurate ones torch's exp / log / sigmoid use.
#include "common.cuh"
#include "row_move.cuh"

namespace {

struct ClassifyParams {
    float max_grad, clone_max_scale, split_min_scale, min_opacity, big_ws_scale, split_shrink;
    int prune_big, grad_column;
};

__global__ void __launch_bounds__(256)
densify_classify_kernel(int64_t N, const float2* __restrict__ accum, const float* __restrict__ denom,
                        const float* __restrict__ scaling, const float* __restrict__ opacity, ClassifyParams P,
                        int32_t* __restrict__ flags, int* __restrict__ n_clone_selected) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    bool clone_sel = false;
    if (i < N) {
        const float2 ac = __ldg(accum + i);
        float g = (P.grad_column ? ac.y : ac.x) / __ldg(denom + i);            // :522 (column 1: the actor / background models' abs-grad option)
        if (g != g) g = 0.0f;                                                  // :523
        const float s0 = __ldg(scaling + 3 * i), s1 = __ldg(scaling + 3 * i + 1), s2 = __ldg(scaling + 3 * i + 2);
        const float e0 = expf(s0), e1 = expf(s1), e2 = expf(s2);
        const float smax = fmaxf(e0, fmaxf(e1, e2));
        const float sig = 1.0f / (1.0f + expf(-__ldg(opacity + i)));
        const bool low = sig < P.min_opacity;                                  // :533
        clone_sel = (fabsf(g) >= P.max_grad) && (smax <= P.clone_max_scale);   // :496-498
        const bool split_sel = (g >= P.max_grad) && (smax > P.split_min_scale);  // :461-463
        const bool pruned_self = low || (P.prune_big && smax > P.big_ws_scale);  // :537-539
        // children carry log(exp(s) / shrink) and are tested on exp() of it (:473, :537)
        const float c0 = expf(logf(e0 / P.split_shrink)), c1 = expf(logf(e1 / P.split_shrink)),
                    c2 = expf(logf(e2 / P.split_shrink));
        const bool pruned_child = low || (P.prune_big && fmaxf(c0, fmaxf(c1, c2)) > P.big_ws_scale);
        flags[i] = (!split_sel && !pruned_self) ? 1 : 0;
        flags[N + i] = (clone_sel && !pruned_self) ? 1 : 0;
        flags[2 * N + i] = (split_sel && !pruned_child) ? 1 : 0;
        flags[3 * N + i] = split_sel ? 1 : 0;
    }
    const unsigned b = __ballot_sync(0xffffffffu, clone_sel);
    if ((threadIdx.x & 31) == 0 && b) atomicAdd(n_clone_selected, __popc(b));
}

struct MoveTable {
    const float* src[SC_COMPACT_MAX_ARRAYS];
    float* dst[SC_COMPACT_MAX_ARRAYS];
    int width[SC_COMPACT_MAX_ARRAYS];
    int kind[SC_COMPACT_MAX_ARRAYS];        // SC_DENSIFY_* | MOVE_VEC
    unsigned block_end[2 * SC_COMPACT_MAX_ARRAYS];   // entry 2a: the `stay` scatter of array a, 2a+1: its appended rows
    int rows_per_cta[SC_COMPACT_MAX_ARRAYS];         // > 0: ragged width, `stay` rows staged through shared memory
    int n;
};
constexpr int MOVE_VEC = 16;        // row width a multiple of 4 floats and both bases 16-byte aligned
constexpr int MOVE_EPT = 4;         // element mode: floats per thread (a CTA covers 1024 consecutive floats)

struct SplitCtx {
    int64_t n_stay, n_clone, n_child, n_split;
    int n_copies;
    const float* noise;
    const float* scaling;
    const float* rotation;
    float shrink;
    const int32_t* clone_src;   // [n_clone]  source row of every surviving clone
    const int32_t* child_src;   // [n_child]  source row of every surviving split row
    const int32_t* child_rank;  // [n_child]  its rank among the split-SELECTED rows (indexes the noise)
};

// Source rows of the appended blocks, in output order (the inverse of the scan): lets the move kernel GATHER the
// clones and children with fully coalesced stores.  (Scattering them from the source side wrote every 12..180-byte
// row at a different time from its neighbours; the L2 filled the partially written sectors from DRAM and the
// kernel read twice its input.)
__global__ void __launch_bounds__(256)
densify_index_kernel(int64_t N, const int32_t* __restrict__ flags, const int64_t* __restrict__ dest, int64_t n_stay,
                     int64_t n_clone, int32_t* __restrict__ clone_src, int32_t* __restrict__ child_src,
                     int32_t* __restrict__ child_rank) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= N) return;
    if (flags[N + i]) clone_src[__ldg(dest + N + i) - n_stay] = (int32_t)i;
    if (flags[2 * N + i]) {
        const int64_t r = __ldg(dest + 2 * N + i) - n_stay - n_clone;
        child_src[r] = (int32_t)i;
        child_rank[r] = (int32_t)(__ldg(dest + 3 * N + i) - __ldg(dest + 3 * N));
    }
}

// c-th child value of one xyz component (row k of R(q) times (z * std), plus the parent's component)
__device__ __forceinline__ float child_xyz(const SplitCtx& S, int64_t row, int64_t rank, int cp, int k, float v) {
    const float qw = __ldg(S.rotation + 4 * row), qx = __ldg(S.rotation + 4 * row + 1),
                qy = __ldg(S.rotation + 4 * row + 2), qz = __ldg(S.rotation + 4 * row + 3);
    const float nrm = sqrtf(qw * qw + qx * qx + qy * qy + qz * qz);                  // general_utils.py:126-128
    const float r = qw / nrm, x = qx / nrm, y = qy / nrm, z = qz / nrm;
    float R0, R1, R2;                                                                // :137-145
    if (k == 0) { R0 = 1.f - 2.f * (y * y + z * z); R1 = 2.f * (x * y - r * z); R2 = 2.f * (x * z + r * y); }
    else if (k == 1) { R0 = 2.f * (x * y + r * z); R1 = 1.f - 2.f * (x * x + z * z); R2 = 2.f * (y * z - r * x); }
    else { R0 = 2.f * (x * z - r * y); R1 = 2.f * (y * z + r * x); R2 = 1.f - 2.f * (x * x + y * y); }
    const float* zn = S.noise + 3 * (cp * S.n_split + rank);
    const float p0 = __ldg(zn) * expf(__ldg(S.scaling + 3 * row)),                   // gaussian_model.py:469-470
                p1 = __ldg(zn + 1) * expf(__ldg(S.scaling + 3 * row + 1)),
                p2 = __ldg(zn + 2) * expf(__ldg(S.scaling + 3 * row + 2));
    return (R0 * p0 + R1 * p1 + R2 * p2) + v;                                        // :472
}

// One flat grid over all arrays, two CTA ranges per array.
//  * `stay` rows: an order-preserving compaction exactly like compact_rows_kernel — a float4 per thread when the
//    row width allows it, otherwise one float per lane access over the flattened array (ragged widths such as
//    f_rest's 45 floats still move whole sectors).
//  * appended rows [clones | children of copy 0 | copy 1 ...]: one float per lane access over the flattened OUTPUT
//    block, source row looked up in the index lists; moments get zeros (no read at all).
__global__ void __launch_bounds__(256)
densify_move_kernel(int64_t N, const int32_t* __restrict__ flags, const int64_t* __restrict__ dest, SplitCtx S,
                    MoveTable T) {
    int t = 0;
    while (t < 2 * T.n - 1 && blockIdx.x >= T.block_end[t]) ++t;
    const unsigned b0 = t ? T.block_end[t - 1] : 0u;
    const int a = t >> 1;
    const int w = T.width[a], kind = T.kind[a] & 15;
    const float* __restrict__ src = T.src[a];
    float* __restrict__ dst = T.dst[a];
    if (t & 1) {    // appended rows, output side
        const int64_t total = (S.n_clone + S.n_copies * S.n_child) * w;
        const int64_t base = (int64_t)(blockIdx.x - b0) * (256 * MOVE_EPT) * (kind == SC_DENSIFY_MOMENT ? 4 : 1);
        const int64_t row0 = base / w;
        const int rem0 = (int)(base - row0 * w);
        float* __restrict__ out = dst + S.n_stay * w;
        if (kind == SC_DENSIFY_MOMENT) {    // cat_optimizer :396-397: zeros — float4 stores on 16-byte boundaries
            const int shift = (int)((((uintptr_t)out) >> 2) & 3);          // floats past the previous boundary
            const int64_t f0 = base + 4 * (int64_t)threadIdx.x - shift;   // base is a multiple of 4096
#pragma unroll
            for (int j = 0; j < MOVE_EPT; ++j) {
                const int64_t f = f0 + (int64_t)j * 1024;                  // this thread's 4 floats: [f, f + 4)
                if (f >= 0 && f + 3 < total) {
                    *reinterpret_cast<float4*>(out + f) = make_float4(0.f, 0.f, 0.f, 0.f);
                } else {
#pragma unroll
                    for (int k = 0; k < 4; ++k)
                        if (f + k >= 0 && f + k < total) out[f + k] = 0.f;
                }
            }
            return;
        }
#pragma unroll
        for (int j = 0; j < MOVE_EPT; ++j) {
            const int64_t e = base + j * 256 + threadIdx.x;
            if (e >= total) break;
            const int L = rem0 + j * 256 + (int)threadIdx.x;
            const int dr = L / w, c = L - dr * w;
            const int64_t orow = row0 + dr;
            if (orow < S.n_clone) {
                out[e] = __ldg(src + (int64_t)__ldg(S.clone_src + orow) * w + c);
            } else {
                const unsigned q = (unsigned)(orow - S.n_clone), nc = (unsigned)S.n_child;   // appended rows < 2^31 (host check)
                const int cp = (int)(q / nc);
                const int64_t r = q - cp * nc;
                const int64_t row = __ldg(S.child_src + r);
                const float v = __ldg(src + row * w + c);
                float o = v;
                if (kind == SC_DENSIFY_XYZ) o = child_xyz(S, row, __ldg(S.child_rank + r), cp, c, v);
                else if (kind == SC_DENSIFY_SCALING) o = logf(expf(v) / S.shrink);   // :473
                out[e] = o;
            }
        }
        return;
    }
    if (T.kind[a] & MOVE_VEC) {
        const int pieces = w >> 2;
        const int64_t idx = (int64_t)(blockIdx.x - b0) * 256 + threadIdx.x;
        const int64_t row = idx / pieces;
        if (row >= N || !flags[row]) return;
        const int c = (int)(idx - row * pieces) * 4;
        *reinterpret_cast<float4*>(dst + __ldg(dest + row) * w + c) =
            __ldg(reinterpret_cast<const float4*>(src + row * w + c));
        return;
    }
    if (T.rows_per_cta[a] > 0) {    // dest[0 .. N] is the scan of the `stay` flags (dest[N] = n_stay)
        rowmove_staged_cta(N, dest, src, dst, w, T.rows_per_cta[a], blockIdx.x - b0);
        return;
    }
    const int64_t base = (int64_t)(blockIdx.x - b0) * (256 * MOVE_EPT);
    const int64_t row0 = base / w, total = N * w;
    const int rem0 = (int)(base - row0 * w);
    // all loads of the four elements are independent of each other (the source is read whether or not its row
    // stays: ~all rows do) and issue before the first store
    float v[MOVE_EPT];
    int64_t shift[MOVE_EPT];
    bool k[MOVE_EPT];
#pragma unroll
    for (int j = 0; j < MOVE_EPT; ++j) {
        const int64_t e = base + j * 256 + threadIdx.x;
        const bool in = e < total;
        const int64_t row = row0 + (rem0 + j * 256 + (int)threadIdx.x) / w;
        v[j] = in ? __ldg(src + e) : 0.f;
        k[j] = in && flags[row];
        shift[j] = in ? (__ldg(dest + row) - row) * w : 0;
    }
#pragma unroll
    for (int j = 0; j < MOVE_EPT; ++j)
        if (k[j]) dst[base + j * 256 + threadIdx.x + shift[j]] = v[j];
}

}  // namespace

extern "C" int sc_densify_classify(int64_t N, const float* xyz_gradient_accum, const float* denom,
                                   const float* scaling, const float* opacity, float max_grad, float clone_max_scale,
                                   float split_min_scale, float min_opacity, int prune_big, float big_ws_scale,
                                   float split_shrink, int grad_column, int32_t* flags, int32_t* n_clone_selected, void* stream) {
    SC_CHECK_ARG(N >= 0 && 4 * N < (1ll << 31) * 256, "sizes");
    SC_CHECK_ARG(n_clone_selected, "null pointer");
    SC_CUDA(cudaMemsetAsync(n_clone_selected, 0, sizeof(int32_t), (cudaStream_t)stream));
    if (N == 0) return SC_OK;
    SC_CHECK_ARG(xyz_gradient_accum && denom && scaling && opacity && flags, "null pointer");
    SC_CHECK_ARG(((uintptr_t)xyz_gradient_accum & 7) == 0, "8-byte alignment of xyz_gradient_accum");
    SC_CHECK_ARG(split_shrink > 0.f && (grad_column == 0 || grad_column == 1), "split_shrink / grad_column");
    ClassifyParams P{max_grad, clone_max_scale, split_min_scale, min_opacity, big_ws_scale, split_shrink, prune_big, grad_column};
    densify_classify_kernel<<<(unsigned)sc_div_up(N, 256), 256, 0, (cudaStream_t)stream>>>(
        N, reinterpret_cast<const float2*>(xyz_gradient_accum), denom, scaling, opacity, P, flags, n_clone_selected);
    SC_CHECK_LAUNCH("densify_classify_kernel");
    return SC_OK;
}

extern "C" int sc_densify_move(int64_t N, const int32_t* flags, const int64_t* dest_index, int64_t n_stay,
                               int64_t n_clone, int64_t n_child, int64_t n_split, int n_copies, const float* noise,
                               const float* scaling, const float* rotation, float split_shrink, int32_t* index_ws,
                               int n_arrays, const float* const* src, float* const* dst, const int* row_floats,
                               const int* kinds, void* stream) {
    SC_CHECK_ARG(N >= 0 && N < (1ll << 31) && n_arrays >= 0 && n_arrays <= SC_COMPACT_MAX_ARRAYS && n_copies >= 1,
                 "sizes");
    SC_CHECK_ARG(n_stay >= 0 && n_clone >= 0 && n_child >= 0 && n_split >= n_child && n_stay + n_split <= N,
                 "stay / clone / child / split counts");
    if (N == 0 || n_arrays == 0) return SC_OK;
    SC_CHECK_ARG(flags && dest_index && src && dst && row_floats && kinds, "null pointer");
    const int64_t n_app = n_clone + (int64_t)n_copies * n_child;
    SC_CHECK_ARG(n_app == 0 || index_ws, "index workspace");
    SC_CHECK_ARG(n_app < (1ll << 31), "appended row count");
    SplitCtx S{n_stay, n_clone, n_child, n_split, n_copies, noise, scaling, rotation, split_shrink,
               index_ws, index_ws + n_clone, index_ws + n_clone + n_child};
    if (n_app > 0) {
        densify_index_kernel<<<(unsigned)sc_div_up(N, 256), 256, 0, (cudaStream_t)stream>>>(
            N, flags, dest_index, n_stay, n_clone, index_ws, index_ws + n_clone, index_ws + n_clone + n_child);
        SC_CHECK_LAUNCH("densify_index_kernel");
    }
    MoveTable T;
    T.n = n_arrays;
    int64_t blocks = 0;
    for (int a = 0; a < SC_COMPACT_MAX_ARRAYS; ++a) {
        if (a < n_arrays) {
            const int w = row_floats[a], k = kinds[a];
            SC_CHECK_ARG(src[a] && dst[a] && w > 0, "array table entry");
            SC_CHECK_ARG(k >= SC_DENSIFY_PARAM && k <= SC_DENSIFY_SCALING, "array kind");
            if (k == SC_DENSIFY_XYZ || k == SC_DENSIFY_SCALING) SC_CHECK_ARG(w == 3, "xyz / scaling rows are 3 floats");
            if (k == SC_DENSIFY_XYZ && n_split > 0) SC_CHECK_ARG(noise && scaling && rotation, "split inputs");
            const bool vec = (w & 3) == 0 && ((((uintptr_t)src[a]) | ((uintptr_t)dst[a])) & 15) == 0;
            T.src[a] = src[a]; T.dst[a] = dst[a]; T.width[a] = w; T.kind[a] = k | (vec ? MOVE_VEC : 0);
            const int R = vec ? 0 : rowmove_rows_per_cta(w);
            T.rows_per_cta[a] = R;
            blocks += n_stay == 0 ? 0 : vec ? sc_div_up(N * (w / 4), 256) : R > 0 ? sc_div_up(N, R) : sc_div_up(N * w, 256 * MOVE_EPT);
            T.block_end[2 * a] = (unsigned)blocks;
            blocks += k == SC_DENSIFY_MOMENT ? sc_div_up(n_app * w + 3, 1024 * MOVE_EPT) : sc_div_up(n_app * w, 256 * MOVE_EPT);
            SC_CHECK_ARG(blocks < (1ll << 31), "N * row width");
        } else {
            T.src[a] = nullptr; T.dst[a] = nullptr; T.width[a] = 1; T.kind[a] = 0; T.rows_per_cta[a] = 0;
            T.block_end[2 * a] = (unsigned)blocks;
        }
        T.block_end[2 * a + 1] = (unsigned)blocks;
    }
    if (blocks == 0) return SC_OK;
    densify_move_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(N, flags, dest_index, S, T);
    SC_CHECK_LAUNCH("densify_move_kernel");
    return SC_OK;
}
