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

namespace {

struct ClassifyParams {
    float max_grad, clone_max_scale, split_min_scale, min_opacity, big_ws_scale, split_shrink;
    int prune_big;
};

__global__ void __launch_bounds__(256)
densify_classify_kernel(int64_t N, const float2* __restrict__ accum, const float* __restrict__ denom,
                        const float* __restrict__ scaling, const float* __restrict__ opacity, ClassifyParams P,
                        int32_t* __restrict__ flags, int* __restrict__ n_clone_selected) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    bool clone_sel = false;
    if (i < N) {
        float g = __ldg(&accum[i].x) / __ldg(denom + i);                       // :522
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
    int kind[SC_COMPACT_MAX_ARRAYS];
    unsigned block_end[SC_COMPACT_MAX_ARRAYS];
    int n;
};

__device__ __forceinline__ void store_piece(float* d, const float v[4], int c, int w, bool vec) {
    if (vec) {
        *reinterpret_cast<float4*>(d) = make_float4(v[0], v[1], v[2], v[3]);
    } else {
#pragma unroll
        for (int k = 0; k < 4; ++k)
            if (c + k < w) d[k] = v[k];
    }
}

// One flat grid over all arrays (as compact_rows_kernel): a thread owns one 4-float piece of one INPUT row and
// writes it to every output row that descends from it.
__global__ void __launch_bounds__(256)
densify_move_kernel(int64_t N, const int32_t* __restrict__ flags, const int64_t* __restrict__ dest, int64_t n_child,
                    int64_t n_split, int n_copies, const float* __restrict__ noise, const float* __restrict__ scaling,
                    const float* __restrict__ rotation, float split_shrink, MoveTable T) {
    int a = 0;
    while (a < T.n - 1 && blockIdx.x >= T.block_end[a]) ++a;
    const unsigned b0 = a ? T.block_end[a - 1] : 0u;
    const int w = T.width[a], kind = T.kind[a];
    const int pieces = (w + 3) >> 2;
    const int64_t idx = (int64_t)(blockIdx.x - b0) * 256 + threadIdx.x;
    const int64_t row = idx / pieces;
    if (row >= N) return;
    const bool stay = flags[row], clone = flags[N + row], child = flags[2 * N + row];
    if (!stay && !clone && !child) return;
    const int c = (int)(idx - row * pieces) * 4;
    const bool vec = ((w & 3) == 0) && ((((uintptr_t)T.src[a]) | ((uintptr_t)T.dst[a])) & 15) == 0;
    const float* s = T.src[a] + row * w + c;
    float v[4] = {0.f, 0.f, 0.f, 0.f};
    if (vec) {
        const float4 t = __ldg(reinterpret_cast<const float4*>(s));
        v[0] = t.x; v[1] = t.y; v[2] = t.z; v[3] = t.w;
    } else {
#pragma unroll
        for (int k = 0; k < 4; ++k)
            if (c + k < w) v[k] = __ldg(s + k);
    }
    const float zero[4] = {0.f, 0.f, 0.f, 0.f};
    const bool moment = kind == SC_DENSIFY_MOMENT;
    if (stay) store_piece(T.dst[a] + __ldg(dest + row) * w + c, v, c, w, vec);
    if (clone) store_piece(T.dst[a] + __ldg(dest + N + row) * w + c, moment ? zero : v, c, w, vec);
    if (child) {
        const int64_t d0 = __ldg(dest + 2 * N + row);
        const int64_t rank = __ldg(dest + 3 * N + row) - __ldg(dest + 3 * N);   // among the split-selected rows
        float R[9], sd[3];
        if (kind == SC_DENSIFY_XYZ) {   // width 3: this thread holds the whole row
            const float qw = __ldg(rotation + 4 * row), qx = __ldg(rotation + 4 * row + 1),
                        qy = __ldg(rotation + 4 * row + 2), qz = __ldg(rotation + 4 * row + 3);
            const float nrm = sqrtf(qw * qw + qx * qx + qy * qy + qz * qz);      // general_utils.py:126-128
            const float r = qw / nrm, x = qx / nrm, y = qy / nrm, z = qz / nrm;
            R[0] = 1.f - 2.f * (y * y + z * z); R[1] = 2.f * (x * y - r * z); R[2] = 2.f * (x * z + r * y);
            R[3] = 2.f * (x * y + r * z); R[4] = 1.f - 2.f * (x * x + z * z); R[5] = 2.f * (y * z - r * x);
            R[6] = 2.f * (x * z - r * y); R[7] = 2.f * (y * z + r * x); R[8] = 1.f - 2.f * (x * x + y * y);
#pragma unroll
            for (int k = 0; k < 3; ++k) sd[k] = expf(__ldg(scaling + 3 * row + k));
        }
        for (int cp = 0; cp < n_copies; ++cp) {
            float o[4] = {v[0], v[1], v[2], v[3]};
            if (moment) {
                o[0] = o[1] = o[2] = o[3] = 0.f;
            } else if (kind == SC_DENSIFY_XYZ) {
                const float* zn = noise + 3 * (cp * n_split + rank);
                const float p0 = __ldg(zn) * sd[0], p1 = __ldg(zn + 1) * sd[1], p2 = __ldg(zn + 2) * sd[2];  // :469-470
#pragma unroll
                for (int k = 0; k < 3; ++k) o[k] = (R[3 * k] * p0 + R[3 * k + 1] * p1 + R[3 * k + 2] * p2) + v[k];  // :472
            } else if (kind == SC_DENSIFY_SCALING) {
#pragma unroll
                for (int k = 0; k < 4; ++k) o[k] = logf(expf(v[k]) / split_shrink);                                  // :473
            }
            store_piece(T.dst[a] + (d0 + cp * n_child) * w + c, o, c, w, vec);
        }
    }
}

}  // namespace

extern "C" int sc_densify_classify(int64_t N, const float* xyz_gradient_accum, const float* denom,
                                   const float* scaling, const float* opacity, float max_grad, float clone_max_scale,
                                   float split_min_scale, float min_opacity, int prune_big, float big_ws_scale,
                                   float split_shrink, int32_t* flags, int32_t* n_clone_selected, void* stream) {
    SC_CHECK_ARG(N >= 0 && 4 * N < (1ll << 31) * 256, "sizes");
    SC_CHECK_ARG(n_clone_selected, "null pointer");
    SC_CUDA(cudaMemsetAsync(n_clone_selected, 0, sizeof(int32_t), (cudaStream_t)stream));
    if (N == 0) return SC_OK;
    SC_CHECK_ARG(xyz_gradient_accum && denom && scaling && opacity && flags, "null pointer");
    SC_CHECK_ARG(((uintptr_t)xyz_gradient_accum & 7) == 0, "8-byte alignment of xyz_gradient_accum");
    SC_CHECK_ARG(split_shrink > 0.f, "split_shrink");
    ClassifyParams P{max_grad, clone_max_scale, split_min_scale, min_opacity, big_ws_scale, split_shrink, prune_big};
    densify_classify_kernel<<<(unsigned)sc_div_up(N, 256), 256, 0, (cudaStream_t)stream>>>(
        N, reinterpret_cast<const float2*>(xyz_gradient_accum), denom, scaling, opacity, P, flags, n_clone_selected);
    SC_CHECK_LAUNCH("densify_classify_kernel");
    return SC_OK;
}

extern "C" int sc_densify_move(int64_t N, const int32_t* flags, const int64_t* dest_index, int64_t n_child,
                               int64_t n_split, int n_copies, const float* noise, const float* scaling,
                               const float* rotation, float split_shrink, int n_arrays, const float* const* src,
                               float* const* dst, const int* row_floats, const int* kinds, void* stream) {
    SC_CHECK_ARG(N >= 0 && n_arrays >= 0 && n_arrays <= SC_COMPACT_MAX_ARRAYS && n_copies >= 1, "sizes");
    SC_CHECK_ARG(n_child >= 0 && n_split >= n_child, "child / split counts");
    if (N == 0 || n_arrays == 0) return SC_OK;
    SC_CHECK_ARG(flags && dest_index && src && dst && row_floats && kinds, "null pointer");
    MoveTable T;
    T.n = n_arrays;
    int64_t blocks = 0;
    for (int a = 0; a < SC_COMPACT_MAX_ARRAYS; ++a) {
        if (a < n_arrays) {
            SC_CHECK_ARG(src[a] && dst[a] && row_floats[a] > 0, "array table entry");
            SC_CHECK_ARG(kinds[a] >= SC_DENSIFY_PARAM && kinds[a] <= SC_DENSIFY_SCALING, "array kind");
            if (kinds[a] == SC_DENSIFY_XYZ || kinds[a] == SC_DENSIFY_SCALING)
                SC_CHECK_ARG(row_floats[a] == 3, "xyz / scaling rows are 3 floats");
            if (kinds[a] == SC_DENSIFY_XYZ && n_split > 0) SC_CHECK_ARG(noise && scaling && rotation, "split inputs");
            T.src[a] = src[a]; T.dst[a] = dst[a]; T.width[a] = row_floats[a]; T.kind[a] = kinds[a];
            blocks += sc_div_up(N * ((row_floats[a] + 3) / 4), 256);
            SC_CHECK_ARG(blocks < (1ll << 31), "N * row width");
        } else {
            T.src[a] = nullptr; T.dst[a] = nullptr; T.width[a] = 1; T.kind[a] = 0;
        }
        T.block_end[a] = (unsigned)blocks;
    }
    densify_move_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(N, flags, dest_index, n_child, n_split,
                                                                          n_copies, noise, scaling, rotation,
                                                                          split_shrink, T);
    SC_CHECK_LAUNCH("densify_move_kernel");
    return SC_OK;
}
