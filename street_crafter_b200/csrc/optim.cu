// Optimiser step and checkpoint row packing on the flat parameter arena.
//
//  * sc_adam_step: torch.optim.Adam(l, lr=0.0, eps=1e-15) as the reference configures it
//    (/root/reference/street_gaussian/models/gaussian_model.py:293-305: one param group per attribute — xyz,
//    f_dc, f_rest, opacity, scaling, rotation, semantic — each with its own learning rate) over ONE flat
//    buffer holding every group, restricted to a [lo, hi) element range: the view-sharded trainer
//    reduce-scatters the gradient arena, every rank steps its 1/world slice and the parameters are
//    all-gathered.  HBM-bound: 16 B read + 12 B written per element.
//  * sc_ply_pack / sc_ply_unpack: the row layout of GaussianModel.make_ply / load_ply
//    (gaussian_model.py:82-99, :107-156, attribute order :328-342): x y z nx ny nz f_dc (3) f_rest
//    (3 (K-1), CHANNEL-major: transpose(1, 2).flatten(1)) opacity scale (3) rot (4) semantic (S).
#include "common.cuh"

namespace {

constexpr int ADAM_MAX_GROUPS = 64;

struct AdamGroups {
    int64_t start[ADAM_MAX_GROUPS + 1];   // element offsets (multiples of 4), ascending; start[n] = end
    float step_size[ADAM_MAX_GROUPS];     // lr / (1 - beta1^t)
    int n;
};

// torch/optim/adam.py _single_tensor_adam (no amsgrad, no weight decay, maximize=False):
//   exp_avg.lerp_(grad, 1 - beta1); exp_avg_sq.mul_(beta2).addcmul_(grad, grad, value=1 - beta2)
//   denom = (exp_avg_sq.sqrt() / sqrt(1 - beta2^t)).add_(eps); param.addcdiv_(exp_avg, denom, value=-lr / (1 - beta1^t))
__global__ void __launch_bounds__(256)
adam_step_kernel(AdamGroups G, int64_t lo4, int64_t hi4, float4* __restrict__ p, const float4* __restrict__ g,
                 float4* __restrict__ m, float4* __restrict__ v, float one_minus_beta1, float beta2,
                 float one_minus_beta2, float bc2_sqrt, float eps) {
    const int64_t i4 = lo4 + (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i4 >= hi4) return;
    const int64_t e = i4 * 4;
    int s = 0;
    while (s < G.n - 1 && e >= G.start[s + 1]) ++s;
    const float step = G.step_size[s];
    const float4 gg = __ldg(g + i4);
    float4 mm = m[i4], vv = v[i4], pp = p[i4];
    const float ga[4] = {gg.x, gg.y, gg.z, gg.w};
    float ma[4] = {mm.x, mm.y, mm.z, mm.w}, va[4] = {vv.x, vv.y, vv.z, vv.w}, pa[4] = {pp.x, pp.y, pp.z, pp.w};
#pragma unroll
    for (int c = 0; c < 4; ++c) {
        ma[c] = ma[c] + one_minus_beta1 * (ga[c] - ma[c]);               // lerp, weight < 0.5
        va[c] = va[c] * beta2 + one_minus_beta2 * (ga[c] * ga[c]);
        const float denom = sqrtf(va[c]) / bc2_sqrt + eps;
        pa[c] = pa[c] - step * (ma[c] / denom);
    }
    m[i4] = make_float4(ma[0], ma[1], ma[2], ma[3]);
    v[i4] = make_float4(va[0], va[1], va[2], va[3]);
    p[i4] = make_float4(pa[0], pa[1], pa[2], pa[3]);
}

struct PlyCols {
    int K, S, width;        // SH coefficients per channel, semantic columns, floats per row
};

__device__ __forceinline__ int64_t ply_src(const PlyCols C, int64_t n, int col, int& which) {
    // which: 0 xyz, 1 zero (normals), 2 f_dc, 3 f_rest, 4 opacity, 5 scaling, 6 rotation, 7 semantic
    const int kr = C.K - 1;
    if (col < 3) { which = 0; return n * 3 + col; }
    if (col < 6) { which = 1; return 0; }
    col -= 6;
    if (col < 3) { which = 2; return n * 3 + col; }                                   // f_dc [N,1,3] -> channel col
    col -= 3;
    if (col < 3 * kr) { which = 3; const int ch = col / kr, k = col - ch * kr; return (n * kr + k) * 3 + ch; }
    col -= 3 * kr;
    if (col < 1) { which = 4; return n; }
    col -= 1;
    if (col < 3) { which = 5; return n * 3 + col; }
    col -= 3;
    if (col < 4) { which = 6; return n * 4 + col; }
    col -= 4;
    which = 7;
    return n * C.S + col;
}

__global__ void __launch_bounds__(256)
ply_pack_kernel(int64_t N, PlyCols C, const float* __restrict__ xyz, const float* __restrict__ f_dc,
                const float* __restrict__ f_rest, const float* __restrict__ opacity, const float* __restrict__ scaling,
                const float* __restrict__ rotation, const float* __restrict__ semantic, float* __restrict__ rows) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= N * C.width) return;
    const int64_t n = i / C.width;
    int which;
    const int64_t j = ply_src(C, n, (int)(i - n * C.width), which);
    const float* src[8] = {xyz, nullptr, f_dc, f_rest, opacity, scaling, rotation, semantic};
    rows[i] = which == 1 ? 0.0f : __ldg(src[which] + j);
}

__global__ void __launch_bounds__(256)
ply_unpack_kernel(int64_t N, PlyCols C, const float* __restrict__ rows, float* __restrict__ xyz, float* __restrict__ f_dc,
                  float* __restrict__ f_rest, float* __restrict__ opacity, float* __restrict__ scaling,
                  float* __restrict__ rotation, float* __restrict__ semantic) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= N * C.width) return;
    const int64_t n = i / C.width;
    int which;
    const int64_t j = ply_src(C, n, (int)(i - n * C.width), which);
    float* dst[8] = {xyz, nullptr, f_dc, f_rest, opacity, scaling, rotation, semantic};
    if (which != 1) dst[which][j] = __ldg(rows + i);
}

int ply_cols(PlyCols& C, int K, int S) {
    SC_CHECK_ARG(K >= 1 && S >= 0, "K, S");
    C.K = K;
    C.S = S;
    C.width = 6 + 3 * K + 1 + 3 + 4 + S;
    return SC_OK;
}

}  // namespace

extern "C" int sc_adam_step(int n_groups, const int64_t* host_group_start /*[n_groups+1]*/, const float* host_lr,
                            int64_t lo, int64_t hi, float* params, const float* grads, float* exp_avg,
                            float* exp_avg_sq, double beta1, double beta2, double eps, int64_t step, void* stream) {
    SC_CHECK_ARG(n_groups >= 1 && n_groups <= ADAM_MAX_GROUPS && host_group_start && host_lr, "groups");
    SC_CHECK_ARG(step >= 1 && lo >= 0 && hi >= lo, "step / range");
    SC_CHECK_ARG((lo & 3) == 0 && (hi & 3) == 0, "range must be a multiple of 4 elements");
    if (hi == lo) return SC_OK;
    SC_CHECK_ARG(params && grads && exp_avg && exp_avg_sq, "null pointer");
    SC_CHECK_ARG((((uintptr_t)params | (uintptr_t)grads | (uintptr_t)exp_avg | (uintptr_t)exp_avg_sq) & 15) == 0,
                 "16-byte alignment");
    AdamGroups G;
    G.n = n_groups;
    // the scalars torch forms in Python doubles before the float32 tensor ops
    const double bc1 = 1.0 - pow(beta1, (double)step);
    const double bc2 = 1.0 - pow(beta2, (double)step);
    for (int s = 0; s <= ADAM_MAX_GROUPS; ++s) G.start[s] = host_group_start[s <= n_groups ? s : n_groups];
    for (int s = 0; s < ADAM_MAX_GROUPS; ++s) G.step_size[s] = s < n_groups ? (float)((double)host_lr[s] / bc1) : 0.f;
    for (int s = 0; s < n_groups; ++s)
        SC_CHECK_ARG((G.start[s] & 3) == 0 && G.start[s + 1] >= G.start[s], "group starts: ascending multiples of 4");
    SC_CHECK_ARG(lo >= G.start[0] && hi <= G.start[n_groups], "range outside the groups");
    const int64_t n4 = (hi - lo) / 4;
    adam_step_kernel<<<(unsigned)sc_div_up(n4, 256), 256, 0, (cudaStream_t)stream>>>(
        G, lo / 4, hi / 4, reinterpret_cast<float4*>(params), reinterpret_cast<const float4*>(grads),
        reinterpret_cast<float4*>(exp_avg), reinterpret_cast<float4*>(exp_avg_sq), (float)(1.0 - beta1), (float)beta2,
        (float)(1.0 - beta2), (float)sqrt(bc2), (float)eps);
    SC_CHECK_LAUNCH("adam_step_kernel");
    return SC_OK;
}

extern "C" int sc_ply_row_width(int K, int S) { return 6 + 3 * K + 1 + 3 + 4 + S; }

extern "C" int sc_ply_pack(int64_t N, int K, int S, const float* xyz, const float* f_dc, const float* f_rest,
                           const float* opacity, const float* scaling, const float* rotation, const float* semantic,
                           float* rows, void* stream) {
    PlyCols C;
    const int rc = ply_cols(C, K, S);
    if (rc != SC_OK) return rc;
    SC_CHECK_ARG(N >= 0, "N");
    if (N == 0) return SC_OK;
    SC_CHECK_ARG(xyz && f_dc && (K == 1 || f_rest) && opacity && scaling && rotation && (S == 0 || semantic) && rows,
                 "null pointer");
    ply_pack_kernel<<<(unsigned)sc_div_up(N * C.width, 256), 256, 0, (cudaStream_t)stream>>>(
        N, C, xyz, f_dc, f_rest, opacity, scaling, rotation, semantic, rows);
    SC_CHECK_LAUNCH("ply_pack_kernel");
    return SC_OK;
}

extern "C" int sc_ply_unpack(int64_t N, int K, int S, const float* rows, float* xyz, float* f_dc, float* f_rest,
                             float* opacity, float* scaling, float* rotation, float* semantic, void* stream) {
    PlyCols C;
    const int rc = ply_cols(C, K, S);
    if (rc != SC_OK) return rc;
    SC_CHECK_ARG(N >= 0, "N");
    if (N == 0) return SC_OK;
    SC_CHECK_ARG(rows && xyz && f_dc && (K == 1 || f_rest) && opacity && scaling && rotation && (S == 0 || semantic),
                 "null pointer");
    ply_unpack_kernel<<<(unsigned)sc_div_up(N * C.width, 256), 256, 0, (cudaStream_t)stream>>>(
        N, C, rows, xyz, f_dc, f_rest, opacity, scaling, rotation, semantic);
    SC_CHECK_LAUNCH("ply_unpack_kernel");
    return SC_OK;
}
