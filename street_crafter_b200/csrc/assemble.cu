// Scene-graph assembly of positions and rotations (geometric half): what the
// reference's `StreetGaussianModel.get_xyz` / `get_rotation` properties
// (/root/reference/street_gaussian/models/street_gaussian_model.py:296-364) build every render call with
// per-model cat / clone / masked-index / quaternion_to_matrix / einsum / quaternion_raw_multiply / normalize
// launches (~40 per pass, plus as many autograd nodes) — as ONE forward and ONE backward kernel over the
// concatenated Gaussian set, which write the projection kernel's `means` / `quats` inputs directly:
//
//   plain segment (background, sky):   xyz_w = xyz                    rot_w = normalize(rot)
//   sky shell (gaussian_model_sky.py:68-76): d = xyz - c, ratio = |d| / (2 R); xyz_w = c + d / ratio where ratio < 1
//   actor segment (rigid pose a, t):   x' = flip_axis-mirrored xyz where flip_mask      (:347-350)
//                                      xyz_w = R(a / |a|) x' + t                         (:351-352)
//                                      q  = normalize(rot); q = f (x) q where flip_mask  (:313-318)
//                                      rot_w = normalize(a (x) q)   (a RAW, as the reference) (:319-320)
//
// normalize = torch.nn.functional.normalize (x / max(|x|, 1e-12)); (x) = quaternion_raw_multiply
// (/root/reference/street_gaussian/utils/general_utils.py:245-264); R() = quaternion_to_matrix (:125-146).
// Backward: per-Gaussian gradients to every segment's own xyz / rotation arrays, and the tracked-pose gradients
// (a, t — optimised by the reference's ActorPose) reduced per warp and accumulated with 7 atomics per warp.
// HBM-bound: 28 B read + 28 B written per Gaussian forward, 56 B + 28 B backward.
#include "common.cuh"

namespace {

struct SegTable {
    int64_t start[SC_ASSEMBLE_MAX_SEGMENTS + 1];
    const float* xyz[SC_ASSEMBLE_MAX_SEGMENTS];
    const float* rot[SC_ASSEMBLE_MAX_SEGMENTS];
    const float* pose_rot[SC_ASSEMBLE_MAX_SEGMENTS];    // null: plain segment
    const float* pose_trans[SC_ASSEMBLE_MAX_SEGMENTS];
    const uint8_t* flip[SC_ASSEMBLE_MAX_SEGMENTS];      // null: no flips
    float shell[SC_ASSEMBLE_MAX_SEGMENTS][4];            // sky shell: centre xyz, 2 R (0 = none)
    int n;
    int flip_axis;
    float fq[4];
};

struct GradTable {
    float* v_xyz[SC_ASSEMBLE_MAX_SEGMENTS];
    float* v_rot[SC_ASSEMBLE_MAX_SEGMENTS];
    float* v_pose_rot[SC_ASSEMBLE_MAX_SEGMENTS];        // null: pose not optimised
    float* v_pose_trans[SC_ASSEMBLE_MAX_SEGMENTS];
};

struct Quat { float w, x, y, z; };

__device__ __forceinline__ Quat qmul(const Quat a, const Quat b) {          // general_utils.py:257-264
    return {a.w * b.w - a.x * b.x - a.y * b.y - a.z * b.z, a.w * b.x + a.x * b.w + a.y * b.z - a.z * b.y,
            a.w * b.y - a.x * b.z + a.y * b.w + a.z * b.x, a.w * b.z + a.x * b.y - a.y * b.x + a.z * b.w};
}
// gradient of (a (x) b) w.r.t. b and a, given g = d/d(a (x) b)
__device__ __forceinline__ Quat qmul_vjp_b(const Quat a, const Quat g) {
    return {a.w * g.w + a.x * g.x + a.y * g.y + a.z * g.z, -a.x * g.w + a.w * g.x + a.z * g.y - a.y * g.z,
            -a.y * g.w - a.z * g.x + a.w * g.y + a.x * g.z, -a.z * g.w + a.y * g.x - a.x * g.y + a.w * g.z};
}
__device__ __forceinline__ Quat qmul_vjp_a(const Quat b, const Quat g) {
    return {b.w * g.w + b.x * g.x + b.y * g.y + b.z * g.z, -b.x * g.w + b.w * g.x - b.z * g.y + b.y * g.z,
            -b.y * g.w + b.z * g.x + b.w * g.y - b.x * g.z, -b.z * g.w - b.y * g.x + b.x * g.y + b.w * g.z};
}
__device__ __forceinline__ float qnorm_f(const Quat q) {                    // F.normalize denominator
    return fmaxf(sqrtf(q.w * q.w + q.x * q.x + q.y * q.y + q.z * q.z), 1e-12f);
}
__device__ __forceinline__ Quat qscale(const Quat q, float s) { return {q.w * s, q.x * s, q.y * s, q.z * s}; }
// y = x / max(|x|, eps) (n = that denominator): VJP given y and g = dL/dy
__device__ __forceinline__ Quat normalize_vjp(const Quat y, const Quat g, float n) {
    const float d = y.w * g.w + y.x * g.x + y.y * g.y + y.z * g.z;
    return {(g.w - y.w * d) / n, (g.x - y.x * d) / n, (g.y - y.y * d) / n, (g.z - y.z * d) / n};
}
__device__ __forceinline__ void quat_to_R(const Quat q, float R[9]) {       // general_utils.py:137-145, q unit
    const float r = q.w, x = q.x, y = q.y, z = q.z;
    R[0] = 1.f - 2.f * (y * y + z * z); R[1] = 2.f * (x * y - r * z); R[2] = 2.f * (x * z + r * y);
    R[3] = 2.f * (x * y + r * z); R[4] = 1.f - 2.f * (x * x + z * z); R[5] = 2.f * (y * z - r * x);
    R[6] = 2.f * (x * z - r * y); R[7] = 2.f * (y * z + r * x); R[8] = 1.f - 2.f * (x * x + y * y);
}
__device__ __forceinline__ Quat ldq(const float* p) { return {__ldg(p), __ldg(p + 1), __ldg(p + 2), __ldg(p + 3)}; }

__device__ __forceinline__ int find_segment(const SegTable& T, int64_t i) {
    int s = 0;
    while (s < T.n - 1 && i >= T.start[s + 1]) ++s;
    return s;
}

__global__ void __launch_bounds__(256)
assemble_fwd_kernel(SegTable T, float* __restrict__ out_xyz, float* __restrict__ out_rot) {
    const int64_t i = T.start[0] + (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= T.start[T.n]) return;
    const int s = find_segment(T, i);
    const int64_t k = i - T.start[s];
    float x[3] = {__ldg(T.xyz[s] + 3 * k), __ldg(T.xyz[s] + 3 * k + 1), __ldg(T.xyz[s] + 3 * k + 2)};
    const Quat raw = ldq(T.rot[s] + 4 * k);
    Quat q = qscale(raw, 1.f / qnorm_f(raw));
    if (T.pose_rot[s] == nullptr) {
        if (T.shell[s][3] > 0.f) {
            const float d[3] = {x[0] - T.shell[s][0], x[1] - T.shell[s][1], x[2] - T.shell[s][2]};
            const float ratio = sqrtf(d[0] * d[0] + d[1] * d[1] + d[2] * d[2]) / T.shell[s][3];
            if (ratio < 1.f) {
#pragma unroll
                for (int c = 0; c < 3; ++c) x[c] = T.shell[s][c] + d[c] / ratio;
            }
        }
        out_xyz[3 * i] = x[0]; out_xyz[3 * i + 1] = x[1]; out_xyz[3 * i + 2] = x[2];
        reinterpret_cast<float4*>(out_rot)[i] = make_float4(q.w, q.x, q.y, q.z);
        return;
    }
    const bool flip = T.flip[s] != nullptr && T.flip[s][k];
    if (flip) {
#pragma unroll
        for (int c = 0; c < 3; ++c) x[c] = c == T.flip_axis ? -x[c] : x[c];
        q = qmul({T.fq[0], T.fq[1], T.fq[2], T.fq[3]}, q);
    }
    const Quat a = ldq(T.pose_rot[s]);
    const float na = sqrtf(a.w * a.w + a.x * a.x + a.y * a.y + a.z * a.z);       // general_utils.py:126-128
    float R[9];
    quat_to_R({a.w / na, a.x / na, a.y / na, a.z / na}, R);
    const float* t = T.pose_trans[s];
#pragma unroll
    for (int r = 0; r < 3; ++r)
        out_xyz[3 * i + r] = (R[3 * r] * x[0] + R[3 * r + 1] * x[1] + R[3 * r + 2] * x[2]) + __ldg(t + r);
    const Quat m = qmul(a, q);
    const Quat o = qscale(m, 1.f / qnorm_f(m));
    reinterpret_cast<float4*>(out_rot)[i] = make_float4(o.w, o.x, o.y, o.z);
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

__global__ void __launch_bounds__(256)
assemble_bwd_kernel(SegTable T, GradTable G, const float* __restrict__ v_out_xyz, const float* __restrict__ v_out_rot) {
    const int64_t i = T.start[0] + (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const bool live = i < T.start[T.n];
    // a warp's 32 points span at most a few segments: pose gradients are reduced per (warp, segment of lane 0 ..)
    int s = -1;
    float pa[4] = {0.f, 0.f, 0.f, 0.f}, pt[3] = {0.f, 0.f, 0.f};
    if (live) {
        s = find_segment(T, i);
        const int64_t k = i - T.start[s];
        const float gx[3] = {__ldg(v_out_xyz + 3 * i), __ldg(v_out_xyz + 3 * i + 1), __ldg(v_out_xyz + 3 * i + 2)};
        const float4 go4 = __ldg(reinterpret_cast<const float4*>(v_out_rot) + i);
        const Quat go = {go4.x, go4.y, go4.z, go4.w};
        const Quat raw = ldq(T.rot[s] + 4 * k);
        const float nq = qnorm_f(raw);
        const Quat q0 = qscale(raw, 1.f / nq);
        if (T.pose_rot[s] == nullptr) {
            float gp[3] = {gx[0], gx[1], gx[2]};
            if (T.shell[s][3] > 0.f) {
                // y = c + d / ratio = c + 2R d / |d|  ->  v_d = (g - dhat (dhat . g)) / ratio
                const float d[3] = {__ldg(T.xyz[s] + 3 * k) - T.shell[s][0], __ldg(T.xyz[s] + 3 * k + 1) - T.shell[s][1],
                                    __ldg(T.xyz[s] + 3 * k + 2) - T.shell[s][2]};
                const float dist = sqrtf(d[0] * d[0] + d[1] * d[1] + d[2] * d[2]);
                const float ratio = dist / T.shell[s][3];
                if (ratio < 1.f) {
                    const float dg = (d[0] * gx[0] + d[1] * gx[1] + d[2] * gx[2]) / (dist * dist);
#pragma unroll
                    for (int c = 0; c < 3; ++c) gp[c] = (gx[c] - d[c] * dg) / ratio;
                }
            }
            if (G.v_xyz[s]) { G.v_xyz[s][3 * k] = gp[0]; G.v_xyz[s][3 * k + 1] = gp[1]; G.v_xyz[s][3 * k + 2] = gp[2]; }
            if (G.v_rot[s]) {
                const Quat d = normalize_vjp(q0, go, nq);
                reinterpret_cast<float4*>(G.v_rot[s])[k] = make_float4(d.w, d.x, d.y, d.z);
            }
            s = -1;   // no pose gradient
        } else {
            const bool flip = T.flip[s] != nullptr && T.flip[s][k];
            const Quat f = {T.fq[0], T.fq[1], T.fq[2], T.fq[3]};
            float x[3] = {__ldg(T.xyz[s] + 3 * k), __ldg(T.xyz[s] + 3 * k + 1), __ldg(T.xyz[s] + 3 * k + 2)};
#pragma unroll
            for (int c = 0; c < 3; ++c) x[c] = (flip && c == T.flip_axis) ? -x[c] : x[c];
            const Quat q = flip ? qmul(f, q0) : q0;
            const Quat a = ldq(T.pose_rot[s]);
            const float na = sqrtf(a.w * a.w + a.x * a.x + a.y * a.y + a.z * a.z);
            const Quat u = {a.w / na, a.x / na, a.y / na, a.z / na};
            float R[9];
            quat_to_R(u, R);
            // positions: xyz_w = R x' + t
            float dx[3];
#pragma unroll
            for (int c = 0; c < 3; ++c) dx[c] = R[c] * gx[0] + R[3 + c] * gx[1] + R[6 + c] * gx[2];    // R^T g
#pragma unroll
            for (int c = 0; c < 3; ++c) dx[c] = (flip && c == T.flip_axis) ? -dx[c] : dx[c];
            if (G.v_xyz[s]) { G.v_xyz[s][3 * k] = dx[0]; G.v_xyz[s][3 * k + 1] = dx[1]; G.v_xyz[s][3 * k + 2] = dx[2]; }
            // rotations: o = normalize(m), m = a (x) q
            const Quat m = qmul(a, q);
            const float nm = qnorm_f(m);
            const Quat gm = normalize_vjp(qscale(m, 1.f / nm), go, nm);
            if (G.v_rot[s]) {
                Quat dq = qmul_vjp_b(a, gm);
                if (flip) dq = qmul_vjp_b(f, dq);
                const Quat d = normalize_vjp(q0, dq, nq);
                reinterpret_cast<float4*>(G.v_rot[s])[k] = make_float4(d.w, d.x, d.y, d.z);
            }
            if (G.v_pose_rot[s]) {
                // dR = g x'^T, chained through quaternion_to_matrix(u) and u = a / |a|; plus the raw a of (x)
                float D[9];
#pragma unroll
                for (int r = 0; r < 3; ++r)
#pragma unroll
                    for (int c = 0; c < 3; ++c) D[3 * r + c] = gx[r] * x[c];
                const float w = u.w, X = u.x, Y = u.y, Z = u.z;
                Quat du;
                du.w = 2.f * (-Z * D[1] + Y * D[2] + Z * D[3] - X * D[5] - Y * D[6] + X * D[7]);
                du.x = 2.f * (Y * D[1] + Z * D[2] + Y * D[3] - 2.f * X * D[4] - w * D[5] + Z * D[6] + w * D[7] - 2.f * X * D[8]);
                du.y = 2.f * (-2.f * Y * D[0] + X * D[1] + w * D[2] + X * D[3] + Z * D[5] - w * D[6] + Z * D[7] - 2.f * Y * D[8]);
                du.z = 2.f * (-2.f * Z * D[0] - w * D[1] + X * D[2] + w * D[3] - 2.f * Z * D[4] + Y * D[5] + X * D[6] + Y * D[7]);
                const Quat da1 = normalize_vjp(u, du, na);
                const Quat da2 = qmul_vjp_a(q, gm);
                pa[0] = da1.w + da2.w; pa[1] = da1.x + da2.x; pa[2] = da1.y + da2.y; pa[3] = da1.z + da2.z;
            }
            pt[0] = gx[0]; pt[1] = gx[1]; pt[2] = gx[2];
        }
    }
    // per-warp reduction of the pose gradients, one round per distinct actor segment present in the warp
    unsigned todo = __ballot_sync(0xffffffffu, s >= 0);
    while (todo) {
        const int leader = __ffs(todo) - 1;
        const int cur = __shfl_sync(0xffffffffu, s, leader);
        const bool mine = s == cur;
        float r[7];
#pragma unroll
        for (int c = 0; c < 4; ++c) r[c] = warp_sum(mine ? pa[c] : 0.f);
#pragma unroll
        for (int c = 0; c < 3; ++c) r[4 + c] = warp_sum(mine ? pt[c] : 0.f);
        if ((threadIdx.x & 31) == leader) {
            if (G.v_pose_rot[cur])
#pragma unroll
                for (int c = 0; c < 4; ++c) atomicAdd(G.v_pose_rot[cur] + c, r[c]);
            if (G.v_pose_trans[cur])
#pragma unroll
                for (int c = 0; c < 3; ++c) atomicAdd(G.v_pose_trans[cur] + c, r[4 + c]);
        }
        todo &= ~__ballot_sync(0xffffffffu, mine);
    }
}

int fill_table(SegTable& T, int n_segments, const int64_t* seg_start, const float* const* xyz, const float* const* rot,
               const float* const* pose_rot, const float* const* pose_trans, const uint8_t* const* flip, int flip_axis,
               const float* flip_quat, const float* shell) {
    SC_CHECK_ARG(n_segments >= 1 && n_segments <= SC_ASSEMBLE_MAX_SEGMENTS, "segment count");
    SC_CHECK_ARG(seg_start && xyz && rot && pose_rot && pose_trans && flip, "null table");
    SC_CHECK_ARG(flip_axis >= 0 && flip_axis < 3 && flip_quat, "flip axis / quaternion");
    T.n = n_segments;
    T.flip_axis = flip_axis;
    for (int c = 0; c < 4; ++c) T.fq[c] = flip_quat[c];
    for (int s = 0; s < SC_ASSEMBLE_MAX_SEGMENTS; ++s) {
        const bool in = s < n_segments;
        if (in) {
            SC_CHECK_ARG(seg_start[s + 1] >= seg_start[s], "segment starts must not decrease");
            SC_CHECK_ARG(seg_start[s + 1] == seg_start[s] || (xyz[s] && rot[s]), "segment arrays");
            SC_CHECK_ARG((pose_rot[s] == nullptr) == (pose_trans[s] == nullptr), "pose rotation and translation go together");
            SC_CHECK_ARG(((uintptr_t)rot[s] & 15) == 0, "16-byte alignment of rotations");
        }
        T.xyz[s] = in ? xyz[s] : nullptr; T.rot[s] = in ? rot[s] : nullptr;
        T.pose_rot[s] = in ? pose_rot[s] : nullptr; T.pose_trans[s] = in ? pose_trans[s] : nullptr;
        T.flip[s] = in ? flip[s] : nullptr;
        for (int c = 0; c < 4; ++c) T.shell[s][c] = (in && shell) ? shell[4 * s + c] : 0.f;
        SC_CHECK_ARG(!(in && shell && shell[4 * s + 3] > 0.f && pose_rot[s]), "the sky shell applies to a plain segment");
    }
    for (int s = 0; s <= SC_ASSEMBLE_MAX_SEGMENTS; ++s) T.start[s] = seg_start[s <= n_segments ? s : n_segments];
    return SC_OK;
}

// ---- activations of the concatenated set: get_scaling (:273-294) and get_opacity (:389-407) -------------------
// scaling_w = min(exp(scaling), clamp_max)   (clamp_max = +inf except for the sky model, gaussian_model_sky.py:63-66)
// opacity_w = sigmoid(opacity)               (gaussian_model.py:215-220, :226-227, :251-252)
struct ActTable {
    int64_t start[SC_ASSEMBLE_MAX_SEGMENTS + 1];
    const float* scaling[SC_ASSEMBLE_MAX_SEGMENTS];
    const float* opacity[SC_ASSEMBLE_MAX_SEGMENTS];
    float* v_scaling[SC_ASSEMBLE_MAX_SEGMENTS];     // backward only
    float* v_opacity[SC_ASSEMBLE_MAX_SEGMENTS];
    float clamp_max[SC_ASSEMBLE_MAX_SEGMENTS];
    int n;
};

__device__ __forceinline__ int find_act_segment(const ActTable& T, int64_t i) {
    int s = 0;
    while (s < T.n - 1 && i >= T.start[s + 1]) ++s;
    return s;
}

__global__ void __launch_bounds__(256)
assemble_act_fwd_kernel(ActTable T, float* __restrict__ out_scaling, float* __restrict__ out_opacity) {
    const int64_t i = T.start[0] + (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= T.start[T.n]) return;
    const int s = find_act_segment(T, i);
    const int64_t k = i - T.start[s];
#pragma unroll
    for (int c = 0; c < 3; ++c) out_scaling[3 * i + c] = fminf(expf(__ldg(T.scaling[s] + 3 * k + c)), T.clamp_max[s]);
    out_opacity[i] = 1.f / (1.f + expf(-__ldg(T.opacity[s] + k)));
}

__global__ void __launch_bounds__(256)
assemble_act_bwd_kernel(ActTable T, const float* __restrict__ v_out_scaling, const float* __restrict__ v_out_opacity) {
    const int64_t i = T.start[0] + (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= T.start[T.n]) return;
    const int s = find_act_segment(T, i);
    const int64_t k = i - T.start[s];
    if (T.v_scaling[s]) {
#pragma unroll
        for (int c = 0; c < 3; ++c) {
            const float e = expf(__ldg(T.scaling[s] + 3 * k + c));
            T.v_scaling[s][3 * k + c] = e <= T.clamp_max[s] ? __ldg(v_out_scaling + 3 * i + c) * e : 0.f;   // clamp passes x <= max
        }
    }
    if (T.v_opacity[s]) {
        const float o = 1.f / (1.f + expf(-__ldg(T.opacity[s] + k)));
        T.v_opacity[s][k] = __ldg(v_out_opacity + i) * ((1.f - o) * o);
    }
}

int fill_act_table(ActTable& T, int n_segments, const int64_t* seg_start, const float* const* scaling,
                   const float* const* opacity, const float* clamp_max, float* const* v_scaling,
                   float* const* v_opacity) {
    SC_CHECK_ARG(n_segments >= 1 && n_segments <= SC_ASSEMBLE_MAX_SEGMENTS, "segment count");
    SC_CHECK_ARG(seg_start && scaling && opacity && clamp_max, "null table");
    T.n = n_segments;
    for (int s = 0; s < SC_ASSEMBLE_MAX_SEGMENTS; ++s) {
        const bool in = s < n_segments;
        if (in) {
            SC_CHECK_ARG(seg_start[s + 1] >= seg_start[s], "segment starts must not decrease");
            SC_CHECK_ARG(seg_start[s + 1] == seg_start[s] || (scaling[s] && opacity[s]), "segment arrays");
        }
        T.scaling[s] = in ? scaling[s] : nullptr; T.opacity[s] = in ? opacity[s] : nullptr;
        T.clamp_max[s] = in ? clamp_max[s] : 0.f;
        T.v_scaling[s] = (in && v_scaling) ? v_scaling[s] : nullptr;
        T.v_opacity[s] = (in && v_opacity) ? v_opacity[s] : nullptr;
    }
    for (int s = 0; s <= SC_ASSEMBLE_MAX_SEGMENTS; ++s) T.start[s] = seg_start[s <= n_segments ? s : n_segments];
    return SC_OK;
}

// ---- features of the concatenated set: get_features (:366-386) ------------------------------------------------
// plain segment: cat(f_dc [n,1,3], f_rest [n,K-1,3]) (gaussian_model.py:238-241); actor segment: the DC colour is the
// inverse Fourier sum over the actor's C coefficients, sum_c f_dc[:,c,:] * idft_base[c] (get_features_fourier,
// gaussian_model_actor.py:67-76; idft_base = IDFT(time, C), sh_utils.py:120-130, evaluated by the host).
// One float of the [N, K, 3] output per lane access, consecutive lanes on consecutive floats.
struct FeatTable {
    int64_t start[SC_ASSEMBLE_MAX_SEGMENTS + 1];
    const float* f_dc[SC_ASSEMBLE_MAX_SEGMENTS];
    const float* f_rest[SC_ASSEMBLE_MAX_SEGMENTS];
    float* v_f_dc[SC_ASSEMBLE_MAX_SEGMENTS];        // backward only
    float* v_f_rest[SC_ASSEMBLE_MAX_SEGMENTS];
    int n_fourier[SC_ASSEMBLE_MAX_SEGMENTS];
    float base[SC_ASSEMBLE_MAX_SEGMENTS][SC_ASSEMBLE_MAX_FOURIER];
    int n;
    int row;                                        // 3 K floats per output row
};
constexpr int FEAT_EPT = 4;

// ROW = 3 K as a compile-time constant for the reference's SH degrees 3 / 1 (row / column split by multiply-shift,
// one 64-bit division per thread instead of one per float), 0 = any K.
template <bool BWD, int ROW>
__global__ void __launch_bounds__(256)
assemble_feat_kernel(FeatTable T, float* __restrict__ out, const float* __restrict__ v_out) {
    const int row = ROW ? ROW : T.row;
    const int64_t e0 = T.start[0] * row, e1 = T.start[T.n] * row;
    const int64_t base = e0 + (int64_t)blockIdx.x * (256 * FEAT_EPT);
    const int64_t i0 = base / row;
    const int rem0 = (int)(base - i0 * row);
    int s = 0;
#pragma unroll
    for (int j = 0; j < FEAT_EPT; ++j) {
        const int64_t e = base + j * 256 + threadIdx.x;
        if (e >= e1) return;
        const unsigned L = (unsigned)(rem0 + j * 256 + (int)threadIdx.x);
        const unsigned di = L / (unsigned)row;
        const int c = (int)(L - di * (unsigned)row);
        const int64_t i = i0 + di;
        while (s < T.n - 1 && i >= T.start[s + 1]) ++s;      // i only grows with j
        const int64_t k = i - T.start[s];
        const int C = T.n_fourier[s];
        if (!BWD) {
            float v;
            if (c < 3) {
                v = 0.f;
                for (int f = 0; f < C; ++f) v += __ldg(T.f_dc[s] + (k * C + f) * 3 + c) * T.base[s][f];
            } else {
                v = __ldg(T.f_rest[s] + k * (row - 3) + (c - 3));
            }
            out[e] = v;
        } else {
            const float g = __ldg(v_out + e);
            if (c < 3) {
                if (T.v_f_dc[s])
                    for (int f = 0; f < C; ++f) T.v_f_dc[s][(k * C + f) * 3 + c] = g * T.base[s][f];
            } else if (T.v_f_rest[s]) {
                T.v_f_rest[s][k * (row - 3) + (c - 3)] = g;
            }
        }
    }
}

// ROW a multiple of 4 floats (K = 16, 4): a thread owns one 16-byte piece of an output row — piece 0 is
// [dc.r, dc.g, dc.b, rest[0]], piece p >= 1 is rest[4p-3 .. 4p] — float4 on the [N, K, 3] side, scalars on the
// model side (rows of 45 / 9 floats are not 16-byte aligned).
template <bool BWD, int ROW>
__global__ void __launch_bounds__(256)
assemble_feat_vec_kernel(FeatTable T, float* __restrict__ out, const float* __restrict__ v_out) {
    constexpr int PIECES = ROW / 4;
    const int64_t p0 = T.start[0] * PIECES, p1 = T.start[T.n] * PIECES;
    int s = 0;
#pragma unroll
    for (int j = 0; j < FEAT_EPT; ++j) {
        const int64_t P = p0 + ((int64_t)blockIdx.x * FEAT_EPT + j) * 256 + threadIdx.x;
        if (P >= p1) return;
        const int64_t i = P / PIECES;
        const int p = (int)(P - i * PIECES);
        while (s < T.n - 1 && i >= T.start[s + 1]) ++s;
        const int64_t k = i - T.start[s];
        const int C = T.n_fourier[s];
        if (!BWD) {
            const float* r = T.f_rest[s] + k * (ROW - 3) + 4 * p - 3;
            float4 v;
            if (p == 0) {
                float d[3] = {0.f, 0.f, 0.f};
                for (int f = 0; f < C; ++f) {
                    const float b = T.base[s][f];
#pragma unroll
                    for (int c = 0; c < 3; ++c) d[c] += __ldg(T.f_dc[s] + (k * C + f) * 3 + c) * b;
                }
                v = make_float4(d[0], d[1], d[2], __ldg(r + 3));
            } else {
                v = make_float4(__ldg(r), __ldg(r + 1), __ldg(r + 2), __ldg(r + 3));
            }
            reinterpret_cast<float4*>(out)[P] = v;
        } else {
            const float4 g = __ldg(reinterpret_cast<const float4*>(v_out) + P);
            float* r = T.v_f_rest[s] ? T.v_f_rest[s] + k * (ROW - 3) + 4 * p - 3 : nullptr;
            if (p == 0) {
                if (T.v_f_dc[s])
                    for (int f = 0; f < C; ++f) {
                        const float b = T.base[s][f];
                        float* d = T.v_f_dc[s] + (k * C + f) * 3;
                        d[0] = g.x * b; d[1] = g.y * b; d[2] = g.z * b;
                    }
                if (r) r[3] = g.w;
            } else if (r) {
                r[0] = g.x; r[1] = g.y; r[2] = g.z; r[3] = g.w;
            }
        }
    }
}

template <bool BWD>
void launch_feat(const FeatTable& T, int64_t n_floats, float* out, const float* v_out, cudaStream_t st) {
    const bool aligned = ((((uintptr_t)out) | ((uintptr_t)v_out)) & 15) == 0;
    const unsigned vb = (unsigned)sc_div_up(n_floats / 4, 256 * FEAT_EPT), eb = (unsigned)sc_div_up(n_floats, 256 * FEAT_EPT);
    if (T.row == 48 && aligned) assemble_feat_vec_kernel<BWD, 48><<<vb, 256, 0, st>>>(T, out, v_out);
    else if (T.row == 12 && aligned) assemble_feat_vec_kernel<BWD, 12><<<vb, 256, 0, st>>>(T, out, v_out);
    else assemble_feat_kernel<BWD, 0><<<eb, 256, 0, st>>>(T, out, v_out);
}

int fill_feat_table(FeatTable& T, int n_segments, const int64_t* seg_start, int K, const float* const* f_dc,
                    const float* const* f_rest, const int* n_fourier, const float* idft_base, float* const* v_f_dc,
                    float* const* v_f_rest) {
    const bool fwd = v_f_dc == nullptr;     // the backward never reads the parameters
    SC_CHECK_ARG(n_segments >= 1 && n_segments <= SC_ASSEMBLE_MAX_SEGMENTS && K >= 1 && K <= 64, "segment count / K");
    SC_CHECK_ARG(seg_start && n_fourier && idft_base && (!fwd || (f_dc && f_rest)), "null table");
    T.n = n_segments;
    T.row = 3 * K;
    for (int s = 0; s < SC_ASSEMBLE_MAX_SEGMENTS; ++s) {
        const bool in = s < n_segments;
        if (in) {
            SC_CHECK_ARG(seg_start[s + 1] >= seg_start[s], "segment starts must not decrease");
            SC_CHECK_ARG(n_fourier[s] >= 1 && n_fourier[s] <= SC_ASSEMBLE_MAX_FOURIER, "fourier dimension");
            if (fwd) SC_CHECK_ARG(seg_start[s + 1] == seg_start[s] || (f_dc[s] && (K == 1 || f_rest[s])), "segment arrays");
        }
        T.f_dc[s] = (in && fwd) ? f_dc[s] : nullptr; T.f_rest[s] = (in && fwd) ? f_rest[s] : nullptr;
        T.v_f_dc[s] = (in && v_f_dc) ? v_f_dc[s] : nullptr; T.v_f_rest[s] = (in && v_f_rest) ? v_f_rest[s] : nullptr;
        T.n_fourier[s] = in ? n_fourier[s] : 1;
        for (int f = 0; f < SC_ASSEMBLE_MAX_FOURIER; ++f)
            T.base[s][f] = (in && f < n_fourier[s]) ? idft_base[s * SC_ASSEMBLE_MAX_FOURIER + f] : 0.f;
    }
    for (int s = 0; s <= SC_ASSEMBLE_MAX_SEGMENTS; ++s) T.start[s] = seg_start[s <= n_segments ? s : n_segments];
    return SC_OK;
}

}  // namespace

extern "C" int sc_assemble_fwd(int n_segments, const int64_t* seg_start, const float* const* xyz,
                               const float* const* rot, const float* const* pose_rot, const float* const* pose_trans,
                               const uint8_t* const* flip, int flip_axis, const float* flip_quat,
                               const float* host_shell, float* out_xyz, float* out_rot, void* stream) {
    SegTable T;
    const int rc = fill_table(T, n_segments, seg_start, xyz, rot, pose_rot, pose_trans, flip, flip_axis, flip_quat,
                              host_shell);
    if (rc != SC_OK) return rc;
    const int64_t n = T.start[T.n] - T.start[0];
    if (n == 0) return SC_OK;
    SC_CHECK_ARG(out_xyz && out_rot && ((uintptr_t)out_rot & 15) == 0, "outputs");
    assemble_fwd_kernel<<<(unsigned)sc_div_up(n, 256), 256, 0, (cudaStream_t)stream>>>(T, out_xyz, out_rot);
    SC_CHECK_LAUNCH("assemble_fwd_kernel");
    return SC_OK;
}

extern "C" int sc_assemble_bwd(int n_segments, const int64_t* seg_start, const float* const* xyz,
                               const float* const* rot, const float* const* pose_rot, const float* const* pose_trans,
                               const uint8_t* const* flip, int flip_axis, const float* flip_quat, const float* host_shell,
                               const float* v_out_xyz, const float* v_out_rot, float* const* v_xyz, float* const* v_rot,
                               float* const* v_pose_rot, float* const* v_pose_trans, void* stream) {
    SegTable T;
    const int rc = fill_table(T, n_segments, seg_start, xyz, rot, pose_rot, pose_trans, flip, flip_axis, flip_quat, host_shell);
    if (rc != SC_OK) return rc;
    const int64_t n = T.start[T.n] - T.start[0];
    if (n == 0) return SC_OK;
    SC_CHECK_ARG(v_out_xyz && v_out_rot && ((uintptr_t)v_out_rot & 15) == 0, "upstream gradients");
    SC_CHECK_ARG(v_xyz && v_rot && v_pose_rot && v_pose_trans, "null gradient table");
    GradTable G;
    for (int s = 0; s < SC_ASSEMBLE_MAX_SEGMENTS; ++s) {
        const bool in = s < n_segments;
        G.v_xyz[s] = in ? v_xyz[s] : nullptr; G.v_rot[s] = in ? v_rot[s] : nullptr;
        G.v_pose_rot[s] = in ? v_pose_rot[s] : nullptr; G.v_pose_trans[s] = in ? v_pose_trans[s] : nullptr;
        if (in) SC_CHECK_ARG(((uintptr_t)G.v_rot[s] & 15) == 0, "16-byte alignment of rotation gradients");
    }
    assemble_bwd_kernel<<<(unsigned)sc_div_up(n, 256), 256, 0, (cudaStream_t)stream>>>(T, G, v_out_xyz, v_out_rot);
    SC_CHECK_LAUNCH("assemble_bwd_kernel");
    return SC_OK;
}

extern "C" int sc_assemble_act_fwd(int n_segments, const int64_t* seg_start, const float* const* scaling,
                                   const float* const* opacity, const float* clamp_max, float* out_scaling,
                                   float* out_opacity, void* stream) {
    ActTable T;
    const int rc = fill_act_table(T, n_segments, seg_start, scaling, opacity, clamp_max, nullptr, nullptr);
    if (rc != SC_OK) return rc;
    const int64_t n = T.start[T.n] - T.start[0];
    if (n == 0) return SC_OK;
    SC_CHECK_ARG(out_scaling && out_opacity, "outputs");
    assemble_act_fwd_kernel<<<(unsigned)sc_div_up(n, 256), 256, 0, (cudaStream_t)stream>>>(T, out_scaling, out_opacity);
    SC_CHECK_LAUNCH("assemble_act_fwd_kernel");
    return SC_OK;
}

extern "C" int sc_assemble_act_bwd(int n_segments, const int64_t* seg_start, const float* const* scaling,
                                   const float* const* opacity, const float* clamp_max, const float* v_out_scaling,
                                   const float* v_out_opacity, float* const* v_scaling, float* const* v_opacity,
                                   void* stream) {
    SC_CHECK_ARG(v_scaling && v_opacity, "null gradient table");
    ActTable T;
    const int rc = fill_act_table(T, n_segments, seg_start, scaling, opacity, clamp_max, v_scaling, v_opacity);
    if (rc != SC_OK) return rc;
    const int64_t n = T.start[T.n] - T.start[0];
    if (n == 0) return SC_OK;
    SC_CHECK_ARG(v_out_scaling && v_out_opacity, "upstream gradients");
    assemble_act_bwd_kernel<<<(unsigned)sc_div_up(n, 256), 256, 0, (cudaStream_t)stream>>>(T, v_out_scaling, v_out_opacity);
    SC_CHECK_LAUNCH("assemble_act_bwd_kernel");
    return SC_OK;
}

extern "C" int sc_assemble_features_fwd(int n_segments, const int64_t* seg_start, int K, const float* const* f_dc,
                                        const float* const* f_rest, const int* n_fourier, const float* idft_base,
                                        float* out_features, void* stream) {
    FeatTable T;
    const int rc = fill_feat_table(T, n_segments, seg_start, K, f_dc, f_rest, n_fourier, idft_base, nullptr, nullptr);
    if (rc != SC_OK) return rc;
    const int64_t n = (T.start[T.n] - T.start[0]) * T.row;
    if (n == 0) return SC_OK;
    SC_CHECK_ARG(out_features, "output");
    const int64_t blocks = sc_div_up(n, 256 * FEAT_EPT);
    SC_CHECK_ARG(blocks < (1ll << 31), "N * K");
    launch_feat<false>(T, n, out_features, nullptr, (cudaStream_t)stream);
    SC_CHECK_LAUNCH("assemble_feat_kernel<fwd>");
    return SC_OK;
}

extern "C" int sc_assemble_features_bwd(int n_segments, const int64_t* seg_start, int K, const int* n_fourier,
                                        const float* idft_base, const float* v_out_features, float* const* v_f_dc,
                                        float* const* v_f_rest, void* stream) {
    SC_CHECK_ARG(v_f_dc && v_f_rest, "null gradient table");
    FeatTable T;
    const int rc = fill_feat_table(T, n_segments, seg_start, K, nullptr, nullptr, n_fourier, idft_base, v_f_dc, v_f_rest);
    if (rc != SC_OK) return rc;
    const int64_t n = (T.start[T.n] - T.start[0]) * T.row;
    if (n == 0) return SC_OK;
    SC_CHECK_ARG(v_out_features, "upstream gradient");
    const int64_t blocks = sc_div_up(n, 256 * FEAT_EPT);
    SC_CHECK_ARG(blocks < (1ll << 31), "N * K");
    launch_feat<true>(T, n, nullptr, v_out_features, (cudaStream_t)stream);
    SC_CHECK_LAUNCH("assemble_feat_kernel<bwd>");
    return SC_OK;
}
