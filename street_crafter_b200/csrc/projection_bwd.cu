// Backward of the fused projection (A8b): v_means2d, v_depths, v_conics, v_compensations
// -> v_means, v_quats, v_scales.  One thread per Gaussian, looping over cameras so the sum
// over cameras needs no atomics.  HBM-bound streaming (≈128 B/Gaussian); recomputes the
// forward intermediates instead of saving them.
//
// Autograd counterpart of gsplat.fully_fused_projection as used at
// /root/reference/street_gaussian/models/street_gaussian_renderer.py:219-232 (viewmat
// gradients are not produced: cameras are not optimised, street_gaussian/config/config.py:126).
#include "common.cuh"

namespace {

constexpr int PB_THREADS = 128;

#ifndef SC_PB_SPAN
#define SC_PB_SPAN 4  // Gaussians per thread in the visibility scan (a CTA owns PB_THREADS * SPAN of them)
#endif
constexpr int PB_SPAN = SC_PB_SPAN;
constexpr int PB_CHUNK = PB_THREADS * PB_SPAN;

// one Gaussian some camera sees
__device__ __forceinline__ void
pb_gaussian(const int64_t g, int64_t N, int C, const float* __restrict__ means, const float* __restrict__ quats,
            const float* __restrict__ scales, const float* __restrict__ viewmats,
            const float* __restrict__ Ks, int width, int height, float eps2d,
            const int32_t* __restrict__ radii, const float* __restrict__ v_means2d,
            const float* __restrict__ v_depths, const float* __restrict__ v_conics,
            const float* __restrict__ v_comps, float* __restrict__ v_means,
            float* __restrict__ v_quats, float* __restrict__ v_scales, int accumulate,
            int vd_stride, const float* __restrict__ opacities, float* __restrict__ v_opacities, int use_comp) {
    // fused-opacity mode (opacities != nullptr): the rasterizer consumed opacity * compensation
    // (plain opacity without compensations), `v_comps` holds the gradient w.r.t. THAT product:
    //   v_compensation = v_prod * opacity,  v_opacity (+)= sum_cam v_prod * compensation
    const float occ = opacities ? __ldg(opacities + g) : 1.0f;
    float v_occ = 0.f;
    // accumulate mode: fetch the running sums now, so their latency hides under the math
    float old_p[3] = {0.f, 0.f, 0.f}, old_s[3] = {0.f, 0.f, 0.f};
    float4 old_q = make_float4(0.f, 0.f, 0.f, 0.f);
    if (accumulate) {
        if (v_means) { old_p[0] = v_means[g * 3 + 0]; old_p[1] = v_means[g * 3 + 1]; old_p[2] = v_means[g * 3 + 2]; }
        if (v_scales) { old_s[0] = v_scales[g * 3 + 0]; old_s[1] = v_scales[g * 3 + 1]; old_s[2] = v_scales[g * 3 + 2]; }
        if (v_quats) old_q = reinterpret_cast<const float4*>(v_quats)[g];
    }
    // camera 0's upstream gradients are fetched NOW (one DRAM round trip together with the
    // parameters) instead of after the forward recompute; further cameras load in the loop
    float pf_conic[3] = {0.f, 0.f, 0.f}, pf_m2[2] = {0.f, 0.f}, pf_z = 0.f, pf_comp = 0.f;
    {  // unconditional (valid rows; unused if camera 0 does not see g): no dependent round trip
        pf_conic[0] = __ldg(v_conics + g * 3 + 0); pf_conic[1] = __ldg(v_conics + g * 3 + 1);
        pf_conic[2] = __ldg(v_conics + g * 3 + 2);
        pf_m2[0] = __ldg(v_means2d + g * 2 + 0); pf_m2[1] = __ldg(v_means2d + g * 2 + 1);
        pf_z = __ldg(v_depths + g * vd_stride);
        if (v_comps) pf_comp = __ldg(v_comps + g);
    }
    // ---- forward recompute: rotation, M = R diag(s), Sigma
    const float4 q = __ldg(reinterpret_cast<const float4*>(quats) + g);
    const float n2 = q.x * q.x + q.y * q.y + q.z * q.z + q.w * q.w;
    const float nrm = fmaxf(sqrtf(n2), 1e-12f);
    const float inv_n = 1.0f / nrm;
    const float w = q.x * inv_n, x = q.y * inv_n, y = q.z * inv_n, z = q.w * inv_n;
    const float R[3][3] = {{1.f - 2.f * (y * y + z * z), 2.f * (x * y - w * z), 2.f * (x * z + w * y)},
                           {2.f * (x * y + w * z), 1.f - 2.f * (x * x + z * z), 2.f * (y * z - w * x)},
                           {2.f * (x * z - w * y), 2.f * (y * z + w * x), 1.f - 2.f * (x * x + y * y)}};
    const float s[3] = {__ldg(scales + g * 3 + 0), __ldg(scales + g * 3 + 1), __ldg(scales + g * 3 + 2)};
    const float p[3] = {__ldg(means + g * 3 + 0), __ldg(means + g * 3 + 1), __ldg(means + g * 3 + 2)};
    float M[3][3], S[3][3];
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
        for (int j = 0; j < 3; ++j) M[i][j] = R[i][j] * s[j];
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
        for (int j = 0; j < 3; ++j) S[i][j] = M[i][0] * M[j][0] + M[i][1] * M[j][1] + M[i][2] * M[j][2];

    float vS[3][3] = {{0.f, 0.f, 0.f}, {0.f, 0.f, 0.f}, {0.f, 0.f, 0.f}};  // dL/dSigma (world)
    float vp[3] = {0.f, 0.f, 0.f};
    bool any = false;

    for (int cam = 0; cam < C; ++cam) {
        const int64_t o = (int64_t)cam * N + g;
        if (__ldg(radii + o) <= 0) continue;
        any = true;
        const float* V = viewmats + cam * 16;
        const float* K = Ks + cam * 9;
        float Wm[3][3], t[3];
#pragma unroll
        for (int i = 0; i < 3; ++i) {
#pragma unroll
            for (int j = 0; j < 3; ++j) Wm[i][j] = __ldg(V + i * 4 + j);
            t[i] = __ldg(V + i * 4 + 3);
        }
        const float fx = __ldg(K + 0), cx = __ldg(K + 2), fy = __ldg(K + 4), cy = __ldg(K + 5);
        const float xc = Wm[0][0] * p[0] + Wm[0][1] * p[1] + Wm[0][2] * p[2] + t[0];
        const float yc = Wm[1][0] * p[0] + Wm[1][1] * p[1] + Wm[1][2] * p[2] + t[1];
        const float zc = Wm[2][0] * p[0] + Wm[2][1] * p[1] + Wm[2][2] * p[2] + t[2];
        float A[3][3], Sc[3][3];
#pragma unroll
        for (int i = 0; i < 3; ++i)
#pragma unroll
            for (int j = 0; j < 3; ++j) A[i][j] = Wm[i][0] * S[0][j] + Wm[i][1] * S[1][j] + Wm[i][2] * S[2][j];
#pragma unroll
        for (int i = 0; i < 3; ++i)
#pragma unroll
            for (int j = 0; j < 3; ++j) Sc[i][j] = A[i][0] * Wm[j][0] + A[i][1] * Wm[j][1] + A[i][2] * Wm[j][2];
        const float Wf = (float)width, Hf = (float)height;
        const float tan_fovx = 0.5f * Wf / fx, tan_fovy = 0.5f * Hf / fy;
        const float lim_x_pos = (Wf - cx) / fx + SC_JAC_PAD * tan_fovx;
        const float lim_x_neg = cx / fx + SC_JAC_PAD * tan_fovx;
        const float lim_y_pos = (Hf - cy) / fy + SC_JAC_PAD * tan_fovy;
        const float lim_y_neg = cy / fy + SC_JAC_PAD * tan_fovy;
        const float rz = 1.0f / zc;
        const float rz2 = rz * rz;
        const float xz = xc * rz, yz = yc * rz;
        const bool x_in = (xz >= -lim_x_neg) && (xz <= lim_x_pos);
        const bool y_in = (yz >= -lim_y_neg) && (yz <= lim_y_pos);
        const float cxz = fminf(fmaxf(xz, -lim_x_neg), lim_x_pos);
        const float cyz = fminf(fmaxf(yz, -lim_y_neg), lim_y_pos);
        const float tx = zc * cxz, ty = zc * cyz;
        const float J00 = fx * rz, J02 = -fx * tx * rz2, J11 = fy * rz, J12 = -fy * ty * rz2;
        // T = J Sc (2x3)
        const float T0[3] = {J00 * Sc[0][0] + J02 * Sc[2][0], J00 * Sc[0][1] + J02 * Sc[2][1],
                             J00 * Sc[0][2] + J02 * Sc[2][2]};
        const float T1[3] = {J11 * Sc[1][0] + J12 * Sc[2][0], J11 * Sc[1][1] + J12 * Sc[2][1],
                             J11 * Sc[1][2] + J12 * Sc[2][2]};
        const float v00 = T0[0] * J00 + T0[2] * J02;
        const float v01 = T0[1] * J11 + T0[2] * J12;
        const float v11 = T1[1] * J11 + T1[2] * J12;
        const float det_orig = v00 * v11 - v01 * v01;
        const float b00 = v00 + eps2d, b11 = v11 + eps2d;
        const float det = b00 * b11 - v01 * v01;
        const float rdet = 1.0f / det;
        const float rdet2 = rdet * rdet;

        // ---- upstream gradients
        const bool c0 = cam == 0;
        const float va = c0 ? pf_conic[0] : __ldg(v_conics + o * 3 + 0), vb = c0 ? pf_conic[1] : __ldg(v_conics + o * 3 + 1),
                    vc = c0 ? pf_conic[2] : __ldg(v_conics + o * 3 + 2);
        const float vmx = c0 ? pf_m2[0] : __ldg(v_means2d + o * 2 + 0), vmy = c0 ? pf_m2[1] : __ldg(v_means2d + o * 2 + 1);
        const float vz_up = c0 ? pf_z : __ldg(v_depths + o * vd_stride);
        // conic = (b11, -v01, b00) / det
        float g_b00 = va * (-b11 * b11 * rdet2) + vb * (v01 * b11 * rdet2) + vc * (rdet - b00 * b11 * rdet2);
        float g_b11 = va * (rdet - b11 * b00 * rdet2) + vb * (v01 * b00 * rdet2) + vc * (-b00 * b00 * rdet2);
        float g_v01 = va * (2.f * b11 * v01 * rdet2) + vb * (-rdet - 2.f * v01 * v01 * rdet2) +
                      vc * (2.f * b00 * v01 * rdet2);
        if (v_comps) {
            float vcomp = c0 ? pf_comp : __ldg(v_comps + o);
            const float ratio = det_orig * rdet;
            if (opacities) {
                const float v_prod = vcomp;
                vcomp = use_comp ? v_prod * occ : 0.f;
                v_occ += use_comp ? (ratio > 0.f ? v_prod * sqrtf(ratio) : 0.f) : v_prod;
            }
            if (ratio > 0.f) {
                const float comp = sqrtf(ratio);
                const float v_r = vcomp * 0.5f / comp;
                const float g_do = v_r * rdet;
                const float g_d = -v_r * det_orig * rdet2;
                g_b00 += g_do * v11 + g_d * b11;
                g_b11 += g_do * v00 + g_d * b00;
                g_v01 += (g_do + g_d) * (-2.f * v01);
            }
        }
        // G2 = [[g_b00, h],[h, g_b11]] with h = g_v01/2
        const float h = 0.5f * g_v01;
        // v_Sc = J^T G2 J ; rows of (G2 J): r0 = g_b00*J0 + h*J1 ; r1 = h*J0 + g_b11*J1
        const float GJ0[3] = {g_b00 * J00, h * J11, g_b00 * J02 + h * J12};
        const float GJ1[3] = {h * J00, g_b11 * J11, h * J02 + g_b11 * J12};
        float vSc[3][3];
#pragma unroll
        for (int j = 0; j < 3; ++j) {
            vSc[0][j] = J00 * GJ0[j];
            vSc[1][j] = J11 * GJ1[j];
            vSc[2][j] = J02 * GJ0[j] + J12 * GJ1[j];
        }
        // v_J = 2 G2 J Sc  (Sc symmetric): v_J[0][k] = 2 * sum_j GJ0[j] Sc[j][k]
        const float vJ00 = 2.f * (GJ0[0] * Sc[0][0] + GJ0[1] * Sc[1][0] + GJ0[2] * Sc[2][0]);
        const float vJ02 = 2.f * (GJ0[0] * Sc[0][2] + GJ0[1] * Sc[1][2] + GJ0[2] * Sc[2][2]);
        const float vJ11 = 2.f * (GJ1[0] * Sc[0][1] + GJ1[1] * Sc[1][1] + GJ1[2] * Sc[2][1]);
        const float vJ12 = 2.f * (GJ1[0] * Sc[0][2] + GJ1[1] * Sc[1][2] + GJ1[2] * Sc[2][2]);
        // camera-space mean gradient
        const float rz3 = rz2 * rz;
        float vxc = 0.f, vyc = 0.f, vzc = vz_up;
        // J00 = fx/z ; J11 = fy/z
        vzc += -fx * rz2 * vJ00 - fy * rz2 * vJ11;
        // J02 = -fx*tx/z^2 with tx = x (in range) or z*lim (clamped)
        if (x_in) {
            vxc += -fx * rz2 * vJ02;
            vzc += 2.f * fx * tx * rz3 * vJ02;
        } else {
            // tx = z*lim -> J02 = -fx*lim/z
            vzc += fx * cxz * rz2 * vJ02;
        }
        if (y_in) {
            vyc += -fy * rz2 * vJ12;
            vzc += 2.f * fy * ty * rz3 * vJ12;
        } else {
            vzc += fy * cyz * rz2 * vJ12;
        }
        // means2d = (fx x/z + cx, fy y/z + cy)
        vxc += vmx * fx * rz;
        vyc += vmy * fy * rz;
        vzc += -(vmx * fx * xc + vmy * fy * yc) * rz2;
        // back to world: v_p += W^T v_pc ; v_S += W^T vSc W
#pragma unroll
        for (int j = 0; j < 3; ++j) vp[j] += Wm[0][j] * vxc + Wm[1][j] * vyc + Wm[2][j] * vzc;
        float B[3][3];  // vSc W
#pragma unroll
        for (int i = 0; i < 3; ++i)
#pragma unroll
            for (int j = 0; j < 3; ++j) B[i][j] = vSc[i][0] * Wm[0][j] + vSc[i][1] * Wm[1][j] + vSc[i][2] * Wm[2][j];
#pragma unroll
        for (int i = 0; i < 3; ++i)
#pragma unroll
            for (int j = 0; j < 3; ++j) vS[i][j] += Wm[0][i] * B[0][j] + Wm[1][i] * B[1][j] + Wm[2][i] * B[2][j];
    }

    float vq[4] = {0.f, 0.f, 0.f, 0.f}, vs[3] = {0.f, 0.f, 0.f};
    if (any) {
        // Sigma = M M^T -> v_M = (vS + vS^T) M
        float vM[3][3];
#pragma unroll
        for (int i = 0; i < 3; ++i)
#pragma unroll
            for (int j = 0; j < 3; ++j)
                vM[i][j] = (vS[i][0] + vS[0][i]) * M[0][j] + (vS[i][1] + vS[1][i]) * M[1][j] +
                           (vS[i][2] + vS[2][i]) * M[2][j];
        float vR[3][3];
#pragma unroll
        for (int j = 0; j < 3; ++j) {
            vs[j] = vM[0][j] * R[0][j] + vM[1][j] * R[1][j] + vM[2][j] * R[2][j];
#pragma unroll
            for (int i = 0; i < 3; ++i) vR[i][j] = vM[i][j] * s[j];
        }
        // R(w,x,y,z) -> gradient wrt the normalised quaternion
        const float vw = 2.f * (x * (vR[2][1] - vR[1][2]) + y * (vR[0][2] - vR[2][0]) + z * (vR[1][0] - vR[0][1]));
        const float vx = 2.f * (-2.f * x * (vR[1][1] + vR[2][2]) + y * (vR[1][0] + vR[0][1]) +
                                z * (vR[2][0] + vR[0][2]) + w * (vR[2][1] - vR[1][2]));
        const float vy = 2.f * (x * (vR[1][0] + vR[0][1]) - 2.f * y * (vR[0][0] + vR[2][2]) +
                                z * (vR[2][1] + vR[1][2]) + w * (vR[0][2] - vR[2][0]));
        const float vzq = 2.f * (x * (vR[2][0] + vR[0][2]) + y * (vR[2][1] + vR[1][2]) -
                                 2.f * z * (vR[0][0] + vR[1][1]) + w * (vR[1][0] - vR[0][1]));
        // normalisation backward: v_q = (v_qn - qn (qn . v_qn)) / |q|
        const float dotq = w * vw + x * vx + y * vy + z * vzq;
        vq[0] = (vw - w * dotq) * inv_n;
        vq[1] = (vx - x * dotq) * inv_n;
        vq[2] = (vy - y * dotq) * inv_n;
        vq[3] = (vzq - z * dotq) * inv_n;
    }
    // (old_* are zero unless accumulate)
    if (v_means) {
        v_means[g * 3 + 0] = old_p[0] + vp[0]; v_means[g * 3 + 1] = old_p[1] + vp[1];
        v_means[g * 3 + 2] = old_p[2] + vp[2];
    }
    if (v_scales) {
        v_scales[g * 3 + 0] = old_s[0] + vs[0]; v_scales[g * 3 + 1] = old_s[1] + vs[1];
        v_scales[g * 3 + 2] = old_s[2] + vs[2];
    }
    if (v_quats)
        reinterpret_cast<float4*>(v_quats)[g] =
            make_float4(old_q.x + vq[0], old_q.y + vq[1], old_q.z + vq[2], old_q.w + vq[3]);
    if (v_opacities) v_opacities[g] = (accumulate ? v_opacities[g] : 0.f) + v_occ;
}

// A CTA owns PB_CHUNK consecutive Gaussians.  Typically a third of them is seen by a camera, so
// a thread-per-Gaussian mapping would run the ~600-instruction VJP with a third of the lanes:
// the CTA first scans the radii, compacts the ids of the visible Gaussians (order preserved, so
// neighbouring threads still touch neighbouring rows) and then runs the VJP with full warps.
__global__ void __launch_bounds__(PB_THREADS, 4)
projection_bwd_kernel(int64_t N, int C, const float* __restrict__ means, const float* __restrict__ quats,
                      const float* __restrict__ scales, const float* __restrict__ viewmats,
                      const float* __restrict__ Ks, int width, int height, float eps2d,
                      const int32_t* __restrict__ radii, const float* __restrict__ v_means2d,
                      const float* __restrict__ v_depths, const float* __restrict__ v_conics,
                      const float* __restrict__ v_comps, float* __restrict__ v_means,
                      float* __restrict__ v_quats, float* __restrict__ v_scales, int accumulate,
                      int vd_stride, const float* __restrict__ opacities, float* __restrict__ v_opacities, int use_comp) {
    __shared__ uint16_t s_list[PB_CHUNK];
    __shared__ int s_wsum[PB_THREADS / 32];
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const int64_t base = (int64_t)blockIdx.x * PB_CHUNK;
    // thread t scans Gaussians base + t, base + t + PB_THREADS, ... (coalesced); list position =
    // round-major, so the list is NOT globally ascending but each round's ids are
    int32_t r[PB_SPAN];
#pragma unroll
    for (int k = 0; k < PB_SPAN; ++k) {
        const int64_t g = base + k * PB_THREADS + tid;
        r[k] = 0;
        if (g < N) {
            r[k] = __ldg(radii + g);
            for (int cam = 1; cam < C; ++cam) r[k] = max(r[k], __ldg(radii + (int64_t)cam * N + g));
        }
    }
    int round_base = 0;
#pragma unroll
    for (int k = 0; k < PB_SPAN; ++k) {
        const int64_t g = base + k * PB_THREADS + tid;
        const bool seen = r[k] > 0;
        if (!seen && g < N && !accumulate) {
            // Gaussians no camera sees: zeros (untouched in accumulate mode)
            if (v_means) { v_means[g * 3 + 0] = 0.f; v_means[g * 3 + 1] = 0.f; v_means[g * 3 + 2] = 0.f; }
            if (v_scales) { v_scales[g * 3 + 0] = 0.f; v_scales[g * 3 + 1] = 0.f; v_scales[g * 3 + 2] = 0.f; }
            if (v_quats) reinterpret_cast<float4*>(v_quats)[g] = make_float4(0.f, 0.f, 0.f, 0.f);
            if (v_opacities) v_opacities[g] = 0.f;
        }
        const unsigned b = __ballot_sync(0xffffffffu, seen);
        if (lane == 0) s_wsum[wid] = __popc(b);
        __syncthreads();
        int off = round_base;
#pragma unroll
        for (int w = 0; w < PB_THREADS / 32; ++w) {
            const int c = s_wsum[w];
            if (w < wid) off += c;
            round_base += c;
        }
        if (seen) s_list[off + __popc(b & ((1u << lane) - 1u))] = (uint16_t)(k * PB_THREADS + tid);
        __syncthreads();
    }
    const int total = round_base;
    for (int i = tid; i < total; i += PB_THREADS)
        pb_gaussian(base + s_list[i], N, C, means, quats, scales, viewmats, Ks, width, height, eps2d, radii,
                    v_means2d, v_depths, v_conics, v_comps, v_means, v_quats, v_scales, accumulate, vd_stride,
                    opacities, v_opacities, use_comp);
}

}  // namespace

static int projection_bwd_launch(int64_t N, int C, const float* means, const float* quats, const float* scales,
                                 const float* viewmats, const float* Ks, int width, int height, float eps2d,
                                 const int32_t* radii, const float* v_means2d, const float* v_depths,
                                 int vd_stride, const float* v_conics, const float* v_compensations,
                                 const float* opacities, int use_comp, float* v_means, float* v_quats,
                                 float* v_scales, float* v_opacities, int accumulate, void* stream) {
    SC_CHECK_ARG(N >= 0 && C >= 1, "N, C");
    if (N == 0) return SC_OK;
    SC_CHECK_ARG(means && quats && scales && viewmats && Ks && radii && v_means2d && v_depths && v_conics,
                 "null pointer");
    SC_CHECK_ARG(((uintptr_t)quats & 15) == 0 && (!v_quats || ((uintptr_t)v_quats & 15) == 0),
                 "16-byte alignment of quats");
    projection_bwd_kernel<<<(unsigned)sc_div_up(N, PB_CHUNK), PB_THREADS, 0, (cudaStream_t)stream>>>(
        N, C, means, quats, scales, viewmats, Ks, width, height, eps2d, radii, v_means2d, v_depths,
        v_conics, v_compensations, v_means, v_quats, v_scales, accumulate, vd_stride, opacities, v_opacities,
        use_comp);
    SC_CHECK_LAUNCH("projection_bwd_kernel");
    return SC_OK;
}

extern "C" int sc_projection_bwd(int64_t N, int C, const float* means, const float* quats,
                                 const float* scales, const float* viewmats, const float* Ks, int width,
                                 int height, float eps2d, const int32_t* radii, const float* v_means2d,
                                 const float* v_depths, const float* v_conics, const float* v_compensations,
                                 float* v_means, float* v_quats, float* v_scales, int accumulate,
                                 void* stream) {
    return projection_bwd_launch(N, C, means, quats, scales, viewmats, Ks, width, height, eps2d, radii,
                                 v_means2d, v_depths, 1, v_conics, v_compensations, nullptr, 1, v_means,
                                 v_quats, v_scales, nullptr, accumulate, stream);
}

extern "C" int sc_projection_bwd_opacity(int64_t N, int C, const float* means, const float* quats,
                                         const float* scales, const float* opacities, const float* viewmats,
                                         const float* Ks, int width, int height, float eps2d,
                                         const int32_t* radii, const float* v_means2d, const float* v_depths,
                                         int v_depths_stride, const float* v_conics,
                                         const float* v_opacities_in, int use_compensations, float* v_means,
                                         float* v_quats, float* v_scales, float* v_opacities,
                                         int accumulate, void* stream) {
    SC_CHECK_ARG(opacities && v_opacities_in && v_opacities && v_depths_stride >= 1, "fused-opacity arguments");
    return projection_bwd_launch(N, C, means, quats, scales, viewmats, Ks, width, height, eps2d, radii,
                                 v_means2d, v_depths, v_depths_stride, v_conics, v_opacities_in, opacities,
                                 use_compensations, v_means, v_quats, v_scales, v_opacities, accumulate, stream);
}
