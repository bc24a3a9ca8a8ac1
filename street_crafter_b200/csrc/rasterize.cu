// Tile-local alpha compositing, forward (A7) and backward (A8).
//
// One CTA per 16x16 tile; every thread owns PPT pixels of ONE column (rows yb, yb+S, ...,
// S = 16/PPT), so the x-dependent half of the conic form (dx, a*dx*dx, b*dx) is computed once
// per (thread, Gaussian) and each pixel costs 3 FMAs more.  Gaussians of the tile's sorted
// range are staged BATCH at a time into shared memory (pre-scaled conic, log2 opacity,
// colours, id).  The per-pair rejection test runs in the log2 domain.  A warp whose pixels
// are all saturated skips the batch body; the CTA leaves when every pixel is saturated.
//
// Backward walks back-to-front starting at the LAST entry any pixel of the tile actually
// reached.  A thread first accumulates the 12 per-Gaussian partials over its own PPT pixels
// in registers (the adds fold into FMAs), then ONE halving butterfly per (warp, Gaussian)
// (16 shuffles instead of 60) leaves one finished sum per lane pair and 12 lanes issue one
// red.global.add each — per-Gaussian accumulation, no per-pixel atomics, and PPT times fewer
// reductions and atomics than a thread-per-pixel layout.  means2d gradients are reduced as
// raw moments and multiplied by the conic once per Gaussian, after the reduction.
//
// FP32-issue + SFU bound; no tensor cores (nothing here is a dense contraction).
// Replaces gsplat.rasterize_to_pixels as called at
// /root/reference/street_gaussian/models/street_gaussian_renderer.py:267-280.
#include <stdlib.h>
#include "common.cuh"

namespace {

constexpr int TILE = 16;
constexpr int BATCH = 128;  // Gaussians staged per round
constexpr float LOG2E = 1.4426950408889634f;
constexpr float LN2 = 0.6931471805599453f;
constexpr float LOG2_ALPHA_FLOOR = -7.994353436858858f;  // log2(1/255)

struct __align__(16) StagedA { float x, y, a2, b2; };          // a2 = .5*a*log2e, b2 = b*log2e
// c2 = .5*c*log2e, lo = log2(opacity), nh = -1/(2*c2) (vertex of s2 along y, times 1/v)
struct __align__(16) StagedB { float c2, lo, nh; int id; };
constexpr float COLUMN_MARGIN = 1e-3f;  // slack of the conservative column pre-test (log2 units)

// x-dependent part of the pair test, shared by the PPT pixels of a thread
struct PairX { float dx, u, v; };
__device__ __forceinline__ PairX pair_x(const StagedA& A, float px) {
    PairX p;
    p.dx = A.x - px;
    p.u = (A.a2 * p.dx) * p.dx;
    p.v = A.b2 * p.dx;
    return p;
}
// log2 of the un-capped alpha and the accept decision; IDENTICAL instruction sequence in the
// forward and the backward kernel, so both take identical accept/reject decisions.
__device__ __forceinline__ bool pair_test(const PairX& X, float c2, float lo, float dy, float& au,
                                          float& alpha) {
    const float s2 = fmaf(fmaf(c2, dy, X.v), dy, X.u);  // sigma * log2e
    au = exp2f(lo - s2);  // opacity * exp(-sigma)
    alpha = fminf(SC_ALPHA_CAP, au);
    return (s2 >= 0.f) && (alpha >= SC_ALPHA_FLOOR);
}
// Conservative pre-test for a thread's whole column segment dy in [dy_lo, dy_hi]: the minimum
// of s2 over the segment (a parabola in dy) against the acceptance threshold.  False only when
// NO pixel of the segment can pass pair_test.
__device__ __forceinline__ bool column_may_pass(const PairX& X, const StagedB& B, float dy_lo, float dy_hi) {
    const float dyc = fminf(fmaxf(X.v * B.nh, dy_lo), dy_hi);
    const float s2min = fmaf(fmaf(B.c2, dyc, X.v), dyc, X.u);
    return !(s2min > B.lo - (LOG2_ALPHA_FLOOR - COLUMN_MARGIN));
}

template <int CH>
__device__ __forceinline__ void stage_entry(int slot, int64_t g, const float2* __restrict__ means2d,
                                            const float* __restrict__ conics, const float* __restrict__ colors,
                                            const float* __restrict__ opacities, StagedA* sA, StagedB* sB,
                                            float* sC) {
    const float2 xy = __ldg(means2d + g);
    const float ca = __ldg(conics + g * 3 + 0), cb = __ldg(conics + g * 3 + 1), cc = __ldg(conics + g * 3 + 2);
    const float op = __ldg(opacities + g);
    sA[slot] = {xy.x, xy.y, 0.5f * LOG2E * ca, LOG2E * cb};
    sB[slot] = {0.5f * LOG2E * cc, __log2f(op), -1.0f / (LOG2E * cc), (int)g};
#pragma unroll
    for (int c = 0; c < CH; ++c) sC[slot * CH + c] = __ldg(colors + g * CH + c);
}

template <int CH>
__device__ __forceinline__ void load_colors(const float* sC, int t, float (&col)[CH]) {
    if constexpr (CH == 4) {
        const float4 v = *reinterpret_cast<const float4*>(sC + t * 4);
        col[0] = v.x; col[1] = v.y; col[2] = v.z; col[3] = v.w;
    } else {
#pragma unroll
        for (int c = 0; c < CH; ++c) col[c] = sC[t * CH + c];
    }
}

// ------------------------------------------------------------------------------- forward
template <int CH, int PPT, bool STATS>
__global__ void __launch_bounds__(TILE * TILE / PPT)
rasterize_fwd_kernel(int64_t N, int64_t n_isects, int width, int height, int tile_w, int tile_h,
                     const float2* __restrict__ means2d, const float* __restrict__ conics,
                     const float* __restrict__ colors, const float* __restrict__ opacities,
                     const float* __restrict__ backgrounds, const int32_t* __restrict__ isect_offsets,
                     const int32_t* __restrict__ flatten_ids, int max_hit, float* __restrict__ render_colors,
                     float* __restrict__ render_alphas, int32_t* __restrict__ last_ids,
                     float* __restrict__ final_T, unsigned long long* __restrict__ stats) {
    constexpr int NT = TILE * TILE / PPT;  // threads per tile
    constexpr int S = TILE / PPT;          // row stride between a thread's pixels
    __shared__ StagedA sA[BATCH];
    __shared__ StagedB sB[BATCH];
    __shared__ __align__(16) float sC[BATCH * CH];
    const int cam = blockIdx.z;
    const int tile_id = blockIdx.y * tile_w + blockIdx.x;
    const int tid = threadIdx.x;
    const int j = blockIdx.x * TILE + (tid & 15);
    const int i0 = blockIdx.y * TILE + (tid >> 4);
    const float px = (float)j + 0.5f, py0 = (float)i0 + 0.5f;

    const int64_t tslot = (int64_t)cam * tile_w * tile_h + tile_id;
    const int range_start = isect_offsets[tslot];
    const int range_end = (tslot == (int64_t)gridDim.z * tile_w * tile_h - 1) ? (int)n_isects
                                                                               : isect_offsets[tslot + 1];
    const int n_batches = (range_end - range_start + BATCH - 1) / BATCH;

    float T[PPT], pix[PPT][CH];
    int last[PPT], hits[PPT];
    unsigned live = 0;  // bit r: pixel r still accumulating
    unsigned visited = 0, contributing = 0;  // STATS only
#pragma unroll
    for (int r = 0; r < PPT; ++r) {
        T[r] = 1.0f;
        last[r] = -1;
        hits[r] = 0;
        if (j < width && i0 + r * S < height) live |= 1u << r;
#pragma unroll
        for (int c = 0; c < CH; ++c) pix[r][c] = 0.f;
    }
    const unsigned inside = live;

    for (int b = 0; b < n_batches; ++b) {
        if (__syncthreads_count(live != 0) == 0) break;
        const int batch_start = range_start + BATCH * b;
        const int bs = min(BATCH, range_end - batch_start);
        for (int k = tid; k < bs; k += NT)
            stage_entry<CH>(k, flatten_ids[batch_start + k], means2d, conics, colors, opacities, sA, sB, sC);
        __syncthreads();
        if (__all_sync(0xffffffffu, live == 0)) continue;
        for (int t = 0; t < bs; ++t) {
            const StagedA A = sA[t];
            const StagedB B = sB[t];
            const PairX X = pair_x(A, px);
            const float dy0 = A.y - py0;
            if (STATS) visited += __popc(live);
            if (!__any_sync(0xffffffffu, live != 0 && column_may_pass(X, B, dy0 - (float)((PPT - 1) * S), dy0)))
                continue;
            float alpha[PPT];
            unsigned ok = 0;
#pragma unroll
            for (int r = 0; r < PPT; ++r) {
                float au;
                if (pair_test(X, B.c2, B.lo, dy0 - (float)(r * S), au, alpha[r])) ok |= 1u << r;
            }
            ok &= live;
            if (!__any_sync(0xffffffffu, ok != 0)) continue;
            float col[CH];
            load_colors<CH>(sC, t, col);
            bool ended = false;
#pragma unroll
            for (int r = 0; r < PPT; ++r) {
                if (ok & (1u << r)) {
                    const float next_T = T[r] * (1.0f - alpha[r]);
                    if (next_T <= SC_T_STOP) {
                        live &= ~(1u << r);
                        ended = true;
                    } else {
                        const float vis = alpha[r] * T[r];
#pragma unroll
                        for (int c = 0; c < CH; ++c) pix[r][c] = fmaf(col[c], vis, pix[r][c]);
                        last[r] = batch_start + t;
                        T[r] = next_T;
                        if (STATS) ++contributing;
                        if (max_hit > 0 && ++hits[r] >= max_hit) {
                            live &= ~(1u << r);
                            ended = true;
                        }
                    }
                }
            }
            if (__any_sync(0xffffffffu, ended) && __all_sync(0xffffffffu, live == 0)) break;
        }
    }
#pragma unroll
    for (int r = 0; r < PPT; ++r) {
        if (inside & (1u << r)) {
            const int64_t pix_id = ((int64_t)cam * height + (i0 + r * S)) * width + j;
            render_alphas[pix_id] = 1.0f - T[r];
            if constexpr (CH == 4) {
                float4 o = make_float4(pix[r][0], pix[r][1], pix[r][2], pix[r][3]);
                if (backgrounds) {
                    o.x += T[r] * __ldg(backgrounds + cam * 4 + 0); o.y += T[r] * __ldg(backgrounds + cam * 4 + 1);
                    o.z += T[r] * __ldg(backgrounds + cam * 4 + 2); o.w += T[r] * __ldg(backgrounds + cam * 4 + 3);
                }
                reinterpret_cast<float4*>(render_colors)[pix_id] = o;
            } else {
#pragma unroll
                for (int c = 0; c < CH; ++c)
                    render_colors[pix_id * CH + c] =
                        backgrounds ? pix[r][c] + T[r] * __ldg(backgrounds + cam * CH + c) : pix[r][c];
            }
            if (last_ids) last_ids[pix_id] = last[r];
            if (final_T) final_T[pix_id] = T[r];
        }
    }
    if (STATS) {
        unsigned long long v = visited, c = contributing;
#pragma unroll
        for (int o = 16; o >= 1; o >>= 1) {
            v += __shfl_xor_sync(0xffffffffu, v, o);
            c += __shfl_xor_sync(0xffffffffu, c, o);
        }
        if ((tid & 31) == 0) {
            atomicAdd(stats + 0, v);
            atomicAdd(stats + 1, c);
        }
    }
}

// ------------------------------------------------------------------------------- backward
// Halving butterfly: 16 per-lane values -> lane l (and its xor-1 partner) ends with the full
// warp sum of value index (l >> 1).
__device__ __forceinline__ float warp_reduce_scatter16(float (&v)[16], int lane) {
    const unsigned full = 0xffffffffu;
    float w8[8], w4[4], w2[2];
    const bool b4 = lane & 16, b3 = lane & 8, b2 = lane & 4, b1 = lane & 2;
#pragma unroll
    for (int k = 0; k < 8; ++k) {
        const float send = b4 ? v[k] : v[k + 8];
        const float keep = b4 ? v[k + 8] : v[k];
        w8[k] = keep + __shfl_xor_sync(full, send, 16);
    }
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        const float send = b3 ? w8[k] : w8[k + 4];
        const float keep = b3 ? w8[k + 4] : w8[k];
        w4[k] = keep + __shfl_xor_sync(full, send, 8);
    }
#pragma unroll
    for (int k = 0; k < 2; ++k) {
        const float send = b2 ? w4[k] : w4[k + 2];
        const float keep = b2 ? w4[k + 2] : w4[k];
        w2[k] = keep + __shfl_xor_sync(full, send, 4);
    }
    const float send = b1 ? w2[0] : w2[1];
    const float keep = b1 ? w2[1] : w2[0];
    float r = keep + __shfl_xor_sync(full, send, 2);
    r += __shfl_xor_sync(full, r, 1);
    return r;  // value index = (b4<<3)|(b3<<2)|(b2<<1)|b1 = lane >> 1
}

// Slots of the reduced vector: 0..4 v_colors, 5..7 second moments (w dx dx, w dx dy, w dy dy),
// 8..9 sum w*h (h = d s2 / d mean), 10..11 the same with |.|, 12 sum w;  w = -au * v_alpha.
template <int CH, int PPT>
__global__ void __launch_bounds__(TILE * TILE / PPT)
rasterize_bwd_kernel(int64_t N, int64_t n_isects, int width, int height, int tile_w, int tile_h,
                     const float2* __restrict__ means2d, const float* __restrict__ conics,
                     const float* __restrict__ colors, const float* __restrict__ opacities,
                     const float* __restrict__ backgrounds, const int32_t* __restrict__ isect_offsets,
                     const int32_t* __restrict__ flatten_ids, const float* __restrict__ final_T,
                     const int32_t* __restrict__ last_ids, const float* __restrict__ v_render_colors,
                     const float* __restrict__ v_render_alphas, float* __restrict__ v_means2d,
                     float* __restrict__ v_means2d_abs, float* __restrict__ v_conics,
                     float* __restrict__ v_colors, float* __restrict__ v_opacities) {
    static_assert(CH <= 5, "reduce-scatter layout holds 16 values");
    constexpr int NT = TILE * TILE / PPT;
    constexpr int NW = NT / 32;
    constexpr int S = TILE / PPT;
    __shared__ StagedA sA[BATCH];
    __shared__ StagedB sB[BATCH];
    __shared__ __align__(16) float sC[BATCH * CH];
    __shared__ int s_max_last[NW];
    const int cam = blockIdx.z;
    const int tile_id = blockIdx.y * tile_w + blockIdx.x;
    const int tid = threadIdx.x;
    const int lane = tid & 31;
    const int j = blockIdx.x * TILE + (tid & 15);
    const int i0 = blockIdx.y * TILE + (tid >> 4);
    const float px = (float)j + 0.5f, py0 = (float)i0 + 0.5f;

    const int64_t tslot = (int64_t)cam * tile_w * tile_h + tile_id;
    const int range_start = isect_offsets[tslot];
    const int range_end = (tslot == (int64_t)gridDim.z * tile_w * tile_h - 1) ? (int)n_isects
                                                                               : isect_offsets[tslot + 1];
    // per-pixel state: T (running transmittance, walking back), R = <v_rc, colour behind> -
    // T_final * v_ra (+ T_final * <bg, v_rc>), so that v_alpha = <v_rc,col> * T - R / (1-alpha)
    float T[PPT], R[PPT], v_rc[PPT][CH];
    int binf[PPT];
    int wmax = -1;
#pragma unroll
    for (int r = 0; r < PPT; ++r) {
        const int i = i0 + r * S;
        const bool inside = (i < height) && (j < width);
        const int64_t pix_id = ((int64_t)cam * height + min(i, height - 1)) * width + min(j, width - 1);
        const float Tf = inside ? final_T[pix_id] : 1.0f;
        binf[r] = inside ? last_ids[pix_id] : -1;
        wmax = max(wmax, binf[r]);
        T[r] = Tf;
        float bg_dot = 0.f;
#pragma unroll
        for (int c = 0; c < CH; ++c) {
            v_rc[r][c] = inside ? v_render_colors[pix_id * CH + c] : 0.f;
            if (backgrounds) bg_dot += __ldg(backgrounds + cam * CH + c) * v_rc[r][c];
        }
        const float v_ra = inside ? v_render_alphas[pix_id] : 0.f;
        R[r] = Tf * (bg_dot - v_ra);
    }
    // the tile never looked past max(last id): start there
#pragma unroll
    for (int o = 16; o >= 1; o >>= 1) wmax = max(wmax, __shfl_xor_sync(0xffffffffu, wmax, o));
    if (lane == 0) s_max_last[tid >> 5] = wmax;
    __syncthreads();
    int tile_last = -1;
#pragma unroll
    for (int w = 0; w < NW; ++w) tile_last = max(tile_last, s_max_last[w]);
    if (tile_last < range_start) return;
    const int eff_end = min(range_end, tile_last + 1);
    const int n_batches = (eff_end - range_start + BATCH - 1) / BATCH;

    for (int b = 0; b < n_batches; ++b) {
        __syncthreads();
        // batch covers [batch_lo, batch_hi), walked from batch_hi-1 down
        const int batch_hi = eff_end - BATCH * b;
        const int batch_lo = max(range_start, batch_hi - BATCH);
        const int bs = batch_hi - batch_lo;
        for (int k = tid; k < bs; k += NT)
            stage_entry<CH>(k, flatten_ids[batch_lo + k], means2d, conics, colors, opacities, sA, sB, sC);
        __syncthreads();
        // warp-uniform trimming: skip entries beyond the warp's own last id
        const int w_hi = min(bs - 1, wmax - batch_lo);
        for (int t = w_hi; t >= 0; --t) {
            const int idx = batch_lo + t;
            const StagedA A = sA[t];
            const StagedB B = sB[t];
            const PairX X = pair_x(A, px);
            const float dy0 = A.y - py0;
            if (!__any_sync(0xffffffffu, column_may_pass(X, B, dy0 - (float)((PPT - 1) * S), dy0))) continue;
            float alpha[PPT], au[PPT];
            unsigned ok = 0;
#pragma unroll
            for (int r = 0; r < PPT; ++r)
                if (pair_test(X, B.c2, B.lo, dy0 - (float)(r * S), au[r], alpha[r]) && idx <= binf[r])
                    ok |= 1u << r;
            if (!__any_sync(0xffffffffu, ok != 0)) continue;
            float col[CH];
            load_colors<CH>(sC, t, col);
            const float hx0 = 2.0f * A.a2 * X.dx;  // d s2 / d mean.x = 2 a2 dx + b2 dy
            const float c2x2 = 2.0f * B.c2;        // d s2 / d mean.y = 2 c2 dy + b2 dx
            float v[16];
#pragma unroll
            for (int k = 0; k < 16; ++k) v[k] = 0.f;
#pragma unroll
            for (int r = 0; r < PPT; ++r) {
                if (ok & (1u << r)) {
                    const float dy = dy0 - (float)(r * S);
                    const float ra = 1.0f / (1.0f - alpha[r]);
                    T[r] *= ra;
                    const float fac = alpha[r] * T[r];
                    float d = 0.f;
#pragma unroll
                    for (int c = 0; c < CH; ++c) {
                        d = fmaf(col[c], v_rc[r][c], d);
                        v[c] = fmaf(fac, v_rc[r][c], v[c]);
                    }
                    const float v_alpha = d * T[r] - R[r] * ra;
                    R[r] = fmaf(d, fac, R[r]);
                    if (au[r] <= SC_ALPHA_CAP) {
                        const float w = -au[r] * v_alpha;
                        const float wdx = w * X.dx, wdy = w * dy;
                        v[5] = fmaf(wdx, X.dx, v[5]);
                        v[6] = fmaf(wdx, dy, v[6]);
                        v[7] = fmaf(wdy, dy, v[7]);
                        const float gx = w * fmaf(A.b2, dy, hx0);
                        const float gy = w * fmaf(c2x2, dy, X.v);
                        v[8] += gx;
                        v[9] += gy;
                        v[10] += fabsf(gx);
                        v[11] += fabsf(gy);
                        v[12] += w;
                    }
                }
            }
            const float red = warp_reduce_scatter16(v, lane);
            if ((lane & 1) == 0) {
                const int k = lane >> 1;
                const int64_t g = B.id;
                float* dst = nullptr;
                float val = red;
                if (k < CH) dst = v_colors + g * CH + k;
                else if (k >= 5 && k < 8) { dst = v_conics + g * 3 + (k - 5); val *= (k == 6) ? 1.0f : 0.5f; }
                else if (k == 8 || k == 9) { dst = v_means2d + g * 2 + (k - 8); val *= LN2; }
                else if ((k == 10 || k == 11) && v_means2d_abs) { dst = v_means2d_abs + g * 2 + (k - 10); val *= LN2; }
                else if (k == 12) { dst = v_opacities + g; val *= -exp2f(-B.lo); }
                if (dst && val != 0.f) atomicAdd(dst, val);
            }
        }
    }
}

// pixels per thread (1, 2, 4 or 8); tunable for measurement through SCB200_FWD_PPT / SCB200_BWD_PPT
int env_ppt(const char* name, int dflt) {
    const char* e = getenv(name);
    if (!e) return dflt;
    const int v = atoi(e);
    return (v == 1 || v == 2 || v == 4 || v == 8) ? v : dflt;
}
int fwd_ppt() { return env_ppt("SCB200_FWD_PPT", 4); }
int bwd_ppt() { return env_ppt("SCB200_BWD_PPT", 4); }

struct FwdArgs {
    int C; int64_t N, n_isects; int width, height;
    const float *means2d, *conics, *colors, *opacities, *backgrounds;
    const int32_t *isect_offsets, *flatten_ids; int max_hit;
    float *render_colors, *render_alphas; int32_t* last_ids; float* final_T; unsigned long long* stats;
    cudaStream_t st;
};

template <int CH, int PPT>
int launch_fwd(const FwdArgs& a) {
    const int tile_w = (a.width + TILE - 1) / TILE, tile_h = (a.height + TILE - 1) / TILE;
    dim3 grid(tile_w, tile_h, a.C), block(TILE * TILE / PPT);
    if (a.stats)
        rasterize_fwd_kernel<CH, PPT, true><<<grid, block, 0, a.st>>>(
            a.N, a.n_isects, a.width, a.height, tile_w, tile_h, reinterpret_cast<const float2*>(a.means2d),
            a.conics, a.colors, a.opacities, a.backgrounds, a.isect_offsets, a.flatten_ids, a.max_hit,
            a.render_colors, a.render_alphas, a.last_ids, a.final_T, a.stats);
    else
        rasterize_fwd_kernel<CH, PPT, false><<<grid, block, 0, a.st>>>(
            a.N, a.n_isects, a.width, a.height, tile_w, tile_h, reinterpret_cast<const float2*>(a.means2d),
            a.conics, a.colors, a.opacities, a.backgrounds, a.isect_offsets, a.flatten_ids, a.max_hit,
            a.render_colors, a.render_alphas, a.last_ids, a.final_T, nullptr);
    SC_CHECK_LAUNCH("rasterize_fwd_kernel");
    return SC_OK;
}

struct BwdArgs {
    int C; int64_t N, n_isects; int width, height;
    const float *means2d, *conics, *colors, *opacities, *backgrounds;
    const int32_t *isect_offsets, *flatten_ids; const float* final_T; const int32_t* last_ids;
    const float *v_render_colors, *v_render_alphas;
    float *v_means2d, *v_means2d_abs, *v_conics, *v_colors, *v_opacities;
    cudaStream_t st;
};

template <int CH, int PPT>
int launch_bwd(const BwdArgs& a) {
    const int tile_w = (a.width + TILE - 1) / TILE, tile_h = (a.height + TILE - 1) / TILE;
    dim3 grid(tile_w, tile_h, a.C), block(TILE * TILE / PPT);
    rasterize_bwd_kernel<CH, PPT><<<grid, block, 0, a.st>>>(
        a.N, a.n_isects, a.width, a.height, tile_w, tile_h, reinterpret_cast<const float2*>(a.means2d), a.conics,
        a.colors, a.opacities, a.backgrounds, a.isect_offsets, a.flatten_ids, a.final_T, a.last_ids,
        a.v_render_colors, a.v_render_alphas, a.v_means2d, a.v_means2d_abs, a.v_conics, a.v_colors,
        a.v_opacities);
    SC_CHECK_LAUNCH("rasterize_bwd_kernel");
    return SC_OK;
}

}  // namespace

extern "C" int sc_rasterize_fwd(int C, int64_t N, int64_t n_isects, int CH, int width, int height,
                                int tile_size, const float* means2d, const float* conics, const float* colors,
                                const float* opacities, const float* backgrounds, const int32_t* isect_offsets,
                                const int32_t* flatten_ids, int max_hit, float* render_colors,
                                float* render_alphas, int32_t* last_ids, float* final_T, uint64_t* stats,
                                void* stream) {
    SC_CHECK_ARG(C >= 1 && N >= 0 && n_isects >= 0 && width > 0 && height > 0, "sizes");
    SC_CHECK_ARG(n_isects < (1ll << 31), "n_isects must fit int32");
    if (tile_size != TILE) {
        sc_set_error("tile_size %d unsupported (this build composites 16x16 tiles)", tile_size);
        return SC_ERR_UNSUPPORTED;
    }
    SC_CHECK_ARG(isect_offsets && render_colors && render_alphas, "null pointer");
    SC_CHECK_ARG(n_isects == 0 || (means2d && conics && colors && opacities && flatten_ids), "null pointer");
    SC_CHECK_ARG(CH != 4 || ((uintptr_t)render_colors & 15) == 0, "16-byte alignment of render_colors");
    const FwdArgs a{C, N, n_isects, width, height, means2d, conics, colors, opacities, backgrounds,
                    isect_offsets, flatten_ids, max_hit, render_colors, render_alphas, last_ids, final_T,
                    reinterpret_cast<unsigned long long*>(stats), (cudaStream_t)stream};
    switch (CH) {
        case 1: return launch_fwd<1, 4>(a);
        case 2: return launch_fwd<2, 4>(a);
        case 3: return launch_fwd<3, 4>(a);
        case 4:
            switch (fwd_ppt()) {
                case 1: return launch_fwd<4, 1>(a);
                case 2: return launch_fwd<4, 2>(a);
                case 8: return launch_fwd<4, 8>(a);
                default: return launch_fwd<4, 4>(a);
            }
        case 5: return launch_fwd<5, 4>(a);
        default: break;
    }
    sc_set_error("unsupported channel count %d (supported 1..5)", CH);
    return SC_ERR_UNSUPPORTED;
}

extern "C" int sc_rasterize_bwd(int C, int64_t N, int64_t n_isects, int CH, int width, int height,
                                int tile_size, const float* means2d, const float* conics, const float* colors,
                                const float* opacities, const float* backgrounds, const int32_t* isect_offsets,
                                const int32_t* flatten_ids, const float* final_T, const int32_t* last_ids,
                                const float* v_render_colors, const float* v_render_alphas, float* v_means2d,
                                float* v_means2d_abs, float* v_conics, float* v_colors, float* v_opacities,
                                void* stream) {
    SC_CHECK_ARG(C >= 1 && N >= 0 && n_isects >= 0 && width > 0 && height > 0, "sizes");
    if (tile_size != TILE) {
        sc_set_error("tile_size %d unsupported (this build composites 16x16 tiles)", tile_size);
        return SC_ERR_UNSUPPORTED;
    }
    if (n_isects == 0 || N == 0) return SC_OK;
    SC_CHECK_ARG(means2d && conics && colors && opacities && isect_offsets && flatten_ids && final_T &&
                     last_ids && v_render_colors && v_render_alphas && v_means2d && v_conics && v_colors &&
                     v_opacities,
                 "null pointer");
    const BwdArgs a{C, N, n_isects, width, height, means2d, conics, colors, opacities, backgrounds,
                    isect_offsets, flatten_ids, final_T, last_ids, v_render_colors, v_render_alphas,
                    v_means2d, v_means2d_abs, v_conics, v_colors, v_opacities, (cudaStream_t)stream};
    switch (CH) {
        case 1: return launch_bwd<1, 4>(a);
        case 2: return launch_bwd<2, 4>(a);
        case 3: return launch_bwd<3, 4>(a);
        case 4:
            switch (bwd_ppt()) {
                case 1: return launch_bwd<4, 1>(a);
                case 2: return launch_bwd<4, 2>(a);
                case 8: return launch_bwd<4, 8>(a);
                default: return launch_bwd<4, 4>(a);
            }
        case 5: return launch_bwd<5, 4>(a);
        default: break;
    }
    sc_set_error("unsupported channel count %d (supported 1..5)", CH);
    return SC_ERR_UNSUPPORTED;
}
