// Tile-local alpha compositing, forward (A7) and backward (A8).
//
// One CTA per 16x16 tile, one thread per pixel.  Gaussians of the tile's sorted range are
// staged 256 at a time into shared memory (pre-scaled conic, log2 opacity, colours), every
// pixel then walks the batch front-to-back.  The per-pair rejection test is done in the log2
// domain (no MUFU for rejected pairs).  A warp whose 32 pixels are all saturated skips the
// batch body; the CTA leaves when every pixel is saturated (__syncthreads_count).
//
// Backward walks back-to-front starting at the LAST batch any pixel of the tile actually
// reached; the 12 per-pixel partials of one Gaussian are reduced across the warp with a
// halving butterfly (16 shuffles instead of 60) that leaves one finished sum per lane pair,
// and 12 lanes issue ONE red.global.add instruction — per-Gaussian accumulation, no per-pixel
// atomics.
//
// FP32-issue + SFU bound; no tensor cores (nothing here is a dense contraction).
// Replaces gsplat.rasterize_to_pixels as called at
// /root/reference/street_gaussian/models/street_gaussian_renderer.py:267-280.
#include "common.cuh"

namespace {

constexpr int TILE = 16;
constexpr int RT = TILE * TILE;  // threads per CTA == batch size
constexpr float LOG2E = 1.4426950408889634f;
constexpr float LN2 = 0.6931471805599453f;
constexpr float LOG2_ALPHA_FLOOR = -7.994353436858858f;  // log2(1/255)

struct __align__(16) StagedA { float x, y, a2, b2; };       // a2 = .5*a*log2e, b2 = b*log2e
struct __align__(16) StagedB { float c2, lo, inv_opac, pad; };  // c2 = .5*c*log2e, lo = log2(opacity)

template <int CH, bool STATS>
__global__ void __launch_bounds__(RT)
rasterize_fwd_kernel(int64_t N, int64_t n_isects, int width, int height, int tile_w, int tile_h,
                     const float2* __restrict__ means2d, const float* __restrict__ conics,
                     const float* __restrict__ colors, const float* __restrict__ opacities,
                     const float* __restrict__ backgrounds, const int32_t* __restrict__ isect_offsets,
                     const int32_t* __restrict__ flatten_ids, int max_hit, float* __restrict__ render_colors,
                     float* __restrict__ render_alphas, int32_t* __restrict__ last_ids,
                     float* __restrict__ final_T, unsigned long long* __restrict__ stats) {
    __shared__ StagedA sA[RT];
    __shared__ StagedB sB[RT];
    __shared__ __align__(16) float sC[RT * CH];
    const int cam = blockIdx.z;
    const int tile_id = blockIdx.y * tile_w + blockIdx.x;
    const int tid = threadIdx.y * TILE + threadIdx.x;
    const int i = blockIdx.y * TILE + threadIdx.y;
    const int j = blockIdx.x * TILE + threadIdx.x;
    const float px = (float)j + 0.5f, py = (float)i + 0.5f;
    const bool inside = (i < height) && (j < width);
    bool done = !inside;

    const int64_t tslot = (int64_t)cam * tile_w * tile_h + tile_id;
    const int range_start = isect_offsets[tslot];
    const int range_end = (tslot == (int64_t)gridDim.z * tile_w * tile_h - 1) ? (int)n_isects
                                                                               : isect_offsets[tslot + 1];
    const int n_batches = (range_end - range_start + RT - 1) / RT;

    float T = 1.0f;
    int cur_idx = -1;
    int hits = 0;
    unsigned visited = 0, contributing = 0;  // STATS only
    float pix[CH];
#pragma unroll
    for (int c = 0; c < CH; ++c) pix[c] = 0.f;

    for (int b = 0; b < n_batches; ++b) {
        if (__syncthreads_count(done) >= RT) break;
        const int batch_start = range_start + RT * b;
        const int idx = batch_start + tid;
        if (idx < range_end) {
            const int64_t g = flatten_ids[idx];  // cam*N + gaussian
            const float2 xy = __ldg(means2d + g);
            const float ca = __ldg(conics + g * 3 + 0), cb = __ldg(conics + g * 3 + 1),
                        cc = __ldg(conics + g * 3 + 2);
            const float op = __ldg(opacities + g);
            sA[tid] = {xy.x, xy.y, 0.5f * LOG2E * ca, LOG2E * cb};
            sB[tid] = {0.5f * LOG2E * cc, __log2f(op), 1.0f / op, 0.f};
#pragma unroll
            for (int c = 0; c < CH; ++c) sC[tid * CH + c] = __ldg(colors + g * CH + c);
        }
        __syncthreads();
        const int bs = min(RT, range_end - batch_start);
        if (!__all_sync(0xffffffffu, done)) {
            for (int t = 0; t < bs && !done; ++t) {
                const StagedA A = sA[t];
                const StagedB B = sB[t];
                if (STATS) ++visited;
                const float dx = A.x - px, dy = A.y - py;
                const float s2 = A.a2 * dx * dx + B.c2 * dy * dy + A.b2 * dx * dy;  // sigma*log2e
                const float e = B.lo - s2;
                if (s2 < 0.f || e < LOG2_ALPHA_FLOOR) continue;
                const float alpha = fminf(SC_ALPHA_CAP, exp2f(e));
                if (alpha < SC_ALPHA_FLOOR) continue;
                const float next_T = T * (1.0f - alpha);
                if (next_T <= SC_T_STOP) {
                    done = true;
                    break;
                }
                const float vis = alpha * T;
#pragma unroll
                for (int c = 0; c < CH; ++c) pix[c] += sC[t * CH + c] * vis;
                cur_idx = batch_start + t;
                T = next_T;
                if (STATS) ++contributing;
                if (max_hit > 0 && ++hits >= max_hit) {
                    done = true;
                    break;
                }
            }
        }
    }
    if (inside) {
        const int64_t pix_id = ((int64_t)cam * height + i) * width + j;
        render_alphas[pix_id] = 1.0f - T;
#pragma unroll
        for (int c = 0; c < CH; ++c)
            render_colors[pix_id * CH + c] = backgrounds ? pix[c] + T * __ldg(backgrounds + cam * CH + c) : pix[c];
        if (last_ids) last_ids[pix_id] = cur_idx;
        if (final_T) final_T[pix_id] = T;
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

template <int CH>
__global__ void __launch_bounds__(RT)
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
    __shared__ int32_t sId[RT];
    __shared__ StagedA sA[RT];
    __shared__ StagedB sB[RT];
    __shared__ __align__(16) float sC[RT * CH];
    __shared__ int s_max_last[RT / 32];
    const int cam = blockIdx.z;
    const int tile_id = blockIdx.y * tile_w + blockIdx.x;
    const int tid = threadIdx.y * TILE + threadIdx.x;
    const int lane = tid & 31;
    const int i = blockIdx.y * TILE + threadIdx.y;
    const int j = blockIdx.x * TILE + threadIdx.x;
    const float px = (float)j + 0.5f, py = (float)i + 0.5f;
    const bool inside = (i < height) && (j < width);
    const int64_t pix_id = ((int64_t)cam * height + min(i, height - 1)) * width + min(j, width - 1);

    const int64_t tslot = (int64_t)cam * tile_w * tile_h + tile_id;
    const int range_start = isect_offsets[tslot];
    const int range_end = (tslot == (int64_t)gridDim.z * tile_w * tile_h - 1) ? (int)n_isects
                                                                               : isect_offsets[tslot + 1];
    const float T_final = inside ? final_T[pix_id] : 1.0f;
    const int bin_final = inside ? last_ids[pix_id] : -1;
    float T = T_final;
    float buffer[CH], v_rc[CH], bg_dot = 0.f;
#pragma unroll
    for (int c = 0; c < CH; ++c) {
        buffer[c] = 0.f;
        v_rc[c] = inside ? v_render_colors[pix_id * CH + c] : 0.f;
        if (backgrounds) bg_dot += __ldg(backgrounds + cam * CH + c) * v_rc[c];
    }
    const float v_ra = inside ? v_render_alphas[pix_id] : 0.f;

    // the tile never looked past max(bin_final): start there
    int wmax = bin_final;
#pragma unroll
    for (int o = 16; o >= 1; o >>= 1) wmax = max(wmax, __shfl_xor_sync(0xffffffffu, wmax, o));
    if (lane == 0) s_max_last[tid >> 5] = wmax;
    __syncthreads();
    int tile_last = -1;
#pragma unroll
    for (int w = 0; w < RT / 32; ++w) tile_last = max(tile_last, s_max_last[w]);
    if (tile_last < range_start) return;
    const int eff_end = min(range_end, tile_last + 1);
    const int n_batches = (eff_end - range_start + RT - 1) / RT;

    for (int b = 0; b < n_batches; ++b) {
        __syncthreads();
        // batch covers [batch_lo, batch_hi), walked from batch_hi-1 down
        const int batch_hi = eff_end - RT * b;
        const int batch_lo = max(range_start, batch_hi - RT);
        const int bs = batch_hi - batch_lo;
        if (tid < bs) {
            const int64_t g = flatten_ids[batch_lo + tid];
            sId[tid] = (int32_t)g;
            const float2 xy = __ldg(means2d + g);
            const float ca = __ldg(conics + g * 3 + 0), cb = __ldg(conics + g * 3 + 1),
                        cc = __ldg(conics + g * 3 + 2);
            const float op = __ldg(opacities + g);
            sA[tid] = {xy.x, xy.y, 0.5f * LOG2E * ca, LOG2E * cb};
            sB[tid] = {0.5f * LOG2E * cc, __log2f(op), 1.0f / op, 0.f};
#pragma unroll
            for (int c = 0; c < CH; ++c) sC[tid * CH + c] = __ldg(colors + g * CH + c);
        }
        __syncthreads();
        // warp-uniform trimming: skip Gaussians beyond the warp's own last id
        const int w_hi = min(bs - 1, wmax - batch_lo);
        for (int t = w_hi; t >= 0; --t) {
            const int idx = batch_lo + t;
            bool valid = inside && (idx <= bin_final);
            float alpha = 0.f, au = 0.f, dx = 0.f, dy = 0.f;
            StagedA A;
            StagedB B;
            if (valid) {
                A = sA[t];
                B = sB[t];
                dx = A.x - px;
                dy = A.y - py;
                // identical arithmetic to the forward kernel -> identical accept/reject decisions
                const float s2 = A.a2 * dx * dx + B.c2 * dy * dy + A.b2 * dx * dy;
                const float e = B.lo - s2;
                if (s2 < 0.f || e < LOG2_ALPHA_FLOOR) {
                    valid = false;
                } else {
                    au = exp2f(e);  // un-capped alpha = opacity * exp(-sigma)
                    alpha = fminf(SC_ALPHA_CAP, au);
                    if (alpha < SC_ALPHA_FLOOR) valid = false;
                }
            }
            if (!__any_sync(0xffffffffu, valid)) continue;
            float v[16];
#pragma unroll
            for (int k = 0; k < 16; ++k) v[k] = 0.f;
            if (valid) {
                const float ra = 1.0f / (1.0f - alpha);
                T *= ra;
                const float fac = alpha * T;
                float v_alpha = 0.f;
#pragma unroll
                for (int c = 0; c < CH; ++c) {
                    const float col = sC[t * CH + c];
                    v[c] = fac * v_rc[c];
                    v_alpha += (col * T - buffer[c] * ra) * v_rc[c];
                    buffer[c] += col * fac;
                }
                v_alpha += T_final * ra * v_ra;
                if (backgrounds) v_alpha += -T_final * ra * bg_dot;
                if (au <= SC_ALPHA_CAP) {
                    const float v_sigma = -au * v_alpha;
                    v[5] = 0.5f * v_sigma * dx * dx;
                    v[6] = v_sigma * dx * dy;
                    v[7] = 0.5f * v_sigma * dy * dy;
                    const float vs2 = v_sigma * LN2;  // staged conic carries a log2e factor
                    const float gx = vs2 * (2.0f * A.a2 * dx + A.b2 * dy);
                    const float gy = vs2 * (A.b2 * dx + 2.0f * B.c2 * dy);
                    v[8] = gx;
                    v[9] = gy;
                    v[10] = fabsf(gx);
                    v[11] = fabsf(gy);
                    v[12] = au * B.inv_opac * v_alpha;  // exp(-sigma) * v_alpha
                }
            }
            const float r = warp_reduce_scatter16(v, lane);
            if ((lane & 1) == 0) {
                const int k = lane >> 1;
                const int64_t g = sId[t];
                float* dst = nullptr;
                if (k < CH) dst = v_colors + g * CH + k;
                else if (k >= 5 && k < 8) dst = v_conics + g * 3 + (k - 5);
                else if (k == 8 || k == 9) dst = v_means2d + g * 2 + (k - 8);
                else if ((k == 10 || k == 11) && v_means2d_abs) dst = v_means2d_abs + g * 2 + (k - 10);
                else if (k == 12) dst = v_opacities + g;
                if (dst && r != 0.f) atomicAdd(dst, r);
            }
        }
    }
}

template <int CH>
int launch_fwd(int C, int64_t N, int64_t n_isects, int width, int height, const float* means2d,
               const float* conics, const float* colors, const float* opacities, const float* backgrounds,
               const int32_t* isect_offsets, const int32_t* flatten_ids, int max_hit, float* render_colors,
               float* render_alphas, int32_t* last_ids, float* final_T, unsigned long long* stats,
               cudaStream_t st) {
    const int tile_w = (width + TILE - 1) / TILE, tile_h = (height + TILE - 1) / TILE;
    dim3 grid(tile_w, tile_h, C), block(TILE, TILE);
    if (stats)
        rasterize_fwd_kernel<CH, true><<<grid, block, 0, st>>>(
            N, n_isects, width, height, tile_w, tile_h, reinterpret_cast<const float2*>(means2d), conics,
            colors, opacities, backgrounds, isect_offsets, flatten_ids, max_hit, render_colors, render_alphas,
            last_ids, final_T, stats);
    else
        rasterize_fwd_kernel<CH, false><<<grid, block, 0, st>>>(
            N, n_isects, width, height, tile_w, tile_h, reinterpret_cast<const float2*>(means2d), conics,
            colors, opacities, backgrounds, isect_offsets, flatten_ids, max_hit, render_colors, render_alphas,
            last_ids, final_T, nullptr);
    SC_CHECK_LAUNCH("rasterize_fwd_kernel");
    return SC_OK;
}

template <int CH>
int launch_bwd(int C, int64_t N, int64_t n_isects, int width, int height, const float* means2d,
               const float* conics, const float* colors, const float* opacities, const float* backgrounds,
               const int32_t* isect_offsets, const int32_t* flatten_ids, const float* final_T,
               const int32_t* last_ids, const float* v_render_colors, const float* v_render_alphas,
               float* v_means2d, float* v_means2d_abs, float* v_conics, float* v_colors, float* v_opacities,
               cudaStream_t st) {
    const int tile_w = (width + TILE - 1) / TILE, tile_h = (height + TILE - 1) / TILE;
    dim3 grid(tile_w, tile_h, C), block(TILE, TILE);
    rasterize_bwd_kernel<CH><<<grid, block, 0, st>>>(
        N, n_isects, width, height, tile_w, tile_h, reinterpret_cast<const float2*>(means2d), conics, colors,
        opacities, backgrounds, isect_offsets, flatten_ids, final_T, last_ids, v_render_colors,
        v_render_alphas, v_means2d, v_means2d_abs, v_conics, v_colors, v_opacities);
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
    cudaStream_t st = (cudaStream_t)stream;
#define ARGS C, N, n_isects, width, height, means2d, conics, colors, opacities, backgrounds, isect_offsets, \
             flatten_ids, max_hit, render_colors, render_alphas, last_ids, final_T,                          \
             reinterpret_cast<unsigned long long*>(stats), st
    switch (CH) {
        case 1: return launch_fwd<1>(ARGS);
        case 2: return launch_fwd<2>(ARGS);
        case 3: return launch_fwd<3>(ARGS);
        case 4: return launch_fwd<4>(ARGS);
        case 5: return launch_fwd<5>(ARGS);
        default: break;
    }
#undef ARGS
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
    cudaStream_t st = (cudaStream_t)stream;
#define ARGS C, N, n_isects, width, height, means2d, conics, colors, opacities, backgrounds, isect_offsets, \
             flatten_ids, final_T, last_ids, v_render_colors, v_render_alphas, v_means2d, v_means2d_abs,   \
             v_conics, v_colors, v_opacities, st
    switch (CH) {
        case 1: return launch_bwd<1>(ARGS);
        case 2: return launch_bwd<2>(ARGS);
        case 3: return launch_bwd<3>(ARGS);
        case 4: return launch_bwd<4>(ARGS);
        case 5: return launch_bwd<5>(ARGS);
        default: break;
    }
#undef ARGS
    sc_set_error("unsupported channel count %d (supported 1..5)", CH);
    return SC_ERR_UNSUPPORTED;
}
