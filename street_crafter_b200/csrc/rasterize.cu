// Tile-local alpha compositing, forward (A7) and backward (A8).
//
// ONE WARP PER 16x16 TILE, 8 pixels per lane: lane l owns column (l & 15) and the rows
// (l >> 4) + 2r, r = 0..7.  Consequences:
//   * no block barriers anywhere: staging, compositing and the gradient reduction are
//     warp-synchronous;
//   * the x-dependent half of the conic form is computed once per (lane, Gaussian), each
//     pixel costs 3 FMAs + ex2 more;
//   * STAGING-TIME CULLING: the tile's sorted list comes from a square 3-sigma box test
//     (gsplat's binning, reproduced bit-exactly upstream).  When a lane stages a Gaussian it
//     tests the actual alpha >= 1/255 ellipse against the tile rectangle (exact convex
//     minimum over the box) and its vertical extent; survivors are compacted into shared
//     memory with one ballot, together with an 8-bit row-pair mask.  A rejected entry costs
//     ~1 issue slot instead of a full per-pixel evaluation by all 32 lanes, and the
//     compositing loop only evaluates the row pairs the ellipse can reach;
//   * backward: a lane first accumulates the 13 per-Gaussian partials over its own 8 pixels
//     in registers (the adds fold into FMAs), then ONE transposed reduction through shared
//     memory per (tile, Gaussian) leaves one finished sum per lane pair and 13 lanes issue one
//     red.global.add each — per-Gaussian accumulation, no per-pixel atomics.  means2d
//     gradients are reduced as raw moments and scaled once per Gaussian after the reduction.
// Forward and backward share the cull and the pair test verbatim, so they take identical
// accept/reject decisions.  The backward starts at the LAST entry any pixel of the tile used.
//
// FP32-issue + SFU bound; no tensor cores (nothing here is a dense contraction).
// Replaces gsplat.rasterize_to_pixels as called at
// /root/reference/street_gaussian/models/street_gaussian_renderer.py:267-280.
#include <stdlib.h>
#include "common.cuh"

#include <algorithm>
#include <type_traits>

namespace {

constexpr int TILE = 16;
constexpr int PPT = 8;       // pixels per lane
constexpr int ROW_STEP = 2;  // row stride between a lane's pixels
constexpr int BATCH = 128;   // capacity of the staged survivor list
constexpr int ROUND = 64;    // candidates examined per staging round (2 per lane)
constexpr unsigned FULL = 0xffffffffu;
constexpr float LOG2E = 1.4426950408889634f;
constexpr float LN2 = 0.6931471805599453f;
constexpr float LOG2_ALPHA_FLOOR = -7.994353436858858f;  // log2(1/255)
constexpr float LO_NOCAP = -0.0016f;  // log2(opacity) below this (opacity < 0.9989): alpha stays below SC_ALPHA_CAP = 0.999
constexpr float CULL_MARGIN = 1e-3f;  // slack of the conservative cull (log2 units of alpha)

constexpr int NRED = 13;         // reduced values per (tile, Gaussian) in the backward pass
constexpr int RED_STRIDE = 36;   // floats per row: 16-byte aligned, conflict-free LDS.128 (see warp_reduce_rows)
struct __align__(16) StagedA { float x, y, a2, b2; };  // a2 = .5*a*log2e, b2 = b*log2e
// c2 = .5*c*log2e, lo = log2(opacity), idx = position in flatten_ids, rmask = row pairs to test
struct __align__(16) StagedB { float c2, lo; int idx; unsigned rmask; };

// x-dependent part of the pair test, shared by the 8 pixels of a lane
struct PairX { float dx, u, v; };
__device__ __forceinline__ PairX pair_x(const StagedA& A, float px) {
    PairX p;
    p.dx = A.x - px;
    p.u = (A.a2 * p.dx) * p.dx;
    p.v = A.b2 * p.dx;
    return p;
}
// The accept decision, IDENTICAL in both kernels, taken in the log domain so that it needs no
// SFU result:  sigma >= 0  and  alpha = opacity * exp(-sigma) >= 1/255
//          <=> 0 <= s2 <= thr,  s2 = sigma * log2(e),  thr = log2(opacity) - log2(1/255)
//          <=> |s2 - h| <= h,   h = thr / 2            (one add + one compare).
// A DEAD pixel (saturated in the forward, not yet reached in the backward, or outside the
// image) holds y = SC_DEAD_Y: s2 becomes huge (or +inf) and fails the window, so liveness costs no
// instruction in the per-pair chain.
__device__ __forceinline__ bool pair_test(const PairX& X, float c2, float h, float dy, float& s2) {
    s2 = fmaf(fmaf(c2, dy, X.v), dy, X.u);  // sigma * log2e
    return fabsf(s2 - h) <= h;
}
// (a large FINITE value, not +inf: dy = -1e18 makes s2 = c2 * 1e36 -> far outside any window (c2 > 0: the 2D
// covariance carries eps2d on its diagonal; non-positive or NaN conics fail the window as well), while every
// product of the backward's moment sums with the pixel's zero weight stays an exact 0 instead of 0 * inf = NaN —
// so the backward needs no per-pixel select on dy)
#define SC_DEAD_Y 1e18f

// ---- Blackwell packed FP32 (fma/add/sub/mul .f32x2 -> SASS FFMA2 / FADD2 / FMUL2) ------------------
// One instruction works on the two halves of a 64-bit register pair: the SAME FLOP rate as scalar FFMA
// (probe.cu: 65.1 vs 71.0 TFLOP/s) in HALF the issue slots, and the compositing kernels are bound by
// instruction issue.  A lane's 8 pixels become 4 pairs (rows 2p and 2p+1 of its column); every
// per-Gaussian operand is broadcast into both halves once per (lane, Gaussian).  Each half is rounded
// exactly like the scalar instruction (.rn, .ftz under --use_fast_math), so the accept / reject
// decisions and the results are those of the scalar build (SC_RASTER_F32X2=0 keeps it for comparison).
#ifndef SC_RASTER_F32X2
#define SC_RASTER_F32X2 1
#endif
typedef unsigned long long f2;
__device__ __forceinline__ f2 pk2(float lo, float hi) { f2 r; asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi)); return r; }
__device__ __forceinline__ f2 bc2(float x) { return pk2(x, x); }
__device__ __forceinline__ float lo2(f2 v) { float a, b; asm("mov.b64 {%0, %1}, %2;" : "=f"(a), "=f"(b) : "l"(v)); (void)b; return a; }
__device__ __forceinline__ float hi2(f2 v) { float a, b; asm("mov.b64 {%0, %1}, %2;" : "=f"(a), "=f"(b) : "l"(v)); (void)a; return b; }
__device__ __forceinline__ float half2of(f2 v, int half) { return half ? hi2(v) : lo2(v); }
__device__ __forceinline__ void set_half(f2& v, int half, float x) { v = half ? pk2(lo2(v), x) : pk2(x, hi2(v)); }
__device__ __forceinline__ f2 fma2(f2 a, f2 b, f2 c) { f2 d; asm("fma.rn.ftz.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c)); return d; }
__device__ __forceinline__ f2 add2(f2 a, f2 b) { f2 d; asm("add.rn.ftz.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b)); return d; }
__device__ __forceinline__ f2 sub2(f2 a, f2 b) { f2 d; asm("sub.rn.ftz.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b)); return d; }
__device__ __forceinline__ f2 mul2(f2 a, f2 b) { f2 d; asm("mul.rn.ftz.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b)); return d; }
// negation of both halves: ptxas folds it into the operand modifier of the consuming packed instruction
// (plain neg.f32: the .ftz form --use_fast_math would pick is kept as a separate flushing FADD)
__device__ __forceinline__ float negf(float x) { float r; asm("neg.f32 %0, %1;" : "=f"(r) : "f"(x)); return r; }
__device__ __forceinline__ f2 neg2(f2 a) { return pk2(negf(lo2(a)), negf(hi2(a))); }
__device__ __forceinline__ float absf_(float x) { float r; asm("abs.f32 %0, %1;" : "=f"(r) : "f"(x)); return r; }
__device__ __forceinline__ f2 abs2(f2 a) { return pk2(absf_(lo2(a)), absf_(hi2(a))); }
__device__ __forceinline__ float rcp_approx(float x) { float r; asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r; }
// per-Gaussian operands of the pair test, broadcast once per (lane, Gaussian)
struct PairX2 { f2 y, c2, v, u, h; };
__device__ __forceinline__ PairX2 pair_x2(const PairX& X, float y, float c2, float h) {
    return {bc2(y), bc2(c2), bc2(X.v), bc2(X.u), bc2(h)};
}
// pair_test on two pixels: the same three roundings per half (dy, inner fma, outer fma, s2 - h)
__device__ __forceinline__ void pair_test2(const PairX2& P, float h, f2 py, f2& dy, f2& s2, bool& ok_lo, bool& ok_hi) {
    dy = sub2(P.y, py);
    s2 = fma2(fma2(P.c2, dy, P.v), dy, P.u);
    const f2 win = sub2(s2, P.h);
    ok_lo = fabsf(lo2(win)) <= h;
    ok_hi = fabsf(hi2(win)) <= h;
}

// minimum of f(t) = p t^2 + q t + r over [lo, hi], p > 0
__device__ __forceinline__ float parabola_min(float p, float q, float r, float lo, float hi) {
    const float t = fminf(fmaxf(-0.5f * q / p, lo), hi);
    return fmaf(fmaf(p, t, q), t, r);
}

// Conservative test of the alpha >= 1/255 ellipse of one Gaussian against the pixel centres
// [X0, X0+15] x [Y0, Y0+15] of a tile.  Returns the mask of row pairs {2r, 2r+1} (bit r) that
// can hold an accepted pixel; 0 = no pixel of the tile can pass pair_test.
__device__ __forceinline__ unsigned tile_cull(float mx, float my, float a2, float b2, float c2, float lo,
                                              float X0, float Y0) {
    const float thr = lo - (LOG2_ALPHA_FLOOR - CULL_MARGIN);  // acceptance needs s2 <= thr
    if (!(thr >= 0.f)) return 0u;
    const float det = a2 * c2 - 0.25f * b2 * b2;
    if (!(det > 0.f) || !(a2 > 0.f) || !(c2 > 0.f)) return 0xffu;  // not positive definite: never cull
    const float qd = thr / det;
    const float hx = sqrtf(qd * c2) * 1.0001f + 1e-3f;  // half extents of {s2 <= thr}
    const float hy = sqrtf(qd * a2) * 1.0001f + 1e-3f;
    const float dxlo = mx - (X0 + 15.0f), dxhi = mx - X0;  // d = mean - pixel over the tile
    const float dylo = my - (Y0 + 15.0f), dyhi = my - Y0;
    if (dxlo > hx || dxhi < -hx || dylo > hy || dyhi < -hy) return 0u;
    const bool inside = dxlo <= 0.f && dxhi >= 0.f && dylo <= 0.f && dyhi >= 0.f;
    if (!inside) {
        // convex form, centre outside the box: the minimum sits on one of the four edges
        float fmin = parabola_min(c2, b2 * dxlo, a2 * dxlo * dxlo, dylo, dyhi);
        fmin = fminf(fmin, parabola_min(c2, b2 * dxhi, a2 * dxhi * dxhi, dylo, dyhi));
        fmin = fminf(fmin, parabola_min(a2, b2 * dylo, c2 * dylo * dylo, dxlo, dxhi));
        fmin = fminf(fmin, parabola_min(a2, b2 * dyhi, c2 * dyhi * dyhi, dxlo, dxhi));
        if (fmin > thr) return 0u;
    }
    // rows whose centre Y0 + row lies within [my - hy, my + hy]  <=>  dyhi - row in [-hy, hy]
    const int row_lo = max(0, (int)ceilf(dyhi - hy));
    const int row_hi = min(15, (int)floorf(dyhi + hy));
    if (row_lo > row_hi) return 0u;
    return ((2u << (row_hi >> 1)) - 1u) & ~((1u << (row_lo >> 1)) - 1u);
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

// One staging round: the warp examines the candidates first, first+step, ... (step = +1
// forward, -1 backward; n_cand <= ROUND of them, lane l takes candidates l and l+32), culls
// them against the tile and appends the survivors, IN CANDIDATE ORDER, to the staged list.
// Returns the new count.
template <int CH, bool WITH_GID>
__device__ __forceinline__ int stage_round(int first, int step, int n_cand, int count, int lane, float X0,
                                           float Y0, const float2* __restrict__ means2d,
                                           const float* __restrict__ conics, const float* __restrict__ colors,
                                           const float* __restrict__ opacities,
                                           const int32_t* __restrict__ flatten_ids, StagedA* sA, StagedB* sB,
                                           float* sC, int* sG) {
    int idx[2];
    int64_t g[2];
    float2 xy[2];
    float ca[2], cb[2], cc[2], op[2];
    bool valid[2];
#pragma unroll
    for (int u = 0; u < 2; ++u) {
        const int k = u * 32 + lane;
        valid[u] = k < n_cand;
        idx[u] = first + step * k;
        g[u] = valid[u] ? (int64_t)__ldg(flatten_ids + idx[u]) : 0;
    }
#pragma unroll
    for (int u = 0; u < 2; ++u) {
        if (valid[u]) {
            xy[u] = __ldg(means2d + g[u]);
            ca[u] = __ldg(conics + g[u] * 3 + 0);
            cb[u] = __ldg(conics + g[u] * 3 + 1);
            cc[u] = __ldg(conics + g[u] * 3 + 2);
            op[u] = __ldg(opacities + g[u]);
        }
    }
#pragma unroll
    for (int u = 0; u < 2; ++u) {
        unsigned rmask = 0u;
        float a2 = 0.f, b2 = 0.f, c2 = 0.f, lo = 0.f;
        if (valid[u]) {
            a2 = 0.5f * LOG2E * ca[u];
            b2 = LOG2E * cb[u];
            c2 = 0.5f * LOG2E * cc[u];
            lo = __log2f(op[u]);
            rmask = tile_cull(xy[u].x, xy[u].y, a2, b2, c2, lo, X0, Y0);
        }
        const unsigned keep = __ballot_sync(FULL, rmask != 0u);
        if (rmask != 0u) {
            const int slot = count + __popc(keep & ((1u << lane) - 1u));
            sA[slot] = {xy[u].x, xy[u].y, a2, b2};
            sB[slot] = {c2, lo, idx[u], rmask};
            if (WITH_GID) sG[slot] = (int)g[u];
#pragma unroll
            for (int c = 0; c < CH; ++c) sC[slot * CH + c] = __ldg(colors + g[u] * CH + c);
        }
        count += __popc(keep);
    }
    return count;
}

// tuning knobs (overridable with -D for the sweeps of tools/raster_variants.sh)
#ifndef SC_FWD_MIN_CTAS
#define SC_FWD_MIN_CTAS 16  // CTAs of ONE warp: register cap 128 -> 16 warps per SM; no spills beats more warps
#endif
#ifndef SC_BWD_MIN_CTAS
#define SC_BWD_MIN_CTAS 16  // register cap 128 -> 16 warps per SM (a 96-register build with 20 warps measured 5 % slower)
#endif
#ifndef SC_WARPS_PER_CTA
#define SC_WARPS_PER_CTA 1  // one tile per CTA: the finest grain for the hardware's CTA scheduler (tile cost varies 100x)
#endif
constexpr int WARPS_PER_CTA = SC_WARPS_PER_CTA;  // independent tiles per CTA (no inter-warp communication)
constexpr int FWD_MIN_CTAS = SC_FWD_MIN_CTAS;
constexpr int BWD_MIN_CTAS = SC_BWD_MIN_CTAS;

#ifndef SC_BWD_BATCH
#define SC_BWD_BATCH 64
#endif
#ifndef SC_FWD_GROUP
#define SC_FWD_GROUP 4   // packed forward: pixel pairs evaluated behind one warp-uniform branch.  4 = all 8 pixels of a
                         // lane at once (one test / vote / branch per Gaussian instead of two, 4 independent chains
                         // to interleave): measured 0.644 -> 0.616 ms per cfg2 step; the same for the backward
                         // (SC_BWD_GROUP 4, 128 registers) measured SLOWER, 1.279 -> 1.309
#endif
#ifndef SC_BWD_GROUP
#define SC_BWD_GROUP 2   // packed backward: pixel pairs evaluated behind one warp-uniform branch (2 = a half tile, 4 = all 8 pixels)
#endif
#ifndef SC_BWD_TEST_FIRST
#define SC_BWD_TEST_FIRST 1   // window tests of all 8 pixels before the two half-tile bodies: 1.275 -> 1.233 ms per cfg2 step
                              // (also hoisting the eight ex2 measured no better: 1.244 -> 1.253; reading entry t + 1's
                              // staged record during entry t in the FORWARD measured worse, 0.616 -> 0.682: registers)
#endif
#ifndef SC_BWD_VRC_SMEM
#define SC_BWD_VRC_SMEM 1  // the backward keeps the per-pixel upstream colour gradient in shared memory (32 registers
                           // freed for the scheduler; one conflict-free LDS.128 per contributing pixel pair)
#endif
constexpr int BATCH_BWD = SC_BWD_BATCH;   // 64 (128 measured slower: 1.281 -> 1.357 ms per cfg2 step)
static_assert(BATCH_BWD >= ROUND && BATCH >= ROUND, "a staging batch holds at least one round of candidates");
constexpr bool BWD_VRC_SMEM = SC_BWD_VRC_SMEM != 0;

template <int CH, bool WITH_GID, int NBATCH = BATCH>
struct WarpSmem {
    StagedA A[NBATCH];
    StagedB B[NBATCH];
    float C[NBATCH * CH];
    int G[WITH_GID ? NBATCH : 1];
    int P[PPT][32];  // per-pixel list positions touched only on rare events (kept out of registers)
    float Rd[WITH_GID ? NRED : 1][RED_STRIDE];  // backward: transposed per-lane partials
    float4 V[(WITH_GID && BWD_VRC_SMEM && CH == 4) ? PPT : 1]
            [(WITH_GID && BWD_VRC_SMEM && CH == 4) ? 32 : 1];  // backward: upstream colour gradients
};

// ------------------------------------------------------------------------------- forward
// Per (pixel, Gaussian) pair the chain is: window test (3 FMA-pipe + 2 compares); then, inside
// one warp-uniform branch per 4 row pairs: ex2, cap, next T, two chained compares (stop / used),
// one select, and the colour FMAs.  Saturation is a RARE event handled in its own warp-uniform
// branch (kill the pixel's y, clear its live bit, record where its backward pass starts).
// last_ids[pixel] = the last list position the backward pass must visit: range_end-1 for a
// pixel that never saturated, else the position just before the entry that saturated it.
template <int CH, bool STATS, bool MAXHIT>
__global__ void __launch_bounds__(WARPS_PER_CTA * 32, FWD_MIN_CTAS)
rasterize_fwd_kernel(int64_t n_isects, int n_cams, int width, int height, int tile_w, int tile_h,
                     const float2* __restrict__ means2d, const float* __restrict__ conics,
                     const float* __restrict__ colors, const float* __restrict__ opacities,
                     const float* __restrict__ backgrounds, const int32_t* __restrict__ isect_offsets,
                     const int32_t* __restrict__ flatten_ids, int max_hit, float* __restrict__ render_colors,
                     float* __restrict__ render_alphas, int32_t* __restrict__ last_ids,
                     float* __restrict__ final_T, unsigned long long* __restrict__ stats,
                     float* __restrict__ depth_norm, const int64_t* __restrict__ n_isects_dev) {
    __shared__ __align__(16) WarpSmem<CH, false> smem[WARPS_PER_CTA];
    // sync-free binning: `n_isects` is the capacity; a list that did not fit was dropped (all offsets 0)
    if (n_isects_dev) n_isects = (*n_isects_dev <= n_isects) ? *n_isects_dev : 0;
    const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
    const int64_t n_slots = (int64_t)n_cams * tile_w * tile_h;
    const int64_t tslot = (int64_t)blockIdx.x * WARPS_PER_CTA + wib;
    if (tslot >= n_slots) return;
    WarpSmem<CH, false>& sm = smem[wib];
    const int cam = (int)(tslot / (tile_w * tile_h));
    const int tile_id = (int)(tslot - (int64_t)cam * tile_w * tile_h);
    const int tyi = tile_id / tile_w, txi = tile_id - tyi * tile_w;
    const int j = txi * TILE + (lane & 15);
    const int i0 = tyi * TILE + (lane >> 4);
    const float px = (float)j + 0.5f;
    const float X0 = (float)(txi * TILE) + 0.5f, Y0 = (float)(tyi * TILE) + 0.5f;

    const int range_start = isect_offsets[tslot];
    const int range_end = (tslot == n_slots - 1) ? (int)n_isects : isect_offsets[tslot + 1];

#if SC_RASTER_F32X2
    // pixel pair p = rows 2p, 2p+1 of the lane's column (low, high half)
    f2 T2[PPT / 2], pix2[PPT / 2][CH], py2[PPT / 2];
#define FWD_T(r) half2of(T2[(r) >> 1], (r) & 1)
#define FWD_PIX(r, c) half2of(pix2[(r) >> 1][c], (r) & 1)
#else
    float T[PPT], pix[PPT][CH], pyv[PPT];
#define FWD_T(r) T[r]
#define FWD_PIX(r, c) pix[r][c]
#endif
    int hits[PPT];
    unsigned live = 0;          // bit r: pixel r still accumulating
    unsigned contributing = 0;  // STATS only
#pragma unroll
    for (int r = 0; r < PPT; ++r) {
        hits[r] = 0;
        const bool in = j < width && i0 + r * ROW_STEP < height;
        if (in) live |= 1u << r;
        const float y = in ? (float)(i0 + r * ROW_STEP) + 0.5f : SC_DEAD_Y;
        sm.P[r][lane] = range_end - 1;
#if SC_RASTER_F32X2
        set_half(py2[r >> 1], r & 1, y);
        if (r & 1) {
            T2[r >> 1] = bc2(1.0f);
#pragma unroll
            for (int c = 0; c < CH; ++c) pix2[r >> 1][c] = 0ull;
        }
#else
        T[r] = 1.0f;
        pyv[r] = y;
#pragma unroll
        for (int c = 0; c < CH; ++c) pix[r][c] = 0.f;
#endif
    }
    const unsigned inside = live;

    int next = range_start;
    bool done = !__any_sync(FULL, live != 0);
    while (next < range_end && !done) {
        int count = 0;
        while (count <= BATCH - ROUND && next < range_end) {
            count = stage_round<CH, false>(next, 1, min(ROUND, range_end - next), count, lane, X0, Y0, means2d,
                                           conics, colors, opacities, flatten_ids, sm.A, sm.B, sm.C, nullptr);
            next += ROUND;
        }
        __syncwarp();
        for (int t = 0; t < count && !done; ++t) {
            const StagedA A = sm.A[t];
            const StagedB B = sm.B[t];
            const PairX X = pair_x(A, px);
            const float hwin = 0.5f * (B.lo - LOG2_ALPHA_FLOOR);
            float col[CH];
            load_colors<CH>(sm.C, t, col);
#if SC_RASTER_F32X2
            // two halves of 4 row pairs (= 2 pixel pairs): one warp-uniform branch each, the chains inside are
            // branch-free so that they interleave
            const PairX2 P2 = pair_x2(X, A.y, B.c2, hwin);
            const f2 lo_2 = bc2(B.lo);
            f2 col2[CH];
#pragma unroll
            for (int c = 0; c < CH; ++c) col2[c] = bc2(col[c]);
            auto halves = [&](auto nocap_tag) {
            constexpr bool NOCAP = decltype(nocap_tag)::value;
            constexpr int G = SC_FWD_GROUP;   // pixel pairs evaluated behind one warp-uniform branch
#pragma unroll
            for (int h = 0; h < 4 / G; ++h) {
                if ((B.rmask >> (2 * G * h)) & ((1u << (2 * G)) - 1u)) {
                    f2 s2[G], dy2;
                    bool ok[2 * G];
                    bool any_ok = false;
#pragma unroll
                    for (int u = 0; u < G; ++u) {
                        pair_test2(P2, hwin, py2[G * h + u], dy2, s2[u], ok[2 * u], ok[2 * u + 1]);
                        any_ok |= ok[2 * u] | ok[2 * u + 1];
                    }
                    if (__any_sync(FULL, any_ok)) {
                        f2 a2[G], next_T2[G];
                        bool any_stop = false;
#pragma unroll
                        for (int u = 0; u < G; ++u) {
                            const int p = G * h + u;
                            // (an opacity below the cap can never reach it: ok implies s2 >= 0; warp-uniform)
                            const f2 e2 = sub2(lo_2, s2[u]);
                            const float au0 = exp2f(lo2(e2)), au1 = exp2f(hi2(e2));
                            a2[u] = pk2(ok[2 * u] ? (NOCAP ? au0 : fminf(SC_ALPHA_CAP, au0)) : 0.f,
                                        ok[2 * u + 1] ? (NOCAP ? au1 : fminf(SC_ALPHA_CAP, au1)) : 0.f);
                            next_T2[u] = fma2(neg2(a2[u]), T2[p], T2[p]);
                            // alpha == 0 leaves T (> 1e-4 for every pixel, live or not) unchanged
                            any_stop |= (lo2(next_T2[u]) <= SC_T_STOP) | (hi2(next_T2[u]) <= SC_T_STOP);
                            if (MAXHIT) any_stop |= (ok[2 * u] && hits[2 * p] + 1 >= max_hit) |
                                                    (ok[2 * u + 1] && hits[2 * p + 1] + 1 >= max_hit);
                        }
                        if (!__any_sync(FULL, any_stop)) {
#pragma unroll
                            for (int u = 0; u < G; ++u) {
                                const int p = G * h + u;
                                const f2 vis = mul2(a2[u], T2[p]);
#pragma unroll
                                for (int c = 0; c < CH; ++c) pix2[p][c] = fma2(col2[c], vis, pix2[p][c]);
                                T2[p] = next_T2[u];
                                if (STATS) contributing += (ok[2 * u] ? 1u : 0u) + (ok[2 * u + 1] ? 1u : 0u);
                                if (MAXHIT) { hits[2 * p] += ok[2 * u] ? 1 : 0; hits[2 * p + 1] += ok[2 * u + 1] ? 1 : 0; }
                            }
                        } else {  // rare: some pixel saturates (or reaches max_hit) on this entry
#pragma unroll
                            for (int q = 0; q < 2 * G; ++q) {
                                const int r = 2 * G * h + q, p = r >> 1, hf = r & 1;
                                const float Tr = half2of(T2[p], hf);
                                const float nT = half2of(next_T2[q >> 1], hf);
                                const float a = half2of(a2[q >> 1], hf);
                                const bool sat = nT <= SC_T_STOP;  // stops BEFORE using the entry
                                const bool used = ok[q] && !sat;
                                const float vis = used ? a * Tr : 0.f;
#pragma unroll
                                for (int c = 0; c < CH; ++c)
                                    set_half(pix2[p][c], hf, fmaf(col[c], vis, half2of(pix2[p][c], hf)));
                                set_half(T2[p], hf, used ? nT : Tr);
                                if (STATS) contributing += used ? 1u : 0u;
                                bool full = false;  // max_hit stops AFTER using the entry
                                if (MAXHIT) {
                                    hits[r] += used ? 1 : 0;
                                    full = used && hits[r] >= max_hit;
                                }
                                if (sat || full) {
                                    set_half(py2[p], hf, SC_DEAD_Y);
                                    live &= ~(1u << r);
                                    sm.P[r][lane] = full ? B.idx : B.idx - 1;
                                }
                            }
                            done = !__any_sync(FULL, live != 0);
                        }
                    }
                }
            }
            };
#else
            // two halves of 4 row pairs: one warp-uniform branch each, the 4 pixel chains inside
            // are branch-free so that they interleave
            auto halves = [&](auto nocap_tag) {
            constexpr bool NOCAP = decltype(nocap_tag)::value;
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                if ((B.rmask >> (4 * h)) & 0xfu) {
                    float s2[4];
                    bool ok[4];
                    bool any_ok = false;
#pragma unroll
                    for (int q = 0; q < 4; ++q) {
                        ok[q] = pair_test(X, B.c2, hwin, A.y - pyv[4 * h + q], s2[q]);
                        any_ok |= ok[q];
                    }
                    if (__any_sync(FULL, any_ok)) {
                        float a[4], next_T[4];
                        bool any_stop = false;
#pragma unroll
                        for (int q = 0; q < 4; ++q) {
                            const int r = 4 * h + q;
                            // (an opacity below the cap can never reach it: ok implies s2 >= 0; warp-uniform)
                            const float au = exp2f(B.lo - s2[q]);
                            a[q] = ok[q] ? (NOCAP ? au : fminf(SC_ALPHA_CAP, au)) : 0.f;
                            next_T[q] = fmaf(-a[q], T[r], T[r]);
                            // a == 0 leaves T (> 1e-4 for every pixel, live or not) unchanged
                            any_stop |= next_T[q] <= SC_T_STOP;
                            if (MAXHIT) any_stop |= ok[q] && (hits[r] + 1 >= max_hit);
                        }
                        if (!__any_sync(FULL, any_stop)) {
#pragma unroll
                            for (int q = 0; q < 4; ++q) {
                                const int r = 4 * h + q;
                                const float vis = a[q] * T[r];
#pragma unroll
                                for (int c = 0; c < CH; ++c) pix[r][c] = fmaf(col[c], vis, pix[r][c]);
                                T[r] = next_T[q];
                                if (STATS) contributing += ok[q] ? 1u : 0u;
                                if (MAXHIT) hits[r] += ok[q] ? 1 : 0;
                            }
                        } else {  // rare: some pixel saturates (or reaches max_hit) on this entry
#pragma unroll
                            for (int q = 0; q < 4; ++q) {
                                const int r = 4 * h + q;
                                const bool sat = next_T[q] <= SC_T_STOP;  // stops BEFORE using the entry
                                const bool used = ok[q] && !sat;
                                const float vis = used ? a[q] * T[r] : 0.f;
#pragma unroll
                                for (int c = 0; c < CH; ++c) pix[r][c] = fmaf(col[c], vis, pix[r][c]);
                                T[r] = used ? next_T[q] : T[r];
                                if (STATS) contributing += used ? 1u : 0u;
                                bool full = false;  // max_hit stops AFTER using the entry
                                if (MAXHIT) {
                                    hits[r] += used ? 1 : 0;
                                    full = used && hits[r] >= max_hit;
                                }
                                if (sat || full) {
                                    pyv[r] = SC_DEAD_Y;
                                    live &= ~(1u << r);
                                    sm.P[r][lane] = full ? B.idx : B.idx - 1;
                                }
                            }
                            done = !__any_sync(FULL, live != 0);
                        }
                    }
                }
            }
            };
#endif
            if (B.lo <= LO_NOCAP) halves(std::true_type{}); else halves(std::false_type{});
        }
        __syncwarp();
    }
#pragma unroll
    for (int r = 0; r < PPT; ++r) {
        if (inside & (1u << r)) {
            const int64_t pix_id = ((int64_t)cam * height + (i0 + r * ROW_STEP)) * width + j;
            const float Tr = FWD_T(r);
            float pixr[CH];
#pragma unroll
            for (int c = 0; c < CH; ++c) pixr[c] = FWD_PIX(r, c);
            render_alphas[pix_id] = 1.0f - Tr;
            if constexpr (CH == 4) {
                float4 o = make_float4(pixr[0], pixr[1], pixr[2], pixr[3]);
                if (backgrounds) {
                    o.x += Tr * __ldg(backgrounds + cam * 4 + 0); o.y += Tr * __ldg(backgrounds + cam * 4 + 1);
                    o.z += Tr * __ldg(backgrounds + cam * 4 + 2); o.w += Tr * __ldg(backgrounds + cam * 4 + 3);
                }
                reinterpret_cast<float4*>(render_colors)[pix_id] = o;
                // A11 fused: the renderer's depth map, render_colors[..., 3] / clamp(alpha, 1e-10)
                // (street_gaussian_renderer.py:283), IEEE division like the torch op it replaces
                if (depth_norm) depth_norm[pix_id] = __fdiv_rn(o.w, fmaxf(1.0f - Tr, 1e-10f));
            } else {
#pragma unroll
                for (int c = 0; c < CH; ++c)
                    render_colors[pix_id * CH + c] =
                        backgrounds ? pixr[c] + Tr * __ldg(backgrounds + cam * CH + c) : pixr[c];
            }
            if (last_ids) last_ids[pix_id] = sm.P[r][lane];
            if (final_T) final_T[pix_id] = Tr;
        }
    }
    if (STATS) {
        // visited = list entries up to and including the one that ended the pixel (the
        // algorithmic count: independent of the culling above); a pixel that never ended
        // visited its whole list
        unsigned long long v = 0, c = contributing;
#pragma unroll
        for (int r = 0; r < PPT; ++r)
            if (inside & (1u << r)) {
                const int lastp = sm.P[r][lane];
                const int visited_to = (live & (1u << r)) ? range_end - 1 : min(lastp + 1, range_end - 1);
                v += (unsigned long long)max(visited_to - range_start + 1, 0);
            }
#pragma unroll
        for (int o = 16; o >= 1; o >>= 1) {
            v += __shfl_xor_sync(FULL, v, o);
            c += __shfl_xor_sync(FULL, c, o);
        }
        if (lane == 0) {
            atomicAdd(stats + 0, v);
            atomicAdd(stats + 1, c);
        }
    }
#undef FWD_T
#undef FWD_PIX
}

// ------------------------------------------------------------------------------- backward
// Sum of NRED per-lane values over the warp, through shared memory: every lane stores its
// values as a column of Rd[NRED][RED_STRIDE]; lane l then sums half (l & 1) of row (l >> 1)
// with four LDS.128 (rows are 36 floats apart, so the 8 lanes of a quarter-warp hit 32
// distinct banks) and one shuffle joins the halves: 13 STS + 4 LDS.128 + 17 FADD (7 FADD2 + 2 FADD packed) + 1 SHFL per
// (tile, Gaussian) instead of 16 SHFL + 31 FADD + 30 selects for the register butterfly.
// Lanes 2k and 2k+1 both return the total of value k (k < NRED; other lanes return 0).
__device__ __forceinline__ float warp_reduce_rows(const float (&v)[NRED], float (*Rd)[RED_STRIDE], int lane) {
#pragma unroll
    for (int k = 0; k < NRED; ++k) Rd[k][lane] = v[k];
    __syncwarp();
    const int k = lane >> 1;
    float r = 0.f;
    if (k < NRED) {
#if SC_RASTER_F32X2
        // 16 floats as 8 register pairs: 7 packed adds + 1 scalar add instead of 15 scalar adds
        const ulonglong2* row = reinterpret_cast<const ulonglong2*>(&Rd[k][(lane & 1) * 16]);
        const ulonglong2 a = row[0], b = row[1], c = row[2], d = row[3];
        const f2 s = add2(add2(add2(a.x, a.y), add2(b.x, b.y)), add2(add2(c.x, c.y), add2(d.x, d.y)));
        r = lo2(s) + hi2(s);
#else
        const float4* row = reinterpret_cast<const float4*>(&Rd[k][(lane & 1) * 16]);
        const float4 a = row[0], b = row[1], c = row[2], d = row[3];
        r = ((a.x + a.y) + (a.z + a.w)) + ((b.x + b.y) + (b.z + b.w)) +
            (((c.x + c.y) + (c.z + c.w)) + ((d.x + d.y) + (d.z + d.w)));
#endif
    }
    r += __shfl_xor_sync(FULL, r, 1);
    __syncwarp();  // the rows are rewritten by the next Gaussian
    return r;
}

// Slots of the reduced vector: 0..4 v_colors, 5..7 second moments (w dx dx, w dx dy, w dy dy),
// 8..9 sum w*h (h = d s2 / d mean), 10..11 the same with |.|, 12 sum w;  w = -au * v_alpha.
// A pixel joins the walk when the list position reaches its last_ids entry (a rare event,
// handled in its own warp-uniform branch: until then its y is SC_DEAD_Y and every test fails).
template <int CH>
__global__ void __launch_bounds__(WARPS_PER_CTA * 32, BWD_MIN_CTAS)
rasterize_bwd_kernel(int64_t n_isects, int n_cams, int width, int height, int tile_w, int tile_h,
                     const float2* __restrict__ means2d, const float* __restrict__ conics,
                     const float* __restrict__ colors, const float* __restrict__ opacities,
                     const float* __restrict__ backgrounds, const int32_t* __restrict__ isect_offsets,
                     const int32_t* __restrict__ flatten_ids, const float* __restrict__ final_T,
                     const int32_t* __restrict__ last_ids, const float* __restrict__ v_render_colors,
                     const float* __restrict__ v_render_alphas, float* __restrict__ v_means2d,
                     float* __restrict__ v_means2d_abs, float* __restrict__ v_conics,
                     float* __restrict__ v_colors, float* __restrict__ v_opacities) {
    static_assert(CH <= 5, "slots 0..4 hold v_colors");
    __shared__ __align__(16) WarpSmem<CH, true, BATCH_BWD> smem[WARPS_PER_CTA];
    const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
    const int64_t n_slots = (int64_t)n_cams * tile_w * tile_h;
    const int64_t tslot = (int64_t)blockIdx.x * WARPS_PER_CTA + wib;
    if (tslot >= n_slots) return;
    WarpSmem<CH, true, BATCH_BWD>& sm = smem[wib];
    constexpr bool VSM = BWD_VRC_SMEM && CH == 4;
    const int cam = (int)(tslot / (tile_w * tile_h));
    const int tile_id = (int)(tslot - (int64_t)cam * tile_w * tile_h);
    const int tyi = tile_id / tile_w, txi = tile_id - tyi * tile_w;
    const int j = txi * TILE + (lane & 15);
    const int i0 = tyi * TILE + (lane >> 4);
    const float px = (float)j + 0.5f;
    const float X0 = (float)(txi * TILE) + 0.5f, Y0 = (float)(tyi * TILE) + 0.5f;
    const int range_start = isect_offsets[tslot];

    // per-pixel state: T (running transmittance, walking back), R = <v_rc, colour behind> -
    // T_final * v_ra (+ T_final * <bg, v_rc>), so that v_alpha = <v_rc,col> * T - R / (1-alpha)
#if SC_RASTER_F32X2
    // pixel pair p = rows 2p, 2p+1 of the lane's column (low, high half).
    // Shared-memory layout of the upstream colour gradients (CH == 4): V[2p][lane] = (c0 lo, c0 hi, c1 lo, c1 hi),
    // V[2p+1][lane] = (c2 lo, c2 hi, c3 lo, c3 hi) — two conflict-free LDS.128 deliver four aligned register pairs.
    f2 T2[PPT / 2], R2[PPT / 2], py2[PPT / 2], v_rc2[VSM ? 1 : PPT / 2][CH];
#else
    float T[PPT], R[PPT], v_rc[VSM ? 1 : PPT][CH], pyv[PPT];
#endif
    int tile_last = -1;
#pragma unroll
    for (int p = 0; p < PPT / 2; ++p) {
        float Tf2[2], Rr2[2], vin2[2][CH];
#pragma unroll
        for (int u = 0; u < 2; ++u) {
            const int r = 2 * p + u;
            const int i = i0 + r * ROW_STEP;
            const bool inside = (i < height) && (j < width);
            const int64_t pix_id = ((int64_t)cam * height + min(i, height - 1)) * width + min(j, width - 1);
            const float Tf = inside ? final_T[pix_id] : 1.0f;
            const int binf = inside ? last_ids[pix_id] : -1;
            sm.P[r][lane] = binf;
            tile_last = max(tile_last, binf);
            float bg_dot = 0.f;
#pragma unroll
            for (int c = 0; c < CH; ++c) {
                vin2[u][c] = inside ? v_render_colors[pix_id * CH + c] : 0.f;
                if (backgrounds) bg_dot += __ldg(backgrounds + cam * CH + c) * vin2[u][c];
            }
            const float v_ra = (inside && v_render_alphas) ? v_render_alphas[pix_id] : 0.f;  // nullable: zeros
            Tf2[u] = Tf;
            Rr2[u] = Tf * (bg_dot - v_ra);
        }
#if SC_RASTER_F32X2
        T2[p] = pk2(Tf2[0], Tf2[1]);
        R2[p] = pk2(Rr2[0], Rr2[1]);
        py2[p] = bc2(SC_DEAD_Y);
        if constexpr (VSM) {
            sm.V[2 * p][lane] = make_float4(vin2[0][0], vin2[1][0], vin2[0][1], vin2[1][1]);
            sm.V[2 * p + 1][lane] = make_float4(vin2[0][2], vin2[1][2], vin2[0][3], vin2[1][3]);
        } else {
#pragma unroll
            for (int c = 0; c < CH; ++c) v_rc2[p][c] = pk2(vin2[0][c], vin2[1][c]);
        }
#else
#pragma unroll
        for (int u = 0; u < 2; ++u) {
            const int r = 2 * p + u;
            T[r] = Tf2[u];
            R[r] = Rr2[u];
            pyv[r] = SC_DEAD_Y;
            if constexpr (VSM) {
                sm.V[r][lane] = make_float4(vin2[u][0], vin2[u][1], vin2[u][2], vin2[u][3]);
            } else {
#pragma unroll
                for (int c = 0; c < CH; ++c) v_rc[r][c] = vin2[u][c];
            }
        }
#endif
    }
    // the tile never used an entry past max(last id): start there
#pragma unroll
    for (int o = 16; o >= 1; o >>= 1) tile_last = max(tile_last, __shfl_xor_sync(FULL, tile_last, o));
    if (tile_last < range_start) return;
    int next_wake = tile_last;  // largest last id among this lane's sleeping pixels (-1: none)
    {
        int m = -1;
#pragma unroll
        for (int r = 0; r < PPT; ++r) m = max(m, sm.P[r][lane]);
        next_wake = m;
    }

    // where this lane's reduced value goes (lane pair k = lane >> 1 ends with slot k)
    const int k_slot = lane >> 1;
    float* my_base = nullptr;
    int my_stride = 0;
    float my_scale = 1.0f;
    if ((lane & 1) == 0) {
        if (k_slot < CH) { my_base = v_colors + k_slot; my_stride = CH; }
        else if (k_slot >= 5 && k_slot < 8) { my_base = v_conics + (k_slot - 5); my_stride = 3; my_scale = (k_slot == 6) ? 1.0f : 0.5f; }
        else if (k_slot == 8 || k_slot == 9) { my_base = v_means2d + (k_slot - 8); my_stride = 2; my_scale = LN2; }
        else if ((k_slot == 10 || k_slot == 11) && v_means2d_abs) { my_base = v_means2d_abs + (k_slot - 10); my_stride = 2; my_scale = LN2; }
        else if (k_slot == 12) { my_base = v_opacities; my_stride = 1; }
    }
    const bool my_is_opacity = k_slot == 12;

    int next = tile_last;  // candidates are taken from here downwards
    while (next >= range_start) {
        int count = 0;
        while (count <= BATCH_BWD - ROUND && next >= range_start) {
            count = stage_round<CH, true>(next, -1, min(ROUND, next - range_start + 1), count, lane, X0, Y0,
                                          means2d, conics, colors, opacities, flatten_ids, sm.A, sm.B, sm.C,
                                          sm.G);
            next -= ROUND;
        }
        __syncwarp();
        for (int t = 0; t < count; ++t) {
            const StagedA A = sm.A[t];
            const StagedB B = sm.B[t];
            if (__any_sync(FULL, B.idx <= next_wake)) {  // rare: a pixel's walk starts at this entry
                int m = -1;
#pragma unroll
                for (int r = 0; r < PPT; ++r) {
                    const int b = sm.P[r][lane];
                    if (b >= B.idx) {
#if SC_RASTER_F32X2
                        set_half(py2[r >> 1], r & 1, (float)(i0 + r * ROW_STEP) + 0.5f);
#else
                        pyv[r] = (float)(i0 + r * ROW_STEP) + 0.5f;
#endif
                        sm.P[r][lane] = -1;
                    } else {
                        m = max(m, b);
                    }
                }
                next_wake = m;
            }
            const PairX X = pair_x(A, px);
            const float hwin = 0.5f * (B.lo - LOG2_ALPHA_FLOOR);
            float col[CH];
            load_colors<CH>(sm.C, t, col);
            const float hx0 = 2.0f * A.a2 * X.dx;  // d s2 / d mean.x = 2 a2 dx + b2 dy
            const float c2x2 = 2.0f * B.c2;        // d s2 / d mean.y = 2 c2 dy + b2 dx
            // dx is constant over a lane's 8 pixels, so of the moments of w = -au * v_alpha only
            // sum w, sum w dy, sum w dy^2 (and the two |.| sums, which are not linear) are
            // accumulated per pixel; the dx factors are applied once per (lane, Gaussian)
            bool any_entry = false;
            // A Gaussian whose opacity is below the cap can never reach it (ok implies s2 >= 0, so
            // alpha <= opacity): its chains drop the clamp and the "not clamped" predicate.  The
            // choice is per Gaussian, i.e. warp-uniform; both instances are the same code otherwise.
#if SC_RASTER_F32X2
            // dx is constant over a lane's 8 pixels, so of the moments of w = -au * v_alpha only sum w, sum w dy,
            // sum w dy^2 (and the two |.| sums, which are not linear) are accumulated per pixel — here per HALF:
            // the low halves sum the even rows, the high halves the odd rows, joined once per (lane, Gaussian).
            const PairX2 P2 = pair_x2(X, A.y, B.c2, hwin);
            const f2 lo_2 = bc2(B.lo), one2 = bc2(1.0f), b22 = bc2(A.b2), hx02 = bc2(hx0), c2x22 = bc2(c2x2);
            f2 col2[CH], vc2[CH], sw2 = 0ull, swy2 = 0ull, swyy2 = 0ull;
            f2 abx2 = 0ull, aby2 = 0ull;
#pragma unroll
            for (int c = 0; c < CH; ++c) { col2[c] = bc2(col[c]); vc2[c] = 0ull; }
            auto halves = [&](auto nocap_tag) {
            constexpr bool NOCAP = decltype(nocap_tag)::value;
            constexpr int G = SC_BWD_GROUP;   // pixel pairs behind one warp-uniform branch (2: a half of the tile)
#if SC_BWD_TEST_FIRST
            // the window tests of ALL pairs first (cheap, independent): the second half's test latency hides under
            // the first half's body
            f2 s2_all[4], dy2_all[4];
            bool ok_all[8];
#pragma unroll
            for (int p = 0; p < 4; ++p) pair_test2(P2, hwin, py2[p], dy2_all[p], s2_all[p], ok_all[2 * p], ok_all[2 * p + 1]);
#endif
#pragma unroll
            for (int h = 0; h < 4 / G; ++h) {
                if ((B.rmask >> (2 * G * h)) & ((1u << (2 * G)) - 1u)) {
                    f2 s2[G], dy2[G];
                    bool ok[2 * G];
                    bool any_ok = false;
#pragma unroll
                    for (int u = 0; u < G; ++u) {
#if SC_BWD_TEST_FIRST
                        s2[u] = s2_all[G * h + u]; dy2[u] = dy2_all[G * h + u];
                        ok[2 * u] = ok_all[2 * (G * h + u)]; ok[2 * u + 1] = ok_all[2 * (G * h + u) + 1];
#else
                        pair_test2(P2, hwin, py2[G * h + u], dy2[u], s2[u], ok[2 * u], ok[2 * u + 1]);
#endif
                        any_ok |= ok[2 * u] | ok[2 * u + 1];
                    }
                    if (__any_sync(FULL, any_ok)) {
                        any_entry = true;
#pragma unroll
                        for (int u = 0; u < G; ++u) {
                            const int p = G * h + u;
                            const bool ok0 = ok[2 * u], ok1 = ok[2 * u + 1];
                            const f2 dy = dy2[u];   // finite for dead pixels too; their weight w is an exact 0
                            const f2 e2 = sub2(lo_2, s2[u]);
                            const float au0 = exp2f(lo2(e2)), au1 = exp2f(hi2(e2));
                            const f2 a = pk2(ok0 ? (NOCAP ? au0 : fminf(SC_ALPHA_CAP, au0)) : 0.f,
                                             ok1 ? (NOCAP ? au1 : fminf(SC_ALPHA_CAP, au1)) : 0.f);
                            const f2 oma = sub2(one2, a);
                            // MUFU.RCP; exactly 1 when the pair is not used
                            const f2 ra = pk2(rcp_approx(lo2(oma)), rcp_approx(hi2(oma)));
                            T2[p] = mul2(T2[p], ra);
                            const f2 fac = mul2(a, T2[p]);
                            f2 vr[CH];
                            if constexpr (VSM) {
                                const ulonglong2 v01 = *reinterpret_cast<const ulonglong2*>(&sm.V[2 * p][lane]);
                                const ulonglong2 v23 = *reinterpret_cast<const ulonglong2*>(&sm.V[2 * p + 1][lane]);
                                vr[0] = v01.x; vr[1] = v01.y; vr[2] = v23.x; vr[3] = v23.y;
                            } else {
#pragma unroll
                                for (int c = 0; c < CH; ++c) vr[c] = v_rc2[p][c];
                            }
                            f2 d = mul2(col2[0], vr[0]);
                            vc2[0] = fma2(fac, vr[0], vc2[0]);
#pragma unroll
                            for (int c = 1; c < CH; ++c) {
                                d = fma2(col2[c], vr[c], d);
                                vc2[c] = fma2(fac, vr[c], vc2[c]);
                            }
                            const f2 v_alpha = fma2(d, T2[p], neg2(mul2(R2[p], ra)));  // d T - R / (1 - alpha)
                            R2[p] = fma2(d, fac, R2[p]);
                            // a clamped alpha passes no gradient; unclamped: -alpha * v_alpha is already 0 when !ok
                            const f2 w = NOCAP ? mul2(neg2(a), v_alpha)
                                               : mul2(neg2(pk2((ok0 && au0 <= SC_ALPHA_CAP) ? au0 : 0.f,
                                                               (ok1 && au1 <= SC_ALPHA_CAP) ? au1 : 0.f)), v_alpha);
                            const f2 wdy = mul2(w, dy);
                            sw2 = add2(sw2, w);
                            swy2 = add2(swy2, wdy);
                            swyy2 = fma2(wdy, dy, swyy2);
                            const f2 ax = mul2(w, fma2(b22, dy, hx02));    // w * d s2 / d mean.x
                            const f2 ay = mul2(w, fma2(c2x22, dy, P2.v));  // w * d s2 / d mean.y
                            abx2 = add2(abx2, abs2(ax));   // |.| is a folded operand modifier of the packed add
                            aby2 = add2(aby2, abs2(ay));
                        }
                    }
                }
            }
            };
#else
            float vc[CH], sw = 0.f, swy = 0.f, swyy = 0.f, abx = 0.f, aby = 0.f;
#pragma unroll
            for (int c = 0; c < CH; ++c) vc[c] = 0.f;
            // two halves of 4 row pairs: one warp-uniform branch each; the 4 pixel chains inside
            // are branch-free so that they interleave
            // A Gaussian whose opacity is below the cap can never reach it (ok implies s2 >= 0, so
            // alpha <= opacity): its chains drop the clamp and the "not clamped" predicate.  The
            // choice is per Gaussian, i.e. warp-uniform; both instances are the same code otherwise.
            auto halves = [&](auto nocap_tag) {
            constexpr bool NOCAP = decltype(nocap_tag)::value;
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                if ((B.rmask >> (4 * h)) & 0xfu) {
                    float s2[4], dyq[4];
                    bool ok[4];
                    bool any_ok = false;
#pragma unroll
                    for (int q = 0; q < 4; ++q) {
                        dyq[q] = A.y - pyv[4 * h + q];
                        ok[q] = pair_test(X, B.c2, hwin, dyq[q], s2[q]);
                        any_ok |= ok[q];
                    }
                    if (__any_sync(FULL, any_ok)) {
                        any_entry = true;
#pragma unroll
                        for (int q = 0; q < 4; ++q) {
                            const int r = 4 * h + q;
                            // (scalar build) keep the pixels that fail the test out of the moment sums
                            const float dy = ok[q] ? dyq[q] : 0.f;
                            const float au = exp2f(B.lo - s2[q]);
                            const float a = ok[q] ? (NOCAP ? au : fminf(SC_ALPHA_CAP, au)) : 0.f;
                            const float ra = __fdividef(1.0f, 1.0f - a);  // MUFU.RCP; exactly 1 when the pair is not used
                            T[r] *= ra;
                            const float fac = a * T[r];
                            float d = 0.f;
                            float vr[CH];
                            if constexpr (VSM) {
                                const float4 v4 = sm.V[r][lane];
                                vr[0] = v4.x; vr[1] = v4.y; vr[2] = v4.z; vr[3] = v4.w;
                            } else {
#pragma unroll
                                for (int c = 0; c < CH; ++c) vr[c] = v_rc[r][c];
                            }
#pragma unroll
                            for (int c = 0; c < CH; ++c) {
                                d = fmaf(col[c], vr[c], d);
                                vc[c] = fmaf(fac, vr[c], vc[c]);
                            }
                            const float v_alpha = d * T[r] - R[r] * ra;
                            R[r] = fmaf(d, fac, R[r]);
                            // a clamped alpha passes no gradient; unclamped: -a * v_alpha is already 0 when !ok
                            const float w = NOCAP ? -a * v_alpha : ((ok[q] && au <= SC_ALPHA_CAP) ? -au * v_alpha : 0.f);
                            const float wdy = w * dy;
                            sw += w;
                            swy += wdy;
                            swyy = fmaf(wdy, dy, swyy);
                            abx = fmaf(fabsf(w), fabsf(fmaf(A.b2, dy, hx0)), abx);
                            aby = fmaf(fabsf(w), fabsf(fmaf(c2x2, dy, X.v)), aby);
                        }
                    }
                }
            }
            };
#endif
            if (B.lo <= LO_NOCAP) halves(std::true_type{}); else halves(std::false_type{});
            if (!any_entry) continue;  // warp-uniform
            float v[NRED];
#if SC_RASTER_F32X2
            // join the halves (even + odd rows)
            const float sw = lo2(sw2) + hi2(sw2), swy = lo2(swy2) + hi2(swy2), swyy = lo2(swyy2) + hi2(swyy2);
            const float abx = lo2(abx2) + hi2(abx2), aby = lo2(aby2) + hi2(aby2);
#pragma unroll
            for (int c = 0; c < CH; ++c) v[c] = lo2(vc2[c]) + hi2(vc2[c]);
#else
#pragma unroll
            for (int c = 0; c < CH; ++c) v[c] = vc[c];
#endif
#pragma unroll
            for (int c = CH; c < 5; ++c) v[c] = 0.f;
            const float swx = sw * X.dx;       // sum w dx
            v[5] = swx * X.dx;                 // sum w dx dx
            v[6] = swy * X.dx;                 // sum w dx dy
            v[7] = swyy;                       // sum w dy dy
            v[8] = fmaf(2.0f * A.a2, swx, A.b2 * swy);  // sum w (2 a2 dx + b2 dy)
            v[9] = fmaf(c2x2, swy, A.b2 * swx);         // sum w (2 c2 dy + b2 dx)
            v[10] = abx;
            v[11] = aby;
            v[12] = sw;
            const float red = warp_reduce_rows(v, sm.Rd, lane);
            const float val = red * (my_is_opacity ? -exp2f(-B.lo) : my_scale);
            if (my_base && val != 0.f) atomicAdd(my_base + (int64_t)sm.G[t] * my_stride, val);
        }
        __syncwarp();
    }
}

// Resident compositing warps per SM (0 = as many as fit, i.e. 16 at 128 registers).  The compositing kernels fill
// the register file; a smaller number leaves registers to the HBM-bound kernels that pipeline.ViewPipeline runs
// BESIDE them on its memory stream (projection, binning, SH of the neighbouring views).  Implemented as a dynamic
// shared-memory request that makes only that many one-warp CTAs fit an SM.
int g_warps_per_sm = 0;
size_t occupancy_pad(size_t static_smem) {
    if (g_warps_per_sm <= 0 || WARPS_PER_CTA != 1) return 0;
    const size_t per_cta = (size_t)(228 * 1024) / (size_t)g_warps_per_sm;   // 1 KB of each CTA is the system's
    const size_t want = (per_cta > static_smem + 1024 + 128) ? per_cta - static_smem - 1024 : 0;
    // static + dynamic stay below the 48 KB a launch may use without opting in (so fewer than 5 warps per SM
    // cannot be enforced this way: the request saturates)
    const size_t cap = (size_t)(48 * 1024) - static_smem - 128;
    return std::min<size_t>(want, cap) / 128 * 128;
}

template <int CH>
int launch_fwd(int C, int64_t n_isects, int width, int height, const float* means2d, const float* conics,
               const float* colors, const float* opacities, const float* backgrounds,
               const int32_t* isect_offsets, const int32_t* flatten_ids, int max_hit, float* render_colors,
               float* render_alphas, int32_t* last_ids, float* final_T, unsigned long long* stats,
               float* depth_norm, const int64_t* n_isects_dev, cudaStream_t st) {
    const int tile_w = (width + TILE - 1) / TILE, tile_h = (height + TILE - 1) / TILE;
    const unsigned grid = (unsigned)sc_div_up((int64_t)C * tile_w * tile_h, WARPS_PER_CTA);
    const size_t pad = occupancy_pad(sizeof(WarpSmem<CH, false>) * WARPS_PER_CTA);
#define FWD_LAUNCH(ST, MH)                                                                              \
    rasterize_fwd_kernel<CH, ST, MH><<<grid, WARPS_PER_CTA * 32, pad, st>>>(                            \
        n_isects, C, width, height, tile_w, tile_h, reinterpret_cast<const float2*>(means2d), conics,    \
        colors, opacities, backgrounds, isect_offsets, flatten_ids, max_hit, render_colors,             \
        render_alphas, last_ids, final_T, stats, depth_norm, n_isects_dev)
    if (stats) {
        if (max_hit > 0) FWD_LAUNCH(true, true); else FWD_LAUNCH(true, false);
    } else {
        if (max_hit > 0) FWD_LAUNCH(false, true); else FWD_LAUNCH(false, false);
    }
#undef FWD_LAUNCH
    SC_CHECK_LAUNCH("rasterize_fwd_kernel");
    return SC_OK;
}

template <int CH>
int launch_bwd(int C, int64_t n_isects, int width, int height, const float* means2d, const float* conics,
               const float* colors, const float* opacities, const float* backgrounds,
               const int32_t* isect_offsets, const int32_t* flatten_ids, const float* final_T,
               const int32_t* last_ids, const float* v_render_colors, const float* v_render_alphas,
               float* v_means2d, float* v_means2d_abs, float* v_conics, float* v_colors, float* v_opacities,
               cudaStream_t st) {
    const int tile_w = (width + TILE - 1) / TILE, tile_h = (height + TILE - 1) / TILE;
    const unsigned grid = (unsigned)sc_div_up((int64_t)C * tile_w * tile_h, WARPS_PER_CTA);
    const size_t pad = occupancy_pad(sizeof(WarpSmem<CH, true, BATCH_BWD>) * WARPS_PER_CTA);
    rasterize_bwd_kernel<CH><<<grid, WARPS_PER_CTA * 32, pad, st>>>(
        n_isects, C, width, height, tile_w, tile_h, reinterpret_cast<const float2*>(means2d), conics, colors,
        opacities, backgrounds, isect_offsets, flatten_ids, final_T, last_ids, v_render_colors, v_render_alphas,
        v_means2d, v_means2d_abs, v_conics, v_colors, v_opacities);
    SC_CHECK_LAUNCH("rasterize_bwd_kernel");
    return SC_OK;
}

}  // namespace

extern "C" int sc_rasterize_limit_warps(int warps_per_sm) {
    SC_CHECK_ARG(warps_per_sm >= 0 && warps_per_sm <= 64, "warps_per_sm");
    g_warps_per_sm = warps_per_sm;
    return SC_OK;
}

static int rasterize_fwd_entry(int C, int64_t N, int64_t n_isects, int CH, int width, int height,
                                int tile_size, const float* means2d, const float* conics, const float* colors,
                                const float* opacities, const float* backgrounds, const int32_t* isect_offsets,
                                const int32_t* flatten_ids, int max_hit, float* render_colors,
                                float* render_alphas, int32_t* last_ids, float* final_T, uint64_t* stats,
                                float* depth_norm, const int64_t* n_isects_dev, void* stream) {
    SC_CHECK_ARG(C >= 1 && N >= 0 && n_isects >= 0 && width > 0 && height > 0, "sizes");
    SC_CHECK_ARG(n_isects < (1ll << 31), "n_isects must fit int32");
    if (tile_size != TILE) {
        sc_set_error("tile_size %d unsupported (this build composites 16x16 tiles)", tile_size);
        return SC_ERR_UNSUPPORTED;
    }
    SC_CHECK_ARG(isect_offsets && render_colors && render_alphas, "null pointer");
    SC_CHECK_ARG(n_isects == 0 || (means2d && conics && colors && opacities && flatten_ids), "null pointer");
    SC_CHECK_ARG(CH != 4 || ((uintptr_t)render_colors & 15) == 0, "16-byte alignment of render_colors");
    SC_CHECK_ARG(depth_norm == nullptr || CH == 4, "depth_norm needs the RGB+depth layout (CH = 4)");
    cudaStream_t st = (cudaStream_t)stream;
#define ARGS C, n_isects, width, height, means2d, conics, colors, opacities, backgrounds, isect_offsets, \
             flatten_ids, max_hit, render_colors, render_alphas, last_ids, final_T,                       \
             reinterpret_cast<unsigned long long*>(stats), depth_norm, n_isects_dev, st
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

extern "C" int sc_rasterize_fwd(int C, int64_t N, int64_t n_isects, int CH, int width, int height,
                                int tile_size, const float* means2d, const float* conics, const float* colors,
                                const float* opacities, const float* backgrounds, const int32_t* isect_offsets,
                                const int32_t* flatten_ids, int max_hit, float* render_colors,
                                float* render_alphas, int32_t* last_ids, float* final_T, uint64_t* stats,
                                void* stream) {
    return rasterize_fwd_entry(C, N, n_isects, CH, width, height, tile_size, means2d, conics, colors, opacities,
                               backgrounds, isect_offsets, flatten_ids, max_hit, render_colors, render_alphas,
                               last_ids, final_T, stats, nullptr, nullptr, stream);
}

extern "C" int sc_rasterize_fwd_depth(int C, int64_t N, int64_t n_isects, int width, int height, int tile_size,
                                      const float* means2d, const float* conics, const float* colors4,
                                      const float* opacities, const int32_t* isect_offsets,
                                      const int32_t* flatten_ids, float* render_colors, float* render_alphas,
                                      float* depth_norm, int32_t* last_ids, float* final_T, uint64_t* stats,
                                      void* stream) {
    SC_CHECK_ARG(depth_norm, "null pointer");
    return rasterize_fwd_entry(C, N, n_isects, 4, width, height, tile_size, means2d, conics, colors4, opacities,
                               nullptr, isect_offsets, flatten_ids, 0, render_colors, render_alphas, last_ids,
                               final_T, stats, depth_norm, nullptr, stream);
}

// sc_rasterize_fwd_depth behind the sync-free binning (sc_bin_finish_dev): `capacity` = length of flatten_ids,
// the actual list length is read from the device (`n_isects_dev`, the second count of the prepare workspace)
extern "C" int sc_rasterize_fwd_depth_dev(int C, int64_t N, int64_t capacity, const int64_t* n_isects_dev, int width,
                                          int height, int tile_size, const float* means2d, const float* conics,
                                          const float* colors4, const float* opacities, const int32_t* isect_offsets,
                                          const int32_t* flatten_ids, float* render_colors, float* render_alphas,
                                          float* depth_norm, int32_t* last_ids, float* final_T, uint64_t* stats,
                                          void* stream) {
    SC_CHECK_ARG(depth_norm && n_isects_dev && capacity >= 1, "null pointer / capacity");
    return rasterize_fwd_entry(C, N, capacity, 4, width, height, tile_size, means2d, conics, colors4, opacities,
                               nullptr, isect_offsets, flatten_ids, 0, render_colors, render_alphas, last_ids,
                               final_T, stats, depth_norm, n_isects_dev, stream);
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
                     last_ids && v_render_colors && v_means2d && v_conics && v_colors &&
                     v_opacities,
                 "null pointer");
    cudaStream_t st = (cudaStream_t)stream;
#define ARGS C, n_isects, width, height, means2d, conics, colors, opacities, backgrounds, isect_offsets, \
             flatten_ids, final_T, last_ids, v_render_colors, v_render_alphas, v_means2d, v_means2d_abs, \
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
