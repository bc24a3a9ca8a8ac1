// Order-preserving compaction of one array with a RAGGED row width (w floats, w not a multiple of 4: the
// reference's xyz 3, f_dc 3, f_rest 45, opacity 1), staged through shared memory — shared by compact_rows_kernel
// and the `stay` half of densify_move_kernel.
//
// A CTA owns R consecutive rows (R a multiple of 4, R * w <= ROWMOVE_MAX_FLOATS): their source span is one
// contiguous, 16-byte aligned run of the flattened array and is read with float4 loads; the rows that stay are
// the ones whose exclusive prefix `dest` steps (dest[row + 1] - dest[row] == 1 — no flag array is read), their
// output span is again one contiguous run, written with float4 stores after a scalar head up to the first
// 16-byte boundary.  One 4-byte access per lane (the previous form) left these kernels LSU-issue bound at
// ~80 % SM throughput and 30-37 % of the DRAM peak.
#pragma once
#include "common.cuh"

constexpr int ROWMOVE_MAX_FLOATS = 3072;   // 12 KB of staged rows
constexpr int ROWMOVE_MAX_ROWS = 2048;     // + 4 KB of kept-row indices

static inline int rowmove_rows_per_cta(int w) {          // 0: row too wide for the staged form
    int r = (ROWMOVE_MAX_FLOATS / w) & ~3;
    return r > ROWMOVE_MAX_ROWS ? ROWMOVE_MAX_ROWS : r;
}

#ifdef __CUDACC__
__device__ __forceinline__ void rowmove_staged_cta(int64_t N, const int64_t* __restrict__ dest,
                                                   const float* __restrict__ src, float* __restrict__ dst, int w,
                                                   int R, int64_t cta) {
    __shared__ __align__(16) float buf[ROWMOVE_MAX_FLOATS];
    __shared__ unsigned short kept[ROWMOVE_MAX_ROWS];
    const int64_t row0 = cta * R;
    if (row0 >= N) return;
    const int nrows = (int)min((int64_t)R, N - row0);
    const int nf = nrows * w;
    const float* s = src + row0 * w;                       // (row0 * w) % 4 == 0
    const int tid = threadIdx.x, nt = blockDim.x;
    if ((((uintptr_t)src) & 15) == 0) {
        const int nv = nf >> 2;
        for (int i = tid; i < nv; i += nt) reinterpret_cast<float4*>(buf)[i] = __ldg(reinterpret_cast<const float4*>(s) + i);
        for (int i = (nv << 2) + tid; i < nf; i += nt) buf[i] = __ldg(s + i);
    } else {
        for (int i = tid; i < nf; i += nt) buf[i] = __ldg(s + i);
    }
    const int64_t d0 = __ldg(dest + row0);
    for (int r = tid; r < nrows; r += nt) {
        const int64_t d = __ldg(dest + row0 + r);
        if (__ldg(dest + row0 + r + 1) != d) kept[d - d0] = (unsigned short)r;
    }
    const int nk = (int)(__ldg(dest + row0 + nrows) - d0);
    __syncthreads();
    const int no = nk * w;
    float* o = dst + d0 * w;
    int head = (int)((16 - (((uintptr_t)o) & 15)) & 15) >> 2;       // floats up to the first 16-byte boundary
    if ((((uintptr_t)dst) & 3) != 0 || head > no) head = no;
    for (int e = tid; e < head; e += nt) {
        const int orow = e / w;
        o[e] = buf[kept[orow] * w + (e - orow * w)];
    }
    const int nv = (no - head) >> 2;
    for (int g = tid; g < nv; g += nt) {
        const int e0 = head + (g << 2);
        int orow = e0 / w, c = e0 - orow * w;
        int base = kept[orow] * w;
        float v[4];
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            v[k] = buf[base + c];
            if (++c == w && k < 3) { c = 0; ++orow; base = (orow < nk) ? kept[orow] * w : 0; }
        }
        *reinterpret_cast<float4*>(o + e0) = make_float4(v[0], v[1], v[2], v[3]);
    }
    for (int e = head + (nv << 2) + tid; e < no; e += nt) {
        const int orow = e / w;
        o[e] = buf[kept[orow] * w + (e - orow * w)];
    }
}
#endif
