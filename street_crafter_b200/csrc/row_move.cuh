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

#ifndef SC_ROWMOVE_BULK
#define SC_ROWMOVE_BULK 1   // 1: the CTA's source span arrives and its output span leaves through ONE cp.async.bulk
                            // each (TMA, SASS UBLKCP): see rowmove_staged_cta_bulk and the measurement in DESIGN.md
#endif

#ifdef __CUDACC__
__device__ __forceinline__ uint32_t rowmove_smem_addr(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

// TMA form: thread 0 issues one bulk copy global -> shared for the (16-byte aligned, 16-byte multiple) body of the
// source span and arms an mbarrier with its byte count; all threads compact the kept rows shared -> shared into a
// second buffer laid out so that the first 16-byte aligned OUTPUT element sits on a 16-byte boundary of shared
// memory; thread 0 issues one bulk copy shared -> global for the aligned body of the output span; the scalar head /
// tail (<= 3 floats each) are written by lanes.  Global-memory instructions per CTA: 2 bulk copies + <= 6 scalar
// stores + the prefix reads, instead of ~768 LDG.128 + ~768 STG.128.
__device__ __forceinline__ void rowmove_staged_cta_bulk(int64_t N, const int64_t* __restrict__ dest,
                                                        const float* __restrict__ src, float* __restrict__ dst, int w,
                                                        int R, int64_t cta) {
    __shared__ __align__(128) float buf[ROWMOVE_MAX_FLOATS];
    __shared__ __align__(128) float obuf[ROWMOVE_MAX_FLOATS + 4];
    __shared__ unsigned short kept[ROWMOVE_MAX_ROWS];
    __shared__ __align__(8) unsigned long long mbar;
    const int64_t row0 = cta * R;
    if (row0 >= N) return;
    const int nrows = (int)min((int64_t)R, N - row0);
    const int nf = nrows * w;
    const float* s = src + row0 * w;                       // (row0 * w) % 4 == 0
    const int tid = threadIdx.x, nt = blockDim.x;
    const bool aligned = ((((uintptr_t)src) | ((uintptr_t)dst)) & 15) == 0;
    const int nbody = aligned ? (nf & ~3) : 0;             // floats that travel in the bulk load
    if (tid == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(rowmove_smem_addr(&mbar)));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    if (tid == 0 && nbody > 0) {
        const uint32_t bytes = (uint32_t)nbody * 4u;
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(rowmove_smem_addr(&mbar)), "r"(bytes)
                     : "memory");
        asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                         rowmove_smem_addr(buf)),
                     "l"(s), "r"(bytes), "r"(rowmove_smem_addr(&mbar))
                     : "memory");
    }
    for (int i = nbody + tid; i < nf; i += nt) buf[i] = __ldg(s + i);      // tail (or everything when misaligned)
    // the kept-row table is built while the bulk copy is in flight
    const int64_t d0 = __ldg(dest + row0);
    for (int r = tid; r < nrows; r += nt) {
        const int64_t d = __ldg(dest + row0 + r);
        if (__ldg(dest + row0 + r + 1) != d) kept[d - d0] = (unsigned short)r;
    }
    const int nk = (int)(__ldg(dest + row0 + nrows) - d0);
    if (nbody > 0) {
        uint32_t done = 0;
        while (!done) {
            asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                         : "=r"(done)
                         : "r"(rowmove_smem_addr(&mbar)), "r"(0u)
                         : "memory");
        }
    }
    __syncthreads();
    const int no = nk * w;
    float* o = dst + d0 * w;
    int head = (int)((16 - (((uintptr_t)o) & 15)) & 15) >> 2;       // floats up to the first 16-byte boundary
    if (!aligned || head > no) head = no;
    const int pad = (4 - head) & 3;                                  // obuf[pad + head] is 16-byte aligned
    for (int e = tid; e < no; e += nt) {
        const int orow = e / w;
        obuf[pad + e] = buf[kept[orow] * w + (e - orow * w)];
    }
    // make the generic-proxy writes to shared memory visible to the async proxy, then one bulk store
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    __syncthreads();
    const int nvb = (no - head) & ~3;                                // floats in the bulk store
    if (tid == 0 && nvb > 0) {
        asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(o + head),
                     "r"(rowmove_smem_addr(obuf + pad + head)), "r"((uint32_t)nvb * 4u)
                     : "memory");
        asm volatile("cp.async.bulk.commit_group;" ::: "memory");
    }
    for (int e = tid; e < head; e += nt) o[e] = obuf[pad + e];
    for (int e = head + nvb + tid; e < no; e += nt) o[e] = obuf[pad + e];
    if (tid == 0 && nvb > 0) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");   // obuf must outlive the read
}

__device__ __forceinline__ void rowmove_staged_cta_ldst(int64_t N, const int64_t* __restrict__ dest,
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

__device__ __forceinline__ void rowmove_staged_cta(int64_t N, const int64_t* __restrict__ dest,
                                                   const float* __restrict__ src, float* __restrict__ dst, int w, int R,
                                                   int64_t cta) {
#if SC_ROWMOVE_BULK
    rowmove_staged_cta_bulk(N, dest, src, dst, w, R, cta);
#else
    rowmove_staged_cta_ldst(N, dest, src, dst, w, R, cta);
#endif
}
#endif
