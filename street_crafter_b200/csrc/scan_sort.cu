// Hand-written device scan and stable LSD radix sort (the cumsum and cub::DeviceRadixSort
// inside gsplat.isect_tiles, /root/reference/street_gaussian/models/street_gaussian_renderer.py:243-252).
//
// Sort design (HBM-bound, integer work; no tensor cores):
//   * one histogram kernel reads the keys once and builds the digit histograms of ALL passes;
//   * each 8-bit pass is ONE kernel ("onesweep"): a tile of 4096 pairs is ranked stably in
//     shared memory (warp match.any multi-split), tile prefixes are resolved with a chained
//     decoupled look-back, and pairs leave through shared memory so that every digit run is
//     written with consecutive, coalesced stores.  Traffic per pass = 12 B read + 12 B write
//     per pair.  Tiles are handed out by an atomic ticket so a waiting tile's predecessors are
//     always resident (forward progress).
//   * only the significant key bits are sorted: 32 depth bits + tile bits (+ camera bits).
#include "common.cuh"
#include "sort_internal.cuh"

namespace {

// ------------------------------------------------------------------------------- scan
constexpr int SCAN_THREADS = 256;
constexpr int SCAN_ITEMS = 8;
constexpr int SCAN_TILE = SCAN_THREADS * SCAN_ITEMS;

__device__ __forceinline__ int64_t block_exclusive_scan_i64(int64_t v, int64_t* s_warp, int64_t& block_total) {
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    int64_t incl = v;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const int64_t t = __shfl_up_sync(0xffffffffu, incl, d);
        if (lane >= d) incl += t;
    }
    if (lane == 31) s_warp[w] = incl;
    __syncthreads();
    int64_t wsum = 0, tot = 0;
    const int nw = blockDim.x >> 5;
    for (int i = 0; i < nw; ++i) {
        const int64_t t = s_warp[i];
        if (i < w) wsum += t;
        tot += t;
    }
    block_total = tot;
    __syncthreads();
    return wsum + incl - v;
}

__global__ void __launch_bounds__(SCAN_THREADS)
scan_block_sums_kernel(int64_t n, const int32_t* __restrict__ in, int64_t* __restrict__ block_sums) {
    __shared__ int64_t s_warp[SCAN_THREADS / 32];
    const int64_t base = (int64_t)blockIdx.x * SCAN_TILE;
    int64_t s = 0;
#pragma unroll
    for (int i = 0; i < SCAN_ITEMS; ++i) {
        const int64_t idx = base + i * SCAN_THREADS + threadIdx.x;
        if (idx < n) s += __ldg(in + idx);
    }
    int64_t tot;
    block_exclusive_scan_i64(s, s_warp, tot);
    if (threadIdx.x == 0) block_sums[blockIdx.x] = tot;
}

__global__ void __launch_bounds__(1024)
scan_spine_kernel(int64_t nb, int64_t* __restrict__ block_sums, int64_t* __restrict__ total) {
    __shared__ int64_t s_warp[32];
    int64_t carry = 0;
    for (int64_t base = 0; base < nb; base += 1024) {
        const int64_t idx = base + threadIdx.x;
        const int64_t v = idx < nb ? block_sums[idx] : 0;
        int64_t tot;
        const int64_t ex = block_exclusive_scan_i64(v, s_warp, tot);
        if (idx < nb) block_sums[idx] = carry + ex;
        carry += tot;
    }
    if (threadIdx.x == 0 && total) *total = carry;
}

__global__ void __launch_bounds__(SCAN_THREADS)
scan_apply_kernel(int64_t n, const int32_t* __restrict__ in, const int64_t* __restrict__ block_sums,
                  int64_t* __restrict__ out) {
    __shared__ int64_t s_warp[SCAN_THREADS / 32];
    // blocked arrangement per thread: thread t owns SCAN_ITEMS consecutive elements
    const int64_t base = (int64_t)blockIdx.x * SCAN_TILE + (int64_t)threadIdx.x * SCAN_ITEMS;
    int32_t v[SCAN_ITEMS];
    int64_t s = 0;
#pragma unroll
    for (int i = 0; i < SCAN_ITEMS; ++i) {
        v[i] = (base + i < n) ? __ldg(in + base + i) : 0;
        s += v[i];
    }
    int64_t tot;
    int64_t ex = block_exclusive_scan_i64(s, s_warp, tot) + block_sums[blockIdx.x];
#pragma unroll
    for (int i = 0; i < SCAN_ITEMS; ++i) {
        if (base + i < n) out[base + i] = ex;
        ex += v[i];
    }
}

// ------------------------------------------------------------------------------- radix sort
constexpr int RADIX_BITS = 8;
constexpr int RADIX = SC_SORT_RADIX;
static_assert(RADIX == (1 << RADIX_BITS), "radix");
constexpr int SORT_THREADS = 256;
constexpr int SORT_WARPS = SORT_THREADS / 32;
// keys per thread (tile = 256 * KPT pairs), per key width; overridable for tools/sort_variants.sh.
// Small tiles = more resident CTAs and a shorter per-tile critical path (the ranking loop is
// KPT dependent steps): the passes are latency-bound, not bandwidth-bound.
#ifndef SC_SORT_KPT_U32
#define SC_SORT_KPT_U32 8
#endif
#ifndef SC_SORT_KPT_U64
#define SC_SORT_KPT_U64 16
#endif
#ifndef SC_SORT_MINB_U32
#define SC_SORT_MINB_U32 3  // resident CTAs per SM the register budget is capped for (no spills beats
                            // more warps: tools/sort_variants.sh)
#endif
#ifndef SC_SORT_MINB_U64
#define SC_SORT_MINB_U64 2
#endif
template <typename KeyT> struct SortCfg {
    static constexpr int KPT = sizeof(KeyT) == 4 ? SC_SORT_KPT_U32 : SC_SORT_KPT_U64;
    static constexpr int MINB = sizeof(KeyT) == 4 ? SC_SORT_MINB_U32 : SC_SORT_MINB_U64;
};
#ifndef SC_SORT_LB_W
#define SC_SORT_LB_W 8
#endif
constexpr int LB_W = SC_SORT_LB_W;  // look-back window: predecessors read per round trip
constexpr int MAX_PASSES = SC_SORT_MAX_PASSES;
constexpr uint32_t FLAG_AGG = 1u << 30;
constexpr uint32_t FLAG_INC = 2u << 30;
constexpr uint32_t FLAG_MASK = 3u << 30;
constexpr uint32_t VAL_MASK = ~FLAG_MASK;

template <typename KeyT>
struct SortSmem {
    static constexpr int SORT_TILE = SORT_THREADS * SortCfg<KeyT>::KPT;
    KeyT keys[SORT_TILE];
    uint32_t vals[SORT_TILE];
    uint32_t warp_hist[SORT_WARPS][RADIX];
    uint32_t digit_start[RADIX];   // exclusive start of each digit inside the tile
    int64_t gbase[RADIX];          // global position of local slot 0 of each digit run
    uint32_t scan_tmp[SORT_WARPS];
    uint32_t tile_id;
};

using PassPlan = ScPassPlan;

template <typename KeyT>
__global__ void __launch_bounds__(256)
sort_histogram_kernel(int64_t n, const int64_t* __restrict__ n_dev, const KeyT* __restrict__ keys, PassPlan pp,
                      uint32_t* __restrict__ hist /*[passes][RADIX]*/) {
    extern __shared__ uint32_t s_hist[];  // [passes][RADIX]
    if (n_dev) n = min(n, *n_dev);
    for (int i = threadIdx.x; i < pp.passes * RADIX; i += blockDim.x) s_hist[i] = 0;
    __syncthreads();
    // warp-aggregated shared atomics: the keys of one warp share most high digits (same tile,
    // similar depth), so un-aggregated atomics would serialise 32-deep on one bank.
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    const int lane = threadIdx.x & 31;
    const uint32_t lt_mask = (1u << lane) - 1u;
    const int64_t n_round = (n + 31) & ~(int64_t)31;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n_round; i += stride) {
        const bool in = i < n;
        const KeyT k = in ? __ldg(keys + i) : (KeyT)0;
        const unsigned act = __ballot_sync(0xffffffffu, in);
        for (int p = 0; p < pp.passes; ++p) {
            const uint32_t d = (uint32_t)(k >> pp.shift[p]) & ((1u << pp.nbits[p]) - 1u);
            const unsigned peers = __match_any_sync(0xffffffffu, in ? d : 0xffffffffu) & act;
            if (in && (peers & lt_mask) == 0) atomicAdd(&s_hist[p * RADIX + d], (uint32_t)__popc(peers));
        }
    }
    __syncthreads();
    for (int i = threadIdx.x; i < pp.passes * RADIX; i += blockDim.x) {
        const uint32_t v = s_hist[i];
        if (v) atomicAdd(&hist[i], v);
    }
}

__global__ void __launch_bounds__(RADIX)
sort_hist_scan_kernel(uint32_t* __restrict__ hist) {
    // exclusive scan of each pass's 256 bins, in place
    __shared__ uint32_t s[RADIX];
    uint32_t* h = hist + blockIdx.x * RADIX;
    s[threadIdx.x] = h[threadIdx.x];
    __syncthreads();
    if (threadIdx.x == 0) {
        uint32_t acc = 0;
        for (int i = 0; i < RADIX; ++i) {
            const uint32_t v = s[i];
            s[i] = acc;
            acc += v;
        }
    }
    __syncthreads();
    h[threadIdx.x] = s[threadIdx.x];
}

template <typename KeyT>
__global__ void __launch_bounds__(SORT_THREADS, SortCfg<KeyT>::MINB)
onesweep_pass_kernel(int64_t n, const int64_t* __restrict__ n_dev, const KeyT* __restrict__ keys_in,
                     const uint32_t* __restrict__ vals_in,
                     KeyT* __restrict__ keys_out, uint32_t* __restrict__ vals_out, int shift, int nbits,
                     const uint32_t* __restrict__ digit_offset /*[RADIX] exclusive, or raw counts*/,
                     volatile uint32_t* __restrict__ lookback /*[tiles][RADIX]*/,
                     uint32_t* __restrict__ ticket, int hist_is_raw) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    SortSmem<KeyT>& sm = *reinterpret_cast<SortSmem<KeyT>*>(smem_raw);
    constexpr int SORT_KPT = SortCfg<KeyT>::KPT;
    constexpr int SORT_TILE = SortSmem<KeyT>::SORT_TILE;
    const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
    const uint32_t dmask = (1u << nbits) - 1u;

    if (n_dev) n = min(n, *n_dev);
    if (tid == 0) sm.tile_id = atomicAdd(ticket, 1u);
    for (int i = tid; i < SORT_WARPS * RADIX; i += SORT_THREADS) (&sm.warp_hist[0][0])[i] = 0;
    __syncthreads();
    const uint32_t tile = sm.tile_id;
    const int64_t tile_base = (int64_t)tile * SORT_TILE;
    if (tile_base >= n) return;  // grid is sized for the maximum count (device-resident n)
    const int n_valid = (int)min((int64_t)SORT_TILE, n - tile_base);

    // ---- load keys (warp-striped: item i of lane l = warp_base + i*32 + l)
    KeyT key[SORT_KPT];
    uint16_t rank[SORT_KPT];
    const int warp_base = w * (32 * SORT_KPT);
#pragma unroll
    for (int i = 0; i < SORT_KPT; ++i) {
        const int local = warp_base + i * 32 + lane;
        key[i] = (local < n_valid) ? __ldg(keys_in + tile_base + local) : (KeyT)~(KeyT)0;
    }
    // ---- stable warp-level ranking: the peer masks of all keys first (independent ballots that
    // pipeline), then the short serial chain through the warp histogram
    const uint32_t lt_mask = (1u << lane) - 1u;
    uint32_t peers_of[SORT_KPT];
#pragma unroll
    for (int i = 0; i < SORT_KPT; ++i) peers_of[i] = sc_match_digit8((uint32_t)(key[i] >> shift) & dmask);
#pragma unroll
    for (int i = 0; i < SORT_KPT; ++i) {
        const uint32_t d = (uint32_t)(key[i] >> shift) & dmask;
        const uint32_t peers = peers_of[i];
        const uint32_t pre = sm.warp_hist[w][d];
        rank[i] = (uint16_t)(pre + __popc(peers & lt_mask));
        __syncwarp();
        if ((peers & lt_mask) == 0) sm.warp_hist[w][d] = pre + __popc(peers);
        __syncwarp();
    }
    // values: issued now so that their DRAM latency hides under the look-back below
    uint32_t val[SORT_KPT];
#pragma unroll
    for (int i = 0; i < SORT_KPT; ++i) {
        const int local = warp_base + i * 32 + lane;
        val[i] = (local < n_valid) ? __ldg(vals_in + tile_base + local) : 0u;
    }
    __syncthreads();

    // ---- per digit (thread == digit): warp-exclusive offsets, tile count, look-back
    {
        const int d = tid;
        uint32_t acc = 0;
#pragma unroll
        for (int ww = 0; ww < SORT_WARPS; ++ww) {
            const uint32_t c = sm.warp_hist[ww][d];
            sm.warp_hist[ww][d] = acc;
            acc += c;
        }
        const uint32_t count = acc;
        // publish aggregate, then resolve exclusive prefix over previous tiles.  The walk reads
        // LB_W predecessors per round trip (independent loads), consumes them in order and
        // re-polls from the first one that is not published yet.  When all tiles of a small sort
        // are resident at once, tile t resolves after ~t / (2 LB_W) round trips: the width of the
        // window, not bandwidth, sets the duration of such a pass.
        volatile uint32_t* my = lookback + (size_t)tile * RADIX + d;
        uint32_t excl = 0;
        if (tile == 0) {
            *my = count | FLAG_INC;
        } else {
            *my = count | FLAG_AGG;
            int64_t t = (int64_t)tile - 1;
            bool done = false;
            while (!done) {
                uint32_t v[LB_W];
#pragma unroll
                for (int u = 0; u < LB_W; ++u) v[u] = (t - u >= 0) ? lookback[(size_t)(t - u) * RADIX + d] : (2u << 30);
                int used = 0;
#pragma unroll
                for (int u = 0; u < LB_W; ++u) {
                    const uint32_t f = v[u] & FLAG_MASK;
                    if (!done && used == u && f != 0) {
                        excl += v[u] & VAL_MASK;
                        used = u + 1;
                        done = f == FLAG_INC;
                    }
                }
                t -= used;
            }
            *my = (excl + count) | FLAG_INC;
        }
        // block exclusive scan of counts over digits
        uint32_t incl = count;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t t2 = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o) incl += t2;
        }
        if (lane == 31) sm.scan_tmp[w] = incl;
        __syncthreads();
        uint32_t wsum = 0;
#pragma unroll
        for (int ww = 0; ww < SORT_WARPS; ++ww)
            if (ww < w) wsum += sm.scan_tmp[ww];
        const uint32_t dstart = wsum + incl - count;
        sm.digit_start[d] = dstart;
        uint32_t goff = digit_offset[d];
        if (hist_is_raw) {
            // the caller handed the raw digit counts: exclusive scan over the digits here (saves the
            // separate scan launch of the fused binning pipeline)
            uint32_t hin = goff;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const uint32_t t2 = __shfl_up_sync(0xffffffffu, hin, o);
                if (lane >= o) hin += t2;
            }
            __syncthreads();  // scan_tmp is re-used
            if (lane == 31) sm.scan_tmp[w] = hin;
            __syncthreads();
            uint32_t hsum = 0;
#pragma unroll
            for (int ww = 0; ww < SORT_WARPS; ++ww)
                if (ww < w) hsum += sm.scan_tmp[ww];
            goff = hsum + hin - goff;
        }
        sm.gbase[d] = (int64_t)goff + (int64_t)excl - (int64_t)dstart;
    }
    __syncthreads();

    // ---- scatter into shared memory in sorted order
#pragma unroll
    for (int i = 0; i < SORT_KPT; ++i) {
        const uint32_t d = (uint32_t)(key[i] >> shift) & dmask;
        const uint32_t pos = sm.digit_start[d] + sm.warp_hist[w][d] + rank[i];
        rank[i] = (uint16_t)pos;
        sm.keys[pos] = key[i];
    }
#pragma unroll
    for (int i = 0; i < SORT_KPT; ++i) {
        const int local = warp_base + i * 32 + lane;
        if (local < n_valid) sm.vals[rank[i]] = val[i];
    }
    __syncthreads();
    // ---- coalesced write-out: consecutive threads -> consecutive slots of a digit run
#pragma unroll
    for (int i = 0; i < SORT_KPT; ++i) {
        const int local = i * SORT_THREADS + tid;
        if (local < n_valid) {
            const KeyT k = sm.keys[local];
            const uint32_t d = (uint32_t)(k >> shift) & dmask;
            const int64_t g = sm.gbase[d] + local;
            keys_out[g] = k;
            vals_out[g] = sm.vals[local];
        }
    }
}

template <typename KeyT>
int sort_histogram(int64_t n_max, const int64_t* n_dev, const KeyT* keys, const ScSortPlan& plan, void* ws,
                   cudaStream_t st) {
    if (n_max == 0) return SC_OK;
    uint32_t* hist = (uint32_t*)((unsigned char*)ws + plan.hist_off);
    const int grid = (int)min(sc_div_up(n_max, 256 * 8), (int64_t)SC_NUM_SMS * 8);
    sort_histogram_kernel<KeyT><<<grid, 256, plan.pp.passes * RADIX * 4, st>>>(n_max, n_dev, keys, plan.pp, hist);
    SC_CHECK_LAUNCH("sort_histogram_kernel");
    return SC_OK;
}

// returns 0 if the sorted pairs are in keys_out/vals_out, 1 if they are in keys_in/vals_in
// (even number of passes), negative on error.
template <typename KeyT>
int sort_run(int64_t n_max, const int64_t* n_dev, KeyT* keys_in, uint32_t* vals_in, KeyT* keys_out,
             uint32_t* vals_out, const ScSortPlan& plan, void* workspace, cudaStream_t st, bool hist_raw = false) {
    if (n_max == 0) return 0;
    unsigned char* ws = (unsigned char*)workspace;
    uint32_t* hist = (uint32_t*)(ws + plan.hist_off);
    uint32_t* tickets = (uint32_t*)(ws + plan.ticket_off);
    uint32_t* lookback = (uint32_t*)(ws + plan.lookback_off);
    const int passes = plan.pp.passes;
    if (!hist_raw) {
        sort_hist_scan_kernel<<<passes, RADIX, 0, st>>>(hist);
        SC_CHECK_LAUNCH("sort_hist_scan_kernel");
    }
    static bool attr_set = false;  // one flag per KeyT instantiation
    if (!attr_set) {
        SC_CUDA(cudaFuncSetAttribute(onesweep_pass_kernel<KeyT>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                     (int)sizeof(SortSmem<KeyT>)));
        attr_set = true;
    }
    KeyT* kin = keys_in;
    uint32_t* vin = vals_in;
    KeyT* kout = keys_out;
    uint32_t* vout = vals_out;
    for (int p = 0; p < passes; ++p) {
        onesweep_pass_kernel<KeyT><<<(unsigned)plan.tiles, SORT_THREADS, sizeof(SortSmem<KeyT>), st>>>(
            n_max, n_dev, kin, vin, kout, vout, plan.pp.shift[p], plan.pp.nbits[p], hist + p * RADIX,
            lookback + (size_t)p * plan.tiles * RADIX, tickets + p, hist_raw ? 1 : 0);
        SC_CHECK_LAUNCH("onesweep_pass_kernel");
        KeyT* tk = kin; kin = kout; kout = tk;
        uint32_t* tv = vin; vin = vout; vout = tv;
    }
    // after the swap `kin` holds the result
    return kin == keys_out ? 0 : 1;
}

template <typename KeyT>
int radix_sort_pairs(int64_t n, KeyT* keys_in, uint32_t* vals_in, KeyT* keys_out, uint32_t* vals_out,
                     int begin_bit, int end_bit, void* workspace, int64_t workspace_bytes, void* stream) {
    SC_CHECK_ARG(n >= 0 && n < (1ll << 30), "n (must be < 2^30)");
    SC_CHECK_ARG(begin_bit >= 0 && end_bit > begin_bit && end_bit <= (int)sizeof(KeyT) * 8, "bit range");
    if (n == 0) return 0;
    SC_CHECK_ARG(keys_in && vals_in && keys_out && vals_out && workspace, "null pointer");
    const ScSortPlan plan = sc_sort_make_plan(n, begin_bit, end_bit, (int)sizeof(KeyT));
    if (workspace_bytes < plan.total) {
        sc_set_error("sort workspace too small: %lld < %lld", (long long)workspace_bytes, (long long)plan.total);
        return SC_ERR_WORKSPACE;
    }
    cudaStream_t st = (cudaStream_t)stream;
    SC_CUDA(cudaMemsetAsync(workspace, 0, (size_t)plan.ctrl_bytes, st));
    const int rc = sort_histogram<KeyT>(n, nullptr, keys_in, plan, workspace, st);
    if (rc < 0) return rc;
    return sort_run<KeyT>(n, nullptr, keys_in, vals_in, keys_out, vals_out, plan, workspace, st);
}

}  // namespace

// passes of <= 8 bits over [begin_bit, end_bit), widths balanced (14 bits -> 7+7, not 8+6)
ScSortPlan sc_sort_make_plan(int64_t n_max, int begin_bit, int end_bit, int key_bytes) {
    ScSortPlan p;
    // key_bytes 0: the smaller tile of the two key widths (workspace upper bound)
    const int kpt = key_bytes == 4 ? SC_SORT_KPT_U32 : key_bytes == 8 ? SC_SORT_KPT_U64
                  : (SC_SORT_KPT_U32 < SC_SORT_KPT_U64 ? SC_SORT_KPT_U32 : SC_SORT_KPT_U64);
    const int64_t SORT_TILE = (int64_t)SORT_THREADS * kpt;
    const int bits = end_bit - begin_bit;
    p.pp.passes = (bits + RADIX_BITS - 1) / RADIX_BITS;
    int shift = begin_bit, left = bits;
    for (int i = 0; i < MAX_PASSES; ++i) {
        const int remaining = p.pp.passes - i;
        const int nb = remaining > 0 ? (left + remaining - 1) / remaining : 0;
        p.pp.shift[i] = shift;
        p.pp.nbits[i] = nb;
        shift += nb;
        left -= nb;
    }
    p.tiles = sc_div_up(n_max, SORT_TILE);
    p.hist_off = 0;
    p.ticket_off = sc_align_up((int64_t)MAX_PASSES * RADIX * 4, 256);
    p.lookback_off = p.ticket_off + 256;
    p.ctrl_bytes = p.lookback_off + (int64_t)p.pp.passes * p.tiles * RADIX * 4;
    p.total = sc_align_up(p.lookback_off + (int64_t)MAX_PASSES * p.tiles * RADIX * 4, 256);
    return p;
}

int sc_sort_histogram_u32(int64_t n_max, const int64_t* n_dev, const uint32_t* keys, const ScSortPlan& plan,
                          void* ws, cudaStream_t st) {
    return sort_histogram<uint32_t>(n_max, n_dev, keys, plan, ws, st);
}
int sc_sort_histogram_u64(int64_t n_max, const int64_t* n_dev, const uint64_t* keys, const ScSortPlan& plan,
                          void* ws, cudaStream_t st) {
    return sort_histogram<uint64_t>(n_max, n_dev, keys, plan, ws, st);
}
int sc_sort_run_u32(int64_t n_max, const int64_t* n_dev, uint32_t* keys_in, uint32_t* vals_in,
                    uint32_t* keys_out, uint32_t* vals_out, const ScSortPlan& plan, void* ws, cudaStream_t st,
                    bool hist_raw) {
    return sort_run<uint32_t>(n_max, n_dev, keys_in, vals_in, keys_out, vals_out, plan, ws, st, hist_raw);
}
int sc_sort_run_u64(int64_t n_max, const int64_t* n_dev, uint64_t* keys_in, uint32_t* vals_in,
                    uint64_t* keys_out, uint32_t* vals_out, const ScSortPlan& plan, void* ws, cudaStream_t st) {
    return sort_run<uint64_t>(n_max, n_dev, keys_in, vals_in, keys_out, vals_out, plan, ws, st);
}

extern "C" int64_t sc_scan_workspace_bytes(int64_t n) { return (sc_div_up(n, SCAN_TILE) + 1) * 8; }

extern "C" int sc_exclusive_scan_i32(int64_t n, const int32_t* in, int64_t* out, int64_t* total,
                                     void* workspace, int64_t workspace_bytes, void* stream) {
    SC_CHECK_ARG(n >= 0, "n");
    cudaStream_t st = (cudaStream_t)stream;
    if (n == 0) {
        if (total) SC_CUDA(cudaMemsetAsync(total, 0, 8, st));
        return SC_OK;
    }
    SC_CHECK_ARG(in && out && workspace, "null pointer");
    if (workspace_bytes < sc_scan_workspace_bytes(n)) {
        sc_set_error("scan workspace too small: %lld < %lld", (long long)workspace_bytes,
                     (long long)sc_scan_workspace_bytes(n));
        return SC_ERR_WORKSPACE;
    }
    const int64_t nb = sc_div_up(n, SCAN_TILE);
    int64_t* block_sums = (int64_t*)workspace;
    scan_block_sums_kernel<<<(unsigned)nb, SCAN_THREADS, 0, st>>>(n, in, block_sums);
    SC_CHECK_LAUNCH("scan_block_sums_kernel");
    scan_spine_kernel<<<1, 1024, 0, st>>>(nb, block_sums, total);
    SC_CHECK_LAUNCH("scan_spine_kernel");
    scan_apply_kernel<<<(unsigned)nb, SCAN_THREADS, 0, st>>>(n, in, block_sums, out);
    SC_CHECK_LAUNCH("scan_apply_kernel");
    return SC_OK;
}

// ---- allocation-free scan -------------------------------------------------------------------------
// No counterpart in the reference: added for callers that cannot
// allocate (CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).  Same three kernels
// as sc_exclusive_scan_i32; the block-sum scratch is this file-scope array instead of a caller workspace.
// Every GPU that loads the library holds the array (64 KiB + 256 B) until the process exits.
#define SC_SCAN_NOALLOC_MAX_N (1ll << 24)
constexpr int64_t SCAN_NOALLOC_MAX_BLOCKS = SC_SCAN_NOALLOC_MAX_N / SCAN_TILE;  // 8192 block sums
static __device__ __align__(256) int64_t g_scan_noalloc_scratch[SCAN_NOALLOC_MAX_BLOCKS + 32];

// In host code the array's NAME is only a handle for the runtime; the device address of the array is looked
// up once per device, with that device current, and cached.
static int scan_noalloc_scratch_ptr(int64_t** out) {
    constexpr int MAX_DEVICES = 64;
    static int64_t* ptr_of[MAX_DEVICES] = {};
    int dev = -1;
    SC_CUDA(cudaGetDevice(&dev));
    SC_CHECK_ARG(dev >= 0 && dev < MAX_DEVICES, "device ordinal beyond 63");
    int64_t* p = __atomic_load_n(&ptr_of[dev], __ATOMIC_ACQUIRE);
    if (!p) {
        void* sym = nullptr;
        const cudaError_t e = cudaGetSymbolAddress(&sym, g_scan_noalloc_scratch);
        if (e != cudaSuccess || !sym) {
            sc_set_error("cudaGetSymbolAddress(g_scan_noalloc_scratch) on device %d: %s", dev, cudaGetErrorString(e));
            return SC_ERR_CUDA;
        }
        p = static_cast<int64_t*>(sym);
        __atomic_store_n(&ptr_of[dev], p, __ATOMIC_RELEASE);  // racing first calls look up the same address
    }
    *out = p;
    return SC_OK;
}

extern "C" int64_t sc_exclusive_scan_i32_noalloc_max_n(void) { return SC_SCAN_NOALLOC_MAX_N; }

extern "C" int sc_exclusive_scan_i32_noalloc(int64_t n, const int32_t* in, int64_t* out, int64_t* total,
                                             void* stream) {
    SC_CHECK_ARG(n >= 0, "n");
    if (n > SC_SCAN_NOALLOC_MAX_N) {
        sc_set_error("sc_exclusive_scan_i32_noalloc: n = %lld exceeds the built-in scratch (at most %lld elements); "
                     "use sc_exclusive_scan_i32 with a caller workspace",
                     (long long)n, (long long)SC_SCAN_NOALLOC_MAX_N);
        return SC_ERR_WORKSPACE;
    }
    cudaStream_t st = (cudaStream_t)stream;
    if (n == 0) {
        if (total) SC_CUDA(cudaMemsetAsync(total, 0, 8, st));
        return SC_OK;
    }
    SC_CHECK_ARG(in && out, "null pointer");
    int64_t* block_sums = nullptr;
    const int rc = scan_noalloc_scratch_ptr(&block_sums);
    if (rc != SC_OK) return rc;
    // nothing in the scratch has to start at zero: the first kernel writes block_sums[0, nb) before any read
    const int64_t nb = sc_div_up(n, SCAN_TILE);
    scan_block_sums_kernel<<<(unsigned)nb, SCAN_THREADS, 0, st>>>(n, in, block_sums);
    SC_CHECK_LAUNCH("scan_block_sums_kernel");
    scan_spine_kernel<<<1, 1024, 0, st>>>(nb, block_sums, total);
    SC_CHECK_LAUNCH("scan_spine_kernel");
    scan_apply_kernel<<<(unsigned)nb, SCAN_THREADS, 0, st>>>(n, in, block_sums, out);
    SC_CHECK_LAUNCH("scan_apply_kernel");
    return SC_OK;
}

extern "C" int64_t sc_sort_workspace_bytes(int64_t n) { return sc_sort_make_plan(n, 0, 64, 0).total; }

extern "C" int sc_radix_sort_pairs(int64_t n, uint64_t* keys_in, uint32_t* vals_in, uint64_t* keys_out,
                                   uint32_t* vals_out, int begin_bit, int end_bit, void* workspace,
                                   int64_t workspace_bytes, void* stream) {
    return radix_sort_pairs<uint64_t>(n, keys_in, vals_in, keys_out, vals_out, begin_bit, end_bit, workspace,
                                      workspace_bytes, stream);
}

extern "C" int sc_radix_sort_pairs_u32(int64_t n, uint32_t* keys_in, uint32_t* vals_in, uint32_t* keys_out,
                                       uint32_t* vals_out, int begin_bit, int end_bit, void* workspace,
                                       int64_t workspace_bytes, void* stream) {
    return radix_sort_pairs<uint32_t>(n, keys_in, vals_in, keys_out, vals_out, begin_bit, end_bit, workspace,
                                      workspace_bytes, stream);
}
