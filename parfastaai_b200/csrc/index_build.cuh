// index_build.cuh -- device index build (K1-K3 of DESIGN.md): Lp scan, segment histogram, T, and the
// stable LSD radix sort that turns the tetramer-major postings F into genome-major tetramer lists.
//
// Replaces (reference, relative to /root/reference):
//   DataStructHelper::parallelPrefixSum  include/pfaai/ds_helper.hpp:112-122   -> k_scan_* (Lc -> Lp)
//   DataStructHelper::constructT         include/pfaai/ds_helper.hpp:46-79     -> k_make_items histogram + k_transpose_t
//   DataStructHelper::constructE + parallelMergeSort  ds_helper.hpp:362-421, psort.hpp:27-53
//        -> NOT materialised; the grouping the sort provided is obtained by sorting the |F| postings on
//           (genome, SCP) (k_radix_*), 40-90x fewer items than |E| and 8 B instead of 12 B each.
// All kernels are HBM-bound streaming passes; grids are sized in multiples of the SM count.
#pragma once
#include "common.cuh"

namespace pfaai {

// ------------------------------------------------------------------------------------------------
// Device-wide exclusive scan int32 -> int64 (three launches: tile sums, scan of sums, final).
// ------------------------------------------------------------------------------------------------
constexpr int SCAN_THREADS = 512;
constexpr int SCAN_ITEMS = 8;
constexpr int SCAN_TILE = SCAN_THREADS * SCAN_ITEMS;

__device__ __forceinline__ int64_t block_exclusive_scan_i64(int64_t v, int64_t *total, int64_t *scratch /*>=33*/) {
    // inclusive scan inside the warp, then across the (<=32) warps
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = (blockDim.x + 31) >> 5;
    int64_t inc = v;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        int64_t o = __shfl_up_sync(0xffffffffu, inc, d);
        if (lane >= d) inc += o;
    }
    __syncthreads();  // scratch may still be read by a previous call
    if (lane == 31) scratch[warp] = inc;
    __syncthreads();
    if (warp == 0) {
        int64_t w = (lane < nwarps) ? scratch[lane] : 0, winc = w;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            int64_t o = __shfl_up_sync(0xffffffffu, winc, d);
            if (lane >= d) winc += o;
        }
        scratch[lane] = winc - w;                 // exclusive offset of warp `lane`
        if (lane == 31) scratch[32] = winc;       // block total
    }
    __syncthreads();
    if (total) *total = scratch[32];
    return scratch[warp] + inc - v;
}

__global__ void __launch_bounds__(SCAN_THREADS) k_scan_tile_sums(const int32_t *__restrict__ in, int64_t n,
                                                                  int64_t *__restrict__ tile_sums) {
    __shared__ int64_t scratch[33];
    const int64_t base = (int64_t)blockIdx.x * SCAN_TILE + (int64_t)threadIdx.x * SCAN_ITEMS;
    int64_t s = 0;
#pragma unroll
    for (int k = 0; k < SCAN_ITEMS; k++)
        if (base + k < n) s += in[base + k];
    int64_t total;
    block_exclusive_scan_i64(s, &total, scratch);
    if (threadIdx.x == 0) tile_sums[blockIdx.x] = total;
}

// single block: exclusive scan of `n` int64 in place (n is small: #tiles)
__global__ void __launch_bounds__(1024) k_scan_small_i64(int64_t *__restrict__ a, int64_t n) {
    __shared__ int64_t scratch[33];
    const int64_t per = (n + blockDim.x - 1) / blockDim.x;
    const int64_t lo = (int64_t)threadIdx.x * per, hi = min(lo + per, n);
    int64_t s = 0;
    for (int64_t i = lo; i < hi; i++) s += a[i];
    int64_t run = block_exclusive_scan_i64(s, nullptr, scratch);
    for (int64_t i = lo; i < hi; i++) { int64_t v = a[i]; a[i] = run; run += v; }
}

// out has n+1 entries: out[n] = grand total.  out32 (optional) gets the same values truncated to int32.
__global__ void __launch_bounds__(SCAN_THREADS) k_scan_final(const int32_t *__restrict__ in, int64_t n,
                                                              const int64_t *__restrict__ tile_offs,
                                                              int64_t *__restrict__ out, int32_t *__restrict__ out32) {
    __shared__ int64_t scratch[33];
    const int64_t base = (int64_t)blockIdx.x * SCAN_TILE + (int64_t)threadIdx.x * SCAN_ITEMS;
    int32_t v[SCAN_ITEMS];
    int64_t s = 0;
#pragma unroll
    for (int k = 0; k < SCAN_ITEMS; k++) { v[k] = (base + k < n) ? in[base + k] : 0; s += v[k]; }
    int64_t run = block_exclusive_scan_i64(s, nullptr, scratch) + tile_offs[blockIdx.x];
#pragma unroll
    for (int k = 0; k < SCAN_ITEMS; k++) {
        if (base + k < n) { out[base + k] = run; if (out32) out32[base + k] = (int32_t)run; }
        run += v[k];
        if (base + k == n - 1) out[n] = run;
    }
}

// ------------------------------------------------------------------------------------------------
// K2a: postings -> sort items.  item = ((genome*P + protein) << 18) | tetramer, where the tetramer of
// posting i is the t with Lp[t] <= i < Lp[t+1].  Also histograms the (genome, SCP) segment lengths
// (K3: this IS the T matrix in the single-DB modes) with one REDG per posting.
// ------------------------------------------------------------------------------------------------
constexpr int MI_THREADS = 256;
constexpr int MI_ITEMS = 8;

__global__ void __launch_bounds__(MI_THREADS) k_make_items(const pfaai_pg_t *__restrict__ F, const int64_t *__restrict__ Lp,
                                                            int64_t nF, int P, int G, uint64_t *__restrict__ items,
                                                            uint32_t *__restrict__ seg_len, int *__restrict__ err) {
    const int64_t i0 = ((int64_t)blockIdx.x * MI_THREADS + threadIdx.x) * MI_ITEMS;
    if (i0 == 0 && Lp[NTETRA] != nF) *err = 2;   // sum(Lc) must equal |F|
    if (i0 >= nF) return;
    // largest t with Lp[t] <= i0 (Lp is non-decreasing, Lp[0] = 0)
    int lo = 0, hi = NTETRA - 1;
    while (lo < hi) {
        int mid = (lo + hi + 1) >> 1;
        if (Lp[mid] <= i0) lo = mid; else hi = mid - 1;
    }
    int t = lo;
    int64_t next = Lp[t + 1];  // Lp has NTETRA+1 entries, Lp[NTETRA] = nF
#pragma unroll
    for (int k = 0; k < MI_ITEMS; k++) {
        const int64_t i = i0 + k;
        if (i >= nF) break;
        while (i >= next && t < NTETRA - 1) { t++; next = Lp[t + 1]; }
        const pfaai_pg_t e = F[i];
        if ((unsigned)e.protein >= (unsigned)P || (unsigned)e.genome >= (unsigned)G) { *err = 1; continue; }
        const uint32_t seg = (uint32_t)e.genome * (uint32_t)P + (uint32_t)e.protein;
        items[i] = ((uint64_t)seg << TET_BITS) | (uint32_t)t;
        atomicAdd(&seg_len[seg], 1u);
    }
}

// T[p*G + g] = seg_len[g*P + p]; also the longest segment (sizes the bit-plane counters of the tile kernel)
__global__ void k_transpose_t(const uint32_t *__restrict__ seg_len, int P, int G, int32_t *__restrict__ T,
                              int *__restrict__ max_len) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    int v = 0;
    if (i < (int64_t)P * G) {
        const int g = (int)(i / P), p = (int)(i % P);
        v = (int)seg_len[i];
        if (T) T[(int64_t)p * G + g] = v;
    }
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) v = max(v, __shfl_xor_sync(0xffffffffu, v, d));
    if ((threadIdx.x & 31) == 0 && v > 0) atomicMax(max_len, v);
}

// ------------------------------------------------------------------------------------------------
// K2b: stable LSD radix sort, 8-bit digits, hand-written (no CUB/Thrust).
// A fixed set of RS_BLOCKS persistent blocks each owns one contiguous chunk of the input, so the
// per-(digit, block) histogram is tiny (256 x RS_BLOCKS) and one small block scans it.
//   pass = k_radix_hist (read 8 B/item) + k_radix_scan + k_radix_scatter (read 8 B, write 8 B / item;
//   the last pass writes only the 4-byte tetramer).  Stability: chunk order x tile order x warp strip
//   order x lane order == input order, ranks by __match_any_sync.
// ------------------------------------------------------------------------------------------------
constexpr int RS_THREADS = 256;
constexpr int RS_WARPS = RS_THREADS / 32;
constexpr int RS_ITEMS = 8;                       // rounds per warp strip
constexpr int RS_TILE = RS_THREADS * RS_ITEMS;    // 2048 items
constexpr int RS_RADIX = 256;

__global__ void __launch_bounds__(RS_THREADS) k_radix_hist(const uint64_t *__restrict__ in, int64_t n, int shift,
                                                            int64_t chunk, uint32_t *__restrict__ block_hist) {
    __shared__ uint32_t h[RS_WARPS][RS_RADIX];
    for (int i = threadIdx.x; i < RS_WARPS * RS_RADIX; i += RS_THREADS) (&h[0][0])[i] = 0;
    __syncthreads();
    const int warp = threadIdx.x >> 5;
    const int64_t lo = (int64_t)blockIdx.x * chunk, hi = min(lo + chunk, n);
    for (int64_t i = lo + threadIdx.x; i < hi; i += RS_THREADS)
        atomicAdd(&h[warp][(uint32_t)(in[i] >> shift) & (RS_RADIX - 1)], 1u);
    __syncthreads();
    for (int d = threadIdx.x; d < RS_RADIX; d += RS_THREADS) {
        uint32_t s = 0;
#pragma unroll
        for (int w = 0; w < RS_WARPS; w++) s += h[w][d];
        block_hist[(int64_t)d * gridDim.x + blockIdx.x] = s;   // digit-major
    }
}

// single block: exclusive scan of the digit-major histogram -> global base of every (digit, block)
__global__ void __launch_bounds__(1024) k_radix_scan(const uint32_t *__restrict__ block_hist, int64_t n,
                                                      int64_t *__restrict__ base) {
    __shared__ int64_t scratch[33];
    const int64_t per = (n + blockDim.x - 1) / blockDim.x;
    const int64_t lo = (int64_t)threadIdx.x * per, hi = min(lo + per, n);
    int64_t s = 0;
    for (int64_t i = lo; i < hi; i++) s += block_hist[i];
    int64_t run = block_exclusive_scan_i64(s, nullptr, scratch);
    for (int64_t i = lo; i < hi; i++) { base[i] = run; run += block_hist[i]; }
}

template <bool LAST>
__global__ void __launch_bounds__(RS_THREADS) k_radix_scatter(const uint64_t *__restrict__ in, int64_t n, int shift,
                                                               int64_t chunk, const int64_t *__restrict__ base,
                                                               uint64_t *__restrict__ out, uint32_t *__restrict__ out_vals) {
    __shared__ int64_t cursor[RS_RADIX];
    __shared__ uint32_t wcount[RS_WARPS][RS_RADIX];
    __shared__ uint32_t tot[RS_RADIX];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const uint32_t lt_mask = (1u << lane) - 1u;
    for (int d = threadIdx.x; d < RS_RADIX; d += RS_THREADS) cursor[d] = base[(int64_t)d * gridDim.x + blockIdx.x];
    const int64_t lo = (int64_t)blockIdx.x * chunk, hi = min(lo + chunk, n);
    for (int64_t tile0 = lo; tile0 < hi; tile0 += RS_TILE) {
        for (int i = threadIdx.x; i < RS_WARPS * RS_RADIX; i += RS_THREADS) (&wcount[0][0])[i] = 0;
        __syncthreads();  // also orders the cursor update of the previous tile
        uint64_t it[RS_ITEMS];
        uint32_t loc[RS_ITEMS];
#pragma unroll
        for (int r = 0; r < RS_ITEMS; r++) {
            const int64_t i = tile0 + (int64_t)warp * (32 * RS_ITEMS) + r * 32 + lane;
            const bool valid = i < hi;
            it[r] = valid ? in[i] : ~0ull;
            const uint32_t d = valid ? ((uint32_t)(it[r] >> shift) & (RS_RADIX - 1)) : RS_RADIX;  // RS_RADIX = "none"
            const uint32_t peers = __match_any_sync(0xffffffffu, d);
            const uint32_t rank = __popc(peers & lt_mask);
            uint32_t old = 0;
            if (valid) old = wcount[warp][d];
            __syncwarp();
            if (valid && rank == 0) wcount[warp][d] = old + __popc(peers);
            __syncwarp();
            loc[r] = old + rank;
        }
        __syncthreads();
        for (int d = threadIdx.x; d < RS_RADIX; d += RS_THREADS) {
            uint32_t run = 0;
#pragma unroll
            for (int w = 0; w < RS_WARPS; w++) { uint32_t c = wcount[w][d]; wcount[w][d] = run; run += c; }
            tot[d] = run;
        }
        __syncthreads();
#pragma unroll
        for (int r = 0; r < RS_ITEMS; r++) {
            const int64_t i = tile0 + (int64_t)warp * (32 * RS_ITEMS) + r * 32 + lane;
            if (i < hi) {
                const uint32_t d = (uint32_t)(it[r] >> shift) & (RS_RADIX - 1);
                const int64_t pos = cursor[d] + wcount[warp][d] + loc[r];
                if (LAST) out_vals[pos] = (uint32_t)it[r] & TET_MASK;
                else out[pos] = it[r];
            }
        }
        __syncthreads();
        for (int d = threadIdx.x; d < RS_RADIX; d += RS_THREADS) cursor[d] += tot[d];
    }
}

}  // namespace pfaai
