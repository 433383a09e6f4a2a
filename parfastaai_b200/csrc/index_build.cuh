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
constexpr int MI_ITEMS = 16;                     // rounds per warp: a warp converts 512 consecutive postings

// largest t in [lo, hi] with Lp[t] <= i   (Lp non-decreasing, Lp[lo] <= i)
__device__ __forceinline__ int tetramer_of(const int64_t *__restrict__ Lp, int lo, int hi, int64_t i) {
    while (lo < hi) {
        const int mid = (lo + hi + 1) >> 1;
        if (Lp[mid] <= i) lo = mid; else hi = mid - 1;
    }
    return lo;
}

// Ownership of a genome in a partitioned build (pfaai_gpu_build_partitioned): owner[genome >> OWN_SHIFT] == shard.  The table has
// one byte per 128 genomes (128 divides the row-block height, so it can express any assignment of row blocks to shards).
constexpr int OWN_SHIFT = 7;
static_assert(TILE_ROWS % (1 << OWN_SHIFT) == 0, "row blocks must be whole ownership granules");

// Partitioned build, pass 1: per 512-posting warp chunk, the number of postings whose genome this shard owns (the
// order-preserving compaction offsets of pass 2 are the exclusive scan of these counts); also the (genome, SCP)
// segment histogram of the owned postings.
__global__ void __launch_bounds__(MI_THREADS) k_count_owned(const pfaai_pg_t *__restrict__ F, int64_t nF, int P, int G,
                                                             const uint8_t *__restrict__ owner, int shard,
                                                             int32_t *__restrict__ warp_cnt, uint32_t *__restrict__ seg_len,
                                                             int *__restrict__ err) {
    const int lane = threadIdx.x & 31;
    const int64_t warp = ((int64_t)blockIdx.x * MI_THREADS + threadIdx.x) >> 5;
    const int64_t w0 = warp * (32 * MI_ITEMS);
    if (w0 >= nF) return;
    int cnt = 0;
#pragma unroll 4
    for (int k = 0; k < MI_ITEMS; k++) {
        const int64_t i = w0 + k * 32 + lane;
        bool own = false;
        if (i < nF) {
            const pfaai_pg_t e = F[i];
            if ((unsigned)e.protein >= (unsigned)P || (unsigned)e.genome >= (unsigned)G) *err = 1;
            else if (__ldg(owner + (e.genome >> OWN_SHIFT)) == (uint8_t)shard) {
                own = true;
                atomicAdd(&seg_len[(uint32_t)e.genome * (uint32_t)P + (uint32_t)e.protein], 1u);
            }
        }
        cnt += __popc(__ballot_sync(0xffffffffu, own));
    }
    if (lane == 0) warp_cnt[warp] = cnt;
}

// owner == NULL: every posting becomes an item at its own index and is counted in seg_len (replicated build).
// owner != NULL: only the postings of genomes owned by `shard` become items, compacted in input order from warp_off[warp]
// on (pass 2 of the partitioned build; seg_len was filled by k_count_owned).
__global__ void __launch_bounds__(MI_THREADS) k_make_items(const pfaai_pg_t *__restrict__ F, const int64_t *__restrict__ Lp,
                                                            int64_t nF, int P, int G, uint64_t *__restrict__ items,
                                                            uint32_t *__restrict__ seg_len, int *__restrict__ err,
                                                            const uint8_t *__restrict__ owner, int shard,
                                                            const int64_t *__restrict__ warp_off) {
    const int lane = threadIdx.x & 31;
    const int64_t warp = ((int64_t)blockIdx.x * MI_THREADS + threadIdx.x) >> 5;
    const int64_t w0 = warp * (32 * MI_ITEMS);
    if (w0 == 0 && lane == 0 && Lp[NTETRA] != nF) *err = 2;   // sum(Lc) must equal |F|
    if (w0 >= nF) return;
    const int64_t last = min(w0 + 32 * MI_ITEMS - 1, nF - 1);
    // one warp-uniform search for the tetramer of the warp's first posting; the next 32 block boundaries are then held
    // one per lane, and (the common case: the warp's 512 postings span fewer than 32 tetramer blocks) every posting finds
    // its tetramer with a 5-step shuffle search instead of touching Lp again.  All F loads / item stores are coalesced.
    const int t0 = tetramer_of(Lp, 0, NTETRA - 1, w0);
    const int tb = min(t0 + 1 + lane, NTETRA);
    const int64_t bnd = Lp[tb];                                       // start of block t0+1+lane (Lp[NTETRA] = nF)
    const bool local = __shfl_sync(0xffffffffu, bnd, 31) > last || t0 + 32 >= NTETRA;
    const int t_hi = local ? 0 : tetramer_of(Lp, t0, NTETRA - 1, last);
    int64_t dst = owner ? warp_off[warp] : 0;
    const uint32_t lt_mask = (1u << lane) - 1u;
#pragma unroll 4
    for (int k = 0; k < MI_ITEMS; k++) {
        const int64_t i = w0 + k * 32 + lane;
        const bool valid = i < nF;
        int t;
        if (local) {           // t = t0 + #{boundaries <= i}; boundaries ascend, so a binary search over the lanes
            int pos = 0;
#pragma unroll
            for (int sft = 16; sft > 0; sft >>= 1) {
                const int64_t b = __shfl_sync(0xffffffffu, bnd, pos + sft - 1);
                if (b <= i) pos += sft;
            }
            t = min(t0 + pos, NTETRA - 1);
        } else {
            t = valid ? tetramer_of(Lp, t0, t_hi, i) : 0;
        }
        if (owner) {
            bool own = false;
            uint32_t seg = 0u;
            if (valid) {
                const pfaai_pg_t e = F[i];
                if ((unsigned)e.protein < (unsigned)P && (unsigned)e.genome < (unsigned)G &&
                    __ldg(owner + (e.genome >> OWN_SHIFT)) == (uint8_t)shard) {
                    own = true;
                    seg = (uint32_t)e.genome * (uint32_t)P + (uint32_t)e.protein;
                }
            }
            const uint32_t bal = __ballot_sync(0xffffffffu, own);
            if (own) items[dst + __popc(bal & lt_mask)] = ((uint64_t)seg << TET_BITS) | (uint32_t)t;
            dst += __popc(bal);
            continue;
        }
        if (!valid) continue;
        const pfaai_pg_t e = F[i];
        uint32_t seg = (uint32_t)e.genome * (uint32_t)P + (uint32_t)e.protein;
        // a bad id fails the build; it is still counted (in segment 0) so that the sort's histograms stay consistent
        if ((unsigned)e.protein >= (unsigned)P || (unsigned)e.genome >= (unsigned)G) { *err = 1; seg = 0u; }
        items[i] = ((uint64_t)seg << TET_BITS) | (uint32_t)t;
        atomicAdd(&seg_len[seg], 1u);
    }
}

// ------------------------------------------------------------------------------------------------
// Routed build (pfaai_gpu_route_prepare): the context holds only a SLICE of F, postings [slice_begin, slice_begin + n) of
// the global array.  Every posting becomes an item tagged with the shard that owns its genome,
//     item = owner << 60 | (genome*P + protein) << 18 | tetramer,
// and the number of items per owner is counted (one ballot round per shard, one atomic per warp and owner).  One onesweep
// pass on the owner digit (shift 60) then groups the slice by destination, in input order; after the all-to-all every shard
// holds exactly the items of its own genomes, in global posting order (sources arrive in slice order).
// ------------------------------------------------------------------------------------------------
constexpr int ROUTE_SHIFT = 60;
constexpr int ROUTE_MAX_SHARDS = 16;
constexpr uint64_t ROUTE_ITEM_MASK = (1ull << ROUTE_SHIFT) - 1;

__global__ void __launch_bounds__(MI_THREADS) k_make_items_routed(const pfaai_pg_t *__restrict__ F, const int64_t *__restrict__ Lp,
                                                                   int64_t slice_begin, int64_t n, int64_t nF_total, int P, int G,
                                                                   const uint8_t *__restrict__ owner, int n_shards,
                                                                   uint64_t *__restrict__ items,
                                                                   unsigned long long *__restrict__ owner_cnt /*[256]*/,
                                                                   int *__restrict__ err) {
    const int lane = threadIdx.x & 31;
    const int64_t warp = ((int64_t)blockIdx.x * MI_THREADS + threadIdx.x) >> 5;
    const int64_t l0 = warp * (32 * MI_ITEMS);                 // local index of the warp's first posting
    if (l0 == 0 && lane == 0 && Lp[NTETRA] != nF_total) *err = 2;   // sum(Lc) must equal |F|
    if (l0 >= n) return;
    const int64_t w0 = slice_begin + l0;                       // global posting index
    const int64_t last = min(w0 + 32 * MI_ITEMS - 1, slice_begin + n - 1);
    const int t0 = tetramer_of(Lp, 0, NTETRA - 1, w0);
    const int tb = min(t0 + 1 + lane, NTETRA);
    const int64_t bnd = Lp[tb];
    const bool local = __shfl_sync(0xffffffffu, bnd, 31) > last || t0 + 32 >= NTETRA;
    const int t_hi = local ? 0 : tetramer_of(Lp, t0, NTETRA - 1, last);
    uint32_t cnt = 0;                                          // lane o accumulates the count of owner o
#pragma unroll 4
    for (int k = 0; k < MI_ITEMS; k++) {
        const int64_t li = l0 + k * 32 + lane, i = slice_begin + li;
        const bool valid = li < n;
        int t;
        if (local) {
            int pos = 0;
#pragma unroll
            for (int sft = 16; sft > 0; sft >>= 1) {
                const int64_t b = __shfl_sync(0xffffffffu, bnd, pos + sft - 1);
                if (b <= i) pos += sft;
            }
            t = min(t0 + pos, NTETRA - 1);
        } else {
            t = valid ? tetramer_of(Lp, t0, t_hi, i) : 0;
        }
        int o = -1;
        if (valid) {
            const pfaai_pg_t e = F[li];
            uint32_t seg = (uint32_t)e.genome * (uint32_t)P + (uint32_t)e.protein;
            o = 0;
            // a bad id fails the build; the item still travels (to shard 0, segment 0) so that all counts stay consistent
            if ((unsigned)e.protein >= (unsigned)P || (unsigned)e.genome >= (unsigned)G) { *err = 1; seg = 0u; }
            else o = __ldg(owner + (e.genome >> OWN_SHIFT));
            items[li] = ((uint64_t)o << ROUTE_SHIFT) | ((uint64_t)seg << TET_BITS) | (uint32_t)t;
        }
        for (int s = 0; s < n_shards; s++) {
            const uint32_t b = __ballot_sync(0xffffffffu, o == s);
            if (lane == s) cnt += __popc(b);
        }
    }
    if (lane < n_shards && cnt) atomicAdd(&owner_cnt[lane], (unsigned long long)cnt);
}

// (genome, SCP) histogram of the routed items a shard received
__global__ void __launch_bounds__(256) k_seg_hist(const uint64_t *__restrict__ items, int64_t n, int64_t nseg,
                                                   uint32_t *__restrict__ seg_len, int *__restrict__ err) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        const uint64_t seg = (items[i] & ROUTE_ITEM_MASK) >> TET_BITS;
        if ((int64_t)seg >= nseg) { *err = 1; continue; }
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
// K2b: stable LSD radix sort, 8-bit digits, hand-written (no CUB/Thrust), one kernel per pass
// ("onesweep": chained scan with decoupled look-back).
//   * The global digit histogram of EVERY pass is derived from the segment histogram seg_len (the sort key is
//     the segment id), so no pass over the items is spent on counting: k_digit_hist + k_digit_scan.
//   * k_onesweep: a tile of 8192 items is ranked in registers (per-warp match_any ranks, stable), its 256 digit
//     counts are published in a [tile][digit] status array (2 flag bits + 62-bit count per word, so |F| >= 2^30
//     is fine) and the exclusive prefix over the preceding tiles is gathered by look-back; the tile is then
//     reordered by digit in shared memory and written out with consecutive threads on consecutive addresses.
//     Per pass and item: read 8 B, write 8 B (last pass: 4 B, the tetramer only) -- the HBM minimum for an
//     out-of-place pass.  Tiles take their index from an atomic counter, so every tile that is waited on is
//     already resident (forward progress without a cooperative launch).
//   Stability: tile order x warp strip order x round order x lane order == input order.
// ------------------------------------------------------------------------------------------------
constexpr int RS_RADIX = 256;
#ifndef PF_OS_THREADS
#define PF_OS_THREADS 384
#endif
#ifndef PF_OS_MATCH
#define PF_OS_MATCH 0                             // 1: MATCH.ANY ranks (fewer instructions, longer latency); 0: 8 ballots per item
#endif
#ifndef PF_OS_CTAS
#define PF_OS_CTAS 2
#endif
constexpr int OS_THREADS = PF_OS_THREADS;         // >= 256 (one thread per digit in the scan / look-back phases)
constexpr int OS_WARPS = OS_THREADS / 32;
constexpr int OS_ITEMS = 16;                          // rounds per warp strip
constexpr int OS_TILE = OS_THREADS * OS_ITEMS;        // 8192 items = 64 KB
constexpr int OS_SMEM = OS_TILE * 8 + OS_WARPS * RS_RADIX * 4 + RS_RADIX * (4 + 8) + 64;
constexpr int MAX_PASSES = 4;
constexpr int OS_LB = 8;                               // look-back window: status words fetched per round trip
constexpr uint64_t OS_FLAG_LOCAL = 1ull << 62, OS_FLAG_INCL = 2ull << 62, OS_COUNT_MASK = (1ull << 62) - 1;

// hist[pass][d] = number of items whose pass-th key digit is d = sum of seg_len over the segments with that digit
__global__ void __launch_bounds__(256) k_digit_hist(const uint32_t *__restrict__ seg_len, int64_t nseg, int passes,
                                                     unsigned long long *__restrict__ hist /*[MAX_PASSES][256]*/) {
    __shared__ unsigned long long h[MAX_PASSES][RS_RADIX];
    for (int i = threadIdx.x; i < MAX_PASSES * RS_RADIX; i += blockDim.x) (&h[0][0])[i] = 0ull;
    __syncthreads();
    for (int64_t s = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; s < nseg; s += (int64_t)gridDim.x * blockDim.x) {
        const unsigned long long n = seg_len[s];
        if (n)
            for (int k = 0; k < passes; k++) atomicAdd(&h[k][((uint32_t)s >> (8 * k)) & 255u], n);
    }
    __syncthreads();
    for (int i = threadIdx.x; i < passes * RS_RADIX; i += blockDim.x)
        if ((&h[0][0])[i]) atomicAdd(&hist[i], (&h[0][0])[i]);
}

// in place: hist[pass][*] -> exclusive prefix (global base of every digit); one block of 256 threads per pass
__global__ void __launch_bounds__(RS_RADIX) k_digit_scan(unsigned long long *__restrict__ hist) {
    __shared__ int64_t scratch[33];
    unsigned long long *h = hist + (size_t)blockIdx.x * RS_RADIX;
    const int64_t v = (int64_t)h[threadIdx.x];
    const int64_t ex = block_exclusive_scan_i64(v, nullptr, scratch);
    h[threadIdx.x] = (unsigned long long)ex;
}

__device__ __forceinline__ uint64_t ld_volatile_u64(const uint64_t *p) {
    uint64_t v;
    asm volatile("ld.volatile.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_volatile_u64(uint64_t *p, uint64_t v) {
    asm volatile("st.volatile.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}

// control: [0] = tile counter, followed (at +8 words) by status[n_tiles][256]; zeroed before every pass
template <bool LAST>
__global__ void __launch_bounds__(OS_THREADS, PF_OS_CTAS) k_onesweep(const uint64_t *__restrict__ in, int64_t n, int shift,
                                                             const unsigned long long *__restrict__ digit_base /*[256]*/,
                                                             uint64_t *__restrict__ control, uint64_t *__restrict__ out,
                                                             uint32_t *__restrict__ out_vals) {
    extern __shared__ __align__(16) unsigned char os_smem[];
    uint64_t *sm_items = reinterpret_cast<uint64_t *>(os_smem);                                   // [OS_TILE]
    uint32_t *wcount = reinterpret_cast<uint32_t *>(os_smem + OS_TILE * 8);                       // [OS_WARPS][256]
    int64_t *gbase = reinterpret_cast<int64_t *>(os_smem + OS_TILE * 8 + OS_WARPS * RS_RADIX * 4);   // [256]
    uint32_t *dstart = reinterpret_cast<uint32_t *>(gbase + RS_RADIX);                            // [256]
    uint32_t *misc = dstart + RS_RADIX;                                                           // [0] tile id; [1..] scan scratch
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint32_t lt_mask = (1u << lane) - 1u;
    uint64_t *status = control + 8;

    if (tid == 0) misc[0] = (uint32_t)atomicAdd(reinterpret_cast<unsigned long long *>(control), 1ull);
    for (int i = tid; i < OS_WARPS * RS_RADIX; i += OS_THREADS) wcount[i] = 0u;
    __syncthreads();
    const int64_t tile = misc[0];
    const int64_t tile0 = tile * OS_TILE;
    const int n_tile = (int)min((int64_t)OS_TILE, n - tile0);

    // ---- load + per-warp stable ranks ----
    // peers of an item = the lanes of its round that hold the same digit, found with 8 ballots (one per digit bit; the 16
    // rounds are independent, so the ballots pipeline -- MATCH.ANY serialised this phase: 40 % of the v1 kernel's stalls);
    // the lowest peer lane then bumps the warp's digit counter and broadcasts the old value.
    uint64_t it[OS_ITEMS];
    uint32_t peers[OS_ITEMS];
    uint16_t loc[OS_ITEMS];
    uint32_t *wc = wcount + warp * RS_RADIX;
#pragma unroll
    for (int r = 0; r < OS_ITEMS; r++) {
        const int li = warp * (32 * OS_ITEMS) + r * 32 + lane;
        it[r] = (li < n_tile) ? in[tile0 + li] : ~0ull;
    }
#pragma unroll
    for (int r = 0; r < OS_ITEMS; r++) {
        const int li = warp * (32 * OS_ITEMS) + r * 32 + lane;
        const bool valid = li < n_tile;
        const uint32_t d = (uint32_t)(it[r] >> shift) & (RS_RADIX - 1);
#if PF_OS_MATCH
        const uint32_t m = __match_any_sync(0xffffffffu, valid ? d : (RS_RADIX + (uint32_t)lane));
#else
        uint32_t m = __ballot_sync(0xffffffffu, valid);
#pragma unroll
        for (int b = 0; b < 8; b++) {
            const bool bit = (d >> b) & 1u;
            const uint32_t bal = __ballot_sync(0xffffffffu, bit);
            m &= bit ? bal : ~bal;
        }
#endif
        peers[r] = valid ? m : 0u;
    }
#pragma unroll
    for (int r = 0; r < OS_ITEMS; r++) {
        const uint32_t d = (uint32_t)(it[r] >> shift) & (RS_RADIX - 1);
        const uint32_t pm = peers[r];
        const int leader = __ffs(pm) - 1;                 // -1 for an invalid item
        uint32_t old = 0;
        if (lane == leader) { old = wc[d]; wc[d] = old + __popc(pm); }
        __syncwarp();
        old = __shfl_sync(0xffffffffu, old, leader & 31);
        loc[r] = (uint16_t)(old + __popc(pm & lt_mask));
    }
    __syncthreads();

    // ---- tile digit counts, look-back, bases (threads 0..255 = digits) ----
    uint32_t cnt = 0;
    if (tid < RS_RADIX) {
#pragma unroll
        for (int w = 0; w < OS_WARPS; w++) { const uint32_t c = wcount[w * RS_RADIX + tid]; wcount[w * RS_RADIX + tid] = cnt; cnt += c; }
        st_volatile_u64(&status[tile * RS_RADIX + tid], (tile == 0 ? OS_FLAG_INCL : OS_FLAG_LOCAL) | cnt);
    }
    // exclusive scan of the 256 digit counts -> start of every digit inside the tile (warps 0..7 hold the digits)
    {
        uint32_t inc = cnt;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) { const uint32_t o = __shfl_up_sync(0xffffffffu, inc, d); if (lane >= d) inc += o; }
        if (tid < RS_RADIX && lane == 31) misc[1 + warp] = inc;
        __syncthreads();
        if (tid < RS_RADIX) {
            uint32_t woff = 0;
            for (int w = 0; w < warp; w++) woff += misc[1 + w];
            dstart[tid] = woff + inc - cnt;
        }
    }
    if (tid < RS_RADIX) {
        // look-back, OS_LB predecessors per round trip: their status words are fetched together (independent loads) and
        // consumed nearest-first; a word that is not published yet is re-polled (its tile is resident: it took its id
        // earlier).  A one-at-a-time walk costs one L2 latency per in-flight predecessor and dominated the v1 kernel.
        uint64_t excl = 0;
        bool done = (tile == 0);
        for (int64_t t = tile - 1; !done; t -= OS_LB) {
            uint64_t v[OS_LB];
#pragma unroll
            for (int u = 0; u < OS_LB; u++)
                v[u] = (t - u >= 0) ? ld_volatile_u64(&status[(t - u) * RS_RADIX + tid]) : OS_FLAG_INCL;   // before tile 0: prefix 0
#pragma unroll
            for (int u = 0; u < OS_LB; u++) {
                if (!done) {
                    while ((v[u] >> 62) == 0) v[u] = ld_volatile_u64(&status[(t - u) * RS_RADIX + tid]);
                    excl += v[u] & OS_COUNT_MASK;
                    done = (v[u] & OS_FLAG_INCL) != 0;
                }
            }
        }
        if (tile > 0) st_volatile_u64(&status[tile * RS_RADIX + tid], OS_FLAG_INCL | (excl + cnt));
        gbase[tid] = (int64_t)digit_base[tid] + (int64_t)excl - (int64_t)dstart[tid];
    }
    __syncthreads();

    // ---- reorder by digit in shared memory, then coalesced write-out ----
#pragma unroll
    for (int r = 0; r < OS_ITEMS; r++) {
        const int li = warp * (32 * OS_ITEMS) + r * 32 + lane;
        if (li < n_tile) {
            const uint32_t d = (uint32_t)(it[r] >> shift) & (RS_RADIX - 1);
            sm_items[dstart[d] + wc[d] + loc[r]] = it[r];
        }
    }
    __syncthreads();
    for (int i = tid; i < n_tile; i += OS_THREADS) {
        const uint64_t v = sm_items[i];
        const uint32_t d = (uint32_t)(v >> shift) & (RS_RADIX - 1);
        const int64_t pos = gbase[d] + i;
        if (LAST) out_vals[pos] = (uint32_t)v & TET_MASK;
        else out[pos] = v;
    }
}

// ------------------------------------------------------------------------------------------------
// K1a: Lc from an explicit tetramer key per posting (constructLc, ds_helper.hpp:82-109 <- scp_db.hpp:121-158): the postings
// arrive tetramer-major, so the count of tetramer t is (last index of its run + 1) - (first index of its run) -- two
// boundary stores per run, no atomics, one coalesced pass over 4 bytes per posting (HBM-bound).  run_beg / Lc must be zeroed.
// err[0] = 1 when the keys are not ascending or out of range.
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_lc_run_begins(const int32_t *__restrict__ tet, int64_t n, int32_t *__restrict__ run_beg_lo,
                                                        int32_t *__restrict__ run_beg_hi, int *__restrict__ err) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        const int32_t t = tet[i];
        if (t < 0 || t >= NTETRA) { *err = 1; continue; }
        const int32_t prev = (i > 0) ? tet[i - 1] : -1;
        if (prev > t) *err = 1;
        if (prev != t) { run_beg_lo[t] = (int32_t)(i & 0x7fffffff); run_beg_hi[t] = (int32_t)(i >> 31); }
    }
}
__global__ void __launch_bounds__(256) k_lc_run_ends(const int32_t *__restrict__ tet, int64_t n, const int32_t *__restrict__ run_beg_lo,
                                                      const int32_t *__restrict__ run_beg_hi, int32_t *__restrict__ Lc) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        const int32_t t = tet[i];
        if (t < 0 || t >= NTETRA) continue;
        if (i + 1 == n || tet[i + 1] != t) Lc[t] = (int32_t)(i + 1 - (((int64_t)run_beg_hi[t] << 31) | (int64_t)run_beg_lo[t]));
    }
}

// ------------------------------------------------------------------------------------------------
// Import of device-ready lists (pfaai_gpu_import_lists: the index cache's fast path): validates what K1-K3 + the sort would
// have guaranteed -- offsets ascending from 0 to nF, tetramers strictly ascending inside a segment and below NTETRA, T(p, g)
// == the segment's length unless the caller supplies its own T -- and returns the longest segment.  One pass, HBM-bound.
// flags[0] = 1 on a violation, flags[1] = max segment length (atomicMax).  With a caller-supplied T (row-major P x G) every
// T(p, g) must be at least the length of segment (g, p): the Jaccard denominator ta + tb - c indexes the epilogue's reciprocal
// table and would wrap around for a T smaller than the lists it describes; flags[2] = 1 when one is.
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_check_lists(const int64_t *__restrict__ seg_off, const uint32_t *__restrict__ vals, int64_t nseg,
                                                      int64_t nF, int *__restrict__ flags, const int32_t *__restrict__ T_or_null, int P,
                                                      int G) {
    const int lane = threadIdx.x & 31;
    const int64_t nwarp = ((int64_t)gridDim.x * blockDim.x) >> 5;
    int mx = 0;
    bool bad = false;
    for (int64_t sg = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5; sg < nseg; sg += nwarp) {
        const int64_t b = seg_off[sg], e = seg_off[sg + 1];
        if (b < 0 || e < b || e > nF || e - b > (1 << 20) || (sg == 0 && b != 0) || (sg == nseg - 1 && e != nF)) { bad = true; continue; }
        mx = max(mx, (int)(e - b));
        if (T_or_null && lane == 0 && (int64_t)T_or_null[(int64_t)(sg % P) * G + sg / P] < e - b) flags[2] = 1;
        for (int64_t k = b + lane; k < e; k += 32) {
            const uint32_t v = vals[k];
            if (v >= (uint32_t)NTETRA || (k > b && vals[k - 1] >= v)) bad = true;
        }
    }
    if (bad) flags[0] = 1;
    for (int o = 16; o > 0; o >>= 1) mx = max(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    if (lane == 0 && mx > 0) atomicMax(flags + 1, mx);
}
// T[p][g] = length of segment (g, p)
__global__ void k_t_from_offsets(const int64_t *__restrict__ seg_off, int P, int G, int32_t *__restrict__ T) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (int64_t)P * G) return;
    const int g = (int)(i / P), p = (int)(i % P);
    T[(int64_t)p * G + g] = (int32_t)(seg_off[i + 1] - seg_off[i]);
}

}  // namespace pfaai
