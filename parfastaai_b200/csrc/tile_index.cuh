// tile_index.cuh -- per-compute-call operand layouts of the tile kernel (K4a/K4b of DESIGN.md):
//
//   row slabs      : the output rows, 16 at a time, as warp-interleaved slices of range-local tetramer ids
//                    (one coalesced 512-byte LDG.128 per warp and chunk in the tile kernel);
//   dictionaries   : for every (block of 64 output columns, SCP, tetramer range) the map
//                    tetramer -> 64-bit column mask, stored as [presence/rank bitmap][masks] blobs that the
//                    tile kernel pulls into shared memory with ONE bulk copy (TMA) per pipeline stage.
//
// Both are derived from the genome-major lists (seg_off / vals) that index_build.cuh produced.  Together
// they play the role of the reference's E array (constructTetramerTuples, ds_helper.hpp:270-357) and of its
// sort (psort.hpp:27-53): they group the postings so that all (pair, SCP) intersections of a tile can be
// counted on-chip -- without ever materialising a pair-level structure.
#pragma once
#include "common.cuh"

namespace pfaai {

// ------------------------------------------------------------------------------------------------
// Row slabs.  Lane l = 2*r + h of group g walks the elements of parity h of row g*16+r's list, per
// tetramer range q:  slab(g,p,q)[c*32 + l] = 4 consecutive range-local ids (PAD_ID padded), each stored
// pre-decoded for the kernel's lookup (slab_encode: bitmap byte offset and shift count).
// n_chunks(g,p,q) = ceil(max_r len_q / 8), rounded up to an even number (the kernel consumes 2 chunks
// per step).  Index of a slab: (g*P + p)*NRANGE + q.
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ int lower_bound_u32(const uint32_t *__restrict__ a, int n, uint32_t key) {
    int lo = 0, hi = n;
    while (lo < hi) {
        const int mid = (lo + hi) >> 1;
        if (a[mid] < key) lo = mid + 1; else hi = mid;
    }
    return lo;
}

__global__ void k_pack_count(const int64_t *__restrict__ seg_off, const uint32_t *__restrict__ vals,
                             const int32_t *__restrict__ rows, int n_rows, int n_groups, int P,
                             int32_t *__restrict__ nch) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (int64_t)n_groups * P) return;
    const int g = (int)(i / P), p = (int)(i % P);
    int mx[NRANGE];
#pragma unroll
    for (int q = 0; q < NRANGE; q++) mx[q] = 0;
    for (int r = 0; r < ROWS_PER_WARP; r++) {
        const int row = g * ROWS_PER_WARP + r;
        if (row >= n_rows) break;
        const int64_t s = (int64_t)rows[row] * P + p;
        const int64_t o = seg_off[s];
        const int len = (int)(seg_off[s + 1] - o);
        int prev = 0;
#pragma unroll
        for (int q = 0; q < NRANGE; q++) {
            const int end = (q == NRANGE - 1) ? len : lower_bound_u32(vals + o, len, (uint32_t)(q + 1) * RANGE_T);
            mx[q] = max(mx[q], end - prev);
            prev = end;
        }
    }
#pragma unroll
    for (int q = 0; q < NRANGE; q++) nch[i * NRANGE + q] = (((mx[q] + 7) >> 3) + 1) & ~1;
}

__global__ void __launch_bounds__(256) k_pack_fill(const int64_t *__restrict__ seg_off, const uint32_t *__restrict__ vals,
                                                    const int32_t *__restrict__ rows, int n_rows, int n_groups, int P,
                                                    const int64_t *__restrict__ ell_off, uint4 *__restrict__ ell) {
    const int lane = threadIdx.x & 31;
    const int64_t warp0 = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
    const int r = lane >> 1, h = lane & 1;
    for (int64_t w = warp0; w < (int64_t)n_groups * P; w += nwarps) {
        const int g = (int)(w / P), p = (int)(w % P);
        const int row = g * ROWS_PER_WARP + r;
        int64_t o = 0; int len = 0;
        if (row < n_rows) {
            const int64_t s = (int64_t)rows[row] * P + p;
            o = seg_off[s]; len = (int)(seg_off[s + 1] - o);
        }
        int prev = 0;
#pragma unroll
        for (int q = 0; q < NRANGE; q++) {
            const int end = (q == NRANGE - 1) ? len : lower_bound_u32(vals + o, len, (uint32_t)(q + 1) * RANGE_T);
            const int lq = end - prev;
            const int64_t c0 = ell_off[w * NRANGE + q];
            const int n = (int)(ell_off[w * NRANGE + q + 1] - c0);
            uint4 *dst = ell + c0 * 32 + lane;
            for (int c = 0; c < n; c++) {
                uint32_t e[4];
#pragma unroll
                for (int u = 0; u < 4; u++) {
                    const int k = 2 * (4 * c + u) + h;
                    e[u] = slab_encode((k < lq) ? vals[o + prev + k] - (uint32_t)q * RANGE_T : PAD_ID);
                }
                dst[(int64_t)c * 32] = make_uint4(e[0], e[1], e[2], e[3]);
            }
            prev = end;
        }
    }
}

// ------------------------------------------------------------------------------------------------
// Dictionaries.  One CTA per (column block, SCP).  For each tetramer range:
//   count : presence bits of the block's column lists (smem atomicOr), exclusive scan of the popcounts
//           -> 16-bit rank per bitmap word; n = number of distinct tetramers;
//   emit  : one 64-bit mask per distinct tetramer (smem atomicOr of the column bit), then the blob
//           [bitmap BM_BYTES][masks 8*(n+1, padded even)] is appended to the dictionary buffer (bump
//           allocation with one atomicAdd) and described by a SubDesc.
// If n+1 exceeds the stage capacity MASK_CAP the columns are split into 2, 4, ... max_sub equal
// sub-ranges, each with its own blob (the tile kernel simply runs one more stage over the same row slab).
// ------------------------------------------------------------------------------------------------
constexpr int DB_THREADS = 512;
constexpr int DB_WARPS = DB_THREADS / 32;
constexpr int DB_ILP = 1;                         // independent posting loads in flight per warp (4 measured slower: 7.2 vs 6.1 ms)
constexpr int DB_WPT = (BM_WORDS + DB_THREADS - 1) / DB_THREADS;   // bitmap words scanned per thread
constexpr int DB_SMEM = BM_BYTES + MASK_CAP * 8 + COLS_PER_TILE * (8 + 4 + 4) + 40 * 4 + 16;

struct __align__(16) SubDesc { int64_t off; uint32_t bytes; uint32_t n; };

struct DictParams {
    const int64_t *seg_off; const uint32_t *vals;
    const int32_t *cols, *col_t; int n_cols;
    const int32_t *T; int t_ld, P, max_sub;
    unsigned char *blob; unsigned long long blob_base, blob_cap; unsigned long long *cursor;   // blobs go to [blob_base, blob_cap)
    const int32_t *cb_list;   // column block of the k-th dictionary group (NULL: k itself); a partitioned build lists the shard's own blocks
    int slot_base;            // slot of group k = slot_base + k (colT / stage_rng are indexed by slot)
    int32_t *nsub;     // [n_groups*P*NRANGE]
    SubDesc *sub;      // [n_groups*P*NRANGE*max_sub]
    int32_t *colT;     // [n_slots*P*64]  T(p, column) of the block's columns (0 beyond the last column)
    int *err;          // set to 1 when the dictionary buffer is too small (cursor still counts the need)
};

struct DictSmem {
    uint32_t *W; uint2 *masks; int64_t *col_beg; int32_t *col_len, *col_split; uint32_t *wsum;
    unsigned long long *s_off;
};

// presence + ranks of columns [c0,c1) restricted to range q; returns the number of distinct tetramers
__device__ __forceinline__ uint32_t dict_count(const DictParams &prm, const DictSmem &s, int q, int c0, int c1) {
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    __syncthreads();
    for (int i = tid; i < BM_WORDS_PAD; i += DB_THREADS) s.W[i] = 0u;
    __syncthreads();
    int maxlen = 0;
    for (int c = c0; c < c1; c++) maxlen = max(maxlen, q ? s.col_len[c] - s.col_split[c] : s.col_split[c]);
    const int nchunk = (maxlen + 31) >> 5, ncs = c1 - c0;
    // DB_ILP independent (column, 32-posting chunk) tasks per warp and step: the loads are issued together
    for (int t0 = warp; t0 < nchunk * ncs; t0 += DB_WARPS * DB_ILP) {
        uint32_t v[DB_ILP];
#pragma unroll
        for (int u = 0; u < DB_ILP; u++) {
            const int t = t0 + u * DB_WARPS;
            v[u] = ~0u;
            if (t < nchunk * ncs) {
                const int c = c0 + t % ncs, k = (t / ncs) * 32 + lane;
                const int beg = q ? s.col_split[c] : 0, cnt = q ? s.col_len[c] - beg : s.col_split[c];
                if (k < cnt) v[u] = __ldg(prm.vals + s.col_beg[c] + beg + k) - (uint32_t)q * RANGE_T;
            }
        }
#pragma unroll
        for (int u = 0; u < DB_ILP; u++)
            if (v[u] != ~0u) atomicOr(&s.W[v[u] >> 4], 1u << (v[u] & 15u));
    }
    __syncthreads();
    const int w0 = tid * DB_WPT, w1 = min(w0 + DB_WPT, BM_WORDS);
    uint32_t sum = 0;
    for (int i = w0; i < w1; i++) sum += __popc(s.W[i]);
    uint32_t inc = sum;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const uint32_t o = __shfl_up_sync(0xffffffffu, inc, d);
        if (lane >= d) inc += o;
    }
    if (lane == 31) s.wsum[warp] = inc;
    __syncthreads();
    if (warp == 0) {
        const uint32_t wv = (lane < DB_WARPS) ? s.wsum[lane] : 0u;
        uint32_t winc = wv;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const uint32_t o = __shfl_up_sync(0xffffffffu, winc, d);
            if (lane >= d) winc += o;
        }
        s.wsum[lane] = winc - wv;
        if (lane == 31) s.wsum[32] = winc;
    }
    __syncthreads();
    const uint32_t n = s.wsum[32];
    if (n + 1 <= (uint32_t)MASK_CAP) {
        uint32_t run = s.wsum[warp] + inc - sum;
        for (int i = w0; i < w1; i++) { const uint32_t w = s.W[i]; s.W[i] = w | (run << 16); run += __popc(w); }
    }
    __syncthreads();
    return n;
}

// masks of columns [c0,c1) in range q (s.W holds the bitmap of exactly these columns); appends the blob
__device__ __forceinline__ void dict_emit(const DictParams &prm, const DictSmem &s, int q, int c0, int c1, uint32_t n,
                                          int64_t slot) {
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint32_t n_pad = (n + 2) & ~1u;                    // slots 0..n, padded to an even count
    for (uint32_t i = tid; i < n_pad; i += DB_THREADS) s.masks[i] = make_uint2(0u, 0u);
    __syncthreads();
    int maxlen = 0;
    for (int c = c0; c < c1; c++) maxlen = max(maxlen, q ? s.col_len[c] - s.col_split[c] : s.col_split[c]);
    const int nchunk = (maxlen + 31) >> 5, ncs = c1 - c0;
    uint32_t *m32 = reinterpret_cast<uint32_t *>(s.masks);
    for (int t0 = warp; t0 < nchunk * ncs; t0 += DB_WARPS * DB_ILP) {
        uint32_t v[DB_ILP];
        int cc[DB_ILP];
#pragma unroll
        for (int u = 0; u < DB_ILP; u++) {
            const int t = t0 + u * DB_WARPS;
            v[u] = ~0u; cc[u] = 0;
            if (t < nchunk * ncs) {
                const int c = c0 + t % ncs, k = (t / ncs) * 32 + lane;
                const int beg = q ? s.col_split[c] : 0, cnt = q ? s.col_len[c] - beg : s.col_split[c];
                cc[u] = c;
                if (k < cnt) v[u] = __ldg(prm.vals + s.col_beg[c] + beg + k) - (uint32_t)q * RANGE_T;
            }
        }
#pragma unroll
        for (int u = 0; u < DB_ILP; u++) {
            if (v[u] != ~0u) {
                const uint32_t w = s.W[v[u] >> 4];
                const uint32_t idx = (w >> 16) + __popc(w & ((2u << (v[u] & 15u)) - 1u) & 0xffffu);   // rank + #present <= v
                atomicOr(&m32[2 * idx + (cc[u] >> 5)], 1u << (cc[u] & 31));
            }
        }
    }
    const uint32_t bytes = BM_BYTES + 8u * n_pad;
    if (tid == 0) {
        unsigned long long off = prm.blob_base + atomicAdd(prm.cursor, (unsigned long long)bytes);
        if (off + bytes > prm.blob_cap) { *prm.err = 1; off = ~0ull; }
        *s.s_off = off;
    }
    __syncthreads();
    const unsigned long long off = *s.s_off;
    if (off != ~0ull) {
        uint4 *dst = reinterpret_cast<uint4 *>(prm.blob + off);
        const uint4 *srcW = reinterpret_cast<const uint4 *>(s.W);
        const uint4 *srcM = reinterpret_cast<const uint4 *>(s.masks);
        for (int i = tid; i < BM_BYTES / 16; i += DB_THREADS) dst[i] = srcW[i];
        dst += BM_BYTES / 16;
        for (uint32_t i = tid; i < n_pad / 2; i += DB_THREADS) dst[i] = srcM[i];
        if (tid == 0) {
            SubDesc d; d.off = (int64_t)off; d.bytes = bytes; d.n = n;
            prm.sub[slot] = d;
        }
    }
    __syncthreads();
}

__global__ void __launch_bounds__(DB_THREADS, 2) k_dict_build(const DictParams prm) {   // 2 CTAs / SM: the phases are latency-bound
    extern __shared__ __align__(16) unsigned char sm[];
    DictSmem s;
    s.W = reinterpret_cast<uint32_t *>(sm);
    s.masks = reinterpret_cast<uint2 *>(sm + BM_BYTES);
    s.col_beg = reinterpret_cast<int64_t *>(sm + BM_BYTES + MASK_CAP * 8);
    s.col_len = reinterpret_cast<int32_t *>(s.col_beg + COLS_PER_TILE);
    s.col_split = s.col_len + COLS_PER_TILE;
    s.wsum = reinterpret_cast<uint32_t *>(s.col_split + COLS_PER_TILE);
    s.s_off = reinterpret_cast<unsigned long long *>(s.wsum + 40);
    const int tid = threadIdx.x;
    const int P = prm.P;
    const int kgrp = blockIdx.x / P, p = blockIdx.x % P;
    const int cb = prm.cb_list ? prm.cb_list[kgrp] : kgrp;
    const int j0 = cb * COLS_PER_TILE, ncol = min(COLS_PER_TILE, prm.n_cols - j0);
    if (tid < COLS_PER_TILE) {
        int64_t beg = 0; int len = 0, split = 0, tv = 0;
        if (tid < ncol) {
            const int64_t sg = (int64_t)prm.cols[j0 + tid] * P + p;
            beg = prm.seg_off[sg];
            len = (int)(prm.seg_off[sg + 1] - beg);
            split = lower_bound_u32(prm.vals + beg, len, (uint32_t)RANGE_T);
            tv = prm.T[(int64_t)p * prm.t_ld + prm.col_t[j0 + tid]];
        }
        s.col_beg[tid] = beg; s.col_len[tid] = len; s.col_split[tid] = split;
        prm.colT[((int64_t)(prm.slot_base + kgrp) * P + p) * COLS_PER_TILE + tid] = tv;
    }
    __syncthreads();
    for (int q = 0; q < NRANGE; q++) {
        const int64_t slot0 = ((int64_t)blockIdx.x * NRANGE + q) * prm.max_sub;
        int nsub = 1;
        for (;;) {
            bool fits = true;
            for (int sidx = 0; sidx < nsub; sidx++) {
                const int c0 = sidx * COLS_PER_TILE / nsub, c1 = (sidx + 1) * COLS_PER_TILE / nsub;
                const uint32_t n = dict_count(prm, s, q, c0, c1);
                if (n + 1 > (uint32_t)MASK_CAP) { fits = false; break; }
                if (nsub == 1) dict_emit(prm, s, q, c0, c1, n, slot0);
            }
            if (fits || nsub >= prm.max_sub) break;   // max_sub is chosen by the host so that it always fits
            nsub *= 2;
        }
        if (nsub > 1) {
            for (int sidx = 0; sidx < nsub; sidx++) {
                const int c0 = sidx * COLS_PER_TILE / nsub, c1 = (sidx + 1) * COLS_PER_TILE / nsub;
                const uint32_t n = dict_count(prm, s, q, c0, c1);
                dict_emit(prm, s, q, c0, c1, n, slot0 + sidx);
            }
        }
        if (tid == 0) prm.nsub[(int64_t)blockIdx.x * NRANGE + q] = nsub;
    }
}

// flatten the sub-dictionaries into the per-column-block stage lists the tile kernel walks in order; group k (the k-th
// column block of this build) owns slot slot_base + k, its stages are stored from stage_base on
__global__ void k_stage_fill(const int32_t *__restrict__ nsub, const int64_t *__restrict__ stage_off, const SubDesc *__restrict__ sub,
                             int n_groups, int P, int max_sub, StageDesc *__restrict__ stages, int64_t *__restrict__ stage_rng,
                             int slot_base, int64_t stage_base, int64_t stage_cap, int *__restrict__ err) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t per = (int64_t)P * NRANGE, n = (int64_t)n_groups * per;
    if (i >= n) return;
    if (i == 0 && stage_base + stage_off[n] > stage_cap) *err = 1;
    if (stage_base + stage_off[i + 1] > stage_cap) return;
    if (i % per == 0) {
        const int64_t k = i / per;
        stage_rng[2 * (slot_base + k)] = stage_base + stage_off[i];
        stage_rng[2 * (slot_base + k) + 1] = stage_base + stage_off[i + per];
    }
    const int q = (int)(i % NRANGE), p = (int)((i / NRANGE) % P);
    const int ns = nsub[i];
    const int64_t base = stage_base + stage_off[i];
    for (int sidx = 0; sidx < ns; sidx++) {
        const SubDesc d = sub[i * max_sub + sidx];
        StageDesc st;
        st.off = d.off; st.bytes = d.bytes; st.p = (uint16_t)p; st.q = (uint8_t)q;
        st.flags = (uint8_t)(((q == 0 && sidx == 0) ? 1 : 0) | ((q == NRANGE - 1 && sidx == ns - 1) ? 2 : 0));
        stages[base + sidx] = st;
    }
}

}  // namespace pfaai
