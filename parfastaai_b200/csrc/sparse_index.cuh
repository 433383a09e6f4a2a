// sparse_index.cuh -- the SPARSE side of the hybrid tile path (K5, aji_mma.cuh): every co-occurrence on a NON-dense tetramer
// (posting list shorter than tau) is counted inside the tile kernel by walking the tile's row items through the column block's
// sparse dictionary -- the reference's pair enumeration (constructTetramerTuples, ds_helper.hpp:270-357) restricted to the
// short lists, without ever writing a pair-level record.
//
//   row items     : per (group of 32 output rows, SCP) the rows' non-dense tetramers as one flat array of 32-bit items
//                   (bitmap byte offset | row | shift), padded to 128-item chunks: perfectly balanced coalesced LDG.128s.
//                       k_sp_count -> k_sp_group -> scan -> k_sp_fill
//   dictionaries  : per (block of 128 output columns, SCP) one blob
//                       [presence bitmap 20,016 B][16-bit rank per bitmap word 10,016 B][16-bit entries]
//                   entry of a tetramer held by ONE column of the block = that column, pre-encoded as the byte offset and the
//                   shift of its 16-bit counter; a tetramer held by several columns points at a list [n, e1 .. en] behind the
//                   entries.  One bulk copy (TMA) per (tile, SCP) brings the blob into shared memory.  Blobs that exceed the
//                   stage are split by columns (2, 4, 8 sub-dictionaries; the tile kernel walks the row items once per part).
//                       k_spdict_build
#pragma once
#include "common.cuh"

namespace pfaai {

constexpr int SP_ROWS = 32;                                   // rows per item group = one tensor-memory lane quadrant of a tile
constexpr int SP_COLS = 128;                                  // columns per dictionary = one tile
constexpr int SP_BM_WORDS = NTETRA / 32;                      // 5000 presence words
constexpr int SP_BM_BYTES = (SP_BM_WORDS + 4) * 4;            // 20016: + 4 always-zero words (padding items land there)
constexpr int SP_RK_BYTES = (((SP_BM_WORDS + 4) * 2) + 15) & ~15;   // 10016: uint16 rank of every bitmap word
constexpr int SP_ENT_OFF = SP_BM_BYTES + SP_RK_BYTES;         // 30032: first entry (slot 0 is never addressed: ranks start at 1)
// entries + list slots per blob (15-bit list offsets, 0x7fff reserved).  Two stage sizes, chosen per compute call by the expected
// number of non-dense tetramers per (column block, SCP): the tile kernel gives the rest of its shared memory to the operand ring.
constexpr int SP_ENT_MAX = 32590;                             // sparse inputs: one 93 KB dictionary stage
constexpr int SP_ENT_MAX_SMALL = 4990;                        // dense inputs: a 39 KB stage
__host__ __device__ constexpr int sp_stage_bytes(int ent_max) { return SP_ENT_OFF + 2 * (ent_max + 2); }
constexpr int SP_STAGE_BYTES = sp_stage_bytes(SP_ENT_MAX);    // 95216
constexpr int SP_MAXSUB = 8;                                  // column sub-dictionaries per (block, SCP)
constexpr int SP_CHUNK = 128;                                 // items per chunk (32 lanes x uint4)
constexpr uint32_t SP_PAD_ITEM = (uint32_t)(SP_BM_WORDS * 4) << 16;   // byte offset 20000 = first always-zero word
constexpr uint16_t SP_NOT_DENSE = 0xffffu;                    // == MM_NOT_DENSE (dense-id table of mma_index.cuh)
static_assert(SP_STAGE_BYTES % 16 == 0 && sp_stage_bytes(SP_ENT_MAX_SMALL) % 16 == 0 && SP_ENT_OFF % 16 == 0, "bulk-copy granularity");

// item of (row r of its group, tetramer v): the kernel loads the bitmap word at byte offset item >> 16, shifts it left by
// item & 31 (presence bit -> bit 31, presence bits of smaller tetramers below it) and takes the row from bits 8..12
__host__ __device__ __forceinline__ uint32_t sp_item(uint32_t r, uint32_t v) { return ((v >> 5) << 18) | (r << 8) | (31u - (v & 31u)); }
// entry of tile column c: byte offset of the 32-bit word that holds its 16-bit counter (bits 0..7) and the shift of the
// counter inside that word (bits 8..12: 0 or 16)
__host__ __device__ __forceinline__ uint32_t sp_entry(uint32_t c) { return ((c >> 1) << 2) | ((c & 1u) << 12); }

struct SpSub { int64_t off; uint32_t bytes; uint32_t n; };   // one sub-dictionary blob
static_assert(sizeof(SpSub) == 16, "SpSub is loaded as one uint4");

// ---- row items ----------------------------------------------------------------------------------------------------------
// one warp per (row, SCP): number of non-dense tetramers -> cnt[(g*P + p)*32 + r]  (g = row / 32, r = row % 32)
__global__ void __launch_bounds__(256) k_sp_count(const int64_t *__restrict__ seg_off, const uint32_t *__restrict__ vals,
                                                   const int32_t *__restrict__ rows, int n_rows, int P,
                                                   const uint16_t *__restrict__ did, const uint8_t *__restrict__ rb_owned,
                                                   int32_t *__restrict__ cnt) {
    const int lane = threadIdx.x & 31;
    const int64_t nwarp = ((int64_t)gridDim.x * blockDim.x) >> 5;
    for (int64_t w = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5; w < (int64_t)n_rows * P; w += nwarp) {
        const int i = (int)(w / P), p = (int)(w % P);
        if (!rb_owned[i >> 7]) continue;                          // row blocks of other shards: no items
        const int64_t s = (int64_t)rows[i] * P + p;
        const int64_t b = seg_off[s], e = seg_off[s + 1];
        const uint16_t *D = did + (int64_t)p * NTETRA;
        int n = 0;
        for (int64_t k = b + lane; k < e; k += 32) n += (D[vals[k]] == SP_NOT_DENSE) ? 1 : 0;
        for (int o = 16; o > 0; o >>= 1) n += __shfl_xor_sync(0xffffffffu, n, o);
        if (lane == 0) cnt[((int64_t)(i >> 5) * P + p) * SP_ROWS + (i & 31)] = n;
    }
}

// one thread per (group, SCP): counts -> exclusive prefix inside the group (in place), chunks of the group's slab, its items
__global__ void k_sp_group(int32_t *__restrict__ cnt, int64_t n_gp, int32_t *__restrict__ nch, int32_t *__restrict__ tot) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_gp) return;
    int32_t *c = cnt + i * SP_ROWS;
    int run = 0;
    for (int r = 0; r < SP_ROWS; r++) { const int v = c[r]; c[r] = run; run += v; }
    tot[i] = run;
    nch[i] = (run + SP_CHUNK - 1) / SP_CHUNK;
}

// one warp per (row, SCP): the row's non-dense tetramers as items at slab(g, p) + prefix; row 0's warp also pads the slab
__global__ void __launch_bounds__(256) k_sp_fill(const int64_t *__restrict__ seg_off, const uint32_t *__restrict__ vals,
                                                  const int32_t *__restrict__ rows, int n_rows, int P,
                                                  const uint16_t *__restrict__ did, const uint8_t *__restrict__ rb_owned,
                                                  const int32_t *__restrict__ cnt, const int32_t *__restrict__ tot, const int64_t *__restrict__ sp_off,
                                                  uint32_t *__restrict__ items) {
    const int lane = threadIdx.x & 31;
    const int64_t nwarp = ((int64_t)gridDim.x * blockDim.x) >> 5;
    for (int64_t w = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5; w < (int64_t)n_rows * P; w += nwarp) {
        const int i = (int)(w / P), p = (int)(w % P);
        if (!rb_owned[i >> 7]) continue;
        const int64_t gp = (int64_t)(i >> 5) * P + p;
        const uint32_t r = (uint32_t)(i & 31);
        uint32_t *dst = items + sp_off[gp] * SP_CHUNK;
        if (r == 0) {
            const int64_t end = (sp_off[gp + 1] - sp_off[gp]) * SP_CHUNK;
            for (int64_t k = tot[gp] + lane; k < end; k += 32) dst[k] = SP_PAD_ITEM;
        }
        const int64_t s = (int64_t)rows[i] * P + p;
        const int64_t b = seg_off[s], e = seg_off[s + 1];
        const uint16_t *D = did + (int64_t)p * NTETRA;
        int at = cnt[gp * SP_ROWS + r];
        for (int64_t k0 = b; k0 < e; k0 += 32) {
            const int64_t k = k0 + lane;
            uint32_t v = 0u;
            bool keep = false;
            if (k < e) { v = vals[k]; keep = D[v] == SP_NOT_DENSE; }
            const uint32_t bal = __ballot_sync(0xffffffffu, keep);
            if (keep) dst[at + __popc(bal & ((1u << lane) - 1u))] = sp_item(r, v);
            at += __popc(bal);
        }
    }
}

// ---- dictionaries ---------------------------------------------------------------------------------------------------------
constexpr int SD_THREADS = 512;
constexpr int SD_WARPS = SD_THREADS / 32;
constexpr int SD_WPT = (SP_BM_WORDS + SD_THREADS - 1) / SD_THREADS;      // bitmap words per thread in the rank scan
// shared memory: [bitmap][ranks][entries (SP_ENT_MAX + 2 uint16)][column begin/length][scan scratch]
constexpr int SD_SMEM = SP_STAGE_BYTES + SP_COLS * (8 + 4) + 48 * 4 + 16;

struct SpDictParams {
    const int64_t *seg_off; const uint32_t *vals;
    const int32_t *cols; int n_cols, P;
    const uint16_t *did;
    int ent_max;          // SP_ENT_MAX or SP_ENT_MAX_SMALL: capacity of the tile kernel's dictionary stage
    unsigned char *blob; unsigned long long blob_cap; unsigned long long *cursor;   // bump allocation; the cursor keeps counting beyond the capacity
    int32_t *nsub;        // [ncb * P]
    SpSub *sub;           // [ncb * P * SP_MAXSUB]
    int *err;             // 1 = the blob buffer is too small (re-run with the counted need), 2 = a dictionary does not fit even at SP_MAXSUB parts
};

struct SdSmem {
    uint32_t *W; uint16_t *rk; uint16_t *ent; int64_t *col_beg; int32_t *col_len; uint32_t *wsum; unsigned long long *s_off;
};

// block-wide exclusive scan of one value per thread; returns the exclusive prefix, *total = the sum (all threads)
__device__ __forceinline__ uint32_t sd_block_scan(const SdSmem &s, uint32_t v, uint32_t *total) {
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    uint32_t inc = v;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) { const uint32_t o = __shfl_up_sync(0xffffffffu, inc, d); if (lane >= d) inc += o; }
    __syncthreads();
    if (lane == 31) s.wsum[warp] = inc;
    __syncthreads();
    if (warp == 0) {
        const uint32_t wv = (lane < SD_WARPS) ? s.wsum[lane] : 0u;
        uint32_t winc = wv;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) { const uint32_t o = __shfl_up_sync(0xffffffffu, winc, d); if (lane >= d) winc += o; }
        s.wsum[lane] = winc - wv;
        if (lane == 31) s.wsum[32] = winc;
    }
    __syncthreads();
    *total = s.wsum[32];
    return s.wsum[warp] + inc - v;
}

// 16-bit add inside a shared-memory array of uint16 (no 16-bit atomics: add into the enclosing 32-bit word); returns the old value
__device__ __forceinline__ uint32_t sd_add16(uint16_t *a, uint32_t i, uint32_t v) {
    uint32_t *w = reinterpret_cast<uint32_t *>(a) + (i >> 1);
    const uint32_t sh = (i & 1u) * 16u;
    return (atomicAdd(w, v << sh) >> sh) & 0xffffu;
}

// visits every non-dense posting (column c, tetramer v) of columns [c0, c1): f(c, v)
template <class Fn>
__device__ __forceinline__ void sd_for_each(const SpDictParams &prm, const SdSmem &s, const uint16_t *D, int c0, int c1, Fn f) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (int c = c0 + warp; c < c1; c += SD_WARPS) {
        const int64_t b = s.col_beg[c];
        const int len = s.col_len[c];
        for (int k = lane; k < len; k += 32) {
            const uint32_t v = __ldg(prm.vals + b + k);
            if (D[v] == SP_NOT_DENSE) f(c, v);
        }
    }
}

// builds and appends the blob of columns [c0, c1) of the block; returns false when it does not fit one stage
__device__ __forceinline__ bool sd_build(const SpDictParams &prm, const SdSmem &s, const uint16_t *D, int c0, int c1, SpSub *out) {
    const int tid = threadIdx.x;
    __syncthreads();
    for (int i = tid; i < SP_BM_BYTES / 4; i += SD_THREADS) s.W[i] = 0u;
    __syncthreads();
    sd_for_each(prm, s, D, c0, c1, [&](int, uint32_t v) { atomicOr(&s.W[v >> 5], 1u << (v & 31u)); });
    __syncthreads();
    // ranks: exclusive prefix of the popcounts
    const int w0 = tid * SD_WPT, w1 = min(w0 + SD_WPT, SP_BM_WORDS);
    uint32_t sum = 0;
    for (int i = w0; i < w1; i++) sum += __popc(s.W[i]);
    uint32_t n;
    uint32_t run = sd_block_scan(s, sum, &n);
    if (n + 1 > (uint32_t)prm.ent_max) return false;
    for (int i = w0; i < w1; i++) { s.rk[i] = (uint16_t)run; run += __popc(s.W[i]); }
    for (int i = SP_BM_WORDS + tid; i < SP_BM_WORDS + 4; i += SD_THREADS) s.rk[i] = 0;
    // multiplicities (number of columns that hold the tetramer) in the entry slots
    uint32_t *e32 = reinterpret_cast<uint32_t *>(s.ent);
    for (uint32_t i = tid; i < (n + 2) / 2 + 1; i += SD_THREADS) e32[i] = 0u;
    __syncthreads();
    auto slot = [&](uint32_t v) -> uint32_t { const uint32_t w = s.W[v >> 5]; return s.rk[v >> 5] + __popc(w & ((2u << (v & 31u)) - 1u)); };
    sd_for_each(prm, s, D, c0, c1, [&](int, uint32_t v) { sd_add16(s.ent, slot(v), 1u); });
    __syncthreads();
    // lists of the multi-column tetramers behind the entries: slot i with m >= 2 columns gets m + 1 list slots
    const uint32_t per = (n + SD_THREADS) / SD_THREADS;                  // slots 1..n, contiguous per thread
    const uint32_t i0 = 1 + tid * per, i1 = min(i0 + per, n + 1);
    uint32_t need = 0;
    for (uint32_t i = i0; i < i1; i++) { const uint32_t m = s.ent[i]; need += (m >= 2u) ? m + 1u : 0u; }
    uint32_t n_list;
    uint32_t at = n + 1 + sd_block_scan(s, need, &n_list);
    const uint32_t n_ent = n + 1 + n_list;
    if (n_ent > (uint32_t)prm.ent_max) return false;
    // single-column slots are cleared, multi-column slots point at their list [cursor/count, e1 .. em], all zero for now
    for (uint32_t i = i0; i < i1; i++) {
        const uint32_t m = s.ent[i];
        if (m >= 2u) {
            s.ent[i] = (uint16_t)(0x8000u | at);
            for (uint32_t k = 0; k <= m; k++) s.ent[at + k] = 0;
            at += m + 1u;
        } else {
            s.ent[i] = 0;
        }
    }
    __syncthreads();
    // the fill pass touches the entry array with 32-bit atomic adds only (a 16-bit value is "stored" by adding it to its zeroed
    // half-word): no plain 16-bit store ever shares a word with a concurrent atomic
    sd_for_each(prm, s, D, c0, c1, [&](int c, uint32_t v) {
        const uint32_t i = slot(v);
        const uint32_t e = reinterpret_cast<volatile uint16_t *>(s.ent)[i];
        if (e & 0x8000u) { const uint32_t l = e & 0x7fffu; const uint32_t k = sd_add16(s.ent, l, 1u); sd_add16(s.ent, l + 1 + k, sp_entry((uint32_t)c)); }
        else sd_add16(s.ent, i, sp_entry((uint32_t)c));
    });
    __syncthreads();
    const uint32_t bytes = (uint32_t)SP_ENT_OFF + ((2u * n_ent + 15u) & ~15u);
    if (tid == 0) {
        unsigned long long off = atomicAdd(prm.cursor, (unsigned long long)bytes);
        if (off + bytes > prm.blob_cap) { atomicMax(prm.err, 1); off = ~0ull; }
        *s.s_off = off;
    }
    __syncthreads();
    const unsigned long long off = *s.s_off;
    if (off != ~0ull) {
        uint4 *dst = reinterpret_cast<uint4 *>(prm.blob + off);
        const uint4 *src = reinterpret_cast<const uint4 *>(s.W);            // bitmap, ranks and entries are contiguous in shared memory
        for (uint32_t i = tid; i < bytes / 16; i += SD_THREADS) dst[i] = src[i];
    }
    if (tid == 0) { SpSub d; d.off = (off == ~0ull) ? 0 : (int64_t)off; d.bytes = bytes; d.n = n; *out = d; }
    __syncthreads();
    return true;
}

// one CTA per (column block, SCP)
__global__ void __launch_bounds__(SD_THREADS, 2) k_spdict_build(const SpDictParams prm) {
    extern __shared__ __align__(16) unsigned char sm[];
    SdSmem s;
    s.W = reinterpret_cast<uint32_t *>(sm);
    s.rk = reinterpret_cast<uint16_t *>(sm + SP_BM_BYTES);
    s.ent = reinterpret_cast<uint16_t *>(sm + SP_ENT_OFF);
    s.col_beg = reinterpret_cast<int64_t *>(sm + SP_STAGE_BYTES);
    s.col_len = reinterpret_cast<int32_t *>(s.col_beg + SP_COLS);
    s.wsum = reinterpret_cast<uint32_t *>(s.col_len + SP_COLS);
    s.s_off = reinterpret_cast<unsigned long long *>(s.wsum + 48);
    const int tid = threadIdx.x, P = prm.P;
    const int cb = blockIdx.x / P, p = blockIdx.x % P;
    const int j0 = cb * SP_COLS, ncol = min(SP_COLS, prm.n_cols - j0);
    const uint16_t *D = prm.did + (int64_t)p * NTETRA;
    if (tid < SP_COLS) {
        int64_t beg = 0; int len = 0;
        if (tid < ncol) {
            const int64_t sg = (int64_t)prm.cols[j0 + tid] * P + p;
            beg = prm.seg_off[sg];
            len = (int)(prm.seg_off[sg + 1] - beg);
        }
        s.col_beg[tid] = beg; s.col_len[tid] = len;
    }
    __syncthreads();
    // any non-dense posting at all?  (dense inputs: most (block, SCP) have none -> no blob, no walk)
    {
        uint32_t any = 0;
        sd_for_each(prm, s, D, 0, ncol, [&](int, uint32_t) { any = 1u; });
        uint32_t tot;
        sd_block_scan(s, any, &tot);
        if (tot == 0) { if (tid == 0) prm.nsub[blockIdx.x] = 0; return; }
    }
    SpSub *sub = prm.sub + (int64_t)blockIdx.x * SP_MAXSUB;
    for (int nsub = 1; nsub <= SP_MAXSUB; nsub *= 2) {
        // the cursor is bumped per blob, so a failed attempt leaks its earlier parts (rare; the buffer is sized with slack)
        bool ok = true;
        for (int k = 0; k < nsub && ok; k++) {
            const int c0 = k * SP_COLS / nsub, c1 = max(c0, min((k + 1) * SP_COLS / nsub, ncol));
            ok = sd_build(prm, s, D, c0, c1, sub + k);
        }
        if (ok) { if (tid == 0) prm.nsub[blockIdx.x] = nsub; return; }
    }
    if (tid == 0) { prm.nsub[blockIdx.x] = 0; atomicMax(prm.err, 2); }
}

}  // namespace pfaai
