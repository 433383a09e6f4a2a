// aji_tile.cuh -- the core kernel (K4 of DESIGN.md): per-(genome pair, SCP) intersection counts accumulated
// on-chip, with Jaccard, the per-pair sum S, the count N and AJI = S/N fused into the same tile kernel.
//
// Replaces (reference, relative to /root/reference):
//   constructTetramerTuples   include/pfaai/ds_helper.hpp:270-357   (pair enumeration -> E)
//   parallelMergeSort         include/pfaai/psort.hpp:27-53         (grouping of E by pair, SCP)
//   findEBlockExtents         include/pfaai/algorithm_impl.hpp:123-219
//   computeEBlockJAC          include/pfaai/algorithm_impl.hpp:222-277  (run length c -> J = c/(Ta+Tb-c), S += J, N += 1)
//   computeAJI                include/pfaai/algorithm_impl.hpp:309-322  (AJI = S/N)
//   printOutput matrix fill   src/main.cpp:143-154                   (mirror across the diagonal, zero diagonal)
//
// Formulation.  One CTA owns a tile of (NW*16 rows) x (64 columns) genome pairs and loops over the SCPs in
// ascending index (same summation order as the reference).  Per SCP:
//   build : the 64 column genomes' tetramer lists are scattered into a shared-memory dictionary
//           tetramer -> 64-bit column mask: a 160k-bit presence bitmap with interleaved 16-bit ranks
//           (40 KB) + one uint2 mask per distinct tetramer (warp-aggregation-free: smem atomicOr only
//           in this build step, |cols|*len operations, amortised over NW*16 rows).
//   walk  : lane (row r, half h) streams its half of row r's sorted list (coalesced LDG.128 from the packed
//           slices), looks every tetramer up (1 LDS.32 + 1 LDS.64) and adds the 64-bit mask into BIT-SLICED
//           counters (carry-save adders, LOP3 only): one mask add advances 64 pair counters at once.
//           No atomics, no E array, no per-pair memory traffic.
//   epilogue: the two half-lanes exchange halves (SHFL), each lane expands 32 counters, forms
//           J = c / (T(p,a) + T(p,b) - c) in fp64, S += J, N += (c > 0)  -- registers only.
// After the last SCP: AJI = S/N written to the dense output (both triangles in symmetric mode).
#pragma once
#include "common.cuh"

namespace pfaai {

struct TileParams {
    // row side: packed slices (k_pack_fill)
    const uint4 *ell;
    const int64_t *ell_off;    // [n_groups * P] in 512-byte chunks
    const int32_t *ell_nch;    // [n_groups * P]
    const int32_t *rows;       // [n_rows] genome id of kernel row i
    const int32_t *row_t;      // [n_rows] T column used for row i
    int n_rows, n_groups;
    // column side: genome-major lists
    const int32_t *cols;       // [n_cols]
    const int32_t *col_t;      // [n_cols]
    int n_cols;
    const int64_t *seg_off;    // [G*P + 1]
    const uint32_t *vals;      // [nF]
    const int32_t *T;          // [P x t_ld]
    int t_ld, P;
    // output
    double *out;               // cell (row i, col j) -> out[i*out_rs + j*out_cs]; symmetric: indexed by genome ids
    int64_t out_rs, out_cs;
    int symmetric;
    double *S_out;             // optional hooks, [n_rows x n_cols]
    int32_t *N_out;
    int32_t *counts_out;       // optional, [n_rows x n_cols x P]
    const int2 *tiles;         // (row block, col block) per CTA
    int *err;                  // device error flag: 1 = a single column list overflows the dictionary
    int mask_cap;              // capacity (entries) of the mask array in shared memory
};

// carry-save adder on one 32-bit slice: (h, l) = a + b + c
#define PF_CSA(h, l, a, b, c)                       \
    do {                                            \
        const uint32_t u__ = (a) ^ (b);             \
        const uint32_t c__ = (c);                   \
        (h) = ((a) & (b)) | (u__ & c__);            \
        (l) = u__ ^ c__;                            \
    } while (0)

struct Mask64 { uint32_t lo, hi; };

__device__ __forceinline__ Mask64 dict_lookup(const uint32_t *__restrict__ bm, const uint2 *__restrict__ masks, uint32_t v) {
    const uint32_t w = bm[v >> 4];
    const uint32_t b = v & 15u;
    const uint32_t below = w & ((1u << b) - 1u);          // presence bits of smaller tetramers in this word
    const uint32_t idx = (w >> 16) + __popc(below) + 1u;    // slot 0 is the all-zero "miss" slot
    const uint32_t slot = ((w >> b) & 1u) ? idx : 0u;
    const uint2 m = masks[slot];
    Mask64 r; r.lo = m.x; r.hi = m.y;
    return r;
}

template <int NW, int NPL>
__global__ void __launch_bounds__(NW * 32, 1) aji_tile_kernel(const TileParams prm) {
    constexpr int NT = NW * 32;
    extern __shared__ __align__(16) uint32_t smem[];
    // ---- shared memory carve-up ----
    uint32_t *bm = smem;                                          // [BM_WORDS] (+pad)
    constexpr int BM_PAD = (BM_WORDS + 3) & ~3;
    int64_t *col_seg = reinterpret_cast<int64_t *>(smem + BM_PAD); // [64] start of the column's list
    int32_t *col_len = reinterpret_cast<int32_t *>(col_seg + COLS_PER_TILE);  // [64]
    int32_t *col_tv = col_len + COLS_PER_TILE;                    // [64] T(p, col)
    int32_t *col_g = col_tv + COLS_PER_TILE;                      // [64] genome id (-1 = none)
    uint32_t *wsum = reinterpret_cast<uint32_t *>(col_g + COLS_PER_TILE);  // [34] scan scratch
    uint2 *masks = reinterpret_cast<uint2 *>(wsum + 36);          // [mask_cap + 1]

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int2 tile = prm.tiles[blockIdx.x];
    const int P = prm.P;
    const int group = tile.x * NW + warp;
    const bool group_ok = group < prm.n_groups;
    const int r = lane >> 1, h = lane & 1;
    const int row = group * ROWS_PER_WARP + r;
    const bool row_ok = row < prm.n_rows;
    const int gi = row_ok ? prm.rows[row] : -1;
    const int ti = row_ok ? prm.row_t[row] : 0;
    const int j0 = tile.y * COLS_PER_TILE;
    const int ncol = min(COLS_PER_TILE, prm.n_cols - j0);

    if (tid < COLS_PER_TILE) col_g[tid] = (tid < ncol) ? prm.cols[j0 + tid] : -1;

    double S[32];
    uint32_t n16[16];
#pragma unroll
    for (int j = 0; j < 32; j++) S[j] = 0.0;
#pragma unroll
    for (int j = 0; j < 16; j++) n16[j] = 0u;

    // words of the bitmap each thread scans; odd so that consecutive lanes hit distinct banks
    constexpr int WPT = ((BM_WORDS + NT - 1) / NT) | 1;

    for (int p = 0; p < P; p++) {
        const int ta = row_ok ? prm.T[(int64_t)p * prm.t_ld + ti] : 0;
        int c0 = 0;
        while (c0 < ncol) {
            int c1 = ncol;
            uint32_t n_distinct = 0;
            // ================= build the dictionary for columns [c0, c1) =================
            for (;;) {
                __syncthreads();   // previous users of bm / masks / col_* are done
                for (int i = tid; i < BM_PAD; i += NT) bm[i] = 0u;
                if (tid < COLS_PER_TILE) {
                    int len = 0; int64_t so = 0; int tv = 0;
                    if (tid >= c0 && tid < c1) {
                        const int64_t s = (int64_t)col_g[tid] * P + p;
                        so = prm.seg_off[s];
                        len = (int)(prm.seg_off[s + 1] - so);
                        tv = prm.T[(int64_t)p * prm.t_ld + prm.col_t[j0 + tid]];
                    }
                    col_seg[tid] = so; col_len[tid] = len; col_tv[tid] = tv;
                }
                __syncthreads();
                // presence bits: task = (32-entry chunk, column); columns of one SCP have similar lengths
                int maxlen = 0;
                for (int c = c0; c < c1; c++) maxlen = max(maxlen, col_len[c]);   // smem broadcast reads
                const int nchunk = (maxlen + 31) >> 5, ncs = c1 - c0;
                for (int t = warp; t < nchunk * ncs; t += NW) {
                    const int c = c0 + t % ncs, k = (t / ncs) * 32 + lane;
                    if (k < col_len[c]) {
                        const uint32_t v = prm.vals[col_seg[c] + k];
                        atomicOr(&bm[v >> 4], 1u << (v & 15u));
                    }
                }
                __syncthreads();
                // ranks: exclusive scan of popc(presence) over the words, stored in the high half-words
                {
                    const int w0 = tid * WPT, w1 = min(w0 + WPT, BM_WORDS);
                    uint32_t s = 0;
                    for (int i = w0; i < w1; i++) s += __popc(bm[i]);
                    uint32_t inc = s;
#pragma unroll
                    for (int d = 1; d < 32; d <<= 1) {
                        const uint32_t o = __shfl_up_sync(0xffffffffu, inc, d);
                        if (lane >= d) inc += o;
                    }
                    if (lane == 31) wsum[warp] = inc;
                    __syncthreads();
                    if (warp == 0) {
                        const uint32_t wv = (lane < NW) ? wsum[lane] : 0u;
                        uint32_t winc = wv;
#pragma unroll
                        for (int d = 1; d < 32; d <<= 1) {
                            const uint32_t o = __shfl_up_sync(0xffffffffu, winc, d);
                            if (lane >= d) winc += o;
                        }
                        wsum[lane] = winc - wv;
                        if (lane == 31) wsum[32] = winc;
                    }
                    __syncthreads();
                    uint32_t run = wsum[warp] + inc - s;
                    n_distinct = wsum[32];
                    if (n_distinct <= (uint32_t)prm.mask_cap) {
                        for (int i = w0; i < w1; i++) { const uint32_t w = bm[i]; bm[i] = w | (run << 16); run += __popc(w); }
                    }
                }
                if (n_distinct <= (uint32_t)prm.mask_cap) break;
                if (c1 - c0 == 1) {           // a single list does not fit: give up on this tile, flag it
                    if (tid == 0) *prm.err = 1;
                    return;                   // uniform: n_distinct is the same in every thread
                }
                c1 = c0 + ((c1 - c0) >> 1);   // halve the column range and rebuild
            }
            // zero the masks (slot 0 = miss), then OR every column's bit into its tetramers' masks
            for (uint32_t i = tid; i <= n_distinct; i += NT) masks[i] = make_uint2(0u, 0u);
            __syncthreads();
            {
                int maxlen = 0;
                for (int c = c0; c < c1; c++) maxlen = max(maxlen, col_len[c]);
                const int nchunk = (maxlen + 31) >> 5, ncs = c1 - c0;
                uint32_t *m32 = reinterpret_cast<uint32_t *>(masks);
                for (int t = warp; t < nchunk * ncs; t += NW) {
                    const int c = c0 + t % ncs, k = (t / ncs) * 32 + lane;
                    if (k < col_len[c]) {
                        const uint32_t v = prm.vals[col_seg[c] + k];
                        const uint32_t w = bm[v >> 4];
                        const uint32_t idx = (w >> 16) + __popc(w & ((1u << (v & 15u)) - 1u)) + 1u;
                        atomicOr(&m32[2 * idx + (c >> 5)], 1u << (c & 31));
                    }
                }
            }
            __syncthreads();

            // ================= walk: bit-sliced accumulation of the row half-lists =================
            uint32_t plo[NPL], phi[NPL];
#pragma unroll
            for (int k = 0; k < NPL; k++) { plo[k] = 0u; phi[k] = 0u; }
            if (group_ok) {
                const int64_t gp = (int64_t)group * P + p;
                const int nch = prm.ell_nch[gp];
                const uint4 *src = prm.ell + prm.ell_off[gp] * 32 + lane;
                int c = 0;
                for (; c + 4 <= nch; c += 4) {       // 16 lookups per iteration
                    uint4 q[4];
#pragma unroll
                    for (int u = 0; u < 4; u++) q[u] = __ldg(src + (int64_t)(c + u) * 32);
                    uint32_t e_lo[2], e_hi[2];       // two "eights" carries
#pragma unroll
                    for (int half = 0; half < 2; half++) {
                        uint32_t f_lo[2], f_hi[2];   // two "fours" carries
#pragma unroll
                        for (int qq = 0; qq < 2; qq++) {
                            const uint4 v = q[half * 2 + qq];
                            const Mask64 m0 = dict_lookup(bm, masks, v.x), m1 = dict_lookup(bm, masks, v.y);
                            const Mask64 m2 = dict_lookup(bm, masks, v.z), m3 = dict_lookup(bm, masks, v.w);
                            uint32_t ta_lo, ta_hi, tb_lo, tb_hi;
                            PF_CSA(ta_lo, plo[0], plo[0], m0.lo, m1.lo); PF_CSA(ta_hi, phi[0], phi[0], m0.hi, m1.hi);
                            PF_CSA(tb_lo, plo[0], plo[0], m2.lo, m3.lo); PF_CSA(tb_hi, phi[0], phi[0], m2.hi, m3.hi);
                            PF_CSA(f_lo[qq], plo[1], plo[1], ta_lo, tb_lo); PF_CSA(f_hi[qq], phi[1], phi[1], ta_hi, tb_hi);
                        }
                        PF_CSA(e_lo[half], plo[2], plo[2], f_lo[0], f_lo[1]); PF_CSA(e_hi[half], phi[2], phi[2], f_hi[0], f_hi[1]);
                    }
                    uint32_t cl, ch;                 // "sixteens" carry, rippled into planes 4..
                    PF_CSA(cl, plo[3], plo[3], e_lo[0], e_lo[1]); PF_CSA(ch, phi[3], phi[3], e_hi[0], e_hi[1]);
#pragma unroll
                    for (int k = 4; k < NPL; k++) {
                        const uint32_t tl = plo[k] & cl, th = phi[k] & ch;
                        plo[k] ^= cl; phi[k] ^= ch; cl = tl; ch = th;
                    }
                }
                for (; c < nch; c++) {               // tail: 4 lookups per iteration
                    const uint4 v = __ldg(src + (int64_t)c * 32);
                    const Mask64 m0 = dict_lookup(bm, masks, v.x), m1 = dict_lookup(bm, masks, v.y);
                    const Mask64 m2 = dict_lookup(bm, masks, v.z), m3 = dict_lookup(bm, masks, v.w);
                    uint32_t ta_lo, ta_hi, tb_lo, tb_hi, cl, ch;
                    PF_CSA(ta_lo, plo[0], plo[0], m0.lo, m1.lo); PF_CSA(ta_hi, phi[0], phi[0], m0.hi, m1.hi);
                    PF_CSA(tb_lo, plo[0], plo[0], m2.lo, m3.lo); PF_CSA(tb_hi, phi[0], phi[0], m2.hi, m3.hi);
                    PF_CSA(cl, plo[1], plo[1], ta_lo, tb_lo); PF_CSA(ch, phi[1], phi[1], ta_hi, tb_hi);
#pragma unroll
                    for (int k = 2; k < NPL; k++) {
                        const uint32_t tl = plo[k] & cl, th = phi[k] & ch;
                        plo[k] ^= cl; phi[k] ^= ch; cl = tl; ch = th;
                    }
                }
            }
            // ================= pair reduce: lane h keeps columns [32h, 32h+32) of both halves =================
            uint32_t R[NPL + 1];
            {
                uint32_t carry = 0u;
#pragma unroll
                for (int k = 0; k < NPL; k++) {
                    const uint32_t mine = h ? phi[k] : plo[k];
                    const uint32_t send = h ? plo[k] : phi[k];
                    const uint32_t theirs = __shfl_xor_sync(0xffffffffu, send, 1);
                    const uint32_t u = mine ^ theirs;
                    R[k] = u ^ carry;
                    carry = (mine & theirs) | (u & carry);
                }
                R[NPL] = carry;
            }
            // ================= epilogue: counts -> Jaccard -> S, N (registers) =================
            const int cbase = 32 * h;
#pragma unroll
            for (int jj = 0; jj < 32; jj++) {
                const int col = cbase + jj;
                uint32_t cnt = 0;
#pragma unroll
                for (int k = 0; k <= NPL; k++) cnt |= ((R[k] >> jj) & 1u) << k;
                const bool active = row_ok && col >= c0 && col < c1;
                if (active) {
                    const int gj = col_g[col];
                    if (gj == gi) cnt = 0;      // a genome is never paired with itself (isValidPair, ds_impl.hpp:89-91)
                    if (prm.counts_out)
                        prm.counts_out[((int64_t)row * prm.n_cols + (j0 + col)) * P + p] = (int32_t)cnt;
                    if (cnt > 0) {
                        const int den = ta + col_tv[col] - (int)cnt;
                        S[jj] += (double)(int)cnt / (double)den;   // algorithm_impl.hpp:250-254
                        n16[jj >> 1] += 1u << (16 * (jj & 1));
                    }
                }
            }
            c0 = c1;
        }
    }

    // ================= output: AJI = S / N (computeAJI) into the dense matrix =================
    const int cbase = 32 * h;
#pragma unroll
    for (int jj = 0; jj < 32; jj++) {
        const int col = cbase + jj;
        if (row_ok && col < ncol) {
            const int gj = col_g[col];
            const int n = (int)((n16[jj >> 1] >> (16 * (jj & 1))) & 0xffffu);
            const double aji = (n > 0) ? S[jj] / (double)n : 0.0;
            const int64_t j = j0 + col;
            if (prm.S_out) prm.S_out[(int64_t)row * prm.n_cols + j] = S[jj];
            if (prm.N_out) prm.N_out[(int64_t)row * prm.n_cols + j] = n;
            if (prm.symmetric) {
                if (gi < gj) {
                    prm.out[(int64_t)gi * prm.out_rs + (int64_t)gj * prm.out_cs] = aji;
                    prm.out[(int64_t)gj * prm.out_rs + (int64_t)gi * prm.out_cs] = aji;
                } else if (gi == gj) {
                    prm.out[(int64_t)gi * prm.out_rs + (int64_t)gj * prm.out_cs] = 0.0;
                }
            } else {
                prm.out[(int64_t)row * prm.out_rs + j * prm.out_cs] = (gi == gj) ? 0.0 : aji;
            }
        }
    }
}

}  // namespace pfaai
