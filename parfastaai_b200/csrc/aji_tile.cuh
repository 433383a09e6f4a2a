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
// Formulation.  A persistent CTA (one per SM, 24 walker warps) owns tiles of 384 rows x 64 columns of genome
// pairs and, per tile, walks the column block's STAGE LIST: for every SCP in ascending index (the reference's
// summation order) and every tetramer range, one dictionary blob
//        [presence/rank bitmap over the range | one 64-bit column mask per distinct tetramer]
// built once per compute call by k_dict_build (tile_index.cuh).
//   pipeline : blobs arrive in a 2-slot shared-memory ring by cp.async.bulk (TMA, SASS UBLKCP) completing on
//              an mbarrier; a slot is refilled by whichever warp is the LAST to finish walking it (an
//              acq_rel shared-memory counter), so there is no producer warp and no block-wide barrier inside
//              a tile: warps drift up to one stage apart and the LDS-bound walk of one warp overlaps the
//              ALU/FP64-bound epilogue of another.
//   walk     : lane (row r, half h) streams its half of row r's sorted list (coalesced LDG.128 from the packed
//              slabs, prefetched one step ahead), looks every tetramer up (1 LDS.32 + 1 LDS.64) and adds the
//              64-bit mask into BIT-SLICED counters (carry-save adders, LOP3 only): one mask add advances
//              64 pair counters at once.  No atomics, no E array, no per-pair memory traffic.
//   epilogue : after the last stage of an SCP the two half-lanes exchange halves (SHFL), each lane expands 32
//              counters, forms J = c / (T(p,a) + T(p,b) - c) in fp64, S += J, N += (c > 0).  The 32 fp64 sums of a
//              lane live in TENSOR MEMORY (tcgen05.ld / tcgen05.st, SASS LDTM / STTM): 64 registers per thread are
//              freed, which is what lets 24 warps (instead of 16) hide the shared-memory latency of the walk.
//              The division is done with a shared-memory reciprocal table + two FMAs (correctly rounded, see
//              jaccard_fast) instead of the ~25-instruction fp64 division sequence.
// After the last SCP: AJI = S/N written to the dense output (both triangles in symmetric mode).
#pragma once
#include "common.cuh"

namespace pfaai {

struct TileParams {
    // row side: packed slabs (k_pack_fill)
    const uint4 *ell;
    const int64_t *ell_off;    // [n_groups * P * NRANGE + 1] in 512-byte chunks
    const int32_t *rows;       // [n_rows] genome id of kernel row i
    const int32_t *row_t;      // [n_rows] T column used for row i
    int n_rows, n_groups;
    // column side: dictionaries (k_dict_build / k_stage_fill)
    const int32_t *cols;       // [n_cols]
    int n_cols;
    const StageDesc *stages;
    const int64_t *stage_rng;  // [n_slots][2]: first / one-past-last stage of the column block stored in slot s
    const unsigned char *blob;
    const int32_t *colT;       // [n_slots * P * 64]; slot of a column block = tiles[].w (== J in an unpartitioned context)
    const int32_t *T;          // [P x t_ld]
    int t_ld, P;
    // output
    double *out;               // cell (row i, col j) -> out[(i - I*TILE_ROWS + tile.z)*out_rs + j*out_cs]
    int64_t out_rs, out_cs;
    int symmetric;             // all-vs-all: only cells with row genome < column genome are computed ...
    int mirror;                // ... and also stored at the transposed position (single-shard contexts)
    double *S_out;             // optional hooks, [n_rows x n_cols]
    int32_t *N_out;
    int32_t *counts_out;       // optional, [n_rows x n_cols x P]
    const int4 *tiles;         // (row block I, col block J, first output row of block I in this shard's matrix, slot of J)
    int n_tiles;
    int *tile_counter;         // next tile to hand out (0 at launch)
    const int4 *tiles2;        // optional second list, taken up when the first one is exhausted (its own counter)
    int n_tiles2;
    int *tile_counter2;
    uint32_t k27, k16;         // 2^27, 2^16 (PF_IMAD_HI: shift-by-multiply constants)
    // streaming output (pfaai_gpu_set_progress): progress[g] counts the finished tiles that store into output rows 64g .. 64g+63
    // (the tile's own 6 granules; in mirror mode also granule J, where the transposed cells land).  NULL = off.
    int *progress;
    int n_gran;
    uint32_t piece;            // a stage blob travels as bulk copies of this many bytes (multiple of 16), all completing on the stage's
                               // mbarrier: one copy keeps only a few bytes per cycle in flight, many overlap their latencies
};

// ---- PTX wrappers: mbarrier + bulk async copy (TMA) ----
__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t *bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t *bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAIT_%=:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra DONE_%=;\n"
        "bra WAIT_%=;\n"
        "DONE_%=:\n"
        "}\n" ::"r"(smem_u32(bar)), "r"(parity) : "memory");
}
__device__ __forceinline__ void bulk_g2s(void *dst_smem, const void *src_gmem, uint32_t bytes, uint64_t *bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(dst_smem)),
                 "l"(src_gmem), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}
__device__ __forceinline__ uint32_t atom_add_acq_rel_smem(uint32_t *p, uint32_t v) {
    uint32_t old;
    asm volatile("atom.acq_rel.cta.shared::cta.add.u32 %0, [%1], %2;" : "=r"(old) : "r"(smem_u32(p)), "r"(v) : "memory");
    return old;
}

// ---- PTX wrappers: tensor memory (tcgen05) used as per-thread accumulator storage ----
__device__ __forceinline__ void tmem_alloc(uint32_t *smem_dst, uint32_t ncols) {   // one full warp
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(smem_dst)), "r"(ncols) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t ncols) {      // the warp that allocated
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}
// 16 consecutive 32-bit columns of this thread's lane <-> 16 registers (= 8 doubles)
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, uint32_t (&v)[16]) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
                 : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]),
                   "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
                 : "r"(taddr));
}
// the registers of a preceding tmem_ld16 may be read only after this wait; passing them through makes the
// dependency visible to the compiler as well
__device__ __forceinline__ void tmem_wait_ld(uint32_t (&v)[16]) {
    asm volatile("tcgen05.wait::ld.sync.aligned;"
                 : "+r"(v[0]), "+r"(v[1]), "+r"(v[2]), "+r"(v[3]), "+r"(v[4]), "+r"(v[5]), "+r"(v[6]), "+r"(v[7]), "+r"(v[8]),
                   "+r"(v[9]), "+r"(v[10]), "+r"(v[11]), "+r"(v[12]), "+r"(v[13]), "+r"(v[14]), "+r"(v[15])::"memory");
}
__device__ __forceinline__ void tmem_st16(uint32_t taddr, const uint32_t (&v)[16]) {
    asm volatile("tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16};" ::"r"(taddr),
                 "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7]), "r"(v[8]), "r"(v[9]),
                 "r"(v[10]), "r"(v[11]), "r"(v[12]), "r"(v[13]), "r"(v[14]), "r"(v[15])
                 : "memory");
}
__device__ __forceinline__ void tmem_wait_st() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }

// J = c / den, correctly rounded (== IEEE fp64 division, algorithm_impl.hpp:250-254).  Fast path: r = RN(1/den) from the
// shared-memory table, q = RN(c*r), q' = RN(q + (c - q*den)*r) with the residual exact in one FMA (Markstein); q' equals
// RN(c/den) for every 0 <= c <= den < 20000 (checked exhaustively in tests/test_division.py with the same operations).
__device__ __forceinline__ double lds_f64(uint32_t a) { double v; asm volatile("ld.shared.f64 %0, [%1];" : "=d"(v) : "r"(a)); return v; }
__device__ __forceinline__ int lds_s32(uint32_t a) { int v; asm volatile("ld.shared.s32 %0, [%1];" : "=r"(v) : "r"(a)); return v; }
// rcp32: 32-bit shared address of the reciprocal table; requires 1 <= den < RCP_N and 0 <= c <= den
__device__ __forceinline__ double jaccard_fast(uint32_t rcp32, int c, int den) {
    const double r = lds_f64(rcp32 + 8u * (uint32_t)den), cd = (double)c, dd = (double)den;
    const double q = __dmul_rn(cd, r);
    return __fma_rn(__fma_rn(-q, dd, cd), r, q);
}

// carry-save adder on one 32-bit slice: (h, l) = a + b + c
#define PF_CSA(h, l, a, b, c)                       \
    do {                                            \
        const uint32_t u__ = (a) ^ (b);             \
        const uint32_t c__ = (c);                   \
        (h) = ((a) & (b)) | (u__ & c__);            \
        (l) = u__ ^ c__;                            \
    } while (0)

struct Mask64 { uint32_t lo, hi; };

#ifndef PF_IMAD_HI
#define PF_IMAD_HI 0
#endif
#ifndef PF_WALK_UNROLL
#define PF_WALK_UNROLL 0     // 1: 16 lookups per walk-loop iteration, ping-pong chunk registers (A/B variant)
#endif

// x = slab_encode(v): bitmap word at byte offset x>>5; shifting it left by x&31 puts v's presence bit in
// bit 31 and the presence bits of smaller tetramers below it (the rank half-word falls off the top).
// A miss (57 % of the lookups on the synthetic family) adds the zero mask and issues no second load.
// bm32 / masks32: 32-bit shared-space addresses of the stage's bitmap and mask array.
__device__ __forceinline__ Mask64 dict_lookup(uint32_t bm32, uint32_t masks32, uint32_t x, uint32_t k27 = 0u, uint32_t k16 = 0u) {
    uint32_t w;
#if PF_IMAD_HI
    // shifts done as the high half of a multiply-add: IMAD.HI runs on the FMA pipe, LEA.HI on the (half-rate, limiting) ALU pipe
    uint32_t a;
    asm("mad.hi.u32 %0, %1, %2, %3;" : "=r"(a) : "r"(x), "r"(k27), "r"(bm32));        // bm32 + (x >> 5), k27 = 2^27 (opaque to ptxas)
    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(w) : "r"(a));
    const uint32_t t = __funnelshift_l(0u, w, x);                   // w << (x & 31)
    uint32_t idx;
    asm("mad.hi.u32 %0, %1, %2, %3;" : "=r"(idx) : "r"(w), "r"(k16), "r"(__popc(t))); // (w >> 16) + popc(t), k16 = 2^16
#else
    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(w) : "r"(bm32 + (x >> 5)));
    const uint32_t t = __funnelshift_l(0u, w, x);                   // w << (x & 31)
    const uint32_t idx = (w >> 16) + __popc(t);                     // rank of the word + #present <= v  (>= 1 on a hit)
#endif
    Mask64 r; r.lo = 0u; r.hi = 0u;
    if ((int32_t)t < 0)
        asm volatile("ld.shared.v2.u32 {%0, %1}, [%2];" : "=r"(r.lo), "=r"(r.hi) : "r"(masks32 + 8u * idx));
    return r;
}

// a StageDesc travels through the kernel as the raw uint4 {off.lo, off.hi, bytes, p | q<<16 | flags<<24}
__device__ __forceinline__ uint4 load_stage(const StageDesc *p) { return __ldg(reinterpret_cast<const uint4 *>(p)); }
__device__ __forceinline__ int st_p(uint32_t w) { return (int)(w & 0xffffu); }
__device__ __forceinline__ int st_q(uint32_t w) { return (int)((w >> 16) & 0xffu); }
__device__ __forceinline__ bool st_first(uint32_t w) { return (w >> 24) & 1u; }
__device__ __forceinline__ bool st_last(uint32_t w) { return (w >> 25) & 1u; }

// NPL = bit planes of a half-list counter, NPN = bit planes of the N counter (7: up to 127 SCPs, 10: up to 1023)
template <int NPL, int NPN>
__global__ void __launch_bounds__(TILE_THREADS, 1) aji_tile_kernel(const TileParams prm) {
    constexpr int N_PLANES_N = NPN;
    extern __shared__ __align__(128) unsigned char smem[];
    // ---- shared memory carve-up: [misc SMEM_MISC][stage 0][stage 1] ----
    uint64_t *full = reinterpret_cast<uint64_t *>(smem);               // [2] mbarriers: blob landed
    uint32_t *done = reinterpret_cast<uint32_t *>(smem + 16);          // [2] warps that finished walking the slot
    int32_t *col_g = reinterpret_cast<int32_t *>(smem + 64);           // [64] genome id of the tile's columns (-1 = none)
    int32_t *colT = reinterpret_cast<int32_t *>(smem + 64 + 256);      // [4][64] ring of T(p, column), slot p & 3
    uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(smem + 32);     // tensor-memory base address (tcgen05.alloc)
    double *rcp = reinterpret_cast<double *>(smem + SMEM_MISC);        // [RCP_N] RN(1/den)
    unsigned char *stage_mem = smem + SMEM_MISC + SMEM_RCP;

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int P = prm.P;
    const int r = lane >> 1, h = lane & 1;
    const int cbase = 32 * h;

    if (tid == 0) {
        mbar_init(&full[0], 1);
        mbar_init(&full[1], 1);
        done[0] = 0u; done[1] = 0u;
        reinterpret_cast<int *>(smem + 40)[1] = 0;   // tile list in use
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 0) tmem_alloc(tmem_slot, TMEM_COLS);
    for (int i = tid; i < RCP_N; i += TILE_THREADS) rcp[i] = (i > 0) ? 1.0 / (double)i : 0.0;
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    uint32_t stage32;   // shared-space address of the stage slots (volatile asm: computed once, never rematerialised in the walk loop)
    asm volatile("{\n.reg .u64 t64;\ncvta.to.shared.u64 t64, %1;\ncvt.u32.u64 %0, t64;\n}" : "=r"(stage32) : "l"(stage_mem));
    // this thread's 64 tensor-memory columns (32 doubles): lane quadrant warp%4, column block warp/4
    const uint32_t taddr = *tmem_slot + ((uint32_t)((warp & 3) * 32) << 16) + (uint32_t)((warp >> 2) * TMEM_COLS_PER_WARP);
    uint32_t k27 = 0u, k16 = 0u;    // shift amounts as multipliers (PF_IMAD_HI), hidden from constant propagation
#if PF_IMAD_HI
    k27 = prm.k27; k16 = prm.k16;   // kernel parameters: ptxas cannot fold them back into shifts
#endif
    uint32_t use_par = 0u;          // bit s = wait parity of slot s (flips per completed fill), identical in every thread

    auto issue = [&](const uint4 st, int slot, int cslot) {
        // runs in ONE thread: refill `slot` with the blob of stage `st` (+ the column T values on an SCP's first stage)
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        const uint32_t extra = st_first(st.w) ? COLS_PER_TILE * 4u : 0u;
        const int64_t off = (int64_t)(((uint64_t)st.y << 32) | st.x);
        mbar_expect_tx(&full[slot], st.z + extra);
        for (uint32_t o = 0; o < st.z; o += prm.piece)
            bulk_g2s(stage_mem + (size_t)slot * STAGE_BYTES + o, prm.blob + off + o, min(prm.piece, st.z - o), &full[slot]);
        if (extra)
            bulk_g2s(colT + (st_p(st.w) & 3) * COLS_PER_TILE, prm.colT + ((int64_t)cslot * P + st_p(st.w)) * COLS_PER_TILE, extra,
                     &full[slot]);
    };

    // tiles are handed out in list order by a global counter (zeroed by the host): CTAs that start late -- e.g. behind the
    // collective kernels of a dictionary exchange that is still running -- simply take fewer tiles
    int *s_tile = reinterpret_cast<int *>(smem + 40);      // [0] tile code: index | list << 30, or -1 = done; [1] current list
    for (;;) {
        if (tid == 0) {
            int code = -1;
            if (s_tile[1] == 0) {
                const int t = atomicAdd(prm.tile_counter, 1);
                if (t < prm.n_tiles) code = t; else s_tile[1] = 1;
            }
            if (code < 0 && prm.n_tiles2 > 0) {
                const int t = atomicAdd(prm.tile_counter2, 1);
                if (t < prm.n_tiles2) code = t | (1 << 30);
            }
            s_tile[0] = code;
        }
        __syncthreads();
        const int code = s_tile[0];     // rewritten only after the barrier that ends this tile
        if (code < 0) break;
        const int4 tile = (code >> 30) ? prm.tiles2[code & 0x3fffffff] : prm.tiles[code];
        const int cb = tile.y, cslot = tile.w;
        const int64_t sb = prm.stage_rng[2 * cslot];
        const int nst = (int)(prm.stage_rng[2 * cslot + 1] - sb);
        const int group = tile.x * TILE_WARPS + warp;
        const bool group_ok = group < prm.n_groups;
        const int row = group * ROWS_PER_WARP + r;
        const bool row_ok = row < prm.n_rows;
        const int gi = row_ok ? prm.rows[row] : -1;
        const int ti = row_ok ? prm.row_t[row] : 0;
        const int j0 = cb * COLS_PER_TILE;
        const int ncol = min(COLS_PER_TILE, prm.n_cols - j0);

        if (tid == 0) {
            if (nst > 0) issue(load_stage(prm.stages + sb), 0, cslot);
            if (nst > 1) issue(load_stage(prm.stages + sb + 1), 1, cslot);
        }
        if (tid < COLS_PER_TILE) col_g[tid] = (tid < ncol) ? prm.cols[j0 + tid] : -1;
        __syncthreads();

        // columns of this lane that are real pairs: inside the block and not the row's own genome (isValidPair)
        uint32_t valid = 0u;
        if (row_ok) {
#pragma unroll
            for (int jj = 0; jj < 32; jj++) {
                const int gj = col_g[cbase + jj];
                if (gj >= 0 && gj != gi) valid |= 1u << jj;
            }
        }

        uint32_t Np[N_PLANES_N];
#pragma unroll
        for (int k = 0; k < N_PLANES_N; k++) Np[k] = 0u;
        {   // S = 0 for the 32 columns of this lane
            uint32_t z[16];
#pragma unroll
            for (int k = 0; k < 16; k++) z[k] = 0u;
            __syncwarp();
#pragma unroll
            for (int g8 = 0; g8 < 4; g8++) tmem_st16(taddr + 16 * g8, z);
            tmem_wait_st();
        }

        uint32_t plo[NPL], phi[NPL];     // bit-sliced counters of this lane's half list: columns 0..31 / 32..63
        uint32_t pe_lo = 0u, pe_hi = 0u; // pending carry of weight 8 (one per two 8-lookup steps)
        bool pending = false;
#pragma unroll
        for (int k = 0; k < NPL; k++) { plo[k] = 0u; phi[k] = 0u; }

        // ---- software pipeline state: descriptors two stages ahead, slab extents one stage ahead ----
        uint32_t st_cur = 0u, st_nxt = 0u;     // p | q << 16 | flags << 24 of stages k and k + 1
        if (nst > 0) st_cur = load_stage(prm.stages + sb).w;
        if (nst > 1) st_nxt = load_stage(prm.stages + sb + 1).w;
        const int64_t slab0 = (int64_t)group * P * NRANGE;
        int64_t o_cur = 0; int n_cur = 0;
        if (group_ok && nst > 0) {
            const int64_t gi0 = slab0 + (int64_t)st_p(st_cur) * NRANGE + st_q(st_cur);
            o_cur = prm.ell_off[gi0];
            n_cur = (int)(prm.ell_off[gi0 + 1] - o_cur);
        }
        // the slabs of a row group lie in stage order, so the walk streams them with ONE running pointer: cur0/cur1 hold the two
        // chunks of the next step, rp points behind them; only a stage that re-walks its slab (column sub-split) repositions it
        uint4 cur0 = make_uint4(0, 0, 0, 0), cur1 = cur0;
        const uint4 *rp = prm.ell + lane;
        int64_t o_fetched = -1;                 // chunk index the running pointer has prefetched up to (start of cur0)
        int ta = (row_ok && nst > 0) ? prm.T[(int64_t)st_p(st_cur) * prm.t_ld + ti] : 0;

        for (int k = 0; k < nst; k++) {
            const int slot = k & 1;
            uint4 st_n2 = make_uint4(0u, 0u, 0u, 0u);
            if (k + 2 < nst) st_n2 = load_stage(prm.stages + sb + k + 2);
            int64_t o_nxt = 0; int n_nxt = 0;
            if (group_ok && k + 1 < nst) {
                const int64_t gi1 = slab0 + (int64_t)st_p(st_nxt) * NRANGE + st_q(st_nxt);
                o_nxt = prm.ell_off[gi1];
                n_nxt = (int)(prm.ell_off[gi1 + 1] - o_nxt);
            }
            int ta_nxt = 0;
            if (st_last(st_cur) && row_ok && k + 1 < nst) ta_nxt = prm.T[(int64_t)st_p(st_nxt) * prm.t_ld + ti];

            mbar_wait(&full[slot], (use_par >> slot) & 1u);
            use_par ^= 1u << slot;

            // ================= walk: bit-sliced accumulation of the row half-lists =================
            {
                const uint32_t bm = stage32 + (uint32_t)slot * STAGE_BYTES;
                const uint32_t masks = bm + BM_BYTES;
                if (n_cur > 0 && o_cur != o_fetched) {
                    rp = prm.ell + o_cur * 32 + lane;
                    cur0 = __ldg(rp); cur1 = __ldg(rp + 32);
                    rp += 64;
                }
                o_fetched = o_cur + n_cur;
#if PF_WALK_UNROLL
                // 8 lookups (two 512-byte chunks v0, v1) -> one carry of weight 8 in (e_lo, e_hi)
#define PF_STEP8(v0, v1, e_lo, e_hi)                                                                                         \
                    do {                                                                                                     \
                        uint32_t f_lo[2], f_hi[2];                                                                           \
                        _Pragma("unroll") for (int qq = 0; qq < 2; qq++) {                                                   \
                            const uint4 v = qq ? (v1) : (v0);                                                                \
                            const Mask64 m0 = dict_lookup(bm, masks, v.x, k27, k16), m1 = dict_lookup(bm, masks, v.y, k27, k16); \
                            const Mask64 m2 = dict_lookup(bm, masks, v.z, k27, k16), m3 = dict_lookup(bm, masks, v.w, k27, k16); \
                            uint32_t ta_lo, ta_hi, tb_lo, tb_hi;                                                             \
                            PF_CSA(ta_lo, plo[0], plo[0], m0.lo, m1.lo); PF_CSA(ta_hi, phi[0], phi[0], m0.hi, m1.hi);        \
                            PF_CSA(tb_lo, plo[0], plo[0], m2.lo, m3.lo); PF_CSA(tb_hi, phi[0], phi[0], m2.hi, m3.hi);        \
                            PF_CSA(f_lo[qq], plo[1], plo[1], ta_lo, tb_lo); PF_CSA(f_hi[qq], phi[1], phi[1], ta_hi, tb_hi);  \
                        }                                                                                                    \
                        PF_CSA(e_lo, plo[2], plo[2], f_lo[0], f_lo[1]); PF_CSA(e_hi, phi[2], phi[2], f_hi[0], f_hi[1]);      \
                    } while (0)
                // two eights -> plane 3, the sixteens carry ripples into planes 4..
#define PF_FOLD16(a_lo, a_hi, b_lo, b_hi)                                                                                    \
                    do {                                                                                                     \
                        uint32_t cl, ch;                                                                                     \
                        PF_CSA(cl, plo[3], plo[3], a_lo, b_lo); PF_CSA(ch, phi[3], phi[3], a_hi, b_hi);                      \
                        _Pragma("unroll") for (int kk = 4; kk < NPL; kk++) {                                                 \
                            const uint32_t tl = plo[kk] & cl, th = phi[kk] & ch;                                             \
                            plo[kk] ^= cl; phi[kk] ^= ch; cl = tl; ch = th;                                                  \
                        }                                                                                                    \
                    } while (0)
                int c = 0;
                if (pending && n_cur > 0) {          // warp-uniform: pair the carry left over by the previous stage
                    const uint4 nx0 = __ldg(rp), nx1 = __ldg(rp + 32);
                    rp += 64;
                    uint32_t e_lo, e_hi;
                    PF_STEP8(cur0, cur1, e_lo, e_hi);
                    PF_FOLD16(pe_lo, pe_hi, e_lo, e_hi);
                    pending = false;
                    cur0 = nx0; cur1 = nx1;
                    c = 2;
                }
                for (; c + 4 <= n_cur; c += 4) {     // 16 lookups per iteration, chunk registers ping-pong (no moves, no flag)
                    const uint4 nx0 = __ldg(rp), nx1 = __ldg(rp + 32);
                    uint32_t e1_lo, e1_hi, e2_lo, e2_hi;
                    PF_STEP8(cur0, cur1, e1_lo, e1_hi);
                    cur0 = __ldg(rp + 64); cur1 = __ldg(rp + 96);
                    rp += 128;
                    PF_STEP8(nx0, nx1, e2_lo, e2_hi);
                    PF_FOLD16(e1_lo, e1_hi, e2_lo, e2_hi);
                }
                if (c < n_cur) {                     // one step left: its carry waits for the next stage (or the SCP's end)
                    const uint4 nx0 = __ldg(rp), nx1 = __ldg(rp + 32);
                    rp += 64;
                    PF_STEP8(cur0, cur1, pe_lo, pe_hi);
                    pending = true;
                    cur0 = nx0; cur1 = nx1;
                }
#undef PF_STEP8
#undef PF_FOLD16
#else
                for (int c = 0; c < n_cur; c += 2) {
                    // prefetch the next step (possibly the next stage's first chunks; the slab buffer is padded at its end)
                    const uint4 nx0 = __ldg(rp), nx1 = __ldg(rp + 32);
                    rp += 64;
                    // 8 lookups -> one carry of weight 8
                    uint32_t f_lo[2], f_hi[2];
#pragma unroll
                    for (int qq = 0; qq < 2; qq++) {
                        const uint4 v = qq ? cur1 : cur0;
                        const Mask64 m0 = dict_lookup(bm, masks, v.x, k27, k16), m1 = dict_lookup(bm, masks, v.y, k27, k16);
                        const Mask64 m2 = dict_lookup(bm, masks, v.z, k27, k16), m3 = dict_lookup(bm, masks, v.w, k27, k16);
                        uint32_t ta_lo, ta_hi, tb_lo, tb_hi;
                        PF_CSA(ta_lo, plo[0], plo[0], m0.lo, m1.lo); PF_CSA(ta_hi, phi[0], phi[0], m0.hi, m1.hi);
                        PF_CSA(tb_lo, plo[0], plo[0], m2.lo, m3.lo); PF_CSA(tb_hi, phi[0], phi[0], m2.hi, m3.hi);
                        PF_CSA(f_lo[qq], plo[1], plo[1], ta_lo, tb_lo); PF_CSA(f_hi[qq], phi[1], phi[1], ta_hi, tb_hi);
                    }
                    uint32_t e_lo, e_hi;
                    PF_CSA(e_lo, plo[2], plo[2], f_lo[0], f_lo[1]); PF_CSA(e_hi, phi[2], phi[2], f_hi[0], f_hi[1]);
                    if (!pending) {           // warp-uniform: keep the eights carry for the next step
                        pe_lo = e_lo; pe_hi = e_hi; pending = true;
                    } else {                  // two eights -> plane 3, the sixteens carry ripples into planes 4..
                        uint32_t cl, ch;
                        PF_CSA(cl, plo[3], plo[3], pe_lo, e_lo); PF_CSA(ch, phi[3], phi[3], pe_hi, e_hi);
#pragma unroll
                        for (int kk = 4; kk < NPL; kk++) {
                            const uint32_t tl = plo[kk] & cl, th = phi[kk] & ch;
                            plo[kk] ^= cl; phi[kk] ^= ch; cl = tl; ch = th;
                        }
                        pending = false;
                    }
                    cur0 = nx0; cur1 = nx1;
                }
#endif
            }
            // ================= release the slot; the last warp to leave refills it =================
            __syncwarp();
            if (lane == 0) {
                const uint32_t old = atom_add_acq_rel_smem(&done[slot], 1u);
                if (old == TILE_WARPS - 1) {
                    done[slot] = 0u;
                    if (k + 2 < nst) issue(st_n2, slot, cslot);
                }
            }

            if (st_last(st_cur)) {
                // ================= SCP finished: fold the pending carry, pair-reduce, epilogue =================
                const int p = st_p(st_cur);
                if (pending) {
                    uint32_t cl = pe_lo, ch = pe_hi;
#pragma unroll
                    for (int kk = 3; kk < NPL; kk++) {
                        const uint32_t tl = plo[kk] & cl, th = phi[kk] & ch;
                        plo[kk] ^= cl; phi[kk] ^= ch; cl = tl; ch = th;
                    }
                    pending = false;
                }
                // lane h keeps columns [32h, 32h+32) of both halves of the row
                uint32_t R[NPL + 1];
                {
                    uint32_t carry = 0u;
#pragma unroll
                    for (int kk = 0; kk < NPL; kk++) {
                        const uint32_t mine = h ? phi[kk] : plo[kk];
                        const uint32_t send = h ? plo[kk] : phi[kk];
                        const uint32_t theirs = __shfl_xor_sync(0xffffffffu, send, 1);
                        const uint32_t u = mine ^ theirs;
                        R[kk] = (u ^ carry) & valid;
                        carry = (mine & theirs) | (u & carry);
                        plo[kk] = 0u; phi[kk] = 0u;
                    }
                    R[NPL] = carry & valid;
                }
                // N += (c > 0), bit-sliced
                {
                    uint32_t nz = 0u;
#pragma unroll
                    for (int kk = 0; kk <= NPL; kk++) nz |= R[kk];
#pragma unroll
                    for (int kk = 0; kk < N_PLANES_N; kk++) { const uint32_t tt = Np[kk] & nz; Np[kk] ^= nz; nz = tt; }
                }
                const uint32_t ct32 = smem_u32(colT) + 4u * (uint32_t)((p & 3) * COLS_PER_TILE + cbase);
                const uint32_t rcp32 = smem_u32(rcp);
                __syncwarp();
                // 8 columns at a time: fetch their sums from tensor memory, expand the bit-sliced counts
                // (nibble -> one bit per byte: (n * 0x00204081) & 0x01010101) while the load is in flight, add J, store back
#pragma unroll
                for (int g8 = 0; g8 < 4; g8++) {
                    uint32_t sv[16];
                    tmem_ld16(taddr + 16 * g8, sv);
                    int cnt[8], den[8];
#pragma unroll
                    for (int hf = 0; hf < 2; hf++) {
                        const int g4 = 2 * g8 + hf;
                        uint32_t lo = 0u, hi = 0u;
#pragma unroll
                        for (int kk = 0; kk <= NPL; kk++) {
                            const uint32_t sp = (((R[kk] >> (4 * g4)) & 15u) * 0x00204081u) & 0x01010101u;
                            if (kk < 8) lo += sp << kk; else hi += sp << (kk - 8);
                        }
#pragma unroll
                        for (int u = 0; u < 4; u++)
                            cnt[4 * hf + u] = (int)(((lo >> (8 * u)) & 0xffu) | (((hi >> (8 * u)) & 0xffu) << 8));
                    }
                    bool slow = false;   // a denominator beyond the table or c > den (only the query-DB label quirk can do that)
#pragma unroll
                    for (int u = 0; u < 8; u++) {
                        den[u] = max(ta + lds_s32(ct32 + 4u * (uint32_t)(8 * g8 + u)) - cnt[u], 1);
                        slow |= (den[u] >= RCP_N) | (cnt[u] > den[u]);
                    }
                    if (prm.counts_out) {                                   // small-N hook, warp-uniform
#pragma unroll
                        for (int u = 0; u < 8; u++) {
                            const int col = cbase + 8 * g8 + u;
                            if (row_ok && col < ncol) prm.counts_out[((int64_t)row * prm.n_cols + (j0 + col)) * P + p] = cnt[u];
                        }
                    }
                    tmem_wait_ld(sv);
                    // J = c / den, algorithm_impl.hpp:250-254 (c == 0 adds +0.0); one warp-uniform branch per 8 columns
                    if (!__any_sync(0xffffffffu, slow)) {
#pragma unroll
                        for (int u = 0; u < 8; u++) {
                            const double sum = __hiloint2double((int)sv[2 * u + 1], (int)sv[2 * u]) + jaccard_fast(rcp32, cnt[u], den[u]);
                            sv[2 * u] = (uint32_t)__double2loint(sum);
                            sv[2 * u + 1] = (uint32_t)__double2hiint(sum);
                        }
                    } else {
#pragma unroll
                        for (int u = 0; u < 8; u++) {
                            const double sum = __hiloint2double((int)sv[2 * u + 1], (int)sv[2 * u]) + (double)cnt[u] / (double)den[u];
                            sv[2 * u] = (uint32_t)__double2loint(sum);
                            sv[2 * u + 1] = (uint32_t)__double2hiint(sum);
                        }
                    }
                    tmem_st16(taddr + 16 * g8, sv);
                }
                tmem_wait_st();
                ta = ta_nxt;
            }
            st_cur = st_nxt; st_nxt = st_n2.w;
            o_cur = o_nxt; n_cur = n_nxt;
        }

        // ================= output: AJI = S / N (computeAJI) into the dense matrix =================
        // The tile (384 x 64 doubles) is staged in the two (now idle) stage slots and written out in contiguous
        // segments: rows of 512 B for the direct cells, columns of 3 KB for the transposed ones (the mirror image in
        // all-vs-all mode, everything in the query modes whose output is the kernel's transpose).  `out` may be
        // device memory or mapped pinned host memory (pfaai_gpu_compute with a pinned matrix: the D2H copy is the
        // kernel's own store stream and overlaps the remaining tiles).
        double *stg = reinterpret_cast<double *>(stage_mem);                          // [TILE_ROWS][STG_PITCH]
        int32_t *row_g = reinterpret_cast<int32_t *>(stage_mem + (size_t)TILE_ROWS * STG_PITCH * 8);   // [TILE_ROWS]
        __syncthreads();   // every warp has finished walking both slots
        {
            const int rl = warp * ROWS_PER_WARP + r;
            if (h == 0) row_g[rl] = gi;
            __syncwarp();
#pragma unroll
            for (int g8 = 0; g8 < 4; g8++) {
                uint32_t sv[16];
                tmem_ld16(taddr + 16 * g8, sv);
                tmem_wait_ld(sv);
#pragma unroll
                for (int u = 0; u < 8; u++) {
                    const int jj = 8 * g8 + u;
                    const int col = cbase + jj;
                    const double Ssum = __hiloint2double((int)sv[2 * u + 1], (int)sv[2 * u]);
                    int n = 0;
#pragma unroll
                    for (int kk = 0; kk < N_PLANES_N; kk++) n |= (int)((Np[kk] >> jj) & 1u) << kk;
                    stg[rl * STG_PITCH + col] = (n > 0) ? Ssum / (double)n : 0.0;
                    if (row_ok && col < ncol) {
                        if (prm.S_out) prm.S_out[(int64_t)row * prm.n_cols + (j0 + col)] = Ssum;
                        if (prm.N_out) prm.N_out[(int64_t)row * prm.n_cols + (j0 + col)] = n;
                    }
                }
            }
        }
        __syncthreads();
        {
            const int row0 = tile.x * TILE_ROWS;
            const int nrow = min(TILE_ROWS, prm.n_rows - row0);
            const int64_t ro0 = (int64_t)tile.z;                       // output row of the tile's first row
            if (prm.out_cs == 1) {
                // direct cells, one 512-B row segment per warp instruction pair; symmetric: only gi < gj, diagonal = 0
                for (int rl = warp; rl < nrow; rl += TILE_WARPS) {
                    const int g_i = row_g[rl];
                    double *dst = prm.out + (ro0 + rl) * prm.out_rs + j0;
#pragma unroll
                    for (int c = lane; c < COLS_PER_TILE; c += 32) {
                        if (c < ncol) {
                            const int g_j = col_g[c];
                            const double v = stg[rl * STG_PITCH + c];
                            if (prm.symmetric) { if (g_i <= g_j) dst[c] = (g_i == g_j) ? 0.0 : v; }
                            else dst[c] = (g_i == g_j) ? 0.0 : v;
                        }
                    }
                }
            }
            if (prm.out_cs != 1 || prm.mirror) {
                // transposed cells: for a fixed column the tile's rows are contiguous in the output
                const int64_t cstride = (prm.out_cs != 1) ? prm.out_cs : prm.out_rs;   // step between columns j
                const int64_t rbase = (prm.out_cs != 1) ? ro0 : (int64_t)row0;          // mirror: indexed by genome id
                for (int c = warp; c < ncol; c += TILE_WARPS) {
                    const int g_j = col_g[c];
                    double *dst = prm.out + (int64_t)(j0 + c) * cstride + rbase;
                    for (int rl = lane; rl < nrow; rl += 32) {
                        const int g_i = row_g[rl];
                        const double v = stg[rl * STG_PITCH + c];
                        if (prm.symmetric) { if (g_i < g_j) dst[rl] = v; }
                        else dst[rl] = (g_i == g_j) ? 0.0 : v;
                    }
                }
            }
        }
        if (prm.progress) __threadfence_system();   // this thread's stores (device or mapped host memory) before the counters move
        __syncthreads();   // staging, col_g and colT are free again before the next tile refills the slots
        if (prm.progress && tid == 0) {
            constexpr int GPR = TILE_ROWS / COLS_PER_TILE;
            for (int k = 0; k < GPR; k++) { const int g = tile.x * GPR + k; if (g < prm.n_gran) atomicAdd(prm.progress + g, 1); }
            if (prm.mirror) atomicAdd(prm.progress + tile.y, 1);
        }
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 0) tmem_dealloc(*tmem_slot, TMEM_COLS);
}

}  // namespace pfaai
