// aji_mma.cuh -- K5 of DESIGN.md: the dense-contraction tile path (tcgen05 tensor cores) of the AJI hot path.
//
// Replaces the same reference loci as aji_tile.cuh (constructTetramerTuples ds_helper.hpp:270-357, parallelMergeSort
// psort.hpp:27-53, findEBlockExtents / computeEBlockJAC / computeAJI algorithm_impl.hpp:123-322, printOutput's matrix fill
// main.cpp:143-154) and is selected instead of the list path where the per-SCP tetramer density makes the intersection count a
// dense contraction (BASELINE north_star; the decision rule and the evidence are in DESIGN.md section 4b).
//
// Formulation.  Per SCP p the tetramers whose posting list is at least `tau` genomes long are the DENSE columns 0..K_p-1.
// With X_p[g][k] = 1 iff genome g holds dense tetramer k of SCP p (one unsigned byte), the intersection count restricted to
// dense tetramers is the integer matrix product  C_p = X_p(rows) * X_p(cols)^T  -- computed exactly by tcgen05.mma kind::i8
// (u8 x u8 -> s32 accumulators in tensor memory).  Every co-occurrence on a NON-dense tetramer (short lists) is counted in the
// same kernel by the SPARSE WALK (sparse_index.cuh): while the tensor cores work on SCP p, the epilogue warps stream the tile
// rows' non-dense tetramers through the column block's sparse dictionary (presence bitmap + rank -> column entry, one bulk copy
// per (tile, SCP)) and bump 16-bit shared-memory counters that the epilogue adds to C_p -- so the counts are bit-exact for any
// input and any tau; the cost model in pfaai_gpu.cu picks tau (and falls back to the list path when that is cheaper).
//
//   tile      : 128 rows x 128 columns of genome pairs, one persistent CTA per SM, SCPs in ascending order (the reference's
//               summation order).
//   operands  : X is stored in HBM as ready-to-use shared-memory images, one per (128-genome block, SCP): K-major, no swizzle,
//               core matrix = 8 genomes x 16 dense tetramers (128 contiguous bytes); byte (r, k) of an image sits at
//               (k/16)*2048 + r*16 + k%16, so any K-range of all 128 genomes is ONE contiguous bulk copy (cp.async.bulk,
//               SASS UBLKCP) -- no tensor maps.  A stage = 64 dense tetramers of the row block + of the column block (16 KB),
//               4 stages.
//   roles     : warp 16 lane 0 = bulk-copy producer of the operand stages and the column T values; warp 17 lane 0 = tcgen05.mma
//               issuer (M = N = 128, K = 32 per instruction, accumulators double-buffered in tensor memory: 2 x 128 columns);
//               warp 18 lane 0 = bulk-copy producer of the sparse dictionaries (one 96 KB stage); warps 0..15 = sparse walk +
//               epilogue.  mbarrier pipelines: full/empty per stage, acc_full/acc_empty per accumulator buffer, ct_full/ct_empty
//               for the ring of column T values, dict_full/dict_empty for the dictionary stage.
//   walk      : the 4 warps of a tensor-memory lane quadrant (32 rows) share the quadrant's row items of the SCP, 128 items per
//               warp step (coalesced LDG.128): item -> LDS.32 bitmap word -> (hit) LDS.U16 rank + POPC -> LDS.U16 entry ->
//               RED.shared on the (row, column) counter.  The dictionary of SCP p+1 arrives while the epilogue of SCP p runs.
//   epilogue  : thread = (row = tensor-memory lane, 32 columns).  tcgen05.ld brings the dense counts, the quadrant's sparse
//               counters are added (and cleared), then J = c/(Ta+Tb-c) (reciprocal table + two FMAs, correctly rounded:
//               tests/test_division.py), S += J in fp64 (tensor memory), N += (c > 0).
//               While the walk and the epilogue of SCP p run, the tensor cores already work on SCP p+1.
//   output    : AJI = S/N staged through shared memory in four 32-column rounds and written as contiguous row segments
//               (and column segments for the mirror image / the transposed query modes), device memory or pinned host memory.
#pragma once
#include "aji_tile.cuh"
#include "sparse_index.cuh"

namespace pfaai {

constexpr int MM_ROWS = 128, MM_COLS = 128;          // tile
constexpr int MM_EPI_WARPS = 16;
constexpr int MM_THREADS = (MM_EPI_WARPS + 3) * 32;   // + operand producer warp + MMA warp + dictionary producer warp
constexpr int MM_STAGE_K = 64;                        // dense tetramers per pipeline stage
constexpr int MM_OP_BYTES = MM_STAGE_K * 128;         // one operand of one stage: 128 genomes x 64 k = 8 KB
constexpr int MM_STAGE_BYTES = 2 * MM_OP_BYTES;
// Shared-memory plan, two variants chosen per compute call (what is not fixed is split between the operand ring, the ring of
// row items and the dictionary stage): sparse inputs walk many items through large dictionaries and contract few dense
// tetramers; dense inputs stream large operands and walk almost nothing.
template <int NSTAGE_, int ENT_MAX_, int RING_D_> struct MmCfg {
    static constexpr int NSTAGE = NSTAGE_;                        // operand pipeline stages
    static constexpr int ENT_MAX = ENT_MAX_;                      // dictionary stage capacity (sparse_index.cuh)
    static constexpr int RING_D = RING_D_;                        // row-item chunks in flight per walker warp (cp.async ring)
    static constexpr int DICT_BYTES = sp_stage_bytes(ENT_MAX_);
    static constexpr int RING_BYTES = 16 * RING_D_ * 512;
    static constexpr int STAGES_BYTES = NSTAGE_ * 2 * 64 * 128;
};
using MmCfgSparse = MmCfg<3, SP_ENT_MAX, 4>;
using MmCfgDense = MmCfg<7, SP_ENT_MAX_SMALL, 2>;
constexpr int MM_CT_RING = 4;                         // ring of column T values (128 x int32 per SCP)
constexpr int MM_CNT_PITCH = 272;                     // bytes per row of the sparse counters (128 x uint16 + pad: conflict-free LDS.128)
constexpr int MM_CNT_QUAD = 32 * MM_CNT_PITCH;        // 8704 B per lane quadrant (32 rows)
constexpr int MM_STG_PITCH = 33;                      // output staging [128][33] doubles (one 32-column round)
constexpr int MM_SMEM_MISC = 1024;
constexpr int MM_SMEM_CT = MM_CT_RING * MM_COLS * 4;
constexpr int MM_SMEM_CNT = 4 * MM_CNT_QUAD;              // 34816 >= 128*33*8 = 33792 (staging shares it)
static_assert(MM_SMEM_CNT >= MM_ROWS * MM_STG_PITCH * 8, "output staging does not fit the sparse-counter area");
constexpr int MM_SMEM_RCP = SMEM_RCP;                   // RN(1/den) per denominator
constexpr int MM_SMEM_KP = 1024;                        // kpad[p] of every SCP (P <= 255)
constexpr int MM_SMEM_FIXED = MM_SMEM_MISC + MM_SMEM_CT + MM_SMEM_RCP + MM_SMEM_KP + MM_SMEM_CNT;
template <class Cfg> constexpr int mm_smem_total() { return MM_SMEM_FIXED + Cfg::STAGES_BYTES + Cfg::RING_BYTES + Cfg::DICT_BYTES; }
static_assert(mm_smem_total<MmCfgSparse>() <= SMEM_TOTAL && mm_smem_total<MmCfgDense>() <= SMEM_TOTAL, "K5 shared memory");
constexpr int MM_TMEM_COLS = 512;                     // 2 accumulator buffers x 128 columns (s32) + 256 columns of fp64 sums S
constexpr int MM_TMEM_S = 256;                        // first column of S: thread (lane, column group cg) owns columns S + 64 cg .. + 63
constexpr uint16_t MM_NOT_DENSE = 0xffffu;

struct MmaParams {
    // operand images: block b of the row / column list, SCP p -> img + b*blk_stride + koff[p]*128, kpad[p]*128 bytes
    const unsigned char *img_rows, *img_cols;
    int64_t blk_stride;                 // bytes per 128-genome block = 128 * sum(kpad)
    const int32_t *kpad;                // [P] dense tetramers of SCP p, padded to a multiple of 32 (0: none)
    const int64_t *koff;                // [P] exclusive prefix sum of kpad
    // sparse side (sparse_index.cuh): row items of (32-row group g, SCP p) = sp_items[sp_off[g*P+p]*128 ...), chunks of 128;
    // dictionaries of (column block J, SCP p) = sp_sub[(J*P + p)*SP_MAXSUB + 0 .. sp_nsub[J*P+p]) inside sp_blob
    const uint32_t *sp_items;
    const int64_t *sp_off;
    const int32_t *sp_nsub;
    const SpSub *sp_sub;
    const unsigned char *sp_blob;
    int has_sparse;
    // rows / columns (kernel index -> genome id, T column)
    const int32_t *rows, *row_t, *cols, *col_t;
    int n_rows, n_cols, ncb;            // ncb = column blocks of 128
    const int32_t *T;                   // [P x t_ld]
    const int32_t *colT;                // [ncb][P][128]: T(p, col_t[column]) (0 beyond n_cols)
    const int32_t *rowTmax, *colTmax;   // [nrb][P], [ncb][P]: max T over the block (fast-path guard: sum < RCP_N)
    int t_ld, P;
    int quirk;                          // 1: T labels are not the genomes' own (query-DB label quirk): always the guarded division
    int all_fast;                       // 1: 2 * max(T) < RCP_N over the whole matrix: no per-block guard loads
    // output (same conventions as TileParams)
    double *out;
    int64_t out_rs, out_cs;
    int symmetric, mirror;
    double *S_out; int32_t *N_out; int32_t *counts_out;
    const int4 *tiles;                  // (row block I, column block J, first output row of block I, unused)
    int n_tiles;
    int *progress;                      // streaming output: finished tiles per 128-row granule (see TileParams::progress); NULL = off
};

__device__ __forceinline__ void mbar_arrive(uint64_t *bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
// the mbarrier receives one arrival when every tcgen05 operation issued so far by this thread has completed
__device__ __forceinline__ void tc_commit(uint64_t *bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
// shared-memory matrix descriptor, K-major, no swizzle: core matrix (8 rows x 16 B) contiguous, next 8 rows +128 B (SBO),
// next 16 bytes of K +2048 B (LBO); descriptor version 1 (sm_100)
__device__ __forceinline__ uint64_t mm_desc(uint32_t saddr) {
    return (uint64_t)((saddr & 0x3ffffu) >> 4) | ((uint64_t)(2048u >> 4) << 16) | ((uint64_t)(128u >> 4) << 32) | (1ull << 46);
}
// instruction descriptor: D = s32 (2 << 4), A = B = unsigned 8-bit (0), both K-major, N = 128 (16 << 17), M = 128 (8 << 24)
constexpr uint32_t MM_IDESC = (2u << 4) | ((uint32_t)(MM_COLS >> 3) << 17) | ((uint32_t)(MM_ROWS >> 4) << 24);
__device__ __forceinline__ void mm_mma(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t accumulate) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "setp.ne.b32 p, %4, 0;\n"
        "tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, p;\n"
        "}\n" ::"r"(d_tmem), "l"(a_desc), "l"(b_desc), "r"(MM_IDESC), "r"(accumulate)
        : "memory");
}
__device__ __forceinline__ double u32_to_f64(uint32_t v) {     // exact, on the fp64 pipe (no conversion unit)
    return __hiloint2double(0x43300000, (int)v) - 4503599627370496.0;
}


// the guarded IEEE division of the list path (rare: denominators beyond the table, or the query-DB label quirk where c may
// exceed the mislabelled denominator); out of line so that its temporaries do not compete with the fp64 sums for registers
__device__ __noinline__ double mm_guarded_div(int c, int ta_tb) {
    const int den = max(ta_tb - c, 1);
    return (double)c / (double)den;            // algorithm_impl.hpp:250-254
}

// Jaccard + S, N for 8 columns of one row: columns C0 .. C0+7 of the thread's 32; sv = their fp64 sums as loaded from tensor
// memory (lo, hi words).  FAST: den < RCP_N and c <= den hold for the whole (tile, SCP), so J = c/den is the table reciprocal +
// two FMAs (== IEEE division, tests/test_division.py); otherwise the guarded IEEE division of the list path.
template <bool FAST, int C0>
__device__ __forceinline__ void mm_epi8(uint32_t (&sv)[16], uint32_t (&N8)[8], const uint32_t *cv, const int ta, const uint32_t ct32,
                                        const uint32_t rcp32) {
#pragma unroll
    for (int g4 = 0; g4 < 2; g4++) {
        int tb[4];
        asm volatile("ld.shared.v4.s32 {%0, %1, %2, %3}, [%4];" : "=r"(tb[0]), "=r"(tb[1]), "=r"(tb[2]), "=r"(tb[3])
                     : "r"(ct32 + 4u * (uint32_t)(C0 + g4 * 4)));
#pragma unroll
        for (int u = 0; u < 4; u++) {
            const int j = C0 + g4 * 4 + u, k = g4 * 4 + u;
            const uint32_t c = cv[k];
            double J;
            if (FAST) {
                const uint32_t den = (uint32_t)(ta + tb[u]) - c;
                const double r = lds_f64(rcp32 + 8u * den), cd = u32_to_f64(c), dd = u32_to_f64(den);
                const double qq = __dmul_rn(cd, r);
                J = __fma_rn(__fma_rn(-qq, dd, cd), r, qq);
            } else {
                J = mm_guarded_div((int)c, ta + tb[u]);
            }
            const double sum = __hiloint2double((int)sv[2 * k + 1], (int)sv[2 * k]) + J;
            sv[2 * k] = (uint32_t)__double2loint(sum);
            sv[2 * k + 1] = (uint32_t)__double2hiint(sum);
            if (c) N8[j >> 2] += 1u << (8 * (j & 3));
        }
    }
}

// one item of the sparse walk: x = sp_item(row, tetramer); dict32 / cq32 = shared-space addresses of the dictionary stage and of
// the quadrant's counters.  A miss costs the bitmap load and the shift; a hit bumps the 16-bit counter(s) of the tetramer's column(s).
__device__ __forceinline__ void sp_lookup(uint32_t dict32, uint32_t cq32, uint32_t x) {
    uint32_t w;
    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(w) : "r"(dict32 + (x >> 16)));
    const uint32_t t = __funnelshift_l(0u, w, x);                   // w << (x & 31): presence bit of the tetramer in bit 31
    if ((int32_t)t < 0) {
        uint32_t rk, e;
        asm volatile("ld.shared.u16 %0, [%1];" : "=r"(rk) : "r"(dict32 + (uint32_t)SP_BM_BYTES + (x >> 17)));
        const uint32_t ent32 = dict32 + (uint32_t)SP_ENT_OFF;
        asm volatile("ld.shared.u16 %0, [%1];" : "=r"(e) : "r"(ent32 + 2u * (rk + __popc(t))));
        const uint32_t row32 = cq32 + ((x >> 8) & 31u) * (uint32_t)MM_CNT_PITCH;
        if (e < 0x8000u) {
            asm volatile("red.shared.add.u32 [%0], %1;" ::"r"(row32 + (e & 0xffu)), "r"(1u << (e >> 8)) : "memory");
        } else {                                                    // several columns of the block hold the tetramer: [n, e1 .. en]
            const uint32_t l32 = ent32 + 2u * (e & 0x7fffu);
            uint32_t n;
            asm volatile("ld.shared.u16 %0, [%1];" : "=r"(n) : "r"(l32));
            for (uint32_t k = 1; k <= n; k++) {
                uint32_t e2;
                asm volatile("ld.shared.u16 %0, [%1];" : "=r"(e2) : "r"(l32 + 2u * k));
                asm volatile("red.shared.add.u32 [%0], %1;" ::"r"(row32 + (e2 & 0xffu)), "r"(1u << (e2 >> 8)) : "memory");
            }
        }
    }
}

// row items travel global -> shared memory by cp.async (LDGSTS), 16 bytes per lane and chunk, into a ring private to the warp:
// RING_D chunks are in flight per warp without holding registers, and the first chunks of SCP p+1 are requested before the
// epilogue of SCP p starts.  Every step commits exactly one group (possibly empty), so "at most RING_D - 1 groups pending"
// means the oldest chunk has landed.
__device__ __forceinline__ void cp_async16(uint32_t dst32, const void *src) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst32), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N> __device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }
// requests this warp's first RING_D chunks (cg, cg + 4, ...) of a slab of nch chunks
template <int RING_D>
__device__ __forceinline__ void sp_prefetch(uint32_t ring32, const uint4 *slab, int nch, int cg) {
#pragma unroll
    for (int d = 0; d < RING_D; d++) {
        if (cg + 4 * d < nch) cp_async16(ring32 + d * 512, slab + (int64_t)(cg + 4 * d) * 32);
        cp_async_commit();
    }
}

template <class Cfg>
__global__ void __launch_bounds__(MM_THREADS, 1) aji_mma_kernel(const MmaParams prm) {
    constexpr int MM_NSTAGE = Cfg::NSTAGE, RING_D = Cfg::RING_D;
    extern __shared__ __align__(128) unsigned char smem[];
    uint64_t *full = reinterpret_cast<uint64_t *>(smem);                  // [NSTAGE]
    uint64_t *empty = full + MM_NSTAGE;                                   // [NSTAGE]
    uint64_t *acc_full = empty + MM_NSTAGE;                               // [2]
    uint64_t *acc_empty = acc_full + 2;                                   // [2]
    uint64_t *ct_full = acc_empty + 2;                                    // [CT_RING]
    uint64_t *ct_empty = ct_full + MM_CT_RING;                            // [CT_RING]
    uint64_t *dict_full = ct_empty + MM_CT_RING;                          // the sparse dictionary stage
    uint64_t *dict_empty = dict_full + 1;
    uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(smem + 240);
    int32_t *col_g = reinterpret_cast<int32_t *>(smem + 256);             // [128] genome id of the tile's columns (output phase)
    int32_t *ctring = reinterpret_cast<int32_t *>(smem + MM_SMEM_MISC);   // [CT_RING][128]
    double *rcp = reinterpret_cast<double *>(smem + MM_SMEM_MISC + MM_SMEM_CT);
    int32_t *kp_s = reinterpret_cast<int32_t *>(smem + MM_SMEM_MISC + MM_SMEM_CT + MM_SMEM_RCP);   // [P] kpad
    unsigned char *cnt_mem = smem + MM_SMEM_MISC + MM_SMEM_CT + MM_SMEM_RCP + MM_SMEM_KP;
    unsigned char *stage_mem = cnt_mem + MM_SMEM_CNT;
    unsigned char *ring_mem = stage_mem + Cfg::STAGES_BYTES;               // [16 warps][RING_D][32 lanes x 16 B]
    unsigned char *dict_mem = ring_mem + Cfg::RING_BYTES;
    static_assert((2 * MM_NSTAGE + 4 + 2 * MM_CT_RING + 2) * 8 <= 240 && 256 + MM_COLS * 4 <= MM_SMEM_MISC, "misc area");

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int P = prm.P;

    if (tid == 0) {
        for (int s = 0; s < MM_NSTAGE; s++) { mbar_init(&full[s], 1); mbar_init(&empty[s], 1); }
        for (int b = 0; b < 2; b++) { mbar_init(&acc_full[b], 1); mbar_init(&acc_empty[b], MM_EPI_WARPS); }
        for (int s = 0; s < MM_CT_RING; s++) { mbar_init(&ct_full[s], 1); mbar_init(&ct_empty[s], MM_EPI_WARPS); }
        mbar_init(dict_full, 1); mbar_init(dict_empty, MM_EPI_WARPS);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 0) tmem_alloc(tmem_slot, MM_TMEM_COLS);
    for (int i = tid; i < RCP_N; i += MM_THREADS) rcp[i] = (i > 0) ? 1.0 / (double)i : 0.0;
    for (int i = tid; i < P; i += MM_THREADS) kp_s[i] = prm.kpad[i];
    for (int i = tid; i < MM_SMEM_CNT / 4; i += MM_THREADS) reinterpret_cast<uint32_t *>(cnt_mem)[i] = 0u;
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == MM_EPI_WARPS) {
        // ===================== producer: bulk copies of the operand images and the column T values =====================
        if (lane == 0) {
            uint32_t it = 0, ci = 0;
            for (int tile_i = blockIdx.x; tile_i < prm.n_tiles; tile_i += gridDim.x) {
                const int4 tile = prm.tiles[tile_i];
                const unsigned char *a_img = prm.img_rows + (int64_t)tile.x * prm.blk_stride;
                const unsigned char *b_img = prm.img_cols + (int64_t)tile.y * prm.blk_stride;
                for (int p = 0; p < P; p++) {
                    {   // column T values of SCP p
                        const uint32_t s = ci % MM_CT_RING, ph = (ci / MM_CT_RING) & 1u;
                        mbar_wait(&ct_empty[s], ph ^ 1u);
                        mbar_expect_tx(&ct_full[s], MM_COLS * 4u);
                        bulk_g2s(ctring + s * MM_COLS, prm.colT + ((int64_t)tile.y * P + p) * MM_COLS, MM_COLS * 4u, &ct_full[s]);
                        ci++;
                    }
                    const int kp = kp_s[p];
                    const int64_t ko = __ldg(prm.koff + p) * 128;
                    for (int k0 = 0; k0 < kp; k0 += MM_STAGE_K) {
                        const uint32_t s = it % MM_NSTAGE, ph = (it / MM_NSTAGE) & 1u;
                        const uint32_t bytes = (uint32_t)min(MM_STAGE_K, kp - k0) * 128u;
                        mbar_wait(&empty[s], ph ^ 1u);
                        mbar_expect_tx(&full[s], 2u * bytes);
                        unsigned char *dst = stage_mem + (size_t)s * MM_STAGE_BYTES;
                        bulk_g2s(dst, a_img + ko + (int64_t)k0 * 128, bytes, &full[s]);
                        bulk_g2s(dst + MM_OP_BYTES, b_img + ko + (int64_t)k0 * 128, bytes, &full[s]);
                        it++;
                    }
                }
            }
        }
    } else if (warp == MM_EPI_WARPS + 1) {
        // ===================== tensor-core issuer =====================
        if (lane == 0) {
            uint32_t it = 0, ai = 0;
            const uint32_t stage32 = smem_u32(stage_mem);
            for (int tile_i = blockIdx.x; tile_i < prm.n_tiles; tile_i += gridDim.x) {
                for (int p = 0; p < P; p++) {
                    const int kp = kp_s[p];
                    if (kp == 0) continue;
                    const uint32_t buf = ai & 1u, aph = (ai >> 1) & 1u;
                    mbar_wait(&acc_empty[buf], aph ^ 1u);        // the epilogue has read this buffer's previous SCP
                    tc_fence_after();
                    const uint32_t d_tmem = tmem_base + buf * MM_COLS;
                    for (int k0 = 0; k0 < kp; k0 += MM_STAGE_K) {
                        const uint32_t s = it % MM_NSTAGE, ph = (it / MM_NSTAGE) & 1u;
                        mbar_wait(&full[s], ph);
                        tc_fence_after();
                        const uint32_t a32 = stage32 + s * MM_STAGE_BYTES, b32 = a32 + MM_OP_BYTES;
                        const int nk = min(MM_STAGE_K, kp - k0) / 32;
                        for (int k = 0; k < nk; k++)
                            mm_mma(d_tmem, mm_desc(a32 + k * 4096), mm_desc(b32 + k * 4096), (k0 | k) ? 1u : 0u);
                        tc_commit(&empty[s]);                    // the stage is free once these MMAs have read it
                        it++;
                    }
                    tc_commit(&acc_full[buf]);                   // SCP p's counts are complete in tensor memory
                    ai++;
                }
            }
        }
    } else if (warp == MM_EPI_WARPS + 2) {
        // ===================== producer of the sparse dictionaries: one bulk copy per (tile, SCP, part) =====================
        if (lane == 0 && prm.has_sparse) {
            uint32_t di = 0;
            for (int tile_i = blockIdx.x; tile_i < prm.n_tiles; tile_i += gridDim.x) {
                const int4 tile = prm.tiles[tile_i];
                const int64_t jp = (int64_t)tile.y * P;
                for (int p = 0; p < P; p++) {
                    const int ns = __ldg(prm.sp_nsub + jp + p);
                    for (int k = 0; k < ns; k++) {
                        const uint4 d = __ldg(reinterpret_cast<const uint4 *>(prm.sp_sub + (jp + p) * SP_MAXSUB + k));   // {off.lo, off.hi, bytes, n}
                        const int64_t off = (int64_t)(((uint64_t)d.y << 32) | d.x);
                        mbar_wait(dict_empty, (di & 1u) ^ 1u);              // every walker warp is done with the previous dictionary
                        mbar_expect_tx(dict_full, d.z);
                        bulk_g2s(dict_mem, prm.sp_blob + off, d.z, dict_full);
                        di++;
                    }
                }
            }
        }
    } else {
        // ===================== walker / epilogue warps: sparse counts, dense counts -> Jaccard -> S, N -> AJI =====================
        const int q = warp & 3, cg = warp >> 2;                  // tensor-memory lane quadrant (rows 32q..), column group (cols 32cg..)
        const int rl = q * 32 + lane;                            // row of the tile
        const uint32_t lane_addr = (uint32_t)(q * 32) << 16;
        const uint32_t s_addr = tmem_base + lane_addr + MM_TMEM_S + cg * 64;       // this thread's 32 fp64 sums
        const uint32_t cq32 = smem_u32(cnt_mem + q * MM_CNT_QUAD);      // the quadrant's 32 x 128 sparse counters (uint16)
        const uint32_t cw32 = cq32 + lane * MM_CNT_PITCH + cg * 64;     // this thread's 32 counters
        const uint32_t dict32 = smem_u32(dict_mem);
        const uint32_t ring32 = smem_u32(ring_mem + warp * (RING_D * 512) + lane * 16);   // this lane's 16 bytes of ring slot 0
        const uint32_t rcp32 = smem_u32(rcp);
        const bool sparse = prm.has_sparse != 0;
        uint32_t ai = 0, ci = 0, di = 0;
        for (int tile_i = blockIdx.x; tile_i < prm.n_tiles; tile_i += gridDim.x) {
            const int4 tile = prm.tiles[tile_i];
            const int row = tile.x * MM_ROWS + rl;
            const bool row_ok = row < prm.n_rows;
            const int gi = row_ok ? prm.rows[row] : -1;
            const int ti = row_ok ? prm.row_t[row] : 0;
            const int j0 = tile.y * MM_COLS;
            const int ncol = min(MM_COLS, prm.n_cols - j0);

            uint32_t N8[8];
#pragma unroll
            for (int j = 0; j < 8; j++) N8[j] = 0u;
            {   // S = 0 for the 32 columns of this thread (tensor memory)
                uint32_t z[16];
#pragma unroll
                for (int k = 0; k < 16; k++) z[k] = 0u;
                __syncwarp();
#pragma unroll
                for (int g = 0; g < 4; g++) tmem_st16(s_addr + 16 * g, z);
                tmem_wait_st();
            }
            // columns of this thread that are real pairs (isValidPair): inside the matrix and not the row's own genome
            uint32_t valid = 0u;
#pragma unroll
            for (int j = 0; j < 32; j++) {
                const int col = cg * 32 + j;
                const int gj = (col < ncol) ? __ldg(prm.cols + j0 + col) : -1;
                if (row_ok && gj >= 0 && gj != gi) valid |= 1u << j;
            }
            // per-SCP scalars are fetched one SCP ahead (their latency hides behind the previous SCP's epilogue): the row's T value,
            // the fast-division guard, the slab extents of the row items and the number of dictionary parts
            const int tileP_r = tile.x * P, tileP_c = tile.y * P;
            const bool never_fast = prm.quirk != 0, all_fast = prm.all_fast != 0;
            int ta = row_ok ? __ldg(prm.T + ti) : 0;
            bool fast = !never_fast && (all_fast || __ldg(prm.rowTmax + tileP_r) + __ldg(prm.colTmax + tileP_c) < RCP_N);
            // sparse side: the slab extents of the quadrant's row items and the number of dictionary parts, one SCP ahead
            const int64_t gp0 = ((int64_t)tile.x * 4 + q) * P, jp0 = (int64_t)tile.y * P;
            int64_t so0 = 0, so1 = 0;
            int ns = 0;
            if (sparse) { so0 = __ldg(prm.sp_off + gp0); so1 = __ldg(prm.sp_off + gp0 + 1); ns = __ldg(prm.sp_nsub + jp0); }
            const uint4 *items = reinterpret_cast<const uint4 *>(prm.sp_items) + lane;
            if (ns > 0) sp_prefetch<RING_D>(ring32, items + so0 * 32, (int)(so1 - so0), cg);   // SCP 0 (later SCPs: one SCP ahead)

            for (int p = 0; p < P; p++) {
                int ta_nxt = 0; bool fast_nxt = false; int64_t so2 = 0; int ns_nxt = 0;
                if (p + 1 < P) {
                    if (row_ok) ta_nxt = __ldg(prm.T + (int64_t)(p + 1) * prm.t_ld + ti);
                    fast_nxt = !never_fast && (all_fast || __ldg(prm.rowTmax + tileP_r + p + 1) + __ldg(prm.colTmax + tileP_c + p + 1) < RCP_N);
                    if (sparse) { so2 = __ldg(prm.sp_off + gp0 + p + 2); ns_nxt = __ldg(prm.sp_nsub + jp0 + p + 1); }
                }
                const int kp = kp_s[p];
                // ---- sparse walk: the quadrant's row items of this SCP through every part of the column block's dictionary ----
                const bool np = ns > 0;                    // the same for the 4 warps of the quadrant (and for the whole CTA)
                if (np) {
                    const uint4 *slab = items + so0 * 32;
                    const int nch = (int)(so1 - so0);
                    for (int k = 0; k < ns; k++) {
                        if (k > 0) sp_prefetch<RING_D>(ring32, slab, nch, cg);                   // further dictionary parts re-stream the items
                        mbar_wait(dict_full, di & 1u);
                        di++;
                        int slot = 0;
                        for (int ch = cg; ch < nch; ch += 4) {
                            cp_async_wait<RING_D - 1>();                                         // the oldest chunk in flight has landed
                            uint4 cur;
                            const uint32_t a = ring32 + slot * 512;
                            asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(cur.x), "=r"(cur.y), "=r"(cur.z), "=r"(cur.w) : "r"(a) : "memory");
                            sp_lookup(dict32, cq32, cur.x); sp_lookup(dict32, cq32, cur.y);
                            sp_lookup(dict32, cq32, cur.z); sp_lookup(dict32, cq32, cur.w);
                            if (ch + 4 * RING_D < nch) cp_async16(a, slab + (int64_t)(ch + 4 * RING_D) * 32);   // refill the slot just read
                            cp_async_commit();
                            slot = (slot + 1 == RING_D) ? 0 : slot + 1;
                        }
                        cp_async_wait<0>();
                        __syncwarp();
                        if (lane == 0) mbar_arrive(dict_empty);
                    }
                    asm volatile("bar.sync %0, 128;" ::"r"(2 + q) : "memory");                   // the quadrant's counters are complete
                }
                // the first chunks of the next SCP's items travel while this SCP's epilogue runs
                if (ns_nxt > 0) sp_prefetch<RING_D>(ring32, items + so1 * 32, (int)(so2 - so1), cg);
                // ---- column T values ----
                const uint32_t cs = ci % MM_CT_RING;
                mbar_wait(&ct_full[cs], (ci / MM_CT_RING) & 1u);
                const uint32_t ct32 = smem_u32(ctring + cs * MM_COLS + cg * 32);
                // ---- counts from tensor memory ----
                uint32_t taddr = 0u, buf = 0u;
                if (kp) {
                    buf = ai & 1u;
                    mbar_wait(&acc_full[buf], (ai >> 1) & 1u);
                    tc_fence_after();
                    taddr = tmem_base + lane_addr + buf * MM_COLS + cg * 32;
                }
#pragma unroll
                for (int hf = 0; hf < 2; hf++) {
                    uint32_t cv[16], sv[16];
                    tmem_ld16(s_addr + hf * 32, sv);        // the sums of columns 16 hf .. 16 hf + 7
                    if (kp) {
                        tmem_ld16(taddr + hf * 16, cv);
                        tmem_wait_ld(cv);
                        if (hf == 1) {                      // both halves are in registers: hand the buffer back to the tensor cores
                            tc_fence_before();
                            __syncwarp();
                            if (lane == 0) mbar_arrive(&acc_empty[buf]);
                        }
                    } else {
#pragma unroll
                        for (int u = 0; u < 16; u++) cv[u] = 0u;
                    }
                    if (np) {                               // uniform over the quadrant: add this row's sparse counters, then clear them
#pragma unroll
                        for (int v4 = 0; v4 < 2; v4++) {
                            uint32_t w[4];
                            const uint32_t a = cw32 + hf * 32 + v4 * 16;
                            asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(w[0]), "=r"(w[1]), "=r"(w[2]), "=r"(w[3]) : "r"(a));
                            asm volatile("st.shared.v4.u32 [%0], {%1, %1, %1, %1};" ::"r"(a), "r"(0u) : "memory");
#pragma unroll
                            for (int u = 0; u < 4; u++) { cv[v4 * 8 + 2 * u] += w[u] & 0xffffu; cv[v4 * 8 + 2 * u + 1] += w[u] >> 16; }
                        }
                    }
                    if (valid != 0xffffffffu) {             // the row's own genome / columns beyond the matrix (diagonal and edge tiles)
#pragma unroll
                        for (int u = 0; u < 16; u++) if (!((valid >> (hf * 16 + u)) & 1u)) cv[u] = 0u;
                    }
                    if (prm.counts_out) {                   // small-N parity hook
#pragma unroll
                        for (int u = 0; u < 16; u++) {
                            const int col = cg * 32 + hf * 16 + u;
                            if (row_ok && col < ncol) prm.counts_out[((int64_t)row * prm.n_cols + (j0 + col)) * P + p] = (int)cv[u];
                        }
                    }
                    tmem_wait_ld(sv);
                    if (hf == 0) { if (fast) mm_epi8<true, 0>(sv, N8, cv, ta, ct32, rcp32); else mm_epi8<false, 0>(sv, N8, cv, ta, ct32, rcp32); }
                    else         { if (fast) mm_epi8<true, 16>(sv, N8, cv, ta, ct32, rcp32); else mm_epi8<false, 16>(sv, N8, cv, ta, ct32, rcp32); }
                    tmem_st16(s_addr + hf * 32, sv);
                    tmem_ld16(s_addr + hf * 32 + 16, sv);   // the sums of columns 16 hf + 8 .. 16 hf + 15
                    tmem_wait_ld(sv);
                    if (hf == 0) { if (fast) mm_epi8<true, 8>(sv, N8, cv + 8, ta, ct32, rcp32); else mm_epi8<false, 8>(sv, N8, cv + 8, ta, ct32, rcp32); }
                    else         { if (fast) mm_epi8<true, 24>(sv, N8, cv + 8, ta, ct32, rcp32); else mm_epi8<false, 24>(sv, N8, cv + 8, ta, ct32, rcp32); }
                    tmem_st16(s_addr + hf * 32 + 16, sv);
                }
                tmem_wait_st();
                if (kp) ai++;
                __syncwarp();
                if (lane == 0) mbar_arrive(&ct_empty[cs]);
                ci++;
                if (np) asm volatile("bar.sync %0, 128;" ::"r"(2 + q) : "memory");              // counters read and cleared: the next walk may add
                ta = ta_nxt; fast = fast_nxt; so0 = so1; so1 = so2; ns = ns_nxt;
            }

            // ================= output: AJI = S / N (computeAJI) into the dense matrix, four 32-column rounds =================
            double *stg = reinterpret_cast<double *>(cnt_mem);                 // [128][MM_STG_PITCH]
            if (cg == 0) col_g[rl] = (rl < ncol) ? prm.cols[j0 + rl] : -1;     // 128 column genomes, written by the 4 warps of group 0
            const int row0 = tile.x * MM_ROWS;
            const int nrow = min(MM_ROWS, prm.n_rows - row0);
            const int64_t ro0 = (int64_t)tile.z;
            for (int round = 0; round < 4; round++) {
                asm volatile("bar.sync 1, %0;" ::"n"(MM_EPI_WARPS * 32) : "memory");      // staging (== sparse counters) is free
                if (cg == round) {
#pragma unroll
                    for (int g = 0; g < 4; g++) {
                        uint32_t sv[16];
                        tmem_ld16(s_addr + 16 * g, sv);
                        tmem_wait_ld(sv);
#pragma unroll
                        for (int u = 0; u < 8; u++) {
                            const int j = 8 * g + u;
                            const double Ssum = __hiloint2double((int)sv[2 * u + 1], (int)sv[2 * u]);
                            const int n = (int)((N8[j >> 2] >> (8 * (j & 3))) & 0xffu);
                            stg[rl * MM_STG_PITCH + j] = n ? Ssum / (double)n : 0.0;
                            const int col = cg * 32 + j;
                            if (row_ok && col < ncol) {
                                if (prm.S_out) prm.S_out[(int64_t)row * prm.n_cols + (j0 + col)] = Ssum;
                                if (prm.N_out) prm.N_out[(int64_t)row * prm.n_cols + (j0 + col)] = n;
                            }
                        }
                    }
                }
                asm volatile("bar.sync 1, %0;" ::"n"(MM_EPI_WARPS * 32) : "memory");
                const int c0 = round * 32, nc = min(32, ncol - c0);
                if (nc > 0) {
                    if (prm.out_cs == 1) {
                        // direct cells: one 256-B row segment per warp instruction; symmetric: only gi <= gj, diagonal = 0
                        for (int r2 = warp; r2 < nrow; r2 += MM_EPI_WARPS) {
                            const int g_i = prm.rows[row0 + r2];
                            double *dst = prm.out + (ro0 + r2) * prm.out_rs + j0 + c0;
                            if (lane < nc) {
                                const int g_j = col_g[c0 + lane];
                                const double v = stg[r2 * MM_STG_PITCH + lane];
                                if (prm.symmetric) { if (g_i <= g_j) dst[lane] = (g_i == g_j) ? 0.0 : v; }
                                else dst[lane] = (g_i == g_j) ? 0.0 : v;
                            }
                        }
                    }
                    if (prm.out_cs != 1 || prm.mirror) {
                        // transposed cells: for a fixed column the tile's rows are contiguous in the output
                        const int64_t cstride = (prm.out_cs != 1) ? prm.out_cs : prm.out_rs;
                        const int64_t rbase = (prm.out_cs != 1) ? ro0 : (int64_t)row0;
                        for (int c = warp; c < nc; c += MM_EPI_WARPS) {
                            const int g_j = col_g[c0 + c];
                            double *dst = prm.out + (int64_t)(j0 + c0 + c) * cstride + rbase;
                            for (int r2 = lane; r2 < nrow; r2 += 32) {
                                const int g_i = prm.rows[row0 + r2];
                                const double v = stg[r2 * MM_STG_PITCH + c];
                                if (prm.symmetric) { if (g_i < g_j) dst[r2] = v; }
                                else dst[r2] = (g_i == g_j) ? 0.0 : v;
                            }
                        }
                    }
                }
            }
            if (prm.progress) __threadfence_system();
            asm volatile("bar.sync 1, %0;" ::"n"(MM_EPI_WARPS * 32) : "memory");
            if (prm.progress && tid == 0) {
                atomicAdd(prm.progress + tile.x, 1);
                if (prm.mirror) atomicAdd(prm.progress + tile.y, 1);
            }
            // the staging area doubles as the sparse counters: clear what the last round left behind
            for (int i = tid; i < MM_SMEM_CNT / 4; i += MM_EPI_WARPS * 32) reinterpret_cast<uint32_t *>(cnt_mem)[i] = 0u;
            asm volatile("bar.sync 1, %0;" ::"n"(MM_EPI_WARPS * 32) : "memory");
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 0) tmem_dealloc(tmem_base, MM_TMEM_COLS);
}

}  // namespace pfaai
