// microbench.cuh -- on-chip ceiling of the list path's access pattern.
//
// The walk loop of aji_tile_kernel is bound by the shared-memory pipe (random LDS.32 into the presence/rank bitmap, predicated
// LDS.64 into the mask array) and the integer pipe (rank + bit-sliced carry-save adds), not by HBM.  This kernel runs EXACTLY
// that loop -- the same dict_lookup() and PF_CSA sequence, the same 24 warps per SM, the same coalesced LDG.128 stream of
// pre-encoded ids -- against a dictionary that never leaves shared memory: no bulk copies, no mbarrier waits, no epilogue, no
// output.  Its lookup rate is what this access pattern can reach on this GPU at this occupancy; the live rate of the tile
// kernel divided by it is the `roofline_onchip.frac` of the bench line.
//   variant 0  ids uniformly random over the range, presence density = the hit rate of the workload (43 % on config 5)
//   variant 1  every lane looks the SAME id up (broadcast: one wavefront per LDS, no bank conflicts) = the issue-bound ceiling
#pragma once
#include "aji_tile.cuh"

namespace pfaai {

__device__ __forceinline__ uint32_t mb_hash(uint32_t x) {
    x ^= x >> 16; x *= 0x7feb352du; x ^= x >> 15; x *= 0x846ca68bu; x ^= x >> 16;
    return x;
}

// ids: [n_chunks][32] uint4 of slab-encoded range-local ids (n_chunks a power of two), L2-resident
template <int NPL>
__global__ void __launch_bounds__(TILE_THREADS, 1) k_walk_ceiling(const uint4 *ids, uint32_t chunk_mask, int steps, uint32_t hit_q16,
                                                                 uint32_t *sink) {
    extern __shared__ __align__(128) unsigned char smem[];
    unsigned char *stage_mem = smem + SMEM_MISC + SMEM_RCP;
    uint32_t *bm = reinterpret_cast<uint32_t *>(stage_mem);
    uint2 *masks = reinterpret_cast<uint2 *>(stage_mem + BM_BYTES);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    // presence bits with density hit_q16 / 65536, ranks by one serial pass (one-time set-up)
    for (int w = tid; w < BM_WORDS_PAD; w += TILE_THREADS) {
        uint32_t bits = 0u;
        if (w < BM_WORDS)
            for (int b = 0; b < 16; b++) bits |= ((mb_hash((uint32_t)(w * 16 + b) ^ 0x9e3779b9u) & 0xffffu) < hit_q16 ? 1u : 0u) << b;
        bm[w] = bits;
    }
    __syncthreads();
    if (tid == 0) {
        uint32_t run = 0u;
        for (int w = 0; w < BM_WORDS_PAD; w++) { const uint32_t b = bm[w]; bm[w] = b | ((run % ((uint32_t)MASK_CAP - 17u)) << 16); run += __popc(b); }   // ranks wrap: mask reads spread over the whole array
    }
    for (int i = tid; i < MASK_CAP; i += TILE_THREADS) masks[i] = i ? make_uint2(mb_hash(2u * i), mb_hash(2u * i + 1u)) : make_uint2(0u, 0u);
    __syncthreads();
    uint32_t stage32;
    asm volatile("{\n.reg .u64 t64;\ncvta.to.shared.u64 t64, %1;\ncvt.u32.u64 %0, t64;\n}" : "=r"(stage32) : "l"(stage_mem));
    const uint32_t bm32 = stage32, masks32 = stage32 + BM_BYTES;

    uint32_t plo[NPL], phi[NPL];
    uint32_t pe_lo = 0u, pe_hi = 0u;
    bool pending = false;
#pragma unroll
    for (int k = 0; k < NPL; k++) { plo[k] = 0u; phi[k] = 0u; }
    uint32_t at = (uint32_t)(blockIdx.x * TILE_WARPS + warp) * 977u;
    const uint4 *rp = ids + lane;
    uint4 cur0 = __ldg(rp + (size_t)((at) & chunk_mask) * 32), cur1 = __ldg(rp + (size_t)((at + 1) & chunk_mask) * 32);
    at += 2;
    for (int c = 0; c < steps; c++) {
        const uint4 nx0 = __ldg(rp + (size_t)(at & chunk_mask) * 32), nx1 = __ldg(rp + (size_t)((at + 1) & chunk_mask) * 32);
        at += 2;
        uint32_t f_lo[2], f_hi[2];
#pragma unroll
        for (int qq = 0; qq < 2; qq++) {
            const uint4 v = qq ? cur1 : cur0;
            const Mask64 m0 = dict_lookup(bm32, masks32, v.x), m1 = dict_lookup(bm32, masks32, v.y);
            const Mask64 m2 = dict_lookup(bm32, masks32, v.z), m3 = dict_lookup(bm32, masks32, v.w);
            uint32_t ta_lo, ta_hi, tb_lo, tb_hi;
            PF_CSA(ta_lo, plo[0], plo[0], m0.lo, m1.lo); PF_CSA(ta_hi, phi[0], phi[0], m0.hi, m1.hi);
            PF_CSA(tb_lo, plo[0], plo[0], m2.lo, m3.lo); PF_CSA(tb_hi, phi[0], phi[0], m2.hi, m3.hi);
            PF_CSA(f_lo[qq], plo[1], plo[1], ta_lo, tb_lo); PF_CSA(f_hi[qq], phi[1], phi[1], ta_hi, tb_hi);
        }
        uint32_t e_lo, e_hi;
        PF_CSA(e_lo, plo[2], plo[2], f_lo[0], f_lo[1]); PF_CSA(e_hi, phi[2], phi[2], f_hi[0], f_hi[1]);
        if (!pending) {
            pe_lo = e_lo; pe_hi = e_hi; pending = true;
        } else {
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
    uint32_t acc = pe_lo ^ pe_hi;
#pragma unroll
    for (int k = 0; k < NPL; k++) acc ^= plo[k] + 3u * phi[k];
    if (acc == 0x12345678u) sink[0] = acc;     // keeps the counters alive
}

// fills the id stream: variant 0 random over [0, RANGE_T), variant 1 one id per chunk for all lanes
__global__ void k_ceiling_ids(uint4 *ids, int64_t n_u4, int variant) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_u4) return;
    uint32_t v[4];
#pragma unroll
    for (int k = 0; k < 4; k++) {
        const uint32_t key = variant ? (uint32_t)((i >> 5) * 4 + k) : (uint32_t)(i * 4 + k);
        v[k] = slab_encode((uint32_t)(((uint64_t)mb_hash(key * 2654435761u + 12345u) * RANGE_T) >> 32));
    }
    ids[i] = make_uint4(v[0], v[1], v[2], v[3]);
}

}  // namespace pfaai
