// mma_index.cuh -- device-side preparation of the dense-contraction path (K5, aji_mma.cuh):
//   k_list_len     posting-list length of every (SCP, tetramer)  [optionally on a genome sample: the planning estimate]
//   k_len_hist     histogram of those lengths + |E| = sum n(n-1)/2   (countTetramerTuples, ds_helper.hpp:206-265, as a work counter)
//   k_dense_ids    per SCP: dense tetramers (length >= tau) numbered 0..K_p-1
//   k_image_fill   0/1 operand images in the tensor cores' shared-memory layout (K-major, no swizzle)
//   k_block_t      per 128-genome block: column T values and the block maxima of T (fast-division guard)
// (the sparse side -- row items and sparse dictionaries of the non-dense tetramers -- is built by sparse_index.cuh)
#pragma once
#include "aji_mma.cuh"

namespace pfaai {

constexpr int MM_HMAX = 4096;     // length histogram bins 0..HMAX (HMAX = "that long or longer": always dense)

// genome-major lists (seg_off / vals): one warp per (genome, SCP) segment; stride > 1 visits every stride-th genome only
__global__ void __launch_bounds__(256) k_list_len(const int64_t *__restrict__ seg_off, const uint32_t *__restrict__ vals, int G, int P,
                                                   int stride, uint32_t *__restrict__ len) {
    const int lane = threadIdx.x & 31;
    const int64_t nwarp = ((int64_t)gridDim.x * blockDim.x) >> 5;
    const int64_t ng = (G + stride - 1) / stride;
    for (int64_t w = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5; w < ng * P; w += nwarp) {
        const int64_t g = (w / P) * stride;
        const int p = (int)(w % P);
        const int64_t b = seg_off[g * P + p], e = seg_off[g * P + p + 1];
        uint32_t *row = len + (int64_t)p * NTETRA;
        for (int64_t i = b + lane; i < e; i += 32) atomicAdd(row + vals[i], 1u);
    }
}

// hist[L] += 1 for every list of length L >= 1 (L capped at HMAX); acc[0] += sum L(L-1)/2 (exact |E| of all-vs-all)
__global__ void __launch_bounds__(256) k_len_hist(const uint32_t *__restrict__ len, int64_t n, unsigned long long *__restrict__ hist,
                                                   unsigned long long *__restrict__ acc) {
    __shared__ uint32_t sh[MM_HMAX + 1];
    for (int i = threadIdx.x; i <= MM_HMAX; i += blockDim.x) sh[i] = 0u;
    __syncthreads();
    unsigned long long e = 0ull;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        const uint32_t L = len[i];
        if (L) { atomicAdd(&sh[min(L, (uint32_t)MM_HMAX)], 1u); e += (unsigned long long)L * (L - 1) / 2; }
    }
    for (int o = 16; o > 0; o >>= 1) e += __shfl_xor_sync(0xffffffffu, e, o);
    if ((threadIdx.x & 31) == 0 && e) atomicAdd(acc, e);
    __syncthreads();
    for (int i = threadIdx.x; i <= MM_HMAX; i += blockDim.x) if (sh[i]) atomicAdd(&hist[i], (unsigned long long)sh[i]);
}

// one CTA per SCP: did[p][t] = rank of t among the SCP's dense tetramers (length >= tau), NOT_DENSE otherwise; kcnt[p] = K_p
__global__ void __launch_bounds__(1024) k_dense_ids(const uint32_t *__restrict__ len, uint32_t tau, uint16_t *__restrict__ did,
                                                     int32_t *__restrict__ kcnt, int kmax, int *__restrict__ err) {
    __shared__ int wsum[32];
    __shared__ int base_s;
    const int p = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint32_t *L = len + (int64_t)p * NTETRA;
    uint16_t *D = did + (int64_t)p * NTETRA;
    if (tid == 0) base_s = 0;
    __syncthreads();
    for (int t0 = 0; t0 < NTETRA; t0 += 1024) {
        const int t = t0 + tid;
        const int f = (t < NTETRA && L[t] >= tau) ? 1 : 0;
        int x = f;
        for (int o = 1; o < 32; o <<= 1) { const int y = __shfl_up_sync(0xffffffffu, x, o); if (lane >= o) x += y; }
        if (lane == 31) wsum[warp] = x;
        __syncthreads();
        if (warp == 0) {
            int s = wsum[lane];
            for (int o = 1; o < 32; o <<= 1) { const int y = __shfl_up_sync(0xffffffffu, s, o); if (lane >= o) s += y; }
            wsum[lane] = s;
        }
        __syncthreads();
        const int id = base_s + (warp ? wsum[warp - 1] : 0) + x - f;
        if (t < NTETRA) D[t] = (f && id < kmax) ? (uint16_t)id : MM_NOT_DENSE;
        __syncthreads();
        if (tid == 0) base_s += wsum[31];
        __syncthreads();
    }
    if (tid == 0) { kcnt[p] = base_s; if (base_s > kmax) *err = 1; }
}

// images of the genome list `list` (kernel row or column index -> genome): one warp per (index, SCP)
__global__ void __launch_bounds__(256) k_image_fill(const int64_t *__restrict__ seg_off, const uint32_t *__restrict__ vals,
                                                     const int32_t *__restrict__ list, int n_list, int P, const uint16_t *__restrict__ did,
                                                     const int64_t *__restrict__ koff, int64_t blk_stride, unsigned char *__restrict__ img) {
    const int lane = threadIdx.x & 31;
    const int64_t nwarp = ((int64_t)gridDim.x * blockDim.x) >> 5;
    for (int64_t w = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5; w < (int64_t)n_list * P; w += nwarp) {
        const int idx = (int)(w / P), p = (int)(w % P);
        const int64_t g = list[idx];
        const int64_t b = seg_off[g * P + p], e = seg_off[g * P + p + 1];
        const uint16_t *D = did + (int64_t)p * NTETRA;
        unsigned char *dst = img + (int64_t)(idx >> 7) * blk_stride + koff[p] * 128 + (idx & 127) * 16;
        for (int64_t i = b + lane; i < e; i += 32) {
            const uint32_t id = D[vals[i]];
            if (id != MM_NOT_DENSE) dst[(int64_t)(id >> 4) * 2048 + (id & 15u)] = 1;
        }
    }
}

// per 128-block of a genome list: T values of its members per SCP ([nblk][P][128], 0 beyond the list) and their maximum
__global__ void __launch_bounds__(128) k_block_t(const int32_t *__restrict__ T, int t_ld, int P, const int32_t *__restrict__ tcol, int n_list,
                                                  int32_t *__restrict__ blkT /*may be NULL*/, int32_t *__restrict__ blkmax) {
    const int blk = blockIdx.x / P, p = blockIdx.x % P, i = blk * 128 + threadIdx.x;
    const int v = (i < n_list) ? T[(int64_t)p * t_ld + tcol[i]] : 0;
    if (blkT) blkT[((int64_t)blk * P + p) * 128 + threadIdx.x] = v;
    int m = v;
    for (int o = 16; o > 0; o >>= 1) m = max(m, __shfl_xor_sync(0xffffffffu, m, o));
    __shared__ int wm[4];
    if ((threadIdx.x & 31) == 0) wm[threadIdx.x >> 5] = m;
    __syncthreads();
    if (threadIdx.x == 0) blkmax[(int64_t)blk * P + p] = max(max(wm[0], wm[1]), max(wm[2], wm[3]));
}

// largest T value of the matrix's rows / columns -> *out (atomicMax); out must be zeroed
__global__ void k_max_t(const int32_t *__restrict__ blkmax, int64_t n, int *__restrict__ out) {
    int m = 0;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) m = max(m, blkmax[i]);
    for (int o = 16; o > 0; o >>= 1) m = max(m, __shfl_xor_sync(0xffffffffu, m, o));
    if ((threadIdx.x & 31) == 0 && m > 0) atomicMax(out, m);
}

}  // namespace pfaai
