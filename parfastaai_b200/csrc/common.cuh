// common.cuh -- shared definitions of the sm_100a AJI path (libpfaai_gpu.so).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include <string>

#include "../../include/pfaai_gpu.h"

namespace pfaai {

constexpr int NTETRA = PFAAI_NTETRAMERS;     // 160000
constexpr int TET_BITS = 18;                 // 2^18 = 262144 > 160000
constexpr uint32_t TET_MASK = (1u << TET_BITS) - 1;
constexpr uint32_t ELL_SENTINEL = NTETRA;    // padding value of packed row slices; its bitmap word is always 0

// ---- tile geometry of the core kernel ----
constexpr int ROWS_PER_WARP = 16;            // one warp = 16 rows x 2 half-lists
constexpr int COLS_PER_TILE = 64;            // one 64-bit column mask per distinct tetramer
constexpr int BM_WORDS = NTETRA / 16 + 1;    // 16 presence bits + 16 rank bits per word; last word = sentinel

struct Ctx;  // pfaai_gpu.cu

#define PF_CUDA(ctx, call)                                                                          \
    do {                                                                                            \
        cudaError_t e__ = (call);                                                                   \
        if (e__ != cudaSuccess) {                                                                   \
            set_error((ctx), std::string(#call) + ": " + cudaGetErrorString(e__) + " (" +           \
                                 __FILE__ + ":" + std::to_string(__LINE__) + ")");                  \
            return PFAAI_ERR_CONSTRUCT;                                                             \
        }                                                                                           \
    } while (0)

void set_error(Ctx *ctx, const std::string &msg);

static inline int64_t ceil_div64(int64_t a, int64_t b) { return (a + b - 1) / b; }

}  // namespace pfaai
