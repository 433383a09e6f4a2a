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

// ---- tile geometry of the core kernel ----
constexpr int ROWS_PER_WARP = 16;            // one warp = 16 rows x 2 half-lists
constexpr int COLS_PER_TILE = 64;            // one 64-bit column mask per distinct tetramer
#ifndef PF_TILE_WARPS
#define PF_TILE_WARPS 24
#endif
constexpr int TILE_WARPS = PF_TILE_WARPS;   // 24 warps = 768 threads, <= 80 registers each (the fp64 sums S live in TMEM)
constexpr int TILE_THREADS = TILE_WARPS * 32;
constexpr int TILE_ROWS = TILE_WARPS * ROWS_PER_WARP;   // 384 rows x 64 columns per tile
// Tensor memory (256 KB per SM, 128 lanes x 512 32-bit columns) holds the per-pair fp64 sums S: a warp may only touch
// lanes 32*(warp%4)..+31, so warp w owns columns [64*(w/4), 64*(w/4)+64) of its lane quadrant = 32 doubles per thread.
constexpr int TMEM_COLS_PER_WARP = 64;
constexpr int TMEM_COLS = 512;               // allocation must be a power of two >= (TILE_WARPS/4) * 64 = 384
static_assert(TILE_WARPS % 4 == 0 && (TILE_WARPS / 4) * TMEM_COLS_PER_WARP <= TMEM_COLS, "S does not fit tensor memory");
// reciprocal table for the Jaccard division: J = c/den via r = RN(1/den), q = c*r, q' = fma(fma(-q,den,c), r, q), which is
// the correctly rounded quotient for every 0 <= c <= den < 20000 (exhaustively checked, tests/test_division.py)
constexpr int RCP_N = 2048;
constexpr int SMEM_RCP = RCP_N * 8;

// The tetramer universe is cut into NRANGE ranges; one pipeline stage holds the dictionary of one
// (column block, SCP, range): a presence/rank bitmap over the range plus one 64-bit column mask per
// distinct tetramer.  Row lists are stored per range with range-local ids.
constexpr int NRANGE = 2;
constexpr int RANGE_T = NTETRA / NRANGE;     // 80000 tetramers per range
constexpr int BM_WORDS = RANGE_T / 16;       // 5000 words: 16 presence bits (low) + 16-bit rank (high)
constexpr int BM_WORDS_PAD = BM_WORDS + 4;   // + 4 always-zero words: the padding id lands there
constexpr int BM_BYTES = BM_WORDS_PAD * 4;   // 20016, a multiple of 16 (bulk-copy granularity)
constexpr uint32_t PAD_ID = RANGE_T;         // padding value of packed row slices (range-local id)
// a range-local id v is stored in the row slabs as  x = (v>>4) << 7 | (31 - (v&15)):  x>>5 is the byte offset of
// its bitmap word, x&31 the left shift that moves its presence bit to bit 31 (and drops the rank half-word)
__host__ __device__ __forceinline__ uint32_t slab_encode(uint32_t v) { return ((v >> 4) << 7) | (31u - (v & 15u)); }
constexpr int MAX_SUB = 64;                   // a (block, SCP, range) dictionary is split into <= 64 column sub-ranges (1 column each)

// shared memory of the tile kernel: [misc 2 KB][reciprocal table 16 KB][stage 0][stage 1]
constexpr int SMEM_TOTAL = 227 * 1024;                       // 232448
constexpr int SMEM_MISC = 2048;
constexpr int STAGE_BYTES = ((SMEM_TOTAL - SMEM_MISC - SMEM_RCP) / 2) & ~127;   // 107008
constexpr int MASK_CAP = ((STAGE_BYTES - BM_BYTES) / 8) & ~1;        // entries incl. the all-zero miss slot 0

// output staging of a finished tile inside the two stage slots: [TILE_ROWS][STG_PITCH] doubles + the rows' genome ids
constexpr int STG_PITCH = COLS_PER_TILE + 1;                           // 65: conflict-free column reads
static_assert((size_t)TILE_ROWS * STG_PITCH * 8 + TILE_ROWS * 4 <= 2 * (size_t)STAGE_BYTES, "tile staging does not fit");

struct __align__(16) StageDesc {             // one pipeline stage = one bulk copy
    int64_t off;                             // byte offset of the blob [bitmap BM_BYTES][masks] in the dictionary buffer
    uint32_t bytes;                          // size of the bulk copy (multiple of 16)
    uint16_t p;                              // SCP index
    uint8_t q;                               // tetramer range
    uint8_t flags;                           // 1 = first stage of SCP p, 2 = last stage of SCP p
};
static_assert(sizeof(StageDesc) == 16, "StageDesc is loaded as one uint4");

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
