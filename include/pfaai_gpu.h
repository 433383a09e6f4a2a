/*
 * pfaai_gpu.h -- C-ABI of the B200-native AJI hot path (libpfaai_gpu.so).
 *
 * This is the drop-in boundary for the one path of AluruLab/ParFastAAI that runs on the GPU:
 *
 *   tetramer inverted index (Lc, F [, T])  ->  genome-major postings (device radix sort + histograms)
 *       ->  per-(genome pair, SCP) intersection counts (on-chip, never materialised)
 *       ->  Jaccard  ->  (S, N)  ->  AJI matrix.
 *
 * Reference interfaces replaced (paths relative to the reference repo):
 *   - DataStructInterface::construct()  = constructL/F/T/E   include/pfaai/interface.hpp:306-327,
 *     implemented by DataStructHelper::constructLc / parallelPrefixSum / constructF / constructT /
 *     constructE  include/pfaai/ds_helper.hpp:46-162,362-421 and parallelMergeSort psort.hpp:27-53;
 *   - ParFAAIImpl::computeJAC / computeAJI / run  include/pfaai/algorithm_impl.hpp:281-329;
 *   - the id maps / predicates of ParFAAIData, ParFAAIQSubData, ParFAAIQryTgtData
 *     include/pfaai/ds_impl.hpp:39-151,159-337,344-490;
 *   - printOutput's dense matrix fill  src/main.cpp:143-154.
 * The reference has no FFI of its own (header-only C++ templates); INTEGRATION.md shows the binding a
 * maintainer adds (a DataStructInterface subclass + ParFAAIImpl replacement that call these functions).
 *
 * Conventions: plain C, host pointers unless a name ends in `_dev`; the caller owns every host buffer,
 * the context owns all device memory.  Every function returns a pfaai_status (the reference's
 * PFAAI_ERROR_CODE values, interface.hpp:39-44); text for the last failure via pfaai_gpu_last_error().
 * One host thread per context; one context drives one CUDA device.  There is NO CPU fallback: without a
 * CUDA device pfaai_gpu_create() fails with PFAAI_ERR_CONSTRUCT.
 */
#ifndef PFAAI_GPU_H
#define PFAAI_GPU_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define PFAAI_NTETRAMERS 160000 /* 20^4, interface.hpp:233 */

typedef struct pfaai_gpu_ctx pfaai_gpu_ctx; /* opaque */

/* == DPair<int,int> (utils.hpp:204-225): one posting of F */
typedef struct { int32_t protein, genome; } pfaai_pg_t;
/* == JACTuple<int,double> fields (interface.hpp:61-75), natural C alignment (24 B) */
typedef struct { int32_t gA, gB; double S; int32_t N; int32_t pad_; } pfaai_jac_t;

/* == PFAAI_ERROR_CODE (interface.hpp:39-44).  A translation unit of the reference itself, which already has that enum with
 * the same enumerator names, defines PFAAI_GPU_REFERENCE_ERROR_CODES before including this header (tests/binding). */
#ifndef PFAAI_GPU_REFERENCE_ERROR_CODES
enum pfaai_status { PFAAI_OK = 0, PFAAI_ERR_SQLITE_DB = 1, PFAAI_ERR_SQLITE_MEM_ALLOC = 2, PFAAI_ERR_CONSTRUCT = 3 };
#endif

/* the three drivers of src/main.cpp:177,234,302 */
enum pfaai_mode { PFAAI_ALL_VS_ALL = 0, PFAAI_QUERY_SUBSET = 1, PFAAI_QUERY_DB = 2 };

/* flags for pfaai_gpu_compute* */
#define PFAAI_FLAG_REF_COMPAT_QT 1u /* query-DB mode: reproduce the reference's T-lookup labels (q, nQ+t),
                                       ds_impl.hpp:434-436; needed to match xdb_qt_{jac,aji}.bin */

/* timings of the last build/compute, milliseconds, CUDA events on the context's stream */
typedef struct {
    float h2d_ms;        /* pfaai_gpu_upload_index                                       */
    float build_ms;      /* pfaai_gpu_build total                                        */
    float sort_ms;       /*   of which: radix sort passes                                */
    float pack_ms;       /* row-slab packing inside compute (list path; excludes plan_ms) */
    float tile_ms;       /* the core tile kernel (aji_tile_kernel)                       */
    float d2h_ms;        /* result download                                              */
    int32_t n_launches;  /* kernels launched by the last build + compute                 */
    int32_t n_tiles;     /* tiles the core kernel processed                              */
    int64_t tile_alg_bytes; /* algorithmic bytes of the core kernel   */
    float dict_ms;       /* column-block dictionary build inside compute (list path)     */
    int32_t reserved_;
    int32_t path;        /* tile path of the last compute: 1 = lists (aji_tile_kernel), 2 = dense contraction (aji_mma_kernel) */
    int32_t dense_tau;   /* dense path: posting lists at least this long are tensor-core columns                */
    int64_t dense_k_sum; /* dense path: dense tetramers over all SCPs (padded to 32 per SCP)                    */
    int64_t sparse_items; /* dense path: postings on non-dense tetramers (walked in-kernel as sparse row items)  */
    int64_t e_size;      /* |E| = sum over posting lists of n(n-1)/2 (the reference's E array, all-vs-all), when known, else -1 */
    float prep_ms;       /* dense path: list lengths, dense ids, operand images, sparse items + dictionaries */
    float plan_ms;       /* choosing the tile path (list-length histogram + cost model; host clock); ~0 once cached for this build */
} pfaai_timings_t;

/* ---- life cycle ------------------------------------------------------------------------------ */
int  pfaai_gpu_create(pfaai_gpu_ctx **out, int device, unsigned flags);
void pfaai_gpu_destroy(pfaai_gpu_ctx *ctx);
const char *pfaai_gpu_last_error(const pfaai_gpu_ctx *ctx); /* ctx may be NULL: error of the last create */
/* run on this cudaStream_t (e.g. torch's current stream); default = a stream the context creates */
int  pfaai_gpu_set_stream(pfaai_gpu_ctx *ctx, void *cuda_stream);
/* Multi-GPU row-block sharding (one context per GPU, every context holds the whole index; no data-path collective,
 * row blocks are independent: SURVEY 8(e)).  The target axis of the output (rows of the all-vs-all matrix, columns =
 * target genomes of the other two modes) is cut into blocks of 384 genomes; block b belongs to shard f(b mod 2n),
 * f(x) = x < n ? x : 2n-1-x (folded, balances the triangular all-vs-all work).  A sharded context computes and stores
 * ONLY its own blocks, compactly and in ascending order:
 *   ALL_VS_ALL    [owned rows x G]; only cells with row < column are computed (the rest stays 0): the gathering host
 *                 mirrors, M[j][i] = M[i][j], so every pair is computed, copied D2H and stored exactly once;
 *   QUERY_SUBSET / QUERY_DB   [n_qry x owned targets].
 * The JAC / counts hooks of pfaai_gpu_compute need an unsharded context. */
int  pfaai_gpu_set_shard(pfaai_gpu_ctx *ctx, int shard, int n_shards);
/* ALL_VS_ALL on a sharded context, alternative output addressing: every shard is handed the SAME full G x G matrix (one
 * page-locked host buffer that all contexts of a process -- or, through shared memory registered with cudaHostRegister, all
 * processes of a box -- can address) and stores its cells AND their mirror images M[j][i] straight into it.  The shards' cell
 * sets are disjoint and together cover the whole matrix (zero diagonal included), so when the last shard returns the matrix is
 * complete in host memory: no gather, no host-side mirror pass (src/main.cpp:143-154 done by the kernels' own stores).
 * pfaai_gpu_matrix_shape then reports G x G.  Off (default): compact per-shard matrices as described above.
 * A PAGEABLE matrix goes through a device copy of the whole G x G matrix (zeros outside the shard's cells) and one download:
 * correct for a matrix private to the shard, but it would overwrite the other shards' cells of a SHARED one -- page-lock it. */
int  pfaai_gpu_set_global_matrix(pfaai_gpu_ctx *ctx, int on);
/* Page-lock (and map for device access) a host range the caller allocated itself -- e.g. the POSIX shared-memory mapping of
 * the global matrix above, once in every process that stores into it.  The range is registered in 1 GiB pieces (one
 * cudaHostRegister call over tens of GB of a not-yet-faulted shared mapping was seen to fail when two processes issue it at
 * once); on failure everything registered so far is released again, the CUDA error state is cleared and PFAAI_ERR_CONSTRUCT
 * is returned, so the caller can fall back to compact per-shard matrices.  ptr and bytes must be multiples of the page size. */
int  pfaai_gpu_host_register(void *ptr, size_t bytes);
int  pfaai_gpu_host_unregister(void *ptr, size_t bytes);
/* Which tile kernel the next computes use (BASELINE north_star: "ncu evidence decides between the two paths"):
 *   PFAAI_PATH_AUTO   (default) a cost model on the measured posting-list length distribution picks per compute call;
 *   PFAAI_PATH_LISTS  the dictionary walk (aji_tile_kernel: bit-sliced counters, no tensor cores) -- sparse inputs;
 *   PFAAI_PATH_MMA    the dense contraction (aji_mma_kernel: 0/1 u8 operands on tcgen05 tensor cores + an exact in-kernel
 *                     sparse walk for the non-dense tetramers) -- real-like inputs whose per-SCP vocabulary is dense.  Needs an
 *                     unpartitioned build and at most 255 SCPs; fails with PFAAI_ERR_CONSTRUCT otherwise.
 * Both paths produce identical counts / S / N / AJI.  tau > 0 fixes the density threshold of the dense path (posting lists of at
 * least tau genomes become tensor-core columns, shorter ones are counted by the sparse walk); 0 = chosen by the cost model. */
enum pfaai_path { PFAAI_PATH_AUTO = 0, PFAAI_PATH_LISTS = 1, PFAAI_PATH_MMA = 2 };
int  pfaai_gpu_set_path(pfaai_gpu_ctx *ctx, int path, int tau);
/* Streaming output (printOutput, main.cpp:156-174): with progress enabled, an unsharded ALL_VS_ALL compute
 * into a PINNED host matrix (the tile kernels store straight into it) maintains device-side counters of finished tiles per block
 * of output rows; every store is fenced (__threadfence_system) before its tile is counted.  pfaai_gpu_progress_poll may be
 * called from ANOTHER host thread while pfaai_gpu_compute blocks its caller: rows [0, rows_final) of the matrix are complete
 * (direct and mirrored cells) and will not be written again, so a CSV writer can format them while later row blocks are still
 * being computed; *finished = 1 once the compute has returned (rows_final = all rows).  Other modes / sharded contexts /
 * pageable matrices report rows_final = 0 until the compute has finished. */
int  pfaai_gpu_set_progress(pfaai_gpu_ctx *ctx, int enable);
int  pfaai_gpu_progress_poll(pfaai_gpu_ctx *ctx, int64_t *rows_final, int *finished);
/* |E| of the built index for all-vs-all (sum over (SCP, tetramer) posting lists of n(n-1)/2: the size of the reference's E
 * array, countTetramerTuples ds_helper.hpp:206-265) -- the layout-independent work unit of SURVEY 8(d). */
int  pfaai_gpu_e_size(pfaai_gpu_ctx *ctx, int64_t *e_size);
/* layout of the last compute: axis (0 = rows, 1 = columns) that is sharded, the number of owned indices along it and
 * (ids != NULL) their global indices in storage order.  Unsharded: n = the full extent, ids = 0..n-1. */
int  pfaai_gpu_shard_layout(const pfaai_gpu_ctx *ctx, int *axis, int64_t *n, int32_t *ids);

/* ---- phase 1/2: index --------------------------------------------------------------------------
 * Upload the reference's arrays exactly as DataBaseInterface delivers them (interface.hpp:153-196):
 *   F  [nF]      postings (protein, genome) in (tetramer, protein, genome) order  -- proteinSetGPPairs
 *   Lc [160000]  postings per tetramer                                           -- tetramerOccCounts
 *   T  [P x G]   row-major tetramer counts, or NULL to derive T on the device as the (p,g) histogram of
 *                F (exact in the single-DB modes; query-DB mode must pass the host T) -- proteinTetramerCounts
 * n_genomes is the TOTAL number of genome ids (query-DB mode: nT + nQ, targets first). */
int  pfaai_gpu_upload_index(pfaai_gpu_ctx *ctx, int n_proteins, int n_genomes,
                            const pfaai_pg_t *F, const int32_t *Lc, int64_t nF, const int32_t *T);
/* The same hand-over with the tetramer key of every posting instead of Lc ("tetramer[nF]"): the triples
 * (tetramer, protein, genome) as the reference's range queries return them (scp_db.hpp:121-158), tetramer-major.  Lc is then
 * built on the device (constructLc, ds_helper.hpp:82-109: one pass over the keys, run boundaries, no atomics) and is bit-identical
 * to tetramerOccCounts; pfaai_gpu_get_arrays returns it.  Keys that are not ascending or outside [0, 160000) fail with
 * PFAAI_ERR_CONSTRUCT. */
int  pfaai_gpu_upload_triples(pfaai_gpu_ctx *ctx, int n_proteins, int n_genomes, const int32_t *tetramer,
                              const pfaai_pg_t *F, int64_t nF, const int32_t *T_or_null);
/* Device-ready hand-over (the index cache's fast path): the genome-major lists that pfaai_gpu_build
 * produces -- seg_off [P*G + 1] (64-bit offsets of segment g*P + p) and seg_vals [nF] (tetramers, ascending inside a segment),
 * exactly what pfaai_gpu_get_arrays returns -- plus Lc and T (NULL: T = segment lengths).  No F is uploaded (half the bytes),
 * no K1-K3, no sort: the context is "built" when the call returns, after one validation pass over the lists (offsets, order and
 * range of the tetramers; PFAAI_ERR_CONSTRUCT on a violation).  Serves every compute mode of an unpartitioned context. */
int  pfaai_gpu_import_lists(pfaai_gpu_ctx *ctx, int n_proteins, int n_genomes, int64_t nF, const int32_t *Lc,
                            const int32_t *T_or_null, const int64_t *seg_off, const uint32_t *seg_vals);
/* The same for a caller that fills the index itself on the device (e.g. a broadcast / all-gather over NVLink from
 * the rank that read the database): sizes the context's F [nF], Lc [160000] and T [P x G] buffers and returns their
 * device pointers; the caller writes them on the context's stream (or synchronises) before pfaai_gpu_build().
 * host_T != 0 declares that T will be provided (query-DB mode), otherwise it is derived on the device. */
int  pfaai_gpu_index_device_buffers(pfaai_gpu_ctx *ctx, int n_proteins, int n_genomes, int64_t nF, int host_T,
                                    void **F_dev, void **Lc_dev, void **T_dev);
/* Device build: Lp (exclusive scan), per-(genome,SCP) segment histogram, T, and the genome-major
 * postings by a stable LSD radix sort of the postings on (genome, SCP). */
int  pfaai_gpu_build(pfaai_gpu_ctx *ctx);
/* Partitioned build for sharded ALL_VS_ALL contexts (the per-GPU share of the index work shrinks with the
 * number of GPUs instead of being repeated on every one).  Every shard still receives the whole F / Lc, but
 *   pfaai_gpu_build_partitioned   sorts only the postings of the genomes of ITS OWN row blocks (order-preserving compaction of F,
 *                                 then the same radix sort on 1/n of the items) and builds the column-block dictionaries of those
 *                                 genomes into its chunk of four exchange buffers (dictionary blobs, stage descriptors, column T
 *                                 values, stage ranges).  Chunk sizes depend on (n_genomes, n_proteins, nF, n_shards) only, so
 *                                 all shards agree on the layout without communicating;
 *   pfaai_gpu_exchange_buffers    describes the buffers: shard s owns bytes [s*chunk, s*chunk + used) of each.  The caller
 *                                 all-gathers them in place across the shards' contexts -- NCCL over NVLink for one process
 *                                 per GPU (bench.py), or pfaai_gpu_exchange_local for contexts that live in one process
 *                                 (peer copies; the C++ driver);
 *   pfaai_gpu_finish_partitioned  declares the exchange done; pfaai_gpu_compute*(PFAAI_ALL_VS_ALL) then skips its own
 *                                 dictionary construction.  Results are identical to the replicated pfaai_gpu_build path.
 * After a partitioned build the T / seg parity hooks of pfaai_gpu_get_arrays cover this shard's genomes only. */
typedef struct {
    void   *buf_dev[4];      /* device pointers: dictionary blobs, stage descriptors, column T values, stage ranges */
    int64_t chunk_bytes[4];  /* size of one shard's chunk of each buffer (identical on every shard)              */
    int64_t used_bytes[4];   /* leading bytes of THIS shard's chunk that carry data                              */
    int32_t shard, n_shards;
} pfaai_exchange_t;
int  pfaai_gpu_build_partitioned(pfaai_gpu_ctx *ctx);
/* Optional: the size of one shard's dictionary-blob chunk for the following partitioned builds (0 = the default upper
 * bound, which is safe but about twice what is used).  A caller that rebuilds the same database repeatedly takes the
 * maximum of used_bytes[0] over the shards after the first build and sets it (plus a margin) on EVERY shard, so that the
 * in-place all-gather moves little padding; a shard whose dictionaries do not fit fails the build with PFAAI_ERR_CONSTRUCT. */
int  pfaai_gpu_set_partition_chunk(pfaai_gpu_ctx *ctx, int64_t blob_chunk_bytes);
int  pfaai_gpu_exchange_buffers(pfaai_gpu_ctx *ctx, pfaai_exchange_t *out);
int  pfaai_gpu_finish_partitioned(pfaai_gpu_ctx *ctx);
/* The same while the exchange is still running on another stream: cuda_event (a cudaEvent_t recorded behind the last exchange
 * operation) marks its completion.  The next pfaai_gpu_compute* first runs the tiles that only need this shard's own
 * dictionaries, then waits for the event on the context's stream and runs the rest: the transfer over NVLink overlaps the
 * computation.  The event must stay valid until that compute returns. */
int  pfaai_gpu_finish_partitioned_async(pfaai_gpu_ctx *ctx, void *cuda_event);
/* exchange + finish for n_shards contexts of THIS process (any mix of devices), after each ran build_partitioned */
int  pfaai_gpu_exchange_local(pfaai_gpu_ctx **ctxs, int n_shards);
/* Routed build (sharded ALL_VS_ALL): no GPU ever holds the whole F.  Shard s is given postings [slice_begin, slice_begin + n_slice)
 * of F (any cut of the array into consecutive slices, in shard order) and the full Lc:
 *   pfaai_gpu_upload_index_slice / index_slice_buffers   host / device hand-over of the slice;
 *   pfaai_gpu_route_prepare   turns the slice into sort items tagged with the shard that owns their genome and groups them by
 *                             destination (one radix pass, input order kept): shard d gets items
 *                             [send_offset[d], send_offset[d+1]) of send_dev (8 bytes each);
 *   the caller moves them -- an all-to-all (NCCL over NVLink; counts from send_offset) into the buffer of
 *   pfaai_gpu_route_recv_buffer, sources concatenated in shard order -- or pfaai_gpu_route_local for contexts of one process;
 *   pfaai_gpu_build_routed    histogram, T, sort and the shard's own dictionaries from the received items; from here on the
 *                             context is in the same state as after pfaai_gpu_build_partitioned (exchange, finish, compute).
 * Per GPU this moves |F|/n postings in and out instead of holding all of F, and removes the filter passes over F. */
typedef struct {
    void   *send_dev;          /* device pointer: the slice's items grouped by destination shard                       */
    int64_t send_offset[17];   /* shard d receives items [send_offset[d], send_offset[d+1])                            */
    int32_t shard, n_shards;
} pfaai_route_t;
int  pfaai_gpu_upload_index_slice(pfaai_gpu_ctx *ctx, int n_proteins, int n_genomes, const pfaai_pg_t *F_slice, const int32_t *Lc,
                                  int64_t nF_total, int64_t slice_begin, int64_t n_slice);
int  pfaai_gpu_index_slice_buffers(pfaai_gpu_ctx *ctx, int n_proteins, int n_genomes, int64_t nF_total, int64_t slice_begin,
                                   int64_t n_slice, void **F_slice_dev, void **Lc_dev);
int  pfaai_gpu_route_prepare(pfaai_gpu_ctx *ctx, pfaai_route_t *out);
int  pfaai_gpu_route_recv_buffer(pfaai_gpu_ctx *ctx, int64_t n_items, void **recv_dev);
int  pfaai_gpu_build_routed(pfaai_gpu_ctx *ctx, int64_t n_items);
/* prepare + peer copies + build_routed for the n_shards contexts (shards 0..n-1, in order) of THIS process */
int  pfaai_gpu_route_local(pfaai_gpu_ctx **ctxs, int n_shards);
/* Parity hooks (any pointer may be NULL): Lc/Lp [160000], T [P x G], and the genome-major index:
 * seg_off [G*P+1] (int64, segment g*P+p) and seg_vals [nF] (tetramers, ascending inside a segment). */
int  pfaai_gpu_get_arrays(pfaai_gpu_ctx *ctx, int32_t *Lc, int32_t *Lp, int32_t *T,
                          int64_t *seg_off, int32_t *seg_vals);

/* ---- phase 3/4: AJI ----------------------------------------------------------------------------
 * Computes the output matrix of the mode on the device (it stays there):
 *   ALL_VS_ALL    G x G, symmetric, zero diagonal            (n_qry, qry_ids ignored)
 *   QUERY_SUBSET  n_qry x G, rows in qry_ids order, self cell 0
 *   QUERY_DB      n_qry x (G - n_qry): rows = query ids G-n_qry.., cols = target ids 0..G-n_qry-1
 * A pair without any shared tetramer gets AJI 0 (N = 0). */
int  pfaai_gpu_compute_device(pfaai_gpu_ctx *ctx, int mode, const int32_t *qry_ids, int n_qry, unsigned flags);
/* shape of the matrix of the last compute */
int  pfaai_gpu_matrix_shape(const pfaai_gpu_ctx *ctx, int64_t *rows, int64_t *cols);
/* device pointer of the row-major fp64 matrix of the last compute (valid until the next compute/destroy) */
const double *pfaai_gpu_matrix_dev(const pfaai_gpu_ctx *ctx);
/* copy the matrix to host memory [rows x cols] */
int  pfaai_gpu_download_matrix(pfaai_gpu_ctx *ctx, double *aji_matrix);
/* One call with host buffers: compute + matrix in host memory, plus the optional small-N outputs in the
 * reference's JAC order (initJAC of the mode): jac [n_pairs] (labels, S, N), aji [n_pairs],
 * counts [n_pairs x P] per-SCP intersection counts.  Any output pointer may be NULL.
 * If aji_matrix is page-locked host memory (cudaHostAlloc / cudaHostRegister) the tile kernel stores every computed
 * cell straight into it (the transfer overlaps the computation; no device copy of the matrix exists afterwards:
 * pfaai_gpu_matrix_dev returns NULL); pageable memory is filled by compute_device + download.  Sharded all-vs-all
 * contexts compute only cells with row < column: with a page-locked matrix the other cells are left UNTOUCHED
 * (the gathering host overwrites them by mirroring), with pageable memory they are zero. */
int  pfaai_gpu_compute(pfaai_gpu_ctx *ctx, int mode, const int32_t *qry_ids, int n_qry, unsigned flags,
                       double *aji_matrix, pfaai_jac_t *jac, double *aji, int32_t *counts);
/* number of JAC entries of a mode (nGenomePairs of the three classes) */
int64_t pfaai_gpu_n_pairs(int mode, int n_genomes, int n_qry);

int  pfaai_gpu_get_timings(const pfaai_gpu_ctx *ctx, pfaai_timings_t *out);

/* Measurement hook: the on-chip ceiling of the list path's access pattern on THIS GPU.  Runs the walk loop of
 * the tile kernel (random LDS.32 bitmap probe + rank + predicated LDS.64 mask read + bit-sliced carry-save adds, 24 warps per SM,
 * ids streamed by coalesced LDG.128 from an L2-resident buffer) against a dictionary that stays in shared memory: no bulk
 * copies, no epilogue, no output.  variant 0: random ids, presence density = hit_rate (bank conflicts as in the real walk);
 * variant 1: all lanes of a warp look the same id up (no bank conflicts: the issue-bound ceiling).  steps = walk-loop iterations
 * (8 lookups per lane each).  Replaces nothing in the reference; it is what turns "% of the shared-memory pipe" into "% of what
 * this access pattern can reach". */
typedef struct { double lookups_per_s; int64_t lookups; float ms; int32_t warps_per_sm; int32_t variant; int32_t reserved_; } pfaai_microbench_t;
int  pfaai_gpu_microbench(pfaai_gpu_ctx *ctx, int variant, double hit_rate, int steps, pfaai_microbench_t *out);

/* Exclusive prefix sum without any allocation.  NO COUNTERPART IN THE REFERENCE (its parallelPrefixSum, ds_helper.hpp:112-122,
 * is what pfaai_gpu_build already replaces); this is the same device scan -- int32 inputs, 64-bit running sums -- offered to
 * callers that cannot let this library allocate device memory:
 *   - a plug-in inside a host application whose own memory pool has reserved the GPU (PyTorch's caching allocator, an RMM
 *     arena, a serving process with a fixed budget): a cudaMalloc behind the pool's back fails or eats into that budget;
 *   - a caller on a latency-critical path: cudaMalloc and cudaFree synchronise the whole device.
 * The build path keeps the scan's per-tile partial sums in a cudaMalloc'ed context buffer.  Here they live in a fixed 8 KiB
 * __device__ array inside the library image, so the call needs no context and performs no cudaMalloc / cudaFree / event or
 * stream creation.  The price: every GPU this library is loaded on holds those 8 KiB until the process exits, and n is limited
 * to PFAAI_GPU_STATIC_SCAN_MAX inputs (one partial sum per 4096 inputs; enough for Lc, 160000 entries, and for the segment
 * histogram of up to 4,194,304 (SCP, genome) lists).  A larger n is refused with PFAAI_ERR_CONSTRUCT before anything is launched.
 *   in_dev    [n]      device pointer, int32 counts
 *   out_dev   [n + 1]  device pointer: out[i] = in[0] + ... + in[i-1], out[n] = the grand total
 *   out32_dev [n]      optional (NULL to skip): the same sums truncated to int32 (the reference's Lp type)
 * All three buffers belong to the caller.  The three kernels are enqueued on cuda_stream (a cudaStream_t; NULL = the default
 * stream) of `device` and the call returns without waiting for them: synchronise the stream before reading the outputs.  The
 * caller's current device is left as it was.
 * Set-up: the FIRST call on each device looks the scratch array's device address up (cudaGetSymbolAddress, with that device
 * current) and remembers it.  Make that first call before any stream capture begins.
 * Sharing: all calls on one device, from every thread and stream of the process, use the same array.  It keeps whatever the
 * previous call left there, which is harmless (each call overwrites the entries it reads before reading them).  But two calls
 * whose kernels can run at the same time on one GPU -- different streams without an ordering between them -- overwrite each
 * other's partial sums and give wrong results: keep such calls on one stream or order them with events.
 * Errors: PFAAI_ERR_CONSTRUCT, text from pfaai_gpu_last_error(NULL). */
#define PFAAI_GPU_STATIC_SCAN_MAX 4194304
int  pfaai_gpu_prefix_sum_static(int device, void *cuda_stream, const int32_t *in_dev, int64_t n, int64_t *out_dev,
                                 int32_t *out32_dev);

#ifdef __cplusplus
}
#endif
#endif /* PFAAI_GPU_H */
