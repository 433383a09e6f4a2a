// pfaai_gpu.cu -- C-ABI (include/pfaai_gpu.h) and host orchestration of the sm_100a AJI path.
// No CPU fallback: every entry point either runs the CUDA kernels or returns PFAAI_ERR_CONSTRUCT.
#include <algorithm>
#include <atomic>
#include <chrono>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "aji_tile.cuh"
#include "common.cuh"
#include "index_build.cuh"
#include "microbench.cuh"
#include "mma_index.cuh"
#include "tile_index.cuh"

namespace pfaai {

static std::string g_create_error;

struct DevBuf {
    void *p = nullptr;
    size_t cap = 0;
};

struct Ctx {
    int device = 0;
    int sm_count = 148;
    cudaStream_t stream = nullptr;
    cudaStream_t stream2 = nullptr;      // second instance of the tile kernel behind a dictionary exchange (compute_internal)
    bool own_stream = false;
    std::string err;
    int shard = 0, n_shards = 1;
    bool global_out = false;              // pfaai_gpu_set_global_matrix: a sharded context addresses the full G x G matrix
    // index
    int P = 0, G = 0;
    int64_t nF = 0;
    bool host_T = false, uploaded = false, built = false;
    int max_len = 0;
    DevBuf F, Lc, Lp64, Lp32, T, seg_len, seg_off, itemsA, itemsB, vals, tile_sums, block_hist, block_base, flags;
    // compute
    DevBuf rows, row_t, cols, col_t, ell_nch, ell_off, ell, tiles, out, S_out, N_out, counts_out;
    DevBuf nsub, sub, colT, blob, stage_off, stages, stage_rng;
    // partitioned build (pfaai_gpu_build_partitioned): this shard built the index part / dictionaries of its own row blocks
    DevBuf owner_tbl, warp_cnt, warp_off, cb_list;
    bool partitioned = false, dict_ready = false;
    int64_t nF_owned = 0;
    int64_t slot_cap = 0;                 // dictionary slots (column blocks) per shard chunk
    void *exch_event = nullptr;           // pfaai_gpu_finish_partitioned_async: the exchange is complete when this cudaEvent_t fires
    DevBuf tile_ctr;
    // streaming output (pfaai_gpu_set_progress / pfaai_gpu_progress_poll): per-granule counters of finished tiles, what they must
    // reach, and a private stream + pinned buffer for polling from another host thread while compute blocks its caller
    DevBuf prog;
    bool lists_only = false;                  // pfaai_gpu_import_lists: no F on the device (no partitioned / routed build possible)
    bool prog_enabled = false;
    std::atomic<int> prog_state{0};          // 0 = no compute under way / not armed, 1 = armed (counters valid), 2 = compute finished
    std::mutex prog_mutex;                   // a poll holds it from its state re-check to its last read of prog_need / prog_host; the computing
                                             // thread takes it to leave state 1, so it never frees or replaces them under a poller
    std::vector<int32_t> prog_need;
    int prog_gran = 0;                        // output rows per granule
    int64_t prog_rows = 0;                    // output rows of the armed compute
    cudaStream_t prog_stream = nullptr;
    int32_t *prog_host = nullptr;
    size_t prog_host_cap = 0;
    // dense-contraction path (K5): planning + operands + sparse row items / dictionaries
    int path_pref = PFAAI_PATH_AUTO, tau_pref = 0;
    DevBuf mm_len, mm_hist, mm_did, mm_kcnt, mm_kpad, mm_koff, mm_img_rows, mm_img_cols, mm_colT, mm_rowmax, mm_colmax,
           mm_owned, sp_cnt, sp_nch, sp_tot, sp_off, sp_items, sp_nsub, sp_sub, sp_blob;
    // planning inputs that depend on the built index only, kept per build (reset wherever e_size is): the posting-list length
    // histograms of mma_length_hist, slot 0 = exact (every genome), slot 1 = the planning sample, and what mm_len holds on the device
    std::vector<unsigned long long> plan_hist[2];
    int plan_hist_stride[2] = {0, 0};        // stride the slot was computed with, 0 = empty
    int mm_len_stride = 0;                   // stride of the lengths in mm_len (the dense path needs 1), 0 = none
    int64_t e_size = -1;                  // |E| of the built index (exact length histogram), -1 = not computed for this build
    // routed build: this context holds postings [slice_begin, slice_begin + n_slice) of F only
    bool sliced = false;
    int64_t slice_begin = 0, n_slice = 0, nF_total = 0;
    int64_t route_off[17] = {};
    float route_ms = 0.f;
    DevBuf route_hist, recv;
    int64_t blob_chunk_hint = 0;          // pfaai_gpu_set_partition_chunk: agreed size of a shard's blob chunk (0: default bound)
    int64_t chunk_bytes[4] = {}, used_bytes[4] = {};   // exchange buffers: blob, stages, colT, stage_rng
    int64_t out_rows = 0, out_cols = 0;
    int last_mode = -1;
    bool out_on_device = true;          // false: the last compute stored straight into the caller's pinned matrix
    std::vector<int32_t> owned_rows;   // kernel rows of this shard, in the order they are stored
    int shard_axis = 0;
    // events / stats
    cudaEvent_t ev[16] = {};
    pfaai_timings_t tm = {};
    int launches = 0;
};

void set_error(Ctx *ctx, const std::string &msg) {
    if (ctx) ctx->err = msg; else g_create_error = msg;
}

static int ensure(Ctx *c, DevBuf &b, size_t bytes) {
    if (bytes == 0) bytes = 16;
    if (b.cap >= bytes) return PFAAI_OK;
    if (b.p) { PF_CUDA(c, cudaFree(b.p)); b.p = nullptr; b.cap = 0; }
    // grow with a little slack so that steady-state calls never reallocate
    size_t want = bytes + bytes / 16 + 256;
    PF_CUDA(c, cudaMalloc(&b.p, want));
    b.cap = want;
    return PFAAI_OK;
}
static void release(DevBuf &b) {
    if (b.p) cudaFree(b.p);
    b.p = nullptr; b.cap = 0;
}
template <typename T> static inline T *ptr(DevBuf &b) { return reinterpret_cast<T *>(b.p); }

#define PF_TRY(call) do { int rc__ = (call); if (rc__ != PFAAI_OK) return rc__; } while (0)
#define PF_LAUNCH_CHECK(c) do { (c)->launches++; PF_CUDA((c), cudaGetLastError()); } while (0)

enum { EV_H2D0, EV_H2D1, EV_B0, EV_S0, EV_S1, EV_B1, EV_P0, EV_P1, EV_T0, EV_T1, EV_D0, EV_D1, EV_K0, EV_K1, EV_FORK, EV_JOIN };

static inline int fold_owner(int block, int n_shards) {   // pfaai_gpu_set_shard: block b belongs to shard f(b mod 2n)
    const int f = block % (2 * n_shards);
    return f < n_shards ? f : 2 * n_shards - 1 - f;
}
constexpr int RB_CBS = TILE_ROWS / COLS_PER_TILE;          // column blocks per row block
static_assert(TILE_ROWS % COLS_PER_TILE == 0, "a row block must be a whole number of column blocks");

static float ev_ms(Ctx *c, int a, int b) {
    float ms = 0.f;
    if (cudaEventElapsedTime(&ms, c->ev[a], c->ev[b]) != cudaSuccess) { cudaGetLastError(); return 0.f; }
    return ms;
}

// int32 -> int64 exclusive scan with n+1 outputs
static int launch_scan(Ctx *c, const int32_t *in, int64_t n, int64_t *out, int32_t *out32) {
    if (n <= 0) { PF_CUDA(c, cudaMemsetAsync(out, 0, sizeof(int64_t), c->stream)); return PFAAI_OK; }
    const int64_t ntiles = ceil_div64(n, SCAN_TILE);
    PF_TRY(ensure(c, c->tile_sums, sizeof(int64_t) * (size_t)ntiles));
    k_scan_tile_sums<<<(unsigned)ntiles, SCAN_THREADS, 0, c->stream>>>(in, n, ptr<int64_t>(c->tile_sums));
    PF_LAUNCH_CHECK(c);
    k_scan_small_i64<<<1, 1024, 0, c->stream>>>(ptr<int64_t>(c->tile_sums), ntiles);
    PF_LAUNCH_CHECK(c);
    k_scan_final<<<(unsigned)ntiles, SCAN_THREADS, 0, c->stream>>>(in, n, ptr<int64_t>(c->tile_sums), out, out32);
    PF_LAUNCH_CHECK(c);
    return PFAAI_OK;
}

static int read_flags(Ctx *c, int *flags_host /*[4]*/) {
    PF_CUDA(c, cudaMemcpyAsync(flags_host, c->flags.p, 4 * sizeof(int), cudaMemcpyDeviceToHost, c->stream));
    PF_CUDA(c, cudaStreamSynchronize(c->stream));
    return PFAAI_OK;
}

// K2b + end of the build: sort the n_items items in `src` (ping-pong buffer `dst`) into c->vals, read the flags back
static int sort_and_finish(Ctx *c, uint64_t *src, uint64_t *dst, int64_t n_items, bool partitioned) {
    const int64_t nseg = (int64_t)c->P * c->G;
    // K2b: stable LSD radix sort on (genome*P + protein), one onesweep kernel per 8-bit digit; the digit histograms
    // of all passes come from the segment histogram; the last pass emits the tetramers only
    PF_CUDA(c, cudaEventRecord(c->ev[EV_S0], c->stream));
    if (n_items > 0) {
        int keybits = 1;
        while (keybits < 31 && ((int64_t)1 << keybits) < nseg) keybits++;
        const int passes = (keybits + 7) / 8;
        const int64_t n_tiles = ceil_div64(n_items, OS_TILE);
        const size_t control_bytes = sizeof(uint64_t) * (size_t)(8 + n_tiles * RS_RADIX);
        PF_TRY(ensure(c, c->block_hist, sizeof(unsigned long long) * MAX_PASSES * RS_RADIX));
        PF_TRY(ensure(c, c->block_base, control_bytes));
        PF_CUDA(c, cudaMemsetAsync(c->block_hist.p, 0, sizeof(unsigned long long) * MAX_PASSES * RS_RADIX, c->stream));
        k_digit_hist<<<(unsigned)std::min<int64_t>(ceil_div64(nseg, 256), (int64_t)c->sm_count * 2), 256, 0, c->stream>>>(
            ptr<uint32_t>(c->seg_len), nseg, passes, ptr<unsigned long long>(c->block_hist));
        PF_LAUNCH_CHECK(c);
        k_digit_scan<<<passes, RS_RADIX, 0, c->stream>>>(ptr<unsigned long long>(c->block_hist));
        PF_LAUNCH_CHECK(c);
        PF_CUDA(c, cudaFuncSetAttribute(k_onesweep<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, OS_SMEM));
        PF_CUDA(c, cudaFuncSetAttribute(k_onesweep<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, OS_SMEM));
        for (int pass = 0; pass < passes; pass++) {
            const int shift = TET_BITS + 8 * pass;
            const unsigned long long *base = ptr<unsigned long long>(c->block_hist) + (size_t)pass * RS_RADIX;
            PF_CUDA(c, cudaMemsetAsync(c->block_base.p, 0, control_bytes, c->stream));
            if (pass == passes - 1)
                k_onesweep<true><<<(unsigned)n_tiles, OS_THREADS, OS_SMEM, c->stream>>>(src, n_items, shift, base, ptr<uint64_t>(c->block_base),
                                                                                        nullptr, ptr<uint32_t>(c->vals));
            else
                k_onesweep<false><<<(unsigned)n_tiles, OS_THREADS, OS_SMEM, c->stream>>>(src, n_items, shift, base, ptr<uint64_t>(c->block_base),
                                                                                         dst, nullptr);
            PF_LAUNCH_CHECK(c);
            std::swap(src, dst);
        }
    }
    PF_CUDA(c, cudaEventRecord(c->ev[EV_S1], c->stream));
    PF_CUDA(c, cudaEventRecord(c->ev[EV_B1], c->stream));
    int fl[4];
    PF_TRY(read_flags(c, fl));
    if (fl[0] == 2) { set_error(c, "build: sum(Lc) != |F|"); return PFAAI_ERR_CONSTRUCT; }
    if (fl[0]) { set_error(c, "build: posting with protein/genome id out of range"); return PFAAI_ERR_CONSTRUCT; }
    c->max_len = fl[1];
    c->tm.build_ms = ev_ms(c, EV_B0, EV_B1);
    c->tm.sort_ms = ev_ms(c, EV_S0, EV_S1);
    c->tm.n_launches = c->launches;
    c->built = true;
    c->partitioned = partitioned;
    return PFAAI_OK;
}


// K1-K3 + K2b.  partitioned: only the postings of the genomes this shard owns (the genomes of its row blocks) are turned into
// sort items (order-preserving compaction in two passes over F), so histogram atomics, sort and the genome-major lists shrink
// by the number of shards; seg_off / T keep their global indexing (non-owned segments are empty, non-owned T columns 0).
static int build_internal(Ctx *c, bool partitioned) {
    if (!c) return PFAAI_ERR_CONSTRUCT;
    if (!c->uploaded) { set_error(c, "build: no index uploaded"); return PFAAI_ERR_CONSTRUCT; }
    if (c->lists_only) { set_error(c, "build: the context holds imported lists, not F (pfaai_gpu_import_lists leaves it built; upload F to rebuild)"); return PFAAI_ERR_CONSTRUCT; }
    if (c->sliced) { set_error(c, "build: the context holds a slice of F (use pfaai_gpu_route_prepare / pfaai_gpu_build_routed)"); return PFAAI_ERR_CONSTRUCT; }
    PF_CUDA(c, cudaSetDevice(c->device));
    const int P = c->P, G = c->G;
    const int64_t nF = c->nF, nseg = (int64_t)P * G;
    c->launches = 0;
    c->built = false; c->partitioned = false; c->dict_ready = false; c->exch_event = nullptr; c->e_size = -1; c->plan_hist_stride[0] = c->plan_hist_stride[1] = 0; c->mm_len_stride = 0;
    PF_TRY(ensure(c, c->Lp64, sizeof(int64_t) * (NTETRA + 1)));
    PF_TRY(ensure(c, c->Lp32, sizeof(int32_t) * NTETRA));
    PF_TRY(ensure(c, c->seg_len, sizeof(uint32_t) * (size_t)nseg));
    PF_TRY(ensure(c, c->seg_off, sizeof(int64_t) * (size_t)(nseg + 1)));
    PF_CUDA(c, cudaEventRecord(c->ev[EV_B0], c->stream));
    PF_CUDA(c, cudaMemsetAsync(c->flags.p, 0, 64, c->stream));
    PF_CUDA(c, cudaMemsetAsync(c->seg_len.p, 0, sizeof(uint32_t) * (size_t)nseg, c->stream));
    // K1: Lp = exclusive scan of Lc (parallelPrefixSum)
    PF_TRY(launch_scan(c, ptr<int32_t>(c->Lc), NTETRA, ptr<int64_t>(c->Lp64), ptr<int32_t>(c->Lp32)));
    // K2a + K3: sort items and the (genome, SCP) histogram
    int64_t n_items = nF;
    const unsigned mi_grid = (unsigned)ceil_div64(nF, (int64_t)MI_THREADS * MI_ITEMS);
    if (partitioned) {
        const int ngran = (G + (1 << OWN_SHIFT) - 1) >> OWN_SHIFT;
        std::vector<uint8_t> own(ngran);
        for (int i = 0; i < ngran; i++) own[i] = (uint8_t)fold_owner((int)(((int64_t)i << OWN_SHIFT) / TILE_ROWS), c->n_shards);
        PF_TRY(ensure(c, c->owner_tbl, (size_t)ngran));
        PF_CUDA(c, cudaMemcpyAsync(c->owner_tbl.p, own.data(), (size_t)ngran, cudaMemcpyHostToDevice, c->stream));
        n_items = 0;
        if (nF > 0) {
            const int64_t nwarp = ceil_div64(nF, 32 * MI_ITEMS);
            PF_TRY(ensure(c, c->warp_cnt, sizeof(int32_t) * (size_t)nwarp));
            PF_TRY(ensure(c, c->warp_off, sizeof(int64_t) * (size_t)(nwarp + 1)));
            k_count_owned<<<mi_grid, MI_THREADS, 0, c->stream>>>(ptr<pfaai_pg_t>(c->F), nF, P, G, ptr<uint8_t>(c->owner_tbl), c->shard,
                                                                 ptr<int32_t>(c->warp_cnt), ptr<uint32_t>(c->seg_len), ptr<int>(c->flags));
            PF_LAUNCH_CHECK(c);
            PF_TRY(launch_scan(c, ptr<int32_t>(c->warp_cnt), nwarp, ptr<int64_t>(c->warp_off), nullptr));
            PF_CUDA(c, cudaMemcpyAsync(&n_items, ptr<int64_t>(c->warp_off) + nwarp, sizeof(int64_t), cudaMemcpyDeviceToHost, c->stream));
            PF_CUDA(c, cudaStreamSynchronize(c->stream));
        }
    }
    c->nF_owned = n_items;
    PF_TRY(ensure(c, c->itemsA, sizeof(uint64_t) * (size_t)n_items));
    PF_TRY(ensure(c, c->itemsB, sizeof(uint64_t) * (size_t)n_items));
    PF_TRY(ensure(c, c->vals, sizeof(uint32_t) * (size_t)n_items));
    if (nF > 0) {
        k_make_items<<<mi_grid, MI_THREADS, 0, c->stream>>>(
            ptr<pfaai_pg_t>(c->F), ptr<int64_t>(c->Lp64), nF, P, G, ptr<uint64_t>(c->itemsA), ptr<uint32_t>(c->seg_len),
            ptr<int>(c->flags), partitioned ? ptr<uint8_t>(c->owner_tbl) : nullptr, c->shard,
            partitioned ? ptr<int64_t>(c->warp_off) : nullptr);
        PF_LAUNCH_CHECK(c);
    }
    PF_TRY(launch_scan(c, ptr<int32_t>(c->seg_len), nseg, ptr<int64_t>(c->seg_off), nullptr));
    k_transpose_t<<<(unsigned)ceil_div64(nseg, 256), 256, 0, c->stream>>>(ptr<uint32_t>(c->seg_len), P, G,
                                                                          c->host_T ? nullptr : ptr<int32_t>(c->T),
                                                                          ptr<int>(c->flags) + 1);
    PF_LAUNCH_CHECK(c);
    return sort_and_finish(c, ptr<uint64_t>(c->itemsA), ptr<uint64_t>(c->itemsB), n_items, partitioned);
}

static int build_own_dictionaries(Ctx *c);

}  // namespace pfaai

using namespace pfaai;

extern "C" {

int pfaai_gpu_create(pfaai_gpu_ctx **out, int device, unsigned /*flags*/) {
    if (!out) return PFAAI_ERR_CONSTRUCT;
    *out = nullptr;
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev <= 0) {
        g_create_error = std::string("no CUDA device: ") + cudaGetErrorString(e) + " (this library has no CPU fallback)";
        cudaGetLastError();
        return PFAAI_ERR_CONSTRUCT;
    }
    if (device < 0 || device >= ndev) { g_create_error = "bad device ordinal"; return PFAAI_ERR_CONSTRUCT; }
    Ctx *c = new Ctx();
    c->device = device;
    if (cudaSetDevice(device) != cudaSuccess) { g_create_error = "cudaSetDevice failed"; delete c; return PFAAI_ERR_CONSTRUCT; }
    cudaDeviceProp prop;
    if (cudaGetDeviceProperties(&prop, device) != cudaSuccess) { g_create_error = "cudaGetDeviceProperties failed"; delete c; return PFAAI_ERR_CONSTRUCT; }
    if (prop.major < 10) {
        g_create_error = std::string("device '") + prop.name + "' is sm_" + std::to_string(prop.major) + std::to_string(prop.minor) +
                         "; this library is built for sm_100a only";
        delete c; return PFAAI_ERR_CONSTRUCT;
    }
    c->sm_count = prop.multiProcessorCount;
    if (cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking) != cudaSuccess) { g_create_error = "cudaStreamCreate failed"; delete c; return PFAAI_ERR_CONSTRUCT; }
    c->own_stream = true;
    if (cudaStreamCreateWithFlags(&c->stream2, cudaStreamNonBlocking) != cudaSuccess) { g_create_error = "cudaStreamCreate failed"; delete c; return PFAAI_ERR_CONSTRUCT; }
    for (auto &ev : c->ev)
        if (cudaEventCreate(&ev) != cudaSuccess) { g_create_error = "cudaEventCreate failed"; delete c; return PFAAI_ERR_CONSTRUCT; }
    if (cudaMalloc(&c->flags.p, 128) != cudaSuccess) { g_create_error = "cudaMalloc failed"; delete c; return PFAAI_ERR_CONSTRUCT; }
    c->flags.cap = 128;
    *out = reinterpret_cast<pfaai_gpu_ctx *>(c);
    return PFAAI_OK;
}

void pfaai_gpu_destroy(pfaai_gpu_ctx *h) {
    Ctx *c = reinterpret_cast<Ctx *>(h);
    if (!c) return;
    cudaSetDevice(c->device);
    cudaStreamSynchronize(c->stream);
    DevBuf *all[] = {&c->F, &c->Lc, &c->Lp64, &c->Lp32, &c->T, &c->seg_len, &c->seg_off, &c->itemsA, &c->itemsB, &c->vals,
                     &c->tile_sums, &c->block_hist, &c->block_base, &c->flags, &c->rows, &c->row_t, &c->cols, &c->col_t,
                     &c->ell_nch, &c->ell_off, &c->ell, &c->tiles, &c->out, &c->S_out, &c->N_out, &c->counts_out,
                     &c->nsub, &c->sub, &c->colT, &c->blob, &c->stage_off, &c->stages, &c->stage_rng,
                     &c->owner_tbl, &c->warp_cnt, &c->warp_off, &c->cb_list, &c->tile_ctr, &c->route_hist, &c->recv,
                     &c->mm_len, &c->mm_hist, &c->mm_did, &c->mm_kcnt, &c->mm_kpad, &c->mm_koff, &c->mm_img_rows, &c->mm_img_cols,
                     &c->mm_colT, &c->mm_rowmax, &c->mm_colmax, &c->mm_owned, &c->sp_cnt, &c->sp_nch, &c->sp_tot, &c->sp_off, &c->sp_items,
                     &c->sp_nsub, &c->sp_sub, &c->sp_blob, &c->prog};
    for (DevBuf *b : all) release(*b);
    if (c->prog_stream) cudaStreamDestroy(c->prog_stream);
    if (c->prog_host) cudaFreeHost(c->prog_host);
    for (auto &ev : c->ev) if (ev) cudaEventDestroy(ev);
    if (c->stream2) { cudaStreamSynchronize(c->stream2); cudaStreamDestroy(c->stream2); }
    if (c->own_stream && c->stream) cudaStreamDestroy(c->stream);
    delete c;
}

const char *pfaai_gpu_last_error(const pfaai_gpu_ctx *h) {
    const Ctx *c = reinterpret_cast<const Ctx *>(h);
    return c ? c->err.c_str() : g_create_error.c_str();
}

int pfaai_gpu_set_stream(pfaai_gpu_ctx *h, void *s) {
    Ctx *c = reinterpret_cast<Ctx *>(h);
    if (!c) return PFAAI_ERR_CONSTRUCT;
    PF_CUDA(c, cudaSetDevice(c->device));
    PF_CUDA(c, cudaStreamSynchronize(c->stream));
    if (c->own_stream && c->stream) cudaStreamDestroy(c->stream);
    c->own_stream = false;
    c->stream = reinterpret_cast<cudaStream_t>(s);
    return PFAAI_OK;
}

int pfaai_gpu_set_shard(pfaai_gpu_ctx *h, int shard, int n_shards) {
    Ctx *c = reinterpret_cast<Ctx *>(h);
    // the ownership table of the partitioned / routed builds holds one byte per 128 genomes
    if (!c || n_shards < 1 || n_shards > 255 || shard < 0 || shard >= n_shards) { set_error(c, "bad shard (1 <= n_shards <= 255)"); return PFAAI_ERR_CONSTRUCT; }
    c->shard = shard; c->n_shards = n_shards;
    return PFAAI_OK;
}

int pfaai_gpu_set_path(pfaai_gpu_ctx *h, int path, int tau) {
    Ctx *c = reinterpret_cast<Ctx *>(h);
    if (!c || path < PFAAI_PATH_AUTO || path > PFAAI_PATH_MMA || tau < 0 || tau == 1) { set_error(c, "set_path: bad path / threshold"); return PFAAI_ERR_CONSTRUCT; }
    c->path_pref = path; c->tau_pref = tau;
    return PFAAI_OK;
}

int pfaai_gpu_set_global_matrix(pfaai_gpu_ctx *h, int on) {
    Ctx *c = reinterpret_cast<Ctx *>(h);
    if (!c) return PFAAI_ERR_CONSTRUCT;
    c->global_out = (on != 0);
    return PFAAI_OK;
}

static constexpr size_t HOST_REG_PIECE = (size_t)1 << 30;

int pfaai_gpu_host_unregister(void *ptr, size_t bytes) {
    int rc = PFAAI_OK;
    for (size_t o = 0; o < bytes; o += HOST_REG_PIECE)
        if (cudaHostUnregister(static_cast<char *>(ptr) + o) != cudaSuccess) { cudaGetLastError(); rc = PFAAI_ERR_CONSTRUCT; }
    return rc;
}

int pfaai_gpu_host_register(void *ptr, size_t bytes) {
    if (!ptr || !bytes || (reinterpret_cast<uintptr_t>(ptr) & 4095u) || (bytes & 4095u)) return PFAAI_ERR_CONSTRUCT;
    for (size_t o = 0; o < bytes; o += HOST_REG_PIECE) {
        const size_t n = bytes - o < HOST_REG_PIECE ? bytes - o : HOST_REG_PIECE;
        if (cudaHostRegister(static_cast<char *>(ptr) + o, n, cudaHostRegisterPortable | cudaHostRegisterMapped) != cudaSuccess) {
            cudaGetLastError();                          // not sticky: leave no error behind for the caller's next CUDA call
            pfaai_gpu_host_unregister(ptr, o);
            return PFAAI_ERR_CONSTRUCT;
        }
    }
    return PFAAI_OK;
}

int pfaai_gpu_upload_index(pfaai_gpu_ctx *h, int P, int G, const pfaai_pg_t *F, const int32_t *Lc, int64_t nF,
                           const int32_t *T) {
    Ctx *c = reinterpret_cast<Ctx *>(h);
    if (!c) return PFAAI_ERR_CONSTRUCT;
    if (P <= 0 || G <= 0 || nF < 0 || !Lc || (nF > 0 && !F)) { set_error(c, "upload_index: bad arguments"); return PFAAI_ERR_CONSTRUCT; }
    if ((int64_t)P * G >= (1ll << 31)) { set_error(c, "upload_index: P*G too large"); return PFAAI_ERR_CONSTRUCT; }
    PF_CUDA(c, cudaSetDevice(c->device));
    c->P = P; c->G = G; c->nF = nF; c->host_T = (T != nullptr); c->built = false; c->uploaded = false; c->sliced = false; c->lists_only = false;
    PF_TRY(ensure(c, c->F, sizeof(pfaai_pg_t) * (size_t)nF));
    PF_TRY(ensure(c, c->Lc, sizeof(int32_t) * NTETRA));
    PF_TRY(ensure(c, c->T, sizeof(int32_t) * (size_t)P * G));
    PF_CUDA(c, cudaEventRecord(c->ev[EV_H2D0], c->stream));
    if (nF > 0) PF_CUDA(c, cudaMemcpyAsync(c->F.p, F, sizeof(pfaai_pg_t) * (size_t)nF, cudaMemcpyHostToDevice, c->stream));
    PF_CUDA(c, cudaMemcpyAsync(c->Lc.p, Lc, sizeof(int32_t) * NTETRA, cudaMemcpyHostToDevice, c->stream));
    if (T) PF_CUDA(c, cudaMemcpyAsync(c->T.p, T, sizeof(int32_t) * (size_t)P * G, cudaMemcpyHostToDevice, c->stream));
    PF_CUDA(c, cudaEventRecord(c->ev[EV_H2D1], c->stream));
    PF_CUDA(c, cudaStreamSynchronize(c->stream));   // the caller may reuse its buffers
    c->tm.h2d_ms = ev_ms(c, EV_H2D0, EV_H2D1);
    c->uploaded = true;
    return PFAAI_OK;
}

int pfaai_gpu_upload_triples(pfaai_gpu_ctx *h, int P, int G, const int32_t *tetramer, const pfaai_pg_t *F, int64_t nF, const int32_t *T) {
    Ctx *c = reinterpret_cast<Ctx *>(h);
    if (!c) return PFAAI_ERR_CONSTRUCT;
    if (P <= 0 || G <= 0 || nF < 0 || (nF > 0 && (!F || !tetramer))) { set_error(c, "upload_triples: bad arguments"); return PFAAI_ERR_CONSTRUCT; }
    if ((int64_t)P * G >= (1ll << 31)) { set_error(c, "upload_triples: P*G too large"); return PFAAI_ERR_CONSTRUCT; }
    PF_CUDA(c, cudaSetDevice(c->device));
    c->P = P; c->G = G; c->nF = nF; c->host_T = (T != nullptr); c->built = false; c->uploaded = false; c->sliced = false; c->lists_only = false;
    PF_TRY(ensure(c, c->F, sizeof(pfaai_pg_t) * (size_t)nF));
    PF_TRY(ensure(c, c->Lc, sizeof(int32_t) * NTETRA));
    PF_TRY(ensure(c, c->T, sizeof(int32_t) * (size_t)P * G));
    // the tetramer keys are staged in the (not yet used) sort buffer; two boundary arrays of NTETRA int32 live in Lp32 / Lp64
    PF_TRY(ensure(c, c->itemsA, sizeof(uint64_t) * (size_t)std::max<int64_t>(nF, 1)));
    PF_TRY(ensure(c, c->Lp64, sizeof(int64_t) * (NTETRA + 1)));
    PF_TRY(ensure(c, c->Lp32, sizeof(int32_t) * NTETRA));
    int32_t *tet_dev = ptr<int32_t>(c->itemsA);
    PF_CUDA(c, cudaEventRecord(c->ev[EV_H2D0], c->stream));
    if (nF > 0) {
        PF_CUDA(c, cudaMemcpyAsync(c->F.p, F, sizeof(pfaai_pg_t) * (size_t)nF, cudaMemcpyHostToDevice, c->stream));
        PF_CUDA(c, cudaMemcpyAsync(tet_dev, tetramer, sizeof(int32_t) * (size_t)nF, cudaMemcpyHostToDevice, c->stream));
    }
    if (T) PF_CUDA(c, cudaMemcpyAsync(c->T.p, T, sizeof(int32_t) * (size_t)P * G, cudaMemcpyHostToDevice, c->stream));
    PF_CUDA(c, cudaEventRecord(c->ev[EV_H2D1], c->stream));
    int *err = ptr<int>(c->flags) + 15;
    PF_CUDA(c, cudaMemsetAsync(err, 0, sizeof(int), c->stream));
    PF_CUDA(c, cudaMemsetAsync(c->Lc.p, 0, sizeof(int32_t) * NTETRA, c->stream));
    PF_CUDA(c, cudaMemsetAsync(c->Lp32.p, 0, sizeof(int32_t) * NTETRA, c->stream));
    PF_CUDA(c, cudaMemsetAsync(c->Lp64.p, 0, sizeof(int32_t) * NTETRA, c->stream));
    if (nF > 0) {
        const unsigned grid = (unsigned)std::min<int64_t>(ceil_div64(nF, 256), (int64_t)c->sm_count * 32);
        k_lc_run_begins<<<grid, 256, 0, c->stream>>>(tet_dev, nF, ptr<int32_t>(c->Lp32), ptr<int32_t>(c->Lp64), err);
        PF_LAUNCH_CHECK(c);
        k_lc_run_ends<<<grid, 256, 0, c->stream>>>(tet_dev, nF, ptr<int32_t>(c->Lp32), ptr<int32_t>(c->Lp64), ptr<int32_t>(c->Lc));
        PF_LAUNCH_CHECK(c);
    }
    int eh = 0;
    PF_CUDA(c, cudaMemcpyAsync(&eh, err, sizeof(int), cudaMemcpyDeviceToHost, c->stream));
    PF_CUDA(c, cudaStreamSynchronize(c->stream));
    if (eh) { set_error(c, "upload_triples: tetramer keys must be ascending and inside [0, 160000)"); return PFAAI_ERR_CONSTRUCT; }
    c->tm.h2d_ms = ev_ms(c, EV_H2D0, EV_H2D1);
    c->uploaded = true;
    return PFAAI_OK;
}

int pfaai_gpu_import_lists(pfaai_gpu_ctx *h, int P, int G, int64_t nF, const int32_t *Lc, const int32_t *T, const int64_t *seg_off,
                           const uint32_t *seg_vals) {
    Ctx *c = reinterpret_cast<Ctx *>(h);
    if (!c) return PFAAI_ERR_CONSTRUCT;
    if (P <= 0 || G <= 0 || nF < 0 || !Lc || !seg_off || (nF > 0 && !seg_vals)) { set_error(c, "import_lists: bad arguments"); return PFAAI_ERR_CONSTRUCT; }
    if ((int64_t)P * G >= (1ll << 31)) { set_error(c, "import_lists: P*G too large"); return PFAAI_ERR_CONSTRUCT; }
    PF_CUDA(c, cudaSetDevice(c->device));
    const int64_t nseg = (int64_t)P * G;
    c->P = P; c->G = G; c->nF = nF; c->nF_owned = nF; c->host_T = (T != nullptr); c->built = false; c->uploaded = false; c->sliced = false; c->lists_only = false;
    c->partitioned = false; c->dict_ready = false; c->exch_event = nullptr; c->e_size = -1; c->plan_hist_stride[0] = c->plan_hist_stride[1] = 0; c->mm_len_stride = 0; c->launches = 0;
    c->lists_only = true;
    PF_TRY(ensure(c, c->Lc, sizeof(int32_t) * NTETRA));
    PF_TRY(ensure(c, c->Lp64, sizeof(int64_t) * (NTETRA + 1)));
    PF_TRY(ensure(c, c->Lp32, sizeof(int32_t) * NTETRA));
    PF_TRY(ensure(c, c->T, sizeof(int32_t) * (size_t)nseg));
    PF_TRY(ensure(c, c->seg_off, sizeof(int64_t) * (size_t)(nseg + 1)));
    PF_TRY(ensure(c, c->vals, sizeof(uint32_t) * (size_t)std::max<int64_t>(nF, 1)));
    PF_CUDA(c, cudaEventRecord(c->ev[EV_H2D0], c->stream));
    PF_CUDA(c, cudaMemcpyAsync(c->Lc.p, Lc, sizeof(int32_t) * NTETRA, cudaMemcpyHostToDevice, c->stream));
    PF_CUDA(c, cudaMemcpyAsync(c->seg_off.p, seg_off, sizeof(int64_t) * (size_t)(nseg + 1), cudaMemcpyHostToDevice, c->stream));
    if (nF > 0) PF_CUDA(c, cudaMemcpyAsync(c->vals.p, seg_vals, sizeof(uint32_t) * (size_t)nF, cudaMemcpyHostToDevice, c->stream));
    if (T) PF_CUDA(c, cudaMemcpyAsync(c->T.p, T, sizeof(int32_t) * (size_t)nseg, cudaMemcpyHostToDevice, c->stream));
    PF_CUDA(c, cudaEventRecord(c->ev[EV_H2D1], c->stream));
    PF_CUDA(c, cudaEventRecord(c->ev[EV_B0], c->stream));
    PF_CUDA(c, cudaMemsetAsync(c->flags.p, 0, 64, c->stream));
    PF_TRY(launch_scan(c, ptr<int32_t>(c->Lc), NTETRA, ptr<int64_t>(c->Lp64), ptr<int32_t>(c->Lp32)));   // Lp parity hook
    k_check_lists<<<(unsigned)std::min<int64_t>(ceil_div64(nseg, 8), (int64_t)c->sm_count * 16), 256, 0, c->stream>>>(
        ptr<int64_t>(c->seg_off), ptr<uint32_t>(c->vals), nseg, nF, ptr<int>(c->flags), T ? ptr<int32_t>(c->T) : nullptr, P, G);
    PF_LAUNCH_CHECK(c);
    if (!T) {
        k_t_from_offsets<<<(unsigned)ceil_div64(nseg, 256), 256, 0, c->stream>>>(ptr<int64_t>(c->seg_off), P, G, ptr<int32_t>(c->T));
        PF_LAUNCH_CHECK(c);
    }
    PF_CUDA(c, cudaEventRecord(c->ev[EV_B1], c->stream));
    int fl[4];
    PF_TRY(read_flags(c, fl));
    if (fl[0]) { set_error(c, "import_lists: offsets must ascend from 0 to nF and tetramers must ascend strictly inside a segment, below 160000"); return PFAAI_ERR_CONSTRUCT; }
    if (fl[2]) { set_error(c, "import_lists: a supplied T(p, g) is smaller than the tetramer list of (SCP p, genome g) it describes"); return PFAAI_ERR_CONSTRUCT; }
    c->max_len = fl[1];
    c->tm.h2d_ms = ev_ms(c, EV_H2D0, EV_H2D1);
    c->tm.build_ms = ev_ms(c, EV_B0, EV_B1);
    c->tm.sort_ms = 0.f;
    c->tm.n_launches = c->launches;
    c->uploaded = true; c->built = true;
    return PFAAI_OK;
}

int pfaai_gpu_index_device_buffers(pfaai_gpu_ctx *h, int P, int G, int64_t nF, int host_T, void **F_dev, void **Lc_dev, void **T_dev) {
    Ctx *c = reinterpret_cast<Ctx *>(h);
    if (!c) return PFAAI_ERR_CONSTRUCT;
    if (P <= 0 || G <= 0 || nF < 0 || (int64_t)P * G >= (1ll << 31)) { set_error(c, "index_device_buffers: bad arguments"); return PFAAI_ERR_CONSTRUCT; }
    PF_CUDA(c, cudaSetDevice(c->device));
    c->P = P; c->G = G; c->nF = nF; c->host_T = (host_T != 0); c->built = false; c->uploaded = false; c->sliced = false; c->lists_only = false;
    PF_TRY(ensure(c, c->F, sizeof(pfaai_pg_t) * (size_t)nF));
    PF_TRY(ensure(c, c->Lc, sizeof(int32_t) * NTETRA));
    PF_TRY(ensure(c, c->T, sizeof(int32_t) * (size_t)P * G));
    if (F_dev) *F_dev = c->F.p;
    if (Lc_dev) *Lc_dev = c->Lc.p;
    if (T_dev) *T_dev = c->T.p;
    c->tm.h2d_ms = 0.f;
    c->uploaded = true;
    return PFAAI_OK;
}

int pfaai_gpu_build(pfaai_gpu_ctx *h) { return build_internal(reinterpret_cast<Ctx *>(h), false); }

int pfaai_gpu_build_partitioned(pfaai_gpu_ctx *h) {
    Ctx *c = reinterpret_cast<Ctx *>(h);
    if (!c) return PFAAI_ERR_CONSTRUCT;
    if (c->n_shards < 2) { set_error(c, "build_partitioned: needs a sharded context (pfaai_gpu_set_shard with n_shards > 1)"); return PFAAI_ERR_CONSTRUCT; }
    if (c->host_T) { set_error(c, "build_partitioned: all-vs-all only (T is derived on the device)"); return PFAAI_ERR_CONSTRUCT; }
    PF_TRY(build_internal(c, true));
    PF_TRY(build_own_dictionaries(c));
    return PFAAI_OK;
}

int pfaai_gpu_set_partition_chunk(pfaai_gpu_ctx *h, int64_t blob_chunk_bytes) {
    Ctx *c = reinterpret_cast<Ctx *>(h);
    if (!c || blob_chunk_bytes < 0) { set_error(c, "set_partition_chunk: bad size"); return PFAAI_ERR_CONSTRUCT; }
    c->blob_chunk_hint = blob_chunk_bytes;
    return PFAAI_OK;
}

int pfaai_gpu_exchange_buffers(pfaai_gpu_ctx *h, pfaai_exchange_t *x) {
    Ctx *c = reinterpret_cast<Ctx *>(h);
    if (!c || !x || !c->partitioned) { set_error(c, "exchange_buffers: no partitioned build"); return PFAAI_ERR_CONSTRUCT; }
    void *bufs[4] = {c->blob.p, c->stages.p, c->colT.p, c->stage_rng.p};
    for (int k = 0; k < 4; k++) { x->buf_dev[k] = bufs[k]; x->chunk_bytes[k] = c->chunk_bytes[k]; x->used_bytes[k] = c->used_bytes[k]; }
    x->shard = c->shard; x->n_shards = c->n_shards;
    return PFAAI_OK;
}

int pfaai_gpu_finish_partitioned(pfaai_gpu_ctx *h) {
    Ctx *c = reinterpret_cast<Ctx *>(h);
    if (!c || !c->partitioned || !c->built) { set_error(c, "finish_partitioned: no partitioned build"); return PFAAI_ERR_CONSTRUCT; }
    c->dict_ready = true;
    c->exch_event = nullptr;
    return PFAAI_OK;
}

int pfaai_gpu_finish_partitioned_async(pfaai_gpu_ctx *h, void *cuda_event) {
    Ctx *c = reinterpret_cast<Ctx *>(h);
    if (!c || !c->partitioned || !c->built || !cuda_event) { set_error(c, "finish_partitioned_async: no partitioned build / no event"); return PFAAI_ERR_CONSTRUCT; }
    c->dict_ready = true;
    c->exch_event = cuda_event;
    return PFAAI_OK;
}

int pfaai_gpu_exchange_local(pfaai_gpu_ctx **hs, int n) {
    if (!hs || n < 2) return PFAAI_ERR_CONSTRUCT;
    std::vector<Ctx *> cs(n);
    for (int i = 0; i < n; i++) {
        cs[i] = reinterpret_cast<Ctx *>(hs[i]);
        if (!cs[i] || !cs[i]->partitioned || cs[i]->n_shards != n) { set_error(cs[i], "exchange_local: every shard's context, after build_partitioned"); return PFAAI_ERR_CONSTRUCT; }
    }
    for (int i = 0; i < n; i++) {                       // the producers' builds must have finished
        PF_CUDA(cs[i], cudaSetDevice(cs[i]->device));
        PF_CUDA(cs[i], cudaStreamSynchronize(cs[i]->stream));
    }
    for (int d = 0; d < n; d++) {
        Ctx *dst = cs[d];
        PF_CUDA(dst, cudaSetDevice(dst->device));
        void *db[4] = {dst->blob.p, dst->stages.p, dst->colT.p, dst->stage_rng.p};
        for (int si = 0; si < n; si++) {
            Ctx *src = cs[si];
            if (src == dst) continue;
            void *sb[4] = {src->blob.p, src->stages.p, src->colT.p, src->stage_rng.p};
            for (int k = 0; k < 4; k++) {
                if (src->chunk_bytes[k] != dst->chunk_bytes[k]) { set_error(dst, "exchange_local: shards disagree on the chunk sizes"); return PFAAI_ERR_CONSTRUCT; }
                const size_t off = (size_t)src->shard * (size_t)src->chunk_bytes[k];
                if (src->used_bytes[k] > 0)
                    PF_CUDA(dst, cudaMemcpyPeerAsync(static_cast<char *>(db[k]) + off, dst->device, static_cast<char *>(sb[k]) + off, src->device,
                                                     (size_t)src->used_bytes[k], dst->stream));
            }
        }
    }
    for (int i = 0; i < n; i++) {
        PF_CUDA(cs[i], cudaSetDevice(cs[i]->device));
        PF_CUDA(cs[i], cudaStreamSynchronize(cs[i]->stream));
        cs[i]->dict_ready = true;
    }
    return PFAAI_OK;
}

// ---- routed build: every shard holds a slice of F; postings travel to the shard that owns their genome ------------------
int pfaai_gpu_index_slice_buffers(pfaai_gpu_ctx *h, int P, int G, int64_t nF_total, int64_t slice_begin, int64_t n_slice,
                                  void **F_slice_dev, void **Lc_dev) {
    Ctx *c = reinterpret_cast<Ctx *>(h);
    if (!c) return PFAAI_ERR_CONSTRUCT;
    if (P <= 0 || G <= 0 || nF_total < 0 || slice_begin < 0 || n_slice < 0 || slice_begin + n_slice > nF_total || (int64_t)P * G >= (1ll << 31)) {
        set_error(c, "index_slice_buffers: bad arguments"); return PFAAI_ERR_CONSTRUCT;
    }
    PF_CUDA(c, cudaSetDevice(c->device));
    c->P = P; c->G = G; c->nF = nF_total; c->nF_total = nF_total; c->host_T = false; c->built = false; c->uploaded = false;
    c->sliced = true; c->lists_only = false; c->slice_begin = slice_begin; c->n_slice = n_slice;
    PF_TRY(ensure(c, c->F, sizeof(pfaai_pg_t) * (size_t)n_slice));
    PF_TRY(ensure(c, c->Lc, sizeof(int32_t) * NTETRA));
    PF_TRY(ensure(c, c->T, sizeof(int32_t) * (size_t)P * G));
    if (F_slice_dev) *F_slice_dev = c->F.p;
    if (Lc_dev) *Lc_dev = c->Lc.p;
    c->tm.h2d_ms = 0.f;
    c->uploaded = true;
    return PFAAI_OK;
}

int pfaai_gpu_upload_index_slice(pfaai_gpu_ctx *h, int P, int G, const pfaai_pg_t *F_slice, const int32_t *Lc, int64_t nF_total,
                                 int64_t slice_begin, int64_t n_slice) {
    Ctx *c = reinterpret_cast<Ctx *>(h);
    if (!c || !Lc || (n_slice > 0 && !F_slice)) { set_error(c, "upload_index_slice: bad arguments"); return PFAAI_ERR_CONSTRUCT; }
    PF_TRY(pfaai_gpu_index_slice_buffers(h, P, G, nF_total, slice_begin, n_slice, nullptr, nullptr));
    c->uploaded = false;
    PF_CUDA(c, cudaEventRecord(c->ev[EV_H2D0], c->stream));
    if (n_slice > 0) PF_CUDA(c, cudaMemcpyAsync(c->F.p, F_slice, sizeof(pfaai_pg_t) * (size_t)n_slice, cudaMemcpyHostToDevice, c->stream));
    PF_CUDA(c, cudaMemcpyAsync(c->Lc.p, Lc, sizeof(int32_t) * NTETRA, cudaMemcpyHostToDevice, c->stream));
    PF_CUDA(c, cudaEventRecord(c->ev[EV_H2D1], c->stream));
    PF_CUDA(c, cudaStreamSynchronize(c->stream));
    c->tm.h2d_ms = ev_ms(c, EV_H2D0, EV_H2D1);
    c->uploaded = true;
    return PFAAI_OK;
}

int pfaai_gpu_route_prepare(pfaai_gpu_ctx *h, pfaai_route_t *out) {
    Ctx *c = reinterpret_cast<Ctx *>(h);
    if (!c || !out) return PFAAI_ERR_CONSTRUCT;
    if (!c->uploaded || !c->sliced) { set_error(c, "route_prepare: no index slice uploaded"); return PFAAI_ERR_CONSTRUCT; }
    if (c->n_shards < 2 || c->n_shards > ROUTE_MAX_SHARDS) { set_error(c, "route_prepare: needs 2..16 shards (pfaai_gpu_set_shard)"); return PFAAI_ERR_CONSTRUCT; }
    PF_CUDA(c, cudaSetDevice(c->device));
    const int P = c->P, G = c->G, n = c->n_shards;
    const int64_t ns = c->n_slice;
    c->launches = 0;
    c->built = false; c->partitioned = false; c->dict_ready = false; c->exch_event = nullptr;
    PF_TRY(ensure(c, c->Lp64, sizeof(int64_t) * (NTETRA + 1)));
    PF_TRY(ensure(c, c->Lp32, sizeof(int32_t) * NTETRA));
    PF_TRY(ensure(c, c->itemsA, sizeof(uint64_t) * (size_t)ns));
    PF_TRY(ensure(c, c->itemsB, sizeof(uint64_t) * (size_t)ns));
    PF_TRY(ensure(c, c->route_hist, sizeof(unsigned long long) * RS_RADIX));
    PF_CUDA(c, cudaEventRecord(c->ev[EV_B0], c->stream));
    PF_CUDA(c, cudaMemsetAsync(c->flags.p, 0, 64, c->stream));
    PF_CUDA(c, cudaMemsetAsync(c->route_hist.p, 0, sizeof(unsigned long long) * RS_RADIX, c->stream));
    PF_TRY(launch_scan(c, ptr<int32_t>(c->Lc), NTETRA, ptr<int64_t>(c->Lp64), ptr<int32_t>(c->Lp32)));
    const int ngran = (G + (1 << OWN_SHIFT) - 1) >> OWN_SHIFT;
    std::vector<uint8_t> own(ngran);
    for (int i = 0; i < ngran; i++) own[i] = (uint8_t)fold_owner((int)(((int64_t)i << OWN_SHIFT) / TILE_ROWS), n);
    PF_TRY(ensure(c, c->owner_tbl, (size_t)ngran));
    PF_CUDA(c, cudaMemcpyAsync(c->owner_tbl.p, own.data(), (size_t)ngran, cudaMemcpyHostToDevice, c->stream));
    unsigned long long base[RS_RADIX] = {};
    if (ns > 0) {
        k_make_items_routed<<<(unsigned)ceil_div64(ns, (int64_t)MI_THREADS * MI_ITEMS), MI_THREADS, 0, c->stream>>>(
            ptr<pfaai_pg_t>(c->F), ptr<int64_t>(c->Lp64), c->slice_begin, ns, c->nF_total, P, G, ptr<uint8_t>(c->owner_tbl), n,
            ptr<uint64_t>(c->itemsA), ptr<unsigned long long>(c->route_hist), ptr<int>(c->flags));
        PF_LAUNCH_CHECK(c);
        k_digit_scan<<<1, RS_RADIX, 0, c->stream>>>(ptr<unsigned long long>(c->route_hist));     // counts -> exclusive bases
        PF_LAUNCH_CHECK(c);
        // stable partition by owner: one onesweep pass on the owner digit
        const int64_t n_tiles = ceil_div64(ns, OS_TILE);
        const size_t control_bytes = sizeof(uint64_t) * (size_t)(8 + n_tiles * RS_RADIX);
        PF_TRY(ensure(c, c->block_base, control_bytes));
        PF_CUDA(c, cudaMemsetAsync(c->block_base.p, 0, control_bytes, c->stream));
        PF_CUDA(c, cudaFuncSetAttribute(k_onesweep<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, OS_SMEM));
        k_onesweep<false><<<(unsigned)n_tiles, OS_THREADS, OS_SMEM, c->stream>>>(ptr<uint64_t>(c->itemsA), ns, ROUTE_SHIFT,
                                                                                 ptr<unsigned long long>(c->route_hist),
                                                                                 ptr<uint64_t>(c->block_base), ptr<uint64_t>(c->itemsB), nullptr);
        PF_LAUNCH_CHECK(c);
        PF_CUDA(c, cudaMemcpyAsync(base, c->route_hist.p, sizeof(base), cudaMemcpyDeviceToHost, c->stream));
    }
    PF_CUDA(c, cudaEventRecord(c->ev[EV_B1], c->stream));
    int fl[4];
    PF_TRY(read_flags(c, fl));
    if (fl[0] == 2) { set_error(c, "route_prepare: sum(Lc) != |F|"); return PFAAI_ERR_CONSTRUCT; }
    if (fl[0]) { set_error(c, "route_prepare: posting with protein/genome id out of range"); return PFAAI_ERR_CONSTRUCT; }
    for (int d = 0; d < n; d++) c->route_off[d] = (int64_t)base[d];
    c->route_off[n] = ns;
    c->route_ms = ev_ms(c, EV_B0, EV_B1);
    out->send_dev = c->itemsB.p; out->n_shards = n; out->shard = c->shard;
    for (int d = 0; d <= ROUTE_MAX_SHARDS; d++) out->send_offset[d] = d <= n ? c->route_off[d] : ns;
    return PFAAI_OK;
}

int pfaai_gpu_route_recv_buffer(pfaai_gpu_ctx *h, int64_t n_items, void **recv_dev) {
    Ctx *c = reinterpret_cast<Ctx *>(h);
    if (!c || n_items < 0 || !recv_dev) { set_error(c, "route_recv_buffer: bad arguments"); return PFAAI_ERR_CONSTRUCT; }
    PF_CUDA(c, cudaSetDevice(c->device));
    PF_TRY(ensure(c, c->recv, sizeof(uint64_t) * (size_t)n_items));
    *recv_dev = c->recv.p;
    return PFAAI_OK;
}

int pfaai_gpu_build_routed(pfaai_gpu_ctx *h, int64_t n_items) {
    Ctx *c = reinterpret_cast<Ctx *>(h);
    if (!c || !c->sliced || !c->uploaded || n_items < 0) { set_error(c, "build_routed: no routed items"); return PFAAI_ERR_CONSTRUCT; }
    if (c->recv.cap < sizeof(uint64_t) * (size_t)n_items) { set_error(c, "build_routed: more items than the receive buffer holds"); return PFAAI_ERR_CONSTRUCT; }
    PF_CUDA(c, cudaSetDevice(c->device));
    const int P = c->P, G = c->G;
    const int64_t nseg = (int64_t)P * G;
    PF_TRY(ensure(c, c->seg_len, sizeof(uint32_t) * (size_t)nseg));
    PF_TRY(ensure(c, c->seg_off, sizeof(int64_t) * (size_t)(nseg + 1)));
    PF_TRY(ensure(c, c->itemsA, sizeof(uint64_t) * (size_t)n_items));
    PF_TRY(ensure(c, c->vals, sizeof(uint32_t) * (size_t)n_items));
    c->nF_owned = n_items;
    PF_CUDA(c, cudaEventRecord(c->ev[EV_B0], c->stream));
    PF_CUDA(c, cudaMemsetAsync(c->flags.p, 0, 64, c->stream));
    PF_CUDA(c, cudaMemsetAsync(c->seg_len.p, 0, sizeof(uint32_t) * (size_t)nseg, c->stream));
    if (n_items > 0) {
        k_seg_hist<<<(unsigned)std::min<int64_t>(ceil_div64(n_items, 256), (int64_t)c->sm_count * 16), 256, 0, c->stream>>>(
            ptr<uint64_t>(c->recv), n_items, nseg, ptr<uint32_t>(c->seg_len), ptr<int>(c->flags));
        PF_LAUNCH_CHECK(c);
    }
    PF_TRY(launch_scan(c, ptr<int32_t>(c->seg_len), nseg, ptr<int64_t>(c->seg_off), nullptr));
    k_transpose_t<<<(unsigned)ceil_div64(nseg, 256), 256, 0, c->stream>>>(ptr<uint32_t>(c->seg_len), P, G, ptr<int32_t>(c->T),
                                                                          ptr<int>(c->flags) + 1);
    PF_LAUNCH_CHECK(c);
    PF_TRY(sort_and_finish(c, ptr<uint64_t>(c->recv), ptr<uint64_t>(c->itemsA), n_items, true));
    c->tm.build_ms += c->route_ms;
    PF_TRY(build_own_dictionaries(c));
    return PFAAI_OK;
}

int pfaai_gpu_route_local(pfaai_gpu_ctx **hs, int n) {
    if (!hs || n < 2 || n > ROUTE_MAX_SHARDS) return PFAAI_ERR_CONSTRUCT;
    std::vector<Ctx *> cs(n);
    std::vector<pfaai_route_t> rt(n);
    for (int i = 0; i < n; i++) {
        cs[i] = reinterpret_cast<Ctx *>(hs[i]);
        if (!cs[i] || cs[i]->n_shards != n || cs[i]->shard != i) { set_error(cs[i], "route_local: contexts must be the shards 0..n-1 in order"); return PFAAI_ERR_CONSTRUCT; }
        PF_TRY(pfaai_gpu_route_prepare(hs[i], &rt[i]));
    }
    // all peer copies first (every destination's copies on its own stream: they overlap across the GPUs) ...
    std::vector<int64_t> totals(n, 0);
    for (int d = 0; d < n; d++) {
        Ctx *dst = cs[d];
        int64_t total = 0;
        for (int si = 0; si < n; si++) total += rt[si].send_offset[d + 1] - rt[si].send_offset[d];
        totals[d] = total;
        void *recv = nullptr;
        PF_TRY(pfaai_gpu_route_recv_buffer(hs[d], total, &recv));
        int64_t at = 0;
        for (int si = 0; si < n; si++) {                    // sources in slice order: global posting order is preserved
            const int64_t cnt = rt[si].send_offset[d + 1] - rt[si].send_offset[d];
            if (cnt > 0)
                PF_CUDA(dst, cudaMemcpyPeerAsync(static_cast<uint64_t *>(recv) + at, dst->device,
                                                 static_cast<uint64_t *>(rt[si].send_dev) + rt[si].send_offset[d], cs[si]->device,
                                                 sizeof(uint64_t) * (size_t)cnt, dst->stream));
            at += cnt;
        }
    }
    // ... then every shard's histogram / sort / dictionary build on its own host thread (pfaai_gpu_build_routed is
    // host-synchronous; one thread per context keeps the N device builds concurrent)
    std::vector<int> rcs(n, PFAAI_OK);
    auto work = [&](int d) {
        if (cudaSetDevice(cs[d]->device) != cudaSuccess || cudaStreamSynchronize(cs[d]->stream) != cudaSuccess) {
            set_error(cs[d], "route_local: peer copies failed"); rcs[d] = PFAAI_ERR_CONSTRUCT; return;
        }
        rcs[d] = pfaai_gpu_build_routed(hs[d], totals[d]);
    };
    {
        std::vector<std::thread> th;
        for (int d = 1; d < n; d++) th.emplace_back(work, d);
        work(0);
        for (auto &t : th) t.join();
    }
    for (int d = 0; d < n; d++) if (rcs[d] != PFAAI_OK) return rcs[d];
    return PFAAI_OK;
}

int pfaai_gpu_get_arrays(pfaai_gpu_ctx *h, int32_t *Lc, int32_t *Lp, int32_t *T, int64_t *seg_off, int32_t *seg_vals) {
    Ctx *c = reinterpret_cast<Ctx *>(h);
    if (!c || !c->built) { set_error(c, "get_arrays: index not built"); return PFAAI_ERR_CONSTRUCT; }
    PF_CUDA(c, cudaSetDevice(c->device));
    const size_t nseg = (size_t)c->P * c->G;
    if (Lc) PF_CUDA(c, cudaMemcpyAsync(Lc, c->Lc.p, sizeof(int32_t) * NTETRA, cudaMemcpyDeviceToHost, c->stream));
    if (Lp) PF_CUDA(c, cudaMemcpyAsync(Lp, c->Lp32.p, sizeof(int32_t) * NTETRA, cudaMemcpyDeviceToHost, c->stream));
    if (T) PF_CUDA(c, cudaMemcpyAsync(T, c->T.p, sizeof(int32_t) * nseg, cudaMemcpyDeviceToHost, c->stream));
    if (seg_off) PF_CUDA(c, cudaMemcpyAsync(seg_off, c->seg_off.p, sizeof(int64_t) * (nseg + 1), cudaMemcpyDeviceToHost, c->stream));
    const int64_t n_vals = c->partitioned ? c->nF_owned : c->nF;   // a partitioned / routed context holds its own genomes' lists only
    if (seg_vals && n_vals > 0)
        PF_CUDA(c, cudaMemcpyAsync(seg_vals, c->vals.p, sizeof(int32_t) * (size_t)n_vals, cudaMemcpyDeviceToHost, c->stream));
    PF_CUDA(c, cudaStreamSynchronize(c->stream));
    return PFAAI_OK;
}

int64_t pfaai_gpu_n_pairs(int mode, int G, int nQ) {
    switch (mode) {
    case PFAAI_ALL_VS_ALL: return (int64_t)G * (G - 1) / 2;                                   // ds_impl.hpp:79-81
    case PFAAI_QUERY_SUBSET: return (int64_t)nQ * (G - nQ) + (int64_t)nQ * (nQ - 1) / 2;     // ds_impl.hpp:244-249
    case PFAAI_QUERY_DB: return (int64_t)nQ * (G - nQ);                                       // ds_impl.hpp:403
    default: return -1;
    }
}

}  // extern "C"

// =====================================================================================================
// Phase 3/4: AJI
// =====================================================================================================
namespace pfaai {

struct ModePlan {
    std::vector<int32_t> rows, row_t, cols, col_t;
    int64_t out_rows = 0, out_cols = 0, rs = 0, cs = 0;
    int symmetric = 0;
};

// The id maps of the three reference mode classes (ds_impl.hpp:39-151,159-337,344-490) turned into the
// kernel's row / column genome lists, T-column labels and output strides.
static int plan_mode(Ctx *c, int mode, const int32_t *qry_ids, int nQ, unsigned flags, ModePlan &m) {
    const int G = c->G;
    if (mode == PFAAI_ALL_VS_ALL) {
        m.rows.resize(G);
        for (int g = 0; g < G; g++) m.rows[g] = g;
        m.cols = m.rows; m.row_t = m.rows; m.col_t = m.rows;
        m.out_rows = G; m.out_cols = G; m.rs = G; m.cs = 1; m.symmetric = 1;
    } else if (mode == PFAAI_QUERY_SUBSET) {
        if (nQ <= 0 || nQ > G || !qry_ids) { set_error(c, "compute: bad query subset"); return PFAAI_ERR_CONSTRUCT; }
        std::vector<char> seen(G, 0);
        for (int q = 0; q < nQ; q++) {
            if (qry_ids[q] < 0 || qry_ids[q] >= G || seen[qry_ids[q]]) { set_error(c, "compute: query id out of range or repeated"); return PFAAI_ERR_CONSTRUCT; }
            seen[qry_ids[q]] = 1;
        }
        // output is nQ x G (rows in query-file order, main.cpp:143-154); the kernel streams all genomes as
        // rows against the (few) query genomes as dictionary columns and writes the transpose
        m.rows.resize(G);
        for (int g = 0; g < G; g++) m.rows[g] = g;
        m.row_t = m.rows;
        m.cols.assign(qry_ids, qry_ids + nQ); m.col_t = m.cols;
        m.out_rows = nQ; m.out_cols = G; m.rs = 1; m.cs = G; m.symmetric = 0;
    } else if (mode == PFAAI_QUERY_DB) {
        const int nT = G - nQ;
        if (nQ <= 0 || nT <= 0) { set_error(c, "compute: bad query-DB split"); return PFAAI_ERR_CONSTRUCT; }
        m.rows.resize(nT); m.row_t.resize(nT); m.cols.resize(nQ); m.col_t.resize(nQ);
        const bool quirk = (flags & PFAAI_FLAG_REF_COMPAT_QT) != 0;
        for (int t = 0; t < nT; t++) { m.rows[t] = t; m.row_t[t] = quirk ? nQ + t : t; }            // ds_impl.hpp:434-436
        for (int q = 0; q < nQ; q++) { m.cols[q] = nT + q; m.col_t[q] = quirk ? q : nT + q; }
        m.out_rows = nQ; m.out_cols = nT; m.rs = 1; m.cs = nT; m.symmetric = 0;
    } else { set_error(c, "compute: unknown mode"); return PFAAI_ERR_CONSTRUCT; }
    return PFAAI_OK;
}

static int upload_i32(Ctx *c, DevBuf &b, const std::vector<int32_t> &v) {
    PF_TRY(ensure(c, b, sizeof(int32_t) * v.size()));
    if (!v.empty()) PF_CUDA(c, cudaMemcpyAsync(b.p, v.data(), sizeof(int32_t) * v.size(), cudaMemcpyHostToDevice, c->stream));
    return PFAAI_OK;
}

// K4b: dictionaries of `n_groups` column blocks of the column list c->cols (cb_list: which blocks, NULL = 0..n_groups-1),
// their T values and stage lists.  Group k gets slot slot_base + k; blobs go to [blob_base, blob_cap) of c->blob (blob_cap == 0:
// the whole buffer, grown on demand), stages to [stage_base, stage_cap) of c->stages (stage_cap == 0: sized here).
static int make_dictionaries(Ctx *c, int n_cols, const std::vector<int32_t> *cb_list, int n_groups, int slot_base, int64_t n_slots,
                             size_t blob_base, size_t blob_cap, int64_t stage_base, int64_t stage_cap) {
    const int P = c->P;
    // the same bound compute_internal applies (bit-plane capacity of the walk); it also guarantees that a dictionary fits a
    // stage at the finest column sub-split (max_len + 1 <= MASK_CAP), which k_dict_build relies on
    if (c->max_len / 2 + 16 > 4095) { set_error(c, "dictionaries: a (SCP, genome) tetramer list exceeds 8158 entries"); return PFAAI_ERR_CONSTRUCT; }
    int max_sub = 1;
    while (max_sub < MAX_SUB && (int64_t)(COLS_PER_TILE / max_sub) * c->max_len + 1 > MASK_CAP) max_sub *= 2;
    const int64_t ndict = (int64_t)n_groups * P * NRANGE;
    const bool growable = (blob_cap == 0);
    PF_TRY(ensure(c, c->nsub, sizeof(int32_t) * (size_t)ndict));
    PF_TRY(ensure(c, c->sub, sizeof(SubDesc) * (size_t)(ndict * max_sub)));
    PF_TRY(ensure(c, c->colT, sizeof(int32_t) * (size_t)n_slots * P * COLS_PER_TILE));
    PF_TRY(ensure(c, c->stage_off, sizeof(int64_t) * (size_t)(ndict + 1)));
    PF_TRY(ensure(c, c->stage_rng, sizeof(int64_t) * 2 * (size_t)n_slots));
    if (stage_cap == 0) {
        PF_TRY(ensure(c, c->stages, sizeof(StageDesc) * (size_t)(ndict * max_sub)));
        stage_cap = stage_base + ndict * max_sub;
    }
    if (growable) {
        const int64_t col_postings = std::min<int64_t>(c->nF, (int64_t)n_cols * P * c->max_len);
        const size_t guess = (size_t)ndict * (BM_BYTES + 16) + 8 * (size_t)col_postings;
        if (c->blob.cap < guess) PF_TRY(ensure(c, c->blob, guess));
    }
    if (cb_list) {
        PF_TRY(ensure(c, c->cb_list, sizeof(int32_t) * cb_list->size()));
        if (!cb_list->empty())
            PF_CUDA(c, cudaMemcpyAsync(c->cb_list.p, cb_list->data(), sizeof(int32_t) * cb_list->size(), cudaMemcpyHostToDevice, c->stream));
    }
    PF_CUDA(c, cudaFuncSetAttribute(k_dict_build, cudaFuncAttributeMaxDynamicSharedMemorySize, DB_SMEM));
    PF_CUDA(c, cudaFuncSetAttribute(k_dict_build, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
    unsigned long long *cursor = reinterpret_cast<unsigned long long *>(reinterpret_cast<char *>(c->flags.p) + 32);
    unsigned long long used = 0;
    for (int attempt = 0;; attempt++) {
        PF_CUDA(c, cudaMemsetAsync(reinterpret_cast<char *>(c->flags.p) + 8, 0, 56, c->stream));   // flags[2..], cursor
        DictParams dp;
        dp.seg_off = ptr<int64_t>(c->seg_off); dp.vals = ptr<uint32_t>(c->vals);
        dp.cols = ptr<int32_t>(c->cols); dp.col_t = ptr<int32_t>(c->col_t); dp.n_cols = n_cols;
        dp.T = ptr<int32_t>(c->T); dp.t_ld = c->G; dp.P = P; dp.max_sub = max_sub;
        dp.blob = ptr<unsigned char>(c->blob); dp.blob_base = blob_base; dp.blob_cap = growable ? c->blob.cap : blob_cap; dp.cursor = cursor;
        dp.cb_list = cb_list ? ptr<int32_t>(c->cb_list) : nullptr; dp.slot_base = slot_base;
        dp.nsub = ptr<int32_t>(c->nsub); dp.sub = ptr<SubDesc>(c->sub); dp.colT = ptr<int32_t>(c->colT);
        dp.err = ptr<int>(c->flags) + 3;
        if (ndict > 0) {
            k_dict_build<<<(unsigned)(n_groups * P), DB_THREADS, DB_SMEM, c->stream>>>(dp);
            PF_LAUNCH_CHECK(c);
        }
        unsigned long long fl[4];   // {flags[2], flags[3]}, -, -, cursor at byte 32
        PF_CUDA(c, cudaMemcpyAsync(fl, reinterpret_cast<char *>(c->flags.p) + 8, 32, cudaMemcpyDeviceToHost, c->stream));
        PF_CUDA(c, cudaStreamSynchronize(c->stream));
        used = fl[3];
        const bool overflow = (fl[0] >> 32) != 0;
        if (!overflow) break;
        if (!growable) {
            set_error(c, "build_partitioned: this shard's dictionaries need " + std::to_string(used) + " bytes, the agreed chunk holds " +
                             std::to_string(blob_cap - blob_base) + " (unbalanced input: use pfaai_gpu_build)");
            return PFAAI_ERR_CONSTRUCT;
        }
        if (attempt > 0) { set_error(c, "compute: dictionary buffer overflow after resize"); return PFAAI_ERR_CONSTRUCT; }
        PF_TRY(ensure(c, c->blob, (size_t)fl[3] + 4096));
    }
    c->used_bytes[0] = (int64_t)used;
    PF_TRY(launch_scan(c, ptr<int32_t>(c->nsub), ndict, ptr<int64_t>(c->stage_off), nullptr));
    if (ndict > 0) {
        k_stage_fill<<<(unsigned)ceil_div64(ndict, 256), 256, 0, c->stream>>>(ptr<int32_t>(c->nsub), ptr<int64_t>(c->stage_off),
                                                                            ptr<SubDesc>(c->sub), n_groups, P, max_sub,
                                                                            ptr<StageDesc>(c->stages), ptr<int64_t>(c->stage_rng), slot_base,
                                                                            stage_base, stage_cap, ptr<int>(c->flags) + 2);
        PF_LAUNCH_CHECK(c);
    }
    return PFAAI_OK;
}

// Partitioned build, second half: the dictionaries of the column blocks inside this shard's own row blocks (all-vs-all: the
// column genomes are the row genomes), written into this shard's chunk of the four exchange buffers.  Chunk sizes are a function
// of (G, P, |F|, n_shards) only, so every shard computes the same layout without talking to the others.
static int build_own_dictionaries(Ctx *c) {
    const int P = c->P, G = c->G, n = c->n_shards;
    const int ncb = (G + COLS_PER_TILE - 1) / COLS_PER_TILE, nrb = (G + TILE_ROWS - 1) / TILE_ROWS;
    std::vector<int> nblk(n, 0);
    std::vector<int32_t> own_cbs;
    int64_t max_rows = 0;
    std::vector<int64_t> rows_of(n, 0);
    for (int I = 0; I < nrb; I++) {
        const int o = fold_owner(I, n);
        nblk[o]++;
        rows_of[o] += std::min(TILE_ROWS, G - I * TILE_ROWS);
        if (o == c->shard)
            for (int j = 0; j < RB_CBS && I * RB_CBS + j < ncb; j++) own_cbs.push_back(I * RB_CBS + j);
    }
    for (int o = 0; o < n; o++) max_rows = std::max(max_rows, rows_of[o]);
    const int max_blk = *std::max_element(nblk.begin(), nblk.end());
    c->slot_cap = (int64_t)max_blk * RB_CBS;
    const int64_t dict_cap = c->slot_cap * P * NRANGE;
    // blob chunk: one bitmap (+ padding) per dictionary, 2x head room for column sub-splits on the first quarter, one mask per
    // posting of the largest shard + 20 % (tetramers shared by several columns of a block need only one mask)
    const int64_t mask_bytes = 8 * (int64_t)((double)c->nF * ((double)max_rows / (double)G) * 1.2) + 4096;
    c->chunk_bytes[0] = ((dict_cap + dict_cap / 4) * (int64_t)(BM_BYTES + 16) + mask_bytes + 255) & ~(int64_t)255;
    if (c->blob_chunk_hint > 0) c->chunk_bytes[0] = (c->blob_chunk_hint + 255) & ~(int64_t)255;
    c->chunk_bytes[1] = (int64_t)sizeof(StageDesc) * dict_cap * 2;
    c->chunk_bytes[2] = (int64_t)sizeof(int32_t) * c->slot_cap * P * COLS_PER_TILE;
    c->chunk_bytes[3] = (int64_t)sizeof(int64_t) * 2 * c->slot_cap;
    PF_TRY(ensure(c, c->blob, (size_t)c->chunk_bytes[0] * n));
    PF_TRY(ensure(c, c->stages, (size_t)c->chunk_bytes[1] * n));
    PF_TRY(ensure(c, c->colT, (size_t)c->chunk_bytes[2] * n));
    PF_TRY(ensure(c, c->stage_rng, (size_t)c->chunk_bytes[3] * n));
    // all-vs-all column plan: column j = genome j
    std::vector<int32_t> ident(G);
    for (int g = 0; g < G; g++) ident[g] = g;
    PF_CUDA(c, cudaEventRecord(c->ev[EV_K0], c->stream));
    PF_TRY(upload_i32(c, c->cols, ident)); PF_TRY(upload_i32(c, c->col_t, ident));
    PF_CUDA(c, cudaMemsetAsync(static_cast<char *>(c->stage_rng.p) + (size_t)c->shard * c->chunk_bytes[3], 0, (size_t)c->chunk_bytes[3], c->stream));
    const int64_t stage_base = (int64_t)c->shard * dict_cap * 2;
    PF_TRY(make_dictionaries(c, G, &own_cbs, (int)own_cbs.size(), (int)(c->shard * c->slot_cap), c->slot_cap * n,
                             (size_t)c->shard * (size_t)c->chunk_bytes[0], (size_t)(c->shard + 1) * (size_t)c->chunk_bytes[0],
                             stage_base, stage_base + dict_cap * 2));
    PF_CUDA(c, cudaEventRecord(c->ev[EV_K1], c->stream));
    int64_t n_stages = 0;
    const int64_t ndict = (int64_t)own_cbs.size() * P * NRANGE;
    int fl[4];
    PF_CUDA(c, cudaMemcpyAsync(&n_stages, ptr<int64_t>(c->stage_off) + ndict, sizeof(int64_t), cudaMemcpyDeviceToHost, c->stream));
    PF_TRY(read_flags(c, fl));
    if (fl[2]) { set_error(c, "build_partitioned: this shard's stage list exceeds the agreed chunk (many column sub-splits: use pfaai_gpu_build)"); return PFAAI_ERR_CONSTRUCT; }
    c->used_bytes[1] = (int64_t)sizeof(StageDesc) * n_stages;
    c->used_bytes[2] = (int64_t)sizeof(int32_t) * (int64_t)own_cbs.size() * P * COLS_PER_TILE;
    c->used_bytes[3] = c->chunk_bytes[3];
    c->tm.dict_ms = ev_ms(c, EV_K0, EV_K1);
    c->tm.build_ms += c->tm.dict_ms;
    c->tm.n_launches = c->launches;
    return PFAAI_OK;
}


// =====================================================================================================
// K5 (dense-contraction path): planning and preparation.  DESIGN.md section 4b.
// =====================================================================================================
struct MmaPlan {
    bool use = false;
    bool exact = false;                   // the length histogram covers every genome (not the planning sample)
    uint32_t tau = 0;
    int64_t k_sum = 0;                    // dense tetramers over all SCPs, padded to 32 per SCP
    int64_t sparse_est = 0;               // postings on non-dense tetramers (estimate)
    int64_t sparse_items = 0;
    bool has_sparse = false;
    bool small_dict = false;              // MmCfgDense: few non-dense tetramers per (column block, SCP) -> small dictionary stage, deep operand ring
    double t_mma = 0, t_lists = 0;        // cost-model seconds
    std::vector<unsigned long long> hist; // [0..HMAX] lists per length (sample units when !exact), [HMAX+1] = sum n(n-1)/2
    int stride = 1;
};

// length histogram of the posting lists; stride > 1: every stride-th genome only (the planning estimate, ~|F|/stride work)
static int mma_length_hist(Ctx *c, int stride, MmaPlan &pl, bool need_device_lengths = false) {
    const int P = c->P, G = c->G;
    const int slot = stride == 1 ? 0 : 1;
    if (c->plan_hist_stride[slot] == stride && (!need_device_lengths || c->mm_len_stride == stride)) {
        // same build, same sample: the histogram (and |E|) cannot have changed -- no memset, no pass over the postings, no host sync
        pl.hist = c->plan_hist[slot];
        pl.stride = stride; pl.exact = (stride == 1);
        return PFAAI_OK;
    }
    const int64_t nlen = (int64_t)P * NTETRA;
    PF_TRY(ensure(c, c->mm_len, sizeof(uint32_t) * (size_t)nlen));
    PF_TRY(ensure(c, c->mm_hist, sizeof(unsigned long long) * (MM_HMAX + 2)));
    PF_CUDA(c, cudaMemsetAsync(c->mm_len.p, 0, sizeof(uint32_t) * (size_t)nlen, c->stream));
    PF_CUDA(c, cudaMemsetAsync(c->mm_hist.p, 0, sizeof(unsigned long long) * (MM_HMAX + 2), c->stream));
    const int64_t nseg = (int64_t)((G + stride - 1) / stride) * P;
    if (nseg > 0 && c->nF > 0) {
        k_list_len<<<(unsigned)std::min<int64_t>(ceil_div64(nseg, 8), (int64_t)c->sm_count * 16), 256, 0, c->stream>>>(
            ptr<int64_t>(c->seg_off), ptr<uint32_t>(c->vals), G, P, stride, ptr<uint32_t>(c->mm_len));
        PF_LAUNCH_CHECK(c);
    }
    k_len_hist<<<(unsigned)(c->sm_count * 4), 256, 0, c->stream>>>(ptr<uint32_t>(c->mm_len), nlen, ptr<unsigned long long>(c->mm_hist),
                                                                   ptr<unsigned long long>(c->mm_hist) + MM_HMAX + 1);
    PF_LAUNCH_CHECK(c);
    pl.hist.assign(MM_HMAX + 2, 0ull);
    PF_CUDA(c, cudaMemcpyAsync(pl.hist.data(), c->mm_hist.p, sizeof(unsigned long long) * (MM_HMAX + 2), cudaMemcpyDeviceToHost, c->stream));
    PF_CUDA(c, cudaStreamSynchronize(c->stream));
    pl.stride = stride; pl.exact = (stride == 1);
    if (pl.exact) c->e_size = (int64_t)pl.hist[MM_HMAX + 1];
    c->plan_hist[slot] = pl.hist; c->plan_hist_stride[slot] = stride; c->mm_len_stride = stride;
    return PFAAI_OK;
}

// The decision rule between the two tile paths (north_star: dense contraction only where the per-SCP tetramer density makes it
// the denser one).  Inputs: the posting-list length histogram, the number of matrix cells this context computes, |F|.
// Constants are measured rates of this repository's kernels on B200:
//   list path    cells * (|F|/G) / 64 dictionary lookups at 8.2e11/s;
//   hybrid path  per (128 x 128 tile, SCP) three things run concurrently on one SM -- (a) the operand stream of the contraction,
//                256 bytes per dense tetramer through a bulk-copy ring whose round trip is ~6,400 cycles (17.5 B/cycle with the
//                112 KB ring of the dense plan, 7.5 B/cycle with the 48 KB ring of the sparse plan); (b) the sparse walk,
//                rows' non-dense postings / 128 columns lookups at PF_WALK_S seconds each (whole GPU) plus PF_INC_S per counter
//                increment (= co-occurrence on a non-dense tetramer: shared-memory atomics, 2 cycles per lane); (c) the fp64
//                epilogue, PF_EPI_S seconds per (cell, SCP).  (b) and (c) share the same warps, (a) overlaps them.
// tau moves tetramers between (a) and (b).
#ifndef PF_WALK_S
#define PF_WALK_S 2.1e-12     // one lookup of the walk: 0.6 SM-cycles (measured, profiles/r2n_*), 148 SMs at 1.9 GHz
#endif
#ifndef PF_EPI_S
#define PF_EPI_S 1.4e-12
#endif
#ifndef PF_INC_S
#define PF_INC_S 7.1e-12      // one counter increment of the walk: a shared-memory atomic costs 2 SM-cycles per lane (measured; the reason
                              // the sparse synthetic family, ~1 non-dense co-occurrence per (pair, SCP), stays on the list path)
#endif
static void mma_choose(const Ctx *c, MmaPlan &pl, double cells, int64_t img_blocks, bool force) {
    static const uint32_t cand[] = {2, 3, 4, 6, 8, 12, 16, 24, 32, 48, 64, 96, 128, 192, 256, 384, 512, 768, 1024, 1536, 2048, 3072, 4096};
    const int P = c->P, G = c->G;
    const double s = (double)pl.stride, nF = (double)c->nF;
    pl.t_lists = cells * (nF / std::max(1, G)) / 64.0 / 8.2e11 + nF * 1.4e-11;
    const double sm_hz = 1.9e9 * c->sm_count;
    double best = 1e300;
    for (uint32_t tau : cand) {
        if (c->tau_pref > 0 && tau != cand[0]) break;
        if (c->tau_pref > 0) tau = (uint32_t)std::min(c->tau_pref, MM_HMAX);   // lists of >= MM_HMAX genomes are always dense
        const uint32_t m0 = (uint32_t)std::max(1.0, std::ceil((double)tau / s));   // sample count at which a list is called dense
        double K = 0, dense_post = 0, e_sparse = 0;
        for (uint32_t m = m0; m <= (uint32_t)MM_HMAX; m++) { K += (double)pl.hist[m]; dense_post += (double)pl.hist[m] * m * s; }
        for (uint32_t m = 2; m < std::min<uint32_t>(m0, MM_HMAX); m++) e_sparse += (double)pl.hist[m] * 0.5 * (m * s) * (m * s);   // pair increments
        const double sparse_post = std::max(0.0, nF - dense_post);                  // includes the lists the sample never saw
        const double k_sum = K + 16.0 * P;                                          // padding to 32 per SCP, on average
        const double img_bytes = (double)img_blocks * 128.0 * k_sum;
        if (c->tau_pref == 0 && img_bytes > 48e9) continue;
        const bool small_dict = 2.0 * sparse_post * SP_COLS / ((double)std::max(1, G) * P) + 64.0 < (double)SP_ENT_MAX_SMALL;
        const double t_op = cells / 16384.0 * 256.0 * k_sum / ((small_dict ? 17.5 : 7.5) * sm_hz);
        const double lookups = cells / 128.0 * sparse_post / std::max(1, G);
        const double incs = e_sparse * std::min(1.0, cells / std::max(1.0, 0.5 * (double)G * ((double)G - 1.0)));
        const double t = std::max(t_op, cells * P * PF_EPI_S + lookups * PF_WALK_S + incs * PF_INC_S) + img_bytes / 3e12 + sparse_post * 4e-11;
        if (t < best) { best = t; pl.tau = tau; pl.k_sum = (int64_t)k_sum; pl.sparse_est = (int64_t)sparse_post; }
    }
    pl.t_mma = best + nF * 2e-11;
    pl.use = force ? (best < 1e299) : (pl.t_mma < pl.t_lists);
}

// dense ids, operand images, column T values, sparse row items and dictionaries (everything aji_mma_kernel reads); sets prm's K5 pointers
static int mma_prepare(Ctx *c, const ModePlan &m, MmaPlan &pl, int n_rows, int n_cols, const std::vector<uint8_t> &rb_owned,
                       std::vector<int32_t> &kpad_host, MmaParams &prm) {
    const int P = c->P, G = c->G;
    const int64_t nlen = (int64_t)P * NTETRA;
    const int nrb = (n_rows + MM_ROWS - 1) / MM_ROWS, ncb = (n_cols + MM_COLS - 1) / MM_COLS;
    constexpr int KMAX = 16384;
    // 0. T values per column block, block maxima (fast-division guard) and their overall maximum
    PF_TRY(ensure(c, c->mm_colT, sizeof(int32_t) * (size_t)ncb * P * MM_COLS));
    PF_TRY(ensure(c, c->mm_colmax, sizeof(int32_t) * (size_t)ncb * P));
    PF_TRY(ensure(c, c->mm_rowmax, sizeof(int32_t) * (size_t)nrb * P));
    int *err = ptr<int>(c->flags) + 12, *tmax = ptr<int>(c->flags) + 13;
    PF_CUDA(c, cudaMemsetAsync(err, 0, 2 * sizeof(int), c->stream));
    if (ncb > 0) {
        k_block_t<<<(unsigned)(ncb * P), 128, 0, c->stream>>>(ptr<int32_t>(c->T), c->G, P, ptr<int32_t>(c->col_t), n_cols, ptr<int32_t>(c->mm_colT),
                                                              ptr<int32_t>(c->mm_colmax));
        PF_LAUNCH_CHECK(c);
        k_max_t<<<64, 256, 0, c->stream>>>(ptr<int32_t>(c->mm_colmax), (int64_t)ncb * P, tmax);
        PF_LAUNCH_CHECK(c);
    }
    if (nrb > 0) {
        k_block_t<<<(unsigned)(nrb * P), 128, 0, c->stream>>>(ptr<int32_t>(c->T), c->G, P, ptr<int32_t>(c->row_t), n_rows, nullptr,
                                                              ptr<int32_t>(c->mm_rowmax));
        PF_LAUNCH_CHECK(c);
        k_max_t<<<64, 256, 0, c->stream>>>(ptr<int32_t>(c->mm_rowmax), (int64_t)nrb * P, tmax);
        PF_LAUNCH_CHECK(c);
    }
    // 1. dense tetramers of every SCP
    PF_TRY(ensure(c, c->mm_did, sizeof(uint16_t) * (size_t)nlen));
    PF_TRY(ensure(c, c->mm_kcnt, sizeof(int32_t) * (size_t)P));
    k_dense_ids<<<(unsigned)P, 1024, 0, c->stream>>>(ptr<uint32_t>(c->mm_len), pl.tau, ptr<uint16_t>(c->mm_did), ptr<int32_t>(c->mm_kcnt), KMAX, err);
    PF_LAUNCH_CHECK(c);
    std::vector<int32_t> kcnt(P);
    int eh[2] = {0, 0};
    PF_CUDA(c, cudaMemcpyAsync(kcnt.data(), c->mm_kcnt.p, sizeof(int32_t) * (size_t)P, cudaMemcpyDeviceToHost, c->stream));
    PF_CUDA(c, cudaMemcpyAsync(eh, err, 2 * sizeof(int), cudaMemcpyDeviceToHost, c->stream));
    PF_CUDA(c, cudaStreamSynchronize(c->stream));
    const int err_h = eh[0];
    prm.all_fast = (2 * (int64_t)eh[1] < RCP_N) ? 1 : 0;
    if (err_h) { set_error(c, "dense path: an SCP has more than 16384 dense tetramers at this threshold"); return PFAAI_ERR_CONSTRUCT; }
    kpad_host.assign(P, 0);
    std::vector<int64_t> koff(P, 0);
    int64_t k_sum = 0;
    for (int p = 0; p < P; p++) { kpad_host[p] = (kcnt[p] + 31) & ~31; koff[p] = k_sum; k_sum += kpad_host[p]; }
    pl.k_sum = k_sum;
    PF_TRY(ensure(c, c->mm_kpad, sizeof(int32_t) * (size_t)P));
    PF_TRY(ensure(c, c->mm_koff, sizeof(int64_t) * (size_t)P));
    PF_CUDA(c, cudaMemcpyAsync(c->mm_kpad.p, kpad_host.data(), sizeof(int32_t) * (size_t)P, cudaMemcpyHostToDevice, c->stream));
    PF_CUDA(c, cudaMemcpyAsync(c->mm_koff.p, koff.data(), sizeof(int64_t) * (size_t)P, cudaMemcpyHostToDevice, c->stream));
    // 2. operand images (all-vs-all: the column list IS the row list)
    const int64_t blk_stride = 128 * k_sum;
    auto fill = [&](DevBuf &img, const DevBuf &list, int n_list, int nblk) -> int {
        PF_TRY(ensure(c, img, (size_t)nblk * (size_t)blk_stride));
        if (blk_stride == 0 || n_list == 0) return PFAAI_OK;
        PF_CUDA(c, cudaMemsetAsync(img.p, 0, (size_t)nblk * (size_t)blk_stride, c->stream));
        k_image_fill<<<(unsigned)std::min<int64_t>(ceil_div64((int64_t)n_list * P, 8), (int64_t)c->sm_count * 16), 256, 0, c->stream>>>(
            ptr<int64_t>(c->seg_off), ptr<uint32_t>(c->vals), reinterpret_cast<const int32_t *>(list.p), n_list, P, ptr<uint16_t>(c->mm_did),
            ptr<int64_t>(c->mm_koff), blk_stride, reinterpret_cast<unsigned char *>(img.p));
        PF_LAUNCH_CHECK(c);
        return PFAAI_OK;
    };
    PF_TRY(fill(c->mm_img_rows, c->rows, n_rows, nrb));
    const bool same = m.symmetric != 0;
    if (!same) PF_TRY(fill(c->mm_img_cols, c->cols, n_cols, ncb));
    // 4. sparse side: row items of the kernel rows and the column blocks' sparse dictionaries (sparse_index.cuh) -- every
    //    co-occurrence on a tetramer whose list is shorter than tau is counted by the tile kernel's walk
    double e_sparse = 0;
    for (uint32_t L = 1; L < std::min<uint32_t>(pl.tau, MM_HMAX); L++) e_sparse += (double)pl.hist[L] * L;   // non-dense postings
    pl.has_sparse = false; pl.sparse_items = 0;
    pl.small_dict = 2.0 * e_sparse * SP_COLS / ((double)std::max(1, G) * P) + 64.0 < (double)SP_ENT_MAX_SMALL;
    if (e_sparse > 0 && nrb > 0 && ncb > 0) {
        const int64_t n_gp = (int64_t)nrb * 4 * P;                   // (32-row group, SCP) slabs; groups beyond the last row stay empty
        PF_TRY(ensure(c, c->sp_cnt, sizeof(int32_t) * (size_t)n_gp * SP_ROWS));
        PF_TRY(ensure(c, c->sp_nch, sizeof(int32_t) * (size_t)n_gp));
        PF_TRY(ensure(c, c->sp_tot, sizeof(int32_t) * (size_t)n_gp));
        PF_TRY(ensure(c, c->sp_off, sizeof(int64_t) * (size_t)(n_gp + 1)));
        PF_CUDA(c, cudaMemsetAsync(c->sp_cnt.p, 0, sizeof(int32_t) * (size_t)n_gp * SP_ROWS, c->stream));
        PF_TRY(ensure(c, c->mm_owned, rb_owned.size()));
        PF_CUDA(c, cudaMemcpyAsync(c->mm_owned.p, rb_owned.data(), rb_owned.size(), cudaMemcpyHostToDevice, c->stream));
        const unsigned wgrid = (unsigned)std::min<int64_t>(ceil_div64((int64_t)n_rows * P, 8), (int64_t)c->sm_count * 16);
        k_sp_count<<<wgrid, 256, 0, c->stream>>>(ptr<int64_t>(c->seg_off), ptr<uint32_t>(c->vals), ptr<int32_t>(c->rows), n_rows, P,
                                                 ptr<uint16_t>(c->mm_did), ptr<uint8_t>(c->mm_owned), ptr<int32_t>(c->sp_cnt));
        PF_LAUNCH_CHECK(c);
        k_sp_group<<<(unsigned)ceil_div64(n_gp, 256), 256, 0, c->stream>>>(ptr<int32_t>(c->sp_cnt), n_gp, ptr<int32_t>(c->sp_nch), ptr<int32_t>(c->sp_tot));
        PF_LAUNCH_CHECK(c);
        PF_TRY(launch_scan(c, ptr<int32_t>(c->sp_nch), n_gp, ptr<int64_t>(c->sp_off), nullptr));
        int64_t n_chunks = 0;
        PF_CUDA(c, cudaMemcpyAsync(&n_chunks, ptr<int64_t>(c->sp_off) + n_gp, sizeof(int64_t), cudaMemcpyDeviceToHost, c->stream));
        PF_CUDA(c, cudaStreamSynchronize(c->stream));
        PF_TRY(ensure(c, c->sp_items, sizeof(uint32_t) * SP_CHUNK * (size_t)std::max<int64_t>(n_chunks, 1)));
        if (n_chunks > 0) {
            k_sp_fill<<<wgrid, 256, 0, c->stream>>>(ptr<int64_t>(c->seg_off), ptr<uint32_t>(c->vals), ptr<int32_t>(c->rows), n_rows, P,
                                                    ptr<uint16_t>(c->mm_did), ptr<uint8_t>(c->mm_owned), ptr<int32_t>(c->sp_cnt), ptr<int32_t>(c->sp_tot),
                                                    ptr<int64_t>(c->sp_off), ptr<uint32_t>(c->sp_items));
            PF_LAUNCH_CHECK(c);
        }
        // dictionaries: bump allocation; when the first estimate is too small the builder still counts the need and we re-run
        const int64_t n_dict = (int64_t)ncb * P;
        PF_TRY(ensure(c, c->sp_nsub, sizeof(int32_t) * (size_t)n_dict));
        PF_TRY(ensure(c, c->sp_sub, sizeof(SpSub) * (size_t)n_dict * SP_MAXSUB));
        int *sp_err = ptr<int>(c->flags) + 14;
        unsigned long long *cursor = reinterpret_cast<unsigned long long *>(ptr<int>(c->flags) + 16);
        size_t cap = (size_t)n_dict * (SP_ENT_OFF + 64) + (size_t)(2.5 * e_sparse * (m.symmetric ? 1.0 : (double)n_cols / std::max(1, G))) + (1u << 20);
        auto kern = k_spdict_build;
        PF_CUDA(c, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, SD_SMEM));
        for (int attempt = 0; attempt < 2; attempt++) {
            PF_TRY(ensure(c, c->sp_blob, cap));
            PF_CUDA(c, cudaMemsetAsync(sp_err, 0, sizeof(int), c->stream));
            PF_CUDA(c, cudaMemsetAsync(cursor, 0, sizeof(unsigned long long), c->stream));
            SpDictParams q;
            q.seg_off = ptr<int64_t>(c->seg_off); q.vals = ptr<uint32_t>(c->vals); q.cols = ptr<int32_t>(c->cols); q.n_cols = n_cols; q.P = P;
            q.did = ptr<uint16_t>(c->mm_did); q.ent_max = pl.small_dict ? SP_ENT_MAX_SMALL : SP_ENT_MAX; q.blob = ptr<unsigned char>(c->sp_blob); q.blob_cap = (unsigned long long)c->sp_blob.cap;
            q.cursor = cursor; q.nsub = ptr<int32_t>(c->sp_nsub); q.sub = ptr<SpSub>(c->sp_sub); q.err = sp_err;
            kern<<<(unsigned)n_dict, SD_THREADS, SD_SMEM, c->stream>>>(q);
            PF_LAUNCH_CHECK(c);
            int eh2 = 0; unsigned long long need = 0;
            PF_CUDA(c, cudaMemcpyAsync(&eh2, sp_err, sizeof(int), cudaMemcpyDeviceToHost, c->stream));
            PF_CUDA(c, cudaMemcpyAsync(&need, cursor, sizeof(need), cudaMemcpyDeviceToHost, c->stream));
            PF_CUDA(c, cudaStreamSynchronize(c->stream));
            if (eh2 == 2) { set_error(c, "dense path: a column block's sparse dictionary does not fit a shared-memory stage at this threshold"); return PFAAI_ERR_CONSTRUCT; }
            if (eh2 == 0) break;
            if (attempt == 1) { set_error(c, "dense path: sparse dictionary buffer too small"); return PFAAI_ERR_CONSTRUCT; }
            cap = (size_t)need + (size_t)need / 16 + (1u << 20);
        }
        pl.has_sparse = true; pl.sparse_items = (int64_t)e_sparse;
    }
    prm.sp_items = ptr<uint32_t>(c->sp_items); prm.sp_off = ptr<int64_t>(c->sp_off); prm.sp_nsub = ptr<int32_t>(c->sp_nsub);
    prm.sp_sub = ptr<SpSub>(c->sp_sub); prm.sp_blob = ptr<unsigned char>(c->sp_blob); prm.has_sparse = pl.has_sparse ? 1 : 0;
    prm.img_rows = reinterpret_cast<unsigned char *>(c->mm_img_rows.p);
    prm.img_cols = same ? prm.img_rows : reinterpret_cast<unsigned char *>(c->mm_img_cols.p);
    prm.blk_stride = blk_stride;
    prm.kpad = ptr<int32_t>(c->mm_kpad); prm.koff = ptr<int64_t>(c->mm_koff);
    prm.colT = ptr<int32_t>(c->mm_colT); prm.rowTmax = ptr<int32_t>(c->mm_rowmax); prm.colTmax = ptr<int32_t>(c->mm_colmax);
    prm.ncb = ncb;
    return PFAAI_OK;
}

template <int NPL, int NPN>
static int launch_tiles(Ctx *c, const TileParams &prm, int n_tiles, int reserve_sms = 0, cudaStream_t stream = nullptr) {
    if (n_tiles <= 0) return PFAAI_OK;
    if (!stream) stream = c->stream;
    auto kern = aji_tile_kernel<NPL, NPN>;
    PF_CUDA(c, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_TOTAL));
    // persistent CTAs, one per SM; CTA b walks tiles b, b + grid, ... (row-block-major order keeps the
    // concurrently processed tiles on the same row slabs, which stay L2-resident)
    const int grid = std::min(n_tiles, std::max(1, c->sm_count - reserve_sms));
    kern<<<(unsigned)grid, TILE_THREADS, SMEM_TOTAL, stream>>>(prm);
    PF_LAUNCH_CHECK(c);
    return PFAAI_OK;
}

// direct_host: non-NULL = the caller's matrix is pinned host memory; the tile kernel stores straight into it (device alias
// direct_dev) and no device copy of the matrix exists afterwards
static int compute_internal(Ctx *c, int mode, const int32_t *qry_ids, int nQ, unsigned flags, bool want_sn, bool want_counts,
                            ModePlan &m, double *direct_dev = nullptr) {
    if (!c->built) { set_error(c, "compute: index not built"); return PFAAI_ERR_CONSTRUCT; }
    PF_CUDA(c, cudaSetDevice(c->device));
    PF_TRY(plan_mode(c, mode, qry_ids, nQ, flags, m));
    const int P = c->P;
    const int n_rows = (int)m.rows.size(), n_cols = (int)m.cols.size();
    const int n_groups = (n_rows + ROWS_PER_WARP - 1) / ROWS_PER_WARP;
    // bit-plane capacity: per SCP a lane performs <= max_len/2 + 16 lookups (two ranges, each padded to 8 per lane)
    const int lane_lookups = c->max_len / 2 + 16;
    if (lane_lookups > 4095) { set_error(c, "compute: a (SCP, genome) tetramer list exceeds 8158 entries"); return PFAAI_ERR_CONSTRUCT; }
    if (P > 1023) { set_error(c, "compute: more than 1023 SCPs"); return PFAAI_ERR_CONSTRUCT; }

    PF_TRY(upload_i32(c, c->rows, m.rows)); PF_TRY(upload_i32(c, c->row_t, m.row_t));
    PF_TRY(upload_i32(c, c->cols, m.cols)); PF_TRY(upload_i32(c, c->col_t, m.col_t));
    // ---- which tile path (include/pfaai_gpu.h, pfaai_gpu_set_path) ----
    const bool sharded = c->n_shards > 1;
    bool use_mma = false;
    MmaPlan pl;
    const auto plan_t0 = std::chrono::steady_clock::now();   // host clock: a planning pass that runs ends in a stream synchronise
    {
        const bool mma_possible = !c->partitioned && P <= 255 && c->nF > 0 && n_rows > 0 && n_cols > 0;
        if (c->path_pref == PFAAI_PATH_MMA && !mma_possible) {
            set_error(c, "compute: the dense path needs an unpartitioned build, a non-empty index and at most 255 SCPs");
            return PFAAI_ERR_CONSTRUCT;
        }
        if (c->path_pref != PFAAI_PATH_LISTS && mma_possible) {
            const bool force = c->path_pref == PFAAI_PATH_MMA;
            const double cells = (double)n_rows * (double)n_cols * (m.symmetric ? 0.5 : 1.0) / (double)c->n_shards;
            const int64_t img_blocks = (n_rows + MM_ROWS - 1) / MM_ROWS + (m.symmetric ? 0 : (n_cols + MM_COLS - 1) / MM_COLS);
            // planning on a genome sample first (~|F| / stride work); the exact lengths only when the dense path is taken
            PF_TRY(mma_length_hist(c, force ? 1 : std::max(1, c->G / 2048), pl));
            mma_choose(c, pl, cells, img_blocks, force);
            if (pl.use && (!pl.exact || c->mm_len_stride != 1)) {
                PF_TRY(mma_length_hist(c, 1, pl, /*need_device_lengths=*/true));   // k_dense_ids reads the exact lengths from mm_len
                mma_choose(c, pl, cells, img_blocks, force);
            }
            if (force && !pl.use) { set_error(c, "compute: the dense path does not fit this input (operand images too large)"); return PFAAI_ERR_CONSTRUCT; }
            use_mma = pl.use;
        }
    }
    c->tm.plan_ms = std::chrono::duration<float, std::milli>(std::chrono::steady_clock::now() - plan_t0).count();
    PF_CUDA(c, cudaEventRecord(c->ev[EV_P0], c->stream));    // pack_ms = row-slab packing only; the planning has its own timer
    const int TR = use_mma ? MM_ROWS : TILE_ROWS, TC = use_mma ? MM_COLS : COLS_PER_TILE;      // tile geometry of the path
    const int nrb = (n_rows + TR - 1) / TR, ncb = (n_cols + TC - 1) / TC;
    if (!use_mma) {
        // ---- K4a: pack the row slabs ----
        const int64_t ngp = (int64_t)n_groups * P, nslab = ngp * NRANGE;
        PF_TRY(ensure(c, c->ell_nch, sizeof(int32_t) * (size_t)nslab));
        PF_TRY(ensure(c, c->ell_off, sizeof(int64_t) * (size_t)(nslab + 1)));
        k_pack_count<<<(unsigned)ceil_div64(ngp, 256), 256, 0, c->stream>>>(ptr<int64_t>(c->seg_off), ptr<uint32_t>(c->vals),
                                                                            ptr<int32_t>(c->rows), n_rows, n_groups, P,
                                                                            ptr<int32_t>(c->ell_nch));
        PF_LAUNCH_CHECK(c);
        PF_TRY(launch_scan(c, ptr<int32_t>(c->ell_nch), nslab, ptr<int64_t>(c->ell_off), nullptr));
        int64_t total_chunks = 0;
        PF_CUDA(c, cudaMemcpyAsync(&total_chunks, ptr<int64_t>(c->ell_off) + nslab, sizeof(int64_t), cudaMemcpyDeviceToHost, c->stream));
        PF_CUDA(c, cudaStreamSynchronize(c->stream));
        PF_TRY(ensure(c, c->ell, sizeof(uint4) * 32 * (size_t)total_chunks + 4096));   // + slack: the walk prefetches one step past the end
        if (total_chunks > 0) {
            k_pack_fill<<<(unsigned)std::min<int64_t>(ceil_div64(ngp, 8), (int64_t)c->sm_count * 16), 256, 0, c->stream>>>(
                ptr<int64_t>(c->seg_off), ptr<uint32_t>(c->vals), ptr<int32_t>(c->rows), n_rows, n_groups, P, ptr<int64_t>(c->ell_off),
                ptr<uint4>(c->ell));
            PF_LAUNCH_CHECK(c);
        }
        PF_CUDA(c, cudaEventRecord(c->ev[EV_P1], c->stream));
        // ---- K4b: column-block dictionaries + stage lists (a partitioned context already holds them: built by the shards that
        //      own the column blocks and exchanged over NVLink before this call) ----
        if (c->partitioned) {
            if (mode != PFAAI_ALL_VS_ALL || c->n_shards < 2) { set_error(c, "compute: a partitioned build serves sharded all-vs-all only (use pfaai_gpu_build)"); return PFAAI_ERR_CONSTRUCT; }
            if (!c->dict_ready) { set_error(c, "compute: partitioned build not finished (exchange the buffers, then pfaai_gpu_finish_partitioned)"); return PFAAI_ERR_CONSTRUCT; }
            PF_CUDA(c, cudaEventRecord(c->ev[EV_K0], c->stream));
            PF_CUDA(c, cudaEventRecord(c->ev[EV_K1], c->stream));
        } else {
            PF_CUDA(c, cudaEventRecord(c->ev[EV_K0], c->stream));
            PF_TRY(make_dictionaries(c, n_cols, nullptr, ncb, 0, ncb, 0, 0, 0, 0));
            PF_CUDA(c, cudaEventRecord(c->ev[EV_K1], c->stream));
        }
    }
    // ---- tile list: row block x column block, I-major so that concurrently resident CTAs share the row
    //      block's slabs in L2; symmetric mode keeps only tiles that contain a cell with row < col.
    //      Sharding: row block I belongs to shard fold(I mod 2n) (folded assignment balances the
    //      triangular work); a sharded context stores only its own row blocks, compactly, in ascending order ----
    // compact: the shard's matrix holds only its own row blocks (upper-triangle cells in all-vs-all).  Not compact
    // (pfaai_gpu_set_global_matrix, all-vs-all): every shard addresses the SAME full G x G matrix and stores its cells and their
    // mirror images into it -- the shards' cell sets are disjoint and together cover the matrix
    const bool compact = sharded && !(c->global_out && m.symmetric);
    if (sharded && (want_sn || want_counts)) { set_error(c, "compute: the JAC / counts hooks need an unsharded context"); return PFAAI_ERR_CONSTRUCT; }
    std::vector<int4> tiles, tiles_late;   // partitioned: tiles on this shard's own dictionaries first (they need no exchange)
    c->owned_rows.clear();
    int local_rows = 0;
    // slot of column block J (where its stage range / T values live): J itself, or -- partitioned -- the position inside the
    // chunk of the shard that built it
    std::vector<int32_t> slot_of(ncb);
    for (int J = 0; J < ncb; J++) slot_of[J] = J;
    if (c->partitioned) {
        std::vector<int> nblk(c->n_shards, 0);
        for (int I = 0; I * RB_CBS < ncb; I++) {
            const int o = fold_owner(I, c->n_shards);
            for (int j = 0; j < RB_CBS && I * RB_CBS + j < ncb; j++)
                slot_of[I * RB_CBS + j] = (int32_t)(o * c->slot_cap + (int64_t)nblk[o] * RB_CBS + j);
            nblk[o]++;
        }
    }
    std::vector<uint8_t> rb_owned(nrb, 0);       // ownership of the path's row blocks (sharding granule: 384 rows, both paths)
    for (int I = 0; I < nrb; I++) {
        if (fold_owner((int)(((int64_t)I * TR) / TILE_ROWS), c->n_shards) != c->shard) continue;
        rb_owned[I] = 1;
        const int row0 = I * TR, nr = std::min(TR, n_rows - row0);
        for (int J = 0; J < ncb; J++) {
            if (m.symmetric) {
                const int col_max = std::min((J + 1) * TC, n_cols) - 1;
                if (row0 > col_max) continue;      // row0 == col_max: the tile still holds the diagonal cell (stored as 0)
            }
            const bool own_dict = !c->partitioned || fold_owner(J / RB_CBS, c->n_shards) == c->shard;
            (own_dict ? tiles : tiles_late).push_back(make_int4(I, J, compact ? local_rows : row0, use_mma ? 0 : slot_of[J]));
        }
        for (int r = 0; r < nr; r++) c->owned_rows.push_back(row0 + r);
        local_rows += nr;
    }
    if (use_mma) {
        // K5 streams column-side data (operand image + sparse dictionary of the column block) as well as row-side data per
        // (tile, SCP): order the tiles in 12 x 12 super-tiles, so that the ~148 tiles in flight share 12 row blocks and 12
        // column blocks and every stream is read from HBM once and from L2 eleven times
        constexpr int ST = 12;
        std::stable_sort(tiles.begin(), tiles.end(), [](const int4 &a, const int4 &b) {
            const int ka = (a.x / ST), kb = (b.x / ST);
            if (ka != kb) return ka < kb;
            const int ja = (a.y / ST), jb = (b.y / ST);
            if (ja != jb) return ja < jb;
            if (a.x != b.x) return a.x < b.x;
            return a.y < b.y;
        });
    }
    const int n_early = (int)tiles.size();
    tiles.insert(tiles.end(), tiles_late.begin(), tiles_late.end());
    PF_TRY(ensure(c, c->tiles, sizeof(int4) * tiles.size()));
    if (!tiles.empty()) PF_CUDA(c, cudaMemcpyAsync(c->tiles.p, tiles.data(), sizeof(int4) * tiles.size(), cudaMemcpyHostToDevice, c->stream));
    // ---- outputs: the kernel's row axis is the output's row axis in all-vs-all and its column axis otherwise ----
    c->shard_axis = m.symmetric ? 0 : 1;
    if (compact) {
        if (m.symmetric) m.out_rows = local_rows;                    // [owned rows x G], row stride G
        else { m.out_cols = local_rows; m.cs = local_rows; }         // [nQ x owned targets]
    }
    c->out_rows = m.out_rows; c->out_cols = m.out_cols;
    const size_t ncell = (size_t)m.out_rows * m.out_cols, nk = (size_t)n_rows * n_cols;
    c->out_on_device = (direct_dev == nullptr);
    if (!direct_dev) {
        PF_TRY(ensure(c, c->out, sizeof(double) * ncell));
        PF_CUDA(c, cudaMemsetAsync(c->out.p, 0, sizeof(double) * ncell, c->stream));
    }
    if (want_sn) {
        PF_TRY(ensure(c, c->S_out, sizeof(double) * nk)); PF_TRY(ensure(c, c->N_out, sizeof(int32_t) * nk));
        PF_CUDA(c, cudaMemsetAsync(c->S_out.p, 0, sizeof(double) * nk, c->stream));
        PF_CUDA(c, cudaMemsetAsync(c->N_out.p, 0, sizeof(int32_t) * nk, c->stream));
    }
    if (want_counts) {
        if ((double)nk * P > 1.0e9) { set_error(c, "compute: counts output is a small-N hook (n_rows*n_cols*P <= 1e9)"); return PFAAI_ERR_CONSTRUCT; }
        PF_TRY(ensure(c, c->counts_out, sizeof(int32_t) * nk * P));
        PF_CUDA(c, cudaMemsetAsync(c->counts_out.p, 0, sizeof(int32_t) * nk * P, c->stream));
    }
    // ---- streaming output: arm the per-granule progress counters (unsharded all-vs-all into a pinned host matrix only) ----
    { std::lock_guard<std::mutex> lk(c->prog_mutex); c->prog_state.store(0, std::memory_order_release); }
    int *prog_dev = nullptr;
    int n_gran = 0;
    if (c->prog_enabled && direct_dev && m.symmetric && !sharded && !tiles.empty() && c->nF > 0) {
        n_gran = (n_cols + TC - 1) / TC;
        const int gpr = TR / TC;
        std::vector<int32_t> need(n_gran, 0);
        for (const int4 &t : tiles) {
            for (int k = 0; k < gpr; k++) if (t.x * gpr + k < n_gran) need[t.x * gpr + k]++;
            need[t.y]++;                                            // the mirror image lands in the rows of column block J
        }
        PF_TRY(ensure(c, c->prog, sizeof(int32_t) * (size_t)n_gran));
        PF_CUDA(c, cudaMemsetAsync(c->prog.p, 0, sizeof(int32_t) * (size_t)n_gran, c->stream));
        if (!c->prog_stream) PF_CUDA(c, cudaStreamCreateWithFlags(&c->prog_stream, cudaStreamNonBlocking));
        if (c->prog_host_cap < (size_t)n_gran) {
            if (c->prog_host) cudaFreeHost(c->prog_host);
            c->prog_host = nullptr; c->prog_host_cap = 0;
            PF_CUDA(c, cudaMallocHost(reinterpret_cast<void **>(&c->prog_host), sizeof(int32_t) * (size_t)n_gran));
            c->prog_host_cap = (size_t)n_gran;
        }
        PF_CUDA(c, cudaStreamSynchronize(c->stream));               // counters are zero before anybody polls
        c->prog_need = std::move(need);
        c->prog_gran = TC; c->prog_rows = n_rows;
        prog_dev = ptr<int>(c->prog);
        c->prog_state.store(1, std::memory_order_release);
    }
    TileParams prm;
    prm.progress = prog_dev; prm.n_gran = n_gran;
    prm.ell = ptr<uint4>(c->ell); prm.ell_off = ptr<int64_t>(c->ell_off);
    prm.rows = ptr<int32_t>(c->rows); prm.row_t = ptr<int32_t>(c->row_t); prm.n_rows = n_rows; prm.n_groups = n_groups;
    prm.cols = ptr<int32_t>(c->cols); prm.n_cols = n_cols;
    prm.stages = ptr<StageDesc>(c->stages); prm.stage_rng = ptr<int64_t>(c->stage_rng);
    prm.blob = ptr<unsigned char>(c->blob); prm.colT = ptr<int32_t>(c->colT);
    prm.T = ptr<int32_t>(c->T); prm.t_ld = c->G; prm.P = P;
    prm.out = direct_dev ? direct_dev : ptr<double>(c->out); prm.out_rs = m.rs; prm.out_cs = m.cs; prm.symmetric = m.symmetric; prm.mirror = (m.symmetric && !compact) ? 1 : 0;
    prm.k27 = 1u << 27; prm.k16 = 1u << 16;
    {   // PFAAI_TILE_PIECE: bytes per bulk copy of a dictionary stage (A/B switch; default 0 = the whole blob in one copy.
        // Measured at config 4: 38.2 ms whole, 41.5 ms in 4 KB pieces, 49.5 ms in 1 KB pieces -- rejected, profiles/r2n_*)
        static const uint32_t piece = [] { const char *e = getenv("PFAAI_TILE_PIECE"); const long v = e ? atol(e) : 0; return (uint32_t)(v <= 0 ? (1 << 30) : ((v + 15) & ~15l)); }();
        prm.piece = piece;
    }
    prm.S_out = want_sn ? ptr<double>(c->S_out) : nullptr;
    prm.N_out = want_sn ? ptr<int32_t>(c->N_out) : nullptr;
    prm.counts_out = want_counts ? ptr<int32_t>(c->counts_out) : nullptr;
    PF_TRY(ensure(c, c->tile_ctr, 2 * sizeof(int)));
    PF_CUDA(c, cudaEventRecord(c->ev[EV_T0], c->stream));
    PF_CUDA(c, cudaMemsetAsync(c->tile_ctr.p, 0, 2 * sizeof(int), c->stream));
    // list 1 = tiles [0, n1), list 2 (optional) = tiles [n1, n1 + n2): a CTA moves on to list 2 when list 1 is exhausted
    if (direct_dev && (c->nF <= 0 || (tiles.empty() && (compact || !sharded))))
        // nothing will be launched: the caller's pinned matrix (no staging buffer, no memset) would stay untouched.  (A shard of a
        // shared global matrix that owns no row block must leave the other shards' cells alone.)
        PF_CUDA(c, cudaMemsetAsync(direct_dev, 0, sizeof(double) * ncell, c->stream));
    auto launch_lists = [&](int n1, int n2, int reserve, cudaStream_t st) -> int {
        if (n1 + n2 <= 0 || c->nF <= 0) return PFAAI_OK;
        prm.tiles = ptr<int4>(c->tiles); prm.n_tiles = n1; prm.tile_counter = ptr<int>(c->tile_ctr);
        prm.tiles2 = ptr<int4>(c->tiles) + n1; prm.n_tiles2 = n2; prm.tile_counter2 = ptr<int>(c->tile_ctr) + 1;
        const int count = n1 + n2;
        if (lane_lookups <= 511) return (P <= 127) ? launch_tiles<9, 7>(c, prm, count, reserve, st) : launch_tiles<9, 10>(c, prm, count, reserve, st);
        return (P <= 127) ? launch_tiles<12, 7>(c, prm, count, reserve, st) : launch_tiles<12, 10>(c, prm, count, reserve, st);
    };
    const int n_late = (int)tiles.size() - n_early;
    if (use_mma) {
        // ---- K5: operands + sparse row items / dictionaries (timed as "prep"), then the tensor-core tile kernel ----
        MmaParams mp;
        std::vector<int32_t> kpad_host;
        PF_CUDA(c, cudaEventRecord(c->ev[EV_K0], c->stream));
        {
            // (PFAAI_TEST_FAIL_MMA_PREP: test hook for the fall-back below)
#ifdef PFAAI_TEST_HOOKS   // only in libpfaai_gpu_hooks.so, which the tests build next to the product library
            const int prc = getenv("PFAAI_TEST_FAIL_MMA_PREP") ? (set_error(c, "dense path: preparation failed (test hook)"), (int)PFAAI_ERR_CONSTRUCT)
                                                               : mma_prepare(c, m, pl, n_rows, n_cols, rb_owned, kpad_host, mp);
#else
            const int prc = mma_prepare(c, m, pl, n_rows, n_cols, rb_owned, kpad_host, mp);
#endif
            if (prc != PFAAI_OK) {
                // the cost model picked the dense path but it cannot be prepared for this input (a sparse dictionary that does not
                // fit a stage, no memory for the operand images, ...): when the choice was ours, serve the call on the list path
                // instead of failing it; a caller who forced PFAAI_PATH_MMA gets the error
                if (c->path_pref != PFAAI_PATH_AUTO) return prc;
                { std::lock_guard<std::mutex> lk(c->prog_mutex); c->prog_state.store(0, std::memory_order_release); }
                c->path_pref = PFAAI_PATH_LISTS;
                ModePlan m2;
                const int rc2 = compute_internal(c, mode, qry_ids, nQ, flags, want_sn, want_counts, m2, direct_dev);
                c->path_pref = PFAAI_PATH_AUTO;
                m = std::move(m2);
                return rc2;
            }
        }
        PF_CUDA(c, cudaEventRecord(c->ev[EV_K1], c->stream));
        mp.rows = prm.rows; mp.row_t = prm.row_t; mp.cols = prm.cols; mp.col_t = ptr<int32_t>(c->col_t);
        mp.n_rows = n_rows; mp.n_cols = n_cols; mp.T = prm.T; mp.t_ld = prm.t_ld; mp.P = P;
        mp.quirk = (mode == PFAAI_QUERY_DB && (flags & PFAAI_FLAG_REF_COMPAT_QT)) ? 1 : 0;
        mp.out = prm.out; mp.out_rs = prm.out_rs; mp.out_cs = prm.out_cs; mp.symmetric = prm.symmetric; mp.mirror = prm.mirror;
        mp.S_out = prm.S_out; mp.N_out = prm.N_out; mp.counts_out = prm.counts_out;
        mp.tiles = ptr<int4>(c->tiles); mp.n_tiles = (int)tiles.size();
        mp.progress = prog_dev;
        PF_CUDA(c, cudaEventRecord(c->ev[EV_T0], c->stream));
        if (!tiles.empty()) {
            const unsigned grid = (unsigned)std::min<int>((int)tiles.size(), c->sm_count);
            if (pl.small_dict) {
                auto kern = aji_mma_kernel<MmCfgDense>;
                PF_CUDA(c, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, mm_smem_total<MmCfgDense>()));
                kern<<<grid, MM_THREADS, mm_smem_total<MmCfgDense>(), c->stream>>>(mp);
            } else {
                auto kern = aji_mma_kernel<MmCfgSparse>;
                PF_CUDA(c, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, mm_smem_total<MmCfgSparse>()));
                kern<<<grid, MM_THREADS, mm_smem_total<MmCfgSparse>(), c->stream>>>(mp);
            }
            PF_LAUNCH_CHECK(c);
        }
    } else if (c->partitioned && c->exch_event && n_late > 0) {
        // The other shards' dictionaries may still be in flight (exchange on the caller's stream).  Instance 1 of the kernel
        // starts now and takes only the tiles on our own column blocks; it leaves some SMs free, because a tile CTA fills its
        // SM (227 KB of shared memory) and would otherwise lock the exchange's collective kernels out.  Instance 2, on a second
        // stream behind the caller's event, helps with whatever is left of those tiles and then runs the rest: both draw from
        // the same counters, so no SM idles at a kernel boundary.
        static const int reserve = [] { const char *e = getenv("PFAAI_EXCHANGE_SMS"); return e ? std::max(0, atoi(e)) : 40; }();
        PF_CUDA(c, cudaEventRecord(c->ev[EV_FORK], c->stream));
        PF_TRY(launch_lists(n_early, 0, reserve, c->stream));
        PF_CUDA(c, cudaStreamWaitEvent(c->stream2, c->ev[EV_FORK], 0));
        PF_CUDA(c, cudaStreamWaitEvent(c->stream2, reinterpret_cast<cudaEvent_t>(c->exch_event), 0));
        PF_TRY(launch_lists(n_early, n_late, 0, c->stream2));
        PF_CUDA(c, cudaEventRecord(c->ev[EV_JOIN], c->stream2));
        PF_CUDA(c, cudaStreamWaitEvent(c->stream, c->ev[EV_JOIN], 0));
    } else {
        if (c->partitioned && c->exch_event)
            PF_CUDA(c, cudaStreamWaitEvent(c->stream, reinterpret_cast<cudaEvent_t>(c->exch_event), 0));
        PF_TRY(launch_lists((int)tiles.size(), 0, 0, c->stream));
    }
    c->exch_event = nullptr;   // consumed: later computes on this build find the dictionaries in place
    PF_CUDA(c, cudaEventRecord(c->ev[EV_T1], c->stream));
    PF_CUDA(c, cudaStreamSynchronize(c->stream));
    if (prog_dev) c->prog_state.store(2, std::memory_order_release);
    c->tm.pack_ms = use_mma ? 0.f : ev_ms(c, EV_P0, EV_P1);
    if (!c->partitioned) c->tm.dict_ms = use_mma ? 0.f : ev_ms(c, EV_K0, EV_K1);
    c->tm.prep_ms = use_mma ? ev_ms(c, EV_K0, EV_K1) : 0.f;
    c->tm.path = use_mma ? PFAAI_PATH_MMA : PFAAI_PATH_LISTS;
    c->tm.dense_tau = use_mma ? (int32_t)pl.tau : 0;
    c->tm.dense_k_sum = use_mma ? pl.k_sum : 0;
    c->tm.sparse_items = use_mma ? pl.sparse_items : 0;
    c->tm.e_size = c->e_size;
    c->tm.tile_ms = ev_ms(c, EV_T0, EV_T1);
    c->tm.n_tiles = (int32_t)tiles.size();
    c->tm.n_launches = c->launches;   // build + this compute
    // algorithmic bytes of the tile kernel: postings once, segment offsets, T, output once
    c->tm.tile_alg_bytes = 4 * c->nF + 8 * ((int64_t)P * c->G + 1) + 4 * (int64_t)P * c->G +
                           8 * (int64_t)ncell;
    c->last_mode = mode;
    return PFAAI_OK;
}

}  // namespace pfaai

extern "C" {

int pfaai_gpu_compute_device(pfaai_gpu_ctx *h, int mode, const int32_t *qry_ids, int n_qry, unsigned flags) {
    Ctx *c = reinterpret_cast<Ctx *>(h);
    if (!c) return PFAAI_ERR_CONSTRUCT;
    ModePlan m;
    return compute_internal(c, mode, qry_ids, n_qry, flags, false, false, m);
}

int pfaai_gpu_matrix_shape(const pfaai_gpu_ctx *h, int64_t *rows, int64_t *cols) {
    const Ctx *c = reinterpret_cast<const Ctx *>(h);
    if (!c) return PFAAI_ERR_CONSTRUCT;
    if (rows) *rows = c->out_rows;
    if (cols) *cols = c->out_cols;
    return PFAAI_OK;
}

int pfaai_gpu_shard_layout(const pfaai_gpu_ctx *h, int *axis, int64_t *n, int32_t *ids) {
    const Ctx *c = reinterpret_cast<const Ctx *>(h);
    if (!c || c->last_mode < 0) return PFAAI_ERR_CONSTRUCT;
    if (axis) *axis = c->shard_axis;
    if (n) *n = (int64_t)c->owned_rows.size();
    if (ids) std::copy(c->owned_rows.begin(), c->owned_rows.end(), ids);
    return PFAAI_OK;
}

const double *pfaai_gpu_matrix_dev(const pfaai_gpu_ctx *h) {
    const Ctx *c = reinterpret_cast<const Ctx *>(h);
    return (c && c->out_on_device) ? reinterpret_cast<const double *>(c->out.p) : nullptr;
}

int pfaai_gpu_download_matrix(pfaai_gpu_ctx *h, double *M) {
    Ctx *c = reinterpret_cast<Ctx *>(h);
    if (!c || !M || c->last_mode < 0) { set_error(c, "download_matrix: nothing computed"); return PFAAI_ERR_CONSTRUCT; }
    if (!c->out_on_device) { set_error(c, "download_matrix: the last compute stored the matrix directly into pinned host memory"); return PFAAI_ERR_CONSTRUCT; }
    PF_CUDA(c, cudaSetDevice(c->device));
    PF_CUDA(c, cudaEventRecord(c->ev[EV_D0], c->stream));
    PF_CUDA(c, cudaMemcpyAsync(M, c->out.p, sizeof(double) * (size_t)c->out_rows * c->out_cols, cudaMemcpyDeviceToHost, c->stream));
    PF_CUDA(c, cudaEventRecord(c->ev[EV_D1], c->stream));
    PF_CUDA(c, cudaStreamSynchronize(c->stream));
    c->tm.d2h_ms = ev_ms(c, EV_D0, EV_D1);
    return PFAAI_OK;
}

int pfaai_gpu_compute(pfaai_gpu_ctx *h, int mode, const int32_t *qry_ids, int n_qry, unsigned flags, double *aji_matrix,
                      pfaai_jac_t *jac, double *aji, int32_t *counts) {
    Ctx *c = reinterpret_cast<Ctx *>(h);
    if (!c) return PFAAI_ERR_CONSTRUCT;
    ModePlan m;
    const bool want_sn = jac || aji;
    // a pinned (page-locked, mapped) host matrix is written by the tile kernel itself: the D2H transfer is the kernel's
    // store stream and overlaps the computation; pageable memory goes through a device buffer + one copy
    double *direct_dev = nullptr;
    // the pointer query needs the context's device current on THIS thread (a fresh host thread -- one per context in the C++
    // driver -- has none yet, and the query then does not recognise page-locked memory: the matrix would silently take the
    // staged path and, for a shared global matrix, overwrite the other shards' cells)
    PF_CUDA(c, cudaSetDevice(c->device));
    if (aji_matrix) {
        cudaPointerAttributes attr;
        if (cudaPointerGetAttributes(&attr, aji_matrix) == cudaSuccess && attr.type == cudaMemoryTypeHost && attr.devicePointer)
            direct_dev = reinterpret_cast<double *>(attr.devicePointer);
        else
            cudaGetLastError();
    }
    PF_TRY(compute_internal(c, mode, qry_ids, n_qry, flags, want_sn, counts != nullptr, m, direct_dev));
    c->tm.d2h_ms = 0.f;
    if (aji_matrix && c->out_on_device) PF_TRY(pfaai_gpu_download_matrix(h, aji_matrix));
    if (!want_sn && !counts) return PFAAI_OK;
    // ---- small-N outputs in the reference's JAC order (initJAC of the mode) ----
    const int G = c->G, P = c->P;
    const int n_rows = (int)m.rows.size(), n_cols = (int)m.cols.size();
    const size_t nk = (size_t)n_rows * n_cols;
    std::vector<double> S; std::vector<int32_t> N, C;
    if (want_sn) {
        S.resize(nk); N.resize(nk);
        PF_CUDA(c, cudaMemcpyAsync(S.data(), c->S_out.p, sizeof(double) * nk, cudaMemcpyDeviceToHost, c->stream));
        PF_CUDA(c, cudaMemcpyAsync(N.data(), c->N_out.p, sizeof(int32_t) * nk, cudaMemcpyDeviceToHost, c->stream));
    }
    if (counts) {
        C.resize(nk * P);
        PF_CUDA(c, cudaMemcpyAsync(C.data(), c->counts_out.p, sizeof(int32_t) * nk * P, cudaMemcpyDeviceToHost, c->stream));
    }
    PF_CUDA(c, cudaStreamSynchronize(c->stream));
    int64_t k = 0;
    auto emit = [&](int labA, int labB, size_t cell) {
        if (jac) { jac[k].gA = labA; jac[k].gB = labB; jac[k].S = S[cell]; jac[k].N = N[cell]; jac[k].pad_ = 0; }
        if (aji) aji[k] = (N[cell] > 0) ? S[cell] / (double)N[cell] : 0.0;   // computeAJI, algorithm_impl.hpp:318
        if (counts) std::memcpy(counts + k * P, C.data() + cell * P, sizeof(int32_t) * P);
        k++;
    };
    if (mode == PFAAI_ALL_VS_ALL) {                       // ds_impl.hpp:99-114
        for (int a = 0; a < G; a++) for (int b = a + 1; b < G; b++) emit(a, b, (size_t)a * n_cols + b);
    } else if (mode == PFAAI_QUERY_SUBSET) {              // ds_impl.hpp:278-305; kernel rows = genomes, cols = queries
        std::vector<char> isq(G, 0);
        for (int q = 0; q < n_qry; q++) isq[qry_ids[q]] = 1;
        for (int q = 0; q < n_qry; q++) for (int g = 0; g < G; g++) if (!isq[g]) emit(qry_ids[q], g, (size_t)g * n_cols + q);
        for (int a = 0; a < n_qry; a++) for (int b = a + 1; b < n_qry; b++) emit(qry_ids[a], qry_ids[b], (size_t)qry_ids[b] * n_cols + a);
    } else {                                              // ds_impl.hpp:426-438; kernel rows = targets, cols = queries
        const int nT = G - n_qry;
        const bool quirk = (flags & PFAAI_FLAG_REF_COMPAT_QT) != 0;
        for (int q = 0; q < n_qry; q++) for (int t = 0; t < nT; t++)
            emit(quirk ? q : nT + q, quirk ? n_qry + t : t, (size_t)t * n_cols + q);
    }
    return PFAAI_OK;
}

int pfaai_gpu_set_progress(pfaai_gpu_ctx *h, int enable) {
    Ctx *c = reinterpret_cast<Ctx *>(h);
    if (!c) return PFAAI_ERR_CONSTRUCT;
    c->prog_enabled = enable != 0;
    { std::lock_guard<std::mutex> lk(c->prog_mutex); c->prog_state.store(0, std::memory_order_release); }
    return PFAAI_OK;
}

// May be called from another host thread while pfaai_gpu_compute blocks its caller: touches only the context's private polling
// stream, its pinned counter copy and data published (release / acquire) before the tile kernel was launched.
int pfaai_gpu_progress_poll(pfaai_gpu_ctx *h, int64_t *rows_final, int *finished) {
    Ctx *c = reinterpret_cast<Ctx *>(h);
    if (!c || !rows_final) return PFAAI_ERR_CONSTRUCT;
    const int st = c->prog_state.load(std::memory_order_acquire);
    if (finished) *finished = (st == 2) ? 1 : 0;
    if (st == 0) { *rows_final = 0; return PFAAI_OK; }
    if (st == 2) { *rows_final = c->prog_rows; return PFAAI_OK; }
    if (cudaSetDevice(c->device) != cudaSuccess) return PFAAI_ERR_CONSTRUCT;
    // The computing thread may leave state 1 at any moment (a dense-path preparation that falls back to the list path re-arms
    // with new counters; the next compute may free prog_host): it does so under prog_mutex, so re-check the state with the lock
    // held and keep it until the last read of the shared arrays.
    std::lock_guard<std::mutex> lk(c->prog_mutex);
    const int st2 = c->prog_state.load(std::memory_order_acquire);
    if (st2 != 1) {
        if (finished) *finished = (st2 == 2) ? 1 : 0;
        *rows_final = (st2 == 2) ? c->prog_rows : 0;
        return PFAAI_OK;
    }
    const int n = (int)c->prog_need.size();
    if (cudaMemcpyAsync(c->prog_host, c->prog.p, sizeof(int32_t) * (size_t)n, cudaMemcpyDeviceToHost, c->prog_stream) != cudaSuccess ||
        cudaStreamSynchronize(c->prog_stream) != cudaSuccess)
        return PFAAI_ERR_CONSTRUCT;
    int g = 0;
    while (g < n && c->prog_host[g] >= c->prog_need[g]) g++;
    *rows_final = std::min<int64_t>(c->prog_rows, (int64_t)g * c->prog_gran);
    return PFAAI_OK;
}

int pfaai_gpu_e_size(pfaai_gpu_ctx *h, int64_t *e_size) {
    Ctx *c = reinterpret_cast<Ctx *>(h);
    if (!c || !e_size || !c->built) { set_error(c, "e_size: index not built"); return PFAAI_ERR_CONSTRUCT; }
    if (c->partitioned) { set_error(c, "e_size: needs an unpartitioned build (every genome's lists on this GPU)"); return PFAAI_ERR_CONSTRUCT; }
    PF_CUDA(c, cudaSetDevice(c->device));
    if (c->e_size < 0) { MmaPlan pl; PF_TRY(mma_length_hist(c, 1, pl)); }
    *e_size = c->e_size;
    return PFAAI_OK;
}

int pfaai_gpu_microbench(pfaai_gpu_ctx *h, int variant, double hit_rate, int steps, pfaai_microbench_t *out) {
    Ctx *c = reinterpret_cast<Ctx *>(h);
    if (!c || !out || steps <= 0 || hit_rate < 0.0 || hit_rate > 1.0 || variant < 0 || variant > 1) { set_error(c, "microbench: bad arguments"); return PFAAI_ERR_CONSTRUCT; }
    PF_CUDA(c, cudaSetDevice(c->device));
    const int64_t n_chunks = 1 << 13;                              // 8192 chunks x 512 B = 4 MB, L2-resident
    DevBuf ids, sink;
    PF_TRY(ensure(c, ids, sizeof(uint4) * 32 * (size_t)n_chunks));
    PF_TRY(ensure(c, sink, 16));
    k_ceiling_ids<<<(unsigned)ceil_div64(n_chunks * 32, 256), 256, 0, c->stream>>>(ptr<uint4>(ids), n_chunks * 32, variant);
    auto kern = k_walk_ceiling<9>;
    PF_CUDA(c, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_TOTAL));
    const uint32_t hit_q16 = (uint32_t)(hit_rate * 65536.0);
    for (int rep = 0; rep < 2; rep++) {                            // rep 0 = warm-up
        PF_CUDA(c, cudaEventRecord(c->ev[EV_T0], c->stream));
        kern<<<(unsigned)c->sm_count, TILE_THREADS, SMEM_TOTAL, c->stream>>>(ptr<uint4>(ids), (uint32_t)(n_chunks - 1), steps, hit_q16, ptr<uint32_t>(sink));
        PF_CUDA(c, cudaGetLastError());
        PF_CUDA(c, cudaEventRecord(c->ev[EV_T1], c->stream));
        PF_CUDA(c, cudaStreamSynchronize(c->stream));
    }
    out->ms = ev_ms(c, EV_T0, EV_T1);
    out->lookups = (int64_t)c->sm_count * TILE_THREADS * (int64_t)steps * 8;
    out->lookups_per_s = out->ms > 0.f ? (double)out->lookups / ((double)out->ms * 1e-3) : 0.0;
    out->warps_per_sm = TILE_WARPS; out->variant = variant;
    release(ids); release(sink);
    return PFAAI_OK;
}

int pfaai_gpu_get_timings(const pfaai_gpu_ctx *h, pfaai_timings_t *out) {
    const Ctx *c = reinterpret_cast<const Ctx *>(h);
    if (!c || !out) return PFAAI_ERR_CONSTRUCT;
    *out = c->tm;
    return PFAAI_OK;
}

}  // extern "C"

// ------------------------------------------------------------------------------------------------
// pfaai_gpu_prefix_sum_static -- the exclusive scan of launch_scan() for callers that must not allocate.
//
// launch_scan() keeps its per-tile partial sums in a context buffer obtained with cudaMalloc.  Some users of this library
// cannot afford that call: a plug-in that lives inside a host application whose own pool has already reserved the GPU's
// memory (a raw cudaMalloc then fails, or steals from the pool's budget), and code on a latency-critical path, where
// cudaMalloc / cudaFree synchronise the whole device.  For them the partial sums live in the array below, which is part of
// the library image: every GPU on which this library is used holds these 8 KiB from module load until the process exits,
// and the call itself allocates nothing -- no cudaMalloc / cudaFree, no context, no events, no library routine that could.
//
// g_static_scan_sums is a __device__ variable: inside host code its name is only the host-side handle of the symbol, not
// an address a kernel can dereference.  The device address is looked up once per device, with that device current, with
// cudaGetSymbolAddress and cached in g_static_scan_ptr; the kernels are then given that pointer like any other buffer.
// ------------------------------------------------------------------------------------------------
namespace pfaai {
constexpr int64_t STATIC_SCAN_TILES = PFAAI_GPU_STATIC_SCAN_MAX / SCAN_TILE;      // 1024 partial sums = 8 KiB
static_assert(PFAAI_GPU_STATIC_SCAN_MAX % SCAN_TILE == 0, "the limit is a whole number of scan tiles");
static __device__ __align__(256) int64_t g_static_scan_sums[STATIC_SCAN_TILES];

constexpr int STATIC_SCAN_MAX_DEVICES = 64;
static int64_t *g_static_scan_ptr[STATIC_SCAN_MAX_DEVICES];                        // device address per ordinal, 0 = not looked up yet
static std::mutex g_static_scan_mutex;                                             // guards the table and the error text

static int static_scan_fail(const std::string &msg, int restore_device) {
    set_error(nullptr, "prefix_sum_static: " + msg);
    if (restore_device >= 0) cudaSetDevice(restore_device);
    return PFAAI_ERR_CONSTRUCT;
}
}  // namespace pfaai

extern "C" int pfaai_gpu_prefix_sum_static(int device, void *cuda_stream, const int32_t *in_dev, int64_t n, int64_t *out_dev,
                                           int32_t *out32_dev) {
    using namespace pfaai;
    std::lock_guard<std::mutex> lock(g_static_scan_mutex);
    // argument checks first: they need no GPU
    if (n < 0 || !out_dev || (n > 0 && !in_dev)) return static_scan_fail("bad arguments (n >= 0, device pointers for input and output)", -1);
    if (n > PFAAI_GPU_STATIC_SCAN_MAX)
        return static_scan_fail("n = " + std::to_string(n) + " exceeds PFAAI_GPU_STATIC_SCAN_MAX = " + std::to_string((long long)PFAAI_GPU_STATIC_SCAN_MAX) +
                                " (the fixed scratch array holds one partial sum per " + std::to_string(SCAN_TILE) +
                                " inputs; larger inputs need the allocating path, pfaai_gpu_build)", -1);
    if (device < 0 || device >= STATIC_SCAN_MAX_DEVICES) return static_scan_fail("device ordinal outside [0, 64)", -1);
    int prev = -1;
    if (cudaGetDevice(&prev) != cudaSuccess) { cudaGetLastError(); return static_scan_fail("no CUDA device (this library has no CPU fallback)", -1); }
    cudaError_t e = cudaSetDevice(device);
    if (e != cudaSuccess) { cudaGetLastError(); return static_scan_fail(std::string("cudaSetDevice: ") + cudaGetErrorString(e), prev); }
    if (!g_static_scan_ptr[device]) {
        // set-up, once per device, with that device current (the FIRST call on a device: make it before any stream capture)
        int64_t *p = nullptr;
        e = cudaGetSymbolAddress((void **)&p, g_static_scan_sums);
        if (e != cudaSuccess || !p) { cudaGetLastError(); return static_scan_fail(std::string("cudaGetSymbolAddress: ") + cudaGetErrorString(e), prev); }
        g_static_scan_ptr[device] = p;
    }
    int64_t *sums = g_static_scan_ptr[device];
    cudaStream_t stream = reinterpret_cast<cudaStream_t>(cuda_stream);
    if (n == 0) {
        e = cudaMemsetAsync(out_dev, 0, sizeof(int64_t), stream);                  // out[0] = grand total = 0
    } else {
        // Nothing in the array has to start at zero: k_scan_tile_sums overwrites sums[0, ntiles) before anything reads them, and
        // entries beyond ntiles (left by an earlier, larger call) are never touched by this one.
        const int64_t ntiles = ceil_div64(n, SCAN_TILE);
        k_scan_tile_sums<<<(unsigned)ntiles, SCAN_THREADS, 0, stream>>>(in_dev, n, sums);
        k_scan_small_i64<<<1, 1024, 0, stream>>>(sums, ntiles);
        k_scan_final<<<(unsigned)ntiles, SCAN_THREADS, 0, stream>>>(in_dev, n, sums, out_dev, out32_dev);
        e = cudaGetLastError();
    }
    if (e != cudaSuccess) return static_scan_fail(std::string("launch: ") + cudaGetErrorString(e), prev);
    if (prev != device) cudaSetDevice(prev);
    return PFAAI_OK;
}
