"""ctypes binding of libpfaai_gpu.so (the C-ABI declared in include/pfaai_gpu.h).

The library is built in-tree by ``__graft_entry__.build()`` (nvcc, sm_100a).  There is no CPU
fallback anywhere in this package: if the shared library is missing or no CUDA device is present,
the calls below raise.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("PFAAI_GPU_LIB") or os.path.join(HERE, "libpfaai_gpu.so")   # override: kernel A/B experiments

PG_DTYPE = np.dtype([("protein", "<i4"), ("genome", "<i4")])
JAC_DTYPE = np.dtype([("gA", "<i4"), ("gB", "<i4"), ("S", "<f8"), ("N", "<i4"), ("pad_", "<i4")])
assert JAC_DTYPE.itemsize == 24

ALL_VS_ALL, QUERY_SUBSET, QUERY_DB = 0, 1, 2
PATH_AUTO, PATH_LISTS, PATH_MMA = 0, 1, 2
FLAG_REF_COMPAT_QT = 1
NTETRA = 160000

EXPORTS = [
    "pfaai_gpu_create", "pfaai_gpu_destroy", "pfaai_gpu_last_error", "pfaai_gpu_set_stream", "pfaai_gpu_set_shard",
    "pfaai_gpu_upload_index", "pfaai_gpu_build", "pfaai_gpu_get_arrays", "pfaai_gpu_compute_device",
    "pfaai_gpu_matrix_shape", "pfaai_gpu_matrix_dev", "pfaai_gpu_download_matrix", "pfaai_gpu_compute",
    "pfaai_gpu_n_pairs", "pfaai_gpu_get_timings", "pfaai_gpu_shard_layout", "pfaai_gpu_index_device_buffers",
    "pfaai_gpu_build_partitioned", "pfaai_gpu_exchange_buffers", "pfaai_gpu_finish_partitioned", "pfaai_gpu_exchange_local",
    "pfaai_gpu_set_partition_chunk", "pfaai_gpu_finish_partitioned_async",
    "pfaai_gpu_upload_index_slice", "pfaai_gpu_index_slice_buffers", "pfaai_gpu_route_prepare", "pfaai_gpu_route_recv_buffer",
    "pfaai_gpu_build_routed", "pfaai_gpu_route_local", "pfaai_gpu_set_global_matrix", "pfaai_gpu_microbench", "pfaai_gpu_set_path", "pfaai_gpu_e_size",
    "pfaai_gpu_host_register", "pfaai_gpu_host_unregister", "pfaai_gpu_upload_triples", "pfaai_gpu_import_lists", "pfaai_gpu_set_progress", "pfaai_gpu_progress_poll",
    "pfaai_gpu_prefix_sum_static",
]


class Timings(C.Structure):
    _fields_ = [("h2d_ms", C.c_float), ("build_ms", C.c_float), ("sort_ms", C.c_float), ("pack_ms", C.c_float),
                ("tile_ms", C.c_float), ("d2h_ms", C.c_float), ("n_launches", C.c_int32), ("n_tiles", C.c_int32),
                ("tile_alg_bytes", C.c_int64), ("dict_ms", C.c_float), ("reserved_", C.c_int32),
                ("path", C.c_int32), ("dense_tau", C.c_int32), ("dense_k_sum", C.c_int64), ("sparse_items", C.c_int64),
                ("e_size", C.c_int64), ("prep_ms", C.c_float), ("plan_ms", C.c_float)]

    def as_dict(self):
        return {k: getattr(self, k) for k, _ in self._fields_}


class Microbench(C.Structure):
    _fields_ = [("lookups_per_s", C.c_double), ("lookups", C.c_int64), ("ms", C.c_float), ("warps_per_sm", C.c_int32),
                ("variant", C.c_int32), ("reserved_", C.c_int32)]


class Exchange(C.Structure):
    """pfaai_exchange_t: the four buffers a partitioned build leaves to be all-gathered across the shards."""
    _fields_ = [("buf_dev", C.c_void_p * 4), ("chunk_bytes", C.c_int64 * 4), ("used_bytes", C.c_int64 * 4),
                ("shard", C.c_int32), ("n_shards", C.c_int32)]


class Route(C.Structure):
    """pfaai_route_t: a slice's sort items grouped by destination shard."""
    _fields_ = [("send_dev", C.c_void_p), ("send_offset", C.c_int64 * 17), ("shard", C.c_int32), ("n_shards", C.c_int32)]


_lib = None


def load():
    """dlopen the C-ABI library; raises (never falls back) when it has not been built."""
    global _lib
    if _lib is None:
        _lib = load_path(LIB_PATH)
    return _lib


def load_path(path):
    """dlopen one build of the C-ABI library and declare its argument types (tests: the -DPFAAI_TEST_HOOKS build)."""
    if not os.path.exists(path):
        raise RuntimeError(f"{path} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                           "(nvcc, sm_100a). There is no CPU fallback.")
    L = C.CDLL(path)
    vp, i32, i64, u32 = C.c_void_p, C.c_int, C.c_int64, C.c_uint
    L.pfaai_gpu_create.argtypes = [C.POINTER(vp), i32, u32]
    L.pfaai_gpu_destroy.argtypes = [vp]
    L.pfaai_gpu_destroy.restype = None
    L.pfaai_gpu_last_error.argtypes = [vp]
    L.pfaai_gpu_last_error.restype = C.c_char_p
    L.pfaai_gpu_set_stream.argtypes = [vp, vp]
    L.pfaai_gpu_set_shard.argtypes = [vp, i32, i32]
    L.pfaai_gpu_upload_index.argtypes = [vp, i32, i32, vp, vp, i64, vp]
    L.pfaai_gpu_upload_triples.argtypes = [vp, i32, i32, vp, vp, i64, vp]
    L.pfaai_gpu_import_lists.argtypes = [vp, i32, i32, i64, vp, vp, vp, vp]
    L.pfaai_gpu_build.argtypes = [vp]
    L.pfaai_gpu_get_arrays.argtypes = [vp, vp, vp, vp, vp, vp]
    L.pfaai_gpu_compute_device.argtypes = [vp, i32, vp, i32, u32]
    L.pfaai_gpu_matrix_shape.argtypes = [vp, C.POINTER(i64), C.POINTER(i64)]
    L.pfaai_gpu_matrix_dev.argtypes = [vp]
    L.pfaai_gpu_matrix_dev.restype = vp
    L.pfaai_gpu_download_matrix.argtypes = [vp, vp]
    L.pfaai_gpu_compute.argtypes = [vp, i32, vp, i32, u32, vp, vp, vp, vp]
    L.pfaai_gpu_n_pairs.argtypes = [i32, i32, i32]
    L.pfaai_gpu_n_pairs.restype = i64
    L.pfaai_gpu_get_timings.argtypes = [vp, C.POINTER(Timings)]
    L.pfaai_gpu_shard_layout.argtypes = [vp, C.POINTER(C.c_int), C.POINTER(i64), vp]
    L.pfaai_gpu_index_device_buffers.argtypes = [vp, i32, i32, i64, i32, C.POINTER(vp), C.POINTER(vp), C.POINTER(vp)]
    L.pfaai_gpu_build_partitioned.argtypes = [vp]
    L.pfaai_gpu_exchange_buffers.argtypes = [vp, C.POINTER(Exchange)]
    L.pfaai_gpu_finish_partitioned.argtypes = [vp]
    L.pfaai_gpu_exchange_local.argtypes = [C.POINTER(vp), i32]
    L.pfaai_gpu_set_partition_chunk.argtypes = [vp, i64]
    L.pfaai_gpu_finish_partitioned_async.argtypes = [vp, vp]
    L.pfaai_gpu_upload_index_slice.argtypes = [vp, i32, i32, vp, vp, i64, i64, i64]
    L.pfaai_gpu_index_slice_buffers.argtypes = [vp, i32, i32, i64, i64, i64, C.POINTER(vp), C.POINTER(vp)]
    L.pfaai_gpu_route_prepare.argtypes = [vp, C.POINTER(Route)]
    L.pfaai_gpu_route_recv_buffer.argtypes = [vp, i64, C.POINTER(vp)]
    L.pfaai_gpu_build_routed.argtypes = [vp, i64]
    L.pfaai_gpu_route_local.argtypes = [C.POINTER(vp), i32]
    L.pfaai_gpu_set_global_matrix.argtypes = [vp, i32]
    L.pfaai_gpu_set_path.argtypes = [vp, i32, i32]
    L.pfaai_gpu_host_register.argtypes = [vp, C.c_size_t]
    L.pfaai_gpu_host_unregister.argtypes = [vp, C.c_size_t]
    L.pfaai_gpu_e_size.argtypes = [vp, C.POINTER(i64)]
    L.pfaai_gpu_set_progress.argtypes = [vp, i32]
    L.pfaai_gpu_progress_poll.argtypes = [vp, C.POINTER(i64), C.POINTER(C.c_int)]
    L.pfaai_gpu_microbench.argtypes = [vp, i32, C.c_double, i32, C.POINTER(Microbench)]
    L.pfaai_gpu_prefix_sum_static.argtypes = [i32, vp, vp, i64, vp, vp]
    return L


def _p(a):
    return None if a is None else C.c_void_p(a.ctypes.data)


class PfaaiError(RuntimeError):
    pass


def host_register(ptr: int, nbytes: int) -> bool:
    """Page-lock + map a caller-allocated host range (pfaai_gpu_host_register); False = the driver refused (no error left behind)."""
    return load().pfaai_gpu_host_register(C.c_void_p(ptr), nbytes) == 0


def host_unregister(ptr: int, nbytes: int) -> bool:
    return load().pfaai_gpu_host_unregister(C.c_void_p(ptr), nbytes) == 0


STATIC_SCAN_MAX = 4194304   # PFAAI_GPU_STATIC_SCAN_MAX


def prefix_sum_static(device: int, stream_ptr: int, in_dev: int, n: int, out_dev: int, out32_dev: int = 0):
    """pfaai_gpu_prefix_sum_static: exclusive scan of n int32 counts into n + 1 int64 sums, for callers that must not
    allocate device memory (the scratch is a fixed array inside the library).  All pointers are DEVICE addresses of buffers
    the caller owns; the kernels are only enqueued on the stream -- synchronise it before reading the outputs."""
    L = load()
    rc = L.pfaai_gpu_prefix_sum_static(int(device), C.c_void_p(stream_ptr or None), C.c_void_p(in_dev or None), int(n),
                                       C.c_void_p(out_dev or None), C.c_void_p(out32_dev or None))
    if rc != 0:
        raise PfaaiError(f"prefix_sum_static rc={rc}: {L.pfaai_gpu_last_error(None).decode()}")


class PfaaiGpu:
    """Thin object wrapper over one pfaai_gpu_ctx (one CUDA device)."""

    def __init__(self, device=0, lib=None):
        self._L = lib if lib is not None else load()
        h = C.c_void_p()
        rc = self._L.pfaai_gpu_create(C.byref(h), device, 0)
        if rc != 0:
            raise PfaaiError(f"pfaai_gpu_create rc={rc}: {self._L.pfaai_gpu_last_error(None).decode()}")
        self._h = h
        self.P = self.G = 0

    def close(self):
        if getattr(self, "_h", None):
            self._L.pfaai_gpu_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _check(self, rc, what):
        if rc != 0:
            raise PfaaiError(f"{what} rc={rc}: {self._L.pfaai_gpu_last_error(self._h).decode()}")

    def set_stream(self, cuda_stream_ptr):
        self._check(self._L.pfaai_gpu_set_stream(self._h, C.c_void_p(cuda_stream_ptr)), "set_stream")

    def set_shard(self, shard, n_shards):
        self._check(self._L.pfaai_gpu_set_shard(self._h, shard, n_shards), "set_shard")
        self._n_shards, self._shard = n_shards, shard

    def set_global_matrix(self, on=True):
        """Sharded all-vs-all: address the full G x G matrix (shared by all shards) and store the mirror images too."""
        self._check(self._L.pfaai_gpu_set_global_matrix(self._h, int(bool(on))), "set_global_matrix")
        self._global = bool(on)

    def set_path(self, path=PATH_AUTO, tau=0):
        """Tile path of the next computes: PATH_AUTO (cost model), PATH_LISTS (dictionary walk), PATH_MMA (tcgen05 dense
        contraction + sparse walk); tau > 0 fixes the dense path's posting-list length threshold."""
        self._check(self._L.pfaai_gpu_set_path(self._h, int(path), int(tau)), "set_path")

    def e_size(self):
        """|E| of the built index (all-vs-all): sum over posting lists of n(n-1)/2."""
        e = C.c_int64()
        self._check(self._L.pfaai_gpu_e_size(self._h, C.byref(e)), "e_size")
        return int(e.value)

    def upload_index(self, P, G, F, Lc, T=None):
        F = np.ascontiguousarray(F, dtype=PG_DTYPE)
        Lc = np.ascontiguousarray(Lc, dtype=np.int32)
        if Lc.shape[0] != NTETRA:
            raise ValueError("Lc must have 160000 entries")
        if T is not None:
            T = np.ascontiguousarray(T, dtype=np.int32)
            if T.shape != (P, G):
                raise ValueError("T must be P x G")
        self.P, self.G, self.nF = P, G, F.shape[0]
        self._check(self._L.pfaai_gpu_upload_index(self._h, P, G, _p(F), _p(Lc), F.shape[0], _p(T)), "upload_index")

    def upload_triples(self, P, G, tetramer, F, T=None):
        """(tetramer, protein, genome) triples in tetramer-major order; Lc is built on the device (pfaai_gpu_upload_triples)."""
        F = np.ascontiguousarray(F, dtype=PG_DTYPE)
        tetramer = np.ascontiguousarray(tetramer, dtype=np.int32)
        if tetramer.shape[0] != F.shape[0]:
            raise ValueError("one tetramer key per posting")
        if T is not None:
            T = np.ascontiguousarray(T, dtype=np.int32)
        self.P, self.G, self.nF = P, G, F.shape[0]
        self._check(self._L.pfaai_gpu_upload_triples(self._h, P, G, _p(tetramer), _p(F), F.shape[0], _p(T)), "upload_triples")

    def import_lists(self, P, G, Lc, seg_off, seg_vals, T=None):
        """Device-ready hand-over (pfaai_gpu_import_lists): the genome-major lists of get_arrays(); no F, no build."""
        Lc = np.ascontiguousarray(Lc, dtype=np.int32)
        seg_off = np.ascontiguousarray(seg_off, dtype=np.int64)
        seg_vals = np.ascontiguousarray(seg_vals).view(np.uint32) if seg_vals.dtype.itemsize == 4 else np.ascontiguousarray(seg_vals, dtype=np.uint32)
        if T is not None:
            T = np.ascontiguousarray(T, dtype=np.int32)
        if seg_off.shape[0] != P * G + 1:
            raise ValueError("seg_off must have P*G + 1 entries")
        self.P, self.G, self.nF = P, G, int(seg_vals.shape[0])
        self._check(self._L.pfaai_gpu_import_lists(self._h, P, G, self.nF, _p(Lc), _p(T), _p(seg_off), _p(seg_vals)), "import_lists")

    def set_progress(self, on=True):
        self._check(self._L.pfaai_gpu_set_progress(self._h, int(on)), "set_progress")

    def progress_poll(self):
        """-> (rows_final, finished); callable from another thread while compute() blocks (pfaai_gpu_progress_poll)."""
        r, f = C.c_int64(), C.c_int()
        self._check(self._L.pfaai_gpu_progress_poll(self._h, C.byref(r), C.byref(f)), "progress_poll")
        return int(r.value), bool(f.value)

    def index_device_buffers(self, P, G, nF, host_T=False):
        """Size the device index buffers and return their pointers (F, Lc, T) for a caller-side fill (NCCL broadcast)."""
        f, lc, t = C.c_void_p(), C.c_void_p(), C.c_void_p()
        self._check(self._L.pfaai_gpu_index_device_buffers(self._h, P, G, nF, int(host_T), C.byref(f), C.byref(lc), C.byref(t)),
                    "index_device_buffers")
        self.P, self.G, self.nF = P, G, nF
        return f.value, lc.value, t.value

    def shard_layout(self):
        """-> (axis, ids): the sharded axis of the last compute's matrix and the global indices stored along it."""
        axis, n = C.c_int(), C.c_int64()
        self._check(self._L.pfaai_gpu_shard_layout(self._h, C.byref(axis), C.byref(n), None), "shard_layout")
        ids = np.zeros(n.value, np.int32)
        self._check(self._L.pfaai_gpu_shard_layout(self._h, C.byref(axis), C.byref(n), _p(ids)), "shard_layout")
        return axis.value, ids

    def build(self):
        self._check(self._L.pfaai_gpu_build(self._h), "build")

    # ---- routed build: the context holds a slice of F, postings travel to the shard that owns their genome ----
    def upload_index_slice(self, P, G, F_slice, Lc, nF_total, slice_begin):
        F_slice = np.ascontiguousarray(F_slice, dtype=PG_DTYPE)
        Lc = np.ascontiguousarray(Lc, dtype=np.int32)
        self.P, self.G, self.nF = P, G, nF_total
        self._check(self._L.pfaai_gpu_upload_index_slice(self._h, P, G, _p(F_slice), _p(Lc), nF_total, slice_begin, F_slice.shape[0]),
                    "upload_index_slice")

    def index_slice_buffers(self, P, G, nF_total, slice_begin, n_slice):
        f, lc = C.c_void_p(), C.c_void_p()
        self._check(self._L.pfaai_gpu_index_slice_buffers(self._h, P, G, nF_total, slice_begin, n_slice, C.byref(f), C.byref(lc)),
                    "index_slice_buffers")
        self.P, self.G, self.nF = P, G, nF_total
        return f.value, lc.value

    def route_prepare(self):
        """-> (send device pointer, offsets [n_shards + 1]): shard d receives items [off[d], off[d+1]) (8 bytes each)."""
        r = Route()
        self._check(self._L.pfaai_gpu_route_prepare(self._h, C.byref(r)), "route_prepare")
        return r.send_dev, [int(r.send_offset[d]) for d in range(r.n_shards + 1)]

    def route_recv_buffer(self, n_items):
        p = C.c_void_p()
        self._check(self._L.pfaai_gpu_route_recv_buffer(self._h, int(n_items), C.byref(p)), "route_recv_buffer")
        return p.value

    def build_routed(self, n_items):
        self._check(self._L.pfaai_gpu_build_routed(self._h, int(n_items)), "build_routed")

    @staticmethod
    def route_local(ctxs):
        """route_prepare + peer copies + build_routed for the shards' contexts of this process (shards 0..n-1 in order)."""
        L = load()
        arr = (C.c_void_p * len(ctxs))(*[c._h for c in ctxs])
        rc = L.pfaai_gpu_route_local(arr, len(ctxs))
        if rc != 0:
            raise PfaaiError(f"route_local rc={rc}: " + "; ".join(L.pfaai_gpu_last_error(c._h).decode() for c in ctxs))

    def build_partitioned(self):
        """Sharded all-vs-all: index part + dictionaries of this shard's own row blocks only (exchange, then finish)."""
        self._check(self._L.pfaai_gpu_build_partitioned(self._h), "build_partitioned")

    def exchange_buffers(self):
        """-> list of (device pointer, chunk bytes, used bytes) for the four buffers to all-gather in place."""
        x = Exchange()
        self._check(self._L.pfaai_gpu_exchange_buffers(self._h, C.byref(x)), "exchange_buffers")
        return [(x.buf_dev[k], int(x.chunk_bytes[k]), int(x.used_bytes[k])) for k in range(4)]

    def set_partition_chunk(self, blob_chunk_bytes):
        self._check(self._L.pfaai_gpu_set_partition_chunk(self._h, int(blob_chunk_bytes)), "set_partition_chunk")

    def finish_partitioned(self):
        self._check(self._L.pfaai_gpu_finish_partitioned(self._h), "finish_partitioned")

    def finish_partitioned_async(self, cuda_event_ptr):
        """The exchange is still in flight on another stream; `cuda_event_ptr` (cudaEvent_t) fires when it is complete."""
        self._check(self._L.pfaai_gpu_finish_partitioned_async(self._h, C.c_void_p(cuda_event_ptr)), "finish_partitioned_async")

    @staticmethod
    def exchange_local(ctxs):
        """Exchange + finish between the shards' contexts of this process (peer copies)."""
        L = load()
        arr = (C.c_void_p * len(ctxs))(*[c._h for c in ctxs])
        rc = L.pfaai_gpu_exchange_local(arr, len(ctxs))
        if rc != 0:
            raise PfaaiError(f"exchange_local rc={rc}: " + "; ".join(L.pfaai_gpu_last_error(c._h).decode() for c in ctxs))

    def get_arrays(self, want_lists=True):
        Lc = np.zeros(NTETRA, np.int32)
        Lp = np.zeros(NTETRA, np.int32)
        T = np.zeros((self.P, self.G), np.int32)
        off = np.zeros(self.P * self.G + 1, np.int64) if want_lists else None
        vals = np.zeros(self.nF, np.int32) if want_lists else None
        self._check(self._L.pfaai_gpu_get_arrays(self._h, _p(Lc), _p(Lp), _p(T), _p(off), _p(vals)), "get_arrays")
        return {"Lc": Lc, "Lp": Lp, "T": T, "seg_off": off, "seg_vals": vals}

    def n_pairs(self, mode, n_qry=0):
        return int(self._L.pfaai_gpu_n_pairs(mode, self.G, n_qry))

    def compute_device(self, mode, qry_ids=None, flags=0):
        q = None if qry_ids is None else np.ascontiguousarray(qry_ids, dtype=np.int32)
        nq = 0 if q is None else q.shape[0]
        self._check(self._L.pfaai_gpu_compute_device(self._h, mode, _p(q), nq, flags), "compute_device")

    def matrix_shape(self):
        r, c = C.c_int64(), C.c_int64()
        self._check(self._L.pfaai_gpu_matrix_shape(self._h, C.byref(r), C.byref(c)), "matrix_shape")
        return int(r.value), int(c.value)

    def matrix_dev_ptr(self):
        return self._L.pfaai_gpu_matrix_dev(self._h)

    def download_matrix(self, out=None):
        r, c = self.matrix_shape()
        if out is None:
            out = np.empty((r, c), np.float64)
        self._check(self._L.pfaai_gpu_download_matrix(self._h, _p(out)), "download_matrix")
        return out

    def compute(self, mode, qry_ids=None, n_qry=None, flags=0, want_jac=True, want_counts=False, matrix_out=None):
        """Host-buffer call. Returns dict(matrix, jac, aji, counts)."""
        q = None if qry_ids is None else np.ascontiguousarray(qry_ids, dtype=np.int32)
        nq = (0 if q is None else q.shape[0]) if n_qry is None else n_qry
        if mode == ALL_VS_ALL:
            shape = (self.G, self.G)
        elif mode == QUERY_SUBSET:
            shape = (nq, self.G)
        else:
            shape = (nq, self.G - nq)
        ns = getattr(self, "_n_shards", 1)
        if ns > 1:
            if want_jac or want_counts:
                raise PfaaiError("the JAC / counts hooks need an unsharded context")
            from . import sharding
            shape = sharding.shard_shape(mode, self.G, nq, self._shard, ns)
            if getattr(self, "_global", False) and mode == ALL_VS_ALL:
                shape = (self.G, self.G)
            M = matrix_out if matrix_out is not None else np.zeros(shape, np.float64)
            if M.shape != tuple(shape):
                raise ValueError(f"matrix_out must have the shard's shape {shape}")
            self._check(self._L.pfaai_gpu_compute(self._h, mode, _p(q), nq, flags, _p(M), None, None, None), "compute")
            return {"matrix": M, "jac": None, "aji": None, "counts": None}
        M = matrix_out if matrix_out is not None else np.empty(shape, np.float64)
        npairs = self.n_pairs(mode, nq)
        jac = np.zeros(npairs, JAC_DTYPE) if want_jac else None
        aji = np.zeros(npairs, np.float64) if want_jac else None
        cnt = np.zeros((npairs, self.P), np.int32) if want_counts else None
        self._check(self._L.pfaai_gpu_compute(self._h, mode, _p(q), nq, flags, _p(M), _p(jac), _p(aji), _p(cnt)), "compute")
        return {"matrix": M, "jac": jac, "aji": aji, "counts": cnt}

    def microbench(self, variant=0, hit_rate=0.43, steps=20000):
        """On-chip ceiling of the list path's walk loop on this GPU (pfaai_gpu_microbench) -> dict."""
        m = Microbench()
        self._check(self._L.pfaai_gpu_microbench(self._h, variant, float(hit_rate), int(steps), C.byref(m)), "microbench")
        return {k: getattr(m, k) for k, _ in m._fields_ if k != "reserved_"}

    def timings(self):
        t = Timings()
        self._check(self._L.pfaai_gpu_get_timings(self._h, C.byref(t)), "get_timings")
        return t.as_dict()
