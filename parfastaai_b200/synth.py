"""Synthetic FastAAI-like inputs (SURVEY 8(d) recipe; BASELINE configs 4/5) -- host tooling.

``generate(N, P, seed)`` returns the genome-major tetramer lists; ``to_reference(...)`` converts them to
the arrays the reference's DataBaseInterface delivers (Lc, F in (tetramer, protein, genome) order).
The heavy lifting is in csrc/synth_host.c (plain C + OpenMP, counter-based RNG, thread-count invariant).
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

from ._lib import NTETRA, PG_DTYPE

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libpfaai_synth.so")
_lib = None


def _load():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(f"{LIB_PATH} missing: run __graft_entry__.build()")
        L = C.CDLL(LIB_PATH)
        L.pfs_count.restype = C.c_int64
        L.pfs_count.argtypes = [C.c_int, C.c_int, C.c_uint64, C.c_int, C.c_void_p]
        L.pfs_fill.restype = None
        L.pfs_fill.argtypes = [C.c_int, C.c_int, C.c_uint64, C.c_int, C.c_void_p, C.c_void_p]
        L.pfs_count_subset.restype = C.c_int64
        L.pfs_count_subset.argtypes = [C.c_int, C.c_int, C.c_uint64, C.c_int, C.c_void_p, C.c_int, C.c_void_p]
        L.pfs_fill_subset.restype = None
        L.pfs_fill_subset.argtypes = [C.c_int, C.c_int, C.c_uint64, C.c_int, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p]
        L.pfs_to_reference.restype = None
        L.pfs_to_reference.argtypes = [C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        L.pfs_protein_length.restype = C.c_int
        L.pfs_protein_length.argtypes = [C.c_uint64, C.c_int]
        _lib = L
    return _lib


def generate(N, P=122, seed=1234, dense=False):
    """-> (seg_off [N*P+1] int64, vals [nF] int32): list (g, p) = vals[seg_off[g*P+p] : seg_off[g*P+p+1]]."""
    L = _load()
    seg_off = np.zeros(N * P + 1, np.int64)
    nF = L.pfs_count(N, P, seed, int(dense), seg_off.ctypes.data)
    vals = np.empty(nF, np.int32)
    L.pfs_fill(N, P, seed, int(dense), seg_off.ctypes.data, vals.ctypes.data)
    return seg_off, vals


def generate_subset(N, P, seed, genomes, dense=False):
    """The lists of `genomes` (ids of the N-genome database generate(N, P, seed, dense) would produce), renumbered 0..n-1:
    -> (seg_off [n*P+1], vals).  Every list is a pure function of (seed, N, dense, genome, SCP)."""
    L = _load()
    ids = np.ascontiguousarray(genomes, dtype=np.int32)
    n = int(ids.shape[0])
    seg_off = np.zeros(n * P + 1, np.int64)
    nF = L.pfs_count_subset(N, P, seed, int(dense), ids.ctypes.data, n, seg_off.ctypes.data)
    vals = np.empty(nF, np.int32)
    L.pfs_fill_subset(N, P, seed, int(dense), ids.ctypes.data, n, seg_off.ctypes.data, vals.ctypes.data)
    return seg_off, vals


def to_reference(N, P, seg_off, vals, out_F=None, out_Lc=None):
    """-> (Lc [160000] int32, F [nF] (protein, genome)) in the reference's order; T = diff(seg_off)."""
    L = _load()
    nF = int(seg_off[-1])
    Lc = out_Lc if out_Lc is not None else np.zeros(NTETRA, np.int32)
    F = out_F if out_F is not None else np.empty(nF, PG_DTYPE)
    L.pfs_to_reference(N, P, seg_off.ctypes.data, vals.ctypes.data, Lc.ctypes.data, F.ctypes.data)
    return Lc, F


def t_matrix(N, P, seg_off):
    """T [P x N]: tetramer count per (SCP, genome) (== proteinTetramerCounts in the single-DB modes)."""
    return np.ascontiguousarray(np.diff(seg_off).reshape(N, P).T.astype(np.int32))


def protein_lengths(P=122, seed=1234):
    L = _load()
    return np.array([L.pfs_protein_length(seed, p) for p in range(P)], np.int32)
