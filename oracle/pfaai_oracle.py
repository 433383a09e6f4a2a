"""ctypes binding of oracle/libpfaai_oracle.so (the CPU restatement, pfaai_oracle.c).

TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / --impl reference legs -- never by the product package ``parfastaai_b200``.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libpfaai_oracle.so")
REF_BIN = os.path.join(HERE, "_ref", "par_fastaai_ref.x")
NTETRA = 160000

PG = np.dtype([("protein", "<i4"), ("genome", "<i4")])
ET = np.dtype([("protein", "<i4"), ("gA", "<i4"), ("gB", "<i4")])
JAC = np.dtype([("gA", "<i4"), ("gB", "<i4"), ("S", "<f8"), ("N", "<i4"), ("pad", "<i4")])  # C layout, 24 B
assert JAC.itemsize == 24

ALL_VS_ALL, QUERY_SUBSET, QUERY_DB = 0, 1, 2

_lib = None


def build():
    """Compile the restatement (and, where /root/reference is mounted, the reference binary)."""
    subprocess.run(["make", "-C", HERE, "libpfaai_oracle.so"], check=True, capture_output=True)
    if os.path.isdir("/root/reference/src") and not os.path.exists(REF_BIN):
        subprocess.run(["make", "-C", HERE, "ref"], check=True, capture_output=True)


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            build()
        L = C.CDLL(LIB_PATH)
        L.pfo_n_pairs.restype = C.c_int64
        L.pfo_n_pairs.argtypes = [C.c_int, C.c_int, C.c_int]
        _lib = L
    return _lib


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def _i32(a):
    return None if a is None else np.ascontiguousarray(a, dtype=np.int32)


def n_pairs(mode, G, nQ=0):
    return int(lib().pfo_n_pairs(mode, G, nQ))


def init_jac(mode, G, nQ=0, qry_ids=None, quirk=False):
    q = _i32(qry_ids)
    jac = np.zeros(n_pairs(mode, G, nQ), dtype=JAC)
    rc = lib().pfo_init_jac(C.c_int(mode), C.c_int(G), C.c_int(nQ), _p(q), C.c_int(int(quirk)), _p(jac))
    if rc:
        raise ValueError(f"pfo_init_jac rc={rc}")
    return jac


def build_index(protein, tetramer, genome, P, G):
    """Unsorted postings -> (Lc, Lp, F, T) exactly as the reference's constructL/F/T produce them."""
    protein, tetramer, genome = _i32(protein), _i32(tetramer), _i32(genome)
    n = protein.shape[0]
    Lc = np.zeros(NTETRA, np.int32)
    Lp = np.zeros(NTETRA, np.int32)
    F = np.zeros(n, PG)
    T = np.zeros((P, G), np.int32)
    rc = lib().pfo_build_index(C.c_int64(n), _p(protein), _p(tetramer), _p(genome), C.c_int(P), C.c_int(G),
                               _p(Lc), _p(Lp), _p(F), _p(T))
    if rc:
        raise ValueError(f"pfo_build_index rc={rc}")
    return Lc, Lp, F, T


def prefix_sum(Lc):
    Lc = _i32(Lc)
    Lp = np.zeros_like(Lc)
    lib().pfo_prefix_sum(_p(Lc), _p(Lp), C.c_int(Lc.shape[0]))
    return Lp


def counts(mode, G, P, Lc, F, nQ=0, qry_ids=None):
    """Per-(pair, SCP) intersection counts [n_pairs, P] in JAC order, and |E|."""
    Lc = _i32(Lc)
    F = np.ascontiguousarray(F, dtype=PG)
    q = _i32(qry_ids)
    npairs = n_pairs(mode, G, nQ)
    cnt = np.zeros((npairs, P), np.int32)
    esz = C.c_int64(0)
    rc = lib().pfo_counts(C.c_int(mode), C.c_int(G), C.c_int(nQ), _p(q), C.c_int(P), _p(Lc), _p(F),
                          C.c_int64(F.shape[0]), _p(cnt), C.byref(esz))
    if rc:
        raise ValueError(f"pfo_counts rc={rc}")
    return cnt, int(esz.value)


def sorted_e(mode, G, P, cnt, nQ=0, qry_ids=None):
    q = _i32(qry_ids)
    cnt = np.ascontiguousarray(cnt, dtype=np.int32)
    E = np.zeros(int(cnt.sum()), ET)
    rc = lib().pfo_sorted_e(C.c_int(mode), C.c_int(G), C.c_int(nQ), _p(q), C.c_int(P), _p(cnt), _p(E))
    if rc:
        raise ValueError(f"pfo_sorted_e rc={rc}")
    return E


def jac_aji(cnt, T, jac):
    """Fills S, N of `jac` (labels from init_jac) and returns (jac, aji)."""
    cnt = np.ascontiguousarray(cnt, dtype=np.int32)
    T = np.ascontiguousarray(T, dtype=np.int32)
    jac = np.ascontiguousarray(jac, dtype=JAC).copy()
    aji = np.zeros(jac.shape[0], np.float64)
    rc = lib().pfo_jac_aji(C.c_int64(jac.shape[0]), C.c_int(cnt.shape[1]), _p(cnt), _p(T), C.c_int(T.shape[1]),
                           _p(jac), _p(aji))
    if rc:
        raise ValueError(f"pfo_jac_aji rc={rc}")
    return jac, aji


def fill_matrix(mode, G, jac_true, aji, nQ=0, qry_ids=None):
    q = _i32(qry_ids)
    jac_true = np.ascontiguousarray(jac_true, dtype=JAC)
    aji = np.ascontiguousarray(aji, dtype=np.float64)
    if mode == ALL_VS_ALL:
        rows, cols = G, G
    elif mode == QUERY_SUBSET:
        rows, cols = nQ, G
    else:
        rows, cols = nQ, G - nQ
    M = np.zeros((rows, cols), np.float64)
    rc = lib().pfo_fill_matrix(C.c_int(mode), C.c_int(G), C.c_int(nQ), _p(q), _p(jac_true), _p(aji), _p(M))
    if rc:
        raise ValueError(f"pfo_fill_matrix rc={rc}")
    return M


def genome_major(P, G, Lc, F):
    Lc = _i32(Lc)
    F = np.ascontiguousarray(F, dtype=PG)
    off = np.zeros(G * P + 1, np.int64)
    vals = np.zeros(F.shape[0], np.int32)
    rc = lib().pfo_genome_major(C.c_int(P), C.c_int(G), _p(Lc), _p(F), C.c_int64(F.shape[0]), _p(off), _p(vals))
    if rc:
        raise ValueError(f"pfo_genome_major rc={rc}")
    return off, vals


def direct_aji(P, G, off, vals, T, rows, cols, row_t=None, col_t=None, want_counts=False):
    """Definition-level rectangle: returns dict(S, N, aji[, counts]) of shape [n_rows, n_cols]."""
    off = np.ascontiguousarray(off, dtype=np.int64)
    vals = _i32(vals)
    T = np.ascontiguousarray(T, dtype=np.int32)
    rows, cols, row_t, col_t = _i32(rows), _i32(cols), _i32(row_t), _i32(col_t)
    nr, nc = rows.shape[0], cols.shape[0]
    S = np.zeros((nr, nc), np.float64)
    N = np.zeros((nr, nc), np.int32)
    aji = np.zeros((nr, nc), np.float64)
    cnt = np.zeros((nr, nc, P), np.int32) if want_counts else None
    rc = lib().pfo_direct_aji(C.c_int(P), C.c_int(G), _p(off), _p(vals), _p(T), C.c_int(T.shape[1]),
                              C.c_int(nr), _p(rows), _p(row_t), C.c_int(nc), _p(cols), _p(col_t),
                              _p(S), _p(N), _p(aji), _p(cnt))
    if rc:
        raise ValueError(f"pfo_direct_aji rc={rc}")
    out = {"S": S, "N": N, "aji": aji}
    if want_counts:
        out["counts"] = cnt
    return out


def full_pipeline(mode, G, P, Lc, F, T, nQ=0, qry_ids=None, quirk=False):
    """Lc/F/T -> counts -> JAC -> AJI -> output matrix, the whole reference path on the CPU."""
    cnt, esz = counts(mode, G, P, Lc, F, nQ, qry_ids)
    jac, aji = jac_aji(cnt, T, init_jac(mode, G, nQ, qry_ids, quirk))
    M = fill_matrix(mode, G, init_jac(mode, G, nQ, qry_ids, False), aji, nQ, qry_ids)
    return {"counts": cnt, "e_size": esz, "jac": jac, "aji": aji, "matrix": M}
