"""Readers/writers for the reference's golden ``.bin`` fixtures.

The reference serialises its test goldens with cereal's ``BinaryOutputArchive``
(little-endian, no header, no padding).  The layouts are fixed by the
``serialize``/``save`` members of the reference PODs:

* ``std::vector<int>``                      -> ``u64 n`` + ``n x i32``
* ``std::vector<DPair<int,int>>``           -> ``u64 n`` + ``n x {i32 first, i32 second}``
  (reference ``include/pfaai/utils.hpp:204-225``)
* ``DMatrix<int>``                          -> ``u64 rows, u64 cols, u64 n`` + ``n x i32`` row-major
  (reference ``include/pfaai/utils.hpp:240-288``)
* ``std::vector<ETriple<int>>``             -> ``u64 n`` + ``n x {i32 protein, i32 gA, i32 gB}``
  (reference ``include/pfaai/interface.hpp:92-121``)
* ``std::vector<JACTuple<int,double>>``     -> ``u64 n`` + ``n x {i32 gA, i32 gB, f64 S, i32 N}`` (20 B, unaligned)
  (reference ``include/pfaai/interface.hpp:61-75``)
* ``std::vector<double>``                   -> ``u64 n`` + ``n x f64``

This module is host-side tooling only (fixtures, tests); nothing on the GPU
path depends on it.
"""
from __future__ import annotations

import numpy as np

PG_DTYPE = np.dtype([("protein", "<i4"), ("genome", "<i4")])
E_DTYPE = np.dtype([("protein", "<i4"), ("gA", "<i4"), ("gB", "<i4")])
JAC_DTYPE = np.dtype([("gA", "<i4"), ("gB", "<i4"), ("S", "<f8"), ("N", "<i4")])  # packed: 20 B
assert JAC_DTYPE.itemsize == 20


def _read_vec(path, dtype):
    raw = np.fromfile(path, dtype=np.uint8)
    n = int(raw[:8].view("<u8")[0])
    body = raw[8:]
    dt = np.dtype(dtype)
    if body.size != n * dt.itemsize:
        raise ValueError(f"{path}: header says {n} x {dt.itemsize} B, body has {body.size} B")
    return body.view(dt).copy()


def read_int_vector(path):
    """``std::vector<int>`` (Lc, Lp, gpe_starts, ...)."""
    return _read_vec(path, "<i4")


def read_double_vector(path):
    """``std::vector<double>`` (AJI)."""
    return _read_vec(path, "<f8")


def read_f_array(path):
    """``std::vector<DPair<int,int>>`` -> structured array (protein, genome)."""
    return _read_vec(path, PG_DTYPE)


def read_e_array(path):
    """``std::vector<ETriple<int>>`` -> structured array (protein, gA, gB)."""
    return _read_vec(path, E_DTYPE)


def read_jac(path):
    """``std::vector<JACTuple<int,double>>`` -> structured array (gA, gB, S, N)."""
    return _read_vec(path, JAC_DTYPE)


def read_t_matrix(path):
    """``DMatrix<int>`` -> 2-D int32 array (rows = proteins, cols = genomes)."""
    raw = np.fromfile(path, dtype=np.uint8)
    rows, cols, n = (int(x) for x in raw[:24].view("<u8"))
    body = raw[24:]
    if n != rows * cols or body.size != 4 * n:
        raise ValueError(f"{path}: bad DMatrix header {rows}x{cols}, n={n}, body {body.size} B")
    return body.view("<i4").reshape(rows, cols).copy()


def _write_vec(path, arr, dtype):
    arr = np.ascontiguousarray(arr, dtype=dtype)
    with open(path, "wb") as fh:
        fh.write(np.uint64(arr.shape[0]).tobytes())
        fh.write(arr.tobytes())


def write_int_vector(path, arr):
    _write_vec(path, arr, "<i4")


def write_double_vector(path, arr):
    _write_vec(path, arr, "<f8")


def write_f_array(path, arr):
    _write_vec(path, arr, PG_DTYPE)


def write_e_array(path, arr):
    _write_vec(path, arr, E_DTYPE)


def write_jac(path, arr):
    _write_vec(path, arr, JAC_DTYPE)


def write_t_matrix(path, mat):
    mat = np.ascontiguousarray(mat, dtype="<i4")
    with open(path, "wb") as fh:
        fh.write(np.array([mat.shape[0], mat.shape[1], mat.size], dtype="<u8").tobytes())
        fh.write(mat.tobytes())
