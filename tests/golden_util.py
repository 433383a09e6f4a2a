"""Helpers shared by the parity tests: load the re-packed reference goldens (tests/golden)."""
import functools
import gzip
import json
import os
import shutil
import tempfile

import numpy as np

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


@functools.lru_cache(maxsize=None)
def load(name):
    with np.load(os.path.join(GOLDEN, name + ".npz")) as z:
        return {k: z[k] for k in z.files}


@functools.lru_cache(maxsize=None)
def names():
    with open(os.path.join(GOLDEN, "names.json")) as fh:
        return json.load(fh)


def csv_bytes(name):
    with open(os.path.join(GOLDEN, name), "rb") as fh:
        return fh.read()


def query_subset_ids():
    """ids (in the 20-genome DB) of the genomes listed in qsub_test_input.txt, in file order."""
    g = names()["xantho_genomes"]
    with open(os.path.join(GOLDEN, "qsub_test_input.txt")) as fh:
        q = fh.read().split()
    return [g.index(x) for x in q]


_dbdir = None


def db_path(name):
    """Decompress tests/golden/<name>.gz into a per-process temp dir and return the path."""
    global _dbdir
    if _dbdir is None:
        _dbdir = tempfile.mkdtemp(prefix="pfaai_golden_")
    dst = os.path.join(_dbdir, name)
    if not os.path.exists(dst):
        with gzip.open(os.path.join(GOLDEN, name + ".gz"), "rb") as fi, open(dst, "wb") as fo:
            shutil.copyfileobj(fi, fo)
    return dst


def run_lengths_from_sorted_e(E, pair_index, n_pairs, P):
    """[n_pairs, P] counts from a sorted-E fixture; pair_index(gA, gB) -> JAC index (vectorised)."""
    cnt = np.zeros((n_pairs, P), np.int32)
    idx = pair_index(E["gA"].astype(np.int64), E["gB"].astype(np.int64))
    np.add.at(cnt, (idx, E["protein"]), 1)
    return cnt


def xantho_db_path():
    """Rebuild the missing 20-genome data/modified_xantho_fastaai2.db from the F/Lc goldens and the name tables
    (SURVEY App. C.3): functionally equivalent for every query the reference and our loader issue."""
    dst_dir = db_path("xdb_subset1.db")  # makes the temp dir
    dst = os.path.join(os.path.dirname(dst_dir), "modified_xantho_fastaai2.db")
    if not os.path.exists(dst):
        from parfastaai_b200 import fastaai_db
        g = load("xanthodb")
        P, G = g["t"].shape
        tet = np.repeat(np.arange(160000, dtype=np.int64), g["lc"])
        seg = g["f"]["genome"].astype(np.int64) * P + g["f"]["protein"]
        order = np.lexsort((tet, seg))
        seg_off = np.zeros(G * P + 1, np.int64)
        np.cumsum(np.bincount(seg, minlength=G * P), out=seg_off[1:])
        n = names()
        fastaai_db.write_db(dst, n["xantho_genomes"], n["xantho_proteins"], seg_off, tet[order].astype(np.int32))
    return dst
