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
