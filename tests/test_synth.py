"""Synthetic generator (tooling) sanity + oracle self-consistency on synthetic inputs. CPU only."""
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as ge  # noqa: E402

ge.build()
from oracle import pfaai_oracle as orc  # noqa: E402
from parfastaai_b200 import synth  # noqa: E402


def test_generator_statistics_and_determinism():
    N, P = 64, 122
    off, vals = synth.generate(N, P, seed=1234)
    off2, vals2 = synth.generate(N, P, seed=1234)
    assert np.array_equal(off, off2) and np.array_equal(vals, vals2)
    lens = np.diff(off)
    assert 280 < lens.mean() < 320 and lens.max() <= 597 and lens.min() >= 60
    assert vals.min() >= 0 and vals.max() < 160000
    # ascending, unique inside every list
    d = np.diff(vals)
    starts = off[1:-1]
    inner = np.ones(vals.shape[0] - 1, bool)
    inner[starts[(starts > 0) & (starts < vals.shape[0])] - 1] = False
    assert np.all(d[inner] > 0)


def test_reference_arrays_roundtrip_and_direct_agrees():
    N, P = 48, 30
    off, vals = synth.generate(N, P, seed=7)
    Lc, F = synth.to_reference(N, P, off, vals)
    T = synth.t_matrix(N, P, off)
    o2, v2 = orc.genome_major(P, N, Lc, F)
    assert np.array_equal(o2, off) and np.array_equal(v2, vals)
    r = orc.full_pipeline(orc.ALL_VS_ALL, N, P, Lc, F, T)
    ids = np.arange(N, dtype=np.int32)
    d = orc.direct_aji(P, N, off, vals, T, ids, ids, want_counts=True)
    iu = np.triu_indices(N, 1)
    assert np.array_equal(d["counts"][iu], r["counts"])
    assert np.array_equal(d["aji"], r["matrix"])
    assert (r["jac"]["N"] == 0).sum() == 0          # no zero-share pairs


def test_subset_generation_reproduces_the_full_database_lists():
    """generate_subset: the lists of a few genomes of a large synthetic database, without generating the database
    (bench.py's per-rank parity check relies on it)."""
    from parfastaai_b200 import synth
    N, P = 300, 5
    off, vals = synth.generate(N, P, seed=7)
    ids = np.array([5, 299, 17, 0], np.int32)
    for dense in (False, True):
        offd, valsd = synth.generate(N, P, seed=7, dense=dense)
        o2, v2 = synth.generate_subset(N, P, 7, ids, dense=dense)
        for k, g in enumerate(ids):
            for p in range(P):
                assert np.array_equal(valsd[offd[g * P + p]:offd[g * P + p + 1]], v2[o2[k * P + p]:o2[k * P + p + 1]])
