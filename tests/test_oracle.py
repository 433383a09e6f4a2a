"""Pins oracle/pfaai_oracle.c against every golden vector the reference ships for the path
(SURVEY 8(c)): Lc/Lp/F/T, sorted E, JAC (S, N), AJI for the five fixture sets.  CPU only."""
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import pfaai_oracle as orc  # noqa: E402
from tests import golden_util as gu  # noqa: E402

NT = 160000


def tetramer_keys(Lc):
    return np.repeat(np.arange(NT, dtype=np.int32), Lc)


@pytest.mark.parametrize("name", ["xanthodb", "xdb_subset1", "xdb_subset2"])
def test_index_build_matches_golden(name):
    g = gu.load(name)
    P, G = g["t"].shape
    tet = tetramer_keys(g["lc"])
    # feed the postings in a scrambled order: the build must restore the reference's F order
    rng = np.random.default_rng(7)
    perm = rng.permutation(tet.shape[0])
    Lc, Lp, F, T = orc.build_index(g["f"]["protein"][perm], tet[perm], g["f"]["genome"][perm], P, G)
    assert np.array_equal(Lc, g["lc"])
    assert np.array_equal(Lp, g["lp"])
    assert np.array_equal(F["protein"], g["f"]["protein"]) and np.array_equal(F["genome"], g["f"]["genome"])
    assert np.array_equal(T, g["t"])  # T == (p,g) histogram of F in single-DB modes


def test_qt_lp_is_prefix_sum_and_t_is_not_histogram():
    g = gu.load("xdb_qt")
    assert np.array_equal(orc.prefix_sum(g["lc"]), g["lp"])
    # query-DB mode: F is inner-joined on tetramer, T is not -> T must come from the host
    assert int(g["t"].sum()) == 123793 and g["f"].shape[0] == 117311


def _check_jac(jac, gold, aji, gold_aji):
    assert np.array_equal(jac["gA"], gold["gA"]) and np.array_equal(jac["gB"], gold["gB"])
    assert np.array_equal(jac["N"], gold["N"])
    assert np.max(np.abs(jac["S"] - gold["S"])) < 1e-7          # JACTuple::operator==, interface.hpp:67-70
    assert np.array_equal(aji, gold_aji)                           # reference tests use exact == on AJI


@pytest.mark.parametrize("name", ["xdb_subset1", "xdb_subset2"])
def test_all_vs_all_small(name):
    g = gu.load(name)
    P, G = g["t"].shape
    r = orc.full_pipeline(orc.ALL_VS_ALL, G, P, g["lc"], g["f"], g["t"])
    assert r["e_size"] == g["sorted_e"].shape[0]
    E = orc.sorted_e(orc.ALL_VS_ALL, G, P, r["counts"])
    for k in ("protein", "gA", "gB"):
        assert np.array_equal(E[k], g["sorted_e"][k])
    _check_jac(r["jac"], g["jac"], r["aji"], g["aji"])


def test_all_vs_all_xantho():
    g = gu.load("xanthodb")
    P, G = g["t"].shape
    r = orc.full_pipeline(orc.ALL_VS_ALL, G, P, g["lc"], g["f"], g["t"])
    assert r["e_size"] == 2608722
    # per-pair block lengths in sorted E (gpe fixtures) == sum_p counts
    assert np.array_equal(r["counts"].sum(axis=1), g["gpe_ends"] - g["gpe_starts"] + 1)
    _check_jac(r["jac"], g["jac"], r["aji"], g["aji"])
    # dense matrix == CSV body
    body = gu.csv_bytes("xanthodb_aji_matrix.csv").decode().strip().split("\n")
    M = np.array([[float(x) for x in line.split(",")] for line in body])
    assert np.array_equal(M, r["matrix"])


def test_query_subset():
    g = gu.load("xanthodb")
    q = gu.load("xdb_qry_subset")
    P, G = g["t"].shape
    qids = gu.query_subset_ids()
    r = orc.full_pipeline(orc.QUERY_SUBSET, G, P, g["lc"], g["f"], g["t"], nQ=len(qids), qry_ids=qids)
    assert r["jac"].shape[0] == 85 and r["e_size"] == 1156769
    _check_jac(r["jac"], q["jac"], r["aji"], q["aji"])
    lines = gu.csv_bytes("qsub_test_output_matrix_wheader.csv").decode().strip().split("\n")
    M = np.array([[float(x) for x in line.split(",")[1:]] for line in lines[1:]])
    assert np.array_equal(M, r["matrix"])


def test_query_db_with_reference_quirk():
    g = gu.load("xdb_qt")
    P, Gu = g["t"].shape
    nQ = 4
    r = orc.full_pipeline(orc.QUERY_DB, Gu, P, g["lc"], g["f"], g["t"], nQ=nQ, quirk=True)
    assert r["e_size"] == g["sorted_e"].shape[0]
    E = orc.sorted_e(orc.QUERY_DB, Gu, P, r["counts"], nQ=nQ)
    for k in ("protein", "gA", "gB"):
        assert np.array_equal(E[k], g["sorted_e"][k])
    _check_jac(r["jac"], g["jac"], r["aji"], g["aji"])
    # without the quirk the values move by < 1e-2 and stay <= 1
    r2 = orc.full_pipeline(orc.QUERY_DB, Gu, P, g["lc"], g["f"], g["t"], nQ=nQ, quirk=False)
    d = np.abs(r2["aji"] - r["aji"]).max()
    assert 0 < d < 1e-2 and r2["aji"].max() <= 1.0 and r["aji"].max() > 1.0


@pytest.mark.parametrize("name", ["xanthodb", "xdb_subset1"])
def test_direct_formulation_agrees(name):
    g = gu.load(name)
    P, G = g["t"].shape
    off, vals = orc.genome_major(P, G, g["lc"], g["f"])
    assert np.array_equal(np.diff(off).reshape(G, P).T, g["t"])
    ids = np.arange(G, dtype=np.int32)
    d = orc.direct_aji(P, G, off, vals, g["t"], ids, ids, want_counts=True)
    r = orc.full_pipeline(orc.ALL_VS_ALL, G, P, g["lc"], g["f"], g["t"])
    iu = np.triu_indices(G, 1)
    assert np.array_equal(d["counts"][iu], r["counts"])
    assert np.array_equal(d["aji"][iu], r["aji"])
    assert np.array_equal(d["aji"], d["aji"].T) and np.all(np.diag(d["aji"]) == 0)
    assert np.array_equal(d["aji"], r["matrix"])
