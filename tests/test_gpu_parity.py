"""GPU parity tests (run on the B200 box: pytest -m gpu).  Everything goes through the C-ABI
(libpfaai_gpu.so via ctypes) and is compared with the reference's goldens and the CPU oracle.

Bars (BASELINE.json north_star): Lc/Lp/T and per-(pair, SCP) intersection counts bit-exact; N exact;
S within 1e-7 (JACTuple::operator==, interface.hpp:67-70); AJI within 1e-6 absolute -- in practice the
kernel sums Jaccards in the reference's order (ascending SCP) with IEEE fp64 division, so the tests
also assert bit-equality of AJI where the reference's own tests do (pfaai_tests.cpp:383,417,450).
"""
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

pytestmark = pytest.mark.gpu

AJI_TOL = 1e-6   # north_star tolerance (absolute)
S_TOL = 1e-7


@pytest.fixture(scope="module")
def pf():
    import __graft_entry__ as ge
    ge.build()
    import parfastaai_b200 as pf
    return pf


@pytest.fixture(scope="module")
def orc():
    from oracle import pfaai_oracle as orc
    return orc


from tests import golden_util as gu  # noqa: E402


def _ctx_for(pf, g, host_T=False):
    P, G = g["t"].shape
    ctx = pf.PfaaiGpu(0)
    ctx.upload_index(P, G, g["f"], g["lc"], g["t"] if host_T else None)
    ctx.build()
    return ctx, P, G


def _check_jac(res, gold_jac, gold_aji):
    assert np.array_equal(res["jac"]["gA"], gold_jac["gA"]) and np.array_equal(res["jac"]["gB"], gold_jac["gB"])
    assert np.array_equal(res["jac"]["N"], gold_jac["N"])
    assert np.abs(res["jac"]["S"] - gold_jac["S"]).max() < S_TOL
    assert np.abs(res["aji"] - gold_aji).max() <= AJI_TOL
    assert np.array_equal(res["aji"], gold_aji), "AJI not bit-identical to the reference golden"


@pytest.mark.parametrize("name", ["xanthodb", "xdb_subset1", "xdb_subset2"])
def test_index_build_golden(pf, orc, name):
    g = gu.load(name)
    ctx, P, G = _ctx_for(pf, g)
    a = ctx.get_arrays()
    assert np.array_equal(a["Lc"], g["lc"])
    assert np.array_equal(a["Lp"], g["lp"])            # *_lp_array.bin
    assert np.array_equal(a["T"], g["t"])              # *_t_matrix.bin, built on device as histogram of F
    off, vals = orc.genome_major(P, G, g["lc"], g["f"])
    assert np.array_equal(a["seg_off"], off)
    assert np.array_equal(a["seg_vals"], vals)          # stable radix sort == stable re-sort of *_f_array.bin
    ctx.close()


@pytest.mark.parametrize("name", ["xanthodb", "xdb_subset1", "xdb_subset2"])
def test_lc_built_on_the_device_from_tetramer_keys(pf, name):
    """constructLc (ds_helper.hpp:82-109) on the device: the (tetramer, protein, genome) triples go in, Lc == *_lc_array.bin
    comes out, and everything downstream (Lp, T, lists, AJI) equals the Lc-fed path; unsorted / out-of-range keys fail loudly."""
    g = gu.load(name)
    P, G = g["t"].shape
    tet = np.repeat(np.arange(pf.NTETRA, dtype=np.int32), g["lc"])
    ctx = pf.PfaaiGpu(0)
    ctx.upload_triples(P, G, tet, g["f"])
    ctx.build()
    a = ctx.get_arrays()
    assert np.array_equal(a["Lc"], g["lc"]) and np.array_equal(a["Lp"], g["lp"]) and np.array_equal(a["T"], g["t"])
    res = ctx.compute(pf.ALL_VS_ALL, want_jac=True)
    assert np.array_equal(res["aji"], g["aji"])
    with pytest.raises(pf.PfaaiError):
        ctx.upload_triples(P, G, tet[::-1].copy(), g["f"])
    bad = tet.copy(); bad[-1] = pf.NTETRA
    with pytest.raises(pf.PfaaiError):
        ctx.upload_triples(P, G, bad, g["f"])
    ctx.close()


@pytest.mark.parametrize("path", ["lists", "mma"])
def test_progress_counters_report_only_finished_rows(pf, path):
    """Streaming output (pfaai_gpu_set_progress / pfaai_gpu_progress_poll): while compute() runs, another thread polls; every
    row reported final is copied at that moment and must equal the finished matrix bit for bit (direct AND mirrored cells);
    the reported prefix never shrinks and ends at G."""
    import threading
    import torch
    from parfastaai_b200 import synth
    N, P = 2600, 40
    off, vals = synth.generate(N, P, seed=11, dense=(path == "mma"))
    Lc, F = synth.to_reference(N, P, off, vals)
    ctx = pf.PfaaiGpu(0)
    ctx.upload_index(P, N, F, Lc)
    ctx.build()
    ctx.set_path(pf.PATH_LISTS if path == "lists" else pf.PATH_MMA)
    want = ctx.compute(pf.ALL_VS_ALL, want_jac=False)["matrix"].copy()
    pin = torch.full((N, N), -7.0, dtype=torch.float64).pin_memory()
    M = pin.numpy()
    ctx.set_progress(True)
    snaps, seen, stop = [], [0], threading.Event()

    def poller():
        while not stop.is_set():
            r, fin = ctx.progress_poll()
            assert r >= seen[-1]
            if r > seen[-1]:
                snaps.append((seen[-1], r, M[seen[-1]:r].copy(), fin))
                seen.append(r)
            if fin:
                break

    th = threading.Thread(target=poller)
    th.start()
    ctx.compute(pf.ALL_VS_ALL, want_jac=False, matrix_out=M)
    th.join(timeout=60)
    stop.set()
    assert not th.is_alive()
    r, fin = ctx.progress_poll()
    assert fin and r == N and seen[-1] == N
    assert np.array_equal(M, want)
    for a, b, rows, _fin in snaps:
        assert np.array_equal(rows, want[a:b]), (a, b)
    assert len(snaps) >= 1
    ctx.set_progress(False)
    assert ctx.progress_poll() == (0, False)
    ctx.close()


def test_device_ready_lists_import_equals_the_built_index(pf, orc):
    """pfaai_gpu_import_lists (the index cache's fast path): the lists exported from a built context, imported into a fresh one
    -- no F, no K1-K3, no sort -- give the same arrays and bit-identical results in all three modes; damaged lists fail loudly."""
    g = gu.load("xanthodb")
    ctx, P, G = _ctx_for(pf, g)
    a = ctx.get_arrays()
    full = ctx.compute(pf.ALL_VS_ALL, want_jac=True, want_counts=True)
    qids = gu.query_subset_ids()
    sub = ctx.compute(pf.QUERY_SUBSET, qry_ids=qids, want_jac=True)
    ctx.close()
    c2 = pf.PfaaiGpu(0)
    c2.import_lists(P, G, a["Lc"], a["seg_off"], a["seg_vals"])          # T derived from the segment lengths
    b = c2.get_arrays()
    for k in ("Lc", "Lp", "T", "seg_off", "seg_vals"):
        assert np.array_equal(a[k], b[k]), k
    for path in (pf.PATH_LISTS, pf.PATH_MMA):
        c2.set_path(path)
        r = c2.compute(pf.ALL_VS_ALL, want_jac=True, want_counts=True)
        assert np.array_equal(r["counts"], full["counts"]) and np.array_equal(r["aji"], full["aji"]) and np.array_equal(r["matrix"], full["matrix"])
    c2.set_path(pf.PATH_AUTO)
    r = c2.compute(pf.QUERY_SUBSET, qry_ids=qids, want_jac=True)
    assert np.array_equal(r["aji"], sub["aji"]) and np.array_equal(r["matrix"], sub["matrix"])
    assert c2.e_size() == int(full["counts"].sum())
    c2.import_lists(P, G, a["Lc"], a["seg_off"], a["seg_vals"], T=a["T"])   # caller-supplied T
    assert np.array_equal(c2.compute(pf.ALL_VS_ALL, want_jac=False)["matrix"], full["matrix"])
    with pytest.raises(pf.PfaaiError):                                    # no F on the device: nothing to rebuild from
        c2.build()
    bad = a["seg_vals"].copy(); bad[[0, 1]] = bad[[1, 0]]
    with pytest.raises(pf.PfaaiError):
        c2.import_lists(P, G, a["Lc"], a["seg_off"], bad)
    bad_off = a["seg_off"].copy(); bad_off[-1] -= 1
    with pytest.raises(pf.PfaaiError):
        c2.import_lists(P, G, a["Lc"], bad_off, a["seg_vals"])
    # a supplied T that is smaller than the list it describes would wrap the Jaccard denominator ta + tb - c: refused up front;
    # a larger one (query-DB mode: T counts tetramers the inner join dropped) is legitimate
    p_, g_ = np.argwhere(a["T"] > 0)[-1]
    small = a["T"].copy(); small[p_, g_] -= 1
    with pytest.raises(pf.PfaaiError, match="smaller than the tetramer list"):
        c2.import_lists(P, G, a["Lc"], a["seg_off"], a["seg_vals"], T=small)
    large = a["T"].copy(); large[p_, g_] += 3
    c2.import_lists(P, G, a["Lc"], a["seg_off"], a["seg_vals"], T=large)
    c2.close()


@pytest.mark.parametrize("name", ["xanthodb", "xdb_subset1", "xdb_subset2"])
def test_all_vs_all_golden(pf, orc, name):
    g = gu.load(name)
    ctx, P, G = _ctx_for(pf, g)
    res = ctx.compute(pf.ALL_VS_ALL, want_jac=True, want_counts=True)
    ref = orc.full_pipeline(orc.ALL_VS_ALL, G, P, g["lc"], g["f"], g["t"])
    assert np.array_equal(res["counts"], ref["counts"])
    if "sorted_e" in g:   # run lengths of the reference's sorted E
        idx = lambda a, b: G * a + b - (a + 2) * (a + 1) // 2
        assert np.array_equal(res["counts"], gu.run_lengths_from_sorted_e(g["sorted_e"], idx, G * (G - 1) // 2, P))
    else:                 # xantho: per-pair block extents of sorted E
        assert np.array_equal(res["counts"].sum(axis=1), g["gpe_ends"] - g["gpe_starts"] + 1)
    _check_jac(res, g["jac"], g["aji"])
    assert np.array_equal(res["matrix"], ref["matrix"])
    assert np.array_equal(res["matrix"], res["matrix"].T) and np.all(np.diag(res["matrix"]) == 0)
    ctx.close()


def test_query_subset_golden(pf, orc):
    g = gu.load("xanthodb")
    q = gu.load("xdb_qry_subset")
    qids = gu.query_subset_ids()
    ctx, P, G = _ctx_for(pf, g)
    res = ctx.compute(pf.QUERY_SUBSET, qry_ids=qids, want_jac=True, want_counts=True)
    ref = orc.full_pipeline(orc.QUERY_SUBSET, G, P, g["lc"], g["f"], g["t"], nQ=len(qids), qry_ids=qids)
    assert np.array_equal(res["counts"], ref["counts"])
    _check_jac(res, q["jac"], q["aji"])
    assert res["matrix"].shape == (5, 20)
    assert np.array_equal(res["matrix"], ref["matrix"])
    lines = gu.csv_bytes("qsub_test_output_matrix_wheader.csv").decode().strip().split("\n")
    M = np.array([[float(x) for x in line.split(",")[1:]] for line in lines[1:]])
    assert np.array_equal(res["matrix"], M)
    ctx.close()


def test_query_db_golden_with_quirk(pf, orc):
    g = gu.load("xdb_qt")
    ctx, P, Gu = _ctx_for(pf, g, host_T=True)
    a = ctx.get_arrays(want_lists=False)
    assert np.array_equal(a["T"], g["t"]) and np.array_equal(a["Lp"], g["lp"])
    nQ = 4
    qids = np.arange(Gu - nQ, Gu, dtype=np.int32)
    res = ctx.compute(pf.QUERY_DB, qry_ids=qids, flags=pf.FLAG_REF_COMPAT_QT, want_jac=True, want_counts=True)
    idx = lambda a, b: (a - (Gu - nQ)) * (Gu - nQ) + b
    assert np.array_equal(res["counts"], gu.run_lengths_from_sorted_e(g["sorted_e"], idx, 16, P))
    _check_jac(res, g["jac"], g["aji"])
    ref = orc.full_pipeline(orc.QUERY_DB, Gu, P, g["lc"], g["f"], g["t"], nQ=nQ, quirk=True)
    assert np.array_equal(res["matrix"], ref["matrix"])
    # corrected labels (flag off) == oracle without the quirk
    res2 = ctx.compute(pf.QUERY_DB, qry_ids=qids, flags=0, want_jac=True)
    ref2 = orc.full_pipeline(orc.QUERY_DB, Gu, P, g["lc"], g["f"], g["t"], nQ=nQ, quirk=False)
    assert np.array_equal(res2["aji"], ref2["aji"]) and res2["aji"].max() <= 1.0
    ctx.close()


@pytest.mark.parametrize("N,P,seed,dense", [(70, 12, 1, False), (200, 9, 2, False), (333, 7, 3, True), (450, 5, 4, False)])
def test_synthetic_all_vs_all_vs_oracle(pf, orc, N, P, seed, dense):
    """Multiple row blocks (192 rows) and column blocks (64 cols), ragged edges, triangular tile skipping."""
    from parfastaai_b200 import synth
    off, vals = synth.generate(N, P, seed=seed, dense=dense)
    Lc, F = synth.to_reference(N, P, off, vals)
    T = synth.t_matrix(N, P, off)
    ctx = pf.PfaaiGpu(0)
    ctx.upload_index(P, N, F, Lc)
    ctx.build()
    a = ctx.get_arrays()
    assert np.array_equal(a["T"], T) and np.array_equal(a["seg_off"], off) and np.array_equal(a["seg_vals"], vals)
    res = ctx.compute(pf.ALL_VS_ALL, want_jac=True, want_counts=True)
    ref = orc.full_pipeline(orc.ALL_VS_ALL, N, P, Lc, F, T)
    assert np.array_equal(res["counts"], ref["counts"])
    assert np.array_equal(res["jac"]["N"], ref["jac"]["N"])
    assert np.array_equal(res["aji"], ref["aji"])
    assert np.array_equal(res["matrix"], ref["matrix"])
    ctx.close()


def test_synthetic_subset_and_query_db_vs_oracle(pf, orc):
    from parfastaai_b200 import synth
    N, P = 260, 8
    off, vals = synth.generate(N, P, seed=11)
    Lc, F = synth.to_reference(N, P, off, vals)
    T = synth.t_matrix(N, P, off)
    ctx = pf.PfaaiGpu(0)
    ctx.upload_index(P, N, F, Lc)
    ctx.build()
    qids = np.array(sorted(np.random.default_rng(5).choice(N, 70, replace=False)), np.int32)
    res = ctx.compute(pf.QUERY_SUBSET, qry_ids=qids, want_jac=True, want_counts=True)
    ref = orc.full_pipeline(orc.QUERY_SUBSET, N, P, Lc, F, T, nQ=len(qids), qry_ids=qids)
    assert np.array_equal(res["counts"], ref["counts"]) and np.array_equal(res["aji"], ref["aji"])
    assert np.array_equal(res["matrix"], ref["matrix"])
    # query-DB split of the same genomes: last 100 are "queries" (nQ != nT: corrected labels only)
    nQ = 100
    q2 = np.arange(N - nQ, N, dtype=np.int32)
    res = ctx.compute(pf.QUERY_DB, qry_ids=q2, flags=0, want_jac=True, want_counts=True)
    ref = orc.full_pipeline(orc.QUERY_DB, N, P, Lc, F, T, nQ=nQ, quirk=False)
    assert np.array_equal(res["counts"], ref["counts"]) and np.array_equal(res["aji"], ref["aji"])
    assert np.array_equal(res["matrix"], ref["matrix"])
    ctx.close()


def _lists_to_arrays(lists, N, P):
    """lists[g][p] = ascending unique tetramers -> genome-major (seg_off, vals)"""
    lens = np.array([len(lists[g][p]) for g in range(N) for p in range(P)], np.int64)
    off = np.concatenate([[0], np.cumsum(lens)]).astype(np.int64)
    vals = np.concatenate([np.asarray(lists[g][p], np.int32) for g in range(N) for p in range(P)] + [np.zeros(0, np.int32)])
    return off, vals.astype(np.int32)


def test_ragged_and_empty_inputs_vs_oracle(pf, orc):
    """Edge cases of the domain: genomes without any posting, SCPs missing from some genomes, a genome that shares no tetramer
    with anyone (N = 0 -> AJI 0), empty lists next to long ones; all three modes against the oracle."""
    from parfastaai_b200 import synth
    rng = np.random.default_rng(5)
    N, P = 90, 6
    lists = [[np.unique(rng.integers(0, 3000, rng.integers(0, 400))).astype(np.int32) for p in range(P)] for g in range(N)]
    for g in (3, 17):
        lists[g] = [np.zeros(0, np.int32)] * P
    for g in range(0, N, 5):
        lists[g][2] = np.zeros(0, np.int32)
    lists[40] = [np.arange(100000 + 10 * p, 100000 + 10 * p + 5, dtype=np.int32) for p in range(P)]
    off, vals = _lists_to_arrays(lists, N, P)
    Lc, F = synth.to_reference(N, P, off, vals)
    T = synth.t_matrix(N, P, off)
    ctx = pf.PfaaiGpu(0)
    ctx.upload_index(P, N, F, Lc)
    ctx.build()
    a = ctx.get_arrays()
    assert np.array_equal(a["T"], T) and np.array_equal(a["seg_off"], off) and np.array_equal(a["seg_vals"], vals)
    res = ctx.compute(pf.ALL_VS_ALL, want_jac=True, want_counts=True)
    ref = orc.full_pipeline(orc.ALL_VS_ALL, N, P, Lc, F, T)
    assert (ref["jac"]["N"] == 0).sum() >= 89 + 88          # the isolated genome and the empty ones
    assert np.array_equal(res["counts"], ref["counts"]) and np.array_equal(res["jac"]["N"], ref["jac"]["N"])
    assert np.array_equal(res["aji"], ref["aji"]) and np.array_equal(res["matrix"], ref["matrix"])
    # isolated, empty and ragged genomes as queries; in database order (the reference's genomePairToIndex, ds_impl.hpp:251-263,
    # is only defined for query lists whose order agrees with the genome ids)
    qids = np.array([0, 3, 5, 40, 89], np.int32)
    res = ctx.compute(pf.QUERY_SUBSET, qry_ids=qids, want_jac=True, want_counts=True)
    ref = orc.full_pipeline(orc.QUERY_SUBSET, N, P, Lc, F, T, nQ=len(qids), qry_ids=qids)
    assert np.array_equal(res["counts"], ref["counts"]) and np.array_equal(res["matrix"], ref["matrix"])
    ctx.close()


def test_empty_index_and_bad_input_fail_or_zero(pf, orc):
    """|F| = 0 gives an all-zero matrix; a posting with an id out of range or sum(Lc) != |F| fails the build loudly."""
    from parfastaai_b200 import synth
    N, P = 10, 3
    ctx = pf.PfaaiGpu(0)
    ctx.upload_index(P, N, np.zeros(0, pf.PG_DTYPE), np.zeros(pf.NTETRA, np.int32))
    ctx.build()
    res = ctx.compute(pf.ALL_VS_ALL, want_jac=True)
    assert not res["matrix"].any() and not res["jac"]["N"].any() and not res["aji"].any()
    off, vals = synth.generate(N, P, seed=2)
    Lc, F = synth.to_reference(N, P, off, vals)
    bad = F.copy(); bad["genome"][7] = N
    ctx.upload_index(P, N, bad, Lc)
    with pytest.raises(pf.PfaaiError, match="out of range"):
        ctx.build()
    bad = F.copy(); bad["protein"][11] = -1
    ctx.upload_index(P, N, bad, Lc)
    with pytest.raises(pf.PfaaiError, match="out of range"):
        ctx.build()
    Lc2 = Lc.copy(); Lc2[int(np.flatnonzero(Lc)[0])] += 1
    ctx.upload_index(P, N, F, Lc2)
    with pytest.raises(pf.PfaaiError, match="sum"):
        ctx.build()
    with pytest.raises(pf.PfaaiError):                       # nothing built: compute refuses
        ctx.compute(pf.ALL_VS_ALL, want_jac=False)
    ctx.upload_index(P, N, F, Lc)                            # and the context recovers
    ctx.build()
    ref = orc.full_pipeline(orc.ALL_VS_ALL, N, P, Lc, F, synth.t_matrix(N, P, off))
    assert np.array_equal(ctx.compute(pf.ALL_VS_ALL, want_jac=False)["matrix"], ref["matrix"])
    ctx.close()


def test_long_lists_and_many_scps_take_the_wide_kernel_variants(pf, orc):
    """Maximum sizes: lists beyond 1022 tetramers select the 12-plane counters, more than 127 SCPs the 10-plane N counter
    (aji_tile_kernel<12,7> / <9,10>); both against the oracle, counts included."""
    from parfastaai_b200 import synth
    rng = np.random.default_rng(9)
    N, P = 70, 2                                             # lists of 1100-2500 tetramers sharing a common core
    core = [np.unique(rng.integers(0, 160000, 2400)) for p in range(P)]
    lists = [[np.unique(np.concatenate([core[p][rng.random(core[p].shape[0]) < 0.8], rng.integers(0, 160000, 300)])).astype(np.int32)
              for p in range(P)] for g in range(N)]
    assert max(len(x) for l in lists for x in l) > 1100
    off, vals = _lists_to_arrays(lists, N, P)
    Lc, F = synth.to_reference(N, P, off, vals)
    T = synth.t_matrix(N, P, off)
    ctx = pf.PfaaiGpu(0)
    ctx.upload_index(P, N, F, Lc)
    ctx.build()
    res = ctx.compute(pf.ALL_VS_ALL, want_jac=True, want_counts=True)
    ref = orc.full_pipeline(orc.ALL_VS_ALL, N, P, Lc, F, T)
    assert ref["counts"].max() > 1022
    assert np.array_equal(res["counts"], ref["counts"]) and np.array_equal(res["aji"], ref["aji"])
    assert np.array_equal(res["matrix"], ref["matrix"])
    N, P = 66, 140                                           # > 127 SCPs
    off, vals = synth.generate(N, P, seed=14)
    Lc, F = synth.to_reference(N, P, off, vals)
    T = synth.t_matrix(N, P, off)
    ctx.upload_index(P, N, F, Lc)
    ctx.build()
    res = ctx.compute(pf.ALL_VS_ALL, want_jac=True)
    ref = orc.full_pipeline(orc.ALL_VS_ALL, N, P, Lc, F, T)
    assert ref["jac"]["N"].max() > 127
    assert np.array_equal(res["jac"]["N"], ref["jac"]["N"]) and np.array_equal(res["aji"], ref["aji"])
    assert np.array_equal(res["matrix"], ref["matrix"])
    ctx.close()


@pytest.mark.parametrize("n_shards", [2, 3])
def test_sharded_row_blocks_assemble_to_the_full_matrix(pf, orc, n_shards):
    """Row-block sharding: every shard stores only its own blocks of 384 target genomes, compactly; the host
    assembly (mirror in all-vs-all) reproduces the unsharded matrix bit for bit; every pair is computed exactly once."""
    from parfastaai_b200 import sharding, synth
    N, P = 1000, 3
    off, vals = synth.generate(N, P, seed=21)
    Lc, F = synth.to_reference(N, P, off, vals)
    ctx = pf.PfaaiGpu(0)
    ctx.upload_index(P, N, F, Lc)
    ctx.build()
    full = ctx.compute(pf.ALL_VS_ALL, want_jac=False)["matrix"].copy()
    axis, ids = ctx.shard_layout()
    assert axis == 0 and np.array_equal(ids, np.arange(N))
    qids = np.array([5, 17, 400, 401, 999], np.int32)
    full_q = ctx.compute(pf.QUERY_SUBSET, qry_ids=qids, want_jac=False)["matrix"].copy()
    parts, lay, parts_q, lay_q, pairs = [], [], [], [], 0
    for s in range(n_shards):
        ctx.set_shard(s, n_shards)
        part = ctx.compute(pf.ALL_VS_ALL, want_jac=False)["matrix"]
        axis, ids = ctx.shard_layout()
        assert axis == 0 and np.array_equal(ids, sharding.owned_indices(N, s, n_shards))
        assert part.shape == sharding.shard_shape(pf.ALL_VS_ALL, N, 0, s, n_shards)
        # only cells with row < column are computed
        assert all(np.all(part[k, :g + 1] == 0) for k, g in enumerate(ids))
        pairs += sharding.pair_count(pf.ALL_VS_ALL, N, 0, s, n_shards)
        parts.append(part.copy()); lay.append(ids)
        pq = ctx.compute(pf.QUERY_SUBSET, qry_ids=qids, want_jac=False)["matrix"]
        axis, idq = ctx.shard_layout()
        assert axis == 1 and pq.shape == (5, idq.shape[0])
        parts_q.append(pq.copy()); lay_q.append(idq)
        with pytest.raises(pf.PfaaiError):
            ctx.compute(pf.ALL_VS_ALL, want_jac=True)
    ctx.set_shard(0, 1)
    assert pairs == N * (N - 1) // 2
    assert np.array_equal(sharding.assemble(pf.ALL_VS_ALL, parts, lay, np.empty((N, N))), full)
    assert np.array_equal(sharding.assemble(pf.QUERY_SUBSET, parts_q, lay_q, np.empty((5, N))), full_q)
    ctx.close()


@pytest.mark.parametrize("n_shards,N,P", [(2, 1000, 3), (3, 1700, 2), (4, 900, 4)])
def test_partitioned_build_matches_the_replicated_build(pf, orc, n_shards, N, P):
    """pfaai_gpu_build_partitioned: every shard sorts only the postings of its own row blocks and builds only their
    dictionaries; after the exchange (here: peer copies between contexts of one process) every shard's matrix must equal the
    replicated-build shard matrix bit for bit, and the assembly the unsharded matrix.  N = 900 / 4 shards leaves a shard
    without any row block (3 blocks of 384)."""
    from parfastaai_b200 import sharding, synth
    off, vals = synth.generate(N, P, seed=77)
    Lc, F = synth.to_reference(N, P, off, vals)
    T = synth.t_matrix(N, P, off)
    ref = pf.PfaaiGpu(0)
    ref.upload_index(P, N, F, Lc)
    ref.build()
    full = ref.compute(pf.ALL_VS_ALL, want_jac=False)["matrix"].copy()
    ctxs = []
    for s in range(n_shards):
        c = pf.PfaaiGpu(0)
        c.set_shard(s, n_shards)
        c.upload_index(P, N, F, Lc)
        c.build_partitioned()
        with pytest.raises(pf.PfaaiError):          # not exchanged yet
            c.compute(pf.ALL_VS_ALL, want_jac=False)
        ctxs.append(c)
    x = [c.exchange_buffers() for c in ctxs]
    assert all([b[1] for b in xs] == [b[1] for b in x[0]] for xs in x)        # same chunk layout on every shard
    assert all(b[2] <= b[1] for xs in x for b in xs)
    pf.PfaaiGpu.exchange_local(ctxs)
    parts, lay = [], []
    for s, c in enumerate(ctxs):
        ids = sharding.owned_indices(N, s, n_shards)
        arr = c.get_arrays(want_lists=False)                                  # T: this shard's genomes only
        assert np.array_equal(arr["T"][:, ids], T[:, ids])
        assert not arr["T"][:, np.setdiff1d(np.arange(N), ids)].any()
        part = c.compute(pf.ALL_VS_ALL, want_jac=False)["matrix"]
        ref.set_shard(s, n_shards)
        assert np.array_equal(part, ref.compute(pf.ALL_VS_ALL, want_jac=False)["matrix"])
        with pytest.raises(pf.PfaaiError):          # a partitioned index serves all-vs-all only
            c.compute(pf.QUERY_SUBSET, qry_ids=np.array([1, 2], np.int32), want_jac=False)
        parts.append(part.copy()); lay.append(ids)
    assert np.array_equal(sharding.assemble(pf.ALL_VS_ALL, parts, lay, np.empty((N, N))), full)
    # a second partitioned build on the same contexts (steady state: no reallocation) gives the same result
    for c in ctxs:
        c.build_partitioned()
    pf.PfaaiGpu.exchange_local(ctxs)
    import torch
    for s, c in enumerate(ctxs):
        assert np.array_equal(c.compute(pf.ALL_VS_ALL, want_jac=False)["matrix"], parts[s])
        # the overlapped variant: own-dictionary tiles first, the rest behind the caller's "exchange done" event
        ev = torch.cuda.Event()
        ev.record()
        c.finish_partitioned_async(ev.cuda_event)
        assert np.array_equal(c.compute(pf.ALL_VS_ALL, want_jac=False)["matrix"], parts[s])
        c.build()                                                              # back to the replicated build
        assert np.array_equal(c.compute(pf.ALL_VS_ALL, want_jac=False)["matrix"], parts[s])
        c.close()
    ref.close()


@pytest.mark.parametrize("n_shards,N,P,cuts", [(2, 1000, 3, None), (3, 1700, 2, [0.1, 0.75]), (4, 900, 4, None)])
def test_routed_build_matches_the_replicated_build(pf, orc, n_shards, N, P, cuts):
    """Routed build: every shard is given a slice of F only; the items are grouped by owner, moved (peer copies here, an NCCL
    all-to-all in bench.py) and sorted by the shard that owns the genome.  The genome-major lists, T and every shard matrix must
    equal the replicated build's, for equal and for ragged slices (incl. an empty shard: 900 genomes / 4 shards)."""
    from parfastaai_b200 import sharding, synth
    off, vals = synth.generate(N, P, seed=91)
    Lc, F = synth.to_reference(N, P, off, vals)
    T = synth.t_matrix(N, P, off)
    nF = F.shape[0]
    bounds = [0] + ([int(c * nF) for c in cuts] if cuts else [nF * (s + 1) // n_shards for s in range(n_shards - 1)]) + [nF]
    ref = pf.PfaaiGpu(0)
    ref.upload_index(P, N, F, Lc)
    ref.build()
    ra = ref.get_arrays()
    ctxs = []
    for s in range(n_shards):
        c = pf.PfaaiGpu(0)
        c.set_shard(s, n_shards)
        c.upload_index_slice(P, N, F[bounds[s]:bounds[s + 1]], Lc, nF, bounds[s])
        with pytest.raises(pf.PfaaiError):          # a slice cannot be built without routing
            c.build()
        ctxs.append(c)
    _, offs = ctxs[0].route_prepare()
    assert offs[0] == 0 and offs[-1] == bounds[1] - bounds[0] and all(a <= b for a, b in zip(offs, offs[1:]))
    pf.PfaaiGpu.route_local(ctxs)
    pf.PfaaiGpu.exchange_local(ctxs)
    parts, lay = [], []
    for s, c in enumerate(ctxs):
        ids = sharding.owned_indices(N, s, n_shards)
        arr = c.get_arrays()
        assert np.array_equal(arr["Lp"], ra["Lp"])
        assert np.array_equal(arr["T"][:, ids], T[:, ids]) and not arr["T"][:, np.setdiff1d(np.arange(N), ids)].any()
        # the shard's genome-major lists: same segments, same tetramers in the same order as the replicated sort
        so, sv = arr["seg_off"], arr["seg_vals"]
        for g in ids[:: max(1, len(ids) // 40)]:
            for p in range(P):
                k = g * P + p
                assert np.array_equal(sv[so[k]:so[k + 1]], ra["seg_vals"][ra["seg_off"][k]:ra["seg_off"][k + 1]])
        assert so[-1] == sum(int(ra["seg_off"][(g + 1) * P] - ra["seg_off"][g * P]) for g in ids)
        part = c.compute(pf.ALL_VS_ALL, want_jac=False)["matrix"]
        ref.set_shard(s, n_shards)
        assert np.array_equal(part, ref.compute(pf.ALL_VS_ALL, want_jac=False)["matrix"])
        parts.append(part.copy()); lay.append(ids)
    ref.set_shard(0, 1)
    full = ref.compute(pf.ALL_VS_ALL, want_jac=False)["matrix"]
    assert np.array_equal(sharding.assemble(pf.ALL_VS_ALL, parts, lay, np.empty((N, N))), full)
    for c in ctxs:
        c.close()
    ref.close()


def test_pinned_host_matrix_is_stored_directly_and_identically(pf, orc):
    """pfaai_gpu_compute with a page-locked matrix: the tile kernel stores straight into host memory (no device copy
    of the matrix, no D2H call); the bytes must equal the staged path's in all three modes, ragged tile edges included."""
    import torch
    from parfastaai_b200 import synth
    N, P = 700, 5
    off, vals = synth.generate(N, P, seed=33)
    Lc, F = synth.to_reference(N, P, off, vals)
    ctx = pf.PfaaiGpu(0)
    ctx.upload_index(P, N, F, Lc)
    ctx.build()
    qids = np.array([699, 3, 384, 64, 63], np.int32)
    for mode, q, shape in [(pf.ALL_VS_ALL, None, (N, N)), (pf.QUERY_SUBSET, qids, (5, N)),
                           (pf.QUERY_DB, np.arange(N - 90, N, dtype=np.int32), (90, N - 90))]:
        staged = ctx.compute(mode, qry_ids=q, want_jac=False)["matrix"].copy()
        assert ctx.matrix_dev_ptr()
        pin = torch.full(shape, -7.0, dtype=torch.float64).pin_memory()   # poisoned: every cell must be written
        direct = ctx.compute(mode, qry_ids=q, want_jac=False, matrix_out=pin.numpy())["matrix"]
        assert not ctx.matrix_dev_ptr()
        assert np.array_equal(direct, staged)
        with pytest.raises(pf.PfaaiError):
            ctx.download_matrix()
    ctx.close()


@pytest.mark.gpu
@pytest.mark.parametrize("N", [385, 769])
def test_pinned_matrix_last_diagonal_cell_is_written(pf, orc, N):
    """G = 384k + 1: the last row block's only tile is (row0 == col_max), i.e. the diagonal cell alone.  With a pinned
    (directly stored, never memset) matrix every cell -- (G-1, G-1) included -- must be written (pfaai_gpu.cu:905)."""
    import torch
    from parfastaai_b200 import synth
    P = 3
    off, vals = synth.generate(N, P, seed=5)
    Lc, F = synth.to_reference(N, P, off, vals)
    ctx = pf.PfaaiGpu(0)
    ctx.upload_index(P, N, F, Lc)
    ctx.build()
    staged = ctx.compute(pf.ALL_VS_ALL, want_jac=False)["matrix"].copy()
    pin = torch.full((N, N), -7.0, dtype=torch.float64).pin_memory()
    direct = ctx.compute(pf.ALL_VS_ALL, want_jac=False, matrix_out=pin.numpy())["matrix"]
    assert direct[N - 1, N - 1] == 0.0
    assert not np.any(direct == -7.0)
    assert np.array_equal(direct, staged)
    assert np.array_equal(direct, direct.T) and not np.any(np.diag(direct))
    ctx.close()


@pytest.mark.gpu
def test_pinned_matrix_of_an_empty_index_is_zeroed(pf):
    """|F| = 0: no tile kernel is launched; a pinned matrix must still come back as zeros, not untouched."""
    import torch
    ctx = pf.PfaaiGpu(0)
    G, P = 9, 2
    ctx.upload_index(P, G, np.zeros(0, pf.PG_DTYPE), np.zeros(pf.NTETRA, np.int32))
    ctx.build()
    pin = torch.full((G, G), -7.0, dtype=torch.float64).pin_memory()
    direct = ctx.compute(pf.ALL_VS_ALL, want_jac=False, matrix_out=pin.numpy())["matrix"]
    assert not np.any(direct)
    ctx.close()


@pytest.mark.parametrize("n_shards,N", [(2, 1000), (3, 1153), (4, 900)])
def test_global_matrix_shards_fill_one_pinned_matrix_completely(pf, orc, n_shards, N):
    """pfaai_gpu_set_global_matrix: all shards store their cells and the mirror images into ONE pinned G x G matrix; their
    cell sets are disjoint and cover the matrix (poisoned buffer), the result equals the unsharded matrix bit for bit
    (N = 1153 = 3 * 384 + 1: the last row block is the diagonal cell alone; N = 900 / 4 shards: one shard owns nothing)."""
    import torch
    from parfastaai_b200 import synth
    P = 3
    off, vals = synth.generate(N, P, seed=91)
    Lc, F = synth.to_reference(N, P, off, vals)
    ctx = pf.PfaaiGpu(0)
    ctx.upload_index(P, N, F, Lc)
    ctx.build()
    full = ctx.compute(pf.ALL_VS_ALL, want_jac=False)["matrix"].copy()
    pin = torch.full((N, N), -7.0, dtype=torch.float64).pin_memory()
    M = pin.numpy()
    ctx.set_global_matrix(True)
    written = np.zeros((N, N), bool)
    for s in range(n_shards):
        ctx.set_shard(s, n_shards)
        before = M.copy()
        ctx.compute(pf.ALL_VS_ALL, want_jac=False, matrix_out=M)
        assert ctx.matrix_shape() == (N, N)
        now = (M != before)
        assert not np.any(now & written), "two shards wrote the same cell"
        written |= now
    assert not np.any(M == -7.0)
    assert np.array_equal(M, full)
    # the staged (device matrix) variant of the same addressing: zeros outside the shard's cells
    ctx.set_shard(0, n_shards)
    dev = ctx.compute(pf.ALL_VS_ALL, want_jac=False)["matrix"]
    own = np.zeros(N, bool); own[ctx.shard_layout()[1]] = True
    iu = np.triu(np.ones((N, N), bool), 1)
    mine = (iu & own[:, None]) | (iu & own[:, None]).T
    assert np.array_equal(dev[mine], full[mine]) and not np.any(dev[~mine])
    ctx.set_global_matrix(False)
    ctx.set_shard(0, 1)
    ctx.close()
