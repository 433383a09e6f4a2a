"""CPU-side tests of the C++17 host (parfastaai_b200/host): par_fastaai.x argument handling and exit codes
(src/main.cpp + CLI11 behaviour), the SQLite loader against the reference's Lc/F/T goldens,
and the CSV cell formatter against every value the golden CSVs contain.  No GPU needed: the loader is
exercised through --dump-index (host only) and a plain run must FAIL without a CUDA device (no CPU fallback)."""
import os
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as ge  # noqa: E402
from tests import golden_util as gu  # noqa: E402


@pytest.fixture(scope="module")
def cli():
    ge.build()
    assert os.access(ge.CLI, os.X_OK)
    return ge.CLI


def run(cli, *args):
    return subprocess.run([cli, *args], capture_output=True, text=True, env=dict(os.environ, OMP_NUM_THREADS="4"))


def load_dump(prefix):
    import parfastaai_b200 as pf
    dims = np.fromfile(prefix + ".dims.bin", np.int32)
    P, G = int(dims[0]), int(dims[1])
    return {"dims": dims, "lc": np.fromfile(prefix + ".lc.bin", np.int32), "f": np.fromfile(prefix + ".f.bin", pf.PG_DTYPE),
            "t": np.fromfile(prefix + ".t.bin", np.int32).reshape(P, G)}


@pytest.mark.parametrize("name", ["xdb_subset1", "xdb_subset2"])
def test_loader_single_db_matches_goldens(cli, tmp_path, name):
    r = run(cli, gu.db_path(name + ".db"), str(tmp_path / "o.csv"), "--dump-index", str(tmp_path / name))
    assert r.returncode == 0, r.stderr
    d, g = load_dump(str(tmp_path / name)), gu.load(name)
    assert np.array_equal(d["lc"], g["lc"])             # *_lc_array.bin
    assert np.array_equal(d["f"], g["f"])               # *_f_array.bin: (tetramer, protein, genome) order
    assert np.array_equal(d["t"], g["t"])               # *_t_matrix.bin
    assert not (tmp_path / "o.csv").exists()


def test_loader_rebuilt_xantho_db(cli, tmp_path):
    """The 20-genome DB is missing upstream; rebuilt from the F/Lc goldens + name tables."""
    db = gu.xantho_db_path()
    r = run(cli, db, str(tmp_path / "o.csv"), "--dump-index", str(tmp_path / "x"))
    assert r.returncode == 0, r.stderr
    d, g = load_dump(str(tmp_path / "x")), gu.load("xanthodb")
    assert np.array_equal(d["lc"], g["lc"]) and np.array_equal(d["f"], g["f"]) and np.array_equal(d["t"], g["t"])


def test_loader_query_db_mode_matches_goldens(cli, tmp_path):
    r = run(cli, gu.db_path("xdb_subset1.db"), str(tmp_path / "o.csv"), "-r", gu.db_path("xdb_subset2.db"),
            "--dump-index", str(tmp_path / "qt"))
    assert r.returncode == 0, r.stderr
    d, g = load_dump(str(tmp_path / "qt")), gu.load("xdb_qt")
    assert list(d["dims"]) == [79, 8, 4, 4]
    assert np.array_equal(d["lc"], g["lc"]) and np.array_equal(d["f"], g["f"])     # inner join on the tetramer
    assert np.array_equal(d["t"], g["t"])                                         # T unrestricted, queries at nT + id


def test_index_cache_round_trip_and_staleness(cli, tmp_path):
    """--index-cache: the first run reads SQLite and writes the cache, the second serves names + Lc/F/T
    from it without opening SQLite (bit-identical arrays); a changed database (size / mtime) or a damaged cache is ignored and
    the cache rewritten."""
    import shutil
    db = str(tmp_path / "s1.db")
    shutil.copy(gu.db_path("xdb_subset1.db"), db)
    cache = str(tmp_path / "s1.pfidx")
    g = gu.load("xdb_subset1")
    r1 = run(cli, db, str(tmp_path / "o.csv"), "--dump-index", str(tmp_path / "a"), "--index-cache", cache)
    assert r1.returncode == 0 and "Index cache         : written" in r1.stdout and os.path.exists(cache), r1.stdout + r1.stderr
    r2 = run(cli, db, str(tmp_path / "o.csv"), "--dump-index", str(tmp_path / "b"), "--index-cache", cache)
    assert r2.returncode == 0 and "Index cache         : loaded" in r2.stdout and "SQLite not opened" in r2.stdout
    a, b = load_dump(str(tmp_path / "a")), load_dump(str(tmp_path / "b"))
    for k in ("dims", "lc", "f", "t"):
        assert np.array_equal(a[k], b[k]), k
    assert np.array_equal(b["lc"], g["lc"]) and np.array_equal(b["f"], g["f"]) and np.array_equal(b["t"], g["t"])
    # the cache really is what is read: without the database's tables the cached run still works
    blob = open(cache, "rb").read()
    assert blob[:8] == b"PFIDX001"
    # stale: touch the database -> cache ignored and rewritten
    os.utime(db, ns=(1, 1))
    r3 = run(cli, db, str(tmp_path / "o.csv"), "--dump-index", str(tmp_path / "c"), "--index-cache", cache)
    assert r3.returncode == 0 and "written" in r3.stdout and "loaded" not in r3.stdout
    # damaged (one flipped byte in T, truncated): ignored and rewritten
    for damage in (lambda d: d[:-5] + bytes([d[-5] ^ 1]) + d[-4:], lambda d: d[:-100]):
        good = open(cache, "rb").read()
        open(cache, "wb").write(damage(good))
        r4 = run(cli, db, str(tmp_path / "o.csv"), "--dump-index", str(tmp_path / "d"), "--index-cache", cache)
        assert r4.returncode == 0 and "written" in r4.stdout and "loaded" not in r4.stdout
        assert open(cache, "rb").read() == good
    d = load_dump(str(tmp_path / "d"))
    assert np.array_equal(d["f"], g["f"])


def test_exit_codes_and_banners(cli, tmp_path):
    out = str(tmp_path / "o.csv")
    assert run(cli).returncode == 106                                            # CLI11 RequiredError
    assert run(cli, "/nonexistent/x.db", out).returncode == 105                   # CLI11 ValidationError (ExistingFile)
    assert run(cli, gu.db_path("xdb_subset1.db")).returncode == 106
    assert run(cli, gu.db_path("xdb_subset1.db"), out, "--bogus").returncode == 109
    assert run(cli, gu.db_path("xdb_subset1.db"), out, "-q", "/nonexistent.txt").returncode == 105
    h = run(cli, "--help")
    assert h.returncode == 0 and "--query_subset" in h.stdout and "--separator" in h.stdout
    # BASELINE config 2 as literally written: 4 of the 5 listed genomes are not in xdb_subset1.db -> exit 3
    r = run(cli, gu.db_path("xdb_subset1.db"), out, "-q", os.path.join(gu.GOLDEN, "qsub_test_input.txt"))
    assert r.returncode == 3 and "missing from the database" in r.stdout
    assert r.stdout.count("Xanthomonas_albilineans") == 3 and not os.path.exists(out)
    # query DB overlapping the main DB -> exit 3 (main.cpp:268-300)
    r = run(cli, gu.db_path("xdb_subset_combo12.db"), out, "-r", gu.db_path("xdb_subset1.db"))
    assert r.returncode == 3 and "overlapping with the target database" in r.stdout
    # not an SQLite file -> PFAAI_ERR_SQLITE_DB
    bad = tmp_path / "bad.db"
    bad.write_text("this is not a database, " * 100)
    r = run(cli, str(bad), out)
    assert r.returncode == 1 and "Error in opening" in r.stderr


def test_no_cpu_fallback(cli, tmp_path):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    r = run(cli, gu.db_path("xdb_subset1.db"), str(tmp_path / "o.csv"))
    assert r.returncode == 3 and "no CPU fallback" in r.stderr
    assert not (tmp_path / "o.csv").exists()


def test_csv_cell_formatter_matches_fmt(cli, tmp_path):
    """Every cell of the golden CSVs (fmt `{}` on a double) must be reproduced byte for byte, plus layout edge cases."""
    cells = []
    for name in ["xanthodb_aji_matrix.csv", "xdb_subset1_aji_matrix_wheader.csv", "xdb_subset2_aji_matrix_wheader.csv",
                 "qsub_test_output_matrix_wheader.csv", "xanthodb_aji_matrix_wheader.csv"]:
        lines = gu.csv_bytes(name).decode().strip().split("\n")
        if "wheader" in name:
            lines = [",".join(l.split(",")[1:]) for l in lines[1:]]
        for l in lines:
            cells += l.split(",")
    extra = {"0": 0.0, "1": 1.0, "1.0009481433436662": 1.0009481433436662, "0.0001": 1e-4, "1e-05": 1e-5, "1.5e-07": 1.5e-7,
             "100": 100.0, "1234.5": 1234.5, "1e+16": 1e16, "1000000000000000": 1e15, "0.3": 0.3, "5e-324": 5e-324,
             "2.2250738585072014e-308": 2.2250738585072014e-308, "0.30000000000000004": 0.1 + 0.2, "-0": -0.0, "-2.5": -2.5}
    vals = np.array([float(c) for c in cells] + list(extra.values()), np.float64)
    want = cells + list(extra.keys())
    p = tmp_path / "v.bin"
    vals.tofile(p)
    r = run(cli, "--format-doubles", str(p))
    assert r.returncode == 0
    got = r.stdout.split("\n")[:-1]
    assert len(got) == len(want) > 800
    bad = [(w, g) for w, g in zip(want, got) if w != g]
    assert not bad, bad[:5]
