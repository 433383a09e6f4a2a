"""GPU end-to-end tests of par_fastaai.x (C++17 host -> C-ABI -> sm_100a kernels -> CSV): output BYTES against the
reference's golden CSVs for the three modes, and against the unmodified reference binary (oracle/_ref, when it
travelled to the box) on a synthetic FastAAI database."""
import os
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as ge  # noqa: E402
from tests import golden_util as gu  # noqa: E402

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def cli():
    ge.build()
    return ge.CLI


def run(cli, *args, threads="4"):
    r = subprocess.run([cli, *args], capture_output=True, text=True, env=dict(os.environ, OMP_NUM_THREADS=threads))
    assert r.returncode == 0, (r.returncode, r.stdout[-500:], r.stderr[-500:])
    return r


def test_all_vs_all_csv_bytes(cli, tmp_path):
    for db, gold in [(gu.db_path("xdb_subset1.db"), "xdb_subset1_aji_matrix_wheader.csv"),
                     (gu.db_path("xdb_subset2.db"), "xdb_subset2_aji_matrix_wheader.csv"),
                     (gu.xantho_db_path(), "xanthodb_aji_matrix_wheader.csv")]:
        out = tmp_path / ("out_" + gold)
        r = run(cli, db, str(out))
        assert out.read_bytes() == gu.csv_bytes(gold), gold
        assert "Writing output with" in r.stdout
        # the reference's phase lines, prefix and number format (interface.hpp:307-325, algorithm_impl.hpp:304, utils.hpp:136-200)
        import re
        for prefix in ("Lc & Lp contruction : ", "F contruction       : ", "T construction      : ", " E construction      : ",
                       "E parallel sorting  : ", "E constr.   (fin)   : ",
                       "JAC Construction    : "):
            assert re.search(r"^" + re.escape(prefix) + r" *[0-9.e+-]+ ms \( *[0-9.e+-]+ sec/ *[0-9.e+-]+ min\);  *\d+ KB \( *\d+ MB/ *\d+ GB\)\.$",
                             r.stdout, re.M), (prefix, r.stdout)


def test_query_subset_csv_bytes(cli, tmp_path):
    out = tmp_path / "qsub.csv"
    run(cli, gu.xantho_db_path(), str(out), "-q", os.path.join(gu.GOLDEN, "qsub_test_input.txt"))
    assert out.read_bytes() == gu.csv_bytes("qsub_test_output_matrix_wheader.csv")


def test_separator_and_long_options(cli, tmp_path):
    out = tmp_path / "semi.csv"
    run(cli, gu.db_path("xdb_subset1.db"), str(out), "--separator=;;")
    assert out.read_bytes() == gu.csv_bytes("xdb_subset1_aji_matrix_wheader.csv").replace(b",", b";;")


def test_query_db_mode_reference_values(cli, tmp_path):
    """No CSV golden is shipped for -r; the reference's own output on these two DBs (reproducing
    xdb_qt_aji.bin incl. the 1.0009 cell of the label quirk) is the expectation."""
    out = tmp_path / "qt.csv"
    run(cli, gu.db_path("xdb_subset1.db"), str(out), "-r", gu.db_path("xdb_subset2.db"))
    lines = out.read_text().strip().split("\n")
    n = gu.names()
    assert lines[0] == "," + ",".join(n["subset1_genomes"])
    assert [l.split(",")[0] for l in lines[1:]] == n["subset2_genomes"]
    M = np.array([[float(x) for x in l.split(",")[1:]] for l in lines[1:]])
    g = gu.load("xdb_qt")
    assert np.array_equal(M.reshape(-1), g["aji"])                                  # bit-exact vs xdb_qt_aji.bin
    assert lines[2].split(",")[4] == "1.0009481433436662"
    # corrected lookup: every value <= 1 and within 1e-2 of the quirk values
    out2 = tmp_path / "qt2.csv"
    run(cli, gu.db_path("xdb_subset1.db"), str(out2), "-r", gu.db_path("xdb_subset2.db"), "--no-ref-compat")
    M2 = np.array([[float(x) for x in l.split(",")[1:]] for l in out2.read_text().strip().split("\n")[1:]])
    assert M2.max() <= 1.0 and 0 < np.abs(M2 - M).max() < 1e-2


def test_query_db_mode_on_the_cached_main_index(cli, tmp_path):
    """The incremental "add query genomes" path: once a single-DB run has left `.pfidx` + `.lists` of the
    main database, `-r query.db --index-cache` reads only the query database, appends its genomes' lists to the cached
    device-ready index and must write the same CSV bytes as the uncached run (incl. the label quirk and --no-ref-compat);
    two contexts (replicated import, sharded compute) too."""
    import shutil
    main_db = str(tmp_path / "main.db")
    shutil.copy(gu.db_path("xdb_subset1.db"), main_db)
    qry_db = gu.db_path("xdb_subset2.db")
    cache = str(tmp_path / "main.pfidx")
    run(cli, main_db, str(tmp_path / "all.csv"), "--index-cache", cache)            # first run: writes the cache and the sidecar
    assert os.path.exists(cache) and os.path.exists(cache + ".lists")
    for extra in ([], ["--no-ref-compat"], ["--gpus", "2", "--same-device"]):
        plain, inc = str(tmp_path / "plain.csv"), str(tmp_path / "inc.csv")
        run(cli, main_db, plain, "-r", qry_db, *extra)
        r = run(cli, main_db, inc, "-r", qry_db, "--index-cache", cache, *extra)
        assert "only the query database is read from SQLite" in r.stdout and "query genomes added to the cached index" in r.stdout
        assert open(plain, "rb").read() == open(inc, "rb").read(), extra
    # stale cache (main database touched): ignored, the ATTACH path answers, same bytes
    os.utime(main_db, (1, 1))
    r = run(cli, main_db, inc, "-r", qry_db, "--index-cache", cache)
    assert "only the query database is read" not in r.stdout
    run(cli, main_db, plain, "-r", qry_db)
    assert open(plain, "rb").read() == open(inc, "rb").read()


def test_against_reference_binary_on_synthetic_db(cli, tmp_path):
    ref = os.path.join(ROOT, "oracle", "_ref", "par_fastaai_ref.x")
    if not os.access(ref, os.X_OK):
        pytest.skip("oracle/_ref/par_fastaai_ref.x not built (needs /root/reference at build time)")
    from parfastaai_b200 import fastaai_db, synth
    N, P = 48, 16
    off, vals = synth.generate(N, P, seed=99)
    gn, sn = fastaai_db.synthetic_names(N, P)
    db = str(tmp_path / "syn.db")
    fastaai_db.write_db(db, gn, sn, off, vals)
    qfile = tmp_path / "q.txt"
    qfile.write_text("\n".join(gn[i] for i in (3, 7, 20, 41)) + "\n")
    for extra in ([], ["-q", str(qfile)], ["-s", "\t"]):
        a, b = str(tmp_path / "ours.csv"), str(tmp_path / "ref.csv")
        run(cli, db, a, *extra)
        r = subprocess.run([ref, db, b, *extra], capture_output=True, text=True, env=dict(os.environ, OMP_NUM_THREADS="4"))
        assert r.returncode == 0, r.stderr[-300:]
        assert open(a, "rb").read() == open(b, "rb").read(), extra


def test_streamed_csv_equals_the_csv_written_afterwards(cli, tmp_path):
    """One GPU, all-vs-all: by default the CSV rows are written by a writer thread while later row blocks are still computed
    (progress counters of the tile kernel); --no-stream-csv writes after the computation.  Same bytes; the streamed run says so."""
    from parfastaai_b200 import fastaai_db, synth
    N, P = 1800, 12
    off, vals = synth.generate(N, P, seed=5)
    gn, sn = fastaai_db.synthetic_names(N, P)
    db = str(tmp_path / "syn.db")
    fastaai_db.write_db(db, gn, sn, off, vals)
    a, b = str(tmp_path / "streamed.csv"), str(tmp_path / "after.csv")
    r1 = run(cli, db, a)
    r2 = run(cli, db, b, "--no-stream-csv")
    assert "streamed:" in r1.stdout and "streamed:" not in r2.stdout
    assert open(a, "rb").read() == open(b, "rb").read()
    assert open(a, "rb").read().count(b"\n") == N + 1


def test_index_cache_gives_identical_csv(cli, tmp_path):
    """--index-cache: a run served from the cache (SQLite never opened) writes the same CSV bytes, in all-vs-all and in
    query-subset mode (the cache holds the whole database)."""
    import shutil
    db = str(tmp_path / "s1.db")
    shutil.copy(gu.db_path("xdb_subset1.db"), db)
    cache = str(tmp_path / "s1.pfidx")
    n = gu.names()
    qfile = tmp_path / "q.txt"
    qfile.write_text("\n".join(n["subset1_genomes"][:2]) + "\n")
    for extra in ([], ["-q", str(qfile)]):
        plain, first, second = (str(tmp_path / f) for f in ("plain.csv", "first.csv", "second.csv"))
        run(cli, db, plain, *extra)
        if os.path.exists(cache):
            os.remove(cache)
        for f in (cache, cache + ".lists"):
            if os.path.exists(f):
                os.remove(f)
        r1 = run(cli, db, first, *extra, "--index-cache", cache)
        r2 = run(cli, db, second, *extra, "--index-cache", cache)
        assert "written" in r1.stdout and "device-ready lists written" in r1.stdout and os.path.exists(cache + ".lists")
        # second run: SQLite never opened, F neither read nor uploaded, no K1-K3 / sort -- the device-ready lists are imported
        assert "SQLite not opened" in r2.stdout and "device-ready lists loaded" in r2.stdout
        assert open(plain, "rb").read() == open(first, "rb").read() == open(second, "rb").read()
        # a damaged sidecar is ignored (the run falls back to F from the cache) and the result is still the same
        blob = bytearray(open(cache + ".lists", "rb").read()); blob[60] ^= 0xff
        open(cache + ".lists", "wb").write(bytes(blob))
        third = str(tmp_path / "third.csv")
        r3 = run(cli, db, third, *extra, "--index-cache", cache)
        assert r3.returncode == 0 and "device-ready lists loaded" not in r3.stdout
        assert open(third, "rb").read() == open(plain, "rb").read()
        # two contexts, replicated import of the cached lists, sharded compute
        os.remove(cache + ".lists")
        run(cli, db, first, *extra, "--index-cache", cache)
        fourth = str(tmp_path / "fourth.csv")
        r4 = run(cli, db, fourth, *extra, "--index-cache", cache, "--gpus", "2", "--same-device")
        assert r4.returncode == 0, r4.stdout + r4.stderr
        assert open(fourth, "rb").read() == open(plain, "rb").read()
    assert open(plain, "rb").read() != b""


def _sidecar_sums(seg_off_bytes, vals):
    """The two checksums of a PFLST002 sidecar (parfastaai_b200/host/index_cache.hpp: fnv1a / words_checksum)."""
    M, FNV0, FNVP = (1 << 64) - 1, 1469598103934665603, 1099511628211

    def fnv(h, data):
        for b in data:
            h = ((h ^ b) * FNVP) & M
        return h
    parts = []
    for lo in range(0, len(vals), 1 << 18):
        h = FNV0
        for w in vals[lo:lo + (1 << 18)].tolist():
            h = ((h ^ w) * FNVP) & M
        parts.append(h)
    return fnv(FNV0, seg_off_bytes), fnv(FNV0 ^ len(vals), np.array(parts, np.uint64).tobytes())


def test_index_cache_sidecar_the_device_refuses(cli, tmp_path):
    """A sidecar whose own checks pass (size, key, both checksums) but whose lists the device refuses -- two tetramers of one
    segment out of order -- must not fail the run, nor every later one: the driver drops it, loads F from the .pfidx cache,
    builds the ordinary way, writes the same CSV bytes and leaves a fresh, valid sidecar behind.  And a flipped bit inside the
    tetramer postings (segment still ascending) is caught by the sidecar's checksum before anything reaches the device."""
    import shutil
    import struct
    db = str(tmp_path / "s1.db")
    shutil.copy(gu.db_path("xdb_subset1.db"), db)
    cache = str(tmp_path / "s1.pfidx")
    plain, first = str(tmp_path / "plain.csv"), str(tmp_path / "first.csv")
    run(cli, db, plain)
    run(cli, db, first, "--index-cache", cache)
    good = open(cache + ".lists", "rb").read()
    assert good[:8] == b"PFLST002"
    P, G, nF = struct.unpack_from("<iiQ", good, 24)
    off_bytes = 8 * (P * G + 1)
    seg_off = np.frombuffer(good, np.int64, P * G + 1, 56)
    vals = np.frombuffer(good, np.uint32, nF, 56 + off_bytes).copy()
    assert _sidecar_sums(good[56:56 + off_bytes], vals) == struct.unpack_from("<QQ", good, 40)   # the Python restatement is right
    s = int(np.argmax(np.diff(seg_off) >= 2))
    a = int(seg_off[s])
    vals[a], vals[a + 1] = vals[a + 1], vals[a]                                  # descending pair inside one segment
    sums = _sidecar_sums(good[56:56 + off_bytes], vals)
    open(cache + ".lists", "wb").write(good[:40] + struct.pack("<QQ", *sums) + good[56:56 + off_bytes] + vals.tobytes())
    out = str(tmp_path / "refused.csv")
    r = run(cli, db, out, "--index-cache", cache)
    assert "device refused the cached device-ready lists" in r.stderr
    assert "device-ready lists written" in r.stdout                              # a fresh sidecar replaces the refused one
    assert open(out, "rb").read() == open(plain, "rb").read()
    assert open(cache + ".lists", "rb").read() == good
    again = str(tmp_path / "again.csv")
    r = run(cli, db, again, "--index-cache", cache)
    assert "device-ready lists loaded" in r.stdout and "refused" not in r.stderr
    assert open(again, "rb").read() == open(plain, "rb").read()
    # one flipped bit in the postings, the segment still ascending: read_lists_cache rejects it (checksum of vals), the run takes F
    flip = bytearray(good)
    flip[56 + off_bytes + 4 * int(seg_off[s + 1]) - 4] ^= 0x01
    open(cache + ".lists", "wb").write(bytes(flip))
    out = str(tmp_path / "flipped.csv")
    r = run(cli, db, out, "--index-cache", cache)
    assert "device-ready lists loaded" not in r.stdout and "refused" not in r.stderr
    assert open(out, "rb").read() == open(plain, "rb").read()
    assert open(cache + ".lists", "rb").read() == good


def test_multi_gpu_sharded_cli_matches_single_gpu(cli, tmp_path):
    """--gpus N: output row blocks sharded over N contexts, host scatter + mirror; bytes must equal the 1-GPU CSV.
    On a single-GPU box the N contexts share the device (--same-device); with >= 2 GPUs both layouts are run."""
    import torch
    from parfastaai_b200 import fastaai_db, synth
    N, P = 900, 3                      # 3 row blocks of 384
    off, vals = synth.generate(N, P, seed=77)
    gn, sn = fastaai_db.synthetic_names(N, P)
    db = str(tmp_path / "syn900.db")
    fastaai_db.write_db(db, gn, sn, off, vals)
    qfile = tmp_path / "q.txt"
    qfile.write_text("\n".join(gn[i] for i in (899, 0, 384, 385, 500)) + "\n")
    variants = [["--gpus", "2", "--same-device"], ["--gpus", "3", "--same-device"]]
    if torch.cuda.device_count() >= 2:
        variants.append(["--gpus", "2"])
    for extra in ([], ["-q", str(qfile)]):
        one = str(tmp_path / "one.csv")
        run(cli, db, one, *extra)
        for v in variants:
            many = str(tmp_path / "many.csv")
            r = run(cli, db, many, *extra, *v)
            assert open(one, "rb").read() == open(many, "rb").read(), (extra, v)
            assert f"{v[1]} GPU(s)" in r.stdout
    # a routed build that cannot be done (skewed input: chunk overflow) must fall back to the replicated build, not fail the run
    many = str(tmp_path / "fallback.csv")
    # (the failure switch exists only in the -DPFAAI_TEST_HOOKS build of the binary; the product binary ignores it)
    r = subprocess.run([ge.CLI_HOOKS, db, many, "--gpus", "2", "--same-device"], capture_output=True, text=True,
                       env=dict(os.environ, OMP_NUM_THREADS="4", PFAAI_TEST_FAIL_ROUTED="1"))
    assert r.returncode == 0 and "falling back to the replicated build" in r.stderr, r.stderr
    run(cli, db, one)
    assert open(one, "rb").read() == open(many, "rb").read()
    r = subprocess.run([cli, db, many, "--gpus", "2", "--same-device"], capture_output=True, text=True,
                       env=dict(os.environ, OMP_NUM_THREADS="4", PFAAI_TEST_FAIL_ROUTED="1"))
    assert r.returncode == 0 and "falling back" not in r.stderr
