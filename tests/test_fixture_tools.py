"""SURVEY 8(f) rank 4 tooling: the streaming synthetic FastAAI-database writer and the cereal fixture dumper.
CPU: writer -> loader round trip against the generator's own reference arrays, cereal write/read round trip.
GPU: tools/dump_fixtures.py on the reference's xdb_subset1 database reproduces the reference's goldens."""
import os
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as ge  # noqa: E402
from tests import golden_util as gu  # noqa: E402


def test_written_database_loads_back_as_the_reference_arrays(tmp_path):
    ge.build()
    from parfastaai_b200 import fastaai_db, synth
    N, P = 150, 9
    off, vals = synth.generate(N, P, seed=3)
    gn, sn = fastaai_db.synthetic_names(N, P)
    db = str(tmp_path / "syn.db")
    fastaai_db.write_db(db, gn, sn, off, vals)
    d = fastaai_db.load_single(db)
    Lc, F = synth.to_reference(N, P, off, vals)
    assert d["genomes"] == gn and d["proteins"] == sn
    assert np.array_equal(d["Lc"], Lc) and np.array_equal(d["F"], F)
    assert np.array_equal(d["T"], synth.t_matrix(N, P, off))


def test_cereal_writers_round_trip(tmp_path):
    ge.build()
    from parfastaai_b200 import cereal_io
    rng = np.random.default_rng(1)
    iv = rng.integers(0, 1000, 77).astype(np.int32)
    dv = rng.random(13)
    T = rng.integers(0, 500, (7, 5)).astype(np.int32)
    F = np.zeros(9, cereal_io.PG_DTYPE); F["protein"] = np.arange(9); F["genome"] = 8 - np.arange(9)
    jac = np.zeros(4, cereal_io.JAC_DTYPE); jac["gA"] = [0, 0, 1, 2]; jac["gB"] = [1, 2, 2, 3]; jac["S"] = dv[:4]; jac["N"] = [3, 0, 9, 1]
    p = lambda n: str(tmp_path / n)
    cereal_io.write_int_vector(p("i"), iv); cereal_io.write_double_vector(p("d"), dv); cereal_io.write_t_matrix(p("t"), T)
    cereal_io.write_f_array(p("f"), F); cereal_io.write_jac(p("j"), jac)
    assert np.array_equal(cereal_io.read_int_vector(p("i")), iv) and np.array_equal(cereal_io.read_double_vector(p("d")), dv)
    assert np.array_equal(cereal_io.read_t_matrix(p("t")), T) and np.array_equal(cereal_io.read_f_array(p("f")), F)
    assert np.array_equal(cereal_io.read_jac(p("j")), jac)
    assert os.path.getsize(p("j")) == 8 + 4 * 20 and os.path.getsize(p("t")) == 24 + 4 * 35      # App. A.2 layouts


@pytest.mark.gpu
def test_fixture_dumper_reproduces_the_reference_goldens(tmp_path):
    ge.build()
    from parfastaai_b200 import cereal_io
    pre = str(tmp_path / "s1")
    r = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "dump_fixtures.py"), gu.db_path("xdb_subset1.db"), pre],
                       capture_output=True, text=True)
    assert r.returncode == 0, r.stderr[-800:]
    g = gu.load("xdb_subset1")
    assert np.array_equal(cereal_io.read_int_vector(pre + "_lc_array.bin"), g["lc"])
    assert np.array_equal(cereal_io.read_int_vector(pre + "_lp_array.bin"), g["lp"])
    assert np.array_equal(cereal_io.read_f_array(pre + "_f_array.bin"), g["f"])
    assert np.array_equal(cereal_io.read_t_matrix(pre + "_t_matrix.bin"), g["t"])
    jac = cereal_io.read_jac(pre + "_jac.bin")
    assert np.array_equal(jac["N"], g["jac"]["N"]) and np.abs(jac["S"] - g["jac"]["S"]).max() <= 1e-7
    assert np.array_equal(cereal_io.read_double_vector(pre + "_aji.bin"), g["aji"])
    assert open(pre + "_aji_matrix_wheader.csv", "rb").read() == gu.csv_bytes("xdb_subset1_aji_matrix_wheader.csv")
