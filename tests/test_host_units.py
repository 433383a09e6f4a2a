"""CPU-only unit checks of host pieces whose callers need a GPU: the device-ready lists sidecar of the index cache and the
streaming CSV writer driven by a scripted progress source, and the incremental query-DB path on the cached main index
(tests/host_units.cpp, compiled here with g++)."""
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from tests import golden_util as gu  # noqa: E402


def test_lists_sidecar_and_streaming_writer(tmp_path):
    exe = str(tmp_path / "host_units.x")
    subprocess.run(["/usr/bin/g++", "-std=c++17", "-O1", "-fopenmp", os.path.join(ROOT, "tests", "host_units.cpp"), "-o", exe,
                    "/usr/lib/x86_64-linux-gnu/libsqlite3.so.0"], check=True)
    # + the incremental query-DB path on the cached main index against the reference's ATTACH path (xdb_subset1 main, xdb_subset2 queries)
    r = subprocess.run([exe, str(tmp_path), gu.db_path("xdb_subset1.db"), gu.db_path("xdb_subset2.db")], capture_output=True, text=True,
                       env=dict(os.environ, OMP_NUM_THREADS="3"))
    assert r.returncode == 0 and r.stdout.strip().endswith("OK"), r.stdout + r.stderr
