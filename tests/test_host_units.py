"""CPU-only unit checks of host pieces whose callers need a GPU: the device-ready lists sidecar of the index cache and the
streaming CSV writer driven by a scripted progress source (tests/host_units.cpp, compiled here with g++)."""
import os
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_lists_sidecar_and_streaming_writer(tmp_path):
    exe = str(tmp_path / "host_units.x")
    subprocess.run(["/usr/bin/g++", "-std=c++17", "-O1", "-fopenmp", os.path.join(ROOT, "tests", "host_units.cpp"), "-o", exe,
                    "/usr/lib/x86_64-linux-gnu/libsqlite3.so.0"], check=True)
    r = subprocess.run([exe, str(tmp_path)], capture_output=True, text=True, env=dict(os.environ, OMP_NUM_THREADS="3"))
    assert r.returncode == 0 and r.stdout.strip() == "OK", r.stdout + r.stderr
