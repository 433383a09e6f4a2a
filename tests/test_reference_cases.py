"""The reference's OWN Catch2 test file (tests/pfaai_tests.cpp of the reference, unmodified) as a parity gate.

oracle/Makefile (`make ref_tests`) compiles that file twice from where it lies under /root/reference:
  oracle/_ref/pfaai_tests_ref.x   with the reference's own ParFAAIImpl (CPU): run here on the data directory that
                                  tests/binding/make_data.py rebuilds from tests/golden/ -- it pins that directory, including the
                                  re-created 20-genome database (absent from the reference repository);
  oracle/_ref/pfaai_tests_gpu.x   with tests/binding/pfaai/algorithm_impl.hpp shadowing the reference's header, i.e. the
                                  reference-side binding of libpfaai_gpu.so: the reference's JAC / AJI cases of all three modes
                                  (tests/pfaai_tests.cpp:355,388,422,455,653) run through the C-ABI on the GPU.
The binaries are built in the container and travel to the GPU box (oracle/_ref is git-ignored, not gpurun-ignored)."""
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
REF_X = os.path.join(ROOT, "oracle", "_ref", "pfaai_tests_ref.x")
GPU_X = os.path.join(ROOT, "oracle", "_ref", "pfaai_tests_gpu.x")
COMPUTE_CASES = "[compute_JAC_AJI],[compute_subset1_JAC_AJI],[compute_subset2_JAC_AJI],[query_subset],[qt_compute_JAC_AJI]"


@pytest.fixture(scope="module")
def data_dir(tmp_path_factory):
    sys.path.insert(0, os.path.join(ROOT, "tests", "binding"))
    import make_data
    out = str(tmp_path_factory.mktemp("refcases"))
    make_data.main(out)
    return out


def _run(exe, spec, cwd):
    r = subprocess.run([exe, spec], cwd=cwd, capture_output=True, text=True, timeout=600)
    tail = (r.stdout + r.stderr)[-2500:]
    assert r.returncode == 0, tail
    assert "All tests passed" in r.stdout, tail
    return r.stdout


def test_unmodified_reference_passes_its_own_cases_on_the_rebuilt_data(data_dir):
    if not os.access(REF_X, os.X_OK):
        pytest.skip("oracle/_ref/pfaai_tests_ref.x not built (needs /root/reference: make -C oracle ref_tests)")
    # every case but the one whose golden (xanthodb_sorted_e_array.bin) is absent from the reference repository itself
    out = _run(REF_X, "~[construct_LFTE]", data_dir)
    assert "16 test cases" in out or "test cases: 16" in out, out[-400:]


@pytest.mark.gpu
def test_reference_catch2_cases_through_the_gpu_binding(data_dir):
    assert os.access(GPU_X, os.X_OK), "oracle/_ref/pfaai_tests_gpu.x missing: run __graft_entry__.build() where /root/reference is mounted"
    out = _run(GPU_X, COMPUTE_CASES, data_dir)
    assert "5 test cases" in out or "test cases: 5" in out, out[-400:]
