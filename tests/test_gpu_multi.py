"""Two-rank NCCL parity test of the multi-GPU path on DISTINCT devices (skipped on a one-GPU box).

bench.py is launched under torchrun exactly as the driver launches it (one process per GPU, 127.0.0.1 rendezvous) on a small
database; every rank's share of the all-vs-all matrix -- a diagonal tile, a tile whose column dictionary was built by the OTHER
rank and travelled through the NCCL exchange, the ragged last row block, the last column block -- is compared with the oracle
(`verify`), and the assembled matrix with an unsharded compute on rank 0 (`verify_unsharded`).  Covered: the partitioned build
with the overlapped dictionary all-gather (default), the routed build (--sliced-index: all-to-all of the postings), the
replicated build, the dense-contraction tile path, and the compact-matrix fallback (--no-shared-matrix)."""
import json
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
pytestmark = pytest.mark.gpu


def _n_devices():
    import torch
    return torch.cuda.device_count() if torch.cuda.is_available() else 0


def _run(extra, port):
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
           "--master-port", str(port), os.path.join(ROOT, "bench.py"), "--gpus", "2", "--steps", "1", "--warmup", "1",
           "--genomes", "2500", "--scps", "16", "--verify", "--no-ceiling"] + extra
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600, cwd=ROOT)
    assert r.returncode == 0, r.stderr[-3000:]
    lines = [ln for ln in r.stdout.splitlines() if ln.startswith("{")]
    assert len(lines) == 1, r.stdout[-2000:]
    return json.loads(lines[0])


@pytest.mark.parametrize("extra", [[], ["--sliced-index"], ["--replicated-build"], ["--no-overlap"], ["--no-shared-matrix"],
                                   ["--dense", "--path", "mma"], ["--dense", "--sliced-index"]],
                         ids=["partitioned_async", "routed", "replicated", "partitioned_sync", "compact_matrices", "dense_mma",
                              "dense_routed"])
def test_two_ranks_over_nccl_match_the_oracle_and_the_unsharded_path(extra):
    if _n_devices() < 2:
        pytest.skip("needs two GPUs")
    line = _run(extra, 29611 + (abs(hash(tuple(extra))) % 200))
    assert line["n_gpus"] == 2
    v = line["verify"]
    assert v["ranks"] == 2 and v["cells"] > 1000
    assert v["max_abs_diff"] == 0.0, v
    names = {r["name"] for rk in v["rects"].values() for r in rk}
    assert {"diagonal", "late_list", "ragged_last_row_block", "last_column_block"} <= names
    assert line["verify_unsharded"]["identical"], line["verify_unsharded"]
    shared = "--no-shared-matrix" not in extra
    G = 2500
    assert line["e2e"]["d2h_bytes_per_step"] == (8 * G * G if shared else 8 * (G * (G - 1) // 2 + G))
