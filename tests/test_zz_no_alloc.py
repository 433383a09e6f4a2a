"""pfaai_gpu_prefix_sum_static: the device prefix sum for callers that must not allocate GPU memory.

Who needs it: code that runs inside a host program which has already handed the GPU's memory to its own pool (a plug-in
under PyTorch's caching allocator or an RMM arena, a server with a fixed memory budget) -- there a cudaMalloc made by this
library fails or takes memory the pool counted on -- and code on a latency-critical path, where cudaMalloc / cudaFree stall
the whole device.  The entry point keeps its partial sums in a fixed 8 KiB array inside the library instead of a
cudaMalloc'ed context buffer; inputs and outputs are device buffers of the caller (here: torch tensors, i.e. memory that
came from the caller's own pool).

The checks: the new entry point against the EXISTING one (pfaai_gpu_upload_index + pfaai_gpu_build + pfaai_gpu_get_arrays,
whose Lp and seg_off are produced by the allocating scan) on the same inputs, bit for bit (integer work: tolerance 0, as in
tests/test_gpu_parity.py), on calls made back to back so that each call finds the previous call's partial sums in the
shared array.  This file sorts last on purpose (pytest -x stops at the first failure); it passes alone in a fresh process
(the first call then does the per-device set-up) and after the rest of the suite.
"""
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from tests import golden_util as gu  # noqa: E402


@pytest.fixture(scope="module")
def pf():
    import __graft_entry__ as ge
    ge.build()
    import parfastaai_b200 as pf
    return pf


def _existing_entry_point(pf, name):
    """Lc, Lp (int32) and the (genome, SCP) segment lengths / offsets as the allocating build path computes them."""
    g = gu.load(name)
    P, G = g["t"].shape
    ctx = pf.PfaaiGpu(0)
    ctx.upload_index(P, G, g["f"], g["lc"])
    ctx.build()
    arr = ctx.get_arrays()
    ctx.close()
    seg_len = np.ascontiguousarray(arr["T"].T).reshape(-1).astype(np.int32)      # segment g*P + p has T[p][g] tetramers
    return {"Lc": arr["Lc"], "Lp": arr["Lp"], "seg_len": seg_len, "seg_off": arr["seg_off"], "nF": int(g["f"].shape[0])}


def _scan_static(pf, torch, counts, want32=True):
    """One call through the C-ABI on torch-owned device buffers; the outputs start poisoned."""
    dev = torch.device("cuda:0")
    n = int(counts.shape[0])
    d_in = torch.from_numpy(counts).to(dev)
    d_out = torch.full((n + 1,), -0x0123456789ABCDEF, dtype=torch.int64, device=dev)
    d_out32 = torch.full((max(n, 1),), -0x1234567, dtype=torch.int32, device=dev) if want32 else None
    torch.cuda.synchronize()
    stream = torch.cuda.current_stream(dev)
    pf.prefix_sum_static(0, stream.cuda_stream, d_in.data_ptr() if n else 0, n, d_out.data_ptr(),
                         d_out32.data_ptr() if want32 else 0)
    stream.synchronize()
    return d_out.cpu().numpy(), (d_out32.cpu().numpy()[:n] if want32 else None)


@pytest.mark.gpu
def test_static_scan_equals_the_allocating_scan_on_consecutive_calls(pf):
    import torch
    big = _existing_entry_point(pf, "xanthodb")
    small = _existing_entry_point(pf, "xdb_subset1")

    # call 1: Lc -> Lp of the large fixture (160000 inputs = 40 scan tiles)
    out, out32 = _scan_static(pf, torch, big["Lc"])
    assert np.array_equal(out32, big["Lp"])
    assert np.array_equal(out[:-1], big["Lp"].astype(np.int64)) and out[-1] == big["nF"]

    # call 2, straight after: a much shorter input (one tile), so entries 1..39 of the shared array still hold call 1's sums
    out, out32 = _scan_static(pf, torch, small["seg_len"])
    assert np.array_equal(out, small["seg_off"])
    assert np.array_equal(out32, small["seg_off"][:-1].astype(np.int32))

    # call 3: back to a long input over the short call's leftovers, this time without the optional int32 output
    out, _ = _scan_static(pf, torch, small["Lc"], want32=False)
    assert np.array_equal(out[:-1], small["Lp"].astype(np.int64)) and out[-1] == small["nF"]

    # call 4: the segment histogram of the large fixture (ragged lengths, several tiles) against seg_off
    out, _ = _scan_static(pf, torch, big["seg_len"])
    assert np.array_equal(out, big["seg_off"])


@pytest.mark.gpu
def test_static_scan_edges_and_the_limit(pf):
    import torch
    # empty input: out[0] = 0 and nothing else is written
    out, _ = _scan_static(pf, torch, np.zeros(0, np.int32))
    assert out.tolist() == [0]
    # exactly the maximum size, sums beyond 32 bits, then a 1-element call on top of its leftovers
    rng = np.random.default_rng(7)
    counts = rng.integers(0, 20000, pf.STATIC_SCAN_MAX, dtype=np.int32)
    expect = np.concatenate([[0], np.cumsum(counts, dtype=np.int64)])
    assert expect[-1] > 2 ** 32
    for _ in range(2):
        out, out32 = _scan_static(pf, torch, counts)
        assert np.array_equal(out, expect)
        assert np.array_equal(out32, expect[:-1].astype(np.int32))              # truncated like the reference's int Lp
    out, _ = _scan_static(pf, torch, np.array([5], np.int32))
    assert out.tolist() == [0, 5]
    # the caller's current device is untouched and the library took no device memory of its own during the calls
    free0, _ = torch.cuda.mem_get_info(0)
    d_in = torch.from_numpy(counts).cuda()
    d_out = torch.empty(counts.shape[0] + 1, dtype=torch.int64, device="cuda:0")
    torch.cuda.synchronize()
    free1, _ = torch.cuda.mem_get_info(0)
    for _ in range(3):
        pf.prefix_sum_static(0, torch.cuda.current_stream().cuda_stream, d_in.data_ptr(), counts.shape[0], d_out.data_ptr(), 0)
    torch.cuda.synchronize()
    free2, _ = torch.cuda.mem_get_info(0)
    assert free2 == free1 <= free0
    assert np.array_equal(d_out.cpu().numpy(), expect)


def test_static_scan_rejects_what_it_cannot_hold(pf):
    """Argument checks come before any CUDA call, so this part also runs on a machine without a GPU."""
    with pytest.raises(pf.PfaaiError, match="exceeds PFAAI_GPU_STATIC_SCAN_MAX"):
        pf.prefix_sum_static(0, 0, 16, pf.STATIC_SCAN_MAX + 1, 16)
    with pytest.raises(pf.PfaaiError, match="bad arguments"):
        pf.prefix_sum_static(0, 0, 16, -1, 16)
    with pytest.raises(pf.PfaaiError, match="bad arguments"):
        pf.prefix_sum_static(0, 0, 0, 8, 16)         # n > 0 without an input
    with pytest.raises(pf.PfaaiError, match="bad arguments"):
        pf.prefix_sum_static(0, 0, 16, 8, 0)         # no output
    with pytest.raises(pf.PfaaiError, match="device ordinal"):
        pf.prefix_sum_static(64, 0, 16, 8, 16)
