"""CPU-side checks of the drop-in boundary: the C-ABI library loads, exports every symbol that
include/pfaai_gpu.h declares, and refuses to run without a CUDA device (no CPU fallback)."""
import ctypes as C
import os
import re
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as ge  # noqa: E402


@pytest.fixture(scope="module")
def lib():
    ge.build()
    from parfastaai_b200 import _lib
    return _lib.load()


def declared_symbols():
    hdr = open(os.path.join(ROOT, "include", "pfaai_gpu.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    return sorted(set(re.findall(r"\b(pfaai_gpu_\w+)\s*\(", hdr)))


def test_header_symbols_exported(lib):
    syms = declared_symbols()
    assert len(syms) >= 15
    for s in syms:
        assert hasattr(lib, s), f"{s} declared in include/pfaai_gpu.h but not exported"
    from parfastaai_b200 import _lib
    assert sorted(_lib.EXPORTS) == syms


def test_n_pairs_matches_reference_formulas(lib):
    # nGenomePairs of the three mode classes (ds_impl.hpp:79-81,244-249,403)
    assert lib.pfaai_gpu_n_pairs(0, 20, 0) == 190
    assert lib.pfaai_gpu_n_pairs(1, 20, 5) == 85
    assert lib.pfaai_gpu_n_pairs(2, 8, 4) == 16
    assert lib.pfaai_gpu_n_pairs(0, 50000, 0) == 1249975000


def test_no_cpu_fallback(lib):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    h = C.c_void_p()
    rc = lib.pfaai_gpu_create(C.byref(h), 0, 0)
    assert rc == 3 and not h.value                      # PFAAI_ERR_CONSTRUCT
    assert b"no CUDA device" in lib.pfaai_gpu_last_error(None)
    import parfastaai_b200 as pf
    with pytest.raises(pf.PfaaiError):
        pf.PfaaiGpu(0)


def test_struct_layouts():
    from parfastaai_b200 import _lib
    assert _lib.PG_DTYPE.itemsize == 8 and _lib.JAC_DTYPE.itemsize == 24
    assert C.sizeof(_lib.Timings) == 88
    # pfaai_exchange_t: 4 pointers + 2 x 4 int64 + 2 int32; pfaai_route_t: pointer + 17 int64 + 2 int32 (include/pfaai_gpu.h)
    assert C.sizeof(_lib.Exchange) == 104 and C.sizeof(_lib.Route) == 152
