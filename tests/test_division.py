"""The tile kernel divides c / (Ta + Tb - c) with a reciprocal table and two FMAs (aji_tile.cuh, jaccard_exact) instead of the
fp64 division sequence.  Bit-exactness with the reference's IEEE division (algorithm_impl.hpp:250-254) rests on this
exhaustive check of the same operation sequence on the CPU: every 0 <= c <= den <= 20000 (the kernel's table covers
den < 2048 and falls back to a true division beyond it)."""
import ctypes as C
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import pfaai_oracle as orc  # noqa: E402


def test_table_division_is_correctly_rounded():
    L = orc.lib()
    L.pfo_check_table_division.restype = C.c_int64
    L.pfo_check_table_division.argtypes = [C.c_int]
    assert L.pfo_check_table_division(20000) == 0
