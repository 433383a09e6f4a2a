"""parfastaai_b200 -- B200-native (sm_100a) implementation of ParFastAAI's AJI hot path.

Only what the path needs lives here: ``csrc/`` (CUDA kernels + the C-ABI of include/pfaai_gpu.h),
``_lib`` (ctypes binding) and small host-side tooling (fixture readers, synthetic generator,
FastAAI-SQLite ingest, CSV writer).  The product path has no CPU fallback.
"""
from ._lib import (ALL_VS_ALL, FLAG_REF_COMPAT_QT, JAC_DTYPE, NTETRA, PATH_AUTO, PATH_LISTS, PATH_MMA, PG_DTYPE,  # noqa: F401
                   QUERY_DB, QUERY_SUBSET, STATIC_SCAN_MAX, PfaaiError, PfaaiGpu, host_register, host_unregister, load,
                   prefix_sum_static)

__all__ = ["PfaaiGpu", "PfaaiError", "load", "ALL_VS_ALL", "QUERY_SUBSET", "QUERY_DB", "FLAG_REF_COMPAT_QT",
           "PG_DTYPE", "JAC_DTYPE", "NTETRA", "PATH_AUTO", "PATH_LISTS", "PATH_MMA", "host_register", "host_unregister",
           "prefix_sum_static", "STATIC_SCAN_MAX"]
