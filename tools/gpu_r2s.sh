#!/bin/bash
set -u
OUT=gpurun_out/${TAG:-r2s}; mkdir -p $OUT
timeout 1500 python -m pytest tests/test_gpu_parity.py tests/test_reference_cases.py tests/test_fixture_tools.py -m gpu -x -q -rs > $OUT/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -8 $OUT/pytest_gpu.log
