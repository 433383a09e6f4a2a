#!/bin/bash
set -u
OUT=gpurun_out/${TAG:-r2x}; mkdir -p $OUT
timeout 600 python -m pytest tests/test_gpu_cli.py tests/test_fixture_tools.py -m gpu -x -q -rs > $OUT/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -5 $OUT/pytest_gpu.log
