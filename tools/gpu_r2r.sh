#!/bin/bash
set -u
OUT=gpurun_out/${TAG:-r2r}; mkdir -p $OUT
timeout 1500 python -m pytest tests -m gpu -x -q -rs > $OUT/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -15 $OUT/pytest_gpu.log
