#!/bin/bash
# K5 bring-up: the dense-path tests only (fast fail), under a timeout so that a hung pipeline cannot hold the box
set -u
OUT=gpurun_out/r2b; mkdir -p $OUT
timeout 600 python -m pytest tests/test_gpu_mma.py -x -q > $OUT/pytest_mma.log 2>&1; echo "pytest mma rc=$?"; tail -40 $OUT/pytest_mma.log
