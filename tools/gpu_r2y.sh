#!/bin/bash
set -u
OUT=gpurun_out/${TAG:-r2y}; mkdir -p $OUT
timeout 900 python -m pytest tests/test_gpu_mma.py tests/test_gpu_parity.py tests/test_gpu_cli.py -m gpu -x -q -rs > $OUT/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -5 $OUT/pytest_gpu.log
timeout 300 python bench.py --genomes 10000 --dense --steps 3 --warmup 3 --no-cpu-baseline --no-ceiling > $OUT/bench_10k_dense_auto.json 2> $OUT/bench_10k_dense_auto.err; echo "dense rc=$?"
python - <<PY
import json
d=json.load(open("$OUT/bench_10k_dense_auto.json")); print("  value %.3e ms %.2f"%(d["value"],d["ms_per_step"]), d["phases_ms"]["tile"], d["path"]["dense_tau"], d["path"]["prep_ms"], "verify", d["verify"]["max_abs_diff"])
PY
