#!/bin/bash
set -u
OUT=gpurun_out/${TAG:-r2s2}; mkdir -p $OUT
timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_mma.py -m gpu -x -q > $OUT/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 $OUT/pytest_gpu.log
timeout 200 python bench.py --genomes 10000 --steps 5 --warmup 3 --no-cpu-baseline --no-ceiling > $OUT/bench_10k.json 2> $OUT/bench_10k.err; echo "bench rc=$?"
python - <<PY
import json
d=json.load(open("$OUT/bench_10k.json")); print("  value %.3e ms %.2f tile %.2f verify"%(d["value"],d["ms_per_step"],d["phases_ms"]["tile"]), d["verify"]["max_abs_diff"])
PY
