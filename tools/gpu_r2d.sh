#!/bin/bash
set -u
OUT=gpurun_out/${TAG:-r2d}; mkdir -p $OUT
timeout 120 python -m pytest tests/test_gpu_mma.py -x -q > $OUT/pytest_mma.log 2>&1; echo "pytest mma rc=$?"; tail -3 $OUT/pytest_mma.log
timeout 240 python bench.py --genomes 10000 --dense --path mma --steps 3 --warmup 3 --no-cpu-baseline > $OUT/bench_10k_dense_mma.json 2> $OUT/bench_10k_dense_mma.err
python - <<PY
import json
try:
    d=json.load(open("$OUT/bench_10k_dense_mma.json"))
    print("  value %.3e e2e %.3e ms %.2f"%(d["value"],d["e2e"]["value"],d["ms_per_step"]), d["phases_ms"], d["path"], "verify", d["verify"]["max_abs_diff"])
except Exception as e: print("  no json", e); print(open("$OUT/bench_10k_dense_mma.err").read()[-1500:])
PY
if [ "${NCU:-1}" = "1" ]; then
timeout 200 python bench.py --genomes 10000 --dense --path mma --steps 1 --warmup 1 --no-cpu-baseline --no-ceiling > $OUT/plain_mma.log 2>&1 &&
timeout 400 ncu --set full --clock-control none --import-source on -k regex:aji_mma -s 1 -c 1 -o $OUT/mma_dense_full \
    python bench.py --genomes 10000 --dense --path mma --steps 1 --warmup 1 --no-cpu-baseline --no-ceiling > $OUT/ncu_mma.log 2>&1
echo "ncu mma rc=$?"
fi
