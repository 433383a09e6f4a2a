#!/bin/bash
# 1 x B200: the hybrid tile path (tensor-core contraction + in-kernel sparse walk): parity tests, then both synthetic families
set -u
OUT=gpurun_out/${TAG:-r2l}; mkdir -p $OUT
timeout 400 python -m pytest tests/test_gpu_mma.py -x -q -rs > $OUT/pytest_mma.log 2>&1; echo "pytest mma rc=$?"; tail -15 $OUT/pytest_mma.log
show() { python - "$1" <<'PY'
import json,sys
try:
    d=json.load(open(sys.argv[1]))
    print("  value %.3e e2e %.3e ms %.2f"%(d["value"],d["e2e"]["value"],d["ms_per_step"]), d["phases_ms"], d.get("path"), "verify", d["verify"]["max_abs_diff"])
except Exception as e: print("  no json", e); print(open(sys.argv[1].replace(".json",".err")).read()[-1500:])
PY
}
if [ "${QUICK:-0}" = "1" ]; then exit 0; fi
timeout 300 python bench.py --genomes 10000 --path mma --steps 3 --warmup 3 --no-cpu-baseline --no-ceiling > $OUT/bench_10k_sparse_mma.json 2> $OUT/bench_10k_sparse_mma.err
echo "sparse 10k mma rc=$?"; show $OUT/bench_10k_sparse_mma.json
timeout 300 python bench.py --genomes 10000 --dense --steps 3 --warmup 3 --no-cpu-baseline --no-ceiling > $OUT/bench_10k_dense_auto.json 2> $OUT/bench_10k_dense_auto.err
echo "dense 10k auto rc=$?"; show $OUT/bench_10k_dense_auto.json
timeout 600 python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-ceiling > $OUT/bench_50k_auto.json 2> $OUT/bench_50k_auto.err
echo "sparse 50k auto rc=$?"; show $OUT/bench_50k_auto.json
