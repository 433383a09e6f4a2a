#!/bin/bash
# 1 x B200: full GPU test suite after the f2/f3/a1 additions (streaming CSV, device-ready cache, device Lc), then the benches
set -u
OUT=gpurun_out/${TAG:-r2p}; mkdir -p $OUT
timeout 1200 python -m pytest tests -m gpu -x -q -rs > $OUT/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -12 $OUT/pytest_gpu.log
timeout 120 python -c "import __graft_entry__ as g; g.smoke()" > $OUT/smoke.log 2>&1; echo "smoke rc=$?"
show() { python - "$1" <<'PY'
import json,sys
try:
    d=json.load(open(sys.argv[1]))
    print("  value %.3e e2e %.3e ms %.2f"%(d["value"],d["e2e"]["value"],d["ms_per_step"]), d["phases_ms"], d.get("path"), "verify", d["verify"]["max_abs_diff"])
except Exception as e: print("  no json", e); print(open(sys.argv[1].replace(".json",".err")).read()[-1500:])
PY
}
run() { tag=$1; shift; timeout 600 python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-ceiling "$@" > $OUT/bench_$tag.json 2> $OUT/bench_$tag.err; echo "$tag rc=$?"; show $OUT/bench_$tag.json; }
run 10k_dense_auto --genomes 10000 --dense
run 10k_sparse_auto --genomes 10000
