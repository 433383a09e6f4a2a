#!/bin/bash
# 1 x B200: bulk copies split into small pieces (hybrid path and, as an A/B switch, the list path)
set -u
OUT=gpurun_out/${TAG:-r2n}; mkdir -p $OUT
timeout 600 python -m pytest tests/test_gpu_mma.py tests/test_gpu_parity.py -x -q -rs > $OUT/pytest.log 2>&1; echo "pytest rc=$?"; tail -6 $OUT/pytest.log
show() { python - "$1" <<'PY'
import json,sys
try:
    d=json.load(open(sys.argv[1]))
    print("  value %.3e e2e %.3e ms %.2f"%(d["value"],d["e2e"]["value"],d["ms_per_step"]), d["phases_ms"], d.get("path"), "verify", d["verify"]["max_abs_diff"])
except Exception as e: print("  no json", e); print(open(sys.argv[1].replace(".json",".err")).read()[-1500:])
PY
}
run() { tag=$1; shift; timeout 600 python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-ceiling "$@" > $OUT/bench_$tag.json 2> $OUT/bench_$tag.err; echo "$tag rc=$?"; show $OUT/bench_$tag.json; }
PFAAI_TILE_PIECE=0 run 10k_lists_piece0 --genomes 10000 --path lists
PFAAI_TILE_PIECE=4096 run 10k_lists_piece4096 --genomes 10000 --path lists
PFAAI_TILE_PIECE=1024 run 10k_lists_piece1024 --genomes 10000 --path lists
run 10k_dense_auto --genomes 10000 --dense
run 10k_sparse_mma_tau1024 --genomes 10000 --path mma --tau 1024
run 10k_sparse_auto --genomes 10000
run 50k_auto
run 50k_mma_tau1024 --path mma --tau 1024
