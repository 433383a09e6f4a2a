#!/bin/bash
# 8 x B200: the 2-rank NCCL parity test, then config 5 at N = 8 with the partitioned build (default) and the routed build
set -u
OUT=gpurun_out/${TAG:-r2t}; mkdir -p $OUT
timeout 300 python -m pytest tests/test_gpu_multi.py -m gpu -x -q -rs > $OUT/pytest_multi.log 2>&1; echo "pytest multi rc=$?"; tail -4 $OUT/pytest_multi.log
run() {  # n, tag, extra...
  n=$1; tag=$2; shift 2
  timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $n --steps 3 --warmup 3 "$@" > $OUT/bench_$tag.json 2> $OUT/bench_$tag.err
  echo "bench $tag rc=$?"
  python - <<PY
import json
try:
    d=json.load(open("$OUT/bench_$tag.json")); print("  value %.3e e2e %.3e ms %.1f e2e_ms %.1f"%(d["value"],d["e2e"]["value"],d["ms_per_step"],d["e2e"]["ms_per_step"]), d["phases_ms"], "d2h", d["e2e"]["d2h_bytes_per_step"], "verify", d["verify"]["max_abs_diff"], d["verify"]["cells"])
except Exception as e: print("  no json", e); print(open("$OUT/bench_$tag.err").read()[-2500:])
PY
  grep -i "shared host matrix" $OUT/bench_$tag.err | head -2
}
run 8 n8
run 8 n8sliced --sliced-index
