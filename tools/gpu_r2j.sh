#!/bin/bash
# 8 x B200: the multi-GPU bench lines at N = 8 (partitioned and routed builds) and N = 4
set -u
OUT=gpurun_out/r2j; mkdir -p $OUT
run() {  # n, tag, extra...
  n=$1; tag=$2; shift 2
  timeout 420 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $n --steps 3 --warmup 3 "$@" > $OUT/bench_$tag.json 2> $OUT/bench_$tag.err
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
run 4 n4
