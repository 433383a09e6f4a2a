#!/bin/bash
# 2 x B200: the 2-rank NCCL parity test + the multi-GPU bench lines (shared host matrix, per-rank oracle verify)
set -u
OUT=gpurun_out/r2i; mkdir -p $OUT
timeout 900 python -m pytest tests/test_gpu_multi.py -x -q > $OUT/pytest_multi.log 2>&1; echo "pytest multi rc=$?"; tail -15 $OUT/pytest_multi.log
for extra in "" "--sliced-index"; do
  tag=n2$(echo $extra | tr -d ' -')
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 3 --warmup 3 $extra > $OUT/bench_$tag.json 2> $OUT/bench_$tag.err
  echo "bench $tag rc=$?"
  python - <<PY
import json
try:
    d=json.load(open("$OUT/bench_$tag.json")); print("  value %.3e e2e %.3e ms %.1f e2e_ms %.1f"%(d["value"],d["e2e"]["value"],d["ms_per_step"],d["e2e"]["ms_per_step"]), d["phases_ms"], "d2h", d["e2e"]["d2h_bytes_per_step"], "verify", d["verify"]["max_abs_diff"], d["verify"]["cells"], d["verify"]["matrix"])
except Exception as e: print("  no json", e); print(open("$OUT/bench_$tag.err").read()[-2500:])
PY
  grep -i "shared host matrix" $OUT/bench_$tag.err | head -3
done
