#!/bin/bash
# 2 x B200: the multi-GPU data path with parity evidence (shared host matrix, per-rank oracle verify), a real 2-rank NCCL test
set -u
OUT=gpurun_out/r2h; mkdir -p $OUT
timeout 120 python -m pytest tests/test_gpu_mma.py -x -q > $OUT/pytest_mma.log 2>&1; echo "pytest mma rc=$?"; tail -2 $OUT/pytest_mma.log
timeout 200 python bench.py --genomes 10000 --dense --path mma --steps 3 --warmup 3 --no-cpu-baseline --no-ceiling > $OUT/bench_10k_dense_mma.json 2> $OUT/bench_10k_dense_mma.err
python - <<PY
import json
try:
    d=json.load(open("$OUT/bench_10k_dense_mma.json")); print("  dense mma value %.3e ms %.2f"%(d["value"],d["ms_per_step"]), d["phases_ms"]["tile"], d["path"]["prep_ms"], "verify", d["verify"]["max_abs_diff"])
except Exception as e: print("  no json", e)
PY
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
done
