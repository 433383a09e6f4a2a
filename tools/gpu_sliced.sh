#!/bin/bash
# routed build (--sliced-index) on N GPUs: tools/gpu_sliced.sh <tag> <N> [genomes]
set -u
TAG=$1; N=$2; G=${3:-50000}
OUT=gpurun_out/$TAG; mkdir -p $OUT
run() {
  local name=$1; shift
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N "$@" > $OUT/$name.json 2> $OUT/$name.err
  echo "$name rc=$?"; python - <<PY
import json
try:
    d=json.load(open("$OUT/$name.json")); print("$name", "value %.4e e2e %.4e ms/step %.2f e2e ms %.2f" % (d["value"], d["e2e"]["value"], d["ms_per_step"], d["e2e"]["ms_per_step"]), d["phases_ms"], d.get("verify"))
except Exception as e: print("no json", e); print(open("$OUT/$name.err").read()[-2500:])
PY
}
run verify_6000 --steps 2 --warmup 3 --genomes 6000 --verify --sliced-index
run sliced_$G --steps 3 --warmup 3 --genomes $G --sliced-index
