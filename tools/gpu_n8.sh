#!/bin/bash
# 8-GPU config-5 points: tools/gpu_n8.sh <tag> <N> "<PFAAI_EXCHANGE_SMS values>"
set -u
TAG=$1; N=$2; VALS=$3
OUT=gpurun_out/$TAG; mkdir -p $OUT
for R in $VALS; do
  PFAAI_EXCHANGE_SMS=$R python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 4 --warmup 3 > $OUT/r$R.json 2> $OUT/r$R.err
  echo "reserve=$R rc=$?"; python - <<PY
import json
try:
    d=json.load(open("$OUT/r$R.json")); print("reserve $R", "value %.4e e2e %.4e ms/step %.2f" % (d["value"], d["e2e"]["value"], d["ms_per_step"]), d["phases_ms"])
except Exception as e: print("no json", e); print(open("$OUT/r$R.err").read()[-2000:])
PY
done
