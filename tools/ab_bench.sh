#!/bin/bash
# A/B of kernel variants on one GPU box: tools/ab_bench.sh <tag> <lib1> <lib2> ...   (libs relative to parfastaai_b200/)
set -u
TAG=$1; shift
OUT=gpurun_out/$TAG; mkdir -p $OUT
python -m pytest tests -m gpu -x -q > $OUT/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 $OUT/pytest_gpu.log
for L in "$@"; do
  PFAAI_GPU_LIB=$PWD/parfastaai_b200/$L python bench.py --no-cpu-baseline --steps 3 --warmup 3 ${BENCH_ARGS:-} > $OUT/bench_$L.json 2> $OUT/bench_$L.err
  echo "$L rc=$?"; python - <<PY
import json
try:
    d=json.load(open("$OUT/bench_$L.json")); print("$L", "value %.3e e2e %.3e ms/step %.2f" % (d["value"], d["e2e"]["value"], d["ms_per_step"]), d["phases_ms"])
except Exception as e: print("no json", e); print(open("$OUT/bench_$L.err").read()[-1500:])
PY
done
