#!/bin/bash
# K5 vs K4 on the dense family (config-4 size) and on the sparse family; full GPU test suite; ncu of the K5 kernel
set -u
OUT=gpurun_out/r2c; mkdir -p $OUT
timeout 900 python -m pytest tests -m gpu -x -q > $OUT/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -4 $OUT/pytest_gpu.log
for fam in dense sparse; do
  for path in mma lists auto; do
    FL=""; [ $fam = dense ] && FL="--dense"
    timeout 600 python bench.py --genomes 10000 $FL --path $path --steps 3 --warmup 3 --no-cpu-baseline > $OUT/bench_10k_${fam}_${path}.json 2> $OUT/bench_10k_${fam}_${path}.err
    echo "$fam $path rc=$?"
    python - <<PY
import json
try:
    d=json.load(open("$OUT/bench_10k_${fam}_${path}.json"))
    print("  value %.3e e2e %.3e ms %.2f"%(d["value"],d["e2e"]["value"],d["ms_per_step"]), d["phases_ms"], d["path"], "verify", d["verify"]["max_abs_diff"])
except Exception as e: print("  no json", e); print(open("$OUT/bench_10k_${fam}_${path}.err").read()[-1500:])
PY
  done
done
python bench.py --genomes 10000 --dense --path mma --steps 1 --warmup 1 --no-cpu-baseline --no-ceiling > $OUT/plain_mma.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:aji_mma -s 1 -c 1 -o $OUT/mma_dense_full \
    python bench.py --genomes 10000 --dense --path mma --steps 1 --warmup 1 --no-cpu-baseline --no-ceiling > $OUT/ncu_mma.log 2>&1
echo "ncu mma rc=$?"
