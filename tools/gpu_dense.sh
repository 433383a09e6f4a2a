#!/bin/bash
# Dense-family visit: the default bench line (with its dense_family side record), then a full ncu capture of the SHIPPED
# tensor-core tile kernel (aji_mma_kernel) on the real-like family at config-4 size.
# Usage (under gpurun): bash tools/gpu_dense.sh <tag> [genomes]
set -u
TAG=${1:-dense}; G=${2:-10000}
OUT=gpurun_out/$TAG; mkdir -p $OUT
python bench.py --steps 3 --warmup 3 --no-cpu-baseline > $OUT/bench_n1.json 2> $OUT/bench_n1.err; echo "bench rc=$?"
python - <<PY
import json
try:
    d = json.load(open("$OUT/bench_n1.json")); print("value %.4e e2e %.4e" % (d["value"], d["e2e"]["value"])); print(json.dumps(d.get("dense_family"), indent=1)[:2500])
except Exception as e: print("no json", e); print(open("$OUT/bench_n1.err").read()[-1500:])
PY
CMD="python bench.py --dense --genomes $G --steps 2 --warmup 1 --no-cpu-baseline --no-ceiling"
$CMD > $OUT/dense_plain.json 2> $OUT/dense_plain.err; echo "dense bench rc=$?"
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $OUT/dense_launches.csv $CMD > $OUT/ncu_launches.log 2>&1
echo "ncu launches rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:aji_mma -s 1 -c 1 -o $OUT/mma_full $CMD > $OUT/ncu_full.log 2>&1
echo "ncu full rc=$?"
ncu -i $OUT/mma_full.ncu-rep --page details > $OUT/mma_details.txt 2>&1
ncu -i $OUT/mma_full.ncu-rep --page source --csv > $OUT/mma_source.csv 2>/dev/null && python tools/ncu_src_summary.py $OUT/mma_source.csv > $OUT/mma_source_summary.txt 2>&1
ls -la $OUT
