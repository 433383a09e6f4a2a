#!/bin/bash
# One GPU-box visit: parity tests, the default bench, the ncu launch list and one full capture of the tile kernel.
# Usage (under gpurun): bash tools/gpu_round.sh <tag>
set -u
TAG=${1:-run}
OUT=gpurun_out/$TAG
mkdir -p $OUT
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $OUT/smi.txt 2>&1
lscpu | grep -E 'Model name|^CPU\(s\)|Socket|Thread' > $OUT/cpu.txt 2>&1
python -m pytest tests -m gpu -x -q > $OUT/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a $OUT/pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke()" > $OUT/smoke.log 2>&1; echo "smoke rc=$?" | tee -a $OUT/smoke.log
python bench.py > $OUT/bench.json 2> $OUT/bench.err; echo "bench rc=$?"
tail -c 3000 $OUT/bench.json
python bench.py --impl reference --steps 2 --warmup 1 > $OUT/bench_ref.json 2> $OUT/bench_ref.err; echo "bench ref rc=$?"
if [ "${NCU:-1}" = "1" ]; then
  python bench.py --steps 2 --warmup 1 --no-cpu-baseline > $OUT/plain.log 2>&1 &&
  ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $OUT/launches.csv \
      python bench.py --steps 2 --warmup 1 --no-cpu-baseline > $OUT/ncu_launches.log 2>&1
  echo "ncu launches rc=$?"
  timeout ${NCU_FULL_TIMEOUT:-900} ncu --set full --clock-control none --import-source on -k regex:aji_tile -s 1 -c 1 -o $OUT/tile_full \
      python bench.py --steps 2 --warmup 1 --no-cpu-baseline > $OUT/ncu_full.log 2>&1
  echo "ncu full rc=$?"
fi
ls -la $OUT
