#!/bin/bash
# config-5 scaling points on one box: tools/gpu_scale.sh <tag> <N list e.g. "2 1">  [genomes]
set -u
TAG=$1; NS=$2; G=${3:-50000}
OUT=gpurun_out/$TAG; mkdir -p $OUT
nvidia-smi -L > $OUT/smi.txt; free -g > $OUT/mem.txt; nproc >> $OUT/mem.txt
for N in $NS; do
  if [ "$N" = "1" ]; then
    python bench.py --gpus 1 --steps 3 --warmup 3 --genomes $G --no-cpu-baseline > $OUT/bench_n1.json 2> $OUT/bench_n1.err
  else
    python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 3 --warmup 3 --genomes $G > $OUT/bench_n$N.json 2> $OUT/bench_n$N.err
  fi
  echo "bench N=$N rc=$?"; tail -c 1800 $OUT/bench_n$N.json; grep -v Warning $OUT/bench_n$N.err | tail -4
  nvidia-smi --query-gpu=memory.used --format=csv >> $OUT/mem.txt
done
