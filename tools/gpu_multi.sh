#!/bin/bash
# N-GPU visit: tools/gpu_multi.sh <tag> <ngpus> [genomes]  -- parity tests, then the sharded bench with --verify
set -u
TAG=$1; N=$2; G=${3:-6000}
OUT=gpurun_out/$TAG; mkdir -p $OUT
nvidia-smi -L > $OUT/smi.txt; free -g > $OUT/mem.txt; nproc >> $OUT/mem.txt
python -m pytest tests -m gpu -x -q > $OUT/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 $OUT/pytest_gpu.log
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 3 --warmup 3 --genomes $G --verify > $OUT/bench_n$N.json 2> $OUT/bench_n$N.err
echo "bench N=$N rc=$?"; tail -c 2500 $OUT/bench_n$N.json; tail -5 $OUT/bench_n$N.err
python bench.py --steps 3 --warmup 3 --genomes $G --no-cpu-baseline --verify > $OUT/bench_n1.json 2> $OUT/bench_n1.err
echo "bench N=1 rc=$?"; tail -c 2500 $OUT/bench_n1.json
