#!/bin/bash
# round 2, first GPU visit: parity tests, default bench (config 5, verify + ceiling), dense family at config-4 size on the list
# kernel (K5 decision evidence) with its ncu capture.
set -u
OUT=gpurun_out/r2a; mkdir -p $OUT
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $OUT/smi.txt 2>&1
lscpu | grep -E 'Model name|^CPU\(s\)|Socket|Thread' > $OUT/cpu.txt 2>&1
df -h /dev/shm > $OUT/shm.txt 2>&1; free -g >> $OUT/shm.txt 2>&1
python -m pytest tests -m gpu -x -q > $OUT/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a $OUT/pytest_gpu.log; tail -5 $OUT/pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke()" > $OUT/smoke.log 2>&1; echo "smoke rc=$?"
python bench.py --genomes 10000 --steps 3 --warmup 3 --no-cpu-baseline > $OUT/bench_10k.json 2> $OUT/bench_10k.err; echo "bench 10k rc=$?"
python bench.py --genomes 10000 --dense --steps 3 --warmup 3 --no-cpu-baseline > $OUT/bench_10k_dense.json 2> $OUT/bench_10k_dense.err; echo "bench 10k dense rc=$?"
python bench.py > $OUT/bench.json 2> $OUT/bench.err; echo "bench rc=$?"
tail -c 2500 $OUT/bench.json
python bench.py --genomes 10000 --dense --steps 1 --warmup 1 --no-cpu-baseline --no-ceiling > $OUT/plain_dense.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:aji_tile -s 1 -c 1 -o $OUT/tile_dense_full \
    python bench.py --genomes 10000 --dense --steps 1 --warmup 1 --no-cpu-baseline --no-ceiling > $OUT/ncu_dense.log 2>&1
echo "ncu dense rc=$?"
ls -la $OUT
