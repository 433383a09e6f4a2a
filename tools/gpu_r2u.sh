#!/bin/bash
set -u
OUT=gpurun_out/${TAG:-r2u}; mkdir -p $OUT
timeout 600 python bench.py > $OUT/bench.json 2> $OUT/bench.err; echo "bench rc=$?"; tail -c 600 $OUT/bench.json; echo
timeout 200 python bench.py --impl reference --steps 1 --warmup 1 > $OUT/bench_ref.json 2> $OUT/bench_ref.err; echo "bench ref rc=$?"; tail -c 700 $OUT/bench_ref.json; echo
