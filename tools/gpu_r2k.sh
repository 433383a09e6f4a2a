#!/bin/bash
# 1 x B200: validation of HEAD (GPU tests incl. the reference's Catch2 cases over the binding, smoke, default bench),
# K5 forced on the SPARSE family (patch cost data point), ncu capture of K5 on the dense family, compute-sanitizer passes.
set -u
OUT=gpurun_out/${TAG:-r2k}; mkdir -p $OUT
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $OUT/smi.txt 2>&1
lscpu | grep -E 'Model name|^CPU\(s\)|Socket|Thread' > $OUT/cpu.txt 2>&1
timeout 900 python -m pytest tests -m gpu -x -q > $OUT/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 $OUT/pytest_gpu.log
timeout 120 python -c "import __graft_entry__ as g; g.smoke()" > $OUT/smoke.log 2>&1; echo "smoke rc=$?"
timeout 600 python bench.py > $OUT/bench.json 2> $OUT/bench.err; echo "bench rc=$?"
show() { python - "$1" <<'PY'
import json,sys
try:
    d=json.load(open(sys.argv[1]))
    print("  value %.3e e2e %.3e ms %.2f"%(d["value"],d["e2e"]["value"],d["ms_per_step"]), d["phases_ms"], d.get("path"), "verify", d["verify"]["max_abs_diff"], "frac_onchip", d.get("roofline_onchip",{}).get("frac"))
except Exception as e: print("  no json", e); print(open(sys.argv[1].replace(".json",".err")).read()[-1500:])
PY
}
show $OUT/bench.json
for tau in 64 512; do
  timeout 300 python bench.py --genomes 10000 --path mma --tau $tau --steps 3 --warmup 3 --no-cpu-baseline --no-ceiling > $OUT/bench_10k_sparse_mma_tau$tau.json 2> $OUT/bench_10k_sparse_mma_tau$tau.err
  echo "sparse mma tau=$tau rc=$?"; show $OUT/bench_10k_sparse_mma_tau$tau.json
done
timeout 240 python bench.py --genomes 10000 --dense --path mma --steps 3 --warmup 3 --no-cpu-baseline > $OUT/bench_10k_dense_mma.json 2> $OUT/bench_10k_dense_mma.err
echo "dense mma rc=$?"; show $OUT/bench_10k_dense_mma.json
timeout 200 python bench.py --genomes 10000 --dense --path mma --steps 1 --warmup 1 --no-cpu-baseline --no-ceiling > $OUT/plain_mma.log 2>&1 &&
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $OUT/launches_dense_mma.csv \
    python bench.py --genomes 10000 --dense --path mma --steps 1 --warmup 1 --no-cpu-baseline --no-ceiling > $OUT/ncu_launches_mma.log 2>&1
echo "ncu launches mma rc=$?"
timeout 400 ncu --set full --clock-control none --import-source on -k regex:aji_mma -s 1 -c 1 -o $OUT/mma_dense_full \
    python bench.py --genomes 10000 --dense --path mma --steps 1 --warmup 1 --no-cpu-baseline --no-ceiling > $OUT/ncu_mma.log 2>&1
echo "ncu mma rc=$?"
# compute-sanitizer: memcheck + racecheck over smoke(), memcheck over the partitioned-build parity test and one K5 test
timeout 600 compute-sanitizer --tool memcheck --log-file $OUT/sanitizer_memcheck_smoke.log python -c "import __graft_entry__ as g; g.smoke()" > $OUT/san1.out 2>&1; echo "memcheck smoke rc=$?"
timeout 600 compute-sanitizer --tool racecheck --log-file $OUT/sanitizer_racecheck_smoke.log python -c "import __graft_entry__ as g; g.smoke()" > $OUT/san2.out 2>&1; echo "racecheck smoke rc=$?"
timeout 900 compute-sanitizer --tool memcheck --log-file $OUT/sanitizer_memcheck_partitioned.log python -m pytest tests/test_gpu_parity.py -x -q -k "partitioned_build_matches" > $OUT/san3.out 2>&1; echo "memcheck partitioned rc=$?"
timeout 600 compute-sanitizer --tool memcheck --log-file $OUT/sanitizer_memcheck_mma.log python -m pytest tests/test_gpu_mma.py -x -q -k "golden" > $OUT/san4.out 2>&1; echo "memcheck mma rc=$?"
for f in $OUT/sanitizer_*.log; do echo $f; tail -2 $f; done
ls -la $OUT
