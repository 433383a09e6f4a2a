#!/bin/bash
# 1 x B200, final validation of HEAD: GPU tests, smoke, default bench + reference arm, ncu launch list + full capture of the tile
# kernel at config 5, and the CLI's own phase timers on a 10,000-genome synthetic database (SQLite run, cached run, no-stream run)
set -u
OUT=gpurun_out/${TAG:-r2z}; mkdir -p $OUT
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $OUT/smi.txt 2>&1
lscpu | grep -E 'Model name|^CPU\(s\)|Socket|Thread' > $OUT/cpu.txt 2>&1
timeout 1500 python -m pytest tests -m gpu -x -q -rs > $OUT/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -6 $OUT/pytest_gpu.log
timeout 120 python -c "import __graft_entry__ as g; g.smoke()" > $OUT/smoke.log 2>&1; echo "smoke rc=$?"
timeout 600 python bench.py > $OUT/bench.json 2> $OUT/bench.err; echo "bench rc=$?"; tail -c 1500 $OUT/bench.json; echo
timeout 300 python bench.py --impl reference --steps 2 --warmup 1 > $OUT/bench_ref.json 2> $OUT/bench_ref.err; echo "bench ref rc=$?"
timeout 300 python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-ceiling > $OUT/plain.log 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $OUT/launches.csv \
    python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-ceiling > $OUT/ncu_launches.log 2>&1
echo "ncu launches rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:aji_tile -s 1 -c 1 -o $OUT/tile_full \
    python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-ceiling > $OUT/ncu_full.log 2>&1
echo "ncu full rc=$?"
# ---- the CLI at config-4 size ----
if [ "${CLI:-1}" = "1" ]; then
python - <<'PY' > $OUT/cli_make_db.log 2>&1
import time, sys
sys.path.insert(0, '.')
from parfastaai_b200 import synth, fastaai_db
N, P = 10000, 122
t = time.time(); off, vals = synth.generate(N, P, seed=1234); print("generate %.1f s, |F| = %d" % (time.time() - t, off[-1]))
gn, sn = fastaai_db.synthetic_names(N, P)
t = time.time(); fastaai_db.write_db('/tmp/syn10k.db', gn, sn, off, vals); print("write_db %.1f s" % (time.time() - t))
PY
echo "make db rc=$?"; tail -2 $OUT/cli_make_db.log
X=parfastaai_b200/par_fastaai.x
( time timeout 600 $X /tmp/syn10k.db /tmp/out_a.csv --index-cache /tmp/syn10k.pfidx ) > $OUT/cli_10k_first_run.log 2>&1; echo "cli first rc=$?"
( time timeout 600 $X /tmp/syn10k.db /tmp/out_b.csv --index-cache /tmp/syn10k.pfidx ) > $OUT/cli_10k_cached_run.log 2>&1; echo "cli cached rc=$?"
( time timeout 600 $X /tmp/syn10k.db /tmp/out_c.csv --index-cache /tmp/syn10k.pfidx --no-stream-csv ) > $OUT/cli_10k_cached_nostream.log 2>&1; echo "cli nostream rc=$?"
cmp /tmp/out_a.csv /tmp/out_b.csv && cmp /tmp/out_a.csv /tmp/out_c.csv && echo "CSV bytes identical (3 runs), $(stat -c %s /tmp/out_a.csv) bytes" | tee $OUT/cli_10k_cmp.txt
grep -E "JAC Construction|CSV output|E constr|Lc & Lp|Index cache|real" $OUT/cli_10k_*.log
fi
ls -la $OUT
