mkdir -p gpurun_out/r3g
for V in "" "--sliced-index"; do
  T=$(echo x$V | tr -d " -")
  python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus 2 --steps 2 --warmup 3 --genomes 20000 $V > gpurun_out/r3g/bench_n2_$T.json 2> gpurun_out/r3g/bench_n2_$T.err
  echo "rc=$?"
  python -c "
import json,sys
d=json.load(open(sys.argv[1])); print(d['value'], d['e2e']['value'], d['ms_per_step'], d['verify']['max_abs_diff'], d['verify']['cells']); print(d['phases_ms'])" gpurun_out/r3g/bench_n2_$T.json || tail -20 gpurun_out/r3g/bench_n2_$T.err
done
