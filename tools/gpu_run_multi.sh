# usage: bash tools/gpu_run_multi.sh N "<configs>"   (configs: "default 3 4 5")
cd $GRAFT_REPO_ROOT
N=$1
for c in $2; do
  if [ "$c" = "default" ]; then flag=""; else flag="--config $c"; fi
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 5 --warmup 3 $flag > gpurun_out/r2_bench_${c}_${N}gpu.json 2> gpurun_out/r2_bench_${c}_${N}gpu.err; echo "N=$N config $c rc=$?"
  python -c "
import json
d=json.load(open('gpurun_out/r2_bench_${c}_${N}gpu.json'))
print({k:d[k] for k in ('value','unit','ms_per_step','scaling','n_gpus')}, 'e2e', d['e2e']['value'], 'check', d.get('check'))" || tail -5 gpurun_out/r2_bench_${c}_${N}gpu.err
done
