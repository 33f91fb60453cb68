cd $GRAFT_REPO_ROOT
for c in 2 3 4 5; do
  timeout 600 python bench.py --config $c --steps 3 --warmup 3 > gpurun_out/r2_bench_config${c}_1gpu.json 2> gpurun_out/r2_bench_config${c}_1gpu.err; echo "config $c rc=$?"
  python -c "
import json
d=json.load(open('gpurun_out/r2_bench_config${c}_1gpu.json'))
print({k:d[k] for k in ('metric','value','unit','ms_per_step','scaling','gpu_launches')}, 'e2e', d['e2e']['value'], 'cpu', d.get('cpu_baseline',{}).get('value'))"
done
