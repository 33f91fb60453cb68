cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x 2>&1 | grep -v "^$" | tail -3
timeout 600 python bench.py --precision split --no-cpu-baseline > gpurun_out/bench_split.json 2> gpurun_out/bench_split.err; echo "bench split rc=$?"
python - <<'PY'
import json
for f in ('gpurun_out/bench_split.json',):
    d=json.load(open(f))
    print(f, d['value'], d['ms_per_step'], d['e2e']['value'], d.get('forward',{}).get('value'), d.get('train_primal_backward',{}).get('value'))
    for k,v in d['roofline']['kernels'].items(): print('  ',k, round(v['ms_per_launch'],3), v['launches'], round(v['share'],3))
PY
