cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/final_tests.log 2>&1; echo "tests rc=$?"; tail -1 gpurun_out/final_tests.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/final_smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/final_smoke.log | cut -c1-300
timeout 600 python bench.py > gpurun_out/final_bench.json 2> gpurun_out/final_bench.err; echo "bench rc=$?"
timeout 600 python bench.py --precision split --no-cpu-baseline > gpurun_out/bench_split.json 2> gpurun_out/bench_split.err; echo "bench split rc=$?"
timeout 600 python bench.py --impl reference > gpurun_out/final_bench_ref.json 2> gpurun_out/final_bench_ref.err; echo "ref rc=$?"
python - <<'PY'
import json
for f in ('gpurun_out/final_bench.json','gpurun_out/bench_split.json'):
    d=json.load(open(f))
    print(f, d['value'], d['ms_per_step'], d['e2e']['value'], d.get('forward',{}).get('value'), d.get('train_primal_backward',{}).get('value'))
    for k,v in d['roofline']['kernels'].items(): print('  ',k, round(v['ms_per_launch'],3), v['launches'], round(v['share'],3), (v.get('own_roof') or {}).get('frac'))
d=json.load(open('gpurun_out/final_bench_ref.json')); print('ref', d['value'], d['ms_per_step'])
PY
