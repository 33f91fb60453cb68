cd $GRAFT_REPO_ROOT
timeout 900 python -m pytest tests -m gpu -q -p no:cacheprovider > gpurun_out/r2_gpu_tests_10.txt 2>&1; echo "pytest rc=$?"
grep -E "passed|failed|^FAILED" gpurun_out/r2_gpu_tests_10.txt | tail -8
timeout 300 python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/r2_bench_10.json 2>/dev/null; python -c "
import json
d=json.load(open('gpurun_out/r2_bench_10.json'))
print('value', d['value'], 'ms', d['ms_per_step'], 'e2e', d['e2e']['value'], 'fwd', d['forward']['value'], 'kept', d['train_primal_backward']['value'])
for k,v in d['roofline']['kernels'].items(): print(k, round(v['ms_per_launch'],3), round(v['share'],3), (v.get('own_roof') or {}).get('frac'))"
