cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 600 python bench.py --precision split --no-cpu-baseline > gpurun_out/bench_split.json 2> gpurun_out/bench_split.err; echo "bench rc=$?"; python - <<'PY'
import json
d=json.load(open('gpurun_out/bench_split.json'))
print(d['value'], d['ms_per_step'], d['e2e']['value'])
for k,v in d['roofline']['kernels'].items(): print(k, round(v['ms_per_launch'],3), v['launches'], round(v['share'],3))
print(d.get('forward',{}).get('value'), d.get('train_primal_backward',{}).get('value'))
PY
python tools/prof_backward.py 32768 2 split > gpurun_out/prof_plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:k_field_backward_tc -s 1 -c 1 -o gpurun_out/r2_prof_bwd_split python tools/prof_backward.py 32768 2 split > gpurun_out/ncu_bwd.log 2>&1; echo "ncu rc=$?"; tail -2 gpurun_out/ncu_bwd.log
