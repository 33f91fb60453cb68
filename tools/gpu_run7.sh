cd $GRAFT_REPO_ROOT
timeout 120 python -m pytest tests -m gpu -q -x -p no:cacheprovider -k "tiny_batch_strict" > gpurun_out/r2_quick.txt 2>&1 || { tail -5 gpurun_out/r2_quick.txt; exit 1; }
timeout 600 python -m pytest tests -m gpu -q -x -p no:cacheprovider -k "fp16 or benched or point_lights or turntable or round2 or full_size" > gpurun_out/r2_gpu_tests_7.txt 2>&1; echo "pytest rc=$?"
grep -E "passed|failed|^FAILED|Error" gpurun_out/r2_gpu_tests_7.txt | tail -8
for v in "" _noagg; do
  echo "variant '$v'"
  NEMIT_LIB=$PWD/nerf_emitter_b200/libnerfemitter$v.so timeout 300 python tools/kernel_times.py 1048576 fwd+bwd fast 3 2>&1 | tail -1 | grep -o "k_field_backward_tc.*"
done
timeout 300 python bench.py --steps 3 --warmup 3 --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); k=d['roofline']['kernels']['k_field_backward_tc']
print('value', d['value'], 'ms', d['ms_per_step'], 'bwd', k['ms_per_launch'], k['own_roof'])"
