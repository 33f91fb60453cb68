cd $GRAFT_REPO_ROOT
for v in "" sg3; do
  lib=""; [ -n "$v" ] && lib=nerf_emitter_b200/libnerfemitter_$v.so
  echo "== lib: ${v:-default}"
  NEMIT_LIB=$lib timeout 300 python tools/kernel_times.py 262144 fwd+bwd fast 3 fp16 2>&1 | tail -1 | sed "s/.*k_field_backward_tc': '\([^']*\)'.*step \(.*\)/  fp16 bwd \1  step \2/"
done
NEMIT_LIB=nerf_emitter_b200/libnerfemitter_sg3.so timeout 600 python -m pytest tests/test_gpu_benched_mode.py tests/test_gpu_backward.py -m gpu -q -x 2>&1 | tail -2
