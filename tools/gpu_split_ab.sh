cd $GRAFT_REPO_ROOT
for lib in "" nerf_emitter_b200/libnerfemitter_iss1.so; do
  echo "== lib: ${lib:-default}"
  for prec in split fp16; do
    samp=exact; [ $prec = fp16 ] && samp=fast
    NEMIT_LIB=$lib timeout 300 python tools/kernel_times.py 262144 fwd+bwd $samp 3 $prec 2>&1 | tail -1
  done
done
timeout 900 python -m pytest tests/test_gpu_backward.py tests/test_gpu_benched_mode.py -m gpu -q -x 2>&1 | grep -v "^$" | tail -4
