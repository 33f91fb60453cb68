cd $GRAFT_REPO_ROOT
for v in "" f4 f1; do
  lib=""; [ -n "$v" ] && lib=nerf_emitter_b200/libnerfemitter_$v.so
  echo "== lib: ${v:-default}"
  NEMIT_LIB=$lib timeout 300 python tools/kernel_times.py 262144 fwd+bwd fast 3 fp16 2>&1 | tail -1 | cut -c1-260
done
