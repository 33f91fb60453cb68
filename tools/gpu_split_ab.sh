cd $GRAFT_REPO_ROOT
for v in "" pair2 pair3; do
  lib=""; [ -n "$v" ] && lib=nerf_emitter_b200/libnerfemitter_$v.so
  echo "== lib: ${v:-default}"
  NEMIT_LIB=$lib timeout 300 python tools/kernel_times.py 1048576 fwd fast 5 fp16 2>&1 | tail -1
done
