cd $GRAFT_REPO_ROOT
for v in "" _sw4 _sw2u1; do
  echo "variant '$v'"
  NEMIT_LIB=$PWD/nerf_emitter_b200/libnerfemitter$v.so timeout 300 python tools/kernel_times.py 1048576 fwd+bwd fast 3 2>&1 | tail -1
done
