cd $GRAFT_REPO_ROOT
for v in "" _sg2; do
  echo "variant '$v'"
  export NEMIT_LIB=$PWD/nerf_emitter_b200/libnerfemitter$v.so
  timeout 200 python -m pytest tests -m gpu -q -x -p no:cacheprovider -k "tiny_batch_strict or test_fp16_mode_backward_tensor_core or benched_mode_backward_given_same" 2>&1 | tail -1
  timeout 300 python tools/kernel_times.py 1048576 fwd+bwd fast 3 2>&1 | tail -1 | grep -o "k_field_backward_tc.*"
done
