cd $GRAFT_REPO_ROOT
for v in _only03 _only47 _only07 _only1215; do
  echo "variant '$v'"
  NEMIT_LIB=$PWD/nerf_emitter_b200/libnerfemitter$v.so timeout 300 python tools/kernel_times.py 1048576 fwd+bwd fast 3 2>&1 | tail -1 | grep -o "k_field_backward_tc[^,]*"
done
python tools/prof_backward.py 65536 2 > gpurun_out/prof_plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:k_field_backward_tc -s 1 -c 1 -o gpurun_out/r2_prof_bwd_ws python tools/prof_backward.py 65536 2 > gpurun_out/ncu_bwd.log 2>&1; echo "ncu rc=$?"; tail -3 gpurun_out/ncu_bwd.log
