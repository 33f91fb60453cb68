cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
python tools/prof_backward.py 262144 1 > gpurun_out/prof_plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:"k_sample_levels|k_field_forward_tc|k_field_backward_tc" -c 5 -o gpurun_out/r2_prof_step python tools/prof_backward.py 262144 1 > gpurun_out/ncu_step.log 2>&1; echo "ncu rc=$?"; tail -2 gpurun_out/ncu_step.log
