cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
python tools/prof_backward.py 65536 2 fp16 > gpurun_out/prof_plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:"k_field_backward_tc" -s 1 -c 1 -o gpurun_out/r2_prof_bwd_fp16 python tools/prof_backward.py 65536 2 fp16 > gpurun_out/ncu_bwd16.log 2>&1; echo "ncu rc=$?"; tail -2 gpurun_out/ncu_bwd16.log
