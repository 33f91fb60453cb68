cd $GRAFT_REPO_ROOT
python tools/prof_sampler.py > gpurun_out/prof_plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:k_sample_levels -s 1 -c 1 -o gpurun_out/r2_prof_sampler2 python tools/prof_sampler.py > gpurun_out/ncu_s.log 2>&1; echo "ncu rc=$?"; tail -2 gpurun_out/ncu_s.log
