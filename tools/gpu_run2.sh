set -x
cd $GRAFT_REPO_ROOT
timeout 1500 python -m pytest tests -m gpu -q -s -rA -p no:cacheprovider > gpurun_out/r2_gpu_tests_2.txt 2>&1; echo "pytest rc=$?"
grep -E "passed|failed" gpurun_out/r2_gpu_tests_2.txt | tail -3
grep -E "^FAILED" gpurun_out/r2_gpu_tests_2.txt
timeout 600 python tools/diag_round2.py > gpurun_out/r2_diag.txt 2>&1; echo "diag rc=$?"; cat gpurun_out/r2_diag.txt | cut -c1-600
timeout 120 tools/tma_gather_bench 1 > gpurun_out/r2_tma_gather_box1.txt 2>&1; echo "tma1 rc=$?"; cat gpurun_out/r2_tma_gather_box1.txt
