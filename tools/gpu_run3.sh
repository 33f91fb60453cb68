set -x
cd $GRAFT_REPO_ROOT
timeout 900 python -m pytest tests -m gpu -q -x -p no:cacheprovider -k "fp16 or benched or point_lights or turntable or round2 or full_size" > gpurun_out/r2_gpu_tests_3.txt 2>&1; echo "pytest rc=$?"
grep -E "passed|failed|^FAILED|Error" gpurun_out/r2_gpu_tests_3.txt | tail -8
timeout 300 python tools/kernel_times.py > gpurun_out/r2_kernel_times_3.txt 2>&1; cat gpurun_out/r2_kernel_times_3.txt
