cd $GRAFT_REPO_ROOT
timeout 120 python -m pytest tests -m gpu -q -x -p no:cacheprovider -k "tiny_batch_strict" > gpurun_out/r2_quick.txt 2>&1 || { tail -5 gpurun_out/r2_quick.txt; exit 1; }
timeout 900 python -m pytest tests -m gpu -q -p no:cacheprovider -k "fp16 or benched or point_lights or turntable or round2 or full_size" > gpurun_out/r2_gpu_tests_11.txt 2>&1; echo "pytest rc=$?"
grep -E "passed|failed|^FAILED" gpurun_out/r2_gpu_tests_11.txt | tail -8
timeout 300 python tools/kernel_times.py 1048576 fwd+bwd fast 3 2>&1 | tail -1
