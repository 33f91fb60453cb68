cd $GRAFT_REPO_ROOT
timeout 600 python -m pytest tests -m gpu -q -p no:cacheprovider -k "forward or benched or round2 or model_api or op" > gpurun_out/r2_gpu_tests_12.txt 2>&1; echo "pytest rc=$?"
grep -E "passed|failed|^FAILED" gpurun_out/r2_gpu_tests_12.txt | tail -8
timeout 300 python tools/kernel_times.py 1048576 fwd+bwd fast 3 2>&1 | tail -1
