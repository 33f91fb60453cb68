cd $GRAFT_REPO_ROOT
timeout 600 python -m pytest tests/test_gpu_rotater.py -q -s -p no:cacheprovider > gpurun_out/r2_gpu_tests_rot.txt 2>&1; echo "pytest rc=$?"
grep -E "passed|failed|^FAILED|PSNR|Error|assert" gpurun_out/r2_gpu_tests_rot.txt | head -30
