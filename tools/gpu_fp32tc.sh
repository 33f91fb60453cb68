cd $GRAFT_REPO_ROOT
timeout 600 python tools/diag_fp32tc.py 2>&1 | grep "vs fp32\|step\|rror"
timeout 900 python -m pytest tests/test_gpu_backward.py tests/test_gpu_forward.py -m gpu -q -x -k split 2>&1 | grep -v "^$" | tail -8
