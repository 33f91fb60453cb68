cd $GRAFT_REPO_ROOT
timeout 600 python tools/diag_fp32tc2.py 2>&1 | grep -A13 "split" | grep -v fp16 | head -60
timeout 600 python tools/diag_fp32tc.py 2>&1 | grep "vs fp32\|step"
timeout 900 python -m pytest tests/test_gpu_backward.py tests/test_gpu_forward.py -m gpu -q -x -k split 2>&1 | grep -v "^$" | tail -30
