cd $GRAFT_REPO_ROOT
timeout 600 python -m pytest tests/test_gpu_round2_api.py tests/test_zz_no_alloc.py -m gpu -q -x 2>&1 | tail -2
