cd $GRAFT_REPO_ROOT
timeout 600 python -m pytest tests/test_gpu_backward.py -m gpu -q -x -s -k multi_chunk 2>&1 | grep -v "^$" | tail -32
