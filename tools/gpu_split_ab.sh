cd $GRAFT_REPO_ROOT
timeout 300 python tools/kernel_times.py 262144 fwd+bwd fast 3 fp16 2>&1 | tail -1
timeout 300 python tools/kernel_times.py 262144 fwd+bwd exact 3 split 2>&1 | tail -1
timeout 1500 python -m pytest tests -m gpu -q -x 2>&1 | tail -2
