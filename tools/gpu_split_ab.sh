cd $GRAFT_REPO_ROOT
timeout 900 python -m pytest tests/test_gpu_turntable.py tests/test_gpu_point_lights.py tests/test_gpu_rotater.py -m gpu -q -x 2>&1 | grep -v "^$" | tail -12
