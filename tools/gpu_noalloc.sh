cd $GRAFT_REPO_ROOT
timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -q -s 2>&1 | grep -v "^$" | tail -12
echo "---- full suite"
timeout 1500 python -m pytest tests -m gpu -q -x 2>&1 | grep -v "^$" | tail -6
