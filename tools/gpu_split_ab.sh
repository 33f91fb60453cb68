cd $GRAFT_REPO_ROOT
for sc in 1.0 1000.0; do
  echo "== table scale $sc"
  timeout 300 python tools/diag_table_grad.py 30000 $sc 2>&1 | grep -v "^max" | grep "split\|level\|non-zero" | head -24
done
