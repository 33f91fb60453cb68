cd $GRAFT_REPO_ROOT
timeout 1500 python -m pytest tests -m gpu -q -x 2>&1 | grep -v "^$" | tail -3
for prec in split fp16; do
  samp=exact; [ $prec = fp16 ] && samp=fast
  timeout 300 python tools/kernel_times.py 262144 fwd+bwd $samp 3 $prec 2>&1 | tail -1 | sed "s/.*k_field_backward_tc': '\([^']*\)'.*step \(.*\)/  $prec bwd \1  step \2/"
done
