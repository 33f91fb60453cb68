cd $GRAFT_REPO_ROOT
for v in "" rb rwb; do
  lib=""; [ -n "$v" ] && lib=nerf_emitter_b200/libnerfemitter_$v.so
  echo "== lib: ${v:-default}"
  for prec in split fp16; do
    samp=exact; [ $prec = fp16 ] && samp=fast
    NEMIT_LIB=$lib timeout 300 python tools/kernel_times.py 262144 fwd+bwd $samp 3 $prec 2>&1 | tail -1 | sed "s/.*k_field_backward_tc': '\([^']*\)'.*step \(.*\)/  $prec bwd \1  step \2/"
  done
done
