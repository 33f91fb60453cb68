"""Diagnostic (GPU box): fp16 tensor-core field kernel vs the fp32 kernel and the CPU oracle (PSNR, timing)."""
import sys, os, math
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from nerf_emitter_b200 import EmitterModelSpec, NerfEmitter, make_params
spec = EmitterModelSpec()
for scale in (1e3, 1.0):
    P = make_params(spec, 0, scale)
    E = NerfEmitter(P, spec)
    E.train(False)
    n = 4096
    g = torch.Generator().manual_seed(1)
    o = torch.rand(3, n, generator=g).cuda(); v = torch.randn(3, n, generator=g); v = (v / v.norm(dim=0, keepdim=True)).cuda()
    ref = E.render(o, v)
    E.set_precision("fp16")
    out = E.render(o, v)
    torch.cuda.synchronize()
    mse = float(((out - ref) ** 2).mean()); peak = float(ref.max())
    print("table_scale %g: max|fp32| %.4e  max abs diff %.3e  rel %.3e  PSNR(peak=max) %.1f dB  nan %d" % (
        scale, peak, float((out - ref).abs().max()), float((out - ref).abs().max()) / peak,
        10 * math.log10(peak * peak / max(mse, 1e-30)), int(torch.isnan(out).sum())), flush=True)
    E.set_precision("fp32")

P = make_params(spec, 0, 1e3)
E = NerfEmitter(P, spec); E.train(False)
n = 1 << 20
g = torch.Generator().manual_seed(1)
o = torch.rand(3, n, generator=g).cuda(); v = torch.randn(3, n, generator=g); v = (v / v.norm(dim=0, keepdim=True)).cuda()
for prec in ("fp32", "fp16"):
    E.set_precision(prec)
    E.render(o, v); torch.cuda.synchronize()
    E.profile(True); E.profile_read()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(5):
        E.render(o, v)
    b.record(); torch.cuda.synchronize()
    ms = a.elapsed_time(b) / 5
    print(prec, "render 2^20 rays: %.2f ms (%.1f Mray/s)" % (ms, n / ms / 1e3), {k: round(t / 5, 2) for k, (t, c) in E.profile_read().items()}, flush=True)
