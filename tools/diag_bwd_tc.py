"""Diagnostic (GPU box): fp16 tensor-core backward vs the fp32 backward, per tensor, plus timing of both."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import torch
from helpers import make_rays, rel_err
from nerf_emitter_b200 import EmitterModelSpec, NerfEmitter, make_params

spec = EmitterModelSpec()
for scale in (1e3, 1.0):
    E = NerfEmitter(make_params(spec, 0, scale), spec)
    E.train(True)
    for N in (2, 1001, 8192):
        o, v, g = make_rays(N, seed=3)
        jit = [torch.rand(N, 1, generator=g) for _ in range(3)]
        gout = torch.randn(N, 3, generator=g)
        res = {}
        for prec in ("fp32", "fp16"):
            E.set_precision(prec)
            E.bind_grads()
            go, gv = E.backward(o.cuda(), v.cuda(), gout.cuda(), jitter=jit, chunk_rays=N, proposal_requires_grad=False)
            torch.cuda.synchronize()
            res[prec] = (go.cpu(), gv.cpu(), {k: t.cpu().clone() for k, t in E.grads.items()})
        a, b = res["fp32"], res["fp16"]
        per_ray = torch.maximum((a[0] - b[0]).abs().max(0).values / a[0].abs().max(),
                                (a[1] - b[1]).abs().max(0).values / a[1].abs().max())
        print("== table_scale %g N %d: rays within 2e-2: %.1f%%, within 1e-2: %.1f%%; grad_o rel %.3e grad_v rel %.3e" % (
            scale, N, 100 * float((per_ray < 2e-2).float().mean()), 100 * float((per_ray < 1e-2).float().mean()),
            rel_err(b[0], a[0]), rel_err(b[1], a[1])))
        for k in a[2]:
            if "proposal" in k:
                continue
            x, y = a[2][k].double().flatten(), b[2][k].double().flatten()
            cos = float((x * y).sum() / (x.norm() * y.norm() + 1e-300))
            print("   %-50s max-rel %.3e  cos %.6f  |fp32|max %.3e nnz %d/%d" % (k, rel_err(y, x), cos, float(x.abs().max()),
                                                                             int((y != 0).sum()), int((x != 0).sum())))
E = NerfEmitter(make_params(spec, 0, 1e3), spec)
E.train(True)
N = 1 << 18
o, v, g = make_rays(N, seed=5)
gout = torch.randn(N, 3, generator=g).cuda()
o, v = o.cuda(), v.cuda()
E.bind_grads()
for prec in ("fp16", "fp32"):
    E.set_precision(prec)
    E.profile(True)
    for _ in range(3):
        E.backward(o, v, gout, chunk_rays=16384, proposal_requires_grad=False)
    torch.cuda.synchronize()
    print(prec, {k: "%.2f ms x%d" % (ms / c, c) for k, (ms, c) in E.profile_read(reset=True).items()})
