"""split (bf16 hi|lo on the tensor cores) against the fp32 FMA path on the GPU: errors and per-kernel times."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import torch
from nerf_emitter_b200 import EmitterModelSpec, NerfEmitter, make_params
from helpers import make_rays, rel_err
spec = EmitterModelSpec()
for scale in (1e3, 1.0):
    E = NerfEmitter(make_params(spec, 0, scale), spec)
    N = 6000
    o, v, g = make_rays(N, seed=4)
    gout = torch.randn(N, 3, generator=g).cuda()
    o, v = o.cuda(), v.cuda()
    res = {}
    for prec in ("fp32", "split", "fp16"):
        E.set_precision(prec).set_sampler("exact")
        E.train(False)
        rgb = E.render(o, v, near=0.01)
        E.train(True)
        E.bind_grads()
        go, gv = E.backward(o, v, gout, near=0.01, chunk_rays=N, proposal_requires_grad=True)
        torch.cuda.synchronize()
        res[prec] = (rgb.clone(), go.clone(), gv.clone(), {k: t.clone() for k, t in E.grads.items()})
    for prec in ("split", "fp16"):
        a, b = res[prec], res["fp32"]
        pr = torch.maximum((a[1] - b[1]).abs().max(0).values / b[1].abs().max(), (a[2] - b[2]).abs().max(0).values / b[2].abs().max())
        print("scale %g %s vs fp32: rgb %.2e  grad_o %.2e grad_v %.2e  rays>1e-3: %.2f%%  rays>1e-4: %.2f%%" % (
            scale, prec, rel_err(a[0], b[0]), rel_err(a[1], b[1]), rel_err(a[2], b[2]), 100 * float((pr > 1e-3).float().mean()),
            100 * float((pr > 1e-4).float().mean())))
        for k in a[3]:
            print("   %-55s %.2e" % (k, rel_err(a[3][k], b[3][k])))
# timing
E = NerfEmitter(make_params(spec, 0, 1e3), spec)
n = 1 << 18
g = torch.Generator().manual_seed(1)
o = torch.rand(3, n, generator=g).cuda(); v = torch.randn(3, n, generator=g); v = (v / v.norm(dim=0, keepdim=True)).cuda()
gout = torch.randn(n, 3, generator=g).cuda()
E.bind_grads()
for prec in ("split", "fp16", "fp32"):
    E.set_precision(prec).set_sampler("exact")
    def step():
        E.train(False); E.render(o, v)
        E.train(True); E.backward(o, v, gout, chunk_rays=16384, proposal_requires_grad=False)
    step(); torch.cuda.synchronize()
    E.profile(True); E.profile_read(reset=True)
    reps = 3
    for _ in range(reps): step()
    torch.cuda.synchronize()
    p = E.profile_read(reset=True); E.profile(False)
    p = {k: v for k, v in p.items() if isinstance(v, tuple)}
    tot = sum(ms for ms, c in p.values()) / reps
    print(prec, {k: "%.3f ms x %d" % (ms / c, c // reps) for k, (ms, c) in p.items()}, "step %.2f ms = %.2f M rays/s" % (tot, n / tot / 1e3))
