"""Where does the split mode's weight-gradient error sit?  Default init, 4096 rays, vs the fp32 FMA kernels."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import torch
from nerf_emitter_b200 import EmitterModelSpec, NerfEmitter, make_params
from helpers import make_rays, rel_err
spec = EmitterModelSpec()
E = NerfEmitter(make_params(spec, 0, 1.0), spec)
for N in (4096, 512, 64):
    o, v, g = make_rays(N, seed=1)
    gout = torch.randn(N, 3, generator=g).cuda()
    o, v = o.cuda(), v.cuda()
    res = {}
    for prec in ("fp32", "split", "fp16"):
        E.set_precision(prec).set_sampler("exact")
        E.train(True)
        E.bind_grads()
        go, gv = E.backward(o, v, gout, chunk_rays=N, proposal_requires_grad=False)
        torch.cuda.synchronize()
        res[prec] = {k: t.clone() for k, t in E.grads.items() if k.startswith("field")}
    for prec in ("split", "fp16"):
        print("N", N, prec)
        for k, a in res[prec].items():
            b = res["fp32"][k]
            if "hash" in k:
                print("  %-45s %.2e" % (k, rel_err(a, b))); continue
            d = (a - b).abs()
            line = "  %-45s %.2e  max|ref| %.2e" % (k, rel_err(a, b), float(b.abs().max()))
            if a.dim() == 2:
                colerr = d.max(0).values / b.abs().max()
                top = torch.topk(colerr, min(4, colerr.numel()))
                line += "  worst cols " + ", ".join("%d:%.1e" % (int(i), float(e)) for e, i in zip(top.values, top.indices))
                rowerr = d.max(1).values / b.abs().max()
                top = torch.topk(rowerr, min(3, rowerr.numel()))
                line += "  worst rows " + ", ".join("%d:%.1e" % (int(i), float(e)) for e, i in zip(top.values, top.indices))
            print(line)
