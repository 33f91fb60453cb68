"""Where do the table gradients of the tensor-core modes differ from the fp32 kernels'?  Default init, same GPU sample edges."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import torch
from nerf_emitter_b200 import EmitterModelSpec, NerfEmitter, make_params
from helpers import make_rays, rel_err
spec = EmitterModelSpec()
SCALE = float(sys.argv[2]) if len(sys.argv) > 2 else 1.0
E = NerfEmitter(make_params(spec, 0, SCALE), spec)
N = int(sys.argv[1]) if len(sys.argv) > 1 else 64
o, v, g = make_rays(N, seed=1)
gout = torch.randn(N, 3, generator=g).cuda()
o, v = o.cuda(), v.cuda()
res = {}
for prec in ("fp32", "split", "fp16"):
    E.set_precision(prec).set_sampler("exact")
    E.train(True)
    E.bind_grads()
    E.backward(o, v, gout, chunk_rays=N, proposal_requires_grad=False)
    torch.cuda.synchronize()
    res[prec] = E.grads["field.mlp_base_grid.hash_table"].clone().reshape(16, -1, 2)
b = res["fp32"]
print("max |ref| per level:", ["%.2e" % float(b[l].abs().max()) for l in range(16)])
for prec in ("split", "fp16"):
    a = res[prec]
    d = (a - b).abs().max(dim=2).values          # (16, T)
    gmax = float(b.abs().max())
    print(prec, "rel err", float(d.max()) / gmax)
    for l in range(16):
        bad = (d[l] > 1e-4 * gmax).nonzero().flatten()
        if bad.numel():
            top = torch.topk(d[l], min(4, d[l].numel()))
            print("  level %2d: %4d entries off by > 1e-4 of max; worst: %s" % (l, bad.numel(), ", ".join(
                "idx %d ref %.3e got %.3e" % (int(i), float(b[l, i].abs().max()), float(a[l, i].abs().max())) for i in top.indices[:3])))
    nz_ref, nz_got = (b.abs().sum(2) != 0), (a.abs().sum(2) != 0)
    print("  entries non-zero in ref only: %d, in %s only: %d" % (int((nz_ref & ~nz_got).sum()), prec, int((nz_got & ~nz_ref).sum())))
if N > 4000:
    sys.exit(0)
# which side agrees with the CPU oracle evaluated at the GPU's own sample edges?
from oracle.nerfacto_oracle import OracleEmitter
O = OracleEmitter(make_params(spec, 0, SCALE), spec)
O.train(True)
E.set_precision("fp32")
aux = E.sample_levels(o, v)
torch.cuda.synchronize()
O.edges_override = [None, aux["s_edges1"].cpu(), aux["s_edges2"].cpu()]
_, _, rg, _ = O.vjp(o.cpu(), v.cpu(), gout.cpu(), proposal_requires_grad=False)
ref = rg["field.mlp_base_grid.hash_table"].reshape(16, -1, 2)
gmax = float(ref.abs().max())
for prec in ("fp32", "split"):
    d = (res[prec].cpu() - ref).abs().max(dim=2).values
    print(prec, "vs oracle: rel err %.3e; entries off by > 1e-4 of max per level:" % (float(d.max()) / gmax),
          [int((d[l] > 1e-4 * gmax).sum()) for l in range(16)])
