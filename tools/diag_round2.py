"""GPU diagnostics of round 2 (not a test): (1) which rays of the 4096-ray default-init strict backward gate differ and why,
(2) what a one-rank NCCL PreMulSum all-reduce does to the buffer."""
import ctypes as C
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import torch
from helpers import make_rays
from nerf_emitter_b200 import EmitterModelSpec, NerfEmitter, make_params
from nerf_emitter_b200 import lib as L
from oracle.nerfacto_oracle import OracleEmitter

spec = EmitterModelSpec()
P = make_params(spec, seed=0, table_scale=1.0)
E = NerfEmitter(P, spec, device=torch.device("cuda:0"))
O = OracleEmitter(P, spec)
N = 4096
o, v, g = make_rays(N, seed=1)
gout = torch.randn(N, 3, generator=g)
for training in (True, False):
    O.train(training); E.train(training)
    ro, rv, rg, _ = O.vjp(o, v, gout, proposal_requires_grad=False)
    E.bind_grads()
    go, gv = E.backward(o.cuda(), v.cuda(), gout.cuda(), chunk_rays=N, proposal_requires_grad=False)
    per_ray = torch.maximum((go.cpu() - ro).abs().max(0).values / ro.abs().max(), (gv.cpu() - rv).abs().max(0).values / rv.abs().max())
    bad = torch.nonzero(per_ray > 1e-3).reshape(-1)
    print("training=%s: %d rays beyond 1e-3: %s" % (training, len(bad), [(int(i), "%.2e" % float(per_ray[i])) for i in bad[:12]]))
    rgb, aux = E.render(o.cuda(), v.cuda(), return_aux=True)
    rrgb, raux = O.render(o, v, return_aux=True)
    for i in bad[:6]:
        i = int(i)
        ds2 = (aux["s_edges2"][i].cpu() - raux["sdist2"][i]).abs().max()
        dw2 = (aux["weights2"][i].cpu() - raux["weights2"][i]).abs().max()
        drgb = (rgb[i].cpu() - rrgb[i]).abs().max()
        print("   ray %d: o=%s v=%s  |d s_edges2| %.2e  |d weights2| %.2e (max w %.2e)  |d rgb| %.2e  grad_o gpu %s ref %s" % (
            i, o[:, i].tolist(), v[:, i].tolist(), float(ds2), float(dw2), float(raux["weights2"][i].max()), float(drgb),
            go[:, i].cpu().tolist(), ro[:, i].tolist()))
        # same samples?
        O.edges_override = [None, aux["s_edges1"][i:i + 1].cpu(), aux["s_edges2"][i:i + 1].cpu()]
        r1o, r1v, _, _ = O.vjp(o[:, i:i + 1].contiguous(), v[:, i:i + 1].contiguous(), gout[i:i + 1], proposal_requires_grad=False)
        O.edges_override = None
        print("      oracle at the GPU's edges (1-ray chunk, reg means differ): grad_o %s" % r1o[:, 0].tolist())

# ---- one-rank NCCL PreMulSum -------------------------------------------------------------------------------------------
ident = (C.c_ubyte * L.UNIQUE_ID_BYTES)()
L.check(E.lib.nemit_comm_unique_id(ident), "id")
L.check(E.lib.nemit_comm_init(E._h, ident, 0, 1), "init")
for n in (1000, 100003, 1 << 20, 19_400_000):
    x = torch.arange(n, device="cuda", dtype=torch.float32) % 4096 + 1
    ref = x * 0.5
    L.check(E.lib.nemit_allreduce_grads(E._h, x.data_ptr(), n, 0.5, E._stream()), "allreduce")
    torch.cuda.synchronize()
    wrong = torch.nonzero(x != ref).reshape(-1)
    print("one-rank PreMulSum n=%d: %d elements not scaled%s" % (n, len(wrong), "" if len(wrong) == 0 else ", first %d last %d" % (int(wrong[0]), int(wrong[-1]))))
L.check(E.lib.nemit_comm_destroy(E._h), "destroy")
