"""Profiling target (GPU box): a few fwd+bwd calls in fp16 mode (one launch per kernel per call for n <= 2^18)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from nerf_emitter_b200 import EmitterModelSpec, NerfEmitter, make_params
spec = EmitterModelSpec()
E = NerfEmitter(make_params(spec, 0, 1e3), spec)
n = int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 16
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 3
E.set_precision(sys.argv[3] if len(sys.argv) > 3 else "fp16")
prg = bool(int(sys.argv[4])) if len(sys.argv) > 4 else False
g = torch.Generator().manual_seed(1)
o = torch.rand(3, n, generator=g).cuda(); v = torch.randn(3, n, generator=g); v = (v / v.norm(dim=0, keepdim=True)).cuda()
gout = torch.randn(n, 3, generator=g).cuda()
E.bind_grads()
for _ in range(reps):
    E.train(False)
    rgb = E.render(o, v)
    E.train(True)
    go, gv = E.backward(o, v, gout, chunk_rays=16384, proposal_requires_grad=prg)
torch.cuda.synchronize()
print("ok", float(rgb.sum()), float(go.sum()))
