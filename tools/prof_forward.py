"""Profiling target (GPU box): a few forward calls (one launch per kernel per call for n <= 2^18)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from nerf_emitter_b200 import EmitterModelSpec, NerfEmitter, make_params
spec = EmitterModelSpec()
E = NerfEmitter(make_params(spec, 0, 1e3), spec)
n = int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 18
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 3
E.set_precision(sys.argv[3] if len(sys.argv) > 3 else "fp32")
g = torch.Generator().manual_seed(1)
o = torch.rand(3, n, generator=g).cuda(); v = torch.randn(3, n, generator=g); v = (v / v.norm(dim=0, keepdim=True)).cuda()
for _ in range(reps):
    rgb = E.render(o, v)
torch.cuda.synchronize()
print("ok", float(rgb.sum()))
