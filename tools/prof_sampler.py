"""Profiling target (GPU box): forward calls with the chosen sampler (argv: rays reps sampler)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from nerf_emitter_b200 import EmitterModelSpec, NerfEmitter, make_params
spec = EmitterModelSpec()
E = NerfEmitter(make_params(spec, 0, 1e3), spec)
n = int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 16
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 3
E.set_precision("fp16")
E.set_sampler(sys.argv[3] if len(sys.argv) > 3 else "fast")
g = torch.Generator().manual_seed(1)
o = torch.rand(3, n, generator=g).cuda(); v = torch.randn(3, n, generator=g); v = (v / v.norm(dim=0, keepdim=True)).cuda()
for _ in range(reps):
    rgb = E.render(o, v)
torch.cuda.synchronize()
print("ok", float(rgb.sum()))
