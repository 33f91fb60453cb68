"""Per-kernel CUDA-event times of the fp16-mode step (GPU box): argv = rays [mode fwd|fwd+bwd] [sampler] [reps]."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from nerf_emitter_b200 import EmitterModelSpec, NerfEmitter, make_params
spec = EmitterModelSpec()
E = NerfEmitter(make_params(spec, 0, 1e3), spec)
n = int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 18
mode = sys.argv[2] if len(sys.argv) > 2 else "fwd"
E.set_precision(sys.argv[5] if len(sys.argv) > 5 else "fp16")
E.set_sampler(sys.argv[3] if len(sys.argv) > 3 else "fast")
reps = int(sys.argv[4]) if len(sys.argv) > 4 else 5
g = torch.Generator().manual_seed(1)
o = torch.rand(3, n, generator=g).cuda(); v = torch.randn(3, n, generator=g); v = (v / v.norm(dim=0, keepdim=True)).cuda()
gout = torch.randn(n, 3, generator=g).cuda()
if mode != "fwd":
    E.bind_grads()
def step():
    E.train(False); rgb = E.render(o, v)
    if mode != "fwd":
        E.train(True); E.backward(o, v, gout, chunk_rays=16384, proposal_requires_grad=False)
    return rgb
for _ in range(2): step()
torch.cuda.synchronize()
E.profile(True); E.profile_read(reset=True)
for _ in range(reps): rgb = step()
torch.cuda.synchronize()
p = E.profile_read()
print({k: "%.3f ms x %d" % (ms / c, c // reps) for k, (ms, c) in p.items()}, "step %.2f ms" % (sum(ms for ms, c in p.values()) / reps), float(rgb.sum()))
