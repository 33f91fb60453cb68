"""Per-kernel CUDA-event times of the light-point-cloud pass (nemit_point_lights) in fp16 mode: its backward kernels run
without the table scatter (null gradient table), which isolates the MMA chain + epilogues of k_field_backward_tc."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from nerf_emitter_b200 import EmitterModelSpec, NerfEmitter, make_params
spec = EmitterModelSpec()
E = NerfEmitter(make_params(spec, 0, 1e3), spec)
n = int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 18
E.set_precision("fp16").set_sampler("fast")
g = torch.Generator().manual_seed(1)
o = torch.rand(3, n, generator=g).cuda(); v = torch.randn(3, n, generator=g); v = (v / v.norm(dim=0, keepdim=True)).cuda()
E.train(False)
for _ in range(2): E.point_lights(o, v)
torch.cuda.synchronize()
E.profile(True); E.profile_read(reset=True)
reps = 5
for _ in range(reps): out = E.point_lights(o, v)
torch.cuda.synchronize()
p = E.profile_read()
print({k: "%.3f ms x %d" % (ms / c, c // reps) for k, (ms, c) in p.items()}, float(out["rgb"].sum()))
