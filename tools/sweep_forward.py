"""Diagnostic (GPU box): CUDA-event time of the forward kernels for several batch sizes."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from nerf_emitter_b200 import EmitterModelSpec, NerfEmitter, make_params
spec = EmitterModelSpec()
E = NerfEmitter(make_params(spec, 0, 1e3), spec)
E.train(False)


def timed(fn, reps=5):
    fn(); torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps):
        fn()
    b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) / reps


for lg in [14, 16, 17, 18, 19, 20]:
    n = 1 << lg
    g = torch.Generator().manual_seed(1)
    o = torch.rand(3, n, generator=g).cuda(); v = torch.randn(3, n, generator=g); v = (v / v.norm(dim=0, keepdim=True)).cuda()
    ts = timed(lambda: E.sample_levels(o, v, edges_only=True))
    tf = timed(lambda: E.render(o, v))
    print("n=2^%d sample %.3f ms (%.1f Mray/s)  render %.3f ms (%.1f Mray/s)" % (lg, ts, n / ts / 1e3, tf, n / tf / 1e3), flush=True)
