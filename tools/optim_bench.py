"""Optimizer-step timing on the GPU box: nemit_adam_step (2 launches over 19.4 M parameters) vs its HBM stream bound."""
import sys, os, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from nerf_emitter_b200 import EmitterModelSpec, NerfEmitter, make_params
from nerf_emitter_b200.optim import EmitterOptimizers
spec = EmitterModelSpec()
E = NerfEmitter(make_params(spec, 0, 1e3), spec)
E.bind_grads()
opt = EmitterOptimizers(E)
n = sum(p.numel() for p in E.params.values())
E.flat_grads.normal_(0, 1e-3)
for zero in (False, True):
    for _ in range(3):
        opt.optimizer_scaler_step_all(1.0, zero_grads=zero)
    torch.cuda.synchronize()
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
    reps = 20
    ev[0].record()
    for _ in range(reps):
        opt.optimizer_scaler_step_all(1.0, zero_grads=zero)
    ev[1].record()
    torch.cuda.synchronize()
    ms = ev[0].elapsed_time(ev[1]) / reps
    bytes_ = n * (4 + 28 + (4 if zero else 0))      # stats pass reads g; step reads g,p,m,v and writes p,m,v (+ g)
    peak = json.load(open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")))["hbm_gbs"] if os.path.exists(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")) else 6650.0
    print("adam step zero_grads=%d: %.4f ms for %d params, %.0f GB/s algorithmic = %.2f of %.0f GB/s" % (zero, ms, n, bytes_ / ms * 1e-6, bytes_ / ms * 1e-6 / peak, peak))
