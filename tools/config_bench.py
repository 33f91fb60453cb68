"""BASELINE.json configs 4 and 5 on the GPU box (one GPU; CUDA events, 1 warm-up + 3 timed repetitions, fp16 mode):
  config 4: vMF-mixture importance sampling + pdf, 64 lobes, 16 M samples, combined with the NeRF radiance eval
  config 5: spherical-sensor environment-map bake 2048 x 1024 x 16 spp from the scene centre (gaussian film)
Prints one line each; the committed copy is profiles/r1_config4_config5.txt."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from nerf_emitter_b200 import EmitterModelSpec, NerfEmitter, VmfMixture, make_params
from nerf_emitter_b200.bake import bake_envmap


def make_lobes(K=64, seed=3):
    """SURVEY.md 8d config 4 synthetic lobe table: positions U([-1,2]^3), std U(0.05,0.5), weights ~ Dirichlet(1)."""
    g = torch.Generator().manual_seed(seed)
    position = torch.rand(K, 3, generator=g) * 3 - 1
    std = torch.rand(K, generator=g) * 0.45 + 0.05
    e = -torch.log(torch.rand(K, generator=g))
    return position, e / e.sum(), std


spec = EmitterModelSpec()
E = NerfEmitter(make_params(spec, 0, 1e3), spec).set_precision("fp16").set_sampler("fast").eval()
dev = E.device


def timed(fn, reps=3):
    fn()
    torch.cuda.synchronize()
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
    ev[0].record()
    for _ in range(reps):
        out = fn()
    ev[1].record()
    torch.cuda.synchronize()
    return ev[0].elapsed_time(ev[1]) / reps, out


# ---- config 4 --------------------------------------------------------------------------------------------------------
N = 1 << 24
pos, w, std = make_lobes(64, seed=3)
V = VmfMixture(pos.to(dev), w.to(dev), std.to(dev))
g = torch.Generator(device=dev).manual_seed(3)
p = torch.rand(N, 3, device=dev, generator=g)
u = torch.rand(N, 2, device=dev, generator=g)
ms_s, (d, pdf) = timed(lambda: V.sample_direction(p, u))
ms_p, pdf2 = timed(lambda: V.pdf_direction(p, d))
o3n, v3n = p.t().contiguous(), d.t().contiguous()
ms_r, rgb = timed(lambda: E.render(o3n, v3n), reps=2)
print("config 4 (1 GPU): vMF sample %.2f ms (%.2f G samples/s), pdf %.2f ms (%.2f G samples/s), NeRF radiance of the "
      "sampled directions %.1f ms (%.1f M rays/s); sample+pdf+radiance %.1f ms for %d samples = %.1f M samples/s"
      % (ms_s, N / ms_s * 1e-6, ms_p, N / ms_p * 1e-6, ms_r, N / ms_r * 1e-3, ms_s + ms_p + ms_r, N,
         N / (ms_s + ms_p + ms_r) * 1e-3))
del p, u, d, pdf, pdf2, o3n, v3n, rgb
# ---- config 5 --------------------------------------------------------------------------------------------------------
W, H, SPP = 2048, 1024, 16
ms_b, img = timed(lambda: bake_envmap(E, W, H, SPP, rows=(0, H)), reps=2)
print("config 5 (1 GPU): %d x %d x %d spp bake = %d rays in %.1f ms = %.1f M rays/s, image %s finite=%s"
      % (W, H, SPP, W * H * SPP, ms_b, W * H * SPP / ms_b * 1e-3, tuple(img.shape), bool(torch.isfinite(img).all())))
