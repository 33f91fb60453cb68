"""GPU parity of the spherical-sensor bake: ray generation and film against the numpy
restatement (oracle/bake_oracle.py; the film is **parity unpinned**, Mitsuba is not in the reference tree), and the
fused bake against its own parts.  fp32 trigonometry / filter weights: tolerance 1e-5 absolute on unit directions,
1e-4 relative on the film."""
import numpy as np
import pytest
import torch

from helpers import rel_err

pytestmark = pytest.mark.gpu
DEV = torch.device("cuda:0")


def test_spherical_rays_match_oracle():
    from nerf_emitter_b200.bake import make_sensor, spherical_rays
    from oracle.bake_oracle import spherical_rays as ref_rays
    W, H, SPP = 64, 32, 4
    s = make_sensor(W, H, SPP, origin=(0.5, 0.4, 0.6), near_clip=0.05)
    first, n = 777, 3000
    jit = torch.rand(n, 2, generator=torch.Generator().manual_seed(3))
    o, v, pos = spherical_rays(s, first, n, DEV, jitter=jit.to(DEV))
    ro, rd, rpos = ref_rays(W, H, SPP, first, n, jit.numpy(), origin=(0.5, 0.4, 0.6), near_clip=0.05)
    assert np.abs(v.cpu().numpy() - rd).max() < 1e-5
    assert np.abs(o.cpu().numpy() - ro).max() < 1e-5
    assert np.abs(pos.cpu().numpy() - rpos).max() < 1e-4
    assert np.abs(np.linalg.norm(v.cpu().numpy(), axis=0) - 1).max() < 1e-5
    # the built-in RNG: positions inside their own pixel, both coordinates roughly uniform
    o2, v2, pos2 = spherical_rays(s, 0, W * H * SPP, DEV)
    p = pos2.cpu().numpy()
    pix = np.arange(W * H * SPP) // SPP
    assert np.array_equal(np.floor(p[:, 0]).astype(int), pix % W) and np.array_equal(np.floor(p[:, 1]).astype(int), pix // W)
    frac = p - np.floor(p)
    assert abs(frac[:, 0].mean() - 0.5) < 0.02 and abs(frac[:, 1].mean() - 0.5) < 0.02
    assert abs(np.corrcoef(frac[:, 0], frac[:, 1])[0, 1]) < 0.05


@pytest.mark.parametrize("filt", ["box", "gaussian"])
def test_film_matches_oracle(filt):
    from nerf_emitter_b200.bake import film_accumulate, film_develop
    from oracle.bake_oracle import film_accumulate as ref_acc, film_develop as ref_dev
    W, H, n = 24, 12, 2500
    g = torch.Generator().manual_seed(4)
    pos = torch.rand(n, 2, generator=g) * torch.tensor([W, H])
    rgb = torch.rand(n, 3, generator=g) * 10
    acc = film_accumulate(torch.zeros(H, W, 4, device=DEV), pos.to(DEV), rgb.to(DEV), filt, 0.25)
    ref = ref_acc(W, H, pos.numpy().astype(np.float64), rgb.numpy().astype(np.float64), filt, 0.25)
    assert rel_err(acc.cpu(), ref) < 1e-4
    img = film_develop(acc)
    assert rel_err(img.cpu(), ref_dev(ref)) < 1e-4
    assert float(img[..., 3].min()) == 1.0


def test_bake_equals_its_parts_and_shards(spec, params_factory):
    """nemit_bake_envmap == raygen -> render -> accumulate -> develop; two row shards sum to the whole film."""
    from nerf_emitter_b200 import NerfEmitter
    from nerf_emitter_b200.bake import bake_envmap, film_accumulate, film_develop, make_sensor, spherical_rays
    E = NerfEmitter(params_factory(0, 1e3), spec, device=DEV).eval()
    W, H, SPP = 64, 32, 2
    whole = bake_envmap(E, W, H, SPP, seed=5, rows=(0, H), develop=False)
    s = make_sensor(W, H, SPP, seed=5)
    o, v, pos = spherical_rays(s, 0, W * H * SPP, DEV)
    rgb = E.render(o, v)
    parts = film_accumulate(torch.zeros(H, W, 4, device=DEV), pos, rgb, "gaussian", 0.25)
    assert rel_err(whole, parts) < 1e-5                      # atomics reorder the sums
    a = bake_envmap(E, W, H, SPP, seed=5, rows=(0, 13), develop=False)
    b = bake_envmap(E, W, H, SPP, seed=5, rows=(13, H), develop=False)
    assert rel_err(a + b, whole) < 1e-5
    img = film_develop(whole)
    assert tuple(img.shape) == (H, W, 4) and bool(torch.isfinite(img).all()) and float(img[..., :3].min()) >= 0.0
    # a pixel's value is a filter-weighted mean of radiance samples around it: inside the range of its neighbourhood
    assert float(img[..., :3].max()) <= float(rgb.max()) * (1 + 1e-5)
