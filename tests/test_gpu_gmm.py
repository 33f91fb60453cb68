"""GPU parity of the light-point-cloud clustering against the float64 numpy restatement
(oracle/gmm_oracle.py; **parity unpinned**: pomegranate is not in the reference tree).  fp32 responsibilities with fp64
sums: 1e-4 relative on the parameters of one EM step, 1e-5 on the weighted log-likelihood."""
import numpy as np
import pytest
import torch

from helpers import rel_err

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _cloud(n=20000, k=8, seed=0):
    g = np.random.default_rng(seed)
    centres = g.uniform(0.1, 0.9, (k, 3))
    stds = g.uniform(0.01, 0.05, k)
    which = g.integers(0, k, n)
    x = centres[which] + g.normal(size=(n, 3)) * stds[which, None]
    w = g.uniform(0.0, 2.0, n) * (g.uniform(size=n) > 0.3)          # 30 % of the points carry zero weight
    return x.astype(np.float32), w.astype(np.float32), centres, stds


def test_compensate_matches_oracle():
    from nerf_emitter_b200.light_clusters import compensate_pc
    from oracle.gmm_oracle import compensate_weights
    g = torch.Generator().manual_seed(1)
    lum = torch.rand(100001, 1, generator=g) ** 3 * 20
    pos = torch.rand(100001, 3, generator=g)
    for thr in (None, 2.5):
        out = compensate_pc(lum.to(DEV), pos.to(DEV), threshold=thr)
        ref = compensate_weights(lum.numpy(), thr)
        assert rel_err(out["weight"].cpu()[:, 0], ref) < 1e-6
        assert int((out["weight"] > 0).sum()) == int((ref > 0).sum())


def test_em_step_matches_oracle():
    from nerf_emitter_b200.light_clusters import em_step
    from oracle.gmm_oracle import em_step as ref_step
    x, w, centres, stds = _cloud()
    K = 16
    g = np.random.default_rng(5)
    mean = x[g.choice(len(x), K, replace=False)].copy()
    var = np.full(K, 0.01, np.float32)
    prior = np.full(K, 1.0 / K, np.float32)
    xs, ws = torch.from_numpy(x).to(DEV), torch.from_numpy(w).to(DEV)
    m, v, p = torch.from_numpy(mean).to(DEV), torch.from_numpy(var).to(DEV), torch.from_numpy(prior).to(DEV)
    last = -np.inf
    for it in range(4):
        m2, v2, p2, logp = em_step(xs, ws, m, v, p)
        rm, rv, rp, rlogp = ref_step(x, w, m.cpu().numpy(), v.cpu().numpy(), p.cpu().numpy())
        assert rel_err(m2.cpu(), rm) < 1e-4 and rel_err(v2.cpu(), rv) < 1e-4 and rel_err(p2.cpu(), rp) < 1e-4, it
        assert abs(float(logp) - rlogp) <= 1e-5 * abs(rlogp), it
        assert float(logp) >= last - 1e-3 * abs(last) if np.isfinite(last) else True      # EM never decreases it
        last = float(logp)
        m, v, p = m2, v2, p2
    assert abs(float(p.sum()) - 1.0) < 1e-5


def test_cluster_light_recovers_planted_lobes():
    from nerf_emitter_b200.light_clusters import cluster_light
    x, w, centres, stds = _cloud(n=40000, k=6, seed=3)
    out = cluster_light(torch.from_numpy(x).to(DEV), torch.from_numpy(w).to(DEV)[:, None], n_cluster=6, n_trials=4, seed=1)
    pos, std, wt = out["position"].cpu().numpy(), out["std"].cpu().numpy()[:, 0], out["weight"].cpu().numpy()
    assert pos.shape == (6, 3) and std.shape == (6,) and abs(wt.sum() - 1) < 1e-5
    for c, s in zip(centres, stds):
        j = np.argmin(((pos - c) ** 2).sum(1))
        assert np.linalg.norm(pos[j] - c) < 3e-3 and abs(std[j] - s) < 0.1 * s, (c, pos[j], s, std[j])
    # the lobe table is what VmfMixture consumes
    from nerf_emitter_b200 import VmfMixture
    V = VmfMixture(out["position"], out["weight"], out["std"][:, 0])
    p = torch.rand(1000, 3, device=DEV)
    d, pdf = V.sample_direction(p, torch.rand(1000, 2, device=DEV))
    assert bool(torch.isfinite(pdf).all()) and float(pdf.min()) > 0


def test_point_lights_to_vmf_lobes_chain(spec, params_factory):
    """rays -> nemit_point_lights -> compensate -> EM -> lobe table -> vMF sampling: the every-10-steps proposal build
    (NS/path_guiding/vmf_guiding.py:21-57) end to end on the device."""
    from helpers import make_rays
    from nerf_emitter_b200 import NerfEmitter, VmfMixture
    from nerf_emitter_b200.light_clusters import build_spherical_gaussian_clusters
    E = NerfEmitter(params_factory(0, 1e3), spec, device=torch.device(DEV)).eval()
    o, v, _ = make_rays(8192, seed=2)
    o = torch.full_like(o, 0.5)                          # probe rays from the scene centre, as the spherical source
    primal, adjoint = build_spherical_gaussian_clusters(E, o.to(DEV), v.to(DEV), n_cluster=16)
    for c in [primal] + adjoint:
        assert tuple(c["position"].shape) == (16, 3) and tuple(c["std"].shape) == (16, 1)
        assert bool(torch.isfinite(c["position"]).all()) and float(c["std"].min()) > 0
        assert abs(float(c["weight"].sum()) - 1.0) < 1e-4
        V = VmfMixture(c["position"], c["weight"], c["std"][:, 0])
        d, pdf = V.sample_direction(torch.rand(512, 3, device=DEV), torch.rand(512, 2, device=DEV))
        assert bool(torch.isfinite(pdf).all())
    assert len(adjoint) == 1


@pytest.mark.gpu
def test_cluster_light_initialisation_edge_cases():
    """the initial means must not depend on torch.multinomial's limits: fewer weighted points than clusters
    (compensate_pc leaves masked points in place with weight 0) and clouds beyond 2^24 points."""
    from nerf_emitter_b200.light_clusters import cluster_light
    g = torch.Generator().manual_seed(0)
    x = torch.rand(5000, 3, generator=g).cuda()
    w = torch.zeros(5000).cuda()
    w[:20] = torch.rand(20, generator=g).cuda() + 0.1            # 20 weighted points, 64 clusters asked for
    out = cluster_light(x, w, n_cluster=64, n_trials=1, max_iter=10)
    assert out["position"].shape == (64, 3) and torch.isfinite(out["position"]).all() and torch.isfinite(out["std"]).all()
    assert abs(float(out["weight"].sum()) - 1.0) < 1e-4
    with pytest.raises(ValueError):
        cluster_light(x, torch.zeros(5000).cuda(), n_cluster=8)
    n = (1 << 24) + 1000
    xb = torch.rand(n, 3, device="cuda")
    wb = torch.rand(n, device="cuda")
    out = cluster_light(xb, wb, n_cluster=16, n_trials=1, max_iter=3)
    assert out["position"].shape == (16, 3) and torch.isfinite(out["position"]).all()
