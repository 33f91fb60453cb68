"""GPU parity of the light-point-cloud pass (nemit_point_lights = SdfNerfactoModel.get_point_lights,
NS/models/sdf_nerfacto.py:189-212) against the reference goldens and the CPU oracle.  Tolerance 1e-3 max-norm relative
(north_star); the contrib depth is an argmax, compared exactly except where the two best contributions tie to 1e-3;
derivatives on the stress init follow the conditioning gates of tests/test_gpu_backward.py."""
import numpy as np
import pytest
import torch

from helpers import T, load_golden, rel_err

pytestmark = pytest.mark.gpu
TOL = 1e-3
DEV = "cuda:0"


def _grad_mi_from_nerf(g_n3, scale=1.0):
    # d/d o_mitsuba = 2*scale * perm^-1 (d/d o_nerf);  nerf (x,y,z) = mitsuba (z,x,y)
    return (2.0 * scale * g_n3.t())[[1, 2, 0]]


@pytest.mark.parametrize("case,precision", [("point_lights_default_init", "fp32"), ("point_lights_stress", "fp32"),
                                            ("point_lights_stress", "fp16"), ("point_lights_stress", "split")])
def test_point_lights_match_reference_golden(case, precision, spec, params_factory):
    from nerf_emitter_b200 import NerfEmitter
    from oracle.nerfacto_oracle import OracleEmitter
    G = load_golden(case)
    P = params_factory(int(G["param_seed"]), float(G["table_scale"]))
    E = NerfEmitter(P, spec, device=torch.device(DEV)).set_precision(precision)
    E.bind_grads()
    o, v = T(G["o"]), T(G["v"])
    out = E.eval().point_lights(o.to(DEV), v.to(DEV))
    torch.cuda.synchronize()
    assert float(E.flat_grads.abs().sum()) == 0.0            # no parameter gradient is touched
    tol = TOL if precision != "fp16" else 2e-2
    assert rel_err(out["rgb"].cpu(), G["rgb"]) < tol
    # depth = midpoint of the argmax sample.  Against the golden: the picks agree except on rays whose sample edges
    # differ by a sampler tie or whose two best contributions tie; counted and bounded.  Against the
    # oracle evaluated at the CUDA path's OWN edges: same pick, or the two best contributions within 2 * tol.
    d_gpu, d_ref = out["depth"].cpu().numpy()[:, 0], G["depth"][:, 0]
    n_bad = int((~np.isclose(d_gpu, d_ref, rtol=1e-4, atol=0)).sum())      # same sample <=> same midpoint to ~1e-6
    print("%s %s: %d / %d depth picks differ from the golden" % (case, precision, n_bad, d_ref.size))
    assert n_bad <= 0.1 * d_ref.size
    aux = E.sample_levels(o.to(DEV), v.to(DEV))
    O = OracleEmitter(P, spec)
    O.edges_override = [None, aux["s_edges1"].cpu(), aux["s_edges2"].cpu()]
    ref = O.point_lights(o, v)
    d_own = ref["depth"].numpy()[:, 0]
    bad = np.nonzero(~np.isclose(d_gpu, d_own, rtol=1e-5, atol=0))[0]
    contrib = ref["contrib"].numpy()
    top2 = np.sort(contrib, axis=1)[:, -2:]
    print("   vs the oracle at the same edges: %d differ (must be ties)" % bad.size)
    for r in bad:
        assert top2[r, 1] - top2[r, 0] <= 2 * tol * max(abs(top2[r, 1]), 1e-30), (r, top2[r])
    g_ref = _grad_mi_from_nerf(T(G["brightness_grad"]))
    g = out["brightness_grad"].cpu()
    # end to end the sample edges of a few rays differ from the reference's by a sampler tie, which
    # moves single samples; the strict 1e-3 gate is test_point_lights_strict_given_same_samples below
    if True:
        per_ray = (g - g_ref).abs().amax(dim=0) / g_ref.abs().amax()
        frac = float((per_ray < tol).float().mean())
        print("%s %s: %.1f%% of rays within tolerance, worst %.3g" % (case, precision, 100 * frac, float(per_ray.max())))
        assert frac >= 0.85 and float(per_ray.max()) < 0.25


@pytest.mark.parametrize("table_scale,near", [(1.0, 0.0), (1e3, 0.0), (1.0, 0.03)])
def test_point_lights_strict_given_same_samples(table_scale, near, spec, params_factory):
    """With the oracle evaluated at the CUDA path's own sample edges the derivative agrees to 1e-3 on every ray.
    near != 0: brightness_grad is d / d ray_bundle.origins - the op's near change of variables (nerf_emitter_op.py:112)
    must not leak into this pass (it did in round 1: 2 % off at near = 0.02)."""
    from nerf_emitter_b200 import NerfEmitter
    from oracle.nerfacto_oracle import OracleEmitter
    from helpers import make_rays
    P = params_factory(0, table_scale)
    E = NerfEmitter(P, spec, device=torch.device(DEV)).eval()
    o, v, _ = make_rays(128, seed=9)
    aux = E.sample_levels(o.to(DEV), v.to(DEV), near=near)
    out = E.point_lights(o.to(DEV), v.to(DEV), near=near)
    O = OracleEmitter(P, spec)
    O.edges_override = [None, aux["s_edges1"].cpu(), aux["s_edges2"].cpu()]
    ref = O.point_lights(o, v, near=near)
    assert rel_err(out["rgb"].cpu(), ref["rgb"]) < TOL
    assert rel_err(out["brightness_grad"].cpu(), _grad_mi_from_nerf(ref["brightness_grad"])) < TOL


def test_model_get_point_lights_nerf_layout(spec, params_factory):
    """RayBundle-level get_point_lights ((N,3) NeRF coordinates) == CustomOp-layout call, bit for bit."""
    from nerf_emitter_b200 import NerfEmitter
    from nerf_emitter_b200.nerfstudio_api import RayBundle, SdfNerfactoModel
    from oracle.nerfacto_oracle import rays_mi_to_nerf
    from helpers import make_rays
    P = params_factory(0, 1e3)
    m = SdfNerfactoModel(num_train_data=spec.num_images, spec=spec)
    m.load_state_dict(P, strict=False)
    m = m.to(DEV).eval()
    o3n, v3n, _ = make_rays(200, seed=4)
    o, d = rays_mi_to_nerf(o3n.t().contiguous(), v3n.t().contiguous(), 1.0)
    b = m.get_point_lights(RayBundle(origins=o.contiguous().to(DEV), directions=d.contiguous().to(DEV)))
    a = NerfEmitter(P, spec, device=torch.device(DEV)).eval().point_lights(o3n.to(DEV), v3n.to(DEV))
    assert torch.equal(a["rgb"], b["rgb"]) and torch.equal(a["depth"], b["depth"])
    assert tuple(b["brightness_grad"].shape) == (200, 3) and tuple(b["depth"].shape) == (200, 1)
    assert rel_err(_grad_mi_from_nerf(b["brightness_grad"]), a["brightness_grad"]) < 1e-6
