"""GPU: the headless NeRFEmitter plugin mirror (nerf_emitter_b200/mitsuba_emitter.py = DS/emitters/nerf.py:20-82):
world -> local ray transform, the outside_only skip distance, m_scale, detach_op, and gradients back to world-space
rays, against a torch-CPU restatement of the same lines composed with the oracle's op (tolerance 1e-3, north_star).
Mitsuba's BoundingBox3f.ray_intersect is third-party: parity unpinned."""
import math

import pytest
import torch

from helpers import make_rays, rel_err

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
TOL = 1e-3


def _to_world():
    c, s = math.cos(0.7), math.sin(0.7)
    rot = torch.tensor([[c, 0, s, 0], [0, 1, 0, 0], [-s, 0, c, 0], [0, 0, 0, 1.0]], dtype=torch.float64)
    tr = torch.eye(4, dtype=torch.float64)
    tr[:3, 3] = torch.tensor([0.2, -0.1, 0.3])
    sc = torch.diag(torch.tensor([1.5, 1.5, 1.5, 1.0], dtype=torch.float64))
    return tr @ rot @ sc


def _ref_local(o, v, to_world, pad, skip, outside_only):
    """nerf.py:69-79 with mi.BoundingBox3f.ray_intersect restated (slab test)."""
    m = torch.linalg.inv(to_world).float()
    ol = m[:3, :3] @ o + m[:3, 3:4]
    vl = m[:3, :3] @ v
    near = torch.full((o.shape[1],), skip)
    if outside_only:
        lo, hi = -pad, 1.0 + pad
        t1, t2 = (lo - ol) / vl, (hi - ol) / vl
        mint, maxt = torch.minimum(t1, t2).max(0).values, torch.maximum(t1, t2).min(0).values
        near = torch.where(maxt >= mint, maxt, near)
    return ol, vl, near


@pytest.mark.parametrize("outside_only", [False, True])
def test_eval_nerf_matches_restatement(outside_only, spec, params_factory):
    from nerf_emitter_b200 import NerfEmitter
    from nerf_emitter_b200.mitsuba_emitter import NeRFEmitter
    from nerf_emitter_b200.op import make_nerf_emitter_op
    from oracle.nerfacto_oracle import OracleEmitter
    P = params_factory(0, 1.0)
    E = NerfEmitter(P, spec, device=torch.device(DEV)).eval()
    O = OracleEmitter(P, spec)
    N = 300
    o, v, g = make_rays(N, seed=31)
    o = o * 2.0 - 0.3                       # world-space points around the transformed cube
    plug = NeRFEmitter(to_world=_to_world(), scale=2.5, padding_size=0.05, skip_distance=0.01, outside_only=outside_only)
    plug.set_op(make_nerf_emitter_op(E))
    plug.set_camera_idx(1)
    od, vd = o.to(DEV).requires_grad_(True), v.to(DEV).requires_grad_(True)
    out = plug.eval_nerf(od, vd)
    ol, vl, near = _ref_local(o, v, _to_world(), 0.05, 0.01, outside_only)
    k_ol, k_vl, k_near = [t.detach().cpu() for t in plug.to_local(o.to(DEV), v.to(DEV))]
    assert rel_err(k_ol, ol) < 1e-6 and rel_err(k_vl, vl) < 1e-6 and rel_err(k_near, near) < 1e-5
    if outside_only:
        assert int((near != 0.01).sum()) > 20 and int((near == 0.01).sum()) > 20      # both branches are exercised
    ref = O.render(ol, vl, near=near[:, None], camera_idx=1) * 2.5
    assert tuple(out.shape) == (N, 3) and rel_err(out.detach().cpu(), ref) < TOL
    # gradients w.r.t. the WORLD rays: the op's local gradients through the transposed linear map
    gout = torch.randn(N, 3, generator=g)
    E.bind_grads()
    out.backward(gout.to(DEV))
    # (the oracle's op at the kernel's own local rays and sample edges: the strict gate of tests/test_gpu_backward.py)
    aux = E.sample_levels(k_ol.to(DEV), k_vl.to(DEV), near=k_near.to(DEV), camera_idx=1)
    O.edges_override = [None, aux["s_edges1"].cpu(), aux["s_edges2"].cpu()]
    go_l, gv_l, _, _ = O.vjp(k_ol, k_vl, gout * 2.5, near=k_near[:, None], camera_idx=1)
    a = torch.linalg.inv(_to_world()).float()[:3, :3]
    assert rel_err(od.grad.cpu(), a.t() @ go_l) < TOL and rel_err(vd.grad.cpu(), a.t() @ gv_l) < TOL
    assert float(plug.dummy_nerf_parameter.grad.abs().sum()) == 0.0


def test_detach_op_and_errors(spec, params_factory):
    from nerf_emitter_b200 import NerfEmitter
    from nerf_emitter_b200.lib import NemitError
    from nerf_emitter_b200.mitsuba_emitter import NeRFEmitter
    from nerf_emitter_b200.op import make_nerf_emitter_op
    E = NerfEmitter(params_factory(0, 1.0), spec, device=torch.device(DEV)).eval()
    o, v, _ = make_rays(64, seed=2)
    plug = NeRFEmitter(detach_op=True)
    with pytest.raises(NemitError):
        plug.eval_nerf(o.to(DEV), v.to(DEV))                 # no op installed
    plug.set_op(make_nerf_emitter_op(E))
    with pytest.raises(NemitError):
        plug.eval_nerf(o, v)                                 # host tensors: no CPU path
    out = plug.eval(o.to(DEV).requires_grad_(True), -v.to(DEV))
    assert not out.requires_grad                             # nerf.py:64-65
    assert torch.equal(out, E.render(o.to(DEV), v.to(DEV)))  # identity transform, skip 0, scale 1
