"""GPU parity of the appearance modes (nemit_set_appearance_mode): the model built with a Rotater for turntable
captures (always_use_camera_indices + camera -> rotation id mapping, NS/models/nerfacto.py:186-206,
NS/models/sdf_nerfacto.py:71-72, NS/field_components/rotater.py:94-100) against goldens made by the reference's own
classes (oracle/make_golden_turntable.py) and the CPU oracle.  Tolerance 1e-3 max-norm relative (north_star)."""
import pytest
import torch

from helpers import T, load_golden, make_rays, rel_err

pytestmark = pytest.mark.gpu
TOL = 1e-3
DEV = "cuda:0"
APP = "field.embedding_appearance.embedding.weight"


@pytest.mark.parametrize("precision", ["fp32", "fp16", "split"])
def test_turntable_forward_and_embedding_gradient_match_reference_golden(precision, spec, params_factory):
    from nerf_emitter_b200 import NerfEmitter
    from oracle.nerfacto_oracle import OracleEmitter
    G = load_golden("turntable_stress")
    P = params_factory(int(G["param_seed"]), float(G["table_scale"]))
    ids = [int(x) for x in G["rotation_ids"]]
    E = NerfEmitter(P, spec, device=torch.device(DEV)).set_precision(precision)
    E.set_appearance_mode(always_use_camera_indices=True, rotation_ids=ids).eval()
    O = OracleEmitter(P, spec, always_use_camera_indices=True, rotation_ids=ids)
    o, v, gout = T(G["o"]), T(G["v"]), T(G["grad_out"])
    tol = TOL if precision != "fp16" else 2e-2      # split = fp32 arithmetic on the tensor cores: the fp32 gates
    for cam in (4, 9):
        rgb = E.render(o.to(DEV), v.to(DEV), camera_idx=cam).cpu()
        assert rel_err(rgb, G["rgb_cam%d" % cam]) < tol, (cam, rel_err(rgb, G["rgb_cam%d" % cam]))
        # reverse mode in eval mode at the CUDA path's own sample edges (the strict gate, tests/test_gpu_backward.py)
        aux = E.sample_levels(o.to(DEV), v.to(DEV), camera_idx=cam)
        O.edges_override = [None, aux["s_edges1"].cpu(), aux["s_edges2"].cpu()]
        ref_go, ref_gv, ref_grads, _ = O.vjp(o, v, gout, camera_idx=cam, proposal_requires_grad=False)
        E.bind_grads()
        go, gv = E.backward(o.to(DEV), v.to(DEV), gout.to(DEV), camera_idx=cam, chunk_rays=o.shape[1],
                            proposal_requires_grad=False)
        torch.cuda.synchronize()
        if precision != "fp16":
            assert rel_err(go.cpu(), ref_go) < tol and rel_err(gv.cpu(), ref_gv) < tol
        else:       # fp16 mode: ReLU masks of single samples move (tests/test_gpu_backward.py)
            per_ray = torch.maximum((go.cpu() - ref_go).abs().max(0).values / ref_go.abs().max(),
                                    (gv.cpu() - ref_gv).abs().max(0).values / ref_gv.abs().max())
            assert float((per_ray < tol).float().mean()) >= 0.90
        g_app = E.grads[APP].cpu()
        row = ids[min(cam, len(ids) - 1)]
        assert rel_err(g_app, ref_grads[APP]) < tol, rel_err(g_app, ref_grads[APP])
        assert float(g_app.abs().sum()) == float(g_app[row].abs().sum()) > 0          # only the rotation's row
        # and against the reference's own gradient of that row (sampler ties move single samples: 5 % gate)
        assert rel_err(g_app, G["grad_app_cam%d" % cam]) < 5e-2, rel_err(g_app, G["grad_app_cam%d" % cam])


def test_appearance_modes_and_errors(spec, params_factory):
    from nerf_emitter_b200 import NerfEmitter
    from nerf_emitter_b200.lib import NemitError
    from oracle.nerfacto_oracle import OracleEmitter
    P = params_factory(0, 1e3)
    E = NerfEmitter(P, spec, device=torch.device(DEV)).eval()
    o, v, _ = make_rays(300, seed=3)
    mean = E.render(o.to(DEV), v.to(DEV), camera_idx=1).cpu()
    assert rel_err(mean, OracleEmitter(P, spec).render(o, v, camera_idx=1)) < TOL
    # eval + always_use_camera_indices without a rotation table: the camera's own row
    E.set_appearance_mode(always_use_camera_indices=True)
    row = E.render(o.to(DEV), v.to(DEV), camera_idx=1).cpu()
    assert rel_err(row, OracleEmitter(P, spec, always_use_camera_indices=True).render(o, v, camera_idx=1)) < TOL
    assert not torch.equal(row, mean)
    with pytest.raises(NemitError):
        E.render(o.to(DEV), v.to(DEV), camera_idx=spec.num_images)          # no such row
    # use_average_appearance_embedding=False: zeros in eval (nerfacto_field.py:281-285)
    E.set_appearance_mode(always_use_camera_indices=False, use_average_appearance_embedding=False)
    zero = E.render(o.to(DEV), v.to(DEV), camera_idx=1).cpu()
    Pz = {k: (torch.zeros_like(t) if k.endswith("embedding_appearance.embedding.weight") else t) for k, t in P.items()}
    assert rel_err(zero, OracleEmitter(Pz, spec).render(o, v, camera_idx=1)) < TOL
    with pytest.raises(NemitError):
        E.set_appearance_mode(always_use_camera_indices=True, rotation_ids=[0, spec.num_images])
    # back to the default
    E.set_appearance_mode()
    assert torch.equal(E.render(o.to(DEV), v.to(DEV), camera_idx=1).cpu(), mean)
