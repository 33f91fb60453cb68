"""CPU, build container only: the standalone oracle against the reference's OWN classes run live
(oracle/ref_harness.py imports /root/reference).  Skipped where the reference tree is absent."""
import pytest
import torch

from helpers import make_rays, rel_err
from oracle.ref_import import reference_available

pytestmark = pytest.mark.skipif(not reference_available(), reason="/root/reference not present on this box")


@pytest.mark.parametrize("training", [False, True])
def test_oracle_equals_reference_live(training, spec, params_factory):
    from oracle.nerfacto_oracle import OracleEmitter
    from oracle.ref_harness import ReferenceEmitter
    P = params_factory(0, 1e3)
    R = ReferenceEmitter(P, spec)
    O = OracleEmitter(P, spec)
    R.train(training)
    O.train(training)
    N = 300
    o, v, g = make_rays(N, seed=11)
    jit = [torch.rand(N, 1, generator=g) for _ in range(3)] if training else None
    _, a = R.render(o, v, near=0.2, camera_idx=2, jitter=jit, return_aux=True)
    _, b = O.render(o, v, near=0.2, camera_idx=2, jitter=jit, return_aux=True)
    assert torch.equal(a["inds1"], b["inds1"]) and torch.equal(a["inds2"], b["inds2"])
    for k in ["rgb", "weights0", "weights1", "weights2", "sdist1", "sdist2", "accumulation"]:
        assert torch.equal(a[k], b[k]), k          # same op order -> same bits on the same CPU
    do, dv = torch.randn(3, N, generator=g), torch.randn(3, N, generator=g)
    assert rel_err(O.jvp(o, v, do, dv, near=0.2, camera_idx=2, jitter=jit),
                   R.jvp(o, v, do, dv, near=0.2, camera_idx=2, jitter=jit)) < 1e-6
    gout = torch.randn(N, 3, generator=g)
    go_r, gv_r, gr, _ = R.vjp(o, v, gout, near=0.2, camera_idx=2, jitter=jit)
    go_o, gv_o, gq, _ = O.vjp(o, v, gout, near=0.2, camera_idx=2, jitter=jit)
    assert rel_err(go_o, go_r) < 1e-5 and rel_err(gv_o, gv_r) < 1e-5
    for k in gr:
        assert rel_err(gq[k], gr[k], floor=1e-30) < 1e-4, k


def test_grid_scalings_match_reference(spec):
    from oracle.ref_import import import_reference
    ns = import_reference()
    for g in [spec.main_grid, *spec.prop_grids]:
        enc = ns.encodings.HashEncoding(num_levels=g.num_levels, min_res=g.base_res, max_res=g.max_res,
                                        log2_hashmap_size=4, implementation="torch")
        assert [float(x) for x in enc.scalings] == g.scalings()


def test_point_lights_oracle_equals_reference_live(spec, params_factory):
    """get_point_lights (sdf_nerfacto.py:189-212): rgb, contrib depth and d luminance / d origins."""
    from oracle.nerfacto_oracle import OracleEmitter
    from oracle.ref_harness import ReferenceEmitter, reference_point_lights
    P = params_factory(0, 1e3)
    R = ReferenceEmitter(P, spec)
    O = OracleEmitter(P, spec)
    o, v, _ = make_rays(200, seed=21)
    a = reference_point_lights(R, o, v)
    b = O.point_lights(o, v)
    assert torch.equal(a["rgb"], b["rgb"]) and torch.equal(a["depth"], b["depth"])
    assert rel_err(b["brightness_grad"], a["brightness_grad"]) < 1e-5


@pytest.mark.parametrize("camera_idx", [4, 9])
def test_turntable_appearance_mode_equals_reference_live(camera_idx, spec, params_factory):
    """Model built with a Rotater (nerfacto.py:186-206): always_use_camera_indices, camera -> rotation id mapping
    (rotater.py:94-100, sdf_nerfacto.py:71-72); the embedding row is used in eval mode too."""
    from oracle.nerfacto_oracle import OracleEmitter
    from oracle.ref_harness import ReferenceEmitter
    P = params_factory(0, 1e3)
    ids = [0, 1, 1, 3, 2, 0]
    R = ReferenceEmitter(P, spec, always_use_camera_indices=True, rotation_ids=ids)
    O = OracleEmitter(P, spec, always_use_camera_indices=True, rotation_ids=ids)
    N = 200
    o, v, g = make_rays(N, seed=13)
    a = R.render(o, v, camera_idx=camera_idx)
    b = O.render(o, v, camera_idx=camera_idx)
    assert torch.equal(a, b)
    assert not torch.equal(b, OracleEmitter(P, spec).render(o, v, camera_idx=camera_idx))     # differs from the mean embedding
    gout = torch.randn(N, 3, generator=g)
    go_r, gv_r, gr, _ = R.vjp(o, v, gout, camera_idx=camera_idx, proposal_requires_grad=False)
    go_o, gv_o, gq, _ = O.vjp(o, v, gout, camera_idx=camera_idx, proposal_requires_grad=False)
    assert rel_err(go_o, go_r) < 1e-5 and rel_err(gv_o, gv_r) < 1e-5
    k = "field.embedding_appearance.embedding.weight"
    row = ids[min(camera_idx, len(ids) - 1)]
    assert rel_err(gq[k], gr[k], floor=1e-30) < 1e-4
    assert gq[k][row].abs().max() > 0 and gq[k].abs().sum() == gq[k][row].abs().sum()          # only that row gets a gradient


def test_sampler_update_state_machine_equals_reference_live(spec, params_factory):
    """ray_samplers.py:598,622-623: which sampler calls carry proposal autograd, over a few
    training steps with several calls (chunks) per step - the mirror's bookkeeping against the reference's sampler."""
    from nerf_emitter_b200.nerfstudio_api import ProposalNetworkSampler
    from oracle.ref_harness import ReferenceEmitter
    R = ReferenceEmitter(params_factory(0, 1.0), spec)
    R.train(True)
    R.sampler._step, R.sampler._steps_since_update = 0, 0
    mine = ProposalNetworkSampler(None, update_sched=lambda step: 1)
    o, v, _ = make_rays(4, seed=1)
    rb = R.collider(R.get_ray_bundle(o.t().contiguous(), v.t().contiguous(), 0))
    seen = []
    for step in list(range(0, 3)) + list(range(9, 16)):
        R.sampler.step_cb(step)
        mine.step_cb(step)
        calls = 1 + step % 3
        ref_flags = []
        for _ in range(calls):
            _, wl, _ = R.sampler(rb, density_fns=R._density_fns())
            ref_flags.append(bool(wl[0].requires_grad))
        assert mine._consume_update(calls) == ref_flags, (step, ref_flags)
        assert mine._steps_since_update == R.sampler._steps_since_update
        seen += ref_flags
    assert True in seen and False in seen
