"""CPU: the oracle's restatement of Rotater.apply_frustums (per-sample frustum rotation of turntable captures,
nerfstudio/field_components/rotater.py:125-149) against goldens made by the reference's own classes
(oracle/make_golden_rotater.py) and, where /root/reference exists, against those classes live."""
import pytest
import torch

from helpers import T, load_golden, rel_err


def _cases(G):
    """Three of the golden's eight cases (the GPU tests walk all of them): both bounds, both modes, both rotations."""
    from oracle.rotater_fixture import rotaters
    R = rotaters()
    for kind, training, cam in (("sphere", False, 1), ("sphere", True, 3), ("box", True, 1)):
        yield kind, R[kind], training, cam, "%s::train%d::cam%d::" % (kind, int(training), cam)


def test_oracle_matches_rotater_golden(spec, params_factory):
    from oracle.nerfacto_oracle import OracleEmitter
    G = load_golden("rotater_stress")
    P = params_factory(int(G["param_seed"]), float(G["table_scale"]))
    ids = [int(x) for x in G["rotation_ids"]]
    o, v, do, dv, gout = (T(G[k]) for k in ("o", "v", "do", "dv", "grad_out"))
    jit = [T(G["jitter"][:, i:i + 1]).contiguous() for i in range(3)]
    for kind, rot, training, cam, tag in _cases(G):
        O = OracleEmitter(P, spec, always_use_camera_indices=True, rotation_ids=ids, rotater=rot)
        O.train(training)
        kw = dict(near=float(G["near"]), camera_idx=cam, jitter=jit if training else None)
        rgb, aux = O.render(o, v, return_aux=True, **kw)
        assert torch.equal(aux["sdist2"], T(G[tag + "sdist2"])), tag            # same samples, bit for bit
        assert rel_err(rgb, G[tag + "rgb"]) < 1e-6, tag
        assert rel_err(O.jvp(o, v, do, dv, **kw), G[tag + "jvp"]) < 1e-5, tag
        go, gv, grads, _ = O.vjp(o, v, gout, proposal_requires_grad=True, **kw)
        assert rel_err(go, G[tag + "grad_o"]) < 1e-5 and rel_err(gv, G[tag + "grad_v"]) < 1e-5, tag
        for k, g in grads.items():
            if k.endswith("hash_table"):
                idx, val = T(G[tag + k + "::pick_idx"]), T(G[tag + k + "::pick_val"])
                if len(idx):
                    assert rel_err(g.reshape(-1)[idx], val) < 1e-4, (tag, k)
            else:
                assert rel_err(g, G[tag + k].reshape(g.shape)) < 1e-4, (tag, k)
        if not training:
            pl = O.point_lights(o, v, near=float(G["near"]), camera_idx=cam)
            assert rel_err(pl["rgb"], G[tag + "pl_rgb"]) < 1e-6 and torch.equal(pl["depth"], T(G[tag + "pl_depth"])), tag
            assert rel_err(pl["brightness_grad"], G[tag + "pl_brightness_grad"]) < 1e-5, tag


def test_rotation_changes_what_it_should(spec, params_factory):
    """The rotation moves the forward value; with the identity transform (rotation id 0) it does not; the reverse pass
    sees rotated proposal levels but unrotated final samples (sdf_nerfacto.py:129-131): its ray gradients equal those of
    an unrotated final stage evaluated at the rotated sampler's edges."""
    from oracle.nerfacto_oracle import OracleEmitter
    from oracle.rotater_fixture import rotaters
    from helpers import make_rays
    P = params_factory(0, 1e3)
    rot = rotaters()["sphere"]
    O = OracleEmitter(P, spec, always_use_camera_indices=True, rotater=rot)
    O0 = OracleEmitter(P, spec, always_use_camera_indices=True)
    o, v, g = make_rays(64, seed=4)
    assert rel_err(O.render(o, v, camera_idx=1), O0.render(o, v, camera_idx=1)) > 1e-3
    assert torch.equal(O.render(o, v, camera_idx=0), O0.render(o, v, camera_idx=0))
    gout = torch.randn(64, 3, generator=g)
    _, aux = O.render(o, v, camera_idx=1, return_aux=True)
    O0.edges_override = [None, aux["sdist1"], aux["sdist2"]]
    a = O.vjp(o, v, gout, camera_idx=1, proposal_requires_grad=False)
    b = O0.vjp(o, v, gout, camera_idx=1, proposal_requires_grad=False)
    assert rel_err(a[0], b[0]) < 1e-6 and rel_err(a[1], b[1]) < 1e-6


def test_oracle_rotater_matches_reference_live(spec, params_factory):
    from oracle.ref_import import reference_available
    if not reference_available():
        pytest.skip("reference checkout not present (GPU box)")
    from oracle.nerfacto_oracle import OracleEmitter
    from oracle.ref_harness import ReferenceEmitter, reference_point_lights
    from oracle.rotater_fixture import rotaters
    from helpers import make_rays
    P = params_factory(0, 1e3)
    rot = rotaters()["box"]
    R = ReferenceEmitter(P, spec, always_use_camera_indices=True, rotater=rot)
    O = OracleEmitter(P, spec, always_use_camera_indices=True, rotater=rot)
    o, v, g = make_rays(48, seed=8)
    do, dv = torch.randn(3, 48, generator=g), torch.randn(3, 48, generator=g)
    for cam in (1, 2):
        assert torch.equal(R.render(o, v, camera_idx=cam), O.render(o, v, camera_idx=cam))
        assert rel_err(O.jvp(o, v, do, dv, camera_idx=cam), R.jvp(o, v, do, dv, camera_idx=cam)) < 1e-6
        a, b = reference_point_lights(R, o, v, camera_idx=cam), O.point_lights(o, v, camera_idx=cam)
        assert rel_err(b["brightness_grad"], a["brightness_grad"]) < 1e-5
