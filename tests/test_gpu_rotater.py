"""GPU parity of Rotater.apply_frustums in the kernels (nemit_set_rotater): forward, forward mode, reverse mode and the
light-point-cloud pass against goldens made by the reference's own classes (oracle/make_golden_rotater.py; only the 15
lines of apply_frustums are restated there, their module imports mitsuba) and against the CPU oracle.
Tolerance 1e-3 max-norm relative in fp32 mode (north_star); the fp16 mode is held to its PSNR bound."""
import pytest
import torch

from helpers import T, load_golden, make_rays, rel_err

pytestmark = pytest.mark.gpu
TOL = 1e-3
DEV = "cuda:0"


def _emitter(P, spec, rot, ids):
    from nerf_emitter_b200 import NerfEmitter
    E = NerfEmitter(P, spec, device=torch.device(DEV))
    E.set_appearance_mode(always_use_camera_indices=True, rotation_ids=ids)
    if rot["kind"] == "sphere":
        E.set_rotater(rot["transforms"], center=rot["center"], radius=float(rot["radius"]))
    else:
        E.set_rotater(rot["transforms"], aabb=rot["aabb"])
    return E


@pytest.mark.parametrize("precision", ["fp32", "split"])
@pytest.mark.parametrize("kind", ["sphere", "box"])
def test_rotated_frustums_match_reference_golden(kind, precision, spec, params_factory):
    from oracle.nerfacto_oracle import OracleEmitter
    from oracle.rotater_fixture import rotaters
    G = load_golden("rotater_stress")
    P = params_factory(int(G["param_seed"]), float(G["table_scale"]))
    ids = [int(x) for x in G["rotation_ids"]]
    rot = rotaters()[kind]
    E = _emitter(P, spec, rot, ids).set_precision(precision)
    O = OracleEmitter(P, spec, always_use_camera_indices=True, rotation_ids=ids, rotater=rot)
    o, v, do, dv, gout = (T(G[k]) for k in ("o", "v", "do", "dv", "grad_out"))
    jit = [T(G["jitter"][:, i:i + 1]).contiguous() for i in range(3)]
    N = o.shape[1]
    for training in (False, True):
        E.train(training)
        O.train(training)
        for cam in (1, 3):
            tag = "%s::train%d::cam%d::" % (kind, int(training), cam)
            kw = dict(near=float(G["near"]), camera_idx=cam, jitter=jit if training else None)
            rgb, aux = E.render(o.to(DEV), v.to(DEV), return_aux=True, **kw)
            # the rotated proposal levels decide the samples: final edges as the reference's (ties aside, tests/test_gpu_forward.py)
            ds = (aux["s_edges2"].cpu() - T(G[tag + "sdist2"])).abs().max(1).values
            assert float((ds < 1e-5).float().mean()) >= 0.97, (tag, float(ds.max()))
            assert rel_err(rgb.cpu(), G[tag + "rgb"]) < TOL, (tag, rel_err(rgb.cpu(), G[tag + "rgb"]))
            # derivatives at the CUDA path's own sample edges (the strict gate of tests/test_gpu_backward.py)
            O.edges_override = [None, aux["s_edges1"].cpu(), aux["s_edges2"].cpu()]
            # The rotated origin R o + t is rounded differently here (explicit FMAs) and in torch's batched matmul, so the
            # sample positions differ in the last bit even at identical edges and single samples change grid cell / ReLU
            # mask: per-ray gates as for the stress goldens, not the every-ray one.
            t = E.jvp(o.to(DEV), v.to(DEV), do.to(DEV), dv.to(DEV), **kw).cpu()
            rt = O.jvp(o, v, do, dv, **kw)
            pr = (t - rt).abs().max(1).values / rt.abs().max()
            print(tag, "jvp: rays within 1e-3: %.1f%%, worst %.2e" % (100 * float((pr < TOL).float().mean()), float(pr.max())))
            assert float((pr < TOL).float().mean()) >= 0.85 and float(pr.max()) < 0.25, (tag, "jvp")
            ref_go, ref_gv, ref_grads, _ = O.vjp(o, v, gout, proposal_requires_grad=True, **kw)
            E.bind_grads()
            go, gv = E.backward(o.to(DEV), v.to(DEV), gout.to(DEV), chunk_rays=N, proposal_requires_grad=True, **kw)
            torch.cuda.synchronize()
            # the reverse pass evaluates UNROTATED final samples: strict again at identical edges
            assert rel_err(go.cpu(), ref_go) < TOL and rel_err(gv.cpu(), ref_gv) < TOL, (tag, "ray grads")
            for k, g in E.grads.items():
                e = rel_err(g.cpu(), ref_grads[k])
                # the proposal levels ARE rotated: their gradients feel the last-bit position differences
                assert e < (2e-2 if "proposal" in k else TOL), (tag, k, e)
            # ... and end to end against the reference's own numbers: most rays within 1e-3 (1-ulp edge noise moves single
            # samples across cell faces), none far off
            per_ray = torch.maximum((go.cpu() - T(G[tag + "grad_o"])).abs().max(0).values / T(G[tag + "grad_o"]).abs().max(),
                                    (gv.cpu() - T(G[tag + "grad_v"])).abs().max(0).values / T(G[tag + "grad_v"]).abs().max())
            assert float((per_ray < TOL).float().mean()) >= 0.85 and float(per_ray.max()) < 0.25, tag
            O.edges_override = None
            if not training:
                pl = E.point_lights(o.to(DEV), v.to(DEV), near=float(G["near"]), camera_idx=cam)
                assert rel_err(pl["rgb"].cpu(), G[tag + "pl_rgb"]) < TOL, tag
                O.edges_override = [None, aux["s_edges1"].cpu(), aux["s_edges2"].cpu()]
                ref = O.point_lights(o, v, near=float(G["near"]), camera_idx=cam)
                O.edges_override = None
                # brightness_grad comes back in the rays' layout ((3,N) Mitsuba coordinates): d/d origins through R^T
                bg = pl["brightness_grad"].cpu()
                ref_bg = ref["brightness_grad"]            # (N,3) NeRF coordinates -> Mitsuba: x_nerf = (z, x, y)_mi * 2 scale
                ref_mi = torch.stack([ref_bg[:, 1], ref_bg[:, 2], ref_bg[:, 0]], 0) * 2.0
                pb = (bg - ref_mi).abs().max(0).values / ref_mi.abs().max()
                print(tag, "brightness_grad: rays within 1e-3: %.1f%%, worst %.2e" % (100 * float((pb < TOL).float().mean()), float(pb.max())))
                assert float((pb < TOL).float().mean()) >= 0.85 and float(pb.max()) < 0.25, (tag, "brightness_grad")
                same = (pl["depth"].cpu() == ref["depth"]).float().mean()
                assert float(same) >= 0.98, (tag, "contrib depth", float(same))


def test_rotater_fp16_mode_and_identity(spec, params_factory):
    from oracle.nerfacto_oracle import OracleEmitter
    from oracle.rotater_fixture import rotaters
    P = params_factory(0, 1e3)
    rot = rotaters()["sphere"]
    E = _emitter(P, spec, rot, None).eval()
    O = OracleEmitter(P, spec, always_use_camera_indices=True, rotater=rot)
    o, v, _ = make_rays(2048, seed=6)
    ref = O.render(o, v, camera_idx=1)
    E.set_precision("fp16").set_sampler("fast")
    rgb = E.render(o.to(DEV), v.to(DEV), camera_idx=1).cpu()
    psnr = 10 * torch.log10(ref.max() ** 2 / ((rgb - ref) ** 2).mean())
    print("rotated frustums, fp16 tensor-core MLP + fast sampler: PSNR %.1f dB" % float(psnr))
    assert float(psnr) >= 70.0
    E.set_precision("fp32").set_sampler("exact")
    # rotation id 0 is the identity transform: same radiance as without a rotater; removing the rotater restores it too
    a = E.render(o.to(DEV), v.to(DEV), camera_idx=0)
    moved = E.render(o.to(DEV), v.to(DEV), camera_idx=1)
    E.set_rotater(None)
    b = E.render(o.to(DEV), v.to(DEV), camera_idx=0)
    c = E.render(o.to(DEV), v.to(DEV), camera_idx=1)
    assert rel_err(a, b) < 1e-6 and rel_err(moved, c) > 1e-3


def test_rotater_argument_errors(spec, params_factory):
    from nerf_emitter_b200 import NerfEmitter
    from nerf_emitter_b200.lib import NemitError
    from oracle.rotater_fixture import rotaters
    P = params_factory(0, 1e3)
    rot = rotaters()["sphere"]
    E = NerfEmitter(P, spec, device=torch.device(DEV)).eval()
    E.set_rotater(rot["transforms"], center=rot["center"], radius=float(rot["radius"]))
    o, v, _ = make_rays(8)
    with pytest.raises(NemitError):          # 3 transforms, camera 3 has none (no rotation id table: identity mapping)
        E.render(o.to(DEV), v.to(DEV), camera_idx=3)
    with pytest.raises(ValueError):
        E.set_rotater(rot["transforms"])
