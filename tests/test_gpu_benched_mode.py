"""GPU parity of the BENCHED mode - fp16 tensor-core MLP (NEMIT_PRECISION_FP16_TC) + fast proposal sampler
(NEMIT_SAMPLER_FAST) - in reverse mode, against the reference: the goldens its own classes produced
(tests/golden/*.npz) and the CPU oracle (pinned to those classes), proposal networks included.

What fp16 can and cannot match: the benched mode is the reference's own mixed-precision mode
(torch.autocast / tiny-cuda-nn: fp16 MLPs, fp16 proposal networks).  Against the fp32 CPU reference two things move:
(i) the fast sampler shifts the final sample edges by ~1e-4 of a bin, (ii) fp16 rounding of the hash encoding flips
first-layer ReLU masks of single samples; with the stress init (tables U(-1,1)) a flipped mask changes one ray's
gradient by percents.  So the end-to-end gates are statistical (share of rays within a bound, bounded worst case,
cosine of the table gradient), and the gates that exclude the sampler - the oracle evaluated at the GPU's own
edges - are tighter.  Every bound below is stated next to what was measured on a B200.
"""
import pytest
import torch

from helpers import golden_inputs, load_golden, make_rays, rel_err

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def emitters(spec, params_factory):
    from nerf_emitter_b200 import NerfEmitter
    cache = {}

    def get(seed, scale):
        if (seed, scale) not in cache:
            cache[(seed, scale)] = NerfEmitter(params_factory(seed, scale), spec, device=torch.device("cuda:0"))
        E = cache[(seed, scale)]
        E.set_precision("fp16").set_sampler("fast")
        return E
    yield get
    for E in cache.values():
        E.set_precision("fp32").set_sampler("exact")


def _cos(a, b):
    a, b = a.double().flatten(), b.double().flatten()
    den = float(a.norm() * b.norm())
    return 1.0 if den == 0.0 else float((a * b).sum()) / den


def _ray_stats(go, gv, ro, rv):
    per_ray = torch.maximum((go - ro).abs().max(0).values / ro.abs().max(),
                            (gv - rv).abs().max(0).values / rv.abs().max())
    return per_ray


def _report(tag, per_ray, grads, ref_grads):
    out = {"rays<2e-2": float((per_ray < 2e-2).float().mean()), "rays<1e-3": float((per_ray < 1e-3).float().mean()),
           "worst_ray": float(per_ray.max())}
    for k, g in grads.items():
        r = ref_grads[k]
        if k.endswith("hash_table"):
            out[k] = ("cos", _cos(g, r))
        else:
            out[k] = ("rel", rel_err(g, r))
    print(tag, {k: (v if not isinstance(v, tuple) else "%s %.3e" % v) for k, v in out.items()})
    return out


def _gate(out, *, rays_ok, worst_ray, mlp_tol, table_cos, prop_tol, prg):
    assert out["rays<2e-2"] >= rays_ok, ("share of rays within 2e-2", out["rays<2e-2"])
    assert out["worst_ray"] < worst_ray, ("worst ray", out["worst_ray"])
    for k, v in out.items():
        if not isinstance(v, tuple):
            continue
        kind, val = v
        is_prop = "proposal" in k
        if is_prop and not prg:
            assert (kind == "cos" and val == 1.0) or (kind == "rel" and val == 0.0), (k, "must stay untouched", val)
            continue
        if kind == "cos":
            assert val >= (table_cos if not is_prop else 1.0 - prop_tol), (k, val)
        else:
            assert val <= (prop_tol if is_prop else mlp_tol), (k, val)


# ---- (a) reference goldens (the reference's own classes, 96 rays) -------------------------------------------------------
@pytest.mark.parametrize("case", ["train_stress", "eval_stress", "eval_default_init"])
@pytest.mark.parametrize("prg", [True, False])
def test_benched_mode_backward_vs_reference_golden(case, prg, spec, params_factory, emitters):
    """Measured on B200 (fp16 + fast sampler vs the fp32 reference golden): see the printed line; bounds are ~2x that."""
    G = load_golden(case)
    I = golden_inputs(G)
    E = emitters(int(G["param_seed"]), float(G["table_scale"]))
    E.train(I["training"])
    E.bind_grads()
    go, gv = E.backward(I["o"].cuda(), I["v"].cuda(), I["grad_out"].cuda(), near=I["near"],
                        camera_idx=I["camera_idx"], jitter=I["jitter"], chunk_rays=I["N"], proposal_requires_grad=prg)
    torch.cuda.synchronize()
    pre = "vjp_prop%d::" % int(prg)
    ro, rv = torch.from_numpy(G[pre + "grad_o"]), torch.from_numpy(G[pre + "grad_v"])
    per_ray = _ray_stats(go.cpu(), gv.cpu(), ro, rv)
    stress = float(G["table_scale"]) > 1.0
    frac = float((per_ray < 2e-2).float().mean())
    print(case, prg, "rays within 2e-2: %.1f%%, within 1e-3: %.1f%%, worst %.2e" % (
        100 * frac, 100 * float((per_ray < 1e-3).float().mean()), float(per_ray.max())))
    # measured on B200 (share of rays within 2e-2 / worst ray): train_stress 94.3 % / 5.6e-2, eval_stress 80.2 % / 0.24,
    # eval_default_init 89.6 % / 9.1e-2 (96 rays each; the reference's own derivatives move by up to 6.5 % on ~13 % of the
    # stress rays under 1 ulp of the origins, tests/test_gpu_backward.py::_stability)
    failures = []
    if not (frac >= (0.75 if stress else 0.85) and float(per_ray.max()) < (0.5 if stress else 0.2)):
        failures.append(("rays", frac, float(per_ray.max())))
    # parameter gradients: dense tensors in full, hash tables through the golden's digests (abs-sum, picked rows).
    # measured: MLP tensors <= 3.0e-2 (stress), table abs-sum <= 2.1e-4, picked-row cosine >= 0.9993, proposal tensors <= 2.7e-3
    for name, g in E.grads.items():
        g = g.cpu()
        is_prop = "proposal" in name
        if is_prop and not prg:
            assert float(g.abs().max()) == 0.0, name
            continue
        tol = 6e-2 if not is_prop else 1e-2
        if name.endswith("hash_table"):
            abssum = float(G[pre + name + "::abssum"])
            if abssum == 0.0:
                assert float(g.abs().sum()) == 0.0, name
                continue
            e = abs(float(g.double().abs().sum()) - abssum) / abssum
            idx, val = torch.from_numpy(G[pre + name + "::pick_idx"]), torch.from_numpy(G[pre + name + "::pick_val"])
            c = _cos(g.reshape(-1)[idx], val)
            print("   %-55s abs-sum err %.2e, picked-row cosine %.5f" % (name, e, c))
            if not (e < 2e-3 and c > 0.998):
                failures.append((name, e, c))
        else:
            e = rel_err(g, G[pre + name].reshape(g.shape))
            print("   %-55s %.2e" % (name, e))
            if not e < tol:
                failures.append((name, e, tol))
    assert not failures, failures


# ---- (b) config-1 shape: 4 096 rays against the oracle ---------------------------------------------------------------------
@pytest.mark.parametrize("prg", [True, False])
def test_benched_mode_backward_vs_oracle_4096(prg, spec, params_factory, emitters):
    from oracle.nerfacto_oracle import OracleEmitter
    P = params_factory(0, 1e3)
    O = OracleEmitter(P, spec)
    E = emitters(0, 1e3)
    O.train(True)
    E.train(True)
    N = 4096
    o, v, g = make_rays(N, seed=1)
    gout = torch.randn(N, 3, generator=g)
    ref_go, ref_gv, ref_grads, _ = O.vjp(o, v, gout, proposal_requires_grad=prg)
    E.bind_grads()
    go, gv = E.backward(o.cuda(), v.cuda(), gout.cuda(), chunk_rays=N, proposal_requires_grad=prg)
    torch.cuda.synchronize()
    out = _report("4096 rays, end to end, prg=%s" % prg, _ray_stats(go.cpu(), gv.cpu(), ref_go, ref_gv),
                  {k: t.cpu() for k, t in E.grads.items()}, ref_grads)
    # measured on B200: 93.6 % of rays within 2e-2, worst ray 0.145, MLP tensors <= 1.05e-2, main-table cosine 0.9997,
    # proposal tensors <= 3.1e-3 (fast-sampler edges differ from the oracle's by ~1e-4 of a bin)
    _gate(out, rays_ok=0.90, worst_ray=0.3, mlp_tol=2.5e-2, table_cos=0.9993, prop_tol=1e-2, prg=prg)


# ---- (c) strict: the oracle at the GPU's own sample edges (sampler noise and its ReLU lottery excluded) ------------------
@pytest.mark.parametrize("scale,N", [(1e3, 2048), (1.0, 1001)])
@pytest.mark.parametrize("prg", [True, False])
def test_benched_mode_backward_given_same_samples(scale, N, prg, spec, params_factory, emitters):
    from oracle.nerfacto_oracle import OracleEmitter
    P = params_factory(0, scale)
    O = OracleEmitter(P, spec)
    E = emitters(0, scale)
    O.train(True)
    E.train(True)
    o, v, g = make_rays(N, seed=17)
    jit = [torch.rand(N, 1, generator=g) for _ in range(3)]
    gout = torch.randn(N, 3, generator=g)
    kw = dict(near=0.01, camera_idx=1, jitter=jit)
    aux = E.sample_levels(o.cuda(), v.cuda(), **kw)           # fast sampler: these are the edges the backward will see
    torch.cuda.synchronize()
    O.edges_override = [None, aux["s_edges1"].cpu(), aux["s_edges2"].cpu()]
    ref_go, ref_gv, ref_grads, _ = O.vjp(o, v, gout, proposal_requires_grad=prg, **kw)
    E.bind_grads()
    go, gv = E.backward(o.cuda(), v.cuda(), gout.cuda(), chunk_rays=N, proposal_requires_grad=prg, **kw)
    torch.cuda.synchronize()
    out = _report("same samples, scale %g, prg=%s" % (scale, prg), _ray_stats(go.cpu(), gv.cpu(), ref_go, ref_gv),
                  {k: t.cpu() for k, t in E.grads.items()}, ref_grads)
    if scale > 1.0:
        # measured on B200: 98.4 % of rays within 2e-2 (72.8 % within 1e-3), worst 0.10, MLP tensors <= 6.8e-3, main-table
        # cosine 0.9999, proposal tensors <= 1.5e-4 (k_prop_backward is fp32 at the same edges: only weights2 carries fp16 noise)
        _gate(out, rays_ok=0.97, worst_ray=0.25, mlp_tol=1.5e-2, table_cos=0.9997, prop_tol=2e-3, prg=prg)
    else:
        # default init (tables U(-1e-3, 1e-3)): 99.7 % within 2e-2 (97.9 % within 1e-3), worst 3.8e-2; the MLP tensors sit at
        # fp16 resolution of O(1e-3) encodings: <= 1.9e-2; proposal tensors <= 2.1e-4
        _gate(out, rays_ok=0.99, worst_ray=0.1, mlp_tol=4e-2, table_cos=0.9997, prop_tol=2e-3, prg=prg)


def test_benched_mode_jvp_and_forward_given_same_samples(spec, params_factory, emitters):
    """Forward value and forward-mode tangent of the benched mode at its own samples vs the oracle (PSNR bound of
    DESIGN.md section 5 restated per call: >= 70 dB)."""
    from oracle.nerfacto_oracle import OracleEmitter
    P = params_factory(0, 1e3)
    O = OracleEmitter(P, spec)
    E = emitters(0, 1e3)
    O.train(False)
    E.train(False)
    N = 2048
    o, v, g = make_rays(N, seed=23)
    aux = E.sample_levels(o.cuda(), v.cuda())
    O.edges_override = [None, aux["s_edges1"].cpu(), aux["s_edges2"].cpu()]
    ref = O.render(o, v)
    rgb = E.render(o.cuda(), v.cuda()).cpu()
    mse = float(((rgb - ref) ** 2).mean())
    psnr = 10 * torch.log10(ref.max() ** 2 / torch.tensor(max(mse, 1e-30)))
    print("benched-mode forward at its own samples: PSNR %.1f dB, max rel err %.2e" % (float(psnr), rel_err(rgb, ref)))
    assert float(psnr) >= 70.0
