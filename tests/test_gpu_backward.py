"""GPU parity: CUDA forward-mode (nemit_jvp) and reverse-mode (nemit_backward) through the C ABI vs the
reference goldens (made by the reference's own torch classes) and the CPU oracle.
Tolerance: 1e-3 max-norm relative per tensor (BASELINE.json north_star, fp32 mode)."""
import numpy as np
import pytest
import torch

from helpers import GOLDEN_CASES, golden_inputs, load_golden, make_rays, rel_err

pytestmark = pytest.mark.gpu
TOL = 1e-3


# every test of this module runs twice: on the fp32 FMA kernels and on the tensor-core kernels with split bf16 operands
# (NEMIT_PRECISION_SPLIT_TC) - the same 1e-3 gates hold for both
@pytest.fixture(scope="module", params=["fp32", "split"])
def emitters(request, spec, params_factory):
    from nerf_emitter_b200 import NerfEmitter
    cache = {}

    def get(seed, scale):
        if (seed, scale) not in cache:
            cache[(seed, scale)] = NerfEmitter(params_factory(seed, scale), spec, device=torch.device("cuda:0"))
        return cache[(seed, scale)].set_precision(request.param).set_sampler("exact")
    get.mode = request.param
    return get


WELL_CONDITIONED = ["eval_default_init"]          # hash tables U(-1e-3,1e-3): strict gate on every ray
STRESS = ["eval_stress", "train_stress"]          # tables x1e3: see _stability below


def _oracle_for(G, spec, params_factory):
    from oracle.nerfacto_oracle import OracleEmitter
    O = OracleEmitter(params_factory(int(G["param_seed"]), float(G["table_scale"])), spec)
    O.train(bool(G["training"]))
    return O


def _stability(O, I, prg=True):
    """The stress init (tables U(-1,1) at resolutions up to 2047) makes the encoding move by ~2.6e-4 per
    ulp of sample position, so first-layer ReLU signs of individual samples flip under 1-ulp input noise and
    the REFERENCE's own per-ray derivatives jump by several percent on ~7 % of the rays (measured, DESIGN.md
    "Conditioning").  This returns, from the oracle alone, which rays are stable under +-1 ulp of the
    origins and how much the parameter gradients move.  End-to-end gates on the stress goldens therefore
    ask for >= 85 % of the rays within 1e-3 (and none beyond 0.25) and bound parameter gradients by the
    reference's own movement; the STRICT every-ray 1e-3 gate is test_derivatives_strict_given_same_samples."""
    kw = dict(near=I["near"], camera_idx=I["camera_idx"], jitter=I["jitter"])
    base = O.vjp(I["o"], I["v"], I["grad_out"], proposal_requires_grad=prg, **kw)
    jb = O.jvp(I["o"], I["v"], I["do"], I["dv"], **kw)
    stable = torch.ones(I["N"], dtype=torch.bool)
    pmove = {k: 0.0 for k in base[2]}
    for direction in (2.0, -2.0):
        o2 = torch.nextafter(I["o"], torch.full_like(I["o"], direction))
        alt = O.vjp(o2, I["v"], I["grad_out"], proposal_requires_grad=prg, **kw)
        ja = O.jvp(o2, I["v"], I["do"], I["dv"], **kw)
        for a, b in [(alt[0], base[0]), (alt[1], base[1])]:
            stable &= (a - b).abs().max(0).values <= 2.5e-4 * b.abs().max()
        stable &= (ja - jb).abs().max(1).values <= 2.5e-4 * jb.abs().max()
        for k in pmove:
            pmove[k] = max(pmove[k], rel_err(alt[2][k], base[2][k]))
    return stable, pmove


@pytest.mark.parametrize("case", GOLDEN_CASES)
def test_jvp_matches_reference_golden(case, spec, params_factory, emitters):
    G = load_golden(case)
    I = golden_inputs(G)
    E = emitters(int(G["param_seed"]), float(G["table_scale"]))
    E.train(I["training"])
    t = E.jvp(I["o"].cuda(), I["v"].cuda(), I["do"].cuda(), I["dv"].cuda(), near=I["near"],
              camera_idx=I["camera_idx"], jitter=I["jitter"]).cpu()
    ref = torch.from_numpy(G["jvp"])
    if case in WELL_CONDITIONED:
        assert rel_err(t, ref) < TOL, rel_err(t, ref)
        return
    stable, _ = _stability(_oracle_for(G, spec, params_factory), I)
    per_ray = (t - ref).abs().max(1).values / ref.abs().max()
    frac_ok = float((per_ray < TOL).float().mean())
    print(case, "jvp: rays within 1e-3: %.1f%% (reference vs itself under +-1 ulp: %.1f%%), worst %.2e" % (
        100 * frac_ok, 100 * float(stable.float().mean()), float(per_ray.max())))
    assert frac_ok >= 0.85 and float(per_ray.max()) < 0.25


def _check_param_grads(tag, grads, G, prefix, tol_of):
    for name, g in grads.items():
        g = g.cpu()
        tol = tol_of(name)
        if name.endswith("hash_table"):
            flat = g.reshape(-1).double()
            abssum = float(G[prefix + name + "::abssum"])
            s = float(G[prefix + name + "::sum"])
            if abssum == 0.0:
                assert float(flat.abs().sum()) == 0.0, (tag, name)
                continue
            assert abs(float(flat.abs().sum()) - abssum) / abssum < tol, (tag, name, "abssum")
            assert abs(float(flat.sum()) - s) / abssum < tol, (tag, name, "sum")
            idx = torch.from_numpy(G[prefix + name + "::pick_idx"])
            val = torch.from_numpy(G[prefix + name + "::pick_val"])
            e = rel_err(flat[idx].float(), val)
            assert e < tol, (tag, name, "picked rows", e, tol)
        else:
            ref = G[prefix + name]
            e = rel_err(g, ref.reshape(g.shape))
            assert e < tol, (tag, name, e, tol)


@pytest.mark.parametrize("case", GOLDEN_CASES)
@pytest.mark.parametrize("prg", [True, False])
def test_backward_matches_reference_golden(case, prg, spec, params_factory, emitters):
    G = load_golden(case)
    I = golden_inputs(G)
    E = emitters(int(G["param_seed"]), float(G["table_scale"]))
    E.train(I["training"])
    E.bind_grads()
    go, gv = E.backward(I["o"].cuda(), I["v"].cuda(), I["grad_out"].cuda(), near=I["near"],
                        camera_idx=I["camera_idx"], jitter=I["jitter"], chunk_rays=I["N"],
                        proposal_requires_grad=prg)
    torch.cuda.synchronize()
    pre = "vjp_prop%d::" % int(prg)
    ro, rv = torch.from_numpy(G[pre + "grad_o"]), torch.from_numpy(G[pre + "grad_v"])
    if case in WELL_CONDITIONED:
        assert rel_err(go.cpu(), ro) < TOL and rel_err(gv.cpu(), rv) < TOL
        _check_param_grads(case, E.grads, G, pre, lambda name: TOL)
        return
    stable, pmove = _stability(_oracle_for(G, spec, params_factory), I, prg)
    per_ray = torch.maximum((go.cpu() - ro).abs().max(0).values / ro.abs().max(),
                            (gv.cpu() - rv).abs().max(0).values / rv.abs().max())
    frac_ok = float((per_ray < TOL).float().mean())
    print(case, prg, "ray grads within 1e-3: %.1f%% of rays (reference vs itself under +-1 ulp: %.1f%%), worst %.2e" % (
        100 * frac_ok, 100 * float(stable.float().mean()), float(per_ray.max())))
    assert frac_ok >= 0.85 and float(per_ray.max()) < 0.25
    # parameter gradients sum over all rays: gate at the reference's own movement under 1-ulp input noise
    _check_param_grads(case, E.grads, G, pre, lambda name: max(TOL, 4.0 * pmove[name]))


@pytest.mark.parametrize("prg", [True, False])
def test_backward_end_to_end_default_init_4096(prg, spec, params_factory, emitters):
    """Config-1 shape (4 096 deterministic rays), default init (tables U(-1e-3, 1e-3)), fp32 mode, exact sampler: the CUDA
    path samples on its own and is compared with the oracle end to end, so the sampler is exercised on thousands of
    rays and not only on the 96-ray golden.  The derivative of the hash encoding w.r.t. position is piecewise constant
    per grid cell: a sample that sits within an ulp of a cell face of the finest level (resolution 2047: about 1e-3 of
    the samples) takes its d enc / d x from the neighbouring cell when the edges differ in the last bit, and the ray
    gradient then moves by percents - for the reference itself just the same (`_stability`: its own derivatives under
    +-1 ulp of the origins).  Measured on B200: 98.3 - 98.5 % of the rays within 1e-3, worst ray 0.11 (training) / 0.02
    (eval).  Gate: >= 97 % within 1e-3, none beyond 0.25, and the share of deviating rays not above twice the
    reference's own; parameter gradients (sums over all rays) within max(1e-3, 4 x the reference's own movement).
    The every-ray strict gate is test_derivatives_strict_given_same_samples.
    The split-operand tensor-core mode runs the same gates."""
    from oracle.nerfacto_oracle import OracleEmitter
    P = params_factory(0, 1.0)
    O = OracleEmitter(P, spec)
    E = emitters(0, 1.0)
    O.train(True)
    E.train(True)
    N = 4096
    o, v, g = make_rays(N, seed=1)
    gout = torch.randn(N, 3, generator=g)
    do, dv = torch.randn(3, N, generator=g), torch.randn(3, N, generator=g)
    ref_go, ref_gv, ref_grads, _ = O.vjp(o, v, gout, proposal_requires_grad=prg)
    E.bind_grads()
    go, gv = E.backward(o.cuda(), v.cuda(), gout.cuda(), chunk_rays=N, proposal_requires_grad=prg)
    torch.cuda.synchronize()
    per_ray = torch.maximum((go.cpu() - ref_go).abs().max(0).values / ref_go.abs().max(),
                            (gv.cpu() - ref_gv).abs().max(0).values / ref_gv.abs().max())
    I = dict(o=o, v=v, do=do, dv=dv, grad_out=gout, near=0.0, camera_idx=0, jitter=None, N=N)
    stable, pmove = _stability(O, I, prg)
    frac_ok = float((per_ray < TOL).float().mean())
    print("default init, 4096 rays, prg=%s: %.2f%% of rays within 1e-3 (reference vs itself under +-1 ulp: %.2f%%), worst %.2e"
          % (prg, 100 * frac_ok, 100 * float(stable.float().mean()), float(per_ray.max())))
    assert frac_ok >= 0.97 and float(per_ray.max()) < 0.25
    assert 1.0 - frac_ok <= 2.0 * (1.0 - float(stable.float().mean())) + 0.005
    for k, gr in E.grads.items():
        e = rel_err(gr.cpu(), ref_grads[k])
        print("%-60s %.3e (reference's own movement %.3e)" % (k, e, pmove[k]))
        assert e < max(TOL, 4.0 * pmove[k]), (k, e, pmove[k])


def _same_samples(O, E, o, v, **kw):
    """Make the oracle evaluate at the GPU's own sample edges (bit-identical sample positions)."""
    aux = E.sample_levels(o.cuda(), v.cuda(), **kw)
    torch.cuda.synchronize()
    O.edges_override = [None, aux["s_edges1"].cpu(), aux["s_edges2"].cpu()]


@pytest.mark.parametrize("training", [False, True])
def test_derivatives_strict_given_same_samples(training, spec, params_factory, emitters):
    """Strict 1e-3 gate on EVERY ray and every parameter with the stress init: the oracle is evaluated at the
    sample edges the CUDA path produced, which removes the 1-ulp position noise (and with it the ReLU-sign
    lottery) from the comparison; everything downstream of the sampler is compared in full."""
    from oracle.nerfacto_oracle import OracleEmitter
    P = params_factory(0, 1e3)
    O = OracleEmitter(P, spec)
    E = emitters(0, 1e3)
    O.train(training)
    E.train(training)
    N = 384
    o, v, g = make_rays(N, seed=91)
    jit = [torch.rand(N, 1, generator=g) for _ in range(3)] if training else None
    do, dv = torch.randn(3, N, generator=g), torch.randn(3, N, generator=g)
    gout = torch.randn(N, 3, generator=g)
    kw = dict(near=0.02, camera_idx=1, jitter=jit)
    _same_samples(O, E, o, v, **kw)
    ref_t = O.jvp(o, v, do, dv, **kw)
    t = E.jvp(o.cuda(), v.cuda(), do.cuda(), dv.cuda(), **kw).cpu()
    assert rel_err(t, ref_t) < TOL, ("jvp", rel_err(t, ref_t))
    ref_go, ref_gv, ref_grads, _ = O.vjp(o, v, gout, **kw)
    E.bind_grads()
    go, gv = E.backward(o.cuda(), v.cuda(), gout.cuda(), chunk_rays=N, **kw)
    torch.cuda.synchronize()
    assert rel_err(go.cpu(), ref_go) < TOL, ("grad_o", rel_err(go.cpu(), ref_go))
    assert rel_err(gv.cpu(), ref_gv) < TOL, ("grad_v", rel_err(gv.cpu(), ref_gv))
    for k, gr in E.grads.items():
        e = rel_err(gr.cpu(), ref_grads[k])
        print("%-60s %.3e" % (k, e))
        assert e < TOL, (k, e)


def test_backward_matches_oracle_chunked(spec, params_factory, emitters):
    """512 rays in backward chunks of 200 (ragged last chunk): regulariser means are per chunk, parameter
    gradients accumulate over chunks like .backward() does (mitsuba_sdf.py:1180-1190); loss_scale scales
    parameter gradients only."""
    from oracle.nerfacto_oracle import OracleEmitter
    P = params_factory(0, 1e3)
    O = OracleEmitter(P, spec)
    E = emitters(0, 1e3)
    O.train(True)
    E.train(True)
    N, C = 512, 200
    o, v, g = make_rays(N, seed=77)
    jit = [torch.rand(N, 1, generator=g) for _ in range(3)]
    gout = torch.randn(N, 3, generator=g)
    near = 0.05
    ref_go, ref_gv = torch.zeros(3, N), torch.zeros(3, N)
    ref_grads = None
    aux = E.sample_levels(o.cuda(), v.cuda(), near=near, camera_idx=2, jitter=jit)
    s1, s2 = aux["s_edges1"].cpu(), aux["s_edges2"].cpu()
    for a in range(0, N, C):
        b = min(a + C, N)
        O.edges_override = [None, s1[a:b], s2[a:b]]
        go, gv, grads, _ = O.vjp(o[:, a:b].contiguous(), v[:, a:b].contiguous(), gout[a:b], near=near, camera_idx=2,
                                 jitter=[j[a:b] for j in jit])
        ref_go[:, a:b], ref_gv[:, a:b] = go, gv
        ref_grads = grads if ref_grads is None else {k: ref_grads[k] + grads[k] for k in grads}
    E.bind_grads()
    go, gv = E.backward(o.cuda(), v.cuda(), gout.cuda(), near=near, camera_idx=2, jitter=jit, chunk_rays=C,
                        loss_scale=4.0)
    torch.cuda.synchronize()
    assert rel_err(go.cpu(), ref_go) < TOL and rel_err(gv.cpu(), ref_gv) < TOL
    for k, gr in E.grads.items():
        e = rel_err(gr.cpu() / 4.0, ref_grads[k])
        print("%-60s %.3e" % (k, e))
        assert e < TOL, (k, e)
    # accumulation: a second call doubles the parameter gradients
    snap = {k: t.clone() for k, t in E.grads.items()}
    E.backward(o.cuda(), v.cuda(), gout.cuda(), near=near, camera_idx=2, jitter=jit, chunk_rays=C, loss_scale=4.0)
    torch.cuda.synchronize()
    for k in snap:
        assert rel_err(E.grads[k].cpu(), 2 * snap[k].cpu()) < 1e-4, k


def test_jvp_matches_oracle_per_ray_near(spec, params_factory, emitters):
    from oracle.nerfacto_oracle import OracleEmitter
    P = params_factory(0, 1e3)
    O = OracleEmitter(P, spec)
    E = emitters(0, 1e3)
    O.train(False)
    E.train(False)
    N = 777
    o, v, g = make_rays(N, seed=5)
    do, dv = torch.randn(3, N, generator=g), torch.randn(3, N, generator=g)
    near = torch.rand(N, generator=g) * 0.3
    aux = E.sample_levels(o.cuda(), v.cuda(), near=near.cuda())
    O.edges_override = [None, aux["s_edges1"].cpu(), aux["s_edges2"].cpu()]
    ref = O.jvp(o, v, do, dv, near=near[:, None])
    out = E.jvp(o.cuda(), v.cuda(), do.cuda(), dv.cuda(), near=near.cuda())
    torch.cuda.synchronize()
    assert rel_err(out.cpu(), ref) < TOL, rel_err(out.cpu(), ref)


def test_backward_requires_bound_grads(spec, params_factory):
    from nerf_emitter_b200 import NerfEmitter
    import ctypes as C
    from nerf_emitter_b200 import lib as L
    E = NerfEmitter(params_factory(0, 1e3), spec, device=torch.device("cuda:0"))
    o, v, _ = make_rays(4)
    R = E._rays(o.cuda(), v.cuda(), 0.0, 0, None)
    g = torch.zeros(4, 3, device="cuda")
    go = torch.zeros(3, 4, device="cuda")
    ws = E._workspace(4)
    rc = E.lib.nemit_backward(E._h, C.byref(R), g.data_ptr(), 4, 1.0, 1, go.data_ptr(), go.data_ptr(), ws.data_ptr(),
                              ws.numel(), None)
    assert rc == -3 and b"not bound" in E.lib.nemit_last_error()


def _fp16_vs_fp32_backward(E, N, seed=3, prg=False):
    o, v, g = make_rays(N, seed=seed)
    jit = [torch.rand(N, 1, generator=g) for _ in range(3)]
    gout = torch.randn(N, 3, generator=g)
    res = {}
    for prec in ("fp32", "fp16"):
        E.set_precision(prec)
        E.bind_grads()
        go, gv = E.backward(o.cuda(), v.cuda(), gout.cuda(), jitter=jit, chunk_rays=N, proposal_requires_grad=prg)
        torch.cuda.synchronize()
        res[prec] = (go.cpu(), gv.cpu(), {k: t.cpu().clone() for k, t in E.grads.items()})
    E.set_precision("fp32")
    return res["fp32"], res["fp16"]


@pytest.mark.parametrize("scale,N", [(1e3, 1024), (1.0, 1001)])
def test_fp16_mode_backward_tensor_core(scale, N, spec, params_factory, emitters):
    """fp16 mode: forward recompute, composite adjoint (k_field_forward_tc<1>) and the whole MLP backward - dgrad, wgrad
    and bias gradients (k_field_backward_tc) - run on the tensor cores with fp16 operands / fp32 accumulation, as the
    reference's autocast path does.  Gates against the fp32 mode: ray gradients within 2e-2 on >= 90 % of rays (a 1e-4
    change of the recomputed activations moves ReLU masks of single samples), MLP parameter
    gradients within 2e-2 max-norm, the hash-table gradient (16.8 M entries, single-sample entries feel mask flips)
    cosine >= 0.9995 and max-norm within 0.1, identical sparsity pattern.  N = 1001 leaves a ragged last tile."""
    E = emitters(0, scale)
    E.train(True)
    a, b = _fp16_vs_fp32_backward(E, N)
    per_ray = torch.maximum((a[0] - b[0]).abs().max(0).values / a[0].abs().max(),
                            (a[1] - b[1]).abs().max(0).values / a[1].abs().max())
    print("fp16 vs fp32 backward: rays within 2e-2: %.1f%%" % (100 * float((per_ray < 2e-2).float().mean())))
    assert float((per_ray < 2e-2).float().mean()) >= 0.90
    for k in a[2]:
        if "proposal" in k:
            continue
        e = rel_err(b[2][k], a[2][k])
        if k.endswith("hash_table"):
            x, y = a[2][k].double().flatten(), b[2][k].double().flatten()
            cos = float((x * y).sum() / (x.norm() * y.norm()))
            assert cos >= 0.9995 and e < 0.1, (k, cos, e)
            assert int((x != 0).sum()) == int((y != 0).sum())
        else:
            assert e < 2e-2, (k, e)


def test_fp16_mode_backward_tiny_batch_strict(spec, params_factory, emitters):
    """Two rays = one 96-sample tile (the second half of the CTA gets no tile): no mask flip happens to matter here and
    every gradient agrees with the fp32 mode within 2e-3."""
    E = emitters(0, 1.0)
    E.train(True)
    a, b = _fp16_vs_fp32_backward(E, 2)
    assert rel_err(b[0], a[0]) < 2e-3 and rel_err(b[1], a[1]) < 2e-3
    for k in a[2]:
        if "proposal" not in k:
            assert rel_err(b[2][k], a[2][k]) < 2e-3, k


def test_fp16_mode_backward_tiny_and_huge_seeds(spec, params_factory, emitters):
    """The fp16 gradient operands are scaled by a power of two derived from the launch's largest seed, so seeds far
    outside the fp16 range (grad_out x 2^-30 and x 2^+30, loss_scale compensating) still give the fp32 mode's
    gradients."""
    E = emitters(0, 1.0)
    E.train(True)
    N = 256
    o, v, g = make_rays(N, seed=7)
    gout = torch.randn(N, 3, generator=g)
    for mul in (2.0 ** -30, 2.0 ** 30):
        res = {}
        for prec in ("fp32", "fp16"):
            E.set_precision(prec)
            E.bind_grads()
            go, gv = E.backward(o.cuda(), v.cuda(), (gout * mul).cuda(), chunk_rays=N, proposal_requires_grad=False)
            torch.cuda.synchronize()
            res[prec] = (go.cpu(), gv.cpu(), {k: t.cpu().clone() for k, t in E.grads.items()})
        E.set_precision("fp32")
        a, b = res["fp32"], res["fp16"]
        assert torch.isfinite(b[0]).all() and torch.isfinite(b[1]).all()
        for k in ("field.mlp_head.layers.2.weight", "field.mlp_head.layers.1.weight", "field.mlp_base_mlp.layers.0.weight"):
            assert float(a[2][k].abs().max()) > 0
            assert rel_err(b[2][k], a[2][k]) < 3e-2, (mul, k, rel_err(b[2][k], a[2][k]))
        x, y = a[2]["field.mlp_base_grid.hash_table"].double().flatten(), b[2]["field.mlp_base_grid.hash_table"].double().flatten()
        assert float((x * y).sum() / (x.norm() * y.norm())) >= 0.999


def test_split_mode_multi_chunk_matches_fp32_kernels(spec, params_factory):
    """70 000 rays = two launch chunks of the backward (2^16 + 4 464 rays), which the split-operand mode walks in sub-chunks of
    2^15 rays with the sample edges produced once per chunk: every gradient against the fp32 FMA kernels on the same rays
    (default init, where the two modes differ by ~1e-5; measured: MLP tensors <= 1e-5, ray gradients 9e-6)."""
    from nerf_emitter_b200 import NerfEmitter
    E = NerfEmitter(params_factory(0, 1.0), spec, device=torch.device("cuda:0"))
    N = 70000
    o, v, g = make_rays(N, seed=31)
    gout = torch.randn(N, 3, generator=g).cuda()
    o, v = o.cuda(), v.cuda()
    res = {}
    for prec in ("fp32", "split"):
        E.set_precision(prec).set_sampler("exact")
        E.train(True)
        E.bind_grads()
        go, gv = E.backward(o, v, gout, chunk_rays=16384, proposal_requires_grad=True)
        torch.cuda.synchronize()
        res[prec] = (go.clone(), gv.clone(), {k: t.clone() for k, t in E.grads.items()})
    a, b = res["split"], res["fp32"]
    per_ray = torch.maximum((a[0] - b[0]).abs().max(0).values / b[0].abs().max(), (a[1] - b[1]).abs().max(0).values / b[1].abs().max())
    frac = float((per_ray < 1e-4).float().mean())
    print("split vs fp32 kernels, %d rays: %.3f%% of ray gradients within 1e-4, worst %.2e" % (N, 100 * frac, float(per_ray.max())))
    assert frac >= 0.995 and float(per_ray.max()) < 0.25
    for k in a[2]:
        e = rel_err(a[2][k], b[2][k])
        print("%-60s %.3e" % (k, e))
        assert e < (5e-3 if k.endswith("mlp_base_grid.hash_table") else TOL), (k, e)


def test_table_scatter_with_samples_on_grid_planes(spec, params_factory):
    """Regression (round 2): the tensor-core backward sums the table-gradient contributions of consecutive samples that fall into
    the same grid cell before issuing the reductions.  A sample that sits EXACTLY on a grid plane has its ceil corner equal to
    its floor corner, i.e. other table indices than the rest of the cell; as the last lane of such a run it sent the run's
    ceil-corner sums to the floor-corner entries.  This 64-ray batch holds one such sample (level 5): before the fix the table
    gradient of both tensor-core modes was off by 4.4e-2 of its largest entry against the fp32 kernels and the oracle."""
    from nerf_emitter_b200 import NerfEmitter
    E = NerfEmitter(params_factory(0, 1.0), spec, device=torch.device("cuda:0"))
    N = 64
    o, v, g = make_rays(N, seed=1)
    gout = torch.randn(N, 3, generator=g).cuda()
    o, v = o.cuda(), v.cuda()
    res = {}
    for prec in ("fp32", "split", "fp16"):
        E.set_precision(prec).set_sampler("exact")
        E.train(True)
        E.bind_grads()
        E.backward(o, v, gout, chunk_rays=N, proposal_requires_grad=False)
        torch.cuda.synchronize()
        res[prec] = E.grads["field.mlp_base_grid.hash_table"].clone()
    e_split, e_fp16 = rel_err(res["split"], res["fp32"]), rel_err(res["fp16"], res["fp32"])
    print("table gradient vs the fp32 kernels: split %.2e, fp16 %.2e" % (e_split, e_fp16))
    assert e_split < 1e-4 and e_fp16 < 3e-3
    assert int(((res["fp32"] != 0).any(1) != (res["split"] != 0).any(1)).sum()) == 0      # the same entries are touched
