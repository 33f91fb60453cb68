"""GPU parity: CUDA emitter forward (through the C ABI) vs the CPU oracle and the reference goldens.
Tolerance: 1e-3 max-norm relative (BASELINE.json north_star, fp32 mode); sample indices compared
exactly with the mismatch count reported (near-ties only)."""
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
    return get


def _check_indices(tag, inds_gpu, inds_ref, cdf_ref, jitter_col=None, tie_tol=2e-6):
    """searchsorted indices must equal the reference's EXCEPT at ties: the deterministic sampler puts
    u_k = (2k+1)/(2 nb) exactly on cdf entries whenever the weights sum to ~1 and leading bins are empty
    (cdf_i = i*0.01/1.96 = i/196 while u_k = (4k+2)/196), so the reference's own index there is decided by
    the last bit of sum(w).  A mismatch is accepted only if it is off by one AND the cdf entry that
    decides it is within tie_tol of u; the count is printed."""
    a = np.asarray(inds_gpu).astype(np.int64)
    b = np.asarray(inds_ref).astype(np.int64)
    cdf = np.asarray(cdf_ref)
    nb = a.shape[1]
    u = np.linspace(0.0, 1.0 - 1.0 / nb, nb, dtype=np.float32)[None, :].repeat(a.shape[0], 0)
    if jitter_col is None:
        u = u + np.float32(1.0 / (2 * nb))
    else:
        u = u + (np.asarray(jitter_col, dtype=np.float32).reshape(-1, 1) / np.float32(nb))
    bad = a != b
    n_bad = int(bad.sum())
    print("%s: %d / %d index mismatches (all must be ties)" % (tag, n_bad, a.size))
    if n_bad:
        assert np.abs(a - b)[bad].max() == 1, tag
        r, k = np.nonzero(bad)
        boundary = cdf[r, np.minimum(a, b)[r, k]]
        assert np.abs(boundary - u[r, k]).max() <= tie_tol, (tag, float(np.abs(boundary - u[r, k]).max()))
    return n_bad


@pytest.mark.parametrize("case", GOLDEN_CASES)
def test_forward_matches_reference_golden(case, spec, emitters):
    G = load_golden(case)
    I = golden_inputs(G)
    E = emitters(int(G["param_seed"]), float(G["table_scale"]))
    E.train(I["training"])
    rgb, aux = E.render(I["o"].cuda(), I["v"].cuda(), near=I["near"], camera_idx=I["camera_idx"],
                        jitter=I["jitter"], return_aux=True)
    torch.cuda.synchronize()
    jit = G["jitter"] if I["training"] else None
    _check_indices(case + " stage1", aux["inds1"].cpu().numpy(), G["inds1"], G["cdf1"], None if jit is None else jit[:, 1])
    _check_indices(case + " stage2", aux["inds2"].cpu().numpy(), G["inds2"], G["cdf2"], None if jit is None else jit[:, 2])
    for k, gk in [("weights0", "weights0"), ("weights1", "weights1"), ("weights2", "weights2"),
                  ("s_edges1", "sdist1"), ("s_edges2", "sdist2"), ("rgb", "rgb"), ("accumulation", "accumulation")]:
        ref = G[gk].reshape(aux[k].shape)
        assert rel_err(aux[k].cpu(), ref) < TOL, (k, rel_err(aux[k].cpu(), ref))


@pytest.mark.parametrize("training", [False, True])
def test_forward_matches_oracle_4096(training, spec, params_factory, emitters):
    """BASELINE config 1 shape: 4096 deterministic rays, 256/96/48, vs the CPU oracle."""
    from oracle.nerfacto_oracle import OracleEmitter
    P = params_factory(0, 1e3)
    O = OracleEmitter(P, spec)
    E = emitters(0, 1e3)
    O.train(training)
    E.train(training)
    N = 4096
    o, v, g = make_rays(N, seed=21)
    jit = [torch.rand(N, 1, generator=g) for _ in range(3)] if training else None
    near = torch.rand(N, generator=g) * 0.2
    rgb_ref, a = O.render(o, v, near=near[:, None], camera_idx=3, jitter=jit, return_aux=True)
    rgb, b = E.render(o.cuda(), v.cuda(), near=near.cuda(), camera_idx=3, jitter=jit, return_aux=True)
    torch.cuda.synchronize()
    _check_indices("4096 stage1", b["inds1"].cpu().numpy(), a["inds1"].numpy(), a["cdf1"].numpy(),
                   None if jit is None else jit[1].numpy())
    _check_indices("4096 stage2", b["inds2"].cpu().numpy(), a["inds2"].numpy(), a["cdf2"].numpy(),
                   None if jit is None else jit[2].numpy())
    for k, rk in [("weights0", "weights0"), ("weights1", "weights1"), ("weights2", "weights2"),
                  ("s_edges1", "sdist1"), ("s_edges2", "sdist2"), ("rgb", "rgb")]:
        assert rel_err(b[k].cpu(), a[rk]) < TOL, (k, rel_err(b[k].cpu(), a[rk]))
    assert rel_err(b["accumulation"].cpu(), a["accumulation"][:, 0]) < TOL


def test_pdf_stage_bit_exact_on_exact_weights(spec, params_factory, emitters):
    """Zero hash tables make every density exp(b1)-independent of position... instead drive the PDF kernel
    with the real pipeline but compare stage-1 indices where the oracle is fed the GPU's own weights:
    given identical fp32 weights the searchsorted indices and the new edges must be BIT-EXACT."""
    from oracle.nerfacto_oracle import pdf_resample, initial_s_edges
    E = emitters(0, 1e3)
    E.train(False)
    N = 2048
    o, v, _ = make_rays(N, seed=33)
    aux = E.sample_levels(o.cuda(), v.cuda())
    torch.cuda.synchronize()
    w0 = aux["weights0"].cpu()
    s0 = initial_s_edges(N, spec.num_proposal_samples[0], None)
    e1, idx1, _ = pdf_resample(w0, s0, spec.num_proposal_samples[1], None)
    assert torch.equal(idx1.to(torch.int32), aux["inds1"].cpu())
    assert torch.equal(e1, aux["s_edges1"].cpu())
    e2, idx2, _ = pdf_resample(aux["weights1"].cpu(), aux["s_edges1"].cpu(), spec.num_nerf_samples, None)
    assert torch.equal(idx2.to(torch.int32), aux["inds2"].cpu())
    assert torch.equal(e2, aux["s_edges2"].cpu())


def test_hash_encode_component(spec, params_factory, emitters):
    from oracle.nerfacto_oracle import hash_encode
    P = params_factory(0, 1e3)
    E = emitters(0, 1e3)
    g = torch.Generator().manual_seed(5)
    x = torch.rand(5000, 3, generator=g)
    x[:10] = 0.0
    x[10:20] = torch.tensor([0.5, 0.25, 0.125])       # exact grid vertices on power-of-two levels
    for which, (name, grid) in enumerate([("field.mlp_base_grid.hash_table", spec.main_grid),
                                          ("proposal_networks.0.encoding.hash_table", spec.prop_grids[0]),
                                          ("proposal_networks.1.encoding.hash_table", spec.prop_grids[1])]):
        ref = hash_encode(x, P[name], grid)
        out = E.hash_encode(x.cuda(), which).cpu()
        assert torch.equal(out, ref), which           # same op order, no contraction: bit-exact


def test_empty_and_ragged_batches(emitters):
    E = emitters(0, 1e3)
    E.train(False)
    z = torch.empty(3, 0, device="cuda")
    assert E.render(z, z).shape == (0, 3)
    for n in (1, 7, 9, 129):
        o, v, _ = make_rays(n, seed=n)
        full = E.render(o.cuda(), v.cuda())
        one = torch.cat([E.render(o[:, i:i + 1].contiguous().cuda(), v[:, i:i + 1].contiguous().cuda()) for i in range(n)])
        assert torch.equal(full, one)                  # per-ray results do not depend on batch shape


def test_errors_are_loud(spec, params_factory):
    from nerf_emitter_b200 import NerfEmitter
    from nerf_emitter_b200.lib import NemitError
    E = NerfEmitter(params_factory(0, 1e3), spec, device=torch.device("cuda:0"))
    E.train(True)
    o, v, _ = make_rays(4)
    with pytest.raises(NemitError):
        E.render(o.cuda(), v.cuda(), camera_idx=99)   # outside the appearance embedding
    with pytest.raises(ValueError):
        E.render(o, v)                                # host tensors are rejected: no CPU path


def test_fp16_tensor_core_mode_psnr(spec, params_factory, emitters):
    """fp16-operand tensor-core MLP (tcgen05, fp32 accumulate) vs the CPU oracle: the stated bound for this mode is
    PSNR >= 70 dB against the oracle radiance (peak = max radiance); measured ~98 dB, i.e. it also meets the
    fp32-mode 1e-3 gate on rgb.  Sampling is the exact path in both modes, so indices are unchanged."""
    import math
    from oracle.nerfacto_oracle import OracleEmitter
    for scale in (1e3, 1.0):
        P = params_factory(0, scale)
        O = OracleEmitter(P, spec)
        E = emitters(0, scale)
        O.train(False)
        E.train(False)
        o, v, _ = make_rays(2048, seed=11)
        ref = O.render(o, v)
        E.set_precision("fp16")
        try:
            out = E.render(o.cuda(), v.cuda()).cpu()
        finally:
            E.set_precision("fp32")
        peak = float(ref.max())
        psnr = 10 * math.log10(peak * peak / max(float(((out - ref) ** 2).mean()), 1e-30))
        print("table_scale %g: fp16-TC PSNR vs oracle %.1f dB, rel err %.2e" % (scale, psnr, rel_err(out, ref)))
        assert psnr >= 70.0 and rel_err(out, ref) < TOL


def test_fast_sampler_mode(spec, params_factory, emitters):
    """NEMIT_SAMPLER_FAST (proposal density fields in packed half precision, the reference's autocast / tcnn behaviour)
    with the fp16 tensor-core field: stated bound PSNR >= 45 dB against the fp32 CPU oracle (peak = max radiance).  The
    final sample edges stay ordered and stay close to the exact sampler's: median |dt| / bin width below 2e-2."""
    import math
    from oracle.nerfacto_oracle import OracleEmitter
    for scale in (1e3, 1.0):
        P = params_factory(0, scale)
        O = OracleEmitter(P, spec)
        E = emitters(0, scale)
        O.train(False)
        E.train(False)
        o, v, _ = make_rays(2048, seed=12)
        ref = O.render(o, v)
        exact = E.sample_levels(o.cuda(), v.cuda(), edges_only=True)["t_edges2"].cpu()
        E.set_precision("fp16")
        E.set_sampler("fast")
        try:
            out = E.render(o.cuda(), v.cuda()).cpu()
            fast = E.sample_levels(o.cuda(), v.cuda(), edges_only=True)["t_edges2"].cpu()
        finally:
            E.set_precision("fp32")
            E.set_sampler("exact")
        peak = float(ref.max())
        psnr = 10 * math.log10(peak * peak / max(float(((out - ref) ** 2).mean()), 1e-30))
        width = (exact[:, 1:] - exact[:, :-1]).clamp_min(1e-12)
        shift = ((fast - exact).abs()[:, :-1] / width)
        print("table_scale %g: fast sampler + fp16-TC PSNR vs oracle %.1f dB, rel err %.2e, median edge shift %.2e bins" % (
            scale, psnr, rel_err(out, ref), float(shift.median())))
        assert torch.isfinite(fast).all() and bool((fast[:, 1:] >= fast[:, :-1]).all())
        assert psnr >= 45.0
        assert float(shift.median()) < 2e-2
