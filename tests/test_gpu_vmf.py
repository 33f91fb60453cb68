"""GPU parity: vMF-mixture sampling / pdf kernels vs the CPU restatement (oracle/vmf_oracle.py).
Floating point: 1e-3 relative on pdf (the mixture sums 64 exponentials), 1e-4 absolute on directions."""
import pytest
import torch

from helpers import rel_err

from oracle.vmf_oracle import make_lobes, mixture_pdf, sample_direction

pytestmark = pytest.mark.gpu


def test_vmf_sample_and_pdf_match_oracle():
    from nerf_emitter_b200 import VmfMixture
    pos, w, std = make_lobes(64, seed=3)
    g = torch.Generator().manual_seed(4)
    N = 200000
    p = torch.rand(N, 3, generator=g)
    u = torch.rand(N, 2, generator=g)
    u[0] = torch.tensor([0.0, 0.0])
    u[1] = torch.tensor([0.999999, 0.5])
    V = VmfMixture(pos.cuda(), w.cuda(), std.cuda())
    d, pdf = V.sample_direction(p.cuda(), u.cuda())
    d_ref, pdf_ref, _ = sample_direction(pos, w, std, p, u)
    # lobe choice is index work: a different lobe would move d by O(1)
    far = (d.cpu() - d_ref).norm(dim=-1) > 1e-3
    assert int(far.sum()) <= N // 20000, int(far.sum())
    ok = ~far
    assert float((pdf.cpu()[ok] - pdf_ref[ok]).abs().max() / pdf_ref[ok].abs().max()) < 1e-3
    pdf2 = V.pdf_direction(p.cuda(), d)
    assert torch.allclose(pdf2, pdf, rtol=1e-3, atol=1e-6)   # kappa up to ~1e3 amplifies the 1e-7 renormalisation of d
    pdf3 = V.pdf_direction(p.cuda(), (d_ref * 2.5).cuda())       # un-normalised input directions
    assert float((pdf3.cpu() - pdf_ref).abs().max() / pdf_ref.abs().max()) < 1e-3


def test_vmf_ragged_and_single_lobe():
    from nerf_emitter_b200 import VmfMixture
    pos = torch.tensor([[0.5, 0.5, 2.0]])
    V = VmfMixture(pos.cuda(), torch.ones(1).cuda(), torch.tensor([0.3]).cuda())
    for n in (0, 1, 33):
        p = torch.full((n, 3), 0.5).cuda()
        u = torch.rand(n, 2).cuda()
        d, pdf = V.sample_direction(p, u)
        assert d.shape == (n, 3) and pdf.shape == (n,)
        if n:
            ref = mixture_pdf(pos, torch.ones(1), torch.tensor([0.3]), p.cpu(), d.cpu())
            assert torch.allclose(pdf.cpu(), ref, rtol=1e-3)


def test_vmf_world_space_direction_sample():
    """V1 in full (vMF.py:24-56): world_transform in and out, dist = 2 max(r_bsphere, |it.p - c|), ds.p, ds.pdf =
    pdf_direction(it, ds); a rigid + non-uniformly scaled to_world, reference points inside and far outside the sphere."""
    import math
    from nerf_emitter_b200 import VmfMixture
    from oracle.vmf_oracle import pdf_direction_world, sample_direction_world
    pos, w, std = make_lobes(64, seed=3)
    a = 0.6
    W = torch.tensor([[math.cos(a) * 1.3, -math.sin(a), 0.0, 0.2], [math.sin(a) * 1.3, math.cos(a), 0.0, -0.1],
                      [0.0, 0.0, 0.8, 0.05], [0.0, 0.0, 0.0, 1.0]])
    c, r = (0.5, 0.5, 0.5), 0.87
    V = VmfMixture(pos.cuda(), w.cuda(), std.cuda(), world_transform=W, bsphere_center=c, bsphere_radius=r)
    g = torch.Generator().manual_seed(5)
    N = 20000
    it_p = torch.rand(N, 3, generator=g) * 4 - 1.5         # some inside the bounding sphere, most outside
    u = torch.rand(N, 2, generator=g)
    ds = V.sample_direction_world(it_p.cuda(), u.cuda())
    ref = sample_direction_world(pos, w, std, it_p, u, W, c, r)
    cosang = (ds["d"].cpu() * ref["d"]).sum(-1)
    assert float((cosang > 1 - 1e-4).float().mean()) > 0.999           # same direction (fp32 trig of a sharp lobe)
    assert torch.allclose(ds["dist"].cpu(), ref["dist"], rtol=1e-6)
    assert float(ds["dist"].min()) >= 2 * r - 1e-6
    assert torch.allclose(ds["p"].cpu(), it_p + ds["d"].cpu() * ds["dist"].cpu()[:, None], atol=1e-5)
    assert torch.equal(ds["n"], -ds["d"])
    # ds.pdf is pdf_direction at the returned direction (vMF.py:52), on the GPU and in the oracle
    pdf2 = V.pdf_direction_world(it_p.cuda(), ds["d"])
    assert torch.allclose(ds["pdf"], pdf2, rtol=1e-5, atol=1e-7)
    ref_pdf = pdf_direction_world(pos, w, std, it_p, ds["d"].cpu(), W)
    assert rel_err(ds["pdf"].cpu(), ref_pdf) < 1e-3
    # identity frame == the emitter-local entries
    V0 = VmfMixture(pos.cuda(), w.cuda(), std.cuda())
    p01 = torch.rand(N, 3, generator=g)
    d0, pdf0 = V0.sample_direction(p01.cuda(), u.cuda())
    ds0 = V0.sample_direction_world(p01.cuda(), u.cuda())
    assert torch.allclose(d0, ds0["d"], atol=1e-6) and torch.allclose(pdf0, ds0["pdf"], rtol=1e-4, atol=1e-7)   # one more normalize


def test_vmf_guiding_state_dict_round_trip():
    """V3: VonMisesFisherGuiding.state_dict / load_pipeline key names and layouts (vmf_guiding.py:66-98): 'vmf.position'
    is dr.ravel of a Point3f array (xyzxyz...), 'vmf.weight', 'vmf.std', and 'adjoint_emitter_i.*' likewise; load_pipeline
    consumes (deletes) the keys it uses and leaves the rest; the hooks the op calls are no-ops."""
    from nerf_emitter_b200.vmf import N_CLUSTER, VmfGuiding
    assert N_CLUSTER == 64
    G = VmfGuiding("cuda:0")
    lobes = {}
    for i, name in enumerate(["vmf", "adjoint_emitter_0", "adjoint_emitter_1"]):
        pos, w, std = make_lobes(N_CLUSTER, seed=10 + i)
        lobes[name] = (pos, w, std)
        G.load_vmf_params(name, {"position": pos, "weight": w, "std": std})
    sd = G.state_dict(prefix="path_guiding.")
    assert sorted(sd) == sorted("path_guiding.%s.%s" % (n, k) for n in lobes for k in ("position", "weight", "std"))
    for n, (pos, w, std) in lobes.items():
        assert tuple(sd["path_guiding.%s.position" % n].shape) == (3 * N_CLUSTER,)         # ravelled Point3f
        assert torch.equal(sd["path_guiding.%s.position" % n].cpu(), pos.reshape(-1))
        assert torch.equal(sd["path_guiding.%s.weight" % n].cpu(), w) and torch.equal(sd["path_guiding.%s.std" % n].cpu(), std)
    loaded = {k[len("path_guiding."):]: v.cpu().clone() for k, v in sd.items()}
    loaded["field.unrelated"] = torch.zeros(1)
    G2 = VmfGuiding("cuda:0")
    G2.load_pipeline(loaded, step=123)
    assert list(loaded) == ["field.unrelated"]                                             # consumed keys are deleted
    assert sorted(G2.mixtures) == sorted(lobes)
    p = torch.rand(1000, 3).cuda()
    u = torch.rand(1000, 2).cuda()
    for n in lobes:
        a, b = G.mixtures[n].sample_direction(p, u), G2.mixtures[n].sample_direction(p, u)
        assert torch.equal(a[0], b[0]) and torch.equal(a[1], b[1])
    assert G2.primal is G2.mixtures["vmf"]
    assert G2.train_primal(None, None, None, None) is None and G2.train_forward() is None and G2.train_backward() is None
