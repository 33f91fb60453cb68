"""GPU parity: vMF-mixture sampling / pdf kernels vs the CPU restatement (oracle/vmf_oracle.py).
Floating point: 1e-3 relative on pdf (the mixture sums 64 exponentials), 1e-4 absolute on directions."""
import pytest
import torch

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
