"""CPU: self-checks of the vMF restatement (parity unpinned by the reference; see oracle/vmf_oracle.py)."""
import math

import torch

from oracle.vmf_oracle import make_lobes, mixture_pdf, sample_direction


def _sphere_grid(n_theta=400, n_phi=800):
    ct = torch.linspace(-1, 1, n_theta + 1)
    ct = (ct[1:] + ct[:-1]) / 2
    ph = torch.linspace(0, 2 * math.pi, n_phi + 1)
    ph = (ph[1:] + ph[:-1]) / 2
    CT, PH = torch.meshgrid(ct, ph, indexing="ij")
    st = torch.sqrt(1 - CT * CT)
    d = torch.stack([st * torch.cos(PH), st * torch.sin(PH), CT], -1).reshape(-1, 3)
    dw = (2.0 / n_theta) * (2 * math.pi / n_phi)
    return d, dw


def test_pdf_integrates_to_one():
    pos, w, std = make_lobes(8, seed=5)
    std = std + 0.3                                   # keep lobes wide enough for the quadrature grid
    d, dw = _sphere_grid()
    p = torch.tensor([[0.4, 0.5, 0.6]]).expand(d.shape[0], 3)
    total = float(mixture_pdf(pos, w, std, p, d).double().sum() * dw)
    assert abs(total - 1.0) < 2e-3


def test_samples_follow_pdf_and_pdf_is_consistent():
    pos, w, std = make_lobes(8, seed=5)
    std = std + 0.3
    g = torch.Generator().manual_seed(9)
    N = 400000
    p = torch.tensor([[0.4, 0.5, 0.6]]).expand(N, 3).contiguous()
    u = torch.rand(N, 2, generator=g)
    d, pdf, idx = sample_direction(pos, w, std, p, u)
    assert torch.allclose(d.norm(dim=-1), torch.ones(N), atol=1e-5)
    assert torch.allclose(pdf, mixture_pdf(pos, w, std, p, d), rtol=1e-5, atol=1e-7)
    # lobe frequencies follow the normalised weights
    freq = torch.bincount(idx, minlength=8).double() / N
    assert torch.allclose(freq, (w / w.sum()).double(), atol=5e-3)
    # histogram over a coarse cos(theta) x phi grid vs integrated pdf
    nb_t, nb_p = 8, 16
    ct = d[:, 2].clamp(-1, 1 - 1e-7)
    ph = torch.atan2(d[:, 1], d[:, 0]) % (2 * math.pi)
    bi = ((ct + 1) / 2 * nb_t).long().clamp(max=nb_t - 1) * nb_p + (ph / (2 * math.pi) * nb_p).long().clamp(max=nb_p - 1)
    hist = torch.bincount(bi, minlength=nb_t * nb_p).double() / N
    gd, dw = _sphere_grid(400, 800)
    gp = mixture_pdf(pos, w, std, p[:1].expand(gd.shape[0], 3), gd).double() * dw
    gbi = ((gd[:, 2] + 1) / 2 * nb_t).long().clamp(max=nb_t - 1) * nb_p + \
        ((torch.atan2(gd[:, 1], gd[:, 0]) % (2 * math.pi)) / (2 * math.pi) * nb_p).long().clamp(max=nb_p - 1)
    expect = torch.zeros(nb_t * nb_p, dtype=torch.float64).index_add_(0, gbi, gp)
    assert float((hist - expect).abs().max()) < 4e-3
