"""TEST INFRASTRUCTURE ONLY (oracle): torch-CPU restatement of the vMF-mixture emitter
(differentiable-sdf-rendering/python/emitters/vMF.py:24-81).

PARITY UNPINNED: the reference composes four Mitsuba 3.4.1 primitives whose C++ source is not in the
reference tree (``mitsuba==3.4.1``, pyproject.toml:60) and ships no test for them; mitsuba/drjit are not
installable here.  The primitives are restated from the published Mitsuba 3 sources:
  * ``warp::square_to_uniform_disk_concentric`` and ``warp::square_to_von_mises_fisher`` (+ ``_pdf``)
    (include/mitsuba/core/warp.h): cos(theta) = 1 + log(s + (1-s) exp(-2k)) / k, s = max(1-r^2, 1e-6);
    pdf = exp(k (z-1)) k / (2 pi (1 - exp(-2k))), 1/(4 pi) at k = 0;
  * ``coordinate_system`` (include/mitsuba/core/vector.h; Duff et al. 2017 branchless ONB);
  * ``DiscreteDistribution::sample_reuse`` (include/mitsuba/core/distr_1d.h): first index with
    cdf[i] >= u * sum; reused sample (u - cdf_norm[i-1]) / pmf_norm[i].
Self-checks in tests/test_vmf_oracle.py: the pdf integrates to 1, samples follow the pdf, and the
pdf returned with a sample equals pdf_direction at that sample.
"""
from __future__ import annotations

import math

import torch


def vmf_pdf(cos_theta, kappa):
    res = torch.exp(kappa * (cos_theta - 1.0)) * (kappa / (2 * math.pi)) / (1.0 - torch.exp(-2.0 * kappa))
    return torch.where(kappa > 0, res, torch.full_like(res, 1 / (4 * math.pi)))


def mixture_pdf(position, weight, std, p, d):
    """vMF.py:64-81. position (K,3), weight (K), std (K); p, d (N,3). Returns (N)."""
    d = d / d.norm(dim=-1, keepdim=True)
    pmf = weight / weight.sum()
    lobe = position[None, :, :] - p[:, None, :]            # (N,K,3)
    ln = lobe.norm(dim=-1)
    mu = lobe / ln[..., None]
    kappa = ln / (std * std)[None, :]
    cos_t = (mu * d[:, None, :]).sum(-1)
    return (vmf_pdf(cos_t, kappa) * pmf[None, :]).sum(-1)


def _disk_concentric(u):
    x = 2 * u[:, 0] - 1
    y = 2 * u[:, 1] - 1
    is_zero = (x == 0) & (y == 0)
    q13 = x.abs() < y.abs()
    r = torch.where(q13, y, x)
    rp = torch.where(q13, x, y)
    phi = 0.25 * math.pi * rp / r
    phi = torch.where(q13, 0.5 * math.pi - phi, phi)
    phi = torch.where(is_zero, torch.zeros_like(phi), phi)
    return r * torch.cos(phi), r * torch.sin(phi)


def _onb(n):
    sign = torch.copysign(torch.ones_like(n[:, 2]), n[:, 2])
    a = -1.0 / (sign + n[:, 2])
    b = n[:, 0] * n[:, 1] * a
    s = torch.stack([1 + sign * n[:, 0] * n[:, 0] * a, sign * b, -sign * n[:, 0]], -1)
    t = torch.stack([b, sign + n[:, 1] * n[:, 1] * a, -n[:, 1]], -1)
    return s, t


def sample_direction(position, weight, std, p, u):
    """vMF.py:24-56 in emitter-local coordinates. Returns d (N,3), pdf (N), lobe index (N)."""
    cdf = torch.cumsum(weight.double(), 0).float()
    total = cdf[-1]
    nz = torch.nonzero(weight > 0).reshape(-1)
    first, last = int(nz[0]), int(nz[-1])
    value = u[:, 0] * total
    idx = torch.searchsorted(cdf[first:last], value.contiguous(), side="left") + first
    pmf = weight / total
    cdf_prev = torch.where(idx > 0, cdf[(idx - 1).clamp(min=0)] / total, torch.zeros_like(value))
    ure = (u[:, 0] - cdf_prev) / pmf[idx]
    lobe = position[idx] - p
    ln = lobe.norm(dim=-1)
    mu = lobe / ln[:, None]
    kappa = ln / (std[idx] ** 2)
    dx, dy = _disk_concentric(torch.stack([ure, u[:, 1]], -1))
    r2 = dx * dx + dy * dy
    sy = torch.clamp(1 - r2, min=1e-6)
    cos_t = 1 + torch.log(sy + (1 - sy) * torch.exp(-2 * kappa)) / kappa
    scale = torch.sqrt(torch.clamp(1 - cos_t * cos_t, min=0) / r2)
    scale = torch.where(r2 > 0, scale, torch.zeros_like(scale))
    local = torch.stack([dx * scale, dy * scale, cos_t], -1)
    s, t = _onb(mu)
    d = s * local[:, 0:1] + t * local[:, 1:2] + mu * local[:, 2:3]
    d = d / d.norm(dim=-1, keepdim=True)
    return d, mixture_pdf(position, weight, std, p, d), idx


def sample_direction_world(position, weight, std, it_p, u, to_world, bsphere_center, bsphere_radius):
    """VonMisesFisherEmitter.sample_direction as the emitter runs it (vMF.py:24-56): world-space it.p, to_world (4,4).
    Returns the DirectionSample fields d, pdf, dist, p."""
    W = torch.as_tensor(to_world, dtype=torch.float32)
    Wi = torch.linalg.inv(W.double()).float()
    p = it_p @ Wi[:3, :3].t() + Wi[:3, 3]                               # transform_affine of a point
    d_local, _, _ = sample_direction(position, weight, std, p, u)
    d = d_local @ W[:3, :3].t()                                        # ... of a vector
    d = d / d.norm(dim=-1, keepdim=True)
    radius = torch.clamp((it_p - torch.as_tensor(bsphere_center, dtype=torch.float32)).norm(dim=-1), min=float(bsphere_radius))
    dist = 2.0 * radius
    return {"d": d, "pdf": pdf_direction_world(position, weight, std, it_p, d, to_world), "dist": dist,
            "p": it_p + d * dist[:, None]}


def pdf_direction_world(position, weight, std, it_p, ds_d, to_world):
    """vMF.py:64-81: d = normalize(W^-1 ds.d), p = W^-1 it.p."""
    Wi = torch.linalg.inv(torch.as_tensor(to_world, dtype=torch.float64)).float()
    d = ds_d @ Wi[:3, :3].t()
    return mixture_pdf(position, weight, std, it_p @ Wi[:3, :3].t() + Wi[:3, 3], d)


def make_lobes(K=64, seed=3):
    """positions U([-1,2]^3), std U(0.05,0.5), weights ~ Dirichlet(1)."""
    g = torch.Generator().manual_seed(seed)
    position = torch.rand(K, 3, generator=g) * 3 - 1
    std = torch.rand(K, generator=g) * 0.45 + 0.05
    e = -torch.log(torch.rand(K, generator=g))
    return position, e / e.sum(), std
