"""Light point cloud -> 64 spherical Gaussians -> vMF lobe table: the host-side mirror of
``compensate_pc`` (NS/model_components/output_light_pc.py:176-191), ``cluster_light``
(NS/model_components/gmm_cluster_light.py:41-85) and the cluster dictionary ``load_vmf_params`` consumes
(NS/path_guiding/vmf_guiding.py:52-57: position (K,3), weight (K), std (K)).  The arithmetic is the library's
(``nemit_compensate_weights``, ``nemit_gmm_em_step``); this file draws the initial means and runs the EM loop.
pomegranate is absent and not in the reference tree: the fit is the textbook weighted EM (parity unpinned)."""
from __future__ import annotations

import ctypes as C
from typing import Dict, Optional

import torch

from . import lib as L
from .emitter import _ptr

N_CLUSTER = 64     # NS/path_guiding/vmf_guiding.py:18


def _stream(dev):
    return C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)


def compensate_pc(lum_samples: torch.Tensor, pos_samples: torch.Tensor, dir_samples=None, output_filename=None,
                  threshold: Optional[float] = None) -> Dict[str, torch.Tensor]:
    """output_light_pc.py:176-191.  Returns the same dictionary; masked-out points stay in place with weight 0
    (the reference compacts them away, which changes no sum of the fit)."""
    if not lum_samples.is_cuda:
        raise L.NemitError("compensate_pc: CUDA tensors required; there is no CPU path")
    lum = lum_samples.reshape(-1).float().contiguous()
    w = torch.empty_like(lum)
    scratch = torch.zeros(1, dtype=torch.float64, device=lum.device)
    L.check(L.load().nemit_compensate_weights(_ptr(lum), lum.numel(), float(threshold or 0.0), int(threshold is not None),
                                              _ptr(w), _ptr(scratch), _stream(lum.device)), "nemit_compensate_weights")
    return {"position": pos_samples.reshape(-1, 3), "weight": w.view(-1, 1)}


def em_step(x, w, mean, var, prior, min_var=1e-12, scratch=None):
    """One EM iteration: (mean', var', prior', logp under the input parameters)."""
    k = mean.shape[0]
    dev = x.device
    mean_o, var_o, prior_o = torch.empty_like(mean), torch.empty_like(var), torch.empty_like(prior)
    logp = torch.empty(1, device=dev)
    if scratch is None:
        scratch = torch.empty(6 * k + 2, dtype=torch.float64, device=dev)
    L.check(L.load().nemit_gmm_em_step(_ptr(x), _ptr(w), x.shape[0], k, _ptr(mean), _ptr(var), _ptr(prior), float(min_var),
                                       _ptr(mean_o), _ptr(var_o), _ptr(prior_o), _ptr(logp), _ptr(scratch), _stream(dev)),
            "nemit_gmm_em_step")
    return mean_o, var_o, prior_o, logp


def cluster_light(position: torch.Tensor, weight: torch.Tensor, n_cluster: int = N_CLUSTER, n_trials: int = 3,
                  max_iter: int = 100, tol: float = 1e-4, ignore_weight: bool = False, seed: int = 0,
                  init_means: Optional[torch.Tensor] = None, check_every: int = 5) -> Dict[str, torch.Tensor]:
    """gmm_cluster_light.py:41-85: best of ``n_trials`` weighted fits of ``n_cluster`` isotropic normals;
    returns {'position' (K,3), 'std' (K,1), 'weight' (K)} as device tensors (``equalize`` only re-splits heavy points
    into equal copies, which leaves every weighted EM sum unchanged; the initial means are drawn by weight either way)."""
    if not position.is_cuda:
        raise L.NemitError("cluster_light: CUDA tensors required; there is no CPU path")
    x = position.reshape(-1, 3).float().contiguous()
    w = weight.reshape(-1).float().contiguous()
    if ignore_weight:
        w = (w > 0).float()
    dev = x.device
    gen = torch.Generator(device=dev).manual_seed(seed)
    scratch = torch.empty(6 * n_cluster + 2, dtype=torch.float64, device=dev)
    best = None
    for trial in range(n_trials if init_means is None else 1):
        if init_means is None:
            # initial means: points drawn in proportion to their weight; one shared variance; equal priors.  Drawn by
            # inverting the cumulative weights (torch.multinomial stops at 2^24 categories, and compensate_pc leaves the
            # masked points in place with weight 0); with fewer weighted points than clusters the draw repeats points,
            # jittered like the reference's position_noisy (gmm_cluster_light.py:60), instead of raising
            cdf = torch.cumsum(w.double(), 0)
            n_pos = int((w > 0).sum())
            if n_pos == 0:
                raise ValueError("cluster_light: every point has zero weight")
            if n_pos >= n_cluster:
                # without replacement: stratified draw over the cumulative weight, one point per stratum, deduplicated
                u = (torch.arange(n_cluster, device=dev, dtype=torch.float64)
                     + torch.rand(n_cluster, device=dev, dtype=torch.float64, generator=gen)) / n_cluster * cdf[-1]
                idx = torch.searchsorted(cdf, u, right=True).clamp(max=x.shape[0] - 1)
                idx = torch.unique(idx)
                if idx.numel() < n_cluster:     # heavy points caught several strata: fill up with the next heaviest unused ones
                    rest = torch.argsort(w, descending=True)
                    rest = rest[~torch.isin(rest, idx)][: n_cluster - idx.numel()]
                    idx = torch.cat([idx, rest])
                mean = x[idx].contiguous()
            else:
                pos_idx = torch.nonzero(w > 0).reshape(-1)
                idx = pos_idx[torch.arange(n_cluster, device=dev) % n_pos]
                mean = (x[idx] + 1e-6 * torch.randn(n_cluster, 3, device=dev, generator=gen)).contiguous()
        else:
            mean = init_means.to(dev).float().contiguous()
        spread = float(((x - (x * w[:, None]).sum(0) / w.sum()).square().sum(1) * w).sum() / w.sum() / 3.0)
        var = torch.full((n_cluster,), max(spread / n_cluster ** (2.0 / 3.0), 1e-12), device=dev)
        prior = torch.full((n_cluster,), 1.0 / n_cluster, device=dev)
        last = None
        for it in range(max_iter):
            mean, var, prior, logp = em_step(x, w, mean, var, prior, scratch=scratch)
            if (it + 1) % check_every == 0:                      # one host read every few iterations
                cur = float(logp)
                if last is not None and abs(cur - last) <= tol * abs(last):
                    break
                last = cur
        _, _, _, logp = em_step(x, w, mean, var, prior, scratch=scratch)      # objective of the final parameters
        score = float(logp)
        if best is None or score > best[0]:
            best = (score, mean, var, prior)
    score, mean, var, prior = best
    return {"position": mean, "std": var.sqrt()[:, None], "weight": prior, "logp": score}


def build_spherical_gaussian_clusters(E, o3n: torch.Tensor, v3n: torch.Tensor, camera_idx: int = 0,
                                      adjoint_sampling_strategy: str = "primal_adjoint",
                                      primal_threshold: Optional[float] = None, adjoint_threshold: Optional[float] = None,
                                      n_cluster: int = N_CLUSTER, n_trials: int = 1, seed: int = 0):
    """build_spherical_gaussian_clusters (NS/path_guiding/vmf_guiding.py:21-49) over extract_light_point_cloud
    (NS/model_components/output_light_pc.py:74-173) for a batch of Mitsuba-space probe rays (3,N): the emitter's
    point-lights pass gives rgb, contrib depth and d luminance / d origin per ray; the primal cloud is weighted by the
    luminance, the adjoint cloud by mean |d luminance / d origin| (already in Mitsuba coordinates here, which is what the
    reference obtains through jac_mi2gl, :137-139); both are compensated and clustered in Mitsuba coordinates.
    Returns (primal_cluster, [adjoint_cluster]) with 'position' (K,3), 'std' (K,1), 'weight' (K) on the device."""
    out = E.point_lights(o3n, v3n, camera_idx=camera_idx)
    rgb_w = torch.tensor([0.299, 0.587, 0.114], device=E.device)
    lum = (out["rgb"] * rgb_w).sum(-1, keepdim=True)
    # x_nerf = o_nerf + t d_nerf with o_nerf = (o_mi * 2 - 1) * scale  =>  x_mi = o_mi + t / (2 scale) * v_mi
    pos = (o3n + v3n * (out["depth"].t() / (2.0 * E.scene_scale))).t().contiguous()
    primal = compensate_pc(lum, pos, threshold=primal_threshold)
    primal_cluster = cluster_light(primal["position"], primal["weight"], n_cluster=n_cluster, n_trials=n_trials, seed=seed)
    adjoint_clusters = []
    if adjoint_sampling_strategy in ("adjoint", "primal_adjoint"):
        gnorm = out["brightness_grad"].abs().mean(dim=0)[:, None]            # (3,N) -> (N,1)
        adj = compensate_pc(gnorm, pos, threshold=adjoint_threshold)
        adjoint_clusters.append(cluster_light(adj["position"], adj["weight"], n_cluster=n_cluster, n_trials=n_trials,
                                              seed=seed + 1))
    return primal_cluster, adjoint_clusters
