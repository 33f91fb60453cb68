"""TEST INFRASTRUCTURE ONLY (oracle). numpy (float64) restatement of the light-point-cloud clustering:
  * compensate_pc: nerfstudio/model_components/output_light_pc.py:176-191
  * one EM iteration of the weighted mixture of isotropic normals that cluster_light fits through pomegranate
    (nerfstudio/model_components/gmm_cluster_light.py:41-85).  PARITY UNPINNED: pomegranate (GeneralMixtureModel,
    Normal(covariance_type='sphere')) is a third-party dependency outside /root/reference and absent here; this is the
    textbook EM those classes implement, self-checked by monotone likelihood and planted-cluster recovery."""
from __future__ import annotations

import numpy as np


def compensate_weights(lum, threshold=None):
    lum = np.asarray(lum, dtype=np.float64).reshape(-1)
    c = np.maximum(lum - lum.mean(), 0.0)
    if threshold is not None:
        c = np.where(lum > threshold, c, 0.0)
    return c


def em_step(x, w, mean, var, prior, min_var=1e-12):
    x, w = np.asarray(x, np.float64), np.asarray(w, np.float64)
    mean, var, prior = np.asarray(mean, np.float64), np.asarray(var, np.float64), np.asarray(prior, np.float64)
    d2 = ((x[:, None, :] - mean[None]) ** 2).sum(-1)                                    # (n,k)
    with np.errstate(divide="ignore"):
        logr = np.log(prior)[None] - 1.5 * np.log(2 * np.pi * var)[None] - 0.5 * d2 / var[None]
    m = logr.max(1, keepdims=True)
    s = np.exp(logr - m).sum(1, keepdims=True)
    logp = float((w * (m + np.log(s))[:, 0]).sum())
    r = np.exp(logr - m) / s * w[:, None]
    nk = r.sum(0)
    new_mean, new_var, new_prior = mean.copy(), var.copy(), np.zeros_like(prior)
    ok = nk > 0
    new_mean[ok] = (r.T @ x)[ok] / nk[ok, None]
    d2n = ((x[:, None, :] - new_mean[None]) ** 2).sum(-1)
    new_var[ok] = np.maximum((r * d2n).sum(0)[ok] / nk[ok] / 3.0, min_var)
    new_prior[ok] = nk[ok] / w.sum()
    return new_mean, new_var, new_prior, logp
