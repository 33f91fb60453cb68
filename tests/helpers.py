"""Shared helpers for the parity tests."""
import os

import numpy as np
import torch

GOLDEN_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
GOLDEN_CASES = ["eval_stress", "train_stress", "eval_default_init"]


def load_golden(name):
    z = np.load(os.path.join(GOLDEN_DIR, name + ".npz"))
    return {k: z[k] for k in z.files}


def T(a):
    return torch.from_numpy(np.asarray(a))


def golden_inputs(G):
    N = G["o"].shape[1]
    training = bool(G["training"])
    jit = [T(G["jitter"][:, i:i + 1]).contiguous() for i in range(3)] if training else None
    return dict(o=T(G["o"]), v=T(G["v"]), do=T(G["do"]), dv=T(G["dv"]), grad_out=T(G["grad_out"]),
                near=float(G["near"]), camera_idx=int(G["camera_idx"]), training=training, jitter=jit, N=N)


def rel_err(a, b, floor=None):
    """max-norm relative error per tensor: max|a-b| / max(max|b|, floor)."""
    a = torch.as_tensor(a).double()
    b = torch.as_tensor(b).double()
    if b.numel() == 0:
        return 0.0
    den = float(b.abs().max())
    if floor is not None:
        den = max(den, floor)
    if den == 0.0:
        return float((a - b).abs().max())
    return float((a - b).abs().max()) / den


def make_rays(N, seed=1):
    g = torch.Generator().manual_seed(seed)
    o = torch.rand(3, N, generator=g)
    v = torch.randn(3, N, generator=g)
    v = v / v.norm(dim=0, keepdim=True)
    return o, v, g
