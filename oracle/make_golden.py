"""TEST INFRASTRUCTURE ONLY (oracle). Generates tests/golden/*.npz by running the REFERENCE's own
torch classes (oracle/ref_harness.py) in the build container.  Run from the repo root:

    python -m oracle.make_golden

Weights are not stored (74 MiB): they are regenerated from ``make_params(seed, table_scale)``,
whose draws come from an explicit CPU generator and are identical on every box with this torch.
Table gradients are stored as (sum, abs-sum, values at 4096 fixed touched rows).
"""
from __future__ import annotations

import os
import sys
import warnings

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from nerf_emitter_b200.params import EmitterModelSpec, make_params  # noqa: E402
from oracle.ref_harness import ReferenceEmitter  # noqa: E402

GOLDEN_DIR = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")

CASES = [
    # name, N, seed_params, table_scale, training, near, camera_idx
    ("eval_stress", 192, 0, 1e3, False, 0.0, 0),
    ("train_stress", 192, 0, 1e3, True, 0.1, 1),
    ("eval_default_init", 96, 0, 1.0, False, 0.0, 0),
]


def make_rays(N: int, seed: int):
    """SURVEY.md 8d synthetic rays: origins U([0,1]^3) in Mitsuba space, directions uniform on S^2."""
    g = torch.Generator().manual_seed(seed)
    o = torch.rand(3, N, generator=g)
    v = torch.randn(3, N, generator=g)
    v = v / v.norm(dim=0, keepdim=True)
    return o, v, g


def table_digest(name: str, grad: torch.Tensor, n_pick: int = 4096):
    flat = grad.reshape(-1)
    nz = torch.nonzero(flat).reshape(-1)
    if nz.numel() == 0:
        pick = torch.zeros(0, dtype=torch.int64)
    else:
        g = torch.Generator().manual_seed(7)
        pick = nz[torch.randint(0, nz.numel(), (n_pick,), generator=g)]
    return {
        name + "::sum": np.float64(flat.double().sum().item()),
        name + "::abssum": np.float64(flat.double().abs().sum().item()),
        name + "::nnz": np.int64(nz.numel()),
        name + "::pick_idx": pick.numpy(),
        name + "::pick_val": flat[pick].numpy(),
    }


def main():
    warnings.filterwarnings("ignore")
    os.makedirs(GOLDEN_DIR, exist_ok=True)
    spec = EmitterModelSpec()
    for name, N, pseed, tscale, training, near, cam in CASES:
        P = make_params(spec, seed=pseed, table_scale=tscale)
        R = ReferenceEmitter(P, spec)
        R.train(training)
        o, v, g = make_rays(N, seed=1)
        jit = [torch.rand(N, 1, generator=g) for _ in range(3)]
        do = torch.randn(3, N, generator=g)
        dv = torch.randn(3, N, generator=g)
        gout = torch.randn(N, 3, generator=g)
        jitter = jit if training else None
        rgb, aux = R.render(o, v, near=near, camera_idx=cam, jitter=jitter, return_aux=True)
        tangent = R.jvp(o, v, do, dv, near=near, camera_idx=cam, jitter=jitter)
        out = dict(o=o.numpy(), v=v.numpy(), do=do.numpy(), dv=dv.numpy(), grad_out=gout.numpy(),
                   jitter=torch.cat(jit, dim=1).numpy(), near=np.float32(near), camera_idx=np.int64(cam),
                   training=np.bool_(training), param_seed=np.int64(pseed), table_scale=np.float64(tscale),
                   jvp=tangent.numpy())
        for k, t in aux.items():
            out[k] = t.numpy()
        for prg in (True, False):
            go, gv, grads, _ = R.vjp(o, v, gout, near=near, camera_idx=cam, jitter=jitter,
                                     proposal_requires_grad=prg)
            tag = "vjp_prop%d::" % int(prg)
            out[tag + "grad_o"] = go.numpy()
            out[tag + "grad_v"] = gv.numpy()
            for pn, gr in grads.items():
                if pn.endswith("hash_table"):
                    for kk, vv in table_digest(tag + pn, gr).items():
                        out[kk] = vv
                else:
                    out[tag + pn] = gr.numpy()
        path = os.path.join(GOLDEN_DIR, name + ".npz")
        np.savez_compressed(path, **out)
        print("wrote", path, "%.1f KiB" % (os.path.getsize(path) / 1024))


if __name__ == "__main__":
    main()
