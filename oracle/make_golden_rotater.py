"""TEST INFRASTRUCTURE ONLY (oracle). Golden vectors of the emitter query with a Rotater's per-sample frustum rotation
(nerfstudio/field_components/rotater.py:125-149, hooked into RayBundle.get_ray_samples, nerfstudio/cameras/rays.py:285-286)
from the REFERENCE's own classes (oracle/ref_harness.py; only apply_frustums is restated there, its module imports
mitsuba).  Build container only:

    python -m oracle.make_golden_rotater
"""
from __future__ import annotations

import os
import sys
import warnings

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from nerf_emitter_b200.params import EmitterModelSpec, make_params  # noqa: E402
from oracle.make_golden import GOLDEN_DIR, make_rays, table_digest  # noqa: E402
from oracle.ref_harness import ReferenceEmitter, reference_point_lights  # noqa: E402

from oracle.rotater_fixture import ROTATION_IDS, rotaters  # noqa: E402


def main():
    warnings.filterwarnings("ignore")
    spec = EmitterModelSpec()
    N = 160
    P = make_params(spec, seed=0, table_scale=1e3)
    o, v, g = make_rays(N, seed=11)
    jit = [torch.rand(N, 1, generator=g) for _ in range(3)]
    do, dv = torch.randn(3, N, generator=g), torch.randn(3, N, generator=g)
    gout = torch.randn(N, 3, generator=g)
    out = dict(o=o.numpy(), v=v.numpy(), do=do.numpy(), dv=dv.numpy(), grad_out=gout.numpy(),
               jitter=torch.cat(jit, dim=1).numpy(), rotation_ids=np.asarray(ROTATION_IDS, np.int32),
               table_scale=np.float64(1e3), param_seed=np.int64(0), near=np.float32(0.02))
    for kind, rot in rotaters().items():
        R = ReferenceEmitter(P, spec, always_use_camera_indices=True, rotation_ids=ROTATION_IDS, rotater=rot)
        for training in (False, True):
            R.train(training)
            for cam in (1, 3):                       # camera 1 -> rotation 2, camera 3 -> rotation 1
                kw = dict(near=0.02, camera_idx=cam, jitter=jit if training else None)
                tag = "%s::train%d::cam%d::" % (kind, int(training), cam)
                rgb, aux = R.render(o, v, return_aux=True, **kw)
                out[tag + "rgb"] = rgb.numpy()
                out[tag + "sdist2"] = aux["sdist2"].numpy()
                out[tag + "weights2"] = aux["weights2"].numpy()
                out[tag + "jvp"] = R.jvp(o, v, do, dv, **kw).numpy()
                go, gv, grads, _ = R.vjp(o, v, gout, proposal_requires_grad=True, **kw)
                out[tag + "grad_o"], out[tag + "grad_v"] = go.numpy(), gv.numpy()
                for pn, gr in grads.items():
                    if pn.endswith("hash_table"):
                        out.update(table_digest(tag + pn, gr, n_pick=512))
                    else:
                        out[tag + pn] = gr.numpy()
                if not training:
                    pl = reference_point_lights(R, o, v, near=0.02, camera_idx=cam)
                    for k, t in pl.items():
                        out[tag + "pl_" + k] = t.numpy()
    path = os.path.join(GOLDEN_DIR, "rotater_stress.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, "%.1f KiB" % (os.path.getsize(path) / 1024))


if __name__ == "__main__":
    main()
