"""TEST INFRASTRUCTURE ONLY (oracle). Golden vectors of the emitter query on a model built with a Rotater (turntable
captures: always_use_camera_indices + camera -> rotation id mapping, nerfstudio/models/nerfacto.py:186-206,
nerfstudio/models/sdf_nerfacto.py:71-72, nerfstudio/field_components/rotater.py:94-100) from the REFERENCE's own
classes (oracle/ref_harness.py), build container only:

    python -m oracle.make_golden_turntable
"""
from __future__ import annotations

import os
import sys
import warnings

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from nerf_emitter_b200.params import EmitterModelSpec, make_params  # noqa: E402
from oracle.make_golden import GOLDEN_DIR, make_rays  # noqa: E402
from oracle.ref_harness import ReferenceEmitter  # noqa: E402

ROTATION_IDS = [0, 1, 1, 3, 2, 0]
APP = "field.embedding_appearance.embedding.weight"


def main():
    warnings.filterwarnings("ignore")
    spec = EmitterModelSpec()
    N = 256
    R = ReferenceEmitter(make_params(spec, seed=0, table_scale=1e3), spec, always_use_camera_indices=True,
                         rotation_ids=ROTATION_IDS)
    o, v, g = make_rays(N, seed=7)
    gout = torch.randn(N, 3, generator=g)
    out = dict(o=o.numpy(), v=v.numpy(), grad_out=gout.numpy(), rotation_ids=np.asarray(ROTATION_IDS, np.int32),
               table_scale=np.float64(1e3), param_seed=np.int64(0))
    for cam in (4, 9):                       # 4 -> rotation 2; 9 -> clamped to the last camera -> rotation 0
        out["rgb_cam%d" % cam] = R.render(o, v, camera_idx=cam).numpy()
        go, gv, grads, _ = R.vjp(o, v, gout, camera_idx=cam, proposal_requires_grad=False)     # eval-mode backward
        out["grad_o_cam%d" % cam] = go.numpy()
        out["grad_v_cam%d" % cam] = gv.numpy()
        out["grad_app_cam%d" % cam] = grads[APP].numpy()
    path = os.path.join(GOLDEN_DIR, "turntable_stress.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, "%.1f KiB" % (os.path.getsize(path) / 1024))


if __name__ == "__main__":
    main()
