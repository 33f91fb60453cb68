"""TEST INFRASTRUCTURE ONLY (oracle). Golden vectors of the light-point-cloud pass (SdfNerfactoModel.get_point_lights,
nerfstudio/models/sdf_nerfacto.py:189-212) from the REFERENCE's own classes (oracle/ref_harness.py), build container only:

    python -m oracle.make_golden_point_lights
"""
from __future__ import annotations

import os
import sys
import warnings

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from nerf_emitter_b200.params import EmitterModelSpec, make_params  # noqa: E402
from oracle.make_golden import GOLDEN_DIR, make_rays  # noqa: E402
from oracle.ref_harness import ReferenceEmitter, reference_point_lights  # noqa: E402


def main():
    warnings.filterwarnings("ignore")
    spec = EmitterModelSpec()
    for name, N, tscale in [("point_lights_stress", 256, 1e3), ("point_lights_default_init", 128, 1.0)]:
        R = ReferenceEmitter(make_params(spec, seed=0, table_scale=tscale), spec)
        o, v, _ = make_rays(N, seed=5)
        out = reference_point_lights(R, o, v)
        path = os.path.join(GOLDEN_DIR, name + ".npz")
        np.savez_compressed(path, o=o.numpy(), v=v.numpy(), table_scale=np.float64(tscale), param_seed=np.int64(0),
                            **{k: t.numpy() for k, t in out.items()})
        print("wrote", path, "%.1f KiB" % (os.path.getsize(path) / 1024))


if __name__ == "__main__":
    main()
