"""TEST INFRASTRUCTURE ONLY (oracle). The synthetic Rotater used by the rotation goldens and tests: three rotations about
axes through (near) the scene centre and the two kinds of rotation bound (rotater.py:60-66)."""
from __future__ import annotations

import math

import torch

ROTATION_IDS = [0, 2, 1, 1, 2]


def rotation(axis, angle, centre):
    """Rotation about `axis` through `centre` as a 4x4 affine (what Rotater builds from mi.Transform4f, rotater.py:49-58)."""
    a = torch.tensor(axis, dtype=torch.float64)
    a = a / a.norm()
    K = torch.tensor([[0, -a[2], a[1]], [a[2], 0, -a[0]], [-a[1], a[0], 0]], dtype=torch.float64)
    R = torch.eye(3, dtype=torch.float64) + math.sin(angle) * K + (1 - math.cos(angle)) * (K @ K)
    T = torch.eye(4, dtype=torch.float64)
    c = torch.tensor(centre, dtype=torch.float64)
    T[:3, :3] = R
    T[:3, 3] = c - R @ c
    return T.float()


def rotaters(scene_scale=1.0):
    Ts = torch.stack([rotation((0, 1, 0), 0.0, (0, 0, 0)), rotation((0, 1, 0), 0.7, (0.05, -0.02, 0.03)),
                      rotation((1, 0.2, 0), -1.1, (0, 0, 0))])
    return {"sphere": dict(transforms=Ts, kind="sphere", center=torch.zeros(3),
                           radius=torch.tensor(math.sqrt(3) * scene_scale)),          # rotater.py:64-66
            "box": dict(transforms=Ts, kind="box", aabb=torch.tensor([[-1.5, -1.2, -1.4], [1.3, 1.6, 1.5]]))}
