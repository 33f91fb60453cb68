"""TEST INFRASTRUCTURE ONLY (oracle). numpy restatement of the spherical sensor and the film for the bake path.

  * spherical_rays: SphericalCamera.sample_ray, differentiable-sdf-rendering/python/sensors/spherical_sensor.py:19-44
    (use_sphere_uv = False; to_world = translation, assets/sensors/spherical_sensor.xml:13-15)
  * film: hdrfilm + gaussian rfilter (assets/sensors/spherical_sensor.xml:3-10).  PARITY UNPINNED: the film lives in
    mitsuba==3.4.1 (pyproject.toml:60), which is not under /root/reference; restated from its published algorithm
    (ImageBlock::put with a separable reconstruction filter of radius 4 stddev, value
    max(0, exp(-x^2 / 2 s^2) - exp(-r^2 / 2 s^2)), weights accumulated in the alpha/weight channel and divided out
    by the develop step; no border, samples splat only into pixels inside the film)."""
from __future__ import annotations

import numpy as np


def spherical_rays(width, height, spp, first, n, jitter, origin=(0.5, 0.5, 0.5), near_clip=0.0):
    s = first + np.arange(n, dtype=np.int64)
    pix = s // spp
    px, py = (pix % width).astype(np.float64), (pix // width).astype(np.float64)
    fx, fy = px + jitter[:, 0].astype(np.float64), py + jitter[:, 1].astype(np.float64)
    u, v = fx / width, fy / height
    sp, cp = np.sin(2 * np.pi * u), np.cos(2 * np.pi * u)
    st, ct = np.sin(np.pi * v), np.cos(np.pi * v)
    d = np.stack([sp * st, ct, -cp * st], axis=0)
    o = np.asarray(origin, dtype=np.float64)[:, None] + d * near_clip
    return o, d, np.stack([fx, fy], axis=1)


def film_accumulate(width, height, pos, rgb, filter="gaussian", stddev=0.25):
    acc = np.zeros((height, width, 4), dtype=np.float64)
    if filter == "box":
        px = np.clip(np.floor(pos[:, 0]).astype(int), 0, width - 1)
        py = np.clip(np.floor(pos[:, 1]).astype(int), 0, height - 1)
        np.add.at(acc, (py, px), np.concatenate([rgb, np.ones((len(rgb), 1))], axis=1))
        return acc
    r = 4.0 * stddev
    alpha = -1.0 / (2 * stddev * stddev)
    bias = np.exp(alpha * r * r)

    def f(x):
        return np.where(np.abs(x) < r, np.maximum(0.0, np.exp(alpha * x * x) - bias), 0.0)
    for i in range(len(pos)):
        x, y = float(pos[i, 0]), float(pos[i, 1])
        for py in range(max(int(np.ceil(y - r - 0.5)), 0), min(int(np.floor(y + r - 0.5)), height - 1) + 1):
            wy = f(py + 0.5 - y)
            for px in range(max(int(np.ceil(x - r - 0.5)), 0), min(int(np.floor(x + r - 0.5)), width - 1) + 1):
                w = wy * f(px + 0.5 - x)
                acc[py, px, :3] += w * rgb[i]
                acc[py, px, 3] += w
    return acc


def film_develop(acc):
    out = np.zeros_like(acc)
    m = acc[..., 3] > 0
    out[m, :3] = acc[m, :3] / acc[m, 3:4]
    out[m, 3] = 1.0
    return out
