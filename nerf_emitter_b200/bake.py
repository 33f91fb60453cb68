"""Headless spherical-sensor bake of the emitter radiance into a device image (BASELINE
config 5): the host-side mirror of ``SphericalCamera`` + hdrfilm (DS/sensors/spherical_sensor.py:19-44,
assets/sensors/spherical_sensor.xml).  Ray generation, the NeRF forward, the reconstruction-filter splat and the
develop step are library kernels (``nemit_spherical_rays`` / ``nemit_forward`` / ``nemit_film_accumulate`` /
``nemit_film_develop``); ranks own film rows and sum their accumulation buffers (one all-reduce of H*W*4 floats)."""
from __future__ import annotations

import ctypes as C
from typing import Optional, Tuple

import torch

from . import lib as L
from .emitter import NerfEmitter, _ptr

FILTERS = {"box": 0, "gaussian": 1}


def make_sensor(width: int, height: int, spp: int, origin=(0.5, 0.5, 0.5), near_clip: float = 0.0, seed: int = 0):
    s = L.NemitSphericalSensor()
    for k in range(3):
        s.origin[k] = float(origin[k])
    s.near_clip, s.width, s.height, s.spp, s.seed = float(near_clip), int(width), int(height), int(spp), int(seed)
    return s


def spherical_rays(sensor, first: int, n: int, device, jitter: Optional[torch.Tensor] = None):
    """SphericalCamera.sample_ray for film samples [first, first+n): (o (3,n), v (3,n), pos (n,2))."""
    o = torch.empty(3, n, device=device)
    v = torch.empty(3, n, device=device)
    pos = torch.empty(n, 2, device=device)
    st = C.c_void_p(torch.cuda.current_stream(device).cuda_stream)
    L.check(L.load().nemit_spherical_rays(C.byref(sensor), first, n, _ptr(jitter), _ptr(o), _ptr(v), _ptr(pos), st),
            "nemit_spherical_rays")
    return o, v, pos


def film_accumulate(accum: torch.Tensor, pos: torch.Tensor, rgb: torch.Tensor, filter: str = "gaussian",
                    stddev: float = 0.25):
    h, w = accum.shape[:2]
    st = C.c_void_p(torch.cuda.current_stream(accum.device).cuda_stream)
    L.check(L.load().nemit_film_accumulate(w, h, FILTERS[filter], float(stddev), _ptr(pos), _ptr(rgb), pos.shape[0],
                                           _ptr(accum), st), "nemit_film_accumulate")
    return accum


def film_develop(accum: torch.Tensor) -> torch.Tensor:
    h, w = accum.shape[:2]
    out = torch.empty_like(accum)
    st = C.c_void_p(torch.cuda.current_stream(accum.device).cuda_stream)
    L.check(L.load().nemit_film_develop(w, h, _ptr(accum), _ptr(out), st), "nemit_film_develop")
    return out


def bake_envmap(E: NerfEmitter, width: int = 2048, height: int = 1024, spp: int = 16, origin=(0.5, 0.5, 0.5),
                filter: str = "gaussian", stddev: float = 0.25, seed: int = 0, rows: Optional[Tuple[int, int]] = None,
                camera_idx: int = 0, group=None, develop: bool = True, reduce: bool = True) -> torch.Tensor:
    """(height, width, 4) rgba radiance seen from ``origin``.  ``rows`` = this rank's film rows (default: the
    ``parallel.shard_range`` split over the process group when torch.distributed is initialised, else all rows).
    With several ranks the accumulation buffers are summed through the library's collective (``reduce``)."""
    import torch.distributed as dist
    from . import parallel
    sensor = make_sensor(width, height, spp, origin, 0.0, seed)
    world = dist.get_world_size(group) if dist.is_available() and dist.is_initialized() else 1
    if rows is None:
        rank = dist.get_rank(group) if world > 1 else 0
        rows = parallel.shard_range(height, rank, world)
    accum = torch.zeros(height, width, 4, device=E.device)
    need = int(E.lib.nemit_bake_workspace_bytes(E._h))
    if E._ws is None or E._ws.numel() < need:
        E._ws = torch.empty(need, dtype=torch.uint8, device=E.device)
    L.check(E.lib.nemit_bake_envmap(E._h, C.byref(sensor), FILTERS[filter], float(stddev), int(rows[0]), int(rows[1]),
                                    int(camera_idx), _ptr(accum), _ptr(E._ws), E._ws.numel(), E._stream()),
            "nemit_bake_envmap")
    if world > 1 and reduce:
        E.allreduce_blob(accum, group=group)
    return film_develop(accum) if develop else accum
