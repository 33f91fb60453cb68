"""Headless mirror of the Mitsuba-side emitter plugin ``NeRFEmitter`` (differentiable-sdf-rendering/python/emitters/
nerf.py:20-101): what sits between the integrator's ``scene.environment().eval(si)`` and the
emitter op.  Dr.Jit arrays become torch CUDA tensors in the same ``(3,N)`` struct-of-arrays layout
(``vector3f_2_tensor``, nerf.py:12-17); the ray transform to emitter-local space and the ``outside_only`` skip distance
are one CUDA launch (``nemit_emitter_local_rays``); everything else is the op (``op.make_nerf_emitter_op``).

Mitsuba's ``BoundingBox3f.ray_intersect`` and ``Transform4f @ Ray3f`` are third-party arithmetic (mitsuba 3.4.1, not
in the reference tree): restated from their published definition, **parity unpinned** (DESIGN.md section 4).
"""
from __future__ import annotations

import ctypes as C
from typing import Callable, Optional

import torch

from . import lib as L


def _m34(matrix4: torch.Tensor):
    m = matrix4.detach().to("cpu", torch.float64)
    return (C.c_float * 12)(*[float(m[r, c]) for r in range(3) for c in range(4)])


def _launch(m34, bmin, bmax, skip, outside_only, linear_only, o, v, want_near):
    lib = L.load()
    n = o.shape[1]
    o_out, v_out = torch.empty_like(o), torch.empty_like(v)
    near = torch.empty(n, device=o.device, dtype=torch.float32) if want_near else None
    f3 = C.c_float * 3
    stream = C.c_void_p(torch.cuda.current_stream(o.device).cuda_stream)
    with torch.cuda.device(o.device):
        L.check(lib.nemit_emitter_local_rays(m34, f3(*bmin), f3(*bmax), float(skip), int(outside_only), int(linear_only),
                                             o.data_ptr(), v.data_ptr(), n, o_out.data_ptr(), v_out.data_ptr(),
                                             near.data_ptr() if want_near else None, stream),
                "nemit_emitter_local_rays")
    return o_out, v_out, near


class _ToLocal(torch.autograd.Function):
    """(o, v) world -> local; the backward applies the transposed 3x3 block to both gradients.  ``near`` carries no
    gradient: the reference hands it to the op outside the CustomOp's differentiable arguments (nerf.py:62-63)."""

    @staticmethod
    def forward(ctx, o, v, emitter):
        ctx.emitter = emitter
        o_l, v_l, near = _launch(_m34(emitter.world_to_local), emitter.bbox_min, emitter.bbox_max, emitter.skip_distance,
                                 emitter.outside_only, False, o, v, True)
        ctx.mark_non_differentiable(near)
        return o_l, v_l, near

    @staticmethod
    def backward(ctx, go, gv, _gnear):
        e = ctx.emitter
        mt = torch.eye(4, dtype=torch.float64)
        mt[:3, :3] = e.world_to_local[:3, :3].to(torch.float64).t()
        go_w, gv_w, _ = _launch(_m34(mt), e.bbox_min, e.bbox_max, 0.0, False, True, go.contiguous(), gv.contiguous(), False)
        return go_w, gv_w, None


class NeRFEmitter:
    """``to_world``: 4x4 world transform of the emitter (``world_transform()``); ``scale``: the env-map emitter's
    ``m_scale`` (envmap.py:97); the other arguments are the plugin's properties (nerf.py:32-40)."""

    def __init__(self, to_world: Optional[torch.Tensor] = None, scale: float = 1.0, padding_size: float = 0.0,
                 skip_distance: float = 0.0, outside_only: bool = False, detach_op: bool = False) -> None:
        self.set_world_transform(torch.eye(4) if to_world is None else to_world)
        self.m_scale = float(scale)
        self.bbox_min = (0.0 - padding_size,) * 3            # nerf.py:33
        self.bbox_max = (1.0 + padding_size,) * 3
        self.skip_distance = float(skip_distance)
        self.outside_only = bool(outside_only)
        self.detach_op = bool(detach_op)
        self.nerf_emitter_op: Optional[Callable] = None
        self.guiding_weight = None
        self.camera_idx = 0
        self.dummy_nerf_parameter: Optional[torch.Tensor] = None          # created on first use on the rays' device

    def set_world_transform(self, to_world: torch.Tensor) -> None:
        """``params['to_world'] = ...`` (Rotater.apply_sdf_scene, rotater.py:118-127)."""
        self.to_world = to_world.detach().to("cpu", torch.float64)
        self.world_to_local = torch.linalg.inv(self.to_world)

    def set_guiding_weight(self, guiding_weight) -> None:
        self.guiding_weight = guiding_weight

    def set_op(self, op: Callable) -> None:
        self.nerf_emitter_op = op

    def set_camera_idx(self, camera_idx: int) -> None:
        self.camera_idx = int(camera_idx)

    def eval_nerf_ray(self, o3n: torch.Tensor, v3n: torch.Tensor, near) -> torch.Tensor:
        """nerf.py:59-67: local rays (3,N) + distance to skip -> radiance (N,3) times ``m_scale``."""
        if self.nerf_emitter_op is None:
            raise L.NemitError("NeRFEmitter: no op installed (set_op)")
        n = o3n.shape[1]
        if self.dummy_nerf_parameter is None or self.dummy_nerf_parameter.device != o3n.device:
            self.dummy_nerf_parameter = torch.full((1,), float("nan"), device=o3n.device, requires_grad=True)  # nerf.py:37
        gw = self.guiding_weight if self.guiding_weight is not None else torch.ones(n, device=o3n.device)
        res = self.nerf_emitter_op(o3n, v3n, near, self.dummy_nerf_parameter, gw, self.camera_idx)
        if self.detach_op:
            res = res.detach()
        return res.reshape(n, 3) * self.m_scale

    def to_local(self, o: torch.Tensor, v: torch.Tensor):
        """nerf.py:70-79: world rays (3,N) -> (local o, local v, distance to skip (N,)); differentiable in o, v."""
        if not (o.is_cuda and v.is_cuda):
            raise L.NemitError("NeRFEmitter needs CUDA tensors; there is no CPU path")
        return _ToLocal.apply(o.to(torch.float32).contiguous(), v.to(torch.float32).contiguous(), self)

    def eval_nerf(self, o: torch.Tensor, v: torch.Tensor) -> torch.Tensor:
        """nerf.py:69-82: world-space points / directions (3,N) -> radiance (N,3)."""
        o_l, v_l, near = self.to_local(o, v)
        return self.eval_nerf_ray(o_l, v_l, near)

    def eval(self, p: torch.Tensor, wi: torch.Tensor) -> torch.Tensor:
        """nerf.py:56-57: ``si.p``, ``si.wi`` as (3,N) tensors."""
        return self.eval_nerf(p, -wi)
