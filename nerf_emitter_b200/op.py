"""Drop-in for ``nerf_emitter_op`` (DS/emitters/nerf_emitter_op.py:18-148).

``make_nerf_emitter_op(emitter, ...)`` returns a callable with the reference's call signature
``op(_o, _v, _near, _dummy_nerf_parameter, _guiding_weight, camera_idx)`` that ``NeRFEmitter.set_op`` accepts
(DS/emitters/nerf.py:50-51, NS/pipelines/mitsuba_sdf.py:411-419).  With Dr.Jit present it wraps the three modes in a
``dr.CustomOp`` exactly like the reference; Dr.Jit is absent on this box, so the same three modes are also exposed
for torch tensors through ``NerfEmitterFunction`` (reverse mode via autograd, forward mode via ``emitter_jvp``).
All arithmetic goes through the C ABI (``NerfEmitter``); nothing here computes radiance.
"""
from __future__ import annotations

from typing import Optional

import torch

from .emitter import NerfEmitter


class _NoGuiding:
    """PathGuiding's no-op hooks (NS/path_guiding/path_guiding.py:65-75)."""

    def train_primal(self, *a):
        pass

    def train_forward(self, *a):
        pass

    def train_backward(self, *a):
        pass


def _near_arg(near):
    if isinstance(near, torch.Tensor):
        return near.reshape(-1)
    return float(near)


def consume_sampler_updates(sampler, num_calls: int):
    """Advance the ``updated`` state machine of a ``ProposalNetworkSampler`` - the reference's own object or the mirror
    in ``nerfstudio_api`` - by ``num_calls`` sampler calls and return the flag of each (ray_samplers.py:598,622-623,
    SURVEY.md 8a quirk 13).  ``None`` (no sampler object kept) means "always updated"."""
    if sampler is None:
        return [True] * int(num_calls)
    flags = []
    for _ in range(int(num_calls)):
        updated = sampler._steps_since_update > sampler.update_sched(sampler._step) or sampler._step < 10
        if updated:
            sampler._steps_since_update = 0
        flags.append(bool(updated))
    return flags


def _chunks(n: int, chunk: Optional[int]) -> int:
    return 1 if not chunk else (int(n) + int(chunk) - 1) // int(chunk)


def emitter_primal(E: NerfEmitter, o3n, v3n, near=0.0, camera_idx=0, jitter=None, keep_edges=False):
    """NeRFEmitterOp.ns_render_ray (nerf_emitter_op.py:38-58): (3,N),(3,N) -> (N,3).
    keep_edges: -> (rgb, kept) with the sample edges the adjoint of the same op instance re-uses instead of sampling again
    (the reference re-derives the same samples: the generator state is restored between primal and adjoint,
    NS/models/base_model.py:299,327-328)."""
    return E.render(o3n.contiguous(), v3n.contiguous(), near=_near_arg(near), camera_idx=camera_idx, jitter=jitter,
                    keep_edges=keep_edges)


def emitter_jvp(E: NerfEmitter, o3n, v3n, grad_o3n, grad_v3n, near=0.0, camera_idx=0, jitter=None, kept=None):
    """NeRFEmitterOp.ns_forward_ray (nerf_emitter_op.py:60-90): tangents (3,N) -> (N,3), NaN tangents zeroed."""
    return E.jvp(o3n.contiguous(), v3n.contiguous(), grad_o3n.contiguous(), grad_v3n.contiguous(),
                 near=_near_arg(near), camera_idx=camera_idx, jitter=jitter, kept=kept)


def emitter_vjp(E: NerfEmitter, o3n, v3n, grad_out_n3, near=0.0, camera_idx=0, jitter=None, chunk_rays=None,
                loss_scale=1.0, proposal_requires_grad=True, kept=None):
    """NeRFEmitterOp.ns_backward_ray (nerf_emitter_op.py:92-117): returns (o.grad, v.grad) as (3,N); parameter
    gradients are accumulated into the emitter's bound gradient buffers (the reference leaves them in
    ``param.grad``)."""
    return E.backward(o3n.contiguous(), v3n.contiguous(), grad_out_n3.contiguous(), near=_near_arg(near),
                      camera_idx=camera_idx, jitter=jitter, chunk_rays=chunk_rays, loss_scale=loss_scale,
                      proposal_requires_grad=proposal_requires_grad, kept=kept)


class NerfEmitterFunction(torch.autograd.Function):
    """torch twin of the CustomOp: ``rgb = NerfEmitterFunction.apply(emitter, o3n, v3n, dummy, near, camera_idx,
    jitter, chunk_rays)``; ``rgb.backward(g)`` fills o/v gradients and accumulates parameter gradients."""

    @staticmethod
    def forward(ctx, E, o3n, v3n, dummy, near, camera_idx, jitter, chunk_rays, sampler=None):
        ctx.E, ctx.near, ctx.camera_idx, ctx.jitter, ctx.chunk_rays = E, near, camera_idx, jitter, chunk_rays
        ctx.sampler = sampler
        ctx.save_for_backward(o3n, v3n, dummy)
        consume_sampler_updates(sampler, _chunks(o3n.shape[1], chunk_rays))      # the no-grad primal chunks count too
        rgb, ctx.kept = emitter_primal(E, o3n.detach(), v3n.detach(), near, camera_idx, jitter, keep_edges=True)
        return rgb

    @staticmethod
    def backward(ctx, grad_out):
        o3n, v3n, dummy = ctx.saved_tensors
        flags = consume_sampler_updates(ctx.sampler, _chunks(o3n.shape[1], ctx.chunk_rays))
        go, gv = emitter_vjp(ctx.E, o3n.detach(), v3n.detach(), grad_out, ctx.near, ctx.camera_idx, ctx.jitter,
                             ctx.chunk_rays, proposal_requires_grad=flags, kept=ctx.kept)
        ctx.kept = None
        gd = torch.zeros_like(dummy) if dummy is not None else None       # nerf_emitter_op.py:116
        return None, go, gv, gd, None, None, None, None, None


def make_nerf_emitter_op(E: NerfEmitter, path_guiding=None, jitter_fn=None, chunk_rays: Optional[int] = 16384,
                         sampler=None, forward_ad_chunk_rays: Optional[int] = 4096):
    """The callable to hand to ``NeRFEmitter.set_op``.  ``jitter_fn(n)`` returns the (n,3) per-ray jitter drawn from
    the samplers' generators in training mode (the reference re-seeds them between primal and adjoint,
    NS/models/base_model.py:299,327-328); ``chunk_rays`` = ``backward_num_rays_per_chunk`` (= ``eval_num_rays_per_chunk``
    in sdf-nerfacto, method_configs.py:315-316).  ``sampler``: the model's ``ProposalNetworkSampler`` object (the
    reference's own, which the trainer keeps stepping through ``step_cb``, or the mirror): its ``updated`` state
    machine then decides per chunk whether the proposal networks get a gradient, and every primal / tangent chunk
    advances it, as in the reference (``consume_sampler_updates``); without it every backward chunk updates them."""
    guiding = path_guiding or _NoGuiding()
    try:
        import drjit as dr          # noqa: F401
        import mitsuba as mi        # noqa: F401
    except ImportError:
        dr = None

    def torch_op(o3n, v3n, near, dummy, guiding_weight, camera_idx):
        jit = jitter_fn(o3n.shape[1]) if (jitter_fn is not None and E.training) else None
        rgb = NerfEmitterFunction.apply(E, o3n, v3n, dummy, near, camera_idx, jit, chunk_rays, sampler)
        guiding.train_primal(o3n.t(), v3n.t(), rgb, guiding_weight)
        return rgb.reshape(-1)

    if dr is None:
        return torch_op

    def drjit_op(_o, _v, _near, _dummy, _guiding_weight, camera_idx):     # pragma: no cover (no Dr.Jit on this box)
        near = _near.torch() if hasattr(_near, "torch") else _near
        gw = _guiding_weight.torch().unsqueeze(-1)
        state = {}

        class NeRFEmitterOp(dr.CustomOp):
            def eval(self, o, v, dummy):
                self.o, self.v = o.torch(), v.torch()
                state["jit"] = jitter_fn(self.o.shape[1]) if (jitter_fn is not None and E.training) else None
                consume_sampler_updates(sampler, _chunks(self.o.shape[1], chunk_rays))
                rgb, state["kept"] = emitter_primal(E, self.o, self.v, near, camera_idx, state["jit"], keep_edges=True)
                guiding.train_primal(self.o.t(), self.v.t(), rgb, gw)
                return type(o)(rgb.reshape(-1))

            def forward(self):
                go, gv, _ = self.grad_in("args")
                consume_sampler_updates(sampler, _chunks(self.o.shape[1], forward_ad_chunk_rays))
                t = emitter_jvp(E, self.o, self.v, go.torch().reshape(3, -1), gv.torch().reshape(3, -1), near,
                                camera_idx, state["jit"], kept=state.get("kept"))
                guiding.train_forward(self.o.t(), self.v.t(), t, gw)
                self.set_grad_out(type(go)(t.reshape(-1)))

            def backward(self):
                g = self.grad_out().torch().reshape(-1, 3)
                flags = consume_sampler_updates(sampler, _chunks(self.o.shape[1], chunk_rays))
                go, gv = emitter_vjp(E, self.o, self.v, g, near, camera_idx, state["jit"], chunk_rays,
                                     proposal_requires_grad=flags, kept=state.get("kept"))
                guiding.train_backward(self.o.t(), self.v.t(), go.t(), gw)
                T = type(self.grad_out())
                self.set_grad_in("args", (T(go), T(gv), T(torch.zeros(1, device=go.device))))

        return dr.custom(NeRFEmitterOp, (_o, _v, _dummy))

    return drjit_op
