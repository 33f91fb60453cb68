"""Host side of the emitter path: owns a libnerfemitter handle and mirrors the three modes of the
reference's ``NeRFEmitterOp`` (DS/emitters/nerf_emitter_op.py:38-117) on torch CUDA tensors.

Parameters stay ordinary torch tensors named as in the reference ``state_dict``; the
library borrows their device pointers.  There is no fallback: every call goes through the C ABI.
"""
from __future__ import annotations

import ctypes as C
from typing import Dict, Optional, Sequence

import torch

from . import lib as L
from .params import EmitterModelSpec, all_param_names, param_shapes, prop_param_names


def _ptr(t: Optional[torch.Tensor]):
    return None if t is None else C.c_void_p(t.data_ptr())


def _check(t: torch.Tensor, shape, name, dtype=torch.float32):
    if t.dtype != dtype or not t.is_cuda or not t.is_contiguous() or tuple(t.shape) != tuple(shape):
        raise ValueError("%s must be a contiguous CUDA %s tensor of shape %s, got %s %s %s" % (
            name, dtype, tuple(shape), t.dtype, tuple(t.shape), t.device))


def sampler_tables(spec: EmitterModelSpec):
    """The samplers' linspace tables, computed by torch on the CPU so that they are bit-identical to
    NS/model_components/ray_samplers.py:104 and :323-331."""
    s0, s1 = spec.num_proposal_samples
    s2 = spec.num_nerf_samples
    lin0 = torch.linspace(0.0, 1.0, s0 + 1)
    u1 = torch.linspace(0.0, 1.0 - (1.0 / (s1 + 1)), steps=s1 + 1)
    u2 = torch.linspace(0.0, 1.0 - (1.0 / (s2 + 1)), steps=s2 + 1)
    return lin0.contiguous(), u1.contiguous(), u2.contiguous()


class NerfEmitter:
    """B200 evaluation of the nerfacto emitter: ``render`` / ``jvp`` / ``backward`` take Mitsuba-space
    rays as (3,N) fp32 CUDA tensors exactly like the dr.CustomOp and return (N,3) radiance."""

    def __init__(self, params: Dict[str, torch.Tensor], spec: Optional[EmitterModelSpec] = None,
                 device: Optional[torch.device] = None, scene_scale: float = 1.0,
                 disable_inside_aabb: bool = True, interlevel_mult: float = 1.0, distortion_mult: float = 7e-3,
                 cpu_sum_width: int = 8):
        self.lib = L.load()
        self.spec = spec = spec or EmitterModelSpec()
        if device is None:
            device = torch.device("cuda", torch.cuda.current_device())
        self.device = torch.device(device)
        if self.device.type != "cuda":
            raise L.NemitError("NerfEmitter needs a CUDA device; there is no CPU path")
        self.scene_scale = float(scene_scale)
        self.training = False
        self.anneal = 1.0
        cfg = L.NemitConfig()
        for g, dst in [(spec.main_grid, cfg.main_grid), (spec.prop_grids[0], cfg.prop_grid[0]),
                       (spec.prop_grids[1], cfg.prop_grid[1])]:
            dst.num_levels = g.num_levels
            dst.log2_table_size = g.log2_hashmap_size
            for i, s in enumerate(g.scalings()):
                dst.scalings[i] = s
        cfg.num_prop_samples[0], cfg.num_prop_samples[1] = spec.num_proposal_samples
        cfg.num_nerf_samples = spec.num_nerf_samples
        cfg.hidden_dim, cfg.geo_feat_dim, cfg.appearance_dim = spec.hidden_dim, spec.geo_feat_dim, spec.appearance_dim
        cfg.prop_hidden_dim, cfg.num_images = spec.prop_hidden, spec.num_images
        cfg.near_plane, cfg.far_plane, cfg.scene_scale = spec.near_plane, spec.far_plane, scene_scale
        for i, v in enumerate([-scene_scale] * 3 + [scene_scale] * 3):
            cfg.aabb[i] = v
        cfg.rgb_bias = spec.rgb_bias
        cfg.disable_inside_aabb = int(disable_inside_aabb)
        cfg.cpu_sum_width = cpu_sum_width
        cfg.interlevel_mult, cfg.distortion_mult = interlevel_mult, distortion_mult
        self._tables = sampler_tables(spec)
        cfg.lin0_host = C.cast(self._tables[0].data_ptr(), L.c_float_p)
        cfg.u1_host = C.cast(self._tables[1].data_ptr(), L.c_float_p)
        cfg.u2_host = C.cast(self._tables[2].data_ptr(), L.c_float_p)
        self._h = C.c_void_p()
        L.check(self.lib.nemit_create(C.byref(cfg), self.device.index or 0, C.byref(self._h)), "nemit_create")
        self._ws = None
        self.params: Dict[str, torch.Tensor] = {}
        self.grads: Dict[str, torch.Tensor] = {}
        self.bind_params(params)

    def __del__(self):
        h = getattr(self, "_h", None)
        if h is not None and h.value:
            self.lib.nemit_destroy(h)
            self._h = C.c_void_p()

    # -- parameters ------------------------------------------------------------------------
    def _pack(self, tensors: Dict[str, torch.Tensor]) -> L.NemitParams:
        shapes = param_shapes(self.spec)
        for n in all_param_names():
            if n not in tensors:
                raise KeyError("missing parameter %s" % n)
            _check(tensors[n], shapes[n], n)
        P = L.NemitParams()
        main = ["field.mlp_base_grid.hash_table",
                "field.mlp_base_mlp.layers.0.weight", "field.mlp_base_mlp.layers.0.bias",
                "field.mlp_base_mlp.layers.1.weight", "field.mlp_base_mlp.layers.1.bias",
                "field.mlp_head.layers.0.weight", "field.mlp_head.layers.0.bias",
                "field.mlp_head.layers.1.weight", "field.mlp_head.layers.1.bias",
                "field.mlp_head.layers.2.weight", "field.mlp_head.layers.2.bias",
                "field.embedding_appearance.embedding.weight"]
        for f, n in zip(L.PARAM_FIELDS, main):
            setattr(P, f, tensors[n].data_ptr())
        for i in range(2):
            names = prop_param_names(i)
            P.prop_table[i] = tensors[names[0]].data_ptr()
            P.prop_w0[i] = tensors[names[1]].data_ptr()
            P.prop_b0[i] = tensors[names[2]].data_ptr()
            P.prop_w1[i] = tensors[names[3]].data_ptr()
            P.prop_b1[i] = tensors[names[4]].data_ptr()
        return P

    def bind_params(self, params: Dict[str, torch.Tensor]):
        """Borrow the parameter tensors (moved to the device once if they are not there yet)."""
        self.params = {k: v.detach().to(self.device, torch.float32).contiguous() for k, v in params.items()}
        self._P = self._pack(self.params)
        L.check(self.lib.nemit_bind_params(self._h, C.byref(self._P)), "nemit_bind_params")

    def bind_grads(self, grads: Optional[Dict[str, torch.Tensor]] = None):
        """Bind gradient accumulators (``param.grad`` tensors). Created zero-filled if not given."""
        if grads is None:
            # one flat fp32 buffer, every gradient tensor a view into it: the multi-GPU all-reduce is one call
            sizes = [(k, v.numel()) for k, v in self.params.items()]
            self.flat_grads = torch.zeros(sum((n + 63) // 64 * 64 for _, n in sizes), device=self.device)
            grads, off = {}, 0
            for k, n in sizes:
                grads[k] = self.flat_grads[off:off + n].view_as(self.params[k])
                off += (n + 63) // 64 * 64
        else:
            self.flat_grads = None
        self.grads = grads
        self._G = self._pack(self.grads)
        L.check(self.lib.nemit_bind_grads(self._h, C.byref(self._G)), "nemit_bind_grads")
        return grads

    def zero_grads(self):
        if getattr(self, "flat_grads", None) is not None:
            self.flat_grads.zero_()
            return
        for g in self.grads.values():
            g.zero_()

    def init_comm(self, group=None):
        """Create the library's own NCCL communicator over the ranks of the (initialised) torch.distributed group:
        rank 0 draws the unique id (nemit_comm_unique_id), torch.distributed only carries its 128 bytes to the others."""
        import torch.distributed as dist
        rank, world = dist.get_rank(group), dist.get_world_size(group)
        ident = torch.zeros(L.UNIQUE_ID_BYTES, dtype=torch.uint8)
        if rank == 0:
            L.check(self.lib.nemit_comm_unique_id(C.c_void_p(ident.data_ptr())), "nemit_comm_unique_id")
        carrier = ident.to(self.device) if dist.get_backend(group) == "nccl" else ident
        dist.broadcast(carrier, src=dist.get_global_rank(group, 0) if group is not None else 0, group=group)
        ident = carrier.cpu().contiguous()
        L.check(self.lib.nemit_comm_init(self._h, C.c_void_p(ident.data_ptr()), rank, world), "nemit_comm_init")
        self.comm_world = world
        return self

    def allreduce_blob(self, blob: torch.Tensor, group=None, scale=None):
        """In-place sum over the ranks of any contiguous fp32 CUDA tensor through nemit_allreduce_grads (the bake's film
        accumulators use the same collective as the gradients)."""
        import torch.distributed as dist
        if blob.dtype != torch.float32 or not blob.is_cuda or not blob.is_contiguous():
            raise ValueError("allreduce_blob needs a contiguous fp32 CUDA tensor")
        if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1 \
                and getattr(self, "comm_world", 0) != dist.get_world_size(group):
            self.init_comm(group)
        L.check(self.lib.nemit_allreduce_grads(self._h, _ptr(blob), blob.numel(), 1.0 if scale is None else float(scale),
                                               self._stream()), "nemit_allreduce_grads")
        return blob

    def allreduce_grads(self, group=None, scale=None):
        """The one collective of the path: sum parameter gradients over the ray-sharded ranks, through the library
        (nemit_allreduce_grads: NCCL all-reduce with the unscale folded in as a pre-multiplied sum, one pass over the
        blob).  The communicator is created on first use when torch.distributed is initialised."""
        if getattr(self, "flat_grads", None) is None:
            raise L.NemitError("allreduce_grads needs the flat gradient blob: call bind_grads() without arguments")
        return self.allreduce_blob(self.flat_grads, group, scale)

    def set_precision(self, precision: str = "fp32"):
        """'fp32' = exact FMA path (parity mode); 'fp16' = fp16-operand tensor-core MLP (tcgen05, fp32 accumulate);
        'split' = fp32 MLP arithmetic on the tensor cores through split bf16 operands (forward: three pieces = 24
        significant bits; reverse: two pieces + the forward's ReLU masks)."""
        code = {"fp32": 0, "fp16": 1, "split": 2}[precision]
        L.check(self.lib.nemit_set_precision(self._h, code), "nemit_set_precision")
        self.precision = precision
        return self

    def set_sampler(self, sampler: str = "exact"):
        """'exact' = the reference's fp32 CPU arithmetic (bit-equal searchsorted indices, default); 'fast' = proposal
        density fields in packed half precision (the reference's autocast / tcnn behaviour)."""
        code = {"exact": 0, "fast": 1}[sampler]
        L.check(self.lib.nemit_set_sampler(self._h, code), "nemit_set_sampler")
        self.sampler = sampler
        return self

    def set_appearance_mode(self, always_use_camera_indices: bool = False,
                            use_average_appearance_embedding: bool = True, rotation_ids=None):
        """The model built with a ``Rotater`` (turntable captures, NS/models/nerfacto.py:186-206): the appearance
        embedding has one row per rotation, the row is used in eval mode too (``always_use_camera_indices``) and
        ``rotation_ids`` (``Rotater.rotation_ids``) maps camera index -> rotation id on every ray-level call exactly
        as ``Rotater.map_rotation_ids`` does (NS/field_components/rotater.py:94-100)."""
        ids = None if rotation_ids is None else [int(x) for x in rotation_ids]
        n = 0 if ids is None else len(ids)
        arr = (C.c_int32 * n)(*ids) if n else None
        L.check(self.lib.nemit_set_appearance_mode(self._h, int(bool(always_use_camera_indices)),
                                                   int(bool(use_average_appearance_embedding)), arr, n),
                "nemit_set_appearance_mode")
        self.always_use_camera_indices = bool(always_use_camera_indices)
        self.use_average_appearance_embedding = bool(use_average_appearance_embedding)
        self.rotation_ids = ids
        return self

    def set_rotater(self, transforms=None, center=None, radius=None, aabb=None):
        """Per-sample frustum rotation of turntable captures, ``Rotater.apply_frustums`` (NS/field_components/rotater.py:
        125-149).  ``transforms``: (R, 4, 4) or (R, 3, 4) = ``Rotater.get_torch_transforms(arange(R))`` in NeRF
        coordinates; the bound is the sphere (``center`` (3,), ``radius``; the reference's default: the scene centre and
        sqrt(3) * scene_scale, rotater.py:64-66) or the box ``aabb`` (2,3).  ``transforms=None`` removes it.  Which
        stages rotate follows the reference (include/nemit.h, nemit_set_rotater)."""
        if transforms is None:
            L.check(self.lib.nemit_set_rotater(self._h, 0, None, 0, None), "nemit_set_rotater")
            self.rotater = None
            return self
        T = torch.as_tensor(transforms, dtype=torch.float32).cpu()[:, :3, :4].contiguous()
        if aabb is not None:
            b = torch.as_tensor(aabb, dtype=torch.float32).reshape(6)
            kind = 2
        else:
            if center is None or radius is None:
                raise ValueError("set_rotater needs either (center, radius) or aabb")
            b = torch.cat([torch.as_tensor(center, dtype=torch.float32).reshape(3), torch.tensor([float(radius)]),
                           torch.zeros(2)])
            kind = 1
        b = b.contiguous()
        L.check(self.lib.nemit_set_rotater(self._h, T.shape[0], C.cast(T.data_ptr(), L.c_float_p), kind,
                                           C.cast(b.data_ptr(), L.c_float_p)), "nemit_set_rotater")
        self.rotater = dict(transforms=T, bound=b, kind=kind)
        return self

    def train(self, mode: bool = True):
        self.training = bool(mode)
        return self

    def eval(self):
        return self.train(False)

    # -- plumbing --------------------------------------------------------------------------
    def _workspace(self, n: int) -> torch.Tensor:
        need = int(self.lib.nemit_workspace_bytes(self._h, n))
        if self._ws is None or self._ws.numel() < need:
            self._ws = torch.empty(need, dtype=torch.uint8, device=self.device)
        return self._ws

    def _stream(self):
        return C.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)

    def _rays(self, o3n, v3n, near, camera_idx, jitter, nerf_layout: bool = False, kept=None) -> L.NemitRays:
        """nerf_layout: o, v are RayBundle.origins / .directions, (N,3) rows in NeRF coordinates (NEMIT_RAYS_NERF_N3)
        instead of the CustomOp's (3,N) Mitsuba unit-cube arrays."""
        n = o3n.shape[0] if nerf_layout else o3n.shape[1]
        _check(o3n, (n, 3) if nerf_layout else (3, n), "o")
        _check(v3n, (n, 3) if nerf_layout else (3, n), "v")
        R = L.NemitRays()
        R.layout = int(nerf_layout)
        R.o, R.v, R.n = o3n.data_ptr(), v3n.data_ptr(), n
        self._keep = [o3n, v3n]
        if isinstance(near, torch.Tensor):
            near = near.reshape(-1).to(self.device, torch.float32).contiguous()
            _check(near, (n,), "near")
            R.near_per_ray = near.data_ptr()
            self._keep.append(near)
        else:
            R.near_per_ray = None
            R.near_scalar = float(near)
        R.camera_idx = int(camera_idx)
        R.training = int(self.training)
        if jitter is not None and self.training:
            if isinstance(jitter, (list, tuple)):
                jitter = torch.cat([j.reshape(n, 1) for j in jitter], dim=1)
            jitter = jitter.to(self.device, torch.float32).contiguous()
            _check(jitter, (n, 3), "jitter")
            R.jitter = jitter.data_ptr()
            self._keep.append(jitter)
        else:
            R.jitter = None
        R.anneal = float(self.anneal)
        if kept is not None:
            # sample edges kept from the primal of the same CustomOp instance (render(..., keep_edges=True))
            s2 = self.spec.num_nerf_samples + 1
            _check(kept["t_edges2"], (n, s2), "kept t_edges2")
            _check(kept["s_edges2"], (n, s2), "kept s_edges2")
            R.kept_t_edges2, R.kept_s_edges2 = kept["t_edges2"].data_ptr(), kept["s_edges2"].data_ptr()
            if kept.get("s_edges1") is not None:
                _check(kept["s_edges1"], (n, self.spec.num_proposal_samples[1] + 1), "kept s_edges1")
                R.kept_s_edges1 = kept["s_edges1"].data_ptr()
            self._keep.append(kept)
        return R

    def _levels_out(self, n: int, want_weights2: bool):
        s0, s1 = self.spec.num_proposal_samples
        s2 = self.spec.num_nerf_samples
        f = dict(device=self.device, dtype=torch.float32)
        out = dict(weights0=torch.empty(n, s0, **f), weights1=torch.empty(n, s1, **f),
                   s_edges1=torch.empty(n, s1 + 1, **f), s_edges2=torch.empty(n, s2 + 1, **f),
                   t_edges2=torch.empty(n, s2 + 1, **f),
                   inds1=torch.empty(n, s1 + 1, device=self.device, dtype=torch.int32),
                   inds2=torch.empty(n, s2 + 1, device=self.device, dtype=torch.int32))
        if want_weights2:
            out["weights2"] = torch.empty(n, s2, **f)
        LO = L.NemitLevelsOut()
        for k, t in out.items():
            setattr(LO, k, t.data_ptr())
        return out, LO

    # -- the three modes -------------------------------------------------------------------
    def render(self, o3n: torch.Tensor, v3n: torch.Tensor, near=0.0, camera_idx: int = 0, jitter=None,
               return_aux: bool = False, nerf_layout: bool = False, keep_edges: bool = False):
        """NeRFEmitterOp.ns_render_ray (nerf_emitter_op.py:38-58): (3,N),(3,N) -> rgb (N,3).
        keep_edges: also return {s_edges1, s_edges2, t_edges2} (780 B per ray) for the adjoint of the same CustomOp
        instance (``backward(..., kept=...)`` then skips the proposal sampling pass; valid while mode, near, anneal,
        jitter and parameters are unchanged - the reference re-derives the same samples from restored generator state)."""
        n = o3n.shape[0] if nerf_layout else o3n.shape[1]
        R = self._rays(o3n, v3n, near, camera_idx, jitter, nerf_layout)
        rgb = torch.empty(n, 3, device=self.device, dtype=torch.float32)
        ws = self._workspace(n)
        if keep_edges and not return_aux:
            s1, s2 = self.spec.num_proposal_samples[1] + 1, self.spec.num_nerf_samples + 1
            kept = dict(s_edges1=torch.empty(n, s1, device=self.device), s_edges2=torch.empty(n, s2, device=self.device),
                        t_edges2=torch.empty(n, s2, device=self.device), training=self.training, anneal=self.anneal)
            LO = L.NemitLevelsOut()
            for k in ("s_edges1", "s_edges2", "t_edges2"):
                setattr(LO, k, kept[k].data_ptr())
            L.check(self.lib.nemit_forward(self._h, C.byref(R), _ptr(rgb), None, C.byref(LO), _ptr(ws), ws.numel(),
                                           self._stream()), "nemit_forward")
            return rgb, kept
        if not return_aux:
            L.check(self.lib.nemit_forward(self._h, C.byref(R), _ptr(rgb), None, None, _ptr(ws), ws.numel(),
                                           self._stream()), "nemit_forward")
            return rgb
        acc = torch.empty(n, device=self.device, dtype=torch.float32)
        aux, LO = self._levels_out(n, True)
        L.check(self.lib.nemit_forward(self._h, C.byref(R), _ptr(rgb), _ptr(acc), C.byref(LO), _ptr(ws), ws.numel(),
                                       self._stream()), "nemit_forward")
        aux["rgb"] = rgb
        aux["accumulation"] = acc
        return rgb, aux

    def render_noalloc(self, o3n: torch.Tensor, v3n: torch.Tensor, rgb_out: torch.Tensor, near=0.0, camera_idx: int = 0,
                       jitter=None):
        """nemit_forward_noalloc: the forward with its scratch in a static device array of the library (at most
        lib.NOALLOC_MAX_RAYS rays); no device allocation in the call - the caller passes the output too."""
        n = o3n.shape[1]
        if not rgb_out.is_cuda or rgb_out.dtype != torch.float32 or not rgb_out.is_contiguous() or rgb_out.shape != (n, 3):
            raise ValueError("rgb_out must be a contiguous fp32 (N, 3) device tensor")
        R = self._rays(o3n, v3n, near, camera_idx, jitter, False)
        L.check(self.lib.nemit_forward_noalloc(self._h, C.byref(R), _ptr(rgb_out), self._stream()), "nemit_forward_noalloc")
        return rgb_out

    def sample_levels(self, o3n, v3n, near=0.0, camera_idx: int = 0, jitter=None, edges_only: bool = False,
                      nerf_layout: bool = False):
        """ProposalNetworkSampler.generate_ray_samples (ray_samplers.py:584-626) for emitter rays.
        edges_only: write just the final 49 euclidean edges (what the field kernel consumes)."""
        n = o3n.shape[0] if nerf_layout else o3n.shape[1]
        R = self._rays(o3n, v3n, near, camera_idx, jitter, nerf_layout)
        if edges_only:
            aux = dict(t_edges2=torch.empty(n, self.spec.num_nerf_samples + 1, device=self.device))
            LO = L.NemitLevelsOut()
            LO.t_edges2 = aux["t_edges2"].data_ptr()
        else:
            aux, LO = self._levels_out(n, False)
        ws = self._workspace(n)
        L.check(self.lib.nemit_sample_levels(self._h, C.byref(R), C.byref(LO), _ptr(ws), ws.numel(), self._stream()),
                "nemit_sample_levels")
        return aux

    def _usable_kept(self, kept):
        if kept is None:
            return None
        if bool(kept.get("training")) != self.training or float(kept.get("anneal", 1.0)) != float(self.anneal):
            return None                      # mode changed since the primal: sample again
        return kept

    def jvp(self, o3n, v3n, do3n, dv3n, near=0.0, camera_idx: int = 0, jitter=None, nerf_layout: bool = False,
            kept=None):
        """NeRFEmitterOp.ns_forward_ray (nerf_emitter_op.py:60-90): tangents (3,N) -> d_rgb (N,3)."""
        n = o3n.shape[0] if nerf_layout else o3n.shape[1]
        R = self._rays(o3n, v3n, near, camera_idx, jitter, nerf_layout, self._usable_kept(kept))
        _check(do3n, tuple(o3n.shape), "d_o")
        _check(dv3n, tuple(o3n.shape), "d_v")
        out = torch.empty(n, 3, device=self.device, dtype=torch.float32)
        ws = self._workspace(n)
        L.check(self.lib.nemit_jvp(self._h, C.byref(R), _ptr(do3n), _ptr(dv3n), _ptr(out), _ptr(ws), ws.numel(),
                                   self._stream()), "nemit_jvp")
        return out

    def backward(self, o3n, v3n, grad_rgb, near=0.0, camera_idx: int = 0, jitter=None,
                 chunk_rays: Optional[int] = None, loss_scale: float = 1.0, proposal_requires_grad: bool = True,
                 nerf_layout: bool = False, kept=None):
        """NeRFEmitterOp.ns_backward_ray (nerf_emitter_op.py:92-117): returns (grad_o, grad_v) as (3,N);
        parameter gradients are accumulated into the bound ``grads``.  ``proposal_requires_grad`` is the sampler's
        ``updated`` flag (ray_samplers.py:598): one bool for the whole call, or one bool per chunk of ``chunk_rays``
        rays (the reference evaluates it per sampler call, i.e. per chunk; runs of equal flags go to the library
        together, the regulariser means being per chunk either way)."""
        kept = self._usable_kept(kept)       # the edges render(..., keep_edges=True) returned, if the mode is unchanged
        if not self.grads:
            self.bind_grads()
        n = o3n.shape[0] if nerf_layout else o3n.shape[1]
        if not isinstance(proposal_requires_grad, (bool, int)):
            flags = [bool(f) for f in proposal_requires_grad]
            chunk = int(chunk_rays or n)
            if len(flags) != (n + chunk - 1) // max(chunk, 1):
                raise ValueError("proposal_requires_grad: one flag per chunk of chunk_rays rays expected")
            if len(set(flags)) <= 1:
                proposal_requires_grad = bool(flags[0]) if flags else True
            else:
                if isinstance(jitter, (list, tuple)):
                    jitter = torch.cat([j.reshape(n, 1) for j in jitter], dim=1)
                go, gv = torch.empty_like(o3n), torch.empty_like(o3n)
                rows = (lambda t, a, b: t[a:b]) if nerf_layout else (lambda t, a, b: t[:, a:b].contiguous())
                c0 = 0
                while c0 < len(flags):
                    c1 = c0 + 1
                    while c1 < len(flags) and flags[c1] == flags[c0]:
                        c1 += 1
                    a, b = c0 * chunk, min(c1 * chunk, n)
                    nr = near[a:b] if isinstance(near, torch.Tensor) and near.numel() > 1 else near
                    kp = None if kept is None else {k: (t[a:b] if isinstance(t, torch.Tensor) else t) for k, t in kept.items()}
                    g_o, g_v = self.backward(rows(o3n, a, b), rows(v3n, a, b), grad_rgb[a:b], near=nr,
                                             camera_idx=camera_idx, jitter=None if jitter is None else jitter[a:b],
                                             chunk_rays=chunk, loss_scale=loss_scale,
                                             proposal_requires_grad=flags[c0], nerf_layout=nerf_layout, kept=kp)
                    if nerf_layout:
                        go[a:b], gv[a:b] = g_o, g_v
                    else:
                        go[:, a:b], gv[:, a:b] = g_o, g_v
                    c0 = c1
                return go, gv
        R = self._rays(o3n, v3n, near, camera_idx, jitter, nerf_layout, kept)
        _check(grad_rgb, (n, 3), "grad_rgb")
        go = torch.empty_like(o3n)
        gv = torch.empty_like(o3n)
        ws = self._workspace(n)
        L.check(self.lib.nemit_backward(self._h, C.byref(R), _ptr(grad_rgb), int(chunk_rays or n), float(loss_scale),
                                        int(proposal_requires_grad), _ptr(go), _ptr(gv), _ptr(ws), ws.numel(),
                                        self._stream()), "nemit_backward")
        return go, gv

    def point_lights(self, o3n, v3n, near=0.0, camera_idx: int = 0, jitter=None, rgb_weights=(0.299, 0.587, 0.114),
                     nerf_layout: bool = False):
        """SdfNerfactoModel.get_point_lights (NS/models/sdf_nerfacto.py:189-212): {rgb (N,3), depth (N,1) of the sample
        with the largest weight * luminance, brightness_grad = d luminance / d origins (layout of ``o3n``)}."""
        n = o3n.shape[0] if nerf_layout else o3n.shape[1]
        R = self._rays(o3n, v3n, near, camera_idx, jitter, nerf_layout)
        rgb = torch.empty(n, 3, device=self.device, dtype=torch.float32)
        depth = torch.empty(n, 1, device=self.device, dtype=torch.float32)
        bg = torch.empty_like(o3n)
        ws = self._workspace(n)
        w = (C.c_float * 3)(*[float(x) for x in rgb_weights])
        L.check(self.lib.nemit_point_lights(self._h, C.byref(R), w, _ptr(rgb), _ptr(depth), _ptr(bg), _ptr(ws), ws.numel(),
                                            self._stream()), "nemit_point_lights")
        return {"rgb": rgb, "depth": depth, "brightness_grad": bg}

    def render_host(self, o_host: torch.Tensor, v_host: torch.Tensor, rgb_host: torch.Tensor, near: float = 0.0,
                    camera_idx: int = 0):
        """End-to-end call with pinned HOST buffers (nemit_forward_host): H2D, forward, D2H on the stream."""
        n = o_host.shape[1]
        for t, nm in [(o_host, "o_host"), (v_host, "v_host"), (rgb_host, "rgb_host")]:
            if t.is_cuda or t.dtype != torch.float32 or not t.is_contiguous():
                raise ValueError("%s must be a contiguous fp32 host tensor" % nm)
        R = L.NemitRays()
        R.o, R.v, R.n = o_host.data_ptr(), v_host.data_ptr(), n
        R.near_per_ray, R.near_scalar, R.camera_idx, R.training = None, float(near), int(camera_idx), int(self.training)
        R.jitter, R.anneal = None, float(self.anneal)
        ws = self._workspace(n)
        L.check(self.lib.nemit_forward_host(self._h, C.byref(R), C.c_void_p(rgb_host.data_ptr()), _ptr(ws), ws.numel(),
                                            self._stream()), "nemit_forward_host")
        return rgb_host

    def backward_host(self, o_host, v_host, grad_rgb_host, rgb_host, grad_o_host, grad_v_host, near: float = 0.0,
                      camera_idx: int = 0, primal_training: bool = False, chunk_rays: Optional[int] = None,
                      loss_scale: float = 1.0, proposal_requires_grad: bool = False):
        """End-to-end fwd+bwd with pinned HOST buffers in one plugin call (nemit_backward_host): H2D of rays and
        grad_out, the primal render (eval mode, or training mode with its edges kept for the adjoint), the
        training-mode backward, D2H of radiance and ray gradients - all queued on the current stream."""
        if not self.grads:
            self.bind_grads()
        n = o_host.shape[1]
        for t, nm in [(o_host, "o_host"), (v_host, "v_host"), (grad_rgb_host, "grad_rgb_host"), (rgb_host, "rgb_host"),
                      (grad_o_host, "grad_o_host"), (grad_v_host, "grad_v_host")]:
            if t.is_cuda or t.dtype != torch.float32 or not t.is_contiguous():
                raise ValueError("%s must be a contiguous fp32 host tensor" % nm)
        R = L.NemitRays()
        R.o, R.v, R.n = o_host.data_ptr(), v_host.data_ptr(), n
        R.near_per_ray, R.near_scalar, R.camera_idx, R.training = None, float(near), int(camera_idx), 1
        R.jitter, R.anneal = None, float(self.anneal)
        need = int(self.lib.nemit_backward_host_workspace_bytes(self._h, n))
        if self._ws is None or self._ws.numel() < need:
            self._ws = torch.empty(need, dtype=torch.uint8, device=self.device)
        ws = self._ws
        L.check(self.lib.nemit_backward_host(self._h, C.byref(R), C.c_void_p(grad_rgb_host.data_ptr()),
                                             int(bool(primal_training)), int(chunk_rays or n), float(loss_scale),
                                             int(bool(proposal_requires_grad)), C.c_void_p(rgb_host.data_ptr()),
                                             C.c_void_p(grad_o_host.data_ptr()), C.c_void_p(grad_v_host.data_ptr()),
                                             _ptr(ws), ws.numel(), self._stream()), "nemit_backward_host")
        return rgb_host, grad_o_host, grad_v_host

    def microbench(self):
        """{name: operations per second} of the ceilings the kernels are reported against (nemit_microbench)."""
        out = {}
        for i, name in enumerate(L.MICROBENCH_KINDS):
            res = (C.c_double * 2)()
            L.check(self.lib.nemit_microbench(self._h, i, res, self._stream()), "nemit_microbench")
            out[name] = float(res[0])
        return out

    def hash_encode(self, x: torch.Tensor, which: int = 0) -> torch.Tensor:
        """HashEncoding.forward (NS/field_components/encodings.py:341-382) on (n,3) points in [0,1]."""
        n = x.shape[0]
        _check(x, (n, 3), "x")
        g = self.spec.main_grid if which == 0 else self.spec.prop_grids[which - 1]
        out = torch.empty(n, g.num_levels * g.features_per_level, device=self.device, dtype=torch.float32)
        L.check(self.lib.nemit_hash_encode(self._h, which, _ptr(x), n, _ptr(out), self._stream()), "nemit_hash_encode")
        return out

    def density_fn(self, positions: torch.Tensor, which: int = 0, return_geo: bool = False):
        """Field.density_fn (NS/fields/base_field.py:49-69) on (n,3) NeRF-coordinate positions: which = 0 the main
        field (nerfacto_field.py:228-256), 1/2 the proposal fields (density_fields.py:95-122).  -> (n,1)."""
        shape = positions.shape[:-1]
        x = positions.reshape(-1, 3).contiguous()
        n = x.shape[0]
        _check(x, (n, 3), "positions")
        den = torch.empty(n, device=self.device, dtype=torch.float32)
        geo = torch.empty(n, self.spec.geo_feat_dim, device=self.device) if (return_geo and which == 0) else None
        L.check(self.lib.nemit_density_fn(self._h, int(which), _ptr(x), n, _ptr(den), _ptr(geo), self._stream()),
                "nemit_density_fn")
        den = den.view(*shape, 1)
        return (den, geo.view(*shape, -1)) if geo is not None else den

    def field_samples(self, positions: torch.Tensor, directions: torch.Tensor, camera_idx: int = 0):
        """NerfactoField.forward on explicit samples: positions, directions (...,3) -> density (...,1), rgb (...,3)."""
        shape = positions.shape[:-1]
        x = positions.reshape(-1, 3).contiguous()
        d = directions.reshape(-1, 3).contiguous()
        n = x.shape[0]
        _check(x, (n, 3), "positions")
        _check(d, (n, 3), "directions")
        den = torch.empty(n, device=self.device, dtype=torch.float32)
        rgb = torch.empty(n, 3, device=self.device, dtype=torch.float32)
        L.check(self.lib.nemit_field_samples(self._h, _ptr(x), _ptr(d), n, int(camera_idx), int(self.training),
                                             _ptr(den), _ptr(rgb), self._stream()), "nemit_field_samples")
        return den.view(*shape, 1), rgb.view(*shape, 3)

    def profile(self, on: bool = True):
        """Per-kernel CUDA-event timing inside the library (on the launching stream)."""
        L.check(self.lib.nemit_profile_enable(self._h, int(on)), "nemit_profile_enable")

    def profile_read(self, reset: bool = True, with_rays: bool = False):
        """{kernel: (total ms, launches)} accumulated since the last reset; with_rays: {kernel: (ms, launches, rays)}
        plus the key "red_lane_ops" (reduction lane-ops the tensor-core backward's table scatter issued)."""
        ms = (C.c_double * len(L.PROFILE_KINDS))()
        cnt = (C.c_int64 * len(L.PROFILE_KINDS))()
        rays = (C.c_int64 * len(L.PROFILE_KINDS))()
        red = C.c_int64(0)
        L.check(self.lib.nemit_profile_read(self._h, ms, cnt, rays, C.byref(red), int(reset)), "nemit_profile_read")
        if not with_rays:
            return {k: (ms[i], int(cnt[i])) for i, k in enumerate(L.PROFILE_KINDS) if cnt[i]}
        out = {k: (ms[i], int(cnt[i]), int(rays[i])) for i, k in enumerate(L.PROFILE_KINDS) if cnt[i]}
        out["red_lane_ops"] = int(red.value)
        return out

    def launch_count(self) -> int:
        return int(self.lib.nemit_launch_count(self._h))
