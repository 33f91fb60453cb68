"""TEST INFRASTRUCTURE ONLY (oracle). The reference's own torch modules wired as the emitter path.

Runs only where /root/reference exists (build container). Used to (a) validate the standalone
restatement ``oracle/nerfacto_oracle.py`` and (b) generate the committed golden vectors under
``tests/golden/`` (``oracle/make_golden.py``).  Nothing in the product path imports this.

Only the glue that lives in un-importable reference files is restated here (they import
mitsuba / torchmetrics); all arithmetic is executed by the reference's own classes:
  * module construction        <- nerfstudio/models/nerfacto.py:169-299 (populate_modules)
  * get_rgb / forward_grad / get_backward_outputs <- nerfstudio/models/sdf_nerfacto.py:64-168
  * backward objective         <- nerfstudio/pipelines/mitsuba_sdf.py:1170-1177
  * get_ray_bundle             <- differentiable-sdf-rendering/python/emitters/nerf_op.py:76-89
  * near change of variables   <- .../emitters/nerf_emitter_op.py:41-44, 66-69, 112
"""
from __future__ import annotations

from typing import Dict, Optional

import numpy as np
import torch
from torch.autograd import forward_ad

from oracle.ref_import import import_reference
from nerf_emitter_b200.params import EmitterModelSpec, prop_param_names

MI2GL_LEFT = torch.tensor([[0., 0., 1., 0.], [1., 0., 0., 0.], [0., 1., 0., 0.], [0., 0., 0., 1.]])


class ReferenceEmitter:
    def __init__(self, params: Dict[str, torch.Tensor], spec: Optional[EmitterModelSpec] = None,
                 scene_scale: float = 1.0, disable_inside_aabb: bool = True,
                 interlevel_mult: float = 1.0, distortion_mult: float = 7e-3,
                 always_use_camera_indices: bool = False, rotation_ids=None, rotater=None):
        # rotater: None or dict(transforms (R,4,4), kind 'sphere' | 'box', center, radius, aabb): the buffers of a
        # nerfstudio.field_components.rotater.Rotater.  That module imports mitsuba at module scope and cannot be imported
        # here, so its apply_frustums / get_rotation_mask (rotater.py:125-149) are RESTATED in _apply_frustums below and
        # hooked into the reference's own RayBundle.get_ray_samples (cameras/rays.py:285-286) exactly as get_ray_bundle
        # does (differentiable-sdf-rendering/python/emitters/nerf_op.py:86-87).
        self.rotater = rotater
        # always_use_camera_indices / rotation_ids: the model as NerfactoModel.populate_modules builds it when the
        # datamanager carries a Rotater (nerfacto.py:186-206).  The Rotater class itself imports mitsuba at module
        # scope (rotater.py:7), so its three-line map_rotation_ids (:94-100) is repeated in _map_rotation_ids below.
        self.rotation_ids = None if rotation_ids is None else torch.tensor([int(x) for x in rotation_ids])
        self.ns = ns = import_reference()
        self.spec = spec = spec or EmitterModelSpec()
        self.scene_scale = scene_scale
        self.interlevel_mult = interlevel_mult
        self.distortion_mult = distortion_mult
        aabb = torch.tensor([[-1., -1., -1.], [1., 1., 1.]]) * scene_scale
        contraction = ns.spatial_distortions.SceneContraction(order=float("inf"))
        mg = spec.main_grid
        self.field = ns.nerfacto_field.NerfactoField(
            aabb, hidden_dim=spec.hidden_dim, num_levels=mg.num_levels, max_res=mg.max_res,
            base_res=mg.base_res, features_per_level=mg.features_per_level,
            log2_hashmap_size=mg.log2_hashmap_size, hidden_dim_color=spec.hidden_dim,
            spatial_distortion=contraction, num_images=spec.num_images, use_pred_normals=False,
            use_average_appearance_embedding=True, color_output_activation="RawNeRF",
            rgb_bias=spec.rgb_bias, appearance_embedding_dim=spec.appearance_dim,
            implementation="torch", always_use_camera_indices=always_use_camera_indices)
        self.props = torch.nn.ModuleList()
        for g in spec.prop_grids:
            self.props.append(ns.density_fields.HashMLPDensityField(
                aabb, spatial_distortion=contraction, hidden_dim=spec.prop_hidden,
                log2_hashmap_size=g.log2_hashmap_size, num_levels=g.num_levels, max_res=g.max_res,
                base_res=g.base_res, features_per_level=g.features_per_level, implementation="torch"))
        self.sampler = ns.ray_samplers.ProposalNetworkSampler(
            num_nerf_samples_per_ray=spec.num_nerf_samples,
            num_proposal_samples_per_ray=spec.num_proposal_samples,
            num_proposal_network_iterations=2, single_jitter=True,
            update_sched=lambda step: 1, initial_sampler=None, train_stratified=True)
        self.collider = ns.scene_colliders.NearFarCollider(near_plane=spec.near_plane, far_plane=spec.far_plane)
        self.renderer = ns.renderers.RGBRenderer(background_color="last_sample", hdr=True)
        self.modules = [self.field, self.props, self.sampler, self.collider, self.renderer]
        self.field.disable_inside_aabb = disable_inside_aabb
        for p in self.props:
            p.disable_inside_aabb = disable_inside_aabb
        self.load(params)
        self.train(False)

    # -- parameters ------------------------------------------------------------------------
    def named_parameters(self) -> Dict[str, torch.nn.Parameter]:
        out = {}
        for n, p in self.field.named_parameters():
            out["field." + n] = p
        for i, pn in enumerate(self.props):
            names = prop_param_names(i)
            out[names[0]] = pn.encoding.hash_table
            for n, p in pn.mlp_base[1].named_parameters():
                out["proposal_networks.%d.mlp_base.1.%s" % (i, n)] = p
        # mlp_base.0.* of the field aliases mlp_base_grid
        return {k: v for k, v in out.items() if ".mlp_base.0." not in k and "field.mlp_base.1." not in k}

    def load(self, params: Dict[str, torch.Tensor]):
        own = self.named_parameters()
        missing = set(params) - set(own)
        assert not missing, "names not in the reference modules: %s" % sorted(missing)
        with torch.no_grad():
            for k, v in params.items():
                assert own[k].shape == v.shape, (k, own[k].shape, v.shape)
                own[k].copy_(v)

    def train(self, mode: bool):
        for m in self.modules:
            m.train(mode)
        self.training = mode

    # -- glue ------------------------------------------------------------------------------
    def get_ray_bundle(self, o, v, camera_idx=0):
        o = (o * 2 - 1) * self.scene_scale                       # mi_gl_conversion.py:40-42
        o = o @ MI2GL_LEFT[:3, :3].t() + MI2GL_LEFT[:3, 3]       # emitters/util.py:31-35
        v = v @ MI2GL_LEFT[:3, :3].t()
        rb = self.ns.rays.RayBundle(origins=o, directions=v, pixel_area=torch.ones_like(o[..., :1]))
        rb.set_camera_indices(camera_idx)
        if self.rotater is not None:
            rb.rotater = self._apply_frustums
        return rb

    def _apply_frustums(self, frustums, camera_indices):
        """RESTATEMENT of Rotater.apply_frustums + get_rotation_mask (rotater.py:125-149) with batch_affine_left
        (nerfstudio/utils/mi_gl_conversion.py:69-73)."""
        R = self.rotater
        positions = frustums.get_positions()
        if R["kind"] == "sphere":
            rotate_mask = torch.linalg.norm(positions - R["center"], dim=-1, keepdim=True) < R["radius"]
        else:
            rotate_mask = ((R["aabb"][0] <= positions).all(dim=-1, keepdim=True) &
                           (positions <= R["aabb"][1]).all(dim=-1, keepdim=True))
        ids = camera_indices[..., 0]
        if self.rotation_ids is not None:
            ids = self._map_rotation_ids(ids)
        transforms = torch.as_tensor(R["transforms"], dtype=torch.float32)[ids]

        def batch_affine_left(M, x, is_pos=True):
            res = (M[..., :3, :3] @ x.unsqueeze(-1))[..., 0]
            if is_pos:
                res = res + M[..., :3, 3]
            return res
        frustums.origins = torch.where(
            rotate_mask, batch_affine_left(transforms, frustums.origins[..., :1, :]).expand(frustums.origins.shape),
            frustums.origins)
        frustums.directions = torch.where(
            rotate_mask,
            batch_affine_left(transforms, frustums.directions[..., :1, :], is_pos=False).expand(frustums.directions.shape),
            frustums.directions)

    def _density_fns(self):
        return [p.density_fn for p in self.props]

    def _map_rotation_ids(self, camera_indices):
        return self.rotation_ids[camera_indices.clamp(max=len(self.rotation_ids) - 1)]

    def _field_and_render(self, ray_samples, grad_scaling=True):
        F = self.ns.FieldHeadNames
        if self.rotation_ids is not None:                        # sdf_nerfacto.py:71-72
            ray_samples.camera_indices = self._map_rotation_ids(ray_samples.camera_indices)
        out = self.field.forward(ray_samples, compute_normals=False)
        if grad_scaling:
            # losses.py:558-577. Identity in the primal; `_GradientScaler` defines no jvp, and on
            # torch 2.11 forward-mode AD through it raises NotImplementedError (the reference
            # targets torch 2.0.1), so the JVP leg skips it: its tangent map is the identity.
            out = self.ns.losses.scale_gradients_by_distance_squared(out, ray_samples)
        weights = ray_samples.get_weights(out[F.DENSITY])
        rgb = self.renderer(rgb=out[F.RGB], weights=weights)
        return rgb, weights, out

    # -- the three modes (torch (3,N) in, like the op) -------------------------------------
    @torch.no_grad()
    def render(self, o3n, v3n, near=0.0, camera_idx=0, jitter=None, return_aux=False):
        o = o3n.t().contiguous()
        v = v3n.t().contiguous()
        o = o + near * v
        rb = self.collider(self.get_ray_bundle(o, v, camera_idx))
        aux = {}
        with _PatchedRand(self, jitter), _CaptureSearchsorted(aux):
            rs, wl, rsl = self.sampler(rb, density_fns=self._density_fns())
        rgb, weights, out = self._field_and_render(rs)
        if not return_aux:
            return rgb
        F = self.ns.FieldHeadNames
        sd = self.ns.losses.ray_samples_to_sdist
        aux.update(dict(
            rgb=rgb, weights0=wl[0][..., 0], weights1=wl[1][..., 0], weights2=weights[..., 0],
            sdist0=sd(rsl[0]), sdist1=sd(rsl[1]), sdist2=sd(rs),
            density2=out[F.DENSITY][..., 0], rgb_samples=out[F.RGB],
            accumulation=torch.sum(weights, dim=-2)))
        return rgb, aux

    @torch.no_grad()
    def jvp(self, o3n, v3n, do3n, dv3n, near=0.0, camera_idx=0, jitter=None):
        o = o3n.t().contiguous()
        v = v3n.t().contiguous()
        o = o + near * v
        go = do3n.t().contiguous() + near * dv3n.t().contiguous()
        gv = dv3n.t().contiguous()
        with forward_ad.dual_level():
            rb = self.get_ray_bundle(forward_ad.make_dual(o, go), forward_ad.make_dual(v, gv), camera_idx)
            rb.origins, go = forward_ad.unpack_dual(rb.origins)
            rb.directions, gv = forward_ad.unpack_dual(rb.directions)
        rb = self.collider(rb)
        with _PatchedRand(self, jitter):
            rs, _, _ = self.sampler(rb, density_fns=self._density_fns())
        with forward_ad.dual_level():
            rs.frustums.origins = forward_ad.make_dual(
                rs.frustums.origins, go.unsqueeze(-2).expand_as(rs.frustums.origins))
            rs.frustums.directions = forward_ad.make_dual(
                rs.frustums.directions, gv.unsqueeze(-2).expand_as(rs.frustums.directions))
            rgb, _, _ = self._field_and_render(rs, grad_scaling=False)
            t = forward_ad.unpack_dual(rgb).tangent
        t = t.clone()
        t[torch.isnan(t)] = 0.
        return t

    def vjp(self, o3n, v3n, grad_out, near=0.0, camera_idx=0, jitter=None, proposal_requires_grad=True,
            loss_scale=1.0):
        """Returns (grad_o (3,N), grad_v (3,N), {param name: grad}) for one chunk of rays."""
        for p in self.named_parameters().values():
            p.grad = None
        oo = (o3n.t().contiguous() + near * v3n.t().contiguous()).detach().requires_grad_(True)
        vv = v3n.t().contiguous().detach().requires_grad_(True)
        rb = self.collider(self.get_ray_bundle(oo, vv, camera_idx))
        origins, directions = rb.origins, rb.directions
        rb.origins, rb.directions = origins.detach(), directions.detach()
        # `updated` state machine (ray_samplers.py:598): force the requested behaviour
        self.sampler._step = 0 if proposal_requires_grad else 100
        self.sampler._steps_since_update = 0
        with _PatchedRand(self, jitter):
            rs, wl, rsl = self.sampler(rb, density_fns=self._density_fns())
        rs = rs[...]
        rs.frustums.origins = origins[..., None, :].expand_as(rs.frustums.origins)
        rs.frustums.directions = directions[..., None, :].expand_as(rs.frustums.directions)
        rgb, weights, _ = self._field_and_render(rs)
        wl.append(weights)
        rsl.append(rs)
        loss = torch.sum(rgb * grad_out)
        loss = loss + self.interlevel_mult * self.ns.losses.interlevel_loss(wl, rsl)
        loss = loss + self.distortion_mult * self.ns.losses.distortion_loss(wl, rsl)
        (loss * loss_scale).backward()
        go = oo.grad / loss_scale
        gv = vv.grad / loss_scale
        go = go - near * gv                                      # nerf_emitter_op.py:112
        grads = {k: (p.grad.detach().clone() if p.grad is not None else torch.zeros_like(p))
                 for k, p in self.named_parameters().items()}
        return go.t().contiguous(), gv.t().contiguous(), grads, rgb.detach()


def reference_point_lights(R: "ReferenceEmitter", o3n, v3n, near=0.0, camera_idx=0, rgb_weights=(0.299, 0.587, 0.114)):
    """SdfNerfactoModel.get_point_lights (nerfstudio/models/sdf_nerfacto.py:189-212) with the reference's own field,
    sampler, renderers (RGBRenderer, DepthRenderer("contrib")) and autograd; only the glue of get_point_lights /
    calc_brightness_grad (nerfstudio/models/nerfacto.py:547-573) is restated (un-importable file)."""
    ns = R.ns
    wts = torch.tensor(rgb_weights, dtype=torch.float32)
    o = o3n.t().contiguous() + near * v3n.t().contiguous()
    rb = R.collider(R.get_ray_bundle(o, v3n.t().contiguous(), camera_idx))
    with torch.no_grad():
        rs, _, _ = R.sampler(rb, density_fns=R._density_fns())
    with torch.enable_grad():
        rb.origins.requires_grad = True
        rs.frustums.origins = rb.origins.unsqueeze(-2).expand_as(rs.frustums.origins)
        rs.frustums.directions = rb.directions.unsqueeze(-2).expand_as(rs.frustums.directions)
        if rb.rotater is not None:                                # nerfacto.py:553-554
            rb.rotater(rs.frustums, rb.camera_indices[..., None])
        rgb, weights, out = R._field_and_render(rs)
        brightness = torch.sum(rgb * wts, dim=-1, keepdim=True)
        (grad,) = torch.autograd.grad(brightness, rb.origins, grad_outputs=torch.ones_like(brightness))
    lum = torch.sum(out[ns.FieldHeadNames.RGB] * wts, dim=-1, keepdim=True)
    depth = ns.renderers.DepthRenderer(method="contrib")(weights=weights.detach(), ray_samples=rs, values=lum.detach())
    return {"rgb": rgb.detach(), "depth": depth.detach(), "brightness_grad": grad.detach()}


class _PatchedRand:
    """Feeds explicit per-ray jitter (3 tensors (N,1): level 0, pdf 1, pdf 2) to the samplers in
    training mode instead of their private generators (ray_samplers.py:107-117, 323-333)."""

    def __init__(self, emitter, jitter):
        self.jitter = list(jitter) if jitter is not None else None
        self.emitter = emitter

    def __enter__(self):
        if self.jitter is None:
            return self
        self._orig = torch.rand
        q = self.jitter

        def fake_rand(*a, **k):
            return q.pop(0).clone()
        torch.rand = fake_rand
        s = self.emitter.sampler
        for m in [s.initial_sampler] + list(s.pdf_samplers):
            m.get_generator = lambda device: None
        return self

    def __exit__(self, *exc):
        if self.jitter is not None:
            torch.rand = self._orig
        return False


class _CaptureSearchsorted:
    """Records the indices returned by the PDF samplers' searchsorted calls."""

    def __init__(self, aux):
        self.aux = aux

    def __enter__(self):
        self._orig = torch.searchsorted
        k = [0]

        def wrapped(*a, **kw):
            r = self._orig(*a, **kw)
            k[0] += 1
            self.aux["inds%d" % k[0]] = r.clone()
            self.aux["cdf%d" % k[0]] = a[0].clone()
            return r
        torch.searchsorted = wrapped
        return self

    def __exit__(self, *exc):
        torch.searchsorted = self._orig
        return False
