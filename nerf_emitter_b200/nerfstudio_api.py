"""Model-side API of the emitter path, kept intact: the nerfstudio classes the path is built from,
with the reference's names, constructor arguments and ``state_dict`` keys, whose ``forward`` dispatches to the C ABI
(``include/nemit.h``).  Nothing here computes radiance or densities in torch: torch holds the parameters
(``nn.Parameter`` so that the reference's Adam param groups and checkpoints work unchanged) and the output buffers.

    reference class (file:line)                                     entry it dispatches to
    HashEncoding          NS/field_components/encodings.py:250-382   nemit_hash_encode_grid
    SHEncoding            NS/field_components/encodings.py:683-720   nemit_sh_encode
    MLP                   NS/field_components/mlp.py:62-195          nemit_mlp
    HashMLPDensityField   NS/fields/density_fields.py:31-125         nemit_density_fn (which = 1, 2)
    NerfactoField         NS/fields/nerfacto_field.py:60-334         nemit_field_samples, nemit_density_fn (which = 0)
    ProposalNetworkSampler NS/model_components/ray_samplers.py:531-626  nemit_sample_levels
    SdfNerfactoModel      NS/models/sdf_nerfacto.py:41-212, NS/models/base_model.py:290-344,
                          NS/pipelines/mitsuba_sdf.py:1144-1219      nemit_forward / nemit_jvp / nemit_backward

The component modules are inference drop-ins (forward values); reverse mode goes through the fused
``SdfNerfactoModel.backward_for_camera_ray_bundle`` exactly as on the reference's emitter path, which leaves the
gradients in ``param.grad`` / ``ray_bundle.origins.grad`` / ``ray_bundle.directions.grad``.
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass
from typing import Callable, Dict, List, Optional, Sequence, Tuple

import torch
from torch import nn

from . import lib as L
from .emitter import NerfEmitter, _ptr
from .params import EmitterModelSpec, GridSpec, all_param_names


def _stream(device):
    return C.c_void_p(torch.cuda.current_stream(device).cuda_stream)


def _cuda_f32(t: torch.Tensor, what: str) -> torch.Tensor:
    if not t.is_cuda:
        raise L.NemitError("%s: CUDA tensor required; there is no CPU path" % what)
    return t.detach().to(torch.float32).contiguous()


# -------------------------------------------------------------------------------------------------------------
# ray containers (NS/cameras/rays.py:33-105, 107-150, 153-260): plain dataclasses with the fields the path reads
# -------------------------------------------------------------------------------------------------------------
@dataclass
class Frustums:
    origins: torch.Tensor       # (*bs, 3)
    directions: torch.Tensor    # (*bs, 3)
    starts: torch.Tensor        # (*bs, 1)
    ends: torch.Tensor          # (*bs, 1)
    pixel_area: Optional[torch.Tensor] = None

    @property
    def shape(self):
        return self.starts.shape[:-1]

    def get_positions(self) -> torch.Tensor:
        """rays.py:49-58: o + d * (starts + ends) / 2 (index arithmetic only)."""
        return self.origins + self.directions * (self.starts + self.ends) / 2


@dataclass
class RaySamples:
    frustums: Frustums
    camera_indices: Optional[torch.Tensor] = None
    deltas: Optional[torch.Tensor] = None
    spacing_starts: Optional[torch.Tensor] = None
    spacing_ends: Optional[torch.Tensor] = None
    spacing_to_euclidean_fn: Optional[Callable] = None


@dataclass
class RayBundle:
    origins: torch.Tensor                   # (N,3) NeRF coordinates
    directions: torch.Tensor                # (N,3)
    pixel_area: Optional[torch.Tensor] = None
    camera_indices: Optional[torch.Tensor] = None
    nears: Optional[torch.Tensor] = None
    fars: Optional[torch.Tensor] = None

    def __len__(self):
        return self.origins.shape[0]

    def camera_index(self) -> int:
        if self.camera_indices is None or self.camera_indices.numel() == 0:
            return 0
        return int(self.camera_indices.reshape(-1)[0])


# -------------------------------------------------------------------------------------------------------------
# field components
# -------------------------------------------------------------------------------------------------------------
class HashEncoding(nn.Module):
    """encodings.py:250-382 (torch implementation: every level hashed, no dense coarse levels)."""

    def __init__(self, num_levels: int = 16, min_res: int = 16, max_res: int = 1024, log2_hashmap_size: int = 19,
                 features_per_level: int = 2, hash_init_scale: float = 0.001, implementation: str = "torch",
                 interpolation: Optional[str] = None) -> None:
        super().__init__()
        if features_per_level != 2 or num_levels > L.MAX_LEVELS:
            raise ValueError("this build supports 2 features per level and up to 16 levels")
        if interpolation not in (None, "Linear"):
            raise ValueError("only linear interpolation (encodings.py:296-298)")
        self.in_dim = 3
        self.num_levels, self.min_res, self.max_res = num_levels, min_res, max_res
        self.features_per_level, self.log2_hashmap_size = features_per_level, log2_hashmap_size
        self.hash_table_size = 2 ** log2_hashmap_size
        self.grid = GridSpec(num_levels, min_res, max_res, log2_hashmap_size, features_per_level)
        table = torch.rand(size=(self.hash_table_size * num_levels, features_per_level)) * 2 - 1
        self.hash_table = nn.Parameter(table * hash_init_scale)                    # encodings.py:309-312
        self._g = L.NemitGrid()
        self._g.num_levels, self._g.log2_table_size = num_levels, log2_hashmap_size
        for i, s in enumerate(self.grid.scalings()):
            self._g.scalings[i] = s

    def get_out_dim(self) -> int:
        return self.num_levels * self.features_per_level

    def forward(self, in_tensor: torch.Tensor) -> torch.Tensor:
        x = _cuda_f32(in_tensor.reshape(-1, 3), "HashEncoding")
        out = torch.empty(x.shape[0], self.get_out_dim(), device=x.device, dtype=torch.float32)
        L.check(L.load().nemit_hash_encode_grid(C.byref(self._g), _ptr(_cuda_f32(self.hash_table, "hash_table")), _ptr(x),
                                                x.shape[0], _ptr(out), _stream(x.device)), "nemit_hash_encode_grid")
        return out.view(*in_tensor.shape[:-1], -1)


class SHEncoding(nn.Module):
    """encodings.py:683-720, torch path (components_from_spherical_harmonics, utils/math.py:27-90)."""

    def __init__(self, levels: int = 4, implementation: str = "torch") -> None:
        super().__init__()
        if levels <= 0 or levels > 4:
            raise ValueError(f"Spherical harmonic encoding only supports 1 to 4 levels, requested {levels}")
        self.in_dim, self.levels = 3, levels

    def get_out_dim(self) -> int:
        return self.levels ** 2

    def forward(self, in_tensor: torch.Tensor) -> torch.Tensor:
        d = _cuda_f32(in_tensor.reshape(-1, 3), "SHEncoding")
        out = torch.empty(d.shape[0], self.get_out_dim(), device=d.device, dtype=torch.float32)
        L.check(L.load().nemit_sh_encode(_ptr(d), d.shape[0], self.levels, _ptr(out), _stream(d.device)), "nemit_sh_encode")
        return out.view(*in_tensor.shape[:-1], -1)


class MLP(nn.Module):
    """mlp.py:62-195, torch path: ``layers`` is a ModuleList of nn.Linear (state_dict keys layers.i.weight / .bias)."""

    def __init__(self, in_dim: int, num_layers: int, layer_width: int, out_dim: Optional[int] = None,
                 skip_connections: Optional[Tuple[int]] = None, activation: Optional[nn.Module] = nn.ReLU(),
                 out_activation: Optional[nn.Module] = None, implementation: str = "torch") -> None:
        super().__init__()
        if skip_connections:
            raise ValueError("skip connections are not on the emitter path")
        for a in (activation, out_activation):
            if a is not None and not isinstance(a, nn.ReLU):
                raise ValueError("activation must be ReLU or None")
        self.in_dim, self.num_layers, self.layer_width = in_dim, num_layers, layer_width
        self.out_dim = out_dim if out_dim is not None else layer_width
        self.activation, self.out_activation = activation, out_activation
        dims = [in_dim] + [layer_width] * (num_layers - 1) + [self.out_dim]
        if max(dims) > L.MLP_MAX_WIDTH or num_layers > L.MLP_MAX_LAYERS:
            raise ValueError("this build supports widths <= 64 and <= 8 layers")
        self.layers = nn.ModuleList([nn.Linear(dims[i], dims[i + 1]) for i in range(num_layers)])   # mlp.py:146-161

    def get_out_dim(self) -> int:
        return self.out_dim

    def forward(self, in_tensor: torch.Tensor) -> torch.Tensor:
        x = _cuda_f32(in_tensor.reshape(-1, self.in_dim), "MLP")
        M = L.NemitMlp()
        M.in_dim, M.num_layers = self.in_dim, self.num_layers
        M.activation, M.out_activation = int(self.activation is not None), int(self.out_activation is not None)
        keep = []
        for i, lin in enumerate(self.layers):
            w, b = _cuda_f32(lin.weight, "MLP weight"), _cuda_f32(lin.bias, "MLP bias")
            keep += [w, b]
            M.out_dim[i], M.weight[i], M.bias[i] = lin.out_features, w.data_ptr(), b.data_ptr()
        out = torch.empty(x.shape[0], self.out_dim, device=x.device, dtype=torch.float32)
        L.check(L.load().nemit_mlp(C.byref(M), _ptr(x), x.shape[0], _ptr(out), _stream(x.device)), "nemit_mlp")
        return out.view(*in_tensor.shape[:-1], -1)


class Embedding(nn.Module):
    """NS/field_components/embedding.py:24-52 (a parameter holder on this path: the field kernels read the row or
    the mean row themselves, nerfacto_field.py:272-285)."""

    def __init__(self, in_dim: int, out_dim: int) -> None:
        super().__init__()
        self.in_dim, self.out_dim = in_dim, out_dim
        self.embedding = nn.Embedding(in_dim, out_dim)

    def mean(self, dim=0):
        return self.embedding.weight.mean(dim)


class _Bound:
    """A field that evaluates through an emitter handle owned by the model it belongs to."""
    _owner = None

    def _emitter(self) -> NerfEmitter:
        if self._owner is None:
            raise L.NemitError("%s evaluates through its model's emitter handle: construct it inside "
                               "SdfNerfactoModel (model.field / model.proposal_networks[i]) as the reference does "
                               "(NS/models/nerfacto.py:169-260)" % type(self).__name__)
        return self._owner()._emitter()


class HashMLPDensityField(nn.Module, _Bound):
    """density_fields.py:31-125 with the proposal-network settings of nerfacto.py:97-102 (hidden 16, 5 levels, 2^17)."""

    def __init__(self, aabb: torch.Tensor, num_layers: int = 2, hidden_dim: int = 16, spatial_distortion=None,
                 use_linear: bool = False, num_levels: int = 5, max_res: int = 128, base_res: int = 16,
                 log2_hashmap_size: int = 17, features_per_level: int = 2, implementation: str = "torch") -> None:
        super().__init__()
        if use_linear or num_layers != 2:
            raise ValueError("this build supports the 2-layer MLP density field")
        self.register_buffer("aabb", aabb)
        self.register_buffer("max_res", torch.tensor(max_res))
        self.register_buffer("num_levels", torch.tensor(num_levels))
        self.register_buffer("log2_hashmap_size", torch.tensor(log2_hashmap_size))
        self.encoding = HashEncoding(num_levels=num_levels, min_res=base_res, max_res=max_res,
                                     log2_hashmap_size=log2_hashmap_size, features_per_level=features_per_level)
        network = MLP(in_dim=self.encoding.get_out_dim(), num_layers=num_layers, layer_width=hidden_dim, out_dim=1,
                      activation=nn.ReLU(), out_activation=None)
        self.mlp_base = nn.Sequential(self.encoding, network)                      # density_fields.py:89
        self._which = 1

    def density_fn(self, positions: torch.Tensor, times=None) -> torch.Tensor:
        """base_field.py:49-69 -> get_density (density_fields.py:95-122): (...,3) -> (...,1)."""
        return self._emitter().density_fn(_cuda_f32(positions, "positions"), which=self._which)

    def get_density(self, ray_samples: RaySamples):
        return self.density_fn(ray_samples.frustums.get_positions()), None

    def forward(self, ray_samples: RaySamples, compute_normals: bool = False) -> Dict[str, torch.Tensor]:
        return {"density": self.get_density(ray_samples)[0]}


class NerfactoField(nn.Module, _Bound):
    """nerfacto_field.py:60-334 with the sdf-nerfacto settings (RawNeRF colour activation, rgb_bias -5)."""

    def __init__(self, aabb: torch.Tensor, num_images: int, num_layers: int = 2, hidden_dim: int = 64,
                 geo_feat_dim: int = 15, num_levels: int = 16, base_res: int = 16, max_res: int = 2048,
                 log2_hashmap_size: int = 19, num_layers_color: int = 3, features_per_level: int = 2,
                 hidden_dim_color: int = 64, appearance_embedding_dim: int = 32, spatial_distortion=None,
                 implementation: str = "torch", rgb_bias: float = -5.0) -> None:
        super().__init__()
        self.register_buffer("aabb", aabb)
        self.register_buffer("max_res", torch.tensor(max_res))
        self.register_buffer("num_levels", torch.tensor(num_levels))
        self.register_buffer("log2_hashmap_size", torch.tensor(log2_hashmap_size))
        self.geo_feat_dim, self.num_images, self.rgb_bias = geo_feat_dim, num_images, rgb_bias
        self.appearance_embedding_dim = appearance_embedding_dim
        self.embedding_appearance = Embedding(num_images, appearance_embedding_dim)
        self.direction_encoding = SHEncoding(levels=4)
        self.mlp_base_grid = HashEncoding(num_levels=num_levels, min_res=base_res, max_res=max_res,
                                          log2_hashmap_size=log2_hashmap_size, features_per_level=features_per_level)
        self.mlp_base_mlp = MLP(in_dim=self.mlp_base_grid.get_out_dim(), num_layers=num_layers, layer_width=hidden_dim,
                                out_dim=1 + geo_feat_dim, activation=nn.ReLU(), out_activation=None)
        self.mlp_base = nn.Sequential(self.mlp_base_grid, self.mlp_base_mlp)       # nerfacto_field.py:168
        self.mlp_head = MLP(in_dim=self.direction_encoding.get_out_dim() + geo_feat_dim + appearance_embedding_dim,
                            num_layers=num_layers_color, layer_width=hidden_dim_color, out_dim=3, activation=nn.ReLU(),
                            out_activation=None)

    def density_fn(self, positions: torch.Tensor, times=None) -> torch.Tensor:
        return self._emitter().density_fn(_cuda_f32(positions, "positions"), which=0)

    def get_density(self, ray_samples: RaySamples):
        """nerfacto_field.py:228-256: (density (...,1), geometry features (...,15))."""
        return self._emitter().density_fn(_cuda_f32(ray_samples.frustums.get_positions(), "positions"), which=0,
                                          return_geo=True)

    def forward(self, ray_samples: RaySamples, compute_normals: bool = False) -> Dict[str, torch.Tensor]:
        """base_field.py:124-143: {DENSITY, RGB} of the samples (FieldHeadNames values 'density' / 'rgb')."""
        if compute_normals:
            raise ValueError("predict_normals is off on the emitter path (SURVEY.md 8 header)")
        E = self._emitter()
        f = ray_samples.frustums
        den, rgb = E.field_samples(_cuda_f32(f.get_positions(), "positions"),
                                   _cuda_f32(f.directions.expand(*f.shape, 3), "directions"),
                                   camera_idx=0 if ray_samples.camera_indices is None
                                   else int(ray_samples.camera_indices.reshape(-1)[0]))
        return {"density": den, "rgb": rgb}


class ProposalNetworkSampler:
    """ray_samplers.py:531-626: the whole proposal loop (uniform-in-disparity initial sampler, two density
    evaluations, two PDF resampling stages) is ONE kernel launch (nemit_sample_levels)."""

    def __init__(self, owner, num_nerf_samples_per_ray: int = 48, num_proposal_samples_per_ray=(256, 96),
                 num_proposal_network_iterations: int = 2, single_jitter: bool = True, update_sched=None,
                 initial_sampler=None) -> None:
        if tuple(num_proposal_samples_per_ray) != (256, 96) or num_nerf_samples_per_ray != 48 or \
                num_proposal_network_iterations != 2:
            raise ValueError("this build supports 256/96 proposal and 48 nerf samples per ray")
        self._owner = owner
        self.num_nerf_samples_per_ray = num_nerf_samples_per_ray
        self.num_proposal_samples_per_ray = tuple(num_proposal_samples_per_ray)
        self.num_proposal_network_iterations = num_proposal_network_iterations
        self.update_sched = update_sched if update_sched is not None else (lambda step: 1)
        self._anneal, self._steps_since_update, self._step = 1.0, 0, 0

    def _consume_update(self, num_calls: int = 1) -> List[bool]:
        """The ``updated`` state machine of ray_samplers.py:598,622-623 for ``num_calls`` consecutive sampler calls
        (one per chunk of the caller's chunk loop): a call is "updated" (proposal densities carry autograd) when
        ``_steps_since_update > update_sched(_step)`` or ``_step < 10``, and every updated call - the no-grad primal
        ones included - resets the counter."""
        flags = []
        for _ in range(int(num_calls)):
            updated = self._steps_since_update > self.update_sched(self._step) or self._step < 10
            if updated:
                self._steps_since_update = 0
            flags.append(bool(updated))
        return flags

    def set_anneal(self, anneal: float) -> None:
        self._anneal = anneal

    def step_cb(self, step: int) -> None:
        self._step = step
        self._steps_since_update += 1

    def __call__(self, ray_bundle: RayBundle, density_fns: Optional[List[Callable]] = None, jitter=None):
        return self.generate_ray_samples(ray_bundle, density_fns, jitter)

    def generate_ray_samples(self, ray_bundle: RayBundle, density_fns=None, jitter=None):
        """-> (ray_samples, weights_list, ray_samples_list) as ray_samplers.py:584-626.  ``density_fns`` is accepted
        for signature compatibility; the densities come from the model's own proposal networks inside the kernel."""
        model = self._owner()
        E = model._emitter()
        E.anneal = float(self._anneal)
        self._consume_update()
        o, d = _cuda_f32(ray_bundle.origins, "origins"), _cuda_f32(ray_bundle.directions, "directions")
        aux = E.sample_levels(o, d, camera_idx=ray_bundle.camera_index(), jitter=jitter, nerf_layout=True)
        near = E.spec.near_plane if E.training else 0.0
        s_near, s_far = _spacing(torch.tensor(near)), _spacing(torch.tensor(E.spec.far_plane))

        def to_euclid(s):                                                          # ray_samplers.py:121-122
            return _spacing_inv(s * s_far + (1 - s) * s_near)

        def samples(s_edges, t_edges=None):
            t_edges = to_euclid(s_edges) if t_edges is None else t_edges
            fr = Frustums(origins=o[:, None, :], directions=d[:, None, :], starts=t_edges[:, :-1, None],
                          ends=t_edges[:, 1:, None], pixel_area=ray_bundle.pixel_area)
            return RaySamples(frustums=fr, camera_indices=ray_bundle.camera_indices,
                              deltas=(t_edges[:, 1:, None] - t_edges[:, :-1, None]),
                              spacing_starts=s_edges[:, :-1, None], spacing_ends=s_edges[:, 1:, None],
                              spacing_to_euclidean_fn=to_euclid)

        n = o.shape[0]
        s0 = E._tables[0].to(o.device)[None, :].expand(n, -1)
        if jitter is not None and E.training:
            # the kernel's level-0 edges under single jitter (ray_samplers.py:107-117), rebuilt for the returned RaySamples
            j0 = (torch.cat([j.reshape(n, 1) for j in jitter], dim=1) if isinstance(jitter, (list, tuple)) else jitter)
            t_rand = j0.to(o.device, torch.float32).reshape(n, -1)[:, :1]
            centers = (s0[..., 1:] + s0[..., :-1]) / 2.0
            upper = torch.cat([centers, s0[..., -1:]], -1)
            lower = torch.cat([s0[..., :1], centers], -1)
            s0 = lower + (upper - lower) * t_rand
        ray_samples_list = [samples(s0), samples(aux["s_edges1"])]
        weights_list = [aux["weights0"][..., None], aux["weights1"][..., None]]
        return samples(aux["s_edges2"], aux["t_edges2"]), weights_list, ray_samples_list


def _spacing(x):         # ray_samplers.py:250
    return torch.where(x < 1, x / 2, 1 - 1 / (2 * x))


def _spacing_inv(x):     # ray_samplers.py:251
    return torch.where(x < 0.5, 2 * x, 1 / (2 - 2 * x))


# -------------------------------------------------------------------------------------------------------------
# the model
# -------------------------------------------------------------------------------------------------------------
def _num_chunks(num_rays: int, chunk: int) -> int:
    return (int(num_rays) + int(chunk) - 1) // int(chunk)


class SdfNerfactoModel(nn.Module):
    """NerfactoModel.populate_modules (NS/models/nerfacto.py:169-299) with the sdf-nerfacto overrides
    (NS/configs/method_configs.py:314-326) + the three emitter entry points of SdfNerfactoModel / Model /
    MitsubaSdfPipeline.  ``state_dict`` keys equal the reference's (field.*, proposal_networks.*)."""

    def __init__(self, num_train_data: int = 4, scene_scale: float = 1.0, spec: Optional[EmitterModelSpec] = None,
                 disable_inside_aabb: bool = True, precision: str = "fp32", sampler: str = "exact",
                 interlevel_loss_mult: float = 1.0, distortion_loss_mult: float = 7e-3,
                 backward_num_rays_per_chunk: int = 16384, eval_num_rays_per_chunk: int = 16384,
                 forward_ad_num_rays_per_chunk: int = 4096, proposal_update_every: int = 1,
                 proposal_warmup: int = 5000, rotation_ids: Optional[Sequence[int]] = None,
                 use_average_appearance_embedding: bool = True) -> None:
        super().__init__()
        import weakref
        # With a Rotater (turntable captures) the reference sizes the appearance embedding by the number of rotations
        # and sets always_use_camera_indices (nerfacto.py:186-206); ``rotation_ids`` is ``Rotater.rotation_ids``
        # (camera -> rotation id, rotater.py:36-37), applied to the final samples by sdf_nerfacto.py:71-72.
        self.rotation_ids = None if rotation_ids is None else [int(x) for x in rotation_ids]
        self.use_average_appearance_embedding = use_average_appearance_embedding
        if self.rotation_ids is not None:
            num_train_data = len(set(self.rotation_ids))
        spec = spec or EmitterModelSpec(num_images=num_train_data)
        self.spec, self.scene_scale = spec, float(scene_scale)
        self.precision, self.sampler_mode = precision, sampler
        self.disable_inside_aabb = disable_inside_aabb
        self.interlevel_loss_mult, self.distortion_loss_mult = interlevel_loss_mult, distortion_loss_mult
        # chunk sizes of the three chunk loops (base_model.py:53-58 as overridden by sdf-nerfacto,
        # method_configs.py:315-316): they only matter for the sampler's per-call state machine and, in the backward,
        # for the per-chunk regulariser means
        self.backward_num_rays_per_chunk = backward_num_rays_per_chunk
        self.eval_num_rays_per_chunk = eval_num_rays_per_chunk
        self.forward_ad_num_rays_per_chunk = forward_ad_num_rays_per_chunk
        aabb = torch.tensor([[-scene_scale] * 3, [scene_scale] * 3], dtype=torch.float32)
        mg = spec.main_grid
        self.field = NerfactoField(aabb, num_images=spec.num_images, hidden_dim=spec.hidden_dim,
                                   geo_feat_dim=spec.geo_feat_dim, num_levels=mg.num_levels, base_res=mg.base_res,
                                   max_res=mg.max_res, log2_hashmap_size=mg.log2_hashmap_size,
                                   appearance_embedding_dim=spec.appearance_dim, rgb_bias=spec.rgb_bias)
        self.proposal_networks = nn.ModuleList()
        for i, g in enumerate(spec.prop_grids):
            net = HashMLPDensityField(aabb, hidden_dim=spec.prop_hidden, num_levels=g.num_levels, max_res=g.max_res,
                                      base_res=g.base_res, log2_hashmap_size=g.log2_hashmap_size)
            net._which = i + 1
            self.proposal_networks.append(net)
        ref = weakref.ref(self)
        self.field._owner = ref
        for net in self.proposal_networks:
            net._owner = ref
        self.density_fns = [net.density_fn for net in self.proposal_networks]      # nerfacto.py:229

        def update_schedule(step):                      # nerfacto.py:237-242
            x = min(max(step / proposal_warmup, 0.0), 1.0) * proposal_update_every if proposal_warmup > 0 \
                else float(proposal_update_every)
            return min(max(x, 1.0), float(proposal_update_every))

        self.proposal_sampler = ProposalNetworkSampler(ref, spec.num_nerf_samples, spec.num_proposal_samples,
                                                       update_sched=update_schedule)
        self._E: Optional[NerfEmitter] = None
        self._bound_ptrs: Tuple[int, ...] = ()
        self._num_processed_rays = 0                    # base_model.py:122 (the trainer's rays/s meter)

    def clear_num_processed_rays(self) -> None:
        """base_model.py:124-125."""
        self._num_processed_rays = 0

    def num_processed_rays(self) -> int:
        """base_model.py:127-128."""
        return self._num_processed_rays

    # -- handle management: the library borrows the module's own parameter / .grad storage -------------------------
    def _named(self) -> Dict[str, torch.Tensor]:
        # named_parameters() reports a shared tensor once, under its first registration, which is the reference's
        # canonical name (field.mlp_base_grid.*, proposal_networks.i.encoding.* / .mlp_base.1.*)
        sd = dict(self.named_parameters())
        return {k: sd[k] for k in all_param_names()}

    def _emitter(self) -> NerfEmitter:
        P = self._named()
        ptrs = tuple(p.data_ptr() for p in P.values())
        dev = next(iter(P.values())).device
        if dev.type != "cuda":
            raise L.NemitError("SdfNerfactoModel: move the model to a CUDA device; there is no CPU path")
        if self._E is None or self._E.device != dev:
            self._E = NerfEmitter({k: p.data for k, p in P.items()}, self.spec, device=dev, scene_scale=self.scene_scale,
                                  disable_inside_aabb=self.disable_inside_aabb,
                                  interlevel_mult=self.interlevel_loss_mult, distortion_mult=self.distortion_loss_mult)
            self._E.set_precision(self.precision).set_sampler(self.sampler_mode)
            self._E.set_appearance_mode(self.rotation_ids is not None, self.use_average_appearance_embedding,
                                        self.rotation_ids)
            self._bound_ptrs = ptrs
        elif ptrs != self._bound_ptrs:
            self._E.bind_params({k: p.data for k, p in P.items()})
            self._bound_ptrs = ptrs
        self._E.train(self.training)
        self._E.anneal = float(self.proposal_sampler._anneal)
        return self._E

    def get_param_groups(self) -> Dict[str, List[nn.Parameter]]:
        """nerfacto.py:313-317."""
        return {"proposal_networks": list(self.proposal_networks.parameters()), "fields": list(self.field.parameters())}

    # -- emitter entry points ---------------------------------------------------------------------------------------
    @torch.no_grad()
    def get_rgb_for_camera_ray_bundle(self, camera_ray_bundle: RayBundle, jitter=None) -> torch.Tensor:
        """base_model.py:290-313 -> SdfNerfactoModel.get_rgb (sdf_nerfacto.py:64-83): (N,3) radiance."""
        E = self._emitter()
        self._num_processed_rays += len(camera_ray_bundle)                       # base_model.py:302
        self.proposal_sampler._consume_update(_num_chunks(len(camera_ray_bundle), self.eval_num_rays_per_chunk))
        return E.render(_cuda_f32(camera_ray_bundle.origins, "origins"), _cuda_f32(camera_ray_bundle.directions, "directions"),
                        camera_idx=camera_ray_bundle.camera_index(), jitter=jitter, nerf_layout=True)

    get_rgb = get_rgb_for_camera_ray_bundle

    @torch.no_grad()
    def forward_grad_for_camera_ray_bundle(self, camera_ray_bundle: RayBundle, grad_o: torch.Tensor,
                                           grad_v: torch.Tensor, jitter=None) -> torch.Tensor:
        """base_model.py:315-344 -> forward_grad (sdf_nerfacto.py:85-113): tangent of the radiance, (N,3)."""
        E = self._emitter()
        self._num_processed_rays += len(camera_ray_bundle)                       # base_model.py:331
        self.proposal_sampler._consume_update(_num_chunks(len(camera_ray_bundle), self.forward_ad_num_rays_per_chunk))
        return E.jvp(_cuda_f32(camera_ray_bundle.origins, "origins"), _cuda_f32(camera_ray_bundle.directions, "directions"),
                     _cuda_f32(grad_o.reshape(-1, 3), "grad_o"), _cuda_f32(grad_v.reshape(-1, 3), "grad_v"),
                     camera_idx=camera_ray_bundle.camera_index(), jitter=jitter, nerf_layout=True)

    forward_grad = forward_grad_for_camera_ray_bundle

    @torch.no_grad()
    def get_point_lights(self, ray_bundle: RayBundle, jitter=None) -> Dict[str, torch.Tensor]:
        """sdf_nerfacto.py:189-212: {"rgb", "depth" (contrib), "brightness_grad"} for the light point cloud
        (NS/model_components/output_light_pc.py:74-173)."""
        E = self._emitter()
        self.proposal_sampler._consume_update()          # called per chunk by its driver (base_model.py:260-288)
        return E.point_lights(_cuda_f32(ray_bundle.origins, "origins"), _cuda_f32(ray_bundle.directions, "directions"),
                              camera_idx=ray_bundle.camera_index(), jitter=jitter, nerf_layout=True)

    def backward_for_camera_ray_bundle(self, camera_ray_bundle: RayBundle, grad_out: torch.Tensor, jitter=None,
                                       loss_scale: float = 1.0) -> None:
        """MitsubaSdfPipeline.backward_for_camera_ray_bundle (mitsuba_sdf.py:1144-1219): accumulates into
        ``param.grad`` and leaves ``origins.grad`` / ``directions.grad`` on the bundle's tensors."""
        E = self._emitter()
        P = self._named()
        for p in P.values():
            if p.grad is None:
                p.grad = torch.zeros_like(p.data)
        gptrs = tuple(p.grad.data_ptr() for p in P.values())
        if getattr(self, "_bound_gptrs", None) != gptrs:
            E.bind_grads({k: p.grad for k, p in P.items()})
            self._bound_gptrs = gptrs
        o = _cuda_f32(camera_ray_bundle.origins, "origins")
        d = _cuda_f32(camera_ray_bundle.directions, "directions")
        g = _cuda_f32(grad_out.reshape(-1, 3), "grad_out")
        n, chunk = o.shape[0], self.backward_num_rays_per_chunk
        self._num_processed_rays += n                                           # mitsuba_sdf.py:1157
        if isinstance(jitter, (list, tuple)):
            jitter = torch.cat([j.reshape(n, 1) for j in jitter], dim=1)
        # one sampler call per chunk: the flag can differ between the first chunk and the rest (quirk 13)
        flags = self.proposal_sampler._consume_update(_num_chunks(n, chunk))
        go, gv = E.backward(o, d, g, camera_idx=camera_ray_bundle.camera_index(), jitter=jitter, chunk_rays=chunk,
                            loss_scale=loss_scale, proposal_requires_grad=flags, nerf_layout=True)
        camera_ray_bundle.origins.grad = go
        camera_ray_bundle.directions.grad = gv
