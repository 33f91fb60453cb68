"""TEST INFRASTRUCTURE ONLY (oracle): CPU restatement of the reference's emitter ray path.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s cpu_baseline / ``--impl reference``
legs may import this file; the product path (``nerf_emitter_b200``) never does.

It is a self-contained, functional torch-CPU restatement (no reference import, so it travels to
the GPU box) of the algorithm the reference executes with ``implementation="torch"``.  Parity is
PINNED: ``tests/test_oracle_vs_reference.py`` runs it against the reference's own classes
(``oracle/ref_harness.py``) in the build container, and ``tests/golden/*.npz`` (made by
``oracle/make_golden.py`` from the reference itself) pin it everywhere else.

Arithmetic is float32 with torch-CPU semantics on purpose: the reference's PDF resampling depends on
ATen's CPU ``cumsum`` (float64 accumulate, rounded per prefix) and ``searchsorted(right=True)``.

Each function cites the reference lines it follows (paths relative to /root/reference).
"""
from __future__ import annotations

import math
from typing import Dict, List, Optional, Sequence

import torch
import torch.nn.functional as F

from nerf_emitter_b200.params import EmitterModelSpec, GridSpec, prop_param_names

PRIME_Y = 2654435761
PRIME_Z = 805459861
EPS_LOSS = 1.0e-7  # nerfstudio/model_components/losses.py:42 (EPS)


# --------------------------------------------------------------------------------------------
# A0  ray transform: Mitsuba unit cube -> NeRF coordinates
# --------------------------------------------------------------------------------------------
def rays_mi_to_nerf(o_n3: torch.Tensor, v_n3: torch.Tensor, scene_scale: float):
    """differentiable-sdf-rendering/python/emitters/nerf_op.py:76-89 with
    nerfstudio/utils/mi_gl_conversion.py:4-9,40-42 and emitters/util.py:31-35.
    The left matrix is the axis permutation (x,y,z) -> (z,x,y); applied as a matmul so that
    forward/reverse AD follow the same path as the reference."""
    perm = torch.tensor([[0., 0., 1.], [1., 0., 0.], [0., 1., 0.]], dtype=o_n3.dtype)
    o = ((o_n3 * 2 - 1) * scene_scale) @ perm.t()
    o = o + torch.zeros(3, dtype=o_n3.dtype)
    v = v_n3 @ perm.t()
    return o, v


# --------------------------------------------------------------------------------------------
# A2  piecewise spacing and the initial 256-bin edges
# --------------------------------------------------------------------------------------------
def spacing_fwd(x):
    """nerfstudio/model_components/ray_samplers.py:250"""
    return torch.where(x < 1, x / 2, 1 - 1 / (2 * x))


def spacing_inv(s):
    """nerfstudio/model_components/ray_samplers.py:251"""
    return torch.where(s < 0.5, 2 * s, 1 / (2 - 2 * s))


def s_to_euclid(s, s_near, s_far):
    """ray_samplers.py:121-122 (closure spacing_to_euclidean_fn)."""
    return spacing_inv(s * s_far + (1 - s) * s_near)


def initial_s_edges(num_rays: int, num_samples: int, jitter: Optional[torch.Tensor]):
    """ray_samplers.py:104-117. jitter: (num_rays,1) in [0,1) for single-jitter stratified
    training sampling, or None for the deterministic edges."""
    e = torch.linspace(0.0, 1.0, num_samples + 1)[None, :]
    if jitter is not None:
        mid = (e[..., 1:] + e[..., :-1]) / 2.0
        hi = torch.cat([mid, e[..., -1:]], -1)
        lo = torch.cat([e[..., :1], mid], -1)
        e = lo + (hi - lo) * jitter
    return e


# --------------------------------------------------------------------------------------------
# A4  contraction, normalisation, selector
# --------------------------------------------------------------------------------------------
def contract_linf(x):
    """nerfstudio/field_components/spatial_distortions.py:67-70 with ord=inf."""
    mag = torch.linalg.norm(x, ord=float("inf"), dim=-1)[..., None]
    return torch.where(mag < 1, x, (2 - (1 / mag)) * (x / mag))


def normalised_positions(x, aabb, disable_inside_aabb: bool):
    """nerfstudio/fields/density_fields.py:96-106 == nerfacto_field.py:230-242;
    aabb test nerfstudio/data/scene_box.py:36-37."""
    p = (contract_linf(x) + 2.0) / 4.0
    sel = ((p > 0.0) & (p < 1.0)).all(dim=-1)
    if disable_inside_aabb:
        inside = torch.all(x > aabb[0], dim=-1) & torch.all(x < aabb[1], dim=-1)
        sel = sel & ~inside
    return p * sel[..., None], sel


# --------------------------------------------------------------------------------------------
# A5  multiresolution hash encoding (torch path: every level hashed)
# --------------------------------------------------------------------------------------------
def hash_index(ix, iy, iz, table_size: int):
    """nerfstudio/field_components/encodings.py:322-339: int64 products, xor, mod T."""
    h = ix.to(torch.int64) ^ (iy.to(torch.int64) * PRIME_Y) ^ (iz.to(torch.int64) * PRIME_Z)
    return h % table_size


def hash_encode(p: torch.Tensor, table: torch.Tensor, grid: GridSpec) -> torch.Tensor:
    """encodings.py:341-382. p (...,3) in [0,1]; table (L*T, F). Output (..., L*F) ordered
    [l0f0, l0f1, l1f0, ...]. Corner pairing: the ceil corner is weighted by the fractional
    offset and the floor corner by (1 - offset), blended along x, then y, then z."""
    L, T = grid.num_levels, grid.table_size
    scal = torch.tensor(grid.scalings(), dtype=p.dtype)
    xs = p[..., None, :] * scal.view(-1, 1)                       # (..., L, 3)
    hi = torch.ceil(xs).to(torch.int32)
    lo = torch.floor(xs).to(torch.int32)
    frac = xs - lo
    lvl_off = torch.arange(L, dtype=torch.int64) * T

    def corner(cx, cy, cz):
        return table[hash_index(cx[..., 0], cy[..., 1], cz[..., 2], T) + lvl_off]   # (..., L, F)

    fx, fy, fz = frac[..., 0:1], frac[..., 1:2], frac[..., 2:3]
    # z = ceil plane
    a = corner(hi, hi, hi) * fx + corner(lo, hi, hi) * (1 - fx)    # y=hi
    b = corner(hi, lo, hi) * fx + corner(lo, lo, hi) * (1 - fx)    # y=lo
    top = a * fy + b * (1 - fy)
    # z = floor plane
    c = corner(hi, hi, lo) * fx + corner(lo, hi, lo) * (1 - fx)    # y=hi
    d = corner(hi, lo, lo) * fx + corner(lo, lo, lo) * (1 - fx)    # y=lo
    bot = c * fy + d * (1 - fy)
    out = top * fz + bot * (1 - fz)
    return out.flatten(-2)


# --------------------------------------------------------------------------------------------
# activations
# --------------------------------------------------------------------------------------------
def trunc_exp(x):
    """nerfstudio/field_components/activations.py:28-48: value exp(x); both the reverse- and the
    forward-mode derivative are exp(clamp(x,-15,15)).  Written with detach so that one expression
    serves autograd and forward AD."""
    slope = torch.exp(x.detach().clamp(-15, 15))
    return torch.exp(x).detach() + slope * (x - x.detach())


def safe_exp(x, bias: float):
    """nerfstudio/fields/nerfacto_field.py:40-57 (SafeExp, SAFE_EXP_MAX=88)."""
    return torch.exp(torch.clamp(x + bias, max=88.0))


class _ScaleGrad(torch.autograd.Function):
    """Identity whose reverse-mode gradient is multiplied by `s`
    (nerfstudio/model_components/losses.py:542-577)."""

    @staticmethod
    def forward(ctx, v, s):
        ctx.save_for_backward(s)
        return v.view_as(v)

    @staticmethod
    def backward(ctx, g):
        (s,) = ctx.saved_tensors
        return g * s, None


# --------------------------------------------------------------------------------------------
# A6  proposal density field
# --------------------------------------------------------------------------------------------
def proposal_density(P: Dict[str, torch.Tensor], i: int, grid: GridSpec, pos, aabb, disable_inside_aabb):
    """nerfstudio/fields/density_fields.py:95-122 via Field.density_fn (base_field.py:49-69);
    MLP with biases (field_components/mlp.py:163-183)."""
    n = prop_param_names(i)
    p, sel = normalised_positions(pos, aabb, disable_inside_aabb)
    h = hash_encode(p.view(-1, 3), P[n[0]], grid)
    h = F.relu(F.linear(h, P[n[1]], P[n[2]]))
    raw = F.linear(h, P[n[3]], P[n[4]]).view(*pos.shape[:-1], 1)
    return trunc_exp(raw) * sel[..., None]


# --------------------------------------------------------------------------------------------
# A7  weights from density
# --------------------------------------------------------------------------------------------
def weights_from_density(sigma, deltas):
    """nerfstudio/cameras/rays.py:128-150. sigma, deltas (..., S, 1)."""
    ds = deltas * sigma
    alpha = 1 - torch.exp(-ds)
    acc = torch.cumsum(ds[..., :-1, :], dim=-2)
    acc = torch.cat([torch.zeros((*acc.shape[:1], 1, 1)), acc], dim=-2)
    return torch.nan_to_num(alpha * torch.exp(-acc))


# --------------------------------------------------------------------------------------------
# A8  PDF resampling
# --------------------------------------------------------------------------------------------
def pdf_resample(w, s_edges, num_new: int, jitter: Optional[torch.Tensor], pad: float = 0.01, eps: float = 1e-5):
    """nerfstudio/model_components/ray_samplers.py:282-380 with include_original=False
    (ray_samplers.py:567-569).  w (N,S) >= 0, s_edges (N or 1, S+1).  Returns the num_new+1 new
    spacing edges, the searchsorted indices and the cdf (for parity tests)."""
    nb = num_new + 1
    w = w + pad
    tot = torch.sum(w, dim=-1, keepdim=True)
    fill = torch.relu(eps - tot)
    w = w + fill / w.shape[-1]
    tot = tot + fill
    pdf = w / tot
    cdf = torch.min(torch.ones_like(pdf), torch.cumsum(pdf, dim=-1))
    cdf = torch.cat([torch.zeros_like(cdf[..., :1]), cdf], dim=-1)
    u = torch.linspace(0.0, 1.0 - (1.0 / nb), steps=nb)
    if jitter is not None:
        u = u.expand(*cdf.shape[:-1], nb) + jitter / nb
    else:
        u = (u + 1.0 / (2 * nb)).expand(*cdf.shape[:-1], nb)
    u = u.contiguous()
    edges = s_edges.expand(cdf.shape[0], -1)
    idx = torch.searchsorted(cdf, u, side="right")
    last = edges.shape[-1] - 1
    i0 = torch.clamp(idx - 1, 0, last)
    i1 = torch.clamp(idx, 0, last)
    c0, c1 = torch.gather(cdf, -1, i0), torch.gather(cdf, -1, i1)
    e0, e1 = torch.gather(edges, -1, i0), torch.gather(edges, -1, i1)
    t = torch.clip(torch.nan_to_num((u - c0) / (c1 - c0), 0), 0, 1)
    return (e0 + t * (e1 - e0)).detach(), idx, cdf


# --------------------------------------------------------------------------------------------
# A3  Rotater.apply_frustums (turntable captures)
# --------------------------------------------------------------------------------------------
def apply_frustums(rot, o, d, t_edges):
    """nerfstudio/field_components/rotater.py:125-149 (get_rotation_mask + apply_frustums) with batch_affine_left
    (nerfstudio/utils/mi_gl_conversion.py:69-73), as hooked into RayBundle.get_ray_samples (cameras/rays.py:285-286).
    RESTATED: rotater.py imports mitsuba at module scope and cannot be imported here; these are its 15 lines of torch.
    rot: None or dict(T (4,4) of the call's rotation id, kind 'sphere' | 'box', center (3,), radius, aabb (2,3)).
    o, d (N,3); t_edges (N,S+1).  Returns per-sample origins and directions (N,S,3)."""
    t0, t1 = t_edges[..., :-1, None], t_edges[..., 1:, None]
    S = t0.shape[1]
    o_s, d_s = o[:, None, :].expand(-1, S, -1), d[:, None, :].expand(-1, S, -1)
    if rot is None:
        return o_s, d_s
    pos = o_s + d_s * (t0 + t1) / 2                                # Frustums.get_positions, cameras/rays.py:49-58
    if rot["kind"] == "sphere":
        mask = torch.linalg.norm(pos - rot["center"], dim=-1, keepdim=True) < rot["radius"]
    else:
        mask = ((rot["aabb"][0] <= pos).all(dim=-1, keepdim=True) & (pos <= rot["aabb"][1]).all(dim=-1, keepdim=True))
    T = rot["T"]
    o2 = (T[:3, :3] @ o.unsqueeze(-1))[..., 0] + T[:3, 3]
    d2 = (T[:3, :3] @ d.unsqueeze(-1))[..., 0]
    return torch.where(mask, o2[:, None, :].expand(-1, S, -1), o_s), torch.where(mask, d2[:, None, :].expand(-1, S, -1), d_s)


# --------------------------------------------------------------------------------------------
# A9  the three-level proposal loop
# --------------------------------------------------------------------------------------------
def sample_levels(P, spec: EmitterModelSpec, o, d, near: float, far: float, aabb, *,
                  disable_inside_aabb=True, jitter: Optional[Sequence[torch.Tensor]] = None,
                  anneal: float = 1.0, proposal_requires_grad: bool = False,
                  edges_override: Optional[Sequence[Optional[torch.Tensor]]] = None, rot=None):
    """nerfstudio/model_components/ray_samplers.py:584-626.  o, d (N,3) NeRF coords (detached by
    the caller for the backward variant, sdf_nerfacto.py:121-126).  Returns per-level spacing edges,
    euclidean edges, weights and the searchsorted indices.
    edges_override (tests only): spacing edges [None, s_edges1, s_edges2] to use instead of this function's
    own PDF-stage outputs, so that a checker can evaluate the field at bit-identical sample positions
    (the path is ill-conditioned w.r.t. 1-ulp position changes, see DESIGN.md "Conditioning")."""
    N = o.shape[0]
    nears = torch.ones_like(o[..., 0:1]) * near
    fars = torch.ones_like(o[..., 0:1]) * far
    s_near, s_far = spacing_fwd(nears), spacing_fwd(fars)
    out = {"s_edges": [], "t_edges": [], "weights": [], "inds": [], "cdf": []}
    counts = list(spec.num_proposal_samples) + [spec.num_nerf_samples]
    w = None
    s_e = None
    for lvl, S in enumerate(counts):
        jit = None if jitter is None else jitter[lvl]
        if lvl == 0:
            s_e = initial_s_edges(N, S, jit)
        else:
            s_e, idx, cdf = pdf_resample(torch.pow(w, anneal), s_e, S, jit)
            out["inds"].append(idx)
            out["cdf"].append(cdf)
            if edges_override is not None and edges_override[lvl] is not None:
                s_e = edges_override[lvl].detach()
        t_e = s_to_euclid(s_e, s_near, s_far)                       # (N, S+1)
        out["s_edges"].append(s_e)
        out["t_edges"].append(t_e)
        if lvl < 2:
            t0, t1 = t_e[..., :-1, None], t_e[..., 1:, None]
            o_s, d_s = apply_frustums(rot, o, d, t_e)                 # every level's get_ray_samples runs the hook
            pos = o_s + d_s * (t0 + t1) / 2                           # cameras/rays.py:49-58
            with torch.set_grad_enabled(proposal_requires_grad and torch.is_grad_enabled()):
                sig = proposal_density(P, lvl, spec.prop_grids[lvl], pos, aabb, disable_inside_aabb)
                wfull = weights_from_density(sig, t1 - t0)
            out["weights"].append(wfull)
            w = wfull[..., 0]
    return out


# --------------------------------------------------------------------------------------------
# A11  spherical harmonics, 4 levels = 16 components
# --------------------------------------------------------------------------------------------
def sh16(v):
    """nerfstudio/utils/math.py:27-90 for levels=4, evaluated at whatever it is given (the torch
    path feeds (d+1)/2, nerfstudio/fields/base_field.py:146-152)."""
    x, y, z = v[..., 0], v[..., 1], v[..., 2]
    xx, yy, zz = x ** 2, y ** 2, z ** 2
    k1 = 0.4886025119029199
    k2 = 1.0925484305920792
    comps = [
        torch.full_like(x, 0.28209479177387814),
        k1 * y, k1 * z, k1 * x,
        k2 * x * y, k2 * y * z, 0.9461746957575601 * zz - 0.31539156525251999, k2 * x * z,
        0.5462742152960396 * (xx - yy),
        0.5900435899266435 * y * (3 * xx - yy), 2.890611442640554 * x * y * z,
        0.4570457994644658 * y * (5 * zz - 1), 0.3731763325901154 * z * (5 * zz - 3),
        0.4570457994644658 * x * (5 * zz - 1), 1.445305721320277 * z * (xx - yy),
        0.5900435899266435 * x * (xx - 3 * yy),
    ]
    return torch.stack(comps, dim=-1)


# --------------------------------------------------------------------------------------------
# A10 + A12 + A13 + A7 + A14  main field on the 48 final samples and compositing
# --------------------------------------------------------------------------------------------
def main_stage(P, spec: EmitterModelSpec, o, d, t_edges, aabb, *, training: bool, camera_idx: int,
               disable_inside_aabb=True, sh_carries_reverse_grad=False, grad_scaling=True,
               always_use_camera_indices=False, o_s=None, d_s=None):
    """nerfstudio/fields/nerfacto_field.py:228-334 (get_density, get_outputs), Field.forward
    (base_field.py:124-143), scale_gradients_by_distance_squared (losses.py:558-577),
    get_weights (cameras/rays.py:128-150), RGBRenderer (model_components/renderers.py:71-121,203-234).
    o, d (N,3); t_edges (N,49).  Returns rgb (N,3), weights (N,48,1), per-sample density and rgb.
    o_s, d_s (N,48,3): per-sample frustum origins / directions when a Rotater has rewritten them (apply_frustums)."""
    N = o.shape[0]
    t0, t1 = t_edges[..., :-1, None], t_edges[..., 1:, None]
    S = t0.shape[1]
    if o_s is None:
        o_s, d_s = o[:, None, :].expand(N, S, 3), d[:, None, :].expand(N, S, 3)
    pos = o_s + d_s * (t0 + t1) / 2
    p, sel = normalised_positions(pos, aabb, disable_inside_aabb)
    enc = hash_encode(p.view(-1, 3), P["field.mlp_base_grid.hash_table"], spec.main_grid)
    h = F.relu(F.linear(enc, P["field.mlp_base_mlp.layers.0.weight"], P["field.mlp_base_mlp.layers.0.bias"]))
    h = F.linear(h, P["field.mlp_base_mlp.layers.1.weight"], P["field.mlp_base_mlp.layers.1.bias"])
    h = h.view(N, S, -1)
    raw_sigma, geo = torch.split(h, [1, spec.geo_feat_dim], dim=-1)
    sigma = trunc_exp(raw_sigma) * sel[..., None]

    dn = (d_s + 1.0) / 2.0
    # encodings.py:708 decorates the torch SH with no_grad: no reverse-mode gradient reaches the
    # directions through SH, while forward-mode tangents do flow (no_grad does not stop them).
    sh = sh16(dn if sh_carries_reverse_grad else _no_reverse(dn)).reshape(-1, 16)
    emb = P["field.embedding_appearance.embedding.weight"]
    if training or always_use_camera_indices:      # nerfacto_field.py:275
        app = emb[camera_idx][None, :].expand(N * S, -1)
    else:
        app = torch.ones(N * S, spec.appearance_dim) * emb.mean(dim=0)
    x = torch.cat([sh, geo.reshape(-1, spec.geo_feat_dim), app], dim=-1)
    x = F.relu(F.linear(x, P["field.mlp_head.layers.0.weight"], P["field.mlp_head.layers.0.bias"]))
    x = F.relu(F.linear(x, P["field.mlp_head.layers.1.weight"], P["field.mlp_head.layers.1.bias"]))
    x = F.linear(x, P["field.mlp_head.layers.2.weight"], P["field.mlp_head.layers.2.bias"])
    rgb_s = safe_exp(x, spec.rgb_bias).view(N, S, 3)

    if grad_scaling:
        scale = torch.square((t0 + t1) / 2).clamp(0, 1)
        rgb_s = _ScaleGrad.apply(rgb_s, scale)
        sigma = _ScaleGrad.apply(sigma, scale)
    w = weights_from_density(sigma, t1 - t0)
    rgb_in = rgb_s if training else torch.nan_to_num(rgb_s)
    comp = torch.sum(w * rgb_in, dim=-2)
    acc = torch.sum(w, dim=-2)
    comp = comp + rgb_in[..., -1, :] * (1.0 - acc)                  # background "last_sample"
    if not training:
        comp = torch.clamp(comp, min=0.0)                          # hdr: no upper clamp
    return comp, w, sigma, rgb_s


def _no_reverse(x):
    """Value and forward-mode tangent of x, but cut from the reverse-mode graph."""
    if torch.is_grad_enabled() and x.requires_grad:
        return x.detach()
    return x


# --------------------------------------------------------------------------------------------
# A18  regularisers of the backward objective
# --------------------------------------------------------------------------------------------
def _outer_bound(t0_lo, t0_hi, t1_lo, t1_hi, y1):
    """nerfstudio/model_components/losses.py:52-81."""
    cy = torch.cat([torch.zeros_like(y1[..., :1]), torch.cumsum(y1, dim=-1)], dim=-1)
    lo = torch.searchsorted(t1_lo.contiguous(), t0_lo.contiguous(), side="right") - 1
    lo = torch.clamp(lo, min=0, max=y1.shape[-1] - 1)
    hi = torch.searchsorted(t1_hi.contiguous(), t0_hi.contiguous(), side="right")
    hi = torch.clamp(hi, min=0, max=y1.shape[-1] - 1)
    return torch.take_along_dim(cy[..., 1:], hi, dim=-1) - torch.take_along_dim(cy[..., :-1], lo, dim=-1)


def interlevel_loss(weights: List[torch.Tensor], s_edges: List[torch.Tensor]):
    """losses.py:84-131: sum over proposal levels of mean(clip(w - w_outer,0)^2/(w+eps))."""
    c = s_edges[-1].detach()
    w = weights[-1][..., 0].detach()
    total = 0.0
    for wp, cp in zip(weights[:-1], s_edges[:-1]):
        cp = cp.expand(c.shape[0], -1)
        wo = _outer_bound(c[..., :-1], c[..., 1:], cp[..., :-1], cp[..., 1:], wp[..., 0])
        total = total + torch.mean(torch.clip(w - wo, min=0) ** 2 / (w + EPS_LOSS))
    return total


def distortion_loss(weights: List[torch.Tensor], s_edges: List[torch.Tensor]):
    """losses.py:134-153 on the final level."""
    t = s_edges[-1]
    w = weights[-1][..., 0]
    mid = (t[..., 1:] + t[..., :-1]) / 2
    gap = torch.abs(mid[..., :, None] - mid[..., None, :])
    inter = torch.sum(w * torch.sum(w[..., None, :] * gap, dim=-1), dim=-1)
    intra = torch.sum(w ** 2 * (t[..., 1:] - t[..., :-1]), dim=-1) / 3
    return torch.mean(inter + intra)


# --------------------------------------------------------------------------------------------
# A15-A19  the three modes of the op, (3,N) in / (N,3) out like NeRFEmitterOp
# --------------------------------------------------------------------------------------------
class OracleEmitter:
    """Functional mirror of NeRFEmitterOp.ns_render_ray / ns_forward_ray / ns_backward_ray
    (differentiable-sdf-rendering/python/emitters/nerf_emitter_op.py:38-117) over
    SdfNerfactoModel.get_rgb / forward_grad / get_backward_outputs
    (nerfstudio/models/sdf_nerfacto.py:64-168)."""

    def __init__(self, params: Dict[str, torch.Tensor], spec: Optional[EmitterModelSpec] = None,
                 scene_scale: float = 1.0, disable_inside_aabb: bool = True,
                 interlevel_mult: float = 1.0, distortion_mult: float = 7e-3,
                 always_use_camera_indices: bool = False, rotation_ids=None, rotater=None):
        # rotater: None or dict(transforms (R,4,4), kind 'sphere' | 'box', center, radius, aabb) = the Rotater's buffers
        self.rotater = rotater
        # a model built with a Rotater has always_use_camera_indices=True and one embedding row per rotation
        # (nerfstudio/models/nerfacto.py:186-206); rotation_ids = Rotater.rotation_ids
        self.always_use_camera_indices = always_use_camera_indices
        self.rotation_ids = None if rotation_ids is None else [int(x) for x in rotation_ids]
        self.P = {k: v.detach().clone().float().cpu() for k, v in params.items()}
        self.spec = spec or EmitterModelSpec()
        self.scene_scale = scene_scale
        self.aabb = torch.tensor([[-1., -1., -1.], [1., 1., 1.]]) * scene_scale
        self.disable_inside_aabb = disable_inside_aabb
        self.interlevel_mult = interlevel_mult
        self.distortion_mult = distortion_mult
        self.training = False
        self.anneal = 1.0

    def train(self, mode: bool = True):
        self.training = mode

    def _cam(self, camera_idx):
        """Rotater.map_rotation_ids (nerfstudio/field_components/rotater.py:94-100) as applied to the final samples'
        camera_indices by sdf_nerfacto.py:71-72,92-93,127-128,198-199."""
        if self.rotation_ids is None:
            return camera_idx
        return self.rotation_ids[min(camera_idx, len(self.rotation_ids) - 1)]

    def _rot(self, camera_idx):
        """The call's rotation (Rotater.get_torch_transforms(map_rotation_ids(camera_indices)), rotater.py:74-100)."""
        if self.rotater is None:
            return None
        r = dict(self.rotater)
        r["T"] = torch.as_tensor(self.rotater["transforms"], dtype=torch.float32)[self._cam(camera_idx)]
        return r

    def _near_plane(self):
        # NearFarCollider (scene_colliders.py:217-222): 0 in eval mode
        return self.spec.near_plane if self.training else 0.0

    def _levels(self, o, d, jitter, proposal_requires_grad=False, camera_idx=0):
        return sample_levels(self.P, self.spec, o, d, self._near_plane(), self.spec.far_plane, self.aabb,
                             rot=self._rot(camera_idx), disable_inside_aabb=self.disable_inside_aabb,
                             jitter=jitter if self.training else None, anneal=self.anneal,
                             proposal_requires_grad=proposal_requires_grad,
                             edges_override=getattr(self, "edges_override", None))

    @torch.no_grad()
    def render(self, o3n, v3n, near=0.0, camera_idx=0, jitter=None, return_aux=False):
        o = o3n.t().contiguous()
        v = v3n.t().contiguous()
        o, d = rays_mi_to_nerf(o + near * v, v, self.scene_scale)
        lv = self._levels(o, d, jitter, camera_idx=camera_idx)
        o_s, d_s = apply_frustums(self._rot(camera_idx), o, d, lv["t_edges"][2])
        rgb, w, sigma, rgb_s = main_stage(self.P, self.spec, o, d, lv["t_edges"][2], self.aabb,
                                          training=self.training, camera_idx=self._cam(camera_idx),
                                          always_use_camera_indices=self.always_use_camera_indices,
                                          disable_inside_aabb=self.disable_inside_aabb, o_s=o_s, d_s=d_s)
        if not return_aux:
            return rgb
        N = o.shape[0]
        aux = dict(rgb=rgb, weights0=lv["weights"][0][..., 0], weights1=lv["weights"][1][..., 0],
                   weights2=w[..., 0], sdist0=lv["s_edges"][0].expand(N, -1), sdist1=lv["s_edges"][1],
                   sdist2=lv["s_edges"][2], inds1=lv["inds"][0], inds2=lv["inds"][1],
                   cdf1=lv["cdf"][0], cdf2=lv["cdf"][1], density2=sigma[..., 0], rgb_samples=rgb_s,
                   accumulation=torch.sum(w, dim=-2), t_edges2=lv["t_edges"][2])
        return rgb, aux

    @torch.no_grad()
    def jvp(self, o3n, v3n, do3n, dv3n, near=0.0, camera_idx=0, jitter=None):
        """nerf_emitter_op.py:62-89 + sdf_nerfacto.py:85-113: sampling on primals; tangents attached to
        the 48 final samples only; NaN tangents zeroed."""
        from torch.autograd import forward_ad
        o = o3n.t().contiguous() + near * v3n.t().contiguous()
        v = v3n.t().contiguous()
        go = do3n.t().contiguous() + near * dv3n.t().contiguous()
        gv = dv3n.t().contiguous()
        on, dn = rays_mi_to_nerf(o, v, self.scene_scale)
        lv = self._levels(on, dn, jitter, camera_idx=camera_idx)
        with forward_ad.dual_level():
            od, dd = rays_mi_to_nerf(forward_ad.make_dual(o, go), forward_ad.make_dual(v, gv), self.scene_scale)
            o_s = d_s = None
            if self.rotater is not None:
                # sdf_nerfacto.py:96-104: the tangents are attached to the frustums the sampler returned, i.e. to the
                # already rotated per-sample origins / directions; the tangents themselves stay those of the ray
                po, pd = apply_frustums(self._rot(camera_idx), on, dn, lv["t_edges"][2])
                to, td = forward_ad.unpack_dual(od).tangent, forward_ad.unpack_dual(dd).tangent
                o_s = forward_ad.make_dual(po.contiguous(), to[:, None, :].expand_as(po).contiguous())
                d_s = forward_ad.make_dual(pd.contiguous(), td[:, None, :].expand_as(pd).contiguous())
            rgb, _, _, _ = main_stage(self.P, self.spec, od, dd, lv["t_edges"][2], self.aabb,
                                      training=self.training, camera_idx=self._cam(camera_idx),
                                      always_use_camera_indices=self.always_use_camera_indices,
                                      disable_inside_aabb=self.disable_inside_aabb, grad_scaling=False, o_s=o_s, d_s=d_s)
            t = forward_ad.unpack_dual(rgb).tangent
        t = t.clone()
        t[torch.isnan(t)] = 0.
        return t

    def vjp(self, o3n, v3n, grad_out, near=0.0, camera_idx=0, jitter=None, proposal_requires_grad=True):
        """nerf_emitter_op.py:92-117 + sdf_nerfacto.py:115-187 + mitsuba_sdf.py:1170-1177 for ONE chunk
        (the reg-loss means are over the rays given).  Returns grad_o (3,N), grad_v (3,N),
        {name: grad}, rgb."""
        P = {k: v.detach().clone().requires_grad_(True) for k, v in self.P.items()}
        saved, self.P = self.P, P
        try:
            oo = (o3n.t().contiguous() + near * v3n.t().contiguous()).detach().requires_grad_(True)
            vv = v3n.t().contiguous().detach().requires_grad_(True)
            o, d = rays_mi_to_nerf(oo, vv, self.scene_scale)
            # the proposal levels are rotated (the hook runs in every get_ray_samples); the final samples are NOT: the
            # ray's own origins / directions are put back before the field is evaluated (sdf_nerfacto.py:129-131)
            lv = self._levels(o.detach(), d.detach(), jitter, proposal_requires_grad=proposal_requires_grad,
                              camera_idx=camera_idx)
            rgb, w, _, _ = main_stage(P, self.spec, o, d, lv["t_edges"][2], self.aabb, training=self.training,
                                      camera_idx=self._cam(camera_idx), disable_inside_aabb=self.disable_inside_aabb,
                                      always_use_camera_indices=self.always_use_camera_indices)
            ws = lv["weights"] + [w]
            loss = torch.sum(rgb * grad_out)
            loss = loss + self.interlevel_mult * interlevel_loss(ws, lv["s_edges"])
            loss = loss + self.distortion_mult * distortion_loss(ws, lv["s_edges"])
            loss.backward()
            gv = vv.grad
            go = oo.grad - near * gv
            grads = {k: (p.grad if p.grad is not None else torch.zeros_like(p)) for k, p in P.items()}
        finally:
            self.P = saved
        return go.t().contiguous(), gv.t().contiguous(), grads, rgb.detach()

    def point_lights(self, o3n, v3n, near=0.0, camera_idx=0, jitter=None, rgb_weights=(0.299, 0.587, 0.114)):
        """SdfNerfactoModel.get_point_lights (nerfstudio/models/sdf_nerfacto.py:189-212) =
        proposal sampling (no grad) + NerfactoModel.calc_brightness_grad (nerfstudio/models/nerfacto.py:547-573:
        field + t^2 gradient scaling + get_weights + RGBRenderer, brightness = sum(rgb * rgb_weights),
        autograd.grad w.r.t. ray_bundle.origins with the samples' directions NOT carrying the origin) +
        DepthRenderer(method="contrib") (nerfstudio/model_components/renderers.py:389-397).
        Returns rgb (N,3), depth (N,1), brightness_grad (N,3) in NeRF coordinates (d / d ray_bundle.origins)."""
        wts = torch.tensor(rgb_weights, dtype=torch.float32)
        o = o3n.t().contiguous()
        v = v3n.t().contiguous()
        o, d = rays_mi_to_nerf(o + near * v, v, self.scene_scale)
        o = o.detach().requires_grad_(True)
        lv = self._levels(o.detach(), d, jitter, camera_idx=camera_idx)
        t_edges = lv["t_edges"][2]
        o_s, d_s = apply_frustums(self._rot(camera_idx), o, d, t_edges)      # calc_brightness_grad, nerfacto.py:553-554
        rgb, w, _, rgb_s = main_stage(self.P, self.spec, o, d, t_edges, self.aabb, training=self.training,
                                      camera_idx=self._cam(camera_idx), disable_inside_aabb=self.disable_inside_aabb,
                                      always_use_camera_indices=self.always_use_camera_indices, o_s=o_s, d_s=d_s)
        brightness = torch.sum(rgb * wts, dim=-1, keepdim=True)
        (grad,) = torch.autograd.grad(brightness, o, grad_outputs=torch.ones_like(brightness))
        lum = torch.sum(rgb_s.detach() * wts, dim=-1, keepdim=True)
        steps = (t_edges[..., :-1, None] + t_edges[..., 1:, None]) / 2
        contrib = torch.sum(w.detach() * lum, dim=-1)
        idx = torch.max(contrib, dim=-1, keepdim=True).indices
        depth = torch.gather(steps[..., 0], dim=-1, index=idx)
        return {"rgb": rgb.detach(), "depth": depth, "brightness_grad": grad, "contrib": contrib}
