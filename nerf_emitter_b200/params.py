"""Parameter set of the emitter NeRF (names follow the reference ``state_dict``).

The library never owns parameter memory: these tensors are ordinary torch tensors that the
caller (a nerfstudio checkpoint, the benchmark, a test) provides; the C-ABI borrows their
device pointers per call (SURVEY.md 8b "Parameter ownership").

Names and shapes (probed from the reference ``named_parameters()``):
  field.mlp_base_grid.hash_table                       (L*T, 2)   nerfacto_field.py:145-152
  field.mlp_base_mlp.layers.{0,1}.{weight,bias}        (64,32),(64),(16,64),(16)   :153-161
  field.mlp_head.layers.{0,1,2}.{weight,bias}          (64,63),(64),(64,64),(64),(3,64),(3)   :215-224
  field.embedding_appearance.embedding.weight          (num_images, 32)   :133-135
  proposal_networks.{i}.encoding.hash_table            (5*2^17, 2)   density_fields.py:67-74
  proposal_networks.{i}.mlp_base.1.layers.{0,1}.{weight,bias}   (16,10),(16),(1,16),(1)   :77-87
"""
from __future__ import annotations

import math
from dataclasses import dataclass, field
from typing import Dict, List, Tuple

import torch


@dataclass
class GridSpec:
    """One multiresolution hash grid (encodings.py:264-312)."""
    num_levels: int
    base_res: int
    max_res: int
    log2_hashmap_size: int
    features_per_level: int = 2

    @property
    def table_size(self) -> int:
        return 1 << self.log2_hashmap_size

    def scalings(self) -> List[float]:
        """floor(min_res * g^l) with g computed in float64 then applied as the reference does
        (encodings.py:281-283): numpy float64 growth factor, torch float32 pow of an int64 arange."""
        L = self.num_levels
        if L > 1:
            g = math.exp((math.log(self.max_res) - math.log(self.base_res)) / (L - 1))
        else:
            g = 1.0
        levels = torch.arange(L)
        # python float ** int64 tensor -> float32 tensor in torch; then * min_res, floor.
        sc = torch.floor(self.base_res * g ** levels)
        return [float(v) for v in sc]


@dataclass
class EmitterModelSpec:
    """Numeric constants of the sdf-nerfacto emitter (SURVEY.md 8 header)."""
    main_grid: GridSpec = field(default_factory=lambda: GridSpec(16, 16, 2048, 19))
    prop_grids: Tuple[GridSpec, GridSpec] = field(
        default_factory=lambda: (GridSpec(5, 16, 128, 17), GridSpec(5, 16, 256, 17)))
    hidden_dim: int = 64
    geo_feat_dim: int = 15
    appearance_dim: int = 32
    prop_hidden: int = 16
    num_images: int = 4
    num_proposal_samples: Tuple[int, int] = (256, 96)
    num_nerf_samples: int = 48
    near_plane: float = 0.05
    far_plane: float = 1000.0
    rgb_bias: float = -5.0


PARAM_NAMES_MAIN = [
    "field.mlp_base_grid.hash_table",
    "field.mlp_base_mlp.layers.0.weight", "field.mlp_base_mlp.layers.0.bias",
    "field.mlp_base_mlp.layers.1.weight", "field.mlp_base_mlp.layers.1.bias",
    "field.mlp_head.layers.0.weight", "field.mlp_head.layers.0.bias",
    "field.mlp_head.layers.1.weight", "field.mlp_head.layers.1.bias",
    "field.mlp_head.layers.2.weight", "field.mlp_head.layers.2.bias",
    "field.embedding_appearance.embedding.weight",
]


def prop_param_names(i: int) -> List[str]:
    p = "proposal_networks.%d." % i
    return [p + "encoding.hash_table",
            p + "mlp_base.1.layers.0.weight", p + "mlp_base.1.layers.0.bias",
            p + "mlp_base.1.layers.1.weight", p + "mlp_base.1.layers.1.bias"]


def all_param_names() -> List[str]:
    return PARAM_NAMES_MAIN + prop_param_names(0) + prop_param_names(1)


def param_shapes(spec: EmitterModelSpec) -> Dict[str, Tuple[int, ...]]:
    mg = spec.main_grid
    enc = mg.num_levels * mg.features_per_level
    head_in = 16 + spec.geo_feat_dim + spec.appearance_dim
    shapes = {
        "field.mlp_base_grid.hash_table": (mg.num_levels * mg.table_size, mg.features_per_level),
        "field.mlp_base_mlp.layers.0.weight": (spec.hidden_dim, enc),
        "field.mlp_base_mlp.layers.0.bias": (spec.hidden_dim,),
        "field.mlp_base_mlp.layers.1.weight": (1 + spec.geo_feat_dim, spec.hidden_dim),
        "field.mlp_base_mlp.layers.1.bias": (1 + spec.geo_feat_dim,),
        "field.mlp_head.layers.0.weight": (spec.hidden_dim, head_in),
        "field.mlp_head.layers.0.bias": (spec.hidden_dim,),
        "field.mlp_head.layers.1.weight": (spec.hidden_dim, spec.hidden_dim),
        "field.mlp_head.layers.1.bias": (spec.hidden_dim,),
        "field.mlp_head.layers.2.weight": (3, spec.hidden_dim),
        "field.mlp_head.layers.2.bias": (3,),
        "field.embedding_appearance.embedding.weight": (spec.num_images, spec.appearance_dim),
    }
    for i, g in enumerate(spec.prop_grids):
        n = prop_param_names(i)
        pe = g.num_levels * g.features_per_level
        shapes[n[0]] = (g.num_levels * g.table_size, g.features_per_level)
        shapes[n[1]] = (spec.prop_hidden, pe)
        shapes[n[2]] = (spec.prop_hidden,)
        shapes[n[3]] = (1, spec.prop_hidden)
        shapes[n[4]] = (1,)
    return shapes


def make_params(spec: EmitterModelSpec | None = None, seed: int = 0, table_scale: float = 1.0,
                mlp_scale: float = 1.0, device: str | torch.device = "cpu") -> Dict[str, torch.Tensor]:
    """Seeded random-init parameters with the reference's init distributions.

    hash tables ~ U(-1e-3, 1e-3) * table_scale (encodings.py:310-312); linear weight and bias
    ~ U(-1/sqrt(fan_in), 1/sqrt(fan_in)) (torch ``nn.Linear`` default); embedding ~ N(0,1).
    ``table_scale=1e3`` is the "stress init" of SURVEY.md 8d (densities / colours that vary).
    Drawn from an explicit CPU generator so the same seed gives the same weights on every box.
    """
    spec = spec or EmitterModelSpec()
    g = torch.Generator(device="cpu")
    g.manual_seed(seed)
    out: Dict[str, torch.Tensor] = {}
    for name, shape in param_shapes(spec).items():
        if name.endswith("hash_table"):
            t = (torch.rand(shape, generator=g) * 2 - 1) * (1e-3 * table_scale)
        elif name.endswith("embedding.weight"):
            t = torch.randn(shape, generator=g)
        else:
            if name.endswith(".weight"):
                fan_in = shape[1]
                last_fan_in = fan_in
            else:
                fan_in = last_fan_in
            bound = mlp_scale / math.sqrt(fan_in)
            t = (torch.rand(shape, generator=g) * 2 - 1) * bound
        out[name] = t.to(device=device, dtype=torch.float32).contiguous()
    return out
