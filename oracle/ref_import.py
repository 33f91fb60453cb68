"""TEST INFRASTRUCTURE ONLY (oracle). Import shim for the reference's own torch modules.

Works only where /root/reference exists (the build container); never on the GPU box.
Two stubs are installed before import:
  * ``nerfacc``  - imported at module scope by ray_samplers.py:24 / renderers.py:33, never called
    on the nerfacto path;
  * ``nerfstudio.configs.base_config`` - the real file fails on Python >= 3.11 dataclasses
    (base_config.py:113-118); base_field.py:28 only needs the ``InstantiateConfig`` name.
"""
import os
import sys
import types

REF_ROOT = os.environ.get("NEMIT_REFERENCE_ROOT", "/root/reference")


def reference_available() -> bool:
    return os.path.isdir(os.path.join(REF_ROOT, "nerfstudio"))


def import_reference():
    """Returns a namespace with the reference modules used on the emitter path."""
    if not reference_available():
        raise RuntimeError("reference tree not present at %s" % REF_ROOT)
    if "nerfacc" not in sys.modules:
        m = types.ModuleType("nerfacc")
        m.OccGridEstimator = object
        m.accumulate_along_rays = None
        sys.modules["nerfacc"] = m
    if "nerfstudio.configs.base_config" not in sys.modules:
        m = types.ModuleType("nerfstudio.configs.base_config")

        class InstantiateConfig:  # bare stand-in
            pass

        class PrintableConfig:
            pass

        m.InstantiateConfig = InstantiateConfig
        m.PrintableConfig = PrintableConfig
        sys.modules["nerfstudio.configs.base_config"] = m
    if REF_ROOT not in sys.path:
        sys.path.insert(0, REF_ROOT)
    ns = types.SimpleNamespace()
    from nerfstudio.cameras import rays
    from nerfstudio.field_components import activations, encodings, mlp, spatial_distortions
    from nerfstudio.fields import density_fields, nerfacto_field
    from nerfstudio.field_components.field_heads import FieldHeadNames
    from nerfstudio.model_components import losses, ray_samplers, renderers, scene_colliders
    ns.rays, ns.activations, ns.encodings, ns.mlp = rays, activations, encodings, mlp
    ns.spatial_distortions, ns.density_fields, ns.nerfacto_field = spatial_distortions, density_fields, nerfacto_field
    ns.losses, ns.ray_samplers, ns.renderers, ns.scene_colliders = losses, ray_samplers, renderers, scene_colliders
    ns.FieldHeadNames = FieldHeadNames
    return ns
