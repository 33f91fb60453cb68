"""nerf_emitter_b200 - B200-native (sm_100a) NeRF-emitter evaluation path of gerwang/nerf-emitter.

Only the emitter hot path lives here: ``csrc/`` holds the CUDA kernels and the C ABI
(``include/nemit.h``); the Python modules are the host-side mirror of the reference interfaces.
"""
from .params import EmitterModelSpec, GridSpec, all_param_names, make_params, param_shapes  # noqa: F401


def __getattr__(name):
    if name == "NerfEmitter":
        from .emitter import NerfEmitter
        return NerfEmitter
    if name == "VmfMixture":
        from .vmf import VmfMixture
        return VmfMixture
    if name == "VmfGuiding":
        from .vmf import VmfGuiding
        return VmfGuiding
    raise AttributeError(name)
