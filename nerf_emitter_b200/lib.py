"""ctypes binding of libnerfemitter.so (include/nemit.h).  Fails loudly when the library is missing:
there is no CPU or torch fallback for the emitter path."""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
# NEMIT_LIB selects another build of the same library (A/B variants of single kernels made by tools/build_variant.sh)
LIB_PATH = os.environ.get("NEMIT_LIB") or os.path.join(_HERE, "libnerfemitter.so")
MAX_LEVELS = 16

c_float_p = C.POINTER(C.c_float)
c_int32_p = C.POINTER(C.c_int32)


class NemitGrid(C.Structure):
    _fields_ = [("num_levels", C.c_int32), ("log2_table_size", C.c_int32), ("scalings", C.c_float * MAX_LEVELS)]


class NemitConfig(C.Structure):
    _fields_ = [
        ("main_grid", NemitGrid), ("prop_grid", NemitGrid * 2),
        ("num_prop_samples", C.c_int32 * 2), ("num_nerf_samples", C.c_int32),
        ("hidden_dim", C.c_int32), ("geo_feat_dim", C.c_int32), ("appearance_dim", C.c_int32),
        ("prop_hidden_dim", C.c_int32), ("num_images", C.c_int32),
        ("near_plane", C.c_float), ("far_plane", C.c_float), ("scene_scale", C.c_float),
        ("aabb", C.c_float * 6), ("rgb_bias", C.c_float),
        ("disable_inside_aabb", C.c_int32), ("cpu_sum_width", C.c_int32),
        ("interlevel_mult", C.c_float), ("distortion_mult", C.c_float),
        ("lin0_host", c_float_p), ("u1_host", c_float_p), ("u2_host", c_float_p),
        ("use_fake_contraction", C.c_int32), ("far_intersect", C.c_int32),
    ]


PARAM_FIELDS = [
    "main_table", "base_w0", "base_b0", "base_w1", "base_b1", "head_w0", "head_b0", "head_w1", "head_b1",
    "head_w2", "head_b2", "appearance",
]


class NemitParams(C.Structure):
    _fields_ = [(n, C.c_void_p) for n in PARAM_FIELDS] + [
        ("prop_table", C.c_void_p * 2), ("prop_w0", C.c_void_p * 2), ("prop_b0", C.c_void_p * 2),
        ("prop_w1", C.c_void_p * 2), ("prop_b1", C.c_void_p * 2)]


class NemitRays(C.Structure):
    _fields_ = [("o", C.c_void_p), ("v", C.c_void_p), ("near_per_ray", C.c_void_p), ("near_scalar", C.c_float),
                ("n", C.c_int64), ("camera_idx", C.c_int32), ("training", C.c_int32), ("jitter", C.c_void_p),
                ("anneal", C.c_float), ("layout", C.c_int32),
                ("kept_s_edges1", C.c_void_p), ("kept_s_edges2", C.c_void_p), ("kept_t_edges2", C.c_void_p)]


class NemitLevelsOut(C.Structure):
    _fields_ = [(n, C.c_void_p) for n in
                ["weights0", "weights1", "s_edges1", "s_edges2", "t_edges2", "inds1", "inds2", "weights2"]]


MLP_MAX_LAYERS = 8
MLP_MAX_WIDTH = 64


class NemitMlp(C.Structure):
    _fields_ = [("in_dim", C.c_int32), ("num_layers", C.c_int32), ("out_dim", C.c_int32 * MLP_MAX_LAYERS),
                ("weight", C.c_void_p * MLP_MAX_LAYERS), ("bias", C.c_void_p * MLP_MAX_LAYERS),
                ("activation", C.c_int32), ("out_activation", C.c_int32)]


class NemitAdamGroup(C.Structure):
    _fields_ = [(n, C.c_double) for n in ["lr", "beta1", "beta2", "eps"]] + [("max_norm", C.c_float), ("max_value", C.c_float),
                                                                              ("enabled", C.c_int32), ("reserved", C.c_int32)]


class NemitSphericalSensor(C.Structure):
    _fields_ = [("origin", C.c_float * 3), ("near_clip", C.c_float), ("width", C.c_int32), ("height", C.c_int32),
                ("spp", C.c_int32), ("seed", C.c_uint32)]


class NemitVmfFrame(C.Structure):
    _fields_ = [("world_to_local", C.c_float * 12), ("local_to_world", C.c_float * 12), ("bsphere_center", C.c_float * 3),
                ("bsphere_radius", C.c_float)]


class NemitVmfLobes(C.Structure):
    _fields_ = [("position", C.c_void_p), ("weight", C.c_void_p), ("std", C.c_void_p), ("num_lobes", C.c_int32)]


# every symbol include/nemit.h declares: (restype, argtypes)
_H = C.c_void_p
SYMBOLS = {
    "nemit_last_error": (C.c_char_p, []),
    "nemit_version": (C.c_int, []),
    "nemit_create": (C.c_int, [C.POINTER(NemitConfig), C.c_int, C.POINTER(_H)]),
    "nemit_destroy": (C.c_int, [_H]),
    "nemit_emitter_local_rays": (C.c_int, [C.POINTER(C.c_float), C.POINTER(C.c_float), C.POINTER(C.c_float), C.c_float,
                                           C.c_int32, C.c_int32, C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p,
                                           C.c_void_p, C.c_void_p, C.c_void_p]),
    "nemit_set_appearance_mode": (C.c_int, [_H, C.c_int32, C.c_int32, C.POINTER(C.c_int32), C.c_int32]),
    "nemit_set_rotater": (C.c_int, [_H, C.c_int32, C.POINTER(C.c_float), C.c_int32, C.POINTER(C.c_float)]),
    "nemit_bind_params": (C.c_int, [_H, C.POINTER(NemitParams)]),
    "nemit_bind_grads": (C.c_int, [_H, C.POINTER(NemitParams)]),
    "nemit_set_precision": (C.c_int, [_H, C.c_int32]),
    "nemit_workspace_bytes": (C.c_size_t, [_H, C.c_int64]),
    "nemit_forward": (C.c_int, [_H, C.POINTER(NemitRays), C.c_void_p, C.c_void_p, C.POINTER(NemitLevelsOut),
                                C.c_void_p, C.c_size_t, C.c_void_p]),
    "nemit_sample_levels": (C.c_int, [_H, C.POINTER(NemitRays), C.POINTER(NemitLevelsOut), C.c_void_p, C.c_size_t,
                                      C.c_void_p]),
    "nemit_jvp": (C.c_int, [_H, C.POINTER(NemitRays), C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t,
                            C.c_void_p]),
    "nemit_backward": (C.c_int, [_H, C.POINTER(NemitRays), C.c_void_p, C.c_int64, C.c_float, C.c_int32, C.c_void_p,
                                 C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p]),
    "nemit_hash_encode": (C.c_int, [_H, C.c_int32, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p]),
    "nemit_point_lights": (C.c_int, [_H, C.POINTER(NemitRays), C.POINTER(C.c_float), C.c_void_p, C.c_void_p, C.c_void_p,
                                     C.c_void_p, C.c_size_t, C.c_void_p]),
    "nemit_adam_step": (C.c_int, [_H, C.POINTER(NemitParams), C.POINTER(NemitParams), C.POINTER(NemitAdamGroup),
                                  C.POINTER(NemitAdamGroup), C.c_void_p, C.c_float, C.c_int32, C.c_void_p, C.c_void_p]),
    "nemit_spherical_rays": (C.c_int, [C.POINTER(NemitSphericalSensor), C.c_int64, C.c_int64, C.c_void_p, C.c_void_p,
                                       C.c_void_p, C.c_void_p, C.c_void_p]),
    "nemit_film_accumulate": (C.c_int, [C.c_int32, C.c_int32, C.c_int32, C.c_float, C.c_void_p, C.c_void_p, C.c_int64,
                                        C.c_void_p, C.c_void_p]),
    "nemit_film_develop": (C.c_int, [C.c_int32, C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p]),
    "nemit_bake_workspace_bytes": (C.c_size_t, [_H]),
    "nemit_bake_envmap": (C.c_int, [_H, C.POINTER(NemitSphericalSensor), C.c_int32, C.c_float, C.c_int32, C.c_int32,
                                    C.c_int32, C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p]),
    "nemit_compensate_weights": (C.c_int, [C.c_void_p, C.c_int64, C.c_float, C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p]),
    "nemit_gmm_em_step": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p,
                                    C.c_float, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    "nemit_hash_encode_grid": (C.c_int, [C.POINTER(NemitGrid), C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p]),
    "nemit_sh_encode": (C.c_int, [C.c_void_p, C.c_int64, C.c_int32, C.c_void_p, C.c_void_p]),
    "nemit_mlp": (C.c_int, [C.POINTER(NemitMlp), C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p]),
    "nemit_density_fn": (C.c_int, [_H, C.c_int32, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p]),
    "nemit_field_samples": (C.c_int, [_H, C.c_void_p, C.c_void_p, C.c_int64, C.c_int32, C.c_int32, C.c_void_p,
                                      C.c_void_p, C.c_void_p]),
    "nemit_vmf_sample": (C.c_int, [C.POINTER(NemitVmfLobes), C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p,
                                   C.c_void_p, C.c_void_p]),
    "nemit_vmf_pdf": (C.c_int, [C.POINTER(NemitVmfLobes), C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p,
                                C.c_void_p]),
    "nemit_vmf_sample_direction": (C.c_int, [C.POINTER(NemitVmfLobes), C.POINTER(NemitVmfFrame), C.c_void_p, C.c_void_p,
                                             C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    "nemit_vmf_pdf_direction": (C.c_int, [C.POINTER(NemitVmfLobes), C.POINTER(NemitVmfFrame), C.c_void_p, C.c_void_p,
                                          C.c_int64, C.c_void_p, C.c_void_p]),
    "nemit_forward_host": (C.c_int, [_H, C.POINTER(NemitRays), C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p]),
    "nemit_forward_noalloc": (C.c_int, [_H, C.POINTER(NemitRays), C.c_void_p, C.c_void_p]),
    "nemit_launch_count": (C.c_int64, [_H]),
    "nemit_set_sampler": (C.c_int, [_H, C.c_int32]),
    "nemit_profile_enable": (C.c_int, [_H, C.c_int32]),
    "nemit_profile_read": (C.c_int, [_H, C.POINTER(C.c_double), C.POINTER(C.c_int64), C.POINTER(C.c_int64),
                                     C.POINTER(C.c_int64), C.c_int32]),
    "nemit_backward_host": (C.c_int, [_H, C.POINTER(NemitRays), C.c_void_p, C.c_int32, C.c_int64, C.c_float, C.c_int32,
                                      C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p]),
    "nemit_backward_host_workspace_bytes": (C.c_size_t, [_H, C.c_int64]),
    "nemit_comm_unique_id": (C.c_int, [C.c_void_p]),
    "nemit_comm_init": (C.c_int, [_H, C.c_void_p, C.c_int32, C.c_int32]),
    "nemit_comm_destroy": (C.c_int, [_H]),
    "nemit_allreduce_grads": (C.c_int, [_H, C.c_void_p, C.c_int64, C.c_float, C.c_void_p]),
    "nemit_microbench": (C.c_int, [_H, C.c_int32, C.POINTER(C.c_double), C.c_void_p]),
}

MICROBENCH_KINDS = ["gather_l1miss", "gather_l1hit", "gather_coherent", "red", "l2_read"]
NOALLOC_MAX_RAYS = 16384        # NEMIT_NOALLOC_MAX_RAYS
UNIQUE_ID_BYTES = 128

PROFILE_KINDS = ["k_sample_levels", "k_field_forward", "k_field_forward<adjoint_seed>", "k_field_backward",
                 "k_prop_backward", "k_finalize_ray_grads", "k_field_jvp", "k_field_forward_tc", "k_field_backward_tc"]

_LIB = None


class NemitError(RuntimeError):
    pass


def load():
    """Loads the shared library (once). Raises if it has not been built: no fallback exists."""
    global _LIB
    if _LIB is not None:
        return _LIB
    if not os.path.exists(LIB_PATH):
        raise NemitError(
            "libnerfemitter.so is not built (%s). Run `python -c 'import __graft_entry__ as g; g.build()'` "
            "or `make -C nerf_emitter_b200/csrc`. There is no CPU/torch fallback for this path." % LIB_PATH)
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in SYMBOLS.items():
        fn = getattr(lib, name)          # AttributeError if the export is missing
        fn.restype = res
        fn.argtypes = args
    _LIB = lib
    return lib


def check(rc: int, what: str):
    if rc != 0:
        msg = load().nemit_last_error()
        raise NemitError("%s failed (%d): %s" % (what, rc, msg.decode() if msg else "?"))
