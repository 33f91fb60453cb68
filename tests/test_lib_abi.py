"""CPU: the C-ABI library builds, loads and exports every symbol include/nemit.h declares
(no compute calls: there is no GPU here)."""
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def built():
    import __graft_entry__ as g
    g.build()
    from nerf_emitter_b200 import lib
    return lib


def test_header_symbols_are_exported(built):
    text = open(os.path.join(ROOT, "include", "nemit.h")).read()
    declared = set(re.findall(r"\b(nemit_[a-z_0-9]+)\s*\(", text))
    assert declared, "no declarations parsed"
    assert declared == set(built.SYMBOLS), declared ^ set(built.SYMBOLS)
    lib = built.load()
    for name in declared:
        assert hasattr(lib, name), name
    assert lib.nemit_version() >= 100


def test_struct_sizes_match_header(built):
    import ctypes as C
    # 22 parameter pointers, 8 level outputs: a layout drift between nemit.h and lib.py shows up here
    assert C.sizeof(built.NemitParams) == 22 * C.sizeof(C.c_void_p)
    assert C.sizeof(built.NemitLevelsOut) == 8 * C.sizeof(C.c_void_p)
    assert C.sizeof(built.NemitGrid) == 8 + 4 * 16


def test_struct_layouts_match_a_c_compiler(built, tmp_path):
    """sizeof / offsetof of every struct of include/nemit.h as gcc sees them vs the ctypes mirrors in lib.py."""
    import ctypes as C
    import subprocess
    structs = {"NemitGrid": built.NemitGrid, "NemitConfig": built.NemitConfig, "NemitParams": built.NemitParams,
               "NemitRays": built.NemitRays, "NemitLevelsOut": built.NemitLevelsOut, "NemitMlp": built.NemitMlp,
               "NemitAdamGroup": built.NemitAdamGroup, "NemitSphericalSensor": built.NemitSphericalSensor,
               "NemitVmfLobes": built.NemitVmfLobes}
    lines = ['#include <stdio.h>', '#include <stddef.h>', '#include "nemit.h"', 'int main(void) {']
    for name, cls in structs.items():
        lines.append('  printf("%s sizeof %%zu\\n", sizeof(%s));' % (name, name))
        for field, _ in cls._fields_:
            lines.append('  printf("%s %s %%zu\\n", offsetof(%s, %s));' % (name, field, name, field))
    lines += ['  return 0;', '}']
    src = tmp_path / "probe.c"
    src.write_text("\n".join(lines))
    exe = tmp_path / "probe"
    subprocess.run(["gcc", "-std=c99", "-I", os.path.join(ROOT, "include"), str(src), "-o", str(exe)], check=True)
    out = subprocess.run([str(exe)], check=True, capture_output=True, text=True).stdout
    seen = 0
    for ln in out.splitlines():
        name, field, val = ln.split()
        cls = structs[name]
        expect = C.sizeof(cls) if field == "sizeof" else getattr(cls, field).offset
        assert int(val) == expect, (name, field, val, expect)
        seen += 1
    assert seen > 60


def test_unsupported_model_variants_are_rejected_loudly(built):
    """nemit_create must not silently evaluate the default model when the config asks for FakeContraction /
    far_intersect (NS/models/nerfacto.py:144-154,263-270, NS/models/sdf_nerfacto.py:193-194)."""
    import ctypes as C
    import torch
    from nerf_emitter_b200 import EmitterModelSpec
    from nerf_emitter_b200.emitter import sampler_tables
    lib = built.load()
    spec = EmitterModelSpec()
    tabs = sampler_tables(spec)
    for flag in ("use_fake_contraction", "far_intersect"):
        cfg = built.NemitConfig()
        for g, dst in [(spec.main_grid, cfg.main_grid), (spec.prop_grids[0], cfg.prop_grid[0]), (spec.prop_grids[1], cfg.prop_grid[1])]:
            dst.num_levels, dst.log2_table_size = g.num_levels, g.log2_hashmap_size
        cfg.num_prop_samples[0], cfg.num_prop_samples[1] = spec.num_proposal_samples
        cfg.num_nerf_samples = spec.num_nerf_samples
        cfg.hidden_dim, cfg.geo_feat_dim, cfg.appearance_dim = spec.hidden_dim, spec.geo_feat_dim, spec.appearance_dim
        cfg.prop_hidden_dim, cfg.num_images = spec.prop_hidden, spec.num_images
        cfg.lin0_host = C.cast(tabs[0].data_ptr(), built.c_float_p)
        cfg.u1_host = C.cast(tabs[1].data_ptr(), built.c_float_p)
        cfg.u2_host = C.cast(tabs[2].data_ptr(), built.c_float_p)
        setattr(cfg, flag, 1)
        h = C.c_void_p()
        assert lib.nemit_create(C.byref(cfg), 0, C.byref(h)) == -1
        assert flag.encode() in lib.nemit_last_error()


def test_argument_errors_without_gpu(built):
    import ctypes as C
    lib = built.load()
    h = C.c_void_p()
    assert lib.nemit_create(None, 0, C.byref(h)) == -1
    assert b"null" in lib.nemit_last_error()
    cfg = built.NemitConfig()
    assert lib.nemit_create(C.byref(cfg), 0, C.byref(h)) == -1      # unsupported (zeroed) shapes
    assert lib.nemit_destroy(None) == 0


def test_product_path_has_no_oracle_import():
    pkg = os.path.join(ROOT, "nerf_emitter_b200")
    for fn in os.listdir(pkg):
        if fn.endswith(".py"):
            src = open(os.path.join(pkg, fn)).read()
            assert "oracle" not in src.replace("# oracle", ""), fn
