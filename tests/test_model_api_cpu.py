"""CPU: the nerfstudio-named modules (nerf_emitter_b200/nerfstudio_api.py) expose the reference's parameter names and
shapes (SURVEY.md 8b "Parameter ownership") and refuse to run without the CUDA library path (no CPU fallback)."""
import pytest
import torch


@pytest.fixture(scope="module")
def model(spec):
    from nerf_emitter_b200.nerfstudio_api import SdfNerfactoModel
    return SdfNerfactoModel(num_train_data=spec.num_images, spec=spec)


def test_parameter_names_and_shapes(model, spec):
    from nerf_emitter_b200 import all_param_names, param_shapes
    named = dict(model.named_parameters())
    shapes = param_shapes(spec)
    assert sorted(named) == sorted(all_param_names())
    for k, p in named.items():
        assert tuple(p.shape) == shapes[k], k
    groups = model.get_param_groups()
    assert len(groups["fields"]) == 12 and len(groups["proposal_networks"]) == 10
    # hash tables start in U(-1e-3, 1e-3) (encodings.py:309-312)
    t = named["field.mlp_base_grid.hash_table"]
    assert 9e-4 < float(t.detach().abs().max()) <= 1.0001e-3


def test_state_dict_round_trip_with_reference_keys(model, params_factory):
    P = params_factory(0, 1e3)
    missing, unexpected = model.load_state_dict(P, strict=False)
    assert not unexpected
    sd = model.state_dict()
    for k, v in P.items():
        assert torch.equal(sd[k], v), k
    assert torch.equal(sd["proposal_networks.1.mlp_base.0.hash_table"], P["proposal_networks.1.encoding.hash_table"])


def test_no_cpu_path(model):
    from nerf_emitter_b200 import lib
    from nerf_emitter_b200.nerfstudio_api import HashEncoding, RayBundle
    with pytest.raises(lib.NemitError):
        model.get_rgb_for_camera_ray_bundle(RayBundle(origins=torch.zeros(4, 3), directions=torch.ones(4, 3)))
    with pytest.raises(lib.NemitError):
        HashEncoding(num_levels=2, min_res=4, max_res=8, log2_hashmap_size=4)(torch.rand(5, 3))
    with pytest.raises(lib.NemitError):
        model.field.density_fn(torch.zeros(2, 3))
