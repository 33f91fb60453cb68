"""GPU parity of the model-side API (nerf_emitter_b200/nerfstudio_api.py, SURVEY.md 8b "Model-side API kept intact"):
the nerfstudio-named modules dispatch to the C ABI and agree with the CPU oracle; the RayBundle-level entry points
((N,3) NeRF-coordinate layout) agree bit for bit with the CustomOp-level ones ((3,N) Mitsuba layout) that the other
GPU tests pin to the reference goldens.  Tolerance: 1e-3 max-norm relative (north_star, fp32 mode)."""
import pytest
import torch
import torch.nn.functional as F

from helpers import make_rays, rel_err

pytestmark = pytest.mark.gpu
TOL = 1e-3
DEV = "cuda:0"


def _nerf_rays(o3n, v3n, scale=1.0):
    from oracle.nerfacto_oracle import rays_mi_to_nerf       # checker only
    o, d = rays_mi_to_nerf(o3n.t().contiguous(), v3n.t().contiguous(), scale)
    return o.contiguous(), d.contiguous()


@pytest.fixture(scope="module")
def model(spec, params_factory):
    from nerf_emitter_b200.nerfstudio_api import SdfNerfactoModel
    m = SdfNerfactoModel(num_train_data=spec.num_images, spec=spec)
    missing, unexpected = m.load_state_dict(params_factory(0, 1e3), strict=False)
    assert not unexpected
    assert all(".mlp_base.0." in k or "field.mlp_base." in k or k.endswith(("aabb", "max_res", "num_levels", "log2_hashmap_size"))
               for k in missing), missing     # only aliases of shared tensors and buffers
    return m.to(DEV)


def test_state_dict_keys_are_the_reference_names(model):
    from nerf_emitter_b200 import all_param_names
    sd = model.state_dict()
    for k in all_param_names():
        assert k in sd, k
    # the Sequential aliases the reference's checkpoints also carry (nerfacto_field.py:168, density_fields.py:89)
    assert "field.mlp_base.0.hash_table" in sd and "proposal_networks.0.mlp_base.0.hash_table" in sd
    assert sd["field.mlp_base.0.hash_table"].data_ptr() == sd["field.mlp_base_grid.hash_table"].data_ptr()


def test_components_match_oracle(spec, params_factory):
    from nerf_emitter_b200.nerfstudio_api import MLP, HashEncoding, SHEncoding
    from oracle.nerfacto_oracle import hash_encode, sh16
    P = params_factory(0, 1e3)
    g = torch.Generator().manual_seed(5)
    x = torch.rand(3000, 3, generator=g)
    pg = spec.prop_grids[1]
    enc = HashEncoding(num_levels=pg.num_levels, min_res=pg.base_res, max_res=pg.max_res,
                       log2_hashmap_size=pg.log2_hashmap_size)
    enc.load_state_dict({"hash_table": P["proposal_networks.1.encoding.hash_table"]})
    got = enc.to(DEV)(x.to(DEV)).cpu()
    ref = hash_encode(x, P["proposal_networks.1.encoding.hash_table"], pg)
    assert got.shape == ref.shape and rel_err(got, ref) < 1e-6

    d = torch.randn(7, 11, 3, generator=g)
    got = SHEncoding(levels=4)(d.to(DEV)).cpu()
    assert got.shape == (7, 11, 16) and rel_err(got, sh16(d)) < 1e-6

    mlp = MLP(in_dim=63, num_layers=3, layer_width=64, out_dim=3)
    mlp.load_state_dict({k[len("field.mlp_head."):]: v for k, v in P.items() if k.startswith("field.mlp_head.")})
    xin = torch.randn(513, 63, generator=g)
    h = xin
    for i in range(3):
        h = F.linear(h, P["field.mlp_head.layers.%d.weight" % i], P["field.mlp_head.layers.%d.bias" % i])
        h = F.relu(h) if i < 2 else h
    got = mlp.to(DEV)(xin.to(DEV)).cpu()
    assert rel_err(got, h) < 1e-5
    with pytest.raises(Exception):
        enc(x)           # host tensor: no CPU path


def test_sampler_field_and_density_fns_match_oracle(model, spec, params_factory):
    from oracle.nerfacto_oracle import main_stage, proposal_density
    from nerf_emitter_b200.nerfstudio_api import RayBundle
    P = params_factory(0, 1e3)
    o3n, v3n, _ = make_rays(96, seed=11)
    o, d = _nerf_rays(o3n, v3n)
    model.eval()
    rb = RayBundle(origins=o.to(DEV), directions=d.to(DEV))
    ray_samples, weights_list, ray_samples_list = model.proposal_sampler(rb, density_fns=model.density_fns)
    assert [tuple(w.shape) for w in weights_list] == [(96, 256, 1), (96, 96, 1)]
    assert tuple(ray_samples.frustums.starts.shape) == (96, 48, 1) and len(ray_samples_list) == 2
    # Field.forward on the final samples vs the oracle's field stage at the same edges
    t_edges = torch.cat([ray_samples.frustums.starts[..., 0], ray_samples.frustums.ends[:, -1:, 0]], dim=1).cpu()
    aabb = torch.tensor([[-1.0] * 3, [1.0] * 3])
    _, w_ref, sigma_ref, rgb_ref = main_stage(P, spec, o, d, t_edges, aabb, training=False, camera_idx=0)
    out = model.field.forward(ray_samples)
    assert rel_err(out["density"].cpu(), sigma_ref) < TOL
    assert rel_err(out["rgb"].cpu(), rgb_ref) < TOL
    den, geo = model.field.get_density(ray_samples)
    assert torch.equal(den, out["density"]) and tuple(geo.shape) == (96, 48, 15)
    # proposal density_fns on the level-0 / level-1 sample positions
    for i, rs in enumerate(ray_samples_list):
        pos = rs.frustums.get_positions()
        got = model.density_fns[i](pos).cpu()
        ref = proposal_density(P, i, spec.prop_grids[i], pos.cpu(), aabb, True)
        assert got.shape == ref.shape and rel_err(got, ref) < TOL, i


def test_ray_bundle_entry_points_equal_the_custom_op_layout(model, spec, params_factory):
    """(N,3) NeRF-coordinate RayBundle in == (3,N) Mitsuba arrays in, bit for bit: the affine map is exact in fp32
    for scene_scale = 1, so both layouts feed identical rays to the same kernels."""
    from nerf_emitter_b200 import NerfEmitter
    from nerf_emitter_b200.nerfstudio_api import RayBundle
    N = 300
    o3n, v3n, g = make_rays(N, seed=3)
    o, d = _nerf_rays(o3n, v3n)
    E = NerfEmitter(params_factory(0, 1e3), spec, device=torch.device(DEV))
    rb = RayBundle(origins=o.to(DEV), directions=d.to(DEV), camera_indices=torch.full((N, 1), 2, device=DEV))
    model.eval()
    rgb_b = model.get_rgb_for_camera_ray_bundle(rb)
    rgb_a = E.eval().render(o3n.to(DEV), v3n.to(DEV))
    assert torch.equal(rgb_a, rgb_b)
    # forward mode: tangents map as d_o_nerf = 2*scale*perm(d_o_mi)
    do3n, dv3n = torch.randn(3, N, generator=g), torch.randn(3, N, generator=g)
    perm = [2, 0, 1]
    jb = model.forward_grad_for_camera_ray_bundle(rb, (2.0 * do3n[perm]).t().contiguous().to(DEV),
                                                  dv3n[perm].t().contiguous().to(DEV))
    ja = E.jvp(o3n.to(DEV), v3n.to(DEV), do3n.to(DEV), dv3n.to(DEV))
    assert torch.equal(ja, jb)
    # reverse mode (training mode, camera 2): ray gradients and every parameter gradient
    gout = torch.randn(N, 3, generator=g).to(DEV)
    model.train()
    model.zero_grad(set_to_none=True)
    model.proposal_sampler._step = 100          # steady state: proposal networks without autograd on this call
    model.proposal_sampler._steps_since_update = 0
    model.backward_for_camera_ray_bundle(rb, gout)
    E.train(True)
    E.bind_grads()
    go, gv = E.backward(o3n.to(DEV), v3n.to(DEV), gout, camera_idx=2, chunk_rays=16384, proposal_requires_grad=False)
    inv = [1, 2, 0]                              # g_mi = 2*scale*perm^-1(g_nerf)
    assert rel_err((2.0 * rb.origins.grad.t())[inv], go) < 1e-6
    assert rel_err(rb.directions.grad.t()[inv], gv) < 1e-6
    named = dict(model.named_parameters())
    for k, gref in E.grads.items():
        # same kernels, same inputs; float atomics reorder the sums
        assert rel_err(named[k].grad, gref, floor=1e-12) < 1e-4, k
    assert float(named["field.mlp_base_grid.hash_table"].grad.abs().sum()) > 0
    assert float(named["proposal_networks.0.encoding.hash_table"].grad.abs().sum()) == 0.0


def test_sampler_training_mode_with_jitter(model, spec, params_factory):
    """Training mode with explicit per-ray jitter: returned level-0 / level-1 / final sample edges follow the oracle."""
    from oracle.nerfacto_oracle import OracleEmitter
    from nerf_emitter_b200.nerfstudio_api import RayBundle
    N = 64
    o3n, v3n, g = make_rays(N, seed=13)
    o, d = _nerf_rays(o3n, v3n)
    jit = [torch.rand(N, 1, generator=g) for _ in range(3)]
    model.train()
    try:
        rs, wl, rsl = model.proposal_sampler(RayBundle(origins=o.to(DEV), directions=d.to(DEV)), jitter=jit)
    finally:
        model.eval()
    O = OracleEmitter(params_factory(0, 1e3), spec)
    O.train(True)
    _, aux = O.render(o3n, v3n, jitter=jit, return_aux=True)
    s0 = torch.cat([rsl[0].spacing_starts[..., 0], rsl[0].spacing_ends[:, -1:, 0]], dim=1).cpu()
    assert rel_err(s0, aux["sdist0"]) < 1e-6
    s1 = torch.cat([rsl[1].spacing_starts[..., 0], rsl[1].spacing_ends[:, -1:, 0]], dim=1).cpu()
    assert rel_err(s1, aux["sdist1"]) < 1e-4
    assert rel_err(wl[0][..., 0].cpu(), aux["weights0"]) < TOL
    t2 = torch.cat([rs.frustums.starts[..., 0], rs.frustums.ends[:, -1:, 0]], dim=1).cpu()
    assert rel_err(t2, aux["t_edges2"]) < 1e-3
    # euclidean edges of level 0 start at the training near plane 0.05 (NearFarCollider), moved inwards by the jitter
    assert spec.near_plane - 1e-6 <= float(rsl[0].frustums.starts.min()) < spec.near_plane + 0.01


def test_backward_chunks_follow_the_sampler_update_state_machine(model, spec, params_factory):
    """ray_samplers.py:598,622-623: with ``_steps_since_update > update_sched`` the FIRST backward chunk gives the
    proposal networks a gradient and resets the counter, the following chunks do not; with ``_step < 10`` every chunk
    does.  The model's chunk loop must equal per-chunk library calls with those flags."""
    from nerf_emitter_b200 import NerfEmitter
    from nerf_emitter_b200.nerfstudio_api import RayBundle
    N, C = 500, 200
    o3n, v3n, g = make_rays(N, seed=17)
    o, d = _nerf_rays(o3n, v3n)
    gout = torch.randn(N, 3, generator=g).to(DEV)
    jit = torch.rand(N, 3, generator=g).to(DEV)
    E = NerfEmitter(params_factory(0, 1e3), spec, device=torch.device(DEV)).train(True)
    old_chunk = model.backward_num_rays_per_chunk
    model.backward_num_rays_per_chunk = C
    model.train()
    try:
        for step, since, flags in [(100, 2, [True, False, False]), (3, 0, [True, True, True]),
                                   (100, 1, [False, False, False])]:
            model.zero_grad(set_to_none=True)
            model.proposal_sampler._step, model.proposal_sampler._steps_since_update = step, since
            rb = RayBundle(origins=o.to(DEV), directions=d.to(DEV), camera_indices=torch.full((N, 1), 1, device=DEV))
            model.backward_for_camera_ray_bundle(rb, gout, jitter=jit)
            assert model.proposal_sampler._steps_since_update == (0 if flags[0] else since)
            E.bind_grads()
            go, gv = torch.empty(N, 3, device=DEV), torch.empty(N, 3, device=DEV)
            for c, flag in enumerate(flags):
                a, b = c * C, min((c + 1) * C, N)
                go[a:b], gv[a:b] = E.backward(o[a:b].to(DEV), d[a:b].to(DEV), gout[a:b], camera_idx=1, jitter=jit[a:b],
                                              chunk_rays=C, proposal_requires_grad=flag, nerf_layout=True)
            assert rel_err(rb.origins.grad, go) < 1e-6 and rel_err(rb.directions.grad, gv) < 1e-6
            named = dict(model.named_parameters())
            for k, gref in E.grads.items():
                assert rel_err(named[k].grad, gref, floor=1e-12) < 1e-4, (step, k)
            prop = float(named["proposal_networks.0.encoding.hash_table"].grad.abs().sum())
            assert (prop > 0) == any(flags), (step, prop)
    finally:
        model.backward_num_rays_per_chunk = old_chunk
