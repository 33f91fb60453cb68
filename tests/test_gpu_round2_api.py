"""GPU tests of the boundary pieces added in round 2: kept sample edges (the adjoint of a CustomOp instance re-uses the
primal's samples), the host-buffer reverse-mode entry, the in-library all-reduce with the folded scale, the
per-kernel timing records (no launch dropped, rays per launch recorded) and the ceiling microbenchmarks."""
import ctypes as C

import pytest
import torch

from helpers import make_rays, rel_err

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def E(spec, params_factory):
    from nerf_emitter_b200 import NerfEmitter
    return NerfEmitter(params_factory(0, 1e3), spec, device=torch.device("cuda:0"))


@pytest.mark.parametrize("precision,sampler", [("fp32", "exact"), ("fp16", "fast"), ("split", "exact")])
@pytest.mark.parametrize("prg", [False, True])
def test_backward_with_kept_edges_equals_resampling(E, precision, sampler, prg):
    """base_model.py:299,327-328: the reference's adjoint re-derives the primal's samples from restored generator state;
    here the primal hands its edges over.  Gradients must be equal up to the order of the atomic accumulation."""
    E.set_precision(precision).set_sampler(sampler)
    E.train(True)
    N = 3000
    o, v, g = make_rays(N, seed=5)
    jit = torch.rand(N, 3, generator=g).cuda()
    gout = torch.randn(N, 3, generator=g).cuda()
    o, v = o.cuda(), v.cuda()
    E.bind_grads()
    go0, gv0 = E.backward(o, v, gout, near=0.03, camera_idx=1, jitter=jit, chunk_rays=1024, proposal_requires_grad=prg)
    torch.cuda.synchronize()
    ref = {k: t.clone() for k, t in E.grads.items()}
    rgb, kept = E.render(o, v, near=0.03, camera_idx=1, jitter=jit, keep_edges=True)
    launches = E.launch_count()
    E.bind_grads()
    go1, gv1 = E.backward(o, v, gout, near=0.03, camera_idx=1, jitter=jit, chunk_rays=1024, proposal_requires_grad=prg,
                          kept=kept)
    torch.cuda.synchronize()
    assert rel_err(go1, go0) < 1e-5 and rel_err(gv1, gv0) < 1e-5      # per-ray sums are atomic accumulations too
    for k, t in E.grads.items():
        assert rel_err(t, ref[k]) < 1e-5, k
    # one launch fewer per sub-chunk: no k_sample_levels
    E.profile(True)
    E.profile_read(reset=True)
    E.backward(o, v, gout, near=0.03, camera_idx=1, jitter=jit, chunk_rays=1024, proposal_requires_grad=prg, kept=kept)
    prof = E.profile_read(reset=True)
    E.profile(False)
    assert "k_sample_levels" not in prof and launches > 0
    # a mode change invalidates the kept edges: the call samples again instead of using stale ones
    E.train(False)
    E.profile(True)
    E.backward(o, v, gout, near=0.03, camera_idx=1, chunk_rays=1024, proposal_requires_grad=prg, kept=kept)
    prof = E.profile_read(reset=True)
    E.profile(False)
    assert "k_sample_levels" in prof
    E.set_precision("fp32").set_sampler("exact")


def test_kept_edges_argument_errors(E):
    from nerf_emitter_b200 import lib as L
    o, v, _ = make_rays(8)
    R = E._rays(o.cuda(), v.cuda(), 0.0, 0, None)
    t = torch.zeros(8, 49, device="cuda")
    R.kept_t_edges2 = t.data_ptr()          # without kept_s_edges2
    rgb = torch.zeros(8, 3, device="cuda")
    ws = E._workspace(8)
    rc = E.lib.nemit_forward(E._h, C.byref(R), rgb.data_ptr(), None, None, ws.data_ptr(), ws.numel(), None)
    assert rc == -1 and b"kept" in E.lib.nemit_last_error()


@pytest.mark.parametrize("primal_training", [False, True])
def test_backward_host_equals_device_path(E, primal_training):
    E.set_precision("fp32").set_sampler("exact")
    N = 2500
    o, v, g = make_rays(N, seed=9)
    gout = torch.randn(N, 3, generator=g)
    E.train(primal_training)
    rgb_d = E.render(o.cuda(), v.cuda(), near=0.02).cpu()
    E.train(True)
    E.bind_grads()
    go_d, gv_d = E.backward(o.cuda(), v.cuda(), gout.cuda(), near=0.02, chunk_rays=1000, loss_scale=2.0,
                            proposal_requires_grad=True)
    torch.cuda.synchronize()
    ref = {k: t.clone() for k, t in E.grads.items()}
    E.bind_grads()
    pin = lambda t: t.contiguous().pin_memory()
    rgb_h, go_h, gv_h = pin(torch.empty(N, 3)), pin(torch.empty(3, N)), pin(torch.empty(3, N))
    E.backward_host(pin(o), pin(v), pin(gout), rgb_h, go_h, gv_h, near=0.02, primal_training=primal_training,
                    chunk_rays=1000, loss_scale=2.0, proposal_requires_grad=True)
    torch.cuda.synchronize()
    assert torch.equal(rgb_h, rgb_d)
    assert rel_err(go_h, go_d.cpu()) < 1e-5 and rel_err(gv_h, gv_d.cpu()) < 1e-5
    for k, t in E.grads.items():
        assert rel_err(t, ref[k]) < 1e-5, k


def test_allreduce_single_rank_applies_the_scale(E):
    E.bind_grads()
    E.flat_grads.copy_(torch.arange(E.flat_grads.numel(), device="cuda", dtype=torch.float32) % 1000)
    before = E.flat_grads.clone()
    E.allreduce_grads(scale=0.25)
    torch.cuda.synchronize()
    assert torch.equal(E.flat_grads, before * 0.25)
    E.allreduce_grads()                      # no scale, one rank: untouched
    torch.cuda.synchronize()
    assert torch.equal(E.flat_grads, before * 0.25)
    E.zero_grads()


def test_nccl_comm_of_one_rank(E):
    """The library's own communicator (ncclGetUniqueId / ncclCommInitRank through dlopen) with world size 1; the scaled
    sum of one rank is served by the library's own pass (see comm.cu), odd element count included."""
    from nerf_emitter_b200 import lib as L
    ident = (C.c_ubyte * L.UNIQUE_ID_BYTES)()
    L.check(E.lib.nemit_comm_unique_id(ident), "nemit_comm_unique_id")
    L.check(E.lib.nemit_comm_init(E._h, ident, 0, 1), "nemit_comm_init")
    x = torch.arange(100003, device="cuda", dtype=torch.float32)
    L.check(E.lib.nemit_allreduce_grads(E._h, x.data_ptr(), x.numel(), 0.5, E._stream()), "nemit_allreduce_grads")
    torch.cuda.synchronize()
    assert torch.equal(x, torch.arange(100003, device="cuda", dtype=torch.float32) * 0.5)
    L.check(E.lib.nemit_comm_destroy(E._h), "nemit_comm_destroy")


def test_profile_records_every_launch_and_its_rays(E):
    """the 512-event cap truncated the records.  5 000 tiny forward calls = 10 000 launches, more than the
    4 096-entry ring: every one must be counted, with the rays it processed."""
    E.set_precision("fp32").set_sampler("exact")
    E.train(False)
    o, v, _ = make_rays(7)
    o, v = o.cuda(), v.cuda()
    E.profile(True)
    E.profile_read(reset=True)
    l0 = E.launch_count()
    calls = 5000
    for _ in range(calls):
        E.render(o, v)
    prof = E.profile_read(reset=True, with_rays=True)
    E.profile(False)
    assert E.launch_count() - l0 == 2 * calls
    assert prof["k_sample_levels"][1] == calls and prof["k_field_forward"][1] == calls
    assert prof["k_sample_levels"][2] == 7 * calls and prof["k_field_forward"][2] == 7 * calls
    assert prof["k_sample_levels"][0] > 0.0


def test_red_lane_ops_are_counted(E):
    E.set_precision("fp16").set_sampler("fast")
    E.train(True)
    N = 4096
    o, v, g = make_rays(N, seed=3)
    gout = torch.randn(N, 3, generator=g).cuda()
    E.bind_grads()
    E.profile(True)
    E.profile_read(reset=True)
    E.backward(o.cuda(), v.cuda(), gout, chunk_rays=N, proposal_requires_grad=False)
    prof = E.profile_read(reset=True, with_rays=True)
    E.profile(False)
    ops = prof["red_lane_ops"]
    # between 4 (all pairs) and 8 (no pair) lane-ops per level of every sample with a non-zero gradient
    assert 0 < ops <= N * 48 * 16 * 8
    nz = int((E.grads["field.mlp_base_grid.hash_table"] != 0).any(dim=1).sum())
    assert ops >= nz / 2                     # a reduction lane-op touches one table row, or an aligned pair of rows
    E.set_precision("fp32").set_sampler("exact")


def test_microbench_ceilings(E):
    c = E.microbench()
    print({k: "%.1f G/s" % (v / 1e9) for k, v in c.items()})
    assert set(c) == {"gather_l1miss", "gather_l1hit", "gather_coherent", "red", "l2_read"}
    assert c["gather_l1hit"] > c["gather_l1miss"] > 50e9
    assert c["gather_coherent"] > c["gather_l1miss"]
    assert c["red"] > 50e9 and c["l2_read"] > 2e12
