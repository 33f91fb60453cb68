"""GPU: the CustomOp-level drop-in (nerf_emitter_b200/op.py = DS/emitters/nerf_emitter_op.py:18-117): call signature,
(3,N) in / flat (N*3) out, the ``near`` change of variables, the zero gradient of the dummy parameter, the PathGuiding
hooks and the sampler's per-chunk ``updated`` state machine.  The kernels underneath are pinned elsewhere; here the
op is compared with direct library calls and with the oracle's op-level methods (tolerance 1e-3, north_star)."""
import pytest
import torch

from helpers import make_rays, rel_err

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
TOL = 1e-3


class _Guiding:
    def __init__(self):
        self.calls = []

    def train_primal(self, o, v, rgb, w):
        self.calls.append(("primal", tuple(o.shape), tuple(rgb.shape)))

    def train_forward(self, *a):
        self.calls.append(("forward",))

    def train_backward(self, *a):
        self.calls.append(("backward",))


def test_op_signature_layouts_and_gradients(spec, params_factory):
    from nerf_emitter_b200 import NerfEmitter
    from nerf_emitter_b200.op import make_nerf_emitter_op
    from oracle.nerfacto_oracle import OracleEmitter
    P = params_factory(0, 1.0)
    E = NerfEmitter(P, spec, device=torch.device(DEV)).eval()
    O = OracleEmitter(P, spec)
    N = 257
    o, v, g = make_rays(N, seed=23)
    near = torch.rand(N, generator=g) * 0.3
    guiding = _Guiding()
    op = make_nerf_emitter_op(E, path_guiding=guiding)
    dummy = torch.full((1,), float("nan"), device=DEV, requires_grad=True)       # nerf.py:37
    od, vd = o.to(DEV).requires_grad_(True), v.to(DEV).requires_grad_(True)
    out = op(od, vd, near.to(DEV), dummy, torch.ones(N, device=DEV), 1)
    assert tuple(out.shape) == (N * 3,)                                              # nerf_emitter_op.py:35-36
    ref = O.render(o, v, near=near[:, None], camera_idx=1)
    assert rel_err(out.detach().cpu().reshape(N, 3), ref) < TOL
    assert guiding.calls == [("primal", (N, 3), (N, 3))]
    gout = torch.randn(N, 3, generator=g)
    E.bind_grads()
    out.backward(gout.reshape(-1).to(DEV))
    # derivatives at the CUDA path's own sample edges: end to end a sampler tie moves single samples of ~1 % of the
    # rays (here 2 of 257), which is gated in tests/test_gpu_backward.py
    aux = E.sample_levels(o.to(DEV), v.to(DEV), near=near.to(DEV), camera_idx=1)
    O.edges_override = [None, aux["s_edges1"].cpu(), aux["s_edges2"].cpu()]
    ref_go, ref_gv, ref_grads, _ = O.vjp(o, v, gout, near=near[:, None], camera_idx=1, proposal_requires_grad=True)
    assert tuple(od.grad.shape) == (3, N) and rel_err(od.grad.cpu(), ref_go) < TOL
    assert rel_err(vd.grad.cpu(), ref_gv) < TOL
    assert float(dummy.grad.abs().sum()) == 0.0                                     # nerf_emitter_op.py:116
    k = "field.mlp_head.layers.2.weight"
    assert rel_err(E.grads[k].cpu(), ref_grads[k]) < TOL


def test_op_follows_the_sampler_state_machine(spec, params_factory):
    """(3,N) layout, 500 rays in chunks of 200, training mode: the op with the sampler object equals library calls with
    the per-chunk flags the reference's state machine gives (first chunk updated, rest not)."""
    from nerf_emitter_b200 import NerfEmitter
    from nerf_emitter_b200.nerfstudio_api import ProposalNetworkSampler
    from nerf_emitter_b200.op import make_nerf_emitter_op
    P = params_factory(0, 1e3)
    E = NerfEmitter(P, spec, device=torch.device(DEV)).train(True)
    N, C = 500, 200
    o, v, g = make_rays(N, seed=29)
    jit = torch.rand(N, 3, generator=g).to(DEV)
    gout = torch.randn(N, 3, generator=g).to(DEV)
    sampler = ProposalNetworkSampler(None, update_sched=lambda step: 1)
    sampler._step, sampler._steps_since_update = 50, 2
    op = make_nerf_emitter_op(E, jitter_fn=lambda n: jit, chunk_rays=C, sampler=sampler)
    od, vd = o.to(DEV).requires_grad_(True), v.to(DEV).requires_grad_(True)
    out = op(od, vd, 0.0, None, torch.ones(N, device=DEV), 2)
    assert sampler._steps_since_update == 0                  # the primal's first chunk consumed the update
    sampler._steps_since_update = 2                          # as after two more trainer steps
    E.bind_grads()
    out.backward(gout.reshape(-1))
    got = {k: t.clone() for k, t in E.grads.items()}
    got_o, got_v = od.grad.clone(), vd.grad.clone()
    E.bind_grads()
    go, gv = torch.empty(3, N, device=DEV), torch.empty(3, N, device=DEV)
    for c, flag in enumerate([True, False, False]):
        a, b = c * C, min((c + 1) * C, N)
        go[:, a:b], gv[:, a:b] = E.backward(o[:, a:b].contiguous().to(DEV), v[:, a:b].contiguous().to(DEV), gout[a:b],
                                            camera_idx=2, jitter=jit[a:b], chunk_rays=C, proposal_requires_grad=flag)
    assert rel_err(got_o, go) < 1e-6 and rel_err(got_v, gv) < 1e-6
    for k, t in E.grads.items():
        assert rel_err(got[k], t, floor=1e-12) < 1e-4, k
    assert float(got["proposal_networks.0.encoding.hash_table"].abs().sum()) > 0
    with pytest.raises(ValueError):
        E.backward(o.to(DEV), v.to(DEV), gout, chunk_rays=C, proposal_requires_grad=[True, False])
