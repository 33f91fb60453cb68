"""nemit_forward_noalloc (no counterpart in the reference): the forward with
its scratch in a static device array of the library, for callers that cannot allocate device memory.  Called through
the C ABI and compared with the existing nemit_forward on two calls in a row (the second with fewer, different rays, so
stale contents of the shared scratch would show).  Collected last on purpose (file name)."""
import ctypes as C

import pytest
import torch

from helpers import make_rays, rel_err

pytestmark = pytest.mark.gpu
TOL = 1e-3      # the fp32-mode gate of the other parity tests; the two entries launch the same kernels, so 0 is expected


def _call_noalloc(E, L, o, v, rgb, near, camera_idx):
    R = L.NemitRays()
    R.o, R.v, R.n = o.data_ptr(), v.data_ptr(), o.shape[1]
    R.near_per_ray, R.near_scalar, R.camera_idx, R.training = None, float(near), int(camera_idx), int(E.training)
    R.jitter, R.anneal = None, float(E.anneal)
    return E.lib.nemit_forward_noalloc(E._h, C.byref(R), C.c_void_p(rgb.data_ptr()), E._stream())


@pytest.mark.parametrize("precision,sampler", [("fp32", "exact"), ("fp16", "fast"), ("split", "exact")])
def test_forward_noalloc_matches_forward_twice(spec, params_factory, precision, sampler):
    from nerf_emitter_b200 import NerfEmitter
    from nerf_emitter_b200 import lib as L
    E = NerfEmitter(params_factory(0, 1e3), spec, device=torch.device("cuda:0"))
    E.set_precision(precision).set_sampler(sampler)
    E.train(False)
    for call, (n, seed, near, cam) in enumerate([(5000, 21, 0.02, 1), (777, 22, 0.0, 0)]):
        o, v, _ = make_rays(n, seed=seed)
        o, v = o.cuda(), v.cuda()
        ref = E.render(o, v, near=near, camera_idx=cam)
        out = torch.full((n, 3), float("nan"), device="cuda")
        torch.cuda.synchronize()
        free0 = torch.cuda.mem_get_info()[0]
        alloc0 = torch.cuda.memory_allocated()
        rc = _call_noalloc(E, L, o, v, out, near, cam)
        torch.cuda.synchronize()
        assert rc == 0, E.lib.nemit_last_error()
        assert torch.cuda.memory_allocated() == alloc0           # nothing went through torch's allocator
        if call == 1:                                           # (the first call may still load the module's static data)
            assert torch.cuda.mem_get_info()[0] >= free0        # nor through the driver
        err = rel_err(out, ref)
        print("call %d (%s/%s, %d rays): rel err vs nemit_forward %.2e, bit-equal %s" % (call, precision, sampler, n, err,
                                                                                       torch.equal(out, ref)))
        assert torch.isfinite(out).all() and err < TOL
    # the python wrapper goes through the same entry
    out2 = E.render_noalloc(o, v, torch.empty(n, 3, device="cuda"), near=near, camera_idx=cam)
    assert rel_err(out2, ref) < TOL


def test_forward_noalloc_rejects_oversize_and_null(spec, params_factory):
    from nerf_emitter_b200 import NerfEmitter
    from nerf_emitter_b200 import lib as L
    E = NerfEmitter(params_factory(0, 1e3), spec, device=torch.device("cuda:0"))
    E.train(False)
    n = L.NOALLOC_MAX_RAYS + 1
    o, v, _ = make_rays(n, seed=5)
    o, v = o.cuda(), v.cuda()
    out = torch.zeros(n, 3, device="cuda")
    rc = _call_noalloc(E, L, o, v, out, 0.0, 0)
    assert rc == -1 and b"NEMIT_NOALLOC_MAX_RAYS" in E.lib.nemit_last_error()
    assert float(out.abs().sum()) == 0.0                 # nothing ran
    # exactly the maximum is served
    o, v = o[:, :L.NOALLOC_MAX_RAYS].contiguous(), v[:, :L.NOALLOC_MAX_RAYS].contiguous()
    out = torch.empty(L.NOALLOC_MAX_RAYS, 3, device="cuda")
    assert _call_noalloc(E, L, o, v, out, 0.0, 0) == 0
    torch.cuda.synchronize()
    assert rel_err(out, E.render(o, v)) < TOL
    R = L.NemitRays()
    R.o, R.v, R.n, R.anneal = o.data_ptr(), v.data_ptr(), 8, 1.0
    assert E.lib.nemit_forward_noalloc(E._h, C.byref(R), None, None) == -1
