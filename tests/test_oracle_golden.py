"""CPU: the standalone oracle (oracle/nerfacto_oracle.py) against the golden vectors that
oracle/make_golden.py produced from the REFERENCE's own classes.  Pins the oracle on any box."""
import numpy as np
import pytest
import torch

from helpers import GOLDEN_CASES, T, golden_inputs, load_golden, rel_err
from oracle.nerfacto_oracle import OracleEmitter, sh16


@pytest.mark.parametrize("case", GOLDEN_CASES)
def test_forward_and_jvp_match_reference_golden(case, spec, params_factory):
    G = load_golden(case)
    I = golden_inputs(G)
    O = OracleEmitter(params_factory(int(G["param_seed"]), float(G["table_scale"])), spec)
    O.train(I["training"])
    rgb, aux = O.render(I["o"], I["v"], near=I["near"], camera_idx=I["camera_idx"], jitter=I["jitter"],
                        return_aux=True)
    # index work: bit-exact
    assert np.array_equal(aux["inds1"].numpy(), G["inds1"])
    assert np.array_equal(aux["inds2"].numpy(), G["inds2"])
    # the restatement keeps the reference's op order: floats agree to the last bit on the same CPU,
    # and to 1e-6 relative across CPUs with a different ATen vector width
    for k in ["sdist0", "sdist1", "sdist2", "weights0", "weights1", "weights2", "rgb", "accumulation",
              "density2", "rgb_samples", "cdf1", "cdf2"]:
        assert rel_err(aux[k], G[k]) < 1e-6, k
    tangent = O.jvp(I["o"], I["v"], I["do"], I["dv"], near=I["near"], camera_idx=I["camera_idx"],
                    jitter=I["jitter"])
    assert rel_err(tangent, G["jvp"]) < 1e-5


@pytest.mark.parametrize("case", ["train_stress", "eval_stress"])
@pytest.mark.parametrize("prg", [True, False])
def test_vjp_matches_reference_golden(case, prg, spec, params_factory):
    G = load_golden(case)
    I = golden_inputs(G)
    O = OracleEmitter(params_factory(int(G["param_seed"]), float(G["table_scale"])), spec)
    O.train(I["training"])
    go, gv, grads, _ = O.vjp(I["o"], I["v"], I["grad_out"], near=I["near"], camera_idx=I["camera_idx"],
                             jitter=I["jitter"], proposal_requires_grad=prg)
    tag = "vjp_prop%d::" % int(prg)
    assert rel_err(go, G[tag + "grad_o"]) < 1e-5
    assert rel_err(gv, G[tag + "grad_v"]) < 1e-5
    for name, g in grads.items():
        if name.endswith("hash_table"):
            flat = g.reshape(-1)
            assert int(torch.count_nonzero(flat)) == int(G[tag + name + "::nnz"]), name
            assert rel_err(flat[T(G[tag + name + "::pick_idx"])], G[tag + name + "::pick_val"], floor=1e-30) < 1e-4, name
            ref_abs = float(G[tag + name + "::abssum"])
            assert abs(float(flat.double().abs().sum()) - ref_abs) <= 1e-5 * max(ref_abs, 1e-30), name
        else:
            assert rel_err(g, G[tag + name], floor=1e-30) < 1e-4, name


def test_known_answer_get_positions():
    """The reference's one exact known-answer test on the path, tests/cameras/test_rays.py:11-30:
    o=(0,1,2), d=(0,1,0), bin [2,3] -> position (0,3.5,2)."""
    o = torch.tensor([[0., 1., 2.]])
    d = torch.tensor([[0., 1., 0.]])
    t0, t1 = torch.tensor([[[2.]]]), torch.tensor([[[3.]]])
    pos = o[:, None, :] + d[:, None, :] * (t0 + t1) / 2
    assert torch.allclose(pos[0, 0], torch.tensor([0., 3.5, 2.]))


def test_known_answer_sh_orthonormal():
    """tests/utils/test_math.py:7-16 of the reference: (sh^T sh)/N * 4pi ~= I, atol 1e-2."""
    torch.manual_seed(0)
    N = 1000000
    d = torch.randn(N, 3)
    d = d / d.norm(dim=-1, keepdim=True)
    sh = sh16(d)
    gram = (sh.t() @ sh) / N * 4 * torch.pi
    assert torch.allclose(gram, torch.eye(16), atol=1e-2)


@pytest.mark.parametrize("case", ["point_lights_stress", "point_lights_default_init"])
def test_point_lights_oracle_matches_reference_golden(case, spec, params_factory):
    """Light-point-cloud pass (sdf_nerfacto.py:189-212) against vectors frozen from the reference's own classes
    (oracle/make_golden_point_lights.py)."""
    import numpy as np
    from helpers import T, load_golden, rel_err
    from oracle.nerfacto_oracle import OracleEmitter
    G = load_golden(case)
    O = OracleEmitter(params_factory(int(G["param_seed"]), float(G["table_scale"])), spec)
    out = O.point_lights(T(G["o"]), T(G["v"]))
    assert rel_err(out["rgb"], G["rgb"]) < 1e-6
    assert np.array_equal(out["depth"].numpy(), G["depth"])
    assert rel_err(out["brightness_grad"], G["brightness_grad"]) < 1e-5
