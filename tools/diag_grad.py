"""Diagnostic (GPU box): per-ray error of jvp / backward against the goldens."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import numpy as np, torch
from helpers import GOLDEN_CASES, golden_inputs, load_golden, rel_err
from nerf_emitter_b200 import EmitterModelSpec, NerfEmitter, make_params
spec = EmitterModelSpec()
np.set_printoptions(precision=4, linewidth=200)
for case in GOLDEN_CASES[:2]:
    G = load_golden(case); I = golden_inputs(G)
    E = NerfEmitter(make_params(spec, int(G["param_seed"]), float(G["table_scale"])), spec)
    E.train(I["training"])
    t = E.jvp(I["o"].cuda(), I["v"].cuda(), I["do"].cuda(), I["dv"].cuda(), near=I["near"], camera_idx=I["camera_idx"], jitter=I["jitter"]).cpu().numpy()
    ref = G["jvp"]
    err = np.abs(t - ref).max(1)
    mag = np.abs(ref).max()
    order = np.argsort(-err)
    print("==", case, "jvp max|ref| %.3e; rays with err > 1e-3*max: %d / %d" % (mag, (err > 1e-3 * mag).sum(), len(err)))
    w2 = G["weights2"]; dens = G["density2"]
    for r in order[:6]:
        print(" ray %d err %.3e ours %s ref %s | acc %.3f max w %.3f argmax %d  dens max %.3e" % (r, err[r], t[r], ref[r], w2[r].sum(), w2[r].max(), w2[r].argmax(), dens[r].max()))
    E.bind_grads()
    go, gv = E.backward(I["o"].cuda(), I["v"].cuda(), I["grad_out"].cuda(), near=I["near"], camera_idx=I["camera_idx"], jitter=I["jitter"], chunk_rays=I["N"], proposal_requires_grad=True)
    pre = "vjp_prop1::"
    for nm, a in [("grad_o", go), ("grad_v", gv)]:
        a = a.cpu().numpy(); b = G[pre + nm]
        e = np.abs(a - b).max(0); m = np.abs(b).max()
        print("  %s rel %.3e ; rays off: %d ; worst rays %s" % (nm, e.max() / m, (e > 1e-3 * m).sum(), np.argsort(-e)[:6]))
    for k, g in E.grads.items():
        if k.endswith("hash_table"):
            flat = g.reshape(-1).double().cpu()
            print("  %-55s abssum ours %.6e ref %.6e nnz %d ref %d" % (k, flat.abs().sum(), G[pre + k + "::abssum"], (flat != 0).sum(), G[pre + k + "::nnz"]))
            idx = G[pre + k + "::pick_idx"]; val = G[pre + k + "::pick_val"]
            if len(idx):
                print("      picked rel %.3e" % rel_err(flat[idx].float(), val))
        else:
            print("  %-55s rel %.3e" % (k, rel_err(g.cpu(), G[pre + k].reshape(g.shape))))
