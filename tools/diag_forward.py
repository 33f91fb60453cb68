"""Diagnostic (GPU box): prints parity numbers of the CUDA forward against the goldens / oracle."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import numpy as np, torch
from helpers import GOLDEN_CASES, golden_inputs, load_golden, rel_err
from nerf_emitter_b200 import EmitterModelSpec, NerfEmitter, make_params
spec = EmitterModelSpec()
for case in GOLDEN_CASES:
    G = load_golden(case); I = golden_inputs(G)
    E = NerfEmitter(make_params(spec, int(G["param_seed"]), float(G["table_scale"])), spec)
    E.train(I["training"])
    rgb, aux = E.render(I["o"].cuda(), I["v"].cuda(), near=I["near"], camera_idx=I["camera_idx"], jitter=I["jitter"], return_aux=True)
    torch.cuda.synchronize()
    print("==", case)
    for k, gk in [("weights0","weights0"),("weights1","weights1"),("weights2","weights2"),("s_edges1","sdist1"),("s_edges2","sdist2"),("rgb","rgb"),("accumulation","accumulation")]:
        ref = G[gk].reshape(aux[k].shape)
        print("  %-12s rel %.3e" % (k, rel_err(aux[k].cpu(), ref)))
    for k in ("inds1","inds2"):
        a = aux[k].cpu().numpy(); b = G[k]
        bad = a != b
        print("  %s mismatches %d / %d ; rays affected %d ; max |diff| %d" % (k, bad.sum(), b.size, bad.any(1).sum(), np.abs(a-b).max()))
        if bad.any():
            r = np.argwhere(bad)[0]
            print("   first at ray %d k %d gpu %d ref %d" % (r[0], r[1], a[r[0], r[1]], b[r[0], r[1]]))
            cdf = G["cdf1" if k=="inds1" else "cdf2"][r[0]]
            print("   cdf around:", cdf[max(0,b[r[0],r[1]]-2): b[r[0],r[1]]+2])
            w = aux["weights0" if k=="inds1" else "weights1"].cpu().numpy()[r[0]]
            wr = G["weights0" if k=="inds1" else "weights1"][r[0]]
            print("   weights row max abs diff %.3e (max %.3e)" % (np.abs(w-wr).max(), np.abs(wr).max()))
