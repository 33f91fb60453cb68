import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import torch
from helpers import make_rays, rel_err
from nerf_emitter_b200 import EmitterModelSpec, NerfEmitter, make_params
from oracle.nerfacto_oracle import OracleEmitter
from oracle.rotater_fixture import rotaters
spec = EmitterModelSpec()
P = make_params(spec, seed=0, table_scale=1e3)
o, v, _ = make_rays(160, seed=11)
def mi(g): return (2.0 * g.t())[[1, 2, 0]]
for use_rot in (False, True):
    for near in (0.0, 0.02):
        for cam in (0, 1):
            rot = rotaters()["sphere"] if use_rot else None
            E = NerfEmitter(P, spec, device=torch.device("cuda:0")).eval()
            E.set_appearance_mode(always_use_camera_indices=True)
            if rot: E.set_rotater(rot["transforms"], center=rot["center"], radius=float(rot["radius"]))
            O = OracleEmitter(P, spec, always_use_camera_indices=True, rotater=rot)
            aux = E.sample_levels(o.cuda(), v.cuda(), near=near, camera_idx=cam)
            O.edges_override = [None, aux["s_edges1"].cpu(), aux["s_edges2"].cpu()]
            ref = O.point_lights(o, v, near=near, camera_idx=cam)
            out = E.point_lights(o.cuda(), v.cuda(), near=near, camera_idx=cam)
            g, r = out["brightness_grad"].cpu(), mi(ref["brightness_grad"])
            pb = (g - r).abs().max(0).values / r.abs().max()
            print("rot=%s near=%.2f cam=%d: rgb %.2e grad rel %.2e, rays within 1e-3 %.1f%%, ratio median %.4f" % (
                use_rot, near, cam, rel_err(out["rgb"].cpu(), ref["rgb"]), rel_err(g, r), 100 * float((pb < 1e-3).float().mean()),
                float((g.norm(dim=0) / r.norm(dim=0)).median())))
