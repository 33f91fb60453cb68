#!/usr/bin/env python
"""bench.py - emitter rays/sec on N B200s (driver contract; see DESIGN.md "Measurement").

Default line (no flags): a step = one pass of the emitter hot path over one batch of synthetic rays resident in HBM:
eval-mode `nemit_forward` + training-mode `nemit_backward` on `--rays` rays per GPU (2^20; BASELINE configs[1] shape:
1 M rays, random-init weights, deterministic unjittered rays, fp16 tensor-core MLP).  Rays shard over ranks with
replicated weights (weak scaling); no data-path collective in the forward, one gradient all-reduce in the backward
(`nemit_allreduce_grads`, NCCL inside the library).  The same JSON line also carries the forward-only rate
(`forward`), the CustomOp-instance variant whose adjoint re-uses the primal's sample edges (`train_primal_backward`),
the per-kernel roofline against re-measured ceilings, and - under --gpus N > 1 - an on-hardware correctness `check`
of the sharded backward + all-reduce against one rank recomputing every shard.

  python bench.py [--gpus N --steps K --warmup W] [--impl reference] [--config {2,3,4,5}]

--config selects one of BASELINE.json's other configurations instead:
  2  forward only, 2^20 rays per GPU (weak)           3  fwd+bwd, 512*512*64 = 16 777 216 rays split over the ranks (strong)
  4  vMF mixture: 64 lobes x 16 M samples -> sample_direction -> pdf_direction -> NeRF radiance, split over the ranks (strong)
  5  spherical-sensor environment-map bake 2048 x 1024 x 16 spp, film rows split over the ranks, films summed (strong)
"""
from __future__ import annotations

import argparse
import contextlib
import io
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import torch  # noqa: E402

# ---- algorithmic work per ray ---------------------------------------------------
LOADS_SAMPLING = (256 + 96) * 5 * 8                 # 8-byte table lane-loads of the two proposal levels per ray: 14 080
LOADS_FIELD = 48 * 16 * 8                           # main-grid lane-loads per ray: 6 144
MISS_FIELD = 48 * 11 * 4                            # of which compulsory L1 misses: 11 hashed levels x 4 (y,z) pairs: 2 112
ALG_BYTES_SAMPLING = LOADS_SAMPLING * 8             # 112 640 B
ALG_BYTES_FIELD = LOADS_FIELD * 8                   # 49 152 B
SEED_KEEP_BYTES = 48 * (64 + 192 + 192 + 16)        # per ray: encoding, Jacobian, activations, seeds kept for the backward
ALG_BYTES = {"k_sample_levels": ALG_BYTES_SAMPLING, "k_field_forward": ALG_BYTES_FIELD,
             "k_field_forward_tc": ALG_BYTES_FIELD,
             "k_field_forward<adjoint_seed>": ALG_BYTES_FIELD + SEED_KEEP_BYTES,   # fp16 mode; fp32 mode keeps 16 B/sample
             "k_field_backward": 3 * ALG_BYTES_FIELD,      # recompute gathers + d/dx gathers + float2 red.add
             "k_field_backward_tc": ALG_BYTES_FIELD + SEED_KEEP_BYTES,   # float2 red.add + streamed tiles, no gathers
             "k_prop_backward": 2 * ALG_BYTES_SAMPLING, "k_field_jvp": ALG_BYTES_FIELD,
             "k_finalize_ray_grads": 24 + 24}
# which measured ceiling a kernel is held against: (bound, ceiling key of nemit_microbench, operations per ray)
OWN_BOUND = {"k_sample_levels": ("l1_gather", "gather_coherent", LOADS_SAMPLING,
                                 "ray-coherent 8-byte lane-loads (a warp = 32 consecutive samples of one ray) vs the "
                                 "warp-coherent gather rate; the kernel is issue-bound before it is gather-bound (ncu)"),
             "k_field_forward_tc": ("l1_gather", "gather_l1miss", MISS_FIELD,
                                    "compulsory L1-missing lane-loads (11 hashed levels x 4 (y,z) corner pairs per "
                                    "sample) vs the random L1-missing lane-load rate; the 4 032 L1-hitting loads per ray "
                                    "ride on top"),
             "k_field_forward": ("l1_gather", "gather_l1miss", MISS_FIELD, "fp32 SIMT path: FMA-bound before gather-bound"),
             "k_field_forward<adjoint_seed>": ("l1_gather", "gather_l1miss", MISS_FIELD,
                                               "as the forward, plus the kept tiles leaving as streaming stores"),
             "k_field_jvp": ("l1_gather", "gather_l1miss", MISS_FIELD, "fp32 SIMT path"),
             "k_field_backward_tc": ("red_issue", "red", None,
                                     "reduction lane-ops the table scatter issued (counted on the device) vs the "
                                     "random red.global.add lane-op rate")}
# dram__bytes_read.sum + dram__bytes_write.sum per launch OF THE SIZE THE BENCH LAUNCHES, from the committed ncu --set full
# captures (profiles/r2_ncu_step_summary.txt): (rays per launch, bytes)
NCU_TRAFFIC = {"k_sample_levels": (1 << 18, 14.5e6), "k_field_forward_tc": (1 << 18, 111.6e6),
               "k_field_forward": (1 << 16, 63.9e6),
               "k_field_forward<adjoint_seed>": (1 << 16, 1.774e9), "k_field_backward_tc": (1 << 16, 1.584e9)}

METRIC = {"fwd": "NeRF-emitter rays/sec fwd (nerfacto 256/96/48 samples)",
          "fwd+bwd": "NeRF-emitter rays/sec fwd+bwd (nerfacto 256/96/48 samples)",
          "vmf": "vMF-mixture emitter samples/sec (64 lobes: sample_direction + pdf_direction + NeRF radiance of the sampled ray)",
          "bake": "NeRF-emitter rays/sec, spherical-sensor environment-map bake (2048x1024x16 spp, gaussian film)"}
CONFIG3_RAYS = 512 * 512 * 64
CONFIG4_SAMPLES = 1 << 24
CONFIG5_FILM = (2048, 1024, 16)


def make_rays(n, seed, device):
    g = torch.Generator().manual_seed(seed)
    o = torch.rand(3, n, generator=g)
    v = torch.randn(3, n, generator=g)
    v = v / v.norm(dim=0, keepdim=True)
    return o.to(device), v.to(device)


def make_lobes(K=64, seed=3):
    """SURVEY.md 8d config 4 synthetic lobe table: positions U([-1,2]^3), std U(0.05,0.5), weights ~ Dirichlet(1)."""
    g = torch.Generator().manual_seed(seed)
    position = torch.rand(K, 3, generator=g) * 3 - 1
    std = torch.rand(K, generator=g) * 0.45 + 0.05
    e = -torch.log(torch.rand(K, generator=g))
    return position, e / e.sum(), std


class ClockSampler:
    """nvidia-smi clocks + throttle reasons during the timed region (B200_PROFILING.md recipe)."""
    Q = ("index,clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.idx = gpu_index
        self.samples = []
        self.proc = None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.idx), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.samples.append([x.strip() for x in line.split(",")])

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        for s in self.samples:
            try:
                sm.append(float(s[1]))
                mx.append(float(s[2]))
            except Exception:
                continue
            for name, val in zip(["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"], s[3:7]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return float(d["hbm_gbs"]), "MEASURED_PEAKS.json hbm_gbs (burst copy)"
    return 6650.0, "fallback of B200_PROFILING.md (MEASURED_PEAKS.json absent)"


# ---------------------------------------------------------------------------------------------------------------------
# CPU side: the reference's torch-CPU algorithm (oracle port) on the host cores - cpu_baseline and --impl reference
# ---------------------------------------------------------------------------------------------------------------------
def _oracle_step_fn(kind, n, proposal_grad=False):
    """One bounded step of the workload `kind` on the CPU: returns (callable, units per call, description)."""
    from nerf_emitter_b200 import EmitterModelSpec, make_params
    from oracle.nerfacto_oracle import OracleEmitter
    torch.set_num_threads(os.cpu_count() or 1)
    spec = EmitterModelSpec()
    O = OracleEmitter(make_params(spec, seed=0, table_scale=1e3), spec)
    if kind in ("fwd", "fwd+bwd"):
        o, v = make_rays(n, 1, "cpu")
        g = torch.randn(n, 3)

        def step():
            O.train(False)
            O.render(o, v)
            if kind == "fwd+bwd":
                O.train(True)
                O.vjp(o, v, g, proposal_requires_grad=proposal_grad)
        return step, n, "%d rays per step of the same workload (oracle/nerfacto_oracle.py, torch CPU fp32)" % n
    if kind == "vmf":
        from oracle import vmf_oracle
        pos, w, std = make_lobes(64, 3)
        gen = torch.Generator().manual_seed(3)
        p, u = torch.rand(n, 3, generator=gen), torch.rand(n, 2, generator=gen)

        def step():
            d, pdf, _ = vmf_oracle.sample_direction(pos, w, std, p, u)
            vmf_oracle.mixture_pdf(pos, w, std, p, d)
            O.train(False)
            O.render(p.t().contiguous(), d.t().contiguous())
        return step, n, ("%d samples per step: oracle/vmf_oracle.py sample + pdf, then oracle/nerfacto_oracle.py radiance "
                         "(torch CPU fp32)" % n)
    if kind == "bake":
        import numpy as np
        from oracle import bake_oracle
        W, H, SPP = 64, 16, 4
        n = W * H * SPP
        rng = np.random.default_rng(0)

        def step():
            jit = rng.random((n, 2))
            o, d, pos = bake_oracle.spherical_rays(W, H, SPP, 0, n, jit)
            O.train(False)
            rgb = O.render(torch.from_numpy(o).float(), torch.from_numpy(d).float()).numpy()
            bake_oracle.film_develop(bake_oracle.film_accumulate(W, H, pos, rgb))
        return step, n, ("a %dx%dx%d spp film = %d rays per step: oracle/bake_oracle.py rays + film around "
                         "oracle/nerfacto_oracle.py (torch CPU fp32 / numpy)" % (W, H, SPP, n))
    raise ValueError(kind)


def cpu_reference_rate(kind, n, repeats, proposal_grad=False):
    step, units, what = _oracle_step_fn(kind, n, proposal_grad)
    step()
    best = 1e30
    for _ in range(repeats):
        t = time.perf_counter()
        step()
        best = min(best, time.perf_counter() - t)
    return {"value": units / best, "unit": UNIT[kind], "cores": torch.get_num_threads(), "kind": "port",
            "sample": "%s, best of %d (%.2f s each)" % (what, repeats, best)}


UNIT = {"fwd": "rays/s", "fwd+bwd": "rays/s", "vmf": "samples/s", "bake": "rays/s"}


def run_reference_arm(args, rank, kind):
    if rank != 0:
        return
    step, units, what = _oracle_step_fn(kind, args.ref_rays, args.proposal_grad)
    for _ in range(args.warmup):
        step()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        step()
    dt = (time.perf_counter() - t0) / args.steps
    val = units / dt
    line = {"impl": "reference", "metric": METRIC[kind], "value": val, "unit": UNIT[kind], "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt * 1e3, "higher_is_better": True,
            "scaling": SCALING[args.config], "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": workload_config(args, kind, units),
            "cpu_baseline": {"value": val, "unit": UNIT[kind], "cores": torch.get_num_threads(), "kind": "port",
                             "sample": what},
            "e2e": {"value": val, "unit": UNIT[kind], "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line))


SCALING = {None: "weak", 2: "weak", 3: "strong", 4: "strong", 5: "strong"}


def workload_config(args, kind, units_per_gpu):
    prec = args.precision
    sampler = args.sampler or ("fast" if prec == "fp16" else "exact")
    base = {"mode": kind, "mlp_precision": prec, "sampler": sampler, "baseline_config": args.config or "1/2 (default line)",
            "parallelism": "ray-sharded x%d, weights replicated" % args.gpus}
    if kind in ("fwd", "fwd+bwd"):
        total = CONFIG3_RAYS if args.config == 3 else None
        base["workload"] = (
            "nerfacto emitter query, %s, 256/96/48 samples, random-init stress weights (hash tables U(-1,1)), eval-mode "
            "forward%s" % (
                ("%d deterministic rays in total split over the ranks (BASELINE configs[2]: 512x512x64 spp)" % total)
                if total else ("%d deterministic rays per GPU" % units_per_gpu),
                (" + training-mode backward (field-param + ray grads, distortion loss, t^2 gradient scaling; proposal "
                 "networks %s)" % ("also get the interlevel-loss gradient" if args.proposal_grad else
                                   "carry no autograd on this call, as in the reference's steady state"))
                if kind == "fwd+bwd" else ""))
        base["rays_per_gpu"] = units_per_gpu
        base["l2_policy"] = "inputs+tables (%.0f MiB/step touched) exceed L2; no explicit flush" % (
            (units_per_gpu * (36 + 49 * 4 * 2) + 74 * 2 ** 20) / 2 ** 20)
    elif kind == "vmf":
        base["workload"] = ("vMF-mixture emitter, 64 lobes (positions U([-1,2]^3), std U(0.05,0.5), Dirichlet weights), "
                            "%d (p, u) samples in total split over the ranks: sample_direction -> pdf_direction -> NeRF "
                            "radiance of (p, d), 256/96/48 samples (BASELINE configs[3])" % CONFIG4_SAMPLES)
        base["samples_per_gpu"] = units_per_gpu
        base["l2_policy"] = "16 M samples x 56 B of inputs/outputs per step exceed L2; no explicit flush"
    else:
        base["workload"] = ("spherical-sensor environment-map bake %dx%dx%d spp from the scene centre, gaussian rfilter "
                            "stddev 0.25, film rows split over the ranks and the films summed (BASELINE configs[4])"
                            % CONFIG5_FILM)
        base["rays_per_gpu"] = units_per_gpu
        base["l2_policy"] = "33.5 M rays per bake exceed L2; no explicit flush"
    return base


# ---------------------------------------------------------------------------------------------------------------------
# GPU side
# ---------------------------------------------------------------------------------------------------------------------
class Timer:
    def __init__(self, dev, world, dist):
        self.dev, self.world, self.dist = dev, world, dist

    def barrier(self):
        torch.cuda.synchronize()
        if self.world > 1:
            self.dist.barrier()
            torch.cuda.synchronize()

    def run(self, step, steps, warmup):
        """ms per step over exactly `steps` steps after `warmup`, CUDA events, max over ranks."""
        for _ in range(warmup):
            step()
        self.barrier()
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
        ev[0].record()
        for _ in range(steps):
            step()
        ev[1].record()
        self.barrier()
        t = torch.tensor([ev[0].elapsed_time(ev[1])], device=self.dev, dtype=torch.float64)
        if self.world > 1:
            self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
        return float(t.item()) / steps


def kernel_roofline(prof, ceilings, ncu_traffic=True):
    """Per-kernel rates from the library's own CUDA-event records of the timed region (events on the launching stream
    around every launch; rays per launch as recorded at launch time) and the roofline object of the dominant kernel."""
    hbm, src = measured_peaks()
    red_ops = prof.pop("red_lane_ops", 0)
    kern = {}
    for k, (ms, c, rays) in prof.items():
        v = {"ms_total": ms, "launches": c, "rays": rays, "ms_per_launch": ms / c, "rays_per_launch": rays / c}
        sec = ms * 1e-3
        v["rays_per_s"] = rays / sec
        v["algorithmic_bytes_per_ray"] = ALG_BYTES[k]
        v["achieved_gbs"] = ALG_BYTES[k] * rays / sec / 1e9
        v["frac_of_hbm_peak"] = v["achieved_gbs"] / hbm
        if ncu_traffic and k in NCU_TRAFFIC:
            r0, b0 = NCU_TRAFFIC[k]
            v["traffic"] = b0 * v["rays_per_launch"] / r0
            v["traffic_over_algorithmic"] = v["traffic"] / (ALG_BYTES[k] * v["rays_per_launch"])
        if k in OWN_BOUND and ceilings:
            bound, key, per_ray, note = OWN_BOUND[k]
            ops = red_ops if per_ray is None else per_ray * rays
            v["own_roof"] = {"bound": bound, "achieved": ops / sec / 1e9, "peak": ceilings[key] / 1e9,
                             "unit": "G lane-ops/s", "frac": ops / sec / ceilings[key], "ceiling": key, "note": note}
            if per_ray is None:
                v["own_roof"]["lane_ops_per_ray"] = ops / max(rays, 1)
        kern[k] = v
    total = sum(v["ms_total"] for v in kern.values()) or 1.0
    for v in kern.values():
        v["share"] = v["ms_total"] / total
    dom = max(kern, key=lambda k: kern[k]["ms_total"])
    d = kern[dom]
    note = ("achieved = algorithmic bytes (8-byte table gathers / float2 reductions / streamed tiles) per launch / average "
            "CUDA-event launch duration in the timed region, rays per launch as recorded by the library.  ")
    if d["frac_of_hbm_peak"] > 1.0 or d.get("traffic_over_algorithmic", 1.0) < 0.5:
        note += ("The tables (74 MiB) are L2-resident: DRAM traffic (ncu) is %.3f of the algorithmic bytes, so a fraction "
                 "above 1 of the HBM roof is an L1/L2-served rate, not an HBM fraction; the kernel's own ceiling is in "
                 "kernels[...].own_roof." % d.get("traffic_over_algorithmic", float("nan")))
    roof = {"bound": "hbm", "kernel": dom, "achieved": d["achieved_gbs"], "peak": hbm, "peak_source": src, "unit": "GB/s",
            "frac": d["frac_of_hbm_peak"], "traffic": d.get("traffic"), "algorithmic_bytes_per_ray": ALG_BYTES[dom],
            "rays_per_launch": d["rays_per_launch"], "ms_per_launch": d["ms_per_launch"], "note": note,
            "own_roof": d.get("own_roof"), "kernels": kern}
    return roof


def sharded_backward_check(E, rank, world, dist, dev):
    """On-hardware correctness of the multi-GPU path: 8 192 rays split over the ranks, each rank's backward over its shard,
    then the library's all-reduce; rank 0 also recomputes every shard alone.  fp32 mode, exact sampler."""
    from nerf_emitter_b200 import parallel
    prec, samp = getattr(E, "precision", "fp32"), getattr(E, "sampler", "exact")
    E.set_precision("fp32").set_sampler("exact")
    n = 8192
    o, v = make_rays(n, 4242, dev)
    g = torch.Generator().manual_seed(4243)
    gout = torch.randn(n, 3, generator=g).to(dev)
    E.train(True)

    def shard_bwd(r):
        a, b = parallel.shard_range(n, r, world)
        return E.backward(o[:, a:b].contiguous(), v[:, a:b].contiguous(), gout[a:b].contiguous(), chunk_rays=b - a,
                          proposal_requires_grad=True)
    E.zero_grads()
    shard_bwd(rank)
    E.allreduce_grads()
    torch.cuda.synchronize()
    reduced = {k: t.clone() for k, t in E.grads.items()}
    out = None
    if rank == 0:
        E.zero_grads()
        for r in range(world):
            shard_bwd(r)
        torch.cuda.synchronize()
        worst, worst_name = 0.0, None
        for k, t in E.grads.items():
            den = float(t.abs().max())
            if den == 0.0:
                err = float(reduced[k].abs().max())
            else:
                err = float((reduced[k] - t).abs().max()) / den
            if err >= worst:
                worst, worst_name = err, k
        out = {"rays": n, "world": world, "mode": "fp32 / exact sampler, proposal networks included",
               "max_rel_err": worst, "worst_tensor": worst_name, "tolerance": 1e-5, "ok": bool(worst <= 1e-5),
               "what": "rank-sharded nemit_backward + nemit_allreduce_grads vs rank 0 recomputing all shards alone, "
                       "max-norm relative error over every gradient tensor"}
    # the folded scale of the collective (ncclRedOpCreatePreMulSum) on an odd-length blob vs torch.distributed + multiply
    x = (torch.arange(100003, device=dev, dtype=torch.float32) % 977) + rank
    y = x.clone()
    E.allreduce_blob(x, scale=0.5)
    dist.all_reduce(y)
    y *= 0.5
    torch.cuda.synchronize()
    if out is not None:
        out["premul_allreduce_max_rel_err"] = float((x - y).abs().max() / y.abs().max())
        out["ok"] = bool(out["ok"] and out["premul_allreduce_max_rel_err"] <= 1e-6)
    E.zero_grads()
    E.set_precision(prec).set_sampler(samp)
    return out


def bench_rays(args, E, T, rank, world, local, dist, dev, kind):
    from nerf_emitter_b200 import parallel
    strong = args.config == 3
    if strong:
        a, b = parallel.shard_range(CONFIG3_RAYS, rank, world)
        n = b - a
        total_units = CONFIG3_RAYS
    else:
        n = args.rays
        total_units = n * world
    o, v = make_rays(n, 1 + rank, dev)
    gout = torch.randn(n, 3, device=dev)
    if kind == "fwd+bwd":
        E.bind_grads()

    def step():
        E.train(False)
        rgb = E.render(o, v)
        if kind == "fwd+bwd":
            E.train(True)
            E.backward(o, v, gout, chunk_rays=16384, proposal_requires_grad=args.proposal_grad)
            if world > 1:
                E.allreduce_grads()          # the one collective: flat parameter-gradient blob over NVLink
        return rgb

    for _ in range(args.warmup):
        step()
    T.barrier()
    clocks = ClockSampler(local)
    if rank == 0:
        clocks.start()
    launches0 = E.launch_count()
    E.profile(True)
    E.profile_read(reset=True)
    ms_per_step = T.run(step, args.steps, 0)
    launches = E.launch_count() - launches0
    prof = E.profile_read(reset=True, with_rays=True)
    E.profile(False)
    clk = clocks.stop() if rank == 0 else None
    value = total_units / (ms_per_step * 1e-3)

    # e2e: the same step through ONE plugin call with pinned HOST buffers, copies inside the timed region
    oh, vh = o.cpu().pin_memory(), v.cpu().pin_memory()
    rh = torch.empty(n, 3).pin_memory()
    if kind == "fwd+bwd":
        gh = gout.cpu().pin_memory()
        goh, gvh = torch.empty(3, n).pin_memory(), torch.empty(3, n).pin_memory()
        h2d, d2h = 36 * n, 36 * n

        def e2e_step():
            E.train(True)
            E.backward_host(oh, vh, gh, rh, goh, gvh, primal_training=False, chunk_rays=16384,
                            proposal_requires_grad=args.proposal_grad)
            if world > 1:
                E.allreduce_grads()
    else:
        h2d, d2h = 24 * n, 12 * n

        def e2e_step():
            E.train(False)
            E.render_host(oh, vh, rh)
    e2e_steps = max(3, args.steps // 2)
    e2e_ms = T.run(e2e_step, e2e_steps, 2)
    e2e_val = total_units / (e2e_ms * 1e-3)

    extras = {}
    ceilings = None
    if world == 1:
        ceilings = E.microbench()
    if not strong and kind == "fwd+bwd":
        # forward-only (the north-star's own target metric) in the same run
        def fwd_step():
            E.train(False)
            E.render(o, v)
        f_ms = T.run(fwd_step, args.steps, 2)

        def fwd_e2e():
            E.train(False)
            E.render_host(oh, vh, rh)
        fe_ms = T.run(fwd_e2e, e2e_steps, 1)
        extras["forward"] = {"metric": METRIC["fwd"], "value": n * world / (f_ms * 1e-3), "unit": "rays/s",
                             "ms_per_step": f_ms, "steps": args.steps,
                             "e2e": {"value": n * world / (fe_ms * 1e-3), "unit": "rays/s", "h2d_bytes_per_step": 24 * n,
                                     "d2h_bytes_per_step": 12 * n, "through": "nemit_forward_host"}}
        # one CustomOp instance as the reference runs it during training: primal and adjoint in the SAME (training) mode, the
        # adjoint re-uses the primal's sample edges instead of sampling again (the reference re-derives identical samples
        # from restored generator state, NS/models/base_model.py:299,327-328)

        def kept_step():
            E.train(True)
            rgb, kept = E.render(o, v, keep_edges=True)
            E.backward(o, v, gout, chunk_rays=16384, proposal_requires_grad=args.proposal_grad, kept=kept)
            if world > 1:
                E.allreduce_grads()
        k_ms = T.run(kept_step, args.steps, 2)
        extras["train_primal_backward"] = {
            "workload": "training-mode primal + backward of one CustomOp instance; the backward takes the primal's kept "
                        "sample edges (780 B/ray) and skips the proposal sampling pass",
            "value": n * world / (k_ms * 1e-3), "unit": "rays/s", "ms_per_step": k_ms, "steps": args.steps}

    check = None
    if world > 1 and kind == "fwd+bwd":
        check = sharded_backward_check(E, rank, world, dist, dev)

    if rank != 0:
        return None
    line = {"metric": METRIC[kind], "value": value, "unit": "rays/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": SCALING[args.config],
            "vs_baseline": None, "dtype": "f16" if args.precision == "fp16" else "f32", "data": "synthetic",
            "config": workload_config(args, kind, n), "clocks": clk, "gpu_launches": int(launches),
            "e2e": {"value": e2e_val, "unit": "rays/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                    "through": "nemit_backward_host" if kind == "fwd+bwd" else "nemit_forward_host",
                    "note": "pinned host rays (+grad_out) in, host radiance (+ray grads) out, one C-ABI call per step, "
                            "copies on the compute stream"},
            "roofline": kernel_roofline(prof, ceilings, ncu_traffic=args.precision == "fp16")}
    if ceilings:
        line["ceilings"] = {k: v for k, v in ceilings.items()}
        line["ceilings"]["unit"] = "lane-ops/s (l2_read: bytes/s), nemit_microbench in this run"
    line.update(extras)
    if check is not None:
        line["check"] = check
    return line


def bench_vmf(args, E, T, rank, world, local, dist, dev):
    from nerf_emitter_b200 import VmfMixture, parallel
    a, b = parallel.shard_range(CONFIG4_SAMPLES, rank, world)
    n = b - a
    pos, w, std = make_lobes(64, 3)
    V = VmfMixture(pos.to(dev), w.to(dev), std.to(dev))
    g = torch.Generator(device=dev).manual_seed(3 + rank)
    p = torch.rand(n, 3, device=dev, generator=g)
    u = torch.rand(n, 2, device=dev, generator=g)
    E.train(False)

    def step():
        d, pdf = V.sample_direction(p, u)
        pdf2 = V.pdf_direction(p, d)
        rgb = E.render(p.t().contiguous(), d.t().contiguous())
        return d, pdf2, rgb
    clocks = ClockSampler(local)
    for _ in range(args.warmup):
        step()
    T.barrier()
    if rank == 0:
        clocks.start()
    launches0 = E.launch_count()
    ms = T.run(step, args.steps, 0)
    launches = E.launch_count() - launches0 + 4 * args.steps      # + sample, pdf and the two transposes
    clk = clocks.stop() if rank == 0 else None
    # e2e: pinned (p, u) in, (d, pdf, rgb) out
    ph, uh = p.cpu().pin_memory(), u.cpu().pin_memory()
    dh, pdfh, rh = torch.empty(n, 3).pin_memory(), torch.empty(n).pin_memory(), torch.empty(n, 3).pin_memory()
    pd, ud = torch.empty_like(p), torch.empty_like(u)

    def e2e_step():
        pd.copy_(ph, non_blocking=True)
        ud.copy_(uh, non_blocking=True)
        d, pdf = V.sample_direction(pd, ud)
        pdf2 = V.pdf_direction(pd, d)
        rgb = E.render(pd.t().contiguous(), d.t().contiguous())
        dh.copy_(d, non_blocking=True)
        pdfh.copy_(pdf2, non_blocking=True)
        rh.copy_(rgb, non_blocking=True)
    e_ms = T.run(e2e_step, max(3, args.steps // 2), 1)
    if rank != 0:
        return None
    hbm, src = measured_peaks()
    return {"metric": METRIC["vmf"], "value": CONFIG4_SAMPLES / (ms * 1e-3), "unit": "samples/s", "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "f16" if args.precision == "fp16" else "f32", "data": "synthetic",
            "config": workload_config(args, "vmf", n), "clocks": clk, "gpu_launches": int(launches),
            "e2e": {"value": CONFIG4_SAMPLES / (e_ms * 1e-3), "unit": "samples/s", "h2d_bytes_per_step": 20 * n,
                    "d2h_bytes_per_step": 28 * n},
            "roofline": {"bound": "hbm", "kernel": "whole step (radiance evaluation dominates; vMF kernels are SFU/ALU-bound)",
                         "achieved": (ALG_BYTES_SAMPLING + ALG_BYTES_FIELD + 92) * n / (ms * 1e-3) / 1e9, "peak": hbm,
                         "peak_source": src, "unit": "GB/s",
                         "frac": (ALG_BYTES_SAMPLING + ALG_BYTES_FIELD + 92) * n / (ms * 1e-3) / 1e9 / hbm, "traffic": None,
                         "note": "algorithmic bytes per sample = the radiance query's table gathers (161 792 B, L2-resident "
                                 "tables) + 92 B of sample I/O; see the default line for the per-kernel roofs"}}


def bench_bake(args, E, T, rank, world, local, dist, dev):
    from nerf_emitter_b200 import parallel
    from nerf_emitter_b200.bake import bake_envmap, film_develop
    W, H, SPP = CONFIG5_FILM
    rows = parallel.shard_range(H, rank, world)
    n = (rows[1] - rows[0]) * W * SPP
    E.train(False)
    img = [None]

    def step():
        # the ranks' films are summed by the library's all-reduce (nemit_allreduce_grads on the accumulation blob), then developed
        img[0] = film_develop(bake_envmap(E, W, H, SPP, rows=rows, develop=False, reduce=True))
    clocks = ClockSampler(local)
    for _ in range(args.warmup):
        step()
    T.barrier()
    if rank == 0:
        clocks.start()
    launches0 = E.launch_count()
    ms = T.run(step, args.steps, 0)
    launches = E.launch_count() - launches0 + args.steps
    clk = clocks.stop() if rank == 0 else None
    host = torch.empty(H, W, 4).pin_memory()

    def e2e_step():
        step()
        host.copy_(img[0], non_blocking=True)
    e_ms = T.run(e2e_step, max(3, args.steps // 2), 1)
    finite = bool(torch.isfinite(img[0]).all())
    if rank != 0:
        return None
    total = W * H * SPP
    hbm, src = measured_peaks()
    return {"metric": METRIC["bake"], "value": total / (ms * 1e-3), "unit": "rays/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
            "dtype": "f16" if args.precision == "fp16" else "f32", "data": "synthetic",
            "config": workload_config(args, "bake", n), "clocks": clk, "gpu_launches": int(launches),
            "bake_seconds": ms * 1e-3, "image_finite": finite,
            "e2e": {"value": total / (e_ms * 1e-3), "unit": "rays/s", "h2d_bytes_per_step": 0,
                    "d2h_bytes_per_step": H * W * 16, "note": "rays are generated on the device (the sensor is part of "
                    "the path); the developed (H,W,4) image is read back to pinned host memory every step"},
            "roofline": {"bound": "hbm", "kernel": "whole bake (forward kernels dominate)",
                         "achieved": (ALG_BYTES_SAMPLING + ALG_BYTES_FIELD + 68) * n / (ms * 1e-3) / 1e9, "peak": hbm,
                         "peak_source": src, "unit": "GB/s",
                         "frac": (ALG_BYTES_SAMPLING + ALG_BYTES_FIELD + 68) * n / (ms * 1e-3) / 1e9 / hbm, "traffic": None,
                         "note": "per-rank algorithmic bytes (table gathers, L2-resident) over the max-over-ranks step time"}}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--config", type=int, default=None, choices=[2, 3, 4, 5],
                    help="BASELINE.json configuration to run instead of the default line (see the module docstring)")
    ap.add_argument("--rays", type=int, default=1 << 20, help="rays per GPU per step (default line and --config 2)")
    ap.add_argument("--ref-rays", type=int, default=4096, help="rays (samples) per step of the CPU reference arm")
    ap.add_argument("--mode", default=None, choices=["fwd", "fwd+bwd"])
    ap.add_argument("--precision", default="fp16", choices=["fp16", "fp32", "split"],
                    help="MLP arithmetic of the main field: fp16 tensor-core (BASELINE configs[1]), exact fp32 FMA kernels, or "
                         "split = fp32 arithmetic on the tensor cores through bf16 operand pieces (ncu traffic constants "
                         "are those of the fp16 mode and are left out of the line in the other modes)")
    ap.add_argument("--sampler", default=None, choices=["exact", "fast"],
                    help="proposal sampler arithmetic: exact = the reference's fp32 CPU arithmetic (bit-equal indices), "
                         "fast = packed-half proposal networks (autocast behaviour); default: fast with --precision fp16")
    ap.add_argument("--proposal-grad", action="store_true",
                    help="also back-propagate the interlevel loss into the proposal networks (the reference's sampler "
                         "only lets them carry autograd on its first call after an update step, SURVEY 8a quirk 13)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup
    if args.config == 2:
        kind = "fwd"
    elif args.config == 4:
        kind = "vmf"
    elif args.config == 5:
        kind = "bake"
    else:
        kind = args.mode or "fwd+bwd"

    if args.impl == "reference":
        run_reference_arm(args, rank, kind)
        return

    import torch.distributed as dist
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
    from nerf_emitter_b200 import EmitterModelSpec, NerfEmitter, make_params
    spec = EmitterModelSpec()
    E = NerfEmitter(make_params(spec, seed=0, table_scale=1e3), spec, device=dev)
    E.set_precision(args.precision)
    if args.sampler is None:
        args.sampler = "fast" if args.precision == "fp16" else "exact"
    E.set_sampler(args.sampler)
    T = Timer(dev, world, dist)
    if kind == "vmf":
        line = bench_vmf(args, E, T, rank, world, local, dist, dev)
    elif kind == "bake":
        line = bench_bake(args, E, T, rank, world, local, dist, dev)
    else:
        line = bench_rays(args, E, T, rank, world, local, dist, dev, kind)
    if rank == 0:
        if world == 1 and not args.no_cpu_baseline:
            line["cpu_baseline"] = cpu_reference_rate(kind, args.ref_rays, 3, args.proposal_grad)
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    # stdout carries exactly one JSON line: anything a library writes to fd 1 meanwhile (NCCL's version banner
    # under NCCL_DEBUG) goes to stderr instead
    sys.stdout.flush()
    _json_fd = os.dup(1)
    os.dup2(2, 1)
    _buf = io.StringIO()
    try:
        with contextlib.redirect_stdout(_buf):
            main()
    finally:
        os.write(_json_fd, _buf.getvalue().encode())
