#!/usr/bin/env python3
"""bench.py -- fp64 RHS cell-updates/s of the HTR hot path on B200 (contract: see DESIGN.md, section Measurement).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload NAME] [--impl reference]

A "step" is one UpdateDerivatives pass (src/prometeo.rg:502-578) over the rank's block:
Conserved_t = 0 (+ chemistry source) + Euler z,y,x + diffusion z,y,x.  Prints ONE JSON line.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests")]

import numpy as np  # noqa: E402

# name -> (mixture config key, n per GPU, scheme, diffusion, chemistry, reference file it is scaled from)
WORKLOADS = {
    "air512":        ("AIRMIX_3DPERIODIC", 512, "Hybrid", True, True,  "solverTests/3DPeriodic_Air/base.json scaled to 512^3 per GPU (north-star target)"),
    "air256":        ("AIRMIX_3DPERIODIC", 256, "Hybrid", True, True,  "solverTests/3DPeriodic_Air/base.json scaled to 256^3"),
    "tg128_hybrid":  ("CONSTPROP_3DPERIODIC", 128, "Hybrid", True, False, "solverTests/3DPeriodic/base.json scaled to 128^3"),
    "tg128_teno":    ("CONSTPROP_3DPERIODIC", 128, "TENO", True, False, "solverTests/3DPeriodic_TENO/base.json scaled to 128^3"),
    "tg512_skew":    ("CONSTPROP_VISCOUS", 512, "SkewSymmetric", True, False, "solverTests/3DPeriodic_SkewSymmetric/base.json scaled to 512^3, Visc=5e-3"),
    # the same AirMix box with a hot tanh temperature/pressure front (the parity tests' "air-hot-front" state): part of the
    # faces is not smooth, so the Hybrid scheme takes the TENO branch there (config.smooth_fraction < 1)
    "air512_front":  ("AIRMIX_3DPERIODIC", 512, "Hybrid", True, True,  "solverTests/3DPeriodic_Air/base.json scaled to 512^3 + tanh front (front=0.3, T0=6, U=0.5)"),
    "air256_front":  ("AIRMIX_3DPERIODIC", 256, "Hybrid", True, True,  "solverTests/3DPeriodic_Air/base.json scaled to 256^3 + tanh front (front=0.3, T0=6, U=0.5)"),
}
    # BASELINE config C5: y stretched with the reference's Tanh map (testcases/scalingTest/WeakScaling/base.json, Stretching
    # 1.3844), metrics by the reference's InitializeMetricTask (fixture tests/golden/tanh_y_metrics.npz), periodic in y for now
WORKLOADS["air512_tanh"] = ("AIRMIX_3DPERIODIC", 512, "Hybrid", True, True, "testcases/scalingTest/WeakScaling/base.json grid (Tanh in y, 512^3 per GPU) with the 3DPeriodic_Air flow")
IC_KW = {"air512_tanh": {"ygrid": "tanh"}, "air512_front": {"front": 0.3, "T0": 6.0, "p0": 1.0, "U": 0.5}, "air256_front": {"front": 0.3, "T0": 6.0, "p0": 1.0, "U": 0.5}}
DEFAULT_WORKLOAD = "air512"


def algorithmic_bytes(nspec, scheme, diffusion, chemistry):
    """Compulsory HBM bytes per RHS cell-update: every needed field read once,
    Conserved_t written once.  B = 8*(nEq*(2+[N>1]) + 3N + 18) + 15 with diffusion and all inputs."""
    neq = nspec + 4
    if diffusion:
        return 8 * (neq * (2 + (1 if nspec > 1 else 0)) + 3 * nspec + 18) + 15
    return 8 * (neq * (2 + (1 if nspec > 1 else 0)) + nspec + 6 + 3) + 15


def kernel_bytes(kind, nspec, scheme, diffusion=True):
    """Algorithmic bytes per cell of ONE launch of a per-direction kernel (fields it must read once +
    Conserved_t read-modify-write), used for the dominant-kernel roofline."""
    neq = nspec + 4
    if kind == "euler":
        b = 8 * (neq + 3 + 3 + nspec + 1) + 4 + 2 * 8 * neq      # Conserved, rho/SoS/p, velocity, MassFracs, m_e, nType, Conserved_t RMW
        if scheme == "Hybrid":
            b += 1
        if nspec > 1 and scheme != "SkewSymmetric":
            b += 8 * neq                                           # Conserved_old (limiter)
        return b
    if kind == "diffusion":
        return 8 * (nspec + 1 + nspec + 3 + nspec + 3 + 4) + 12 + 2 * 8 * neq   # rhoYi, T, Xi, rho/mu/lam, Di, velocity, m_s+3 m_d, 3 nType, RMW
    # staged kernels of the fast path (csrc/htr_fast.cuh): every field they need read once per cell
    if kind == "euler_z_chem":      # rho, p, velocity, Y, rhoE, T, m_e; nType, sensor; Conserved_t written
        return 8 * (2 + 3 + nspec + 1 + 1 + 1) + 4 + (1 if scheme == "Hybrid" else 0) + 8 * neq
    if kind == "euler_xy":          # same state, two metrics, two nType, two sensors; Conserved_t read + written
        return 8 * (2 + 3 + nspec + 1 + 2) + 8 + (2 if scheme == "Hybrid" else 0) + 2 * 8 * neq
    if kind == "diffusion_z":       # rho, mu, lam, T, velocity, Xi, Di, rhoYi, m_s, 3 m_d, 3 nType; Conserved_t RMW
        return 8 * (4 + 3 + 3 * nspec + 4) + 12 + 2 * 8 * neq
    if kind == "diffusion_xy":
        return 8 * (4 + 3 + 3 * nspec + 5) + 12 + 2 * 8 * neq
    # z-marching two-pass form (csrc/htr_march.cuh)
    if kind == "euler_z_march":     # = euler_z_chem: rho, p, velocity, Y, rhoE, T, m_e; nType, sensor; Conserved_t written
        return 8 * (2 + 3 + nspec + 1 + 1 + 1) + 4 + (1 if scheme == "Hybrid" else 0) + 8 * neq
    if kind == "plane_march":       # rho, p, velocity, Y, rhoE, two m_e, [mu, lam, T, Xi, Di, rhoYi, 3 m_d, 3 m_s], 3 nType, two sensors; Conserved_t RMW
        b = 8 * (2 + 3 + nspec + 1 + 2) + 12 + (2 if scheme == "Hybrid" else 0) + 2 * 8 * neq
        if diffusion:
            b += 8 * (3 + 3 * nspec + 6)
        return b
    if kind == "diffusion_xyz":     # the fused three-direction kernel: state once, m_s x 3, m_d x 3, 3 nType; Conserved_t RMW
        return 8 * (4 + 3 + 3 * nspec + 6) + 12 + 2 * 8 * neq
    if kind == "chem":
        return 8 * (3 + nspec) + 2 * 8 * nspec
    if kind == "zero":
        return 8 * neq
    raise KeyError(kind)


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index=0):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.th = threading.Thread(target=self._read, daemon=True)
            self.th.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm = [float(r[0]) for r in self.rows if len(r) >= 7 and r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in self.rows if len(r) >= 7 and r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(len(r) >= 7 and r[3 + i].lower().startswith("active") for r in self.rows)]
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(sm)}


def tanh_ygrid(gny):
    """Cell centres and y metrics of the Tanh-stretched direction from the committed fixture (generated with the reference's
    own metric code by tests/golden/gen_tanh_metrics.py)."""
    d = np.load(os.path.join(ROOT, "tests", "golden", "tanh_y_metrics.npz"))
    if f"ny{gny}_centers" not in d:
        raise SystemExit(f"no Tanh fixture for {gny} cells in y (tests/golden/gen_tanh_metrics.py makes 40, 512, 1024)")
    return {k: d[f"ny{gny}_{k}"] for k in ("centers", "deta_e", "deta_d", "deta_s")}


def fill_block(solver, fl, cfg, n, offset, global_n, halo, slab=32, ic_kw=None):
    """Synthetic Taylor-Green primitives generated slab by slab on the host, uploaded into the block
    (interior + halo, which for a periodic global box is just more of the same analytic field), then
    properties and conserved variables computed by the product kernels over the whole allocation."""
    import torch
    from htr_solver_b200 import synthetic
    nx = ny = nz = n
    h = halo
    tot = n + 2 * h
    ic_kw = dict(ic_kw or {})
    if ic_kw.get("ygrid") == "tanh":
        ic_kw["ygrid"] = tanh_ygrid(global_n[1])
    for k0 in range(0, tot, slab):
        k1 = min(tot, k0 + slab)
        f = synthetic.taylor_green((tot, tot, tot), cfg, zslab=(k0, k1), species_wave=1.0 if cfg["type"] != "ConstPropMix" else 0.0,
                                   offset=(offset[0] - h, offset[1] - h, offset[2] - h), global_n=global_n, **(ic_kw or {}))
        for name, arr in f.items():
            if name in fl.t:
                fl.t[name][k0:k1].copy_(torch.from_numpy(np.ascontiguousarray(arr)))
    from htr_solver_b200 import Rect
    o = fl.origin
    full = Rect.make((o[0] - h, o[1] - h, o[2] - h), (o[0] + n - 1 + h, o[1] + n - 1 + h, o[2] + n - 1 + h))
    solver.UpdatePropertiesFromPrimitive(fl, tile=full)
    solver.UpdateConservedFromPrimitive(fl, tile=full)
    fl.t["Conserved_old"].copy_(fl.t["Conserved"])
    torch.cuda.synchronize()


def cpu_reference_arm(cfgname, scheme, diffusion, chemistry, budget_s=20.0, n=None, ic_kw=None, keep=None):
    """Times the reference's own CPU tasks (oracle/_ref, `-O3 -fopenmp -DREALM_USE_OPENMP`, all host
    threads) on a bounded sample of the workload: the same Taylor-Green box at n^3.
    Returns (cell-updates/s, description dict).  TEST/BASELINE infrastructure only."""
    from oracle_lib import Oracle, alloc_fields
    from htr_solver_b200 import synthetic
    cfg = getattr(synthetic, cfgname)
    mix = cfg["type"]
    try:
        o = Oracle(mix, omp=True, kind="reference")
        kind = "reference"
    except FileNotFoundError:
        o = Oracle(mix, omp=False)
        kind = o.kind
    o.init_mix(cfg)
    if n is None:
        n = 160 if mix != "ConstPropMix" else 192
    f = alloc_fields(n, n, n, o.nspec)
    ic_kw = dict(ic_kw or {})
    if ic_kw.get("ygrid") == "tanh":          # the sample box gets its own Tanh grid, metrics by the reference's code
        from oracle_lib import tanh_y_metrics
        ic_kw["ygrid"] = tanh_y_metrics(o, n)
    prim = synthetic.taylor_green(n, cfg, species_wave=1.0 if mix != "ConstPropMix" else 0.0, **ic_kw)
    for k, v in prim.items():
        f[k][...] = v
    o.update_properties_from_primitive(f)
    o.update_conserved_from_primitive(f)
    f["Conserved_old"][...] = f["Conserved"]
    o.ducros_sensor(f, 1.0)
    for d in range(3):
        o.shock_sensor(f, d)
    o.rhs(f, scheme, diffusion, chemistry, 1.0, 0.0, 1e-3)          # warm-up
    times = []
    t_end = time.time() + budget_s
    while len(times) < 2 or (time.time() < t_end and len(times) < 50):
        t0 = time.perf_counter()
        o.rhs(f, scheme, diffusion, chemistry, 1.0, 0.0, 1e-3)
        times.append(time.perf_counter() - t0)
    best = min(times)
    if keep is not None:
        keep["fields"], keep["oracle"], keep["cfg"] = f, o, cfg
    return n ** 3 / best, {"kind": kind, "cores": o.threads, "sample": f"{n}^3 box of the same workload, best of {len(times)} RHS evaluations after 1 warm-up",
                           "host_cpus": os.cpu_count(), "ms_per_rhs": best * 1e3, "n": n}


def sample_parity(solver, keep, scheme, diffusion, chemistry):
    """The GPU path (htr_rhs resident and htr_host_rhs from host buffers) on the SAME sample box the CPU baseline just
    evaluated with the reference's own tasks: max error of Conserved_t relative to the flux-operand scale
    (tests/parity.py), relative to max|ref| per component (what scripts/compare_hdf.py + solverTests/testAll.py use), and
    the number of shock-sensor decisions that differ."""
    import torch
    from parity import euler_scale, scaled_err, rel_err
    f, o = keep["fields"], keep["oracle"]
    ref = f["Conserved_t"].copy()
    nz, ny, nx = f["rho"].shape
    fl = solver.new_fluid(nx, ny, nz)
    fl.upload({k: v for k, v in f.items() if k != "Conserved_t"})
    flips = 0
    if scheme == "Hybrid":
        keep_s = {k: f[k].copy() for k in ("shockSensorX", "shockSensorY", "shockSensorZ")}
        solver.UpdateDucrosSensor(fl, 1.0)
        for d in range(3):
            solver.UpdateShockSensor(fl, d)
        got_s = fl.download(list(keep_s))
        flips = int(sum((got_s[k] != keep_s[k]).sum() for k in keep_s))
        fl.upload(keep_s)                      # same decisions for the flux comparison
    solver.UpdateDerivatives(fl, scheme, diffusion, chemistry, (1.0, 0.0), 1e-3)
    got = fl.download(["Conserved_t"])["Conserved_t"]
    scale = np.maximum.reduce([euler_scale(f, d) for d in range(3)])
    if chemistry:
        f["Conserved_t"][...] = 0.0
        o.add_chemistry_sources(f)
        scale = scale + np.abs(f["Conserved_t"]).reshape(-1, scale.size).max(axis=0)
        f["Conserved_t"][...] = ref
    out = {"sample_grid": [nx, ny, nz], "max_scaled_err": scaled_err(got, ref, scale), "max_rel_err_by_maxref": rel_err(got, ref),
           "sensor_flips": flips, "tolerance": 1e-10 if chemistry else 1e-12,
           "against": "the reference's CPU tasks (oracle/_ref, -O3 -fopenmp build) on the same inputs"}
    # the host-buffer entry point on the same sample
    neq = got.shape[-1]
    hC = torch.from_numpy(np.ascontiguousarray(f["Conserved"])).pin_memory()
    hT = torch.from_numpy(np.ascontiguousarray(f["temperature"])).pin_memory()
    hCt = torch.empty((nz, ny, nx, neq), dtype=torch.float64).pin_memory()
    solver.HostRHS(fl, scheme, hC, hT, hCt, diffusion=diffusion, chemistry=chemistry, deltaTime=1e-3)
    torch.cuda.synchronize()
    out["host_rhs_max_scaled_err"] = scaled_err(hCt.numpy(), ref, scale)
    out["ok"] = bool(out["max_scaled_err"] <= out["tolerance"] and out["host_rhs_max_scaled_err"] <= out["tolerance"])
    del fl
    return out


def run_reference(args, rank, world):
    name = args.workload
    cfgname, n, scheme, diffusion, chemistry, src = WORKLOADS[name]
    if rank != 0:
        return
    # each "step" = one RHS evaluation of the bounded sample; steps/warmup are honoured on the sample
    from oracle_lib import Oracle  # noqa: F401
    val, desc = cpu_reference_arm(cfgname, scheme, diffusion, chemistry, budget_s=max(5.0, min(60.0, 4.0 * args.steps)), ic_kw=IC_KW.get(name))
    line = {"impl": "reference", "metric": "fp64 RHS cell-updates/s", "value": val, "unit": "cell-updates/s",
            "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": desc["ms_per_rhs"],
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": name, "source": src, "scheme": scheme, "diffusion": diffusion, "chemistry": chemistry,
                       "grid_per_gpu": [n, n, n], "sample_grid": [desc["n"]] * 3},
            "cpu_baseline": {"value": val, "unit": "cell-updates/s", "cores": desc["cores"], "kind": desc["kind"], "sample": desc["sample"]},
            "e2e": {"value": val, "unit": "cell-updates/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


_JSON_FD = None


def emit(text):
    """The one JSON line of this run, on the real stdout."""
    if _JSON_FD is None:
        print(text, flush=True)
    else:
        sys.stdout.flush()
        os.write(_JSON_FD, (text + "\n").encode())


def tiles_for(world):
    return {1: (1, 1, 1), 2: (1, 1, 2), 4: (1, 2, 2), 8: (2, 2, 2)}.get(world) or (1, 1, world)


def run_gpu(args, rank, world, local_rank):
    import torch
    import torch.distributed as dist
    import htr_solver_b200 as htr
    from htr_solver_b200 import synthetic

    name = args.workload
    cfgname, n, scheme, diffusion, chemistry, src = WORKLOADS[name]
    if args.n:
        n = args.n
    cfg = getattr(synthetic, cfgname)
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    solver = htr.Solver(cfg, variant=args.variant)
    nspec, neq = solver.nspec, solver.neq
    for kv in args.opt:                      # library switches for A/B runs, e.g. --opt bulk=0
        k, v = kv.split("=")
        solver.set_option(k, int(v))
    tiles = tiles_for(world)
    gN = (n * tiles[0], n * tiles[1], n * tiles[2])
    tc = (rank % tiles[0], (rank // tiles[0]) % tiles[1], rank // (tiles[0] * tiles[1]))
    origin = (tc[0] * n, tc[1] * n, tc[2] * n)

    exchanger = None
    if world == 1:
        # one tile = the whole periodic box: no ghost cells, stencils wrap by index like the reference
        fl = solver.new_fluid(n, n, n, halo=0, origin=origin)
        fill_block(solver, fl, cfg, n, origin, gN, 0, ic_kw=IC_KW.get(name))
        tile, bounds = fl.tile(), fl.tile()
    else:
        from htr_solver_b200 import halo as halo_mod
        fl = solver.new_fluid(n, n, n, halo=halo_mod.HALO, origin=origin)
        fill_block(solver, fl, cfg, n, origin, gN, halo_mod.HALO, ic_kw=IC_KW.get(name))
        exchanger = halo_mod.HaloExchanger(solver, fl, tiles, tc, rank, world)
        tile = fl.tile()
        bounds = htr.Rect.make((-2 ** 40,) * 3, (2 ** 40,) * 3)     # ghosts are resident: never wrap
    # shock sensors once (per time step in the reference, src/prometeo.rg:1356)
    if scheme == "Hybrid":
        if exchanger is not None:
            exchanger.sensors(1.0)
        else:
            solver.UpdateDucrosSensor(fl, 1.0)
            for d in range(3):
                solver.UpdateShockSensor(fl, d)
    torch.cuda.synchronize()
    smooth = float(fl.interior("shockSensorX").float().mean().item()) if scheme == "Hybrid" else None

    def step():
        if exchanger is not None:
            exchanger.rhs(scheme, diffusion, chemistry, (1.0, 0.0), 1e-3)
        else:
            solver.UpdateDerivatives(fl, scheme, diffusion, chemistry, (1.0, 0.0), 1e-3, tile=tile, bounds=bounds)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(max(args.warmup, 3)):
        step()
    barrier()
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    ev0.record()
    for _ in range(args.steps):
        step()
    ev1.record()
    barrier()
    ms = ev0.elapsed_time(ev1)
    clocks = sampler.stop() if rank == 0 else None
    if world > 1:
        t = torch.tensor([ms], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    cells_rank = n ** 3
    value = cells_rank * world * args.steps / (ms * 1e-3)

    # ---- multi-GPU: where the step goes (each phase alone on the main stream, max over ranks) ----
    # In the timed step the exchange (NCCL + ghost recovery) runs on a side stream under the inner box; the shells follow.
    phases = None
    if exchanger is not None:
        def phase_ms(fn, k=5):
            fn()
            barrier()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            for _ in range(k):
                fn()
            b.record()
            barrier()
            t = torch.tensor([a.elapsed_time(b) / k], dtype=torch.float64, device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            return float(t.item())

        def ghost_recovery():
            for box in exchanger.ghost_boxes:
                solver.UpdateGhostPrimitiveFromConserved(fl, tile=box)
                solver.UpdatePropertiesFromPrimitive(fl, tile=box)

        def boxes(bs):
            for box in bs:
                solver.UpdateDerivatives(fl, scheme, diffusion, chemistry, (1.0, 0.0), 1e-3, tile=box, bounds=exchanger.nowrap)
        try:
            cells = lambda r: int(np.prod([r.hi[i] - r.lo[i] + 1 for i in range(3)]))
            phases = {"nccl_exchange_ms": phase_ms(lambda: exchanger.nccl.exchange(fl, ["Conserved", "temperature"])),
                      "ghost_recovery_ms": phase_ms(ghost_recovery),
                      "inner_box_ms": phase_ms(lambda: boxes([exchanger.inner])),
                      "shell_boxes_ms": phase_ms(lambda: boxes(exchanger.shells)),
                      "whole_tile_one_call_ms": phase_ms(lambda: boxes([tile])),
                      "inner_cells": cells(exchanger.inner), "shell_cells": [cells(b) for b in exchanger.shells],
                      "ghost_cells": [cells(b) for b in exchanger.ghost_boxes],
                      "overlap": bool(exchanger.overlap),
                      "how": "each phase alone on the main stream, CUDA events, mean of 5, max over ranks; in the timed step " +
                             ("the exchange and the ghost recovery run on a side stream under the inner box, the shells follow"
                              if exchanger.overlap else
                              "the exchange and the ghost recovery run first and the block is one htr_rhs call (one remote direction)")}
        except Exception as ex:
            phases = {"error": repr(ex)[:200]}

    # ---- per-kernel durations (CUDA events on the launching stream) for the dominant-kernel roofline ----
    # htr_rhs_parts runs one staged kernel (+ its masked generic pass, a few microseconds when nothing is deferred)
    kern = {}
    staged = scheme in ("Hybrid", "SkewSymmetric")
    march = 1 if staged else 0          # library default (htr_set_option "march"): bit 0 Euler-z march, bit 1 plane march
    for kv in args.opt:
        if staged and kv.replace(" ", "").startswith("march="):
            march = int(kv.split("=")[1])
    if rank == 0:
        def timed(label, fn, reps=5):
            ts = []
            for _ in range(reps):
                a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                a.record(); fn(); b.record(); torch.cuda.synchronize()
                ts.append(a.elapsed_time(b))
            kern[label] = float(np.mean(ts[1:]))
        if staged and (march & 2):
            parts = [("euler_z_march", 1), ("plane_march", 2)]
        elif staged:
            diff3 = any(kv.replace(" ", "") == "diff3=1" for kv in args.opt)          # library default: off
            dparts = ([("diffusion_xyz", 8)] if diff3 else [("diffusion_z", 4), ("diffusion_xy", 8)]) if diffusion else []
            parts = [("euler_z_march" if march & 1 else "euler_z_chem", 1), ("euler_xy", 2)] + dparts
        if staged:
            for label, m in parts:
                timed(label, lambda m=m: solver.UpdateDerivativesParts(m, fl, scheme, diffusion, chemistry, (1.0, 0.0), 1e-3, tile=tile, bounds=bounds))
        else:
            for d in (2, 1, 0):
                timed(f"euler_{'xyz'[d]}", lambda d=d: solver.UpdateUsingEulerFlux(fl, scheme, d, (1.0, 0.0), 1e-3, tile=tile, bounds=bounds))
            if diffusion:
                for d in (2, 1, 0):
                    timed(f"diffusion_{'xyz'[d]}", lambda d=d: solver.UpdateUsingDiffusionFlux(fl, d, tile=tile, bounds=bounds))
            if chemistry:
                timed("chem_source", lambda: solver.AddChemistrySources(fl, tile=tile))
    if world > 1:
        dist.barrier()

    # ---- end to end through the host-buffer C ABI (pinned host state in, Conserved_t out) ----
    e2e = None
    if rank == 0 or world > 1:
        try:
            hl = fl if world == 1 else solver.new_fluid(n, n, n, halo=0, origin=(0, 0, 0))
            if world > 1:      # e2e at N>1: every rank drives its own periodic block through the host API
                fill_block(solver, hl, cfg, n, (0, 0, 0), (n, n, n), 0, ic_kw=IC_KW.get(name))
            hC = torch.empty((n, n, n, neq), dtype=torch.float64).pin_memory()
            hT = torch.empty((n, n, n), dtype=torch.float64).pin_memory()
            hCt = torch.empty((n, n, n, neq), dtype=torch.float64).pin_memory()
            hC.copy_(hl.t["Conserved"]); hT.copy_(hl.t["temperature"])
            torch.cuda.synchronize()
            ke = max(1, min(args.steps, args.e2e_steps))
            solver.HostRHS(hl, scheme, hC, hT, hCt, diffusion=diffusion, chemistry=chemistry, deltaTime=1e-3)   # warm-up
            torch.cuda.synchronize()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            barrier()
            a.record()
            for _ in range(ke):
                solver.HostRHS(hl, scheme, hC, hT, hCt, diffusion=diffusion, chemistry=chemistry, deltaTime=1e-3)
            b.record()
            barrier()
            ems = a.elapsed_time(b)
            if world > 1:
                t = torch.tensor([ems], dtype=torch.float64, device=dev)
                dist.all_reduce(t, op=dist.ReduceOp.MAX)
                ems = float(t.item())
            h2d = 8 * cells_rank * (neq + 1)      # Conserved + temperature (Conserved_old = Conserved is copied on the device)
            d2h = 8 * cells_rank * (neq + 1)
            e2e = {"value": cells_rank * world * ke / (ems * 1e-3), "unit": "cell-updates/s",
                   "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h, "steps": ke, "ms_per_step": ems / ke,
                   "checksum": float(hCt.view(-1)[:: max(1, hCt.numel() // 4096)].abs().sum().item()),
                   "includes": "H2D Conserved+T (z-slab pipeline), UpdatePrimitiveFromConserved, UpdatePropertiesFromPrimitive, sensors, UpdateDerivatives, D2H Conserved_t+T"}
        except Exception as ex:   # report, never silently drop
            e2e = {"value": None, "unit": "cell-updates/s", "error": repr(ex)[:200]}
    # ---- whole RK3 time step (sensors + 3 x (UpdateDerivatives, UpdateVars + primitive/property recovery)) ----
    # extra information beside the headline metric; run last because it advances the state
    tstep = None
    if rank == 0 and world == 1:
        try:
            dts = 1e-7
            solver.TimeStep(fl, scheme, dts, diffusion, chemistry, 1.0, tile=tile, bounds=bounds)
            torch.cuda.synchronize()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            for _ in range(3):
                solver.TimeStep(fl, scheme, dts, diffusion, chemistry, 1.0, tile=tile, bounds=bounds)
            b.record(); torch.cuda.synchronize()
            tms = a.elapsed_time(b) / 3
            a.record()
            for st in (1, 2, 3):
                solver.UpdateVarsAndRecover(fl, st, dts, False, tile=tile)
            b.record(); torch.cuda.synchronize()
            tstep = {"ms_per_time_step": tms, "cell_time_steps_per_s": cells_rank / (tms * 1e-3),
                     "rk_stage_ms": a.elapsed_time(b) / 3, "finite": bool(torch.isfinite(fl.t["Conserved"]).all().item()),
                     "includes": "UpdateShockSensors, InitializeTemporaries, 3 x (htr_rhs, htr_rk_stage)"}
        except Exception as ex:
            tstep = {"error": repr(ex)[:200]}
    if world > 1:
        # the block form of the same step: sensors and RHS with their NCCL ghost exchanges, fused RK stage per block
        try:
            dts = 1e-7
            exchanger.time_step(scheme, dts, diffusion, chemistry, 1.0)
            barrier()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            for _ in range(3):
                exchanger.time_step(scheme, dts, diffusion, chemistry, 1.0)
            b.record()
            barrier()
            t = torch.tensor([a.elapsed_time(b) / 3], dtype=torch.float64, device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            fin = torch.tensor([float(torch.isfinite(fl.t["Conserved"]).all().item())], dtype=torch.float64, device=dev)
            dist.all_reduce(fin, op=dist.ReduceOp.MIN)
            tms = float(t.item())
            tstep = {"ms_per_time_step": tms, "cell_time_steps_per_s": cells_rank * world / (tms * 1e-3), "finite": bool(fin.item()),
                     "includes": "HaloExchanger.time_step: sensors + 3 x (ghost exchange, htr_rhs " +
                                 ("on inner box and shells" if exchanger.overlap else "on the block in one call") + ", htr_rk_stage)"}
        except Exception as ex:
            tstep = {"error": repr(ex)[:200]}
    if rank != 0:
        return

    # ---- roofline of the dominant kernel ----
    peaks_path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(peaks_path):
        peak, peak_src = float(json.load(open(peaks_path))["hbm_gbs"]), "MEASURED_PEAKS.json hbm_gbs (of measured)"
    else:
        peak, peak_src = 6650.0, "B200_PROFILING.md fallback (of fallback)"
    roof = None
    if kern:
        top = max(kern, key=kern.get)
        kind = top if staged else (top.split("_")[0] if not top.startswith("chem") else "chem")
        kb = kernel_bytes(kind, nspec, scheme, diffusion)
        ach = kb * cells_rank / (kern[top] * 1e-3) / 1e9
        B = algorithmic_bytes(nspec, scheme, diffusion, chemistry)
        # measured DRAM bytes of that kernel (ncu --set full capture committed under profiles/), per launch
        traffic = None
        tpath = os.path.join(ROOT, "profiles", "r2_dram_traffic.json")
        if not os.path.exists(tpath):
            tpath = os.path.join(ROOT, "profiles", "r1_dram_traffic.json")
        if os.path.exists(tpath):
            t = json.load(open(tpath)).get(cfg["type"], {}).get(top)
            if t:
                traffic = t["dram_bytes_per_cell"] * cells_rank
        roof = {"bound": "hbm", "kernel": top, "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak,
                "traffic": traffic, "peak_source": peak_src, "kernel_ms": kern[top], "kernel_bytes_per_cell": kb,
                "kernels_ms": kern, "rhs_bytes_per_cell": B,
                "rhs_achieved_gbs": B * cells_rank * world * args.steps / (ms * 1e-3) / 1e9 / world,
                "rhs_frac": B * cells_rank * args.steps / (ms * 1e-3) / 1e9 / peak}
    listed = not any(kv.replace(" ", "") == "list=0" for kv in args.opt)
    launches_per_step = htr.core.rhs_launches(scheme, diffusion, chemistry, march=bool(march & 2), listed=listed)
    if exchanger is not None:
        launches_per_step = exchanger.launches_per_step(diffusion, chemistry, scheme)

    # fp64 pipe peak of this device (DFMA chains, csrc/htr_kernels.cu: htr_probe_fp64): the other roofline of the path
    if roof is not None:
        try:
            import ctypes
            tf = (ctypes.c_double * 2)()
            solver.L.htr_probe_fp64.argtypes = [ctypes.c_double, ctypes.c_void_p, ctypes.c_void_p]
            if solver.L.htr_probe_fp64(1.5, tf, solver._stream()) == 0:
                roof["fp64_peak_tflops"] = {"sustained": tf[0], "burst": tf[1], "how": "16 independent DFMA chains per thread, 8 CTAs x 256 threads per SM, CUDA events"}
        except Exception as ex:
            roof["fp64_peak_tflops"] = {"error": repr(ex)[:120]}

    cpu = None
    parity = None
    if world == 1 and not args.no_cpu:
        try:
            keep = {}
            v, desc = cpu_reference_arm(cfgname, scheme, diffusion, chemistry, budget_s=args.cpu_budget, ic_kw=IC_KW.get(name), keep=keep)
            cpu = {"value": v, "unit": "cell-updates/s", "cores": desc["cores"], "kind": desc["kind"], "sample": desc["sample"],
                   "host_cpus": desc["host_cpus"]}
            parity = sample_parity(solver, keep, scheme, diffusion, chemistry)
        except Exception as ex:
            cpu = cpu or {"value": None, "unit": "cell-updates/s", "error": repr(ex)[:200]}
            parity = parity or {"error": repr(ex)[:200]}

    line = {"metric": "fp64 RHS cell-updates/s", "value": value, "unit": "cell-updates/s", "n_gpus": world,
            "steps": args.steps, "warmup": max(args.warmup, 3), "ms_per_step": ms / args.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": name, "source": src, "mixture": cfg["type"], "scheme": scheme, "diffusion": diffusion,
                       "chemistry": chemistry, "grid_per_gpu": [n, n, n], "global_grid": list(gN), "tiles": list(tiles),
                       "smooth_fraction": smooth, "l2": "inputs larger than L2 (no flush needed)" if n >= 128 else "inputs may fit L2",
                       "halo": "none (periodic wrap by index)" if world == 1 else
                               ("3 ghost planes, NCCL send/recv per step, " +
                                ("overlapped with the inner box (shells follow)" if exchanger.overlap else
                                 "up front, then the block in one htr_rhs call (one remote direction: see halo.py)"))},
            "clocks": clocks, "e2e": e2e, "gpu_launches": launches_per_step * args.steps,
            "roofline": roof, "cpu_baseline": cpu, "parity": parity, "time_step": tstep, "phases": phases, "build": solver.L.htr_build_info().decode()}
    emit(json.dumps(line))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--workload", default=DEFAULT_WORKLOAD, choices=sorted(WORKLOADS))
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--n", type=int, default=0, help="override cells per direction per GPU (debugging only)")
    ap.add_argument("--e2e-steps", type=int, default=3)
    ap.add_argument("--cpu-budget", type=float, default=15.0)
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--opt", action="append", default=[], help="htr_set_option name=value (A/B runs)")
    ap.add_argument("--variant", default="", help="library build variant (tuning experiments; default = production)")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference(args, rank, world)
        return
    if world > 1:
        # stdout must carry exactly ONE JSON line: NCCL prints its version banner to fd 1 when the first communicator is
        # created, so fd 1 points at stderr until the line is emitted through the saved descriptor
        global _JSON_FD
        sys.stdout.flush()
        _JSON_FD = os.dup(1)
        os.dup2(2, 1)
        import torch
        import torch.distributed as dist
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        torch.cuda.set_device(local_rank)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    try:
        run_gpu(args, rank, world, local_rank)
    finally:
        if world > 1:
            import torch.distributed as dist
            dist.destroy_process_group()


if __name__ == "__main__":
    main()
