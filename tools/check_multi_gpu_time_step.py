"""Multi-GPU parity of the block time step: N ranks (torchrun) advance their blocks of a periodic box by two RK3
steps with HaloExchanger.time_step (NCCL ghost exchange); rank 0 also advances the whole box as one periodic tile
(Solver.TimeStep) and compares its own block.  Strict build: bit-identical.
    python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 tools/check_multi_gpu_time_step.py"""
import os, sys
import numpy as np
import torch
import torch.distributed as dist
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT]
import bench
import htr_solver_b200 as htr
from htr_solver_b200 import synthetic, halo

rank, world, lr = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(lr)
dist.init_process_group("nccl", device_id=torch.device("cuda", lr))
n = 48
tiles = bench.tiles_for(world)
gN = (n*tiles[0], n*tiles[1], n*tiles[2])
tc = (rank % tiles[0], (rank // tiles[0]) % tiles[1], rank // (tiles[0]*tiles[1]))
origin = (tc[0]*n, tc[1]*n, tc[2]*n)
ok = True
for variant in ("strict", ""):
    for cfg, scheme, chem in ((synthetic.AIRMIX_3DPERIODIC, "Hybrid", True), (synthetic.CONSTPROP_VISCOUS, "Hybrid", False)):
        s = htr.Solver(cfg, variant=variant)
        fb = s.new_fluid(n, n, n, halo=halo.HALO, origin=origin)
        bench.fill_block(s, fb, cfg, n, origin, gN, halo.HALO)
        # start from a fixed point of the device's own primitive/property recovery (the analytic velocity is not
        # bit-equal to (rho u)/rho, and ghost copies are always recovered from the exchanged conserved state)
        h = halo.HALO
        full = htr.Rect.make(tuple(o - h for o in origin), tuple(o + n - 1 + h for o in origin))
        for _ in range(2):
            s.UpdatePrimitiveFromConserved(fb, tile=full); s.UpdatePropertiesFromPrimitive(fb, tile=full)
        ex = halo.HaloExchanger(s, fb, tiles, tc, rank, world)
        dt = 1e-4 if cfg["type"] == "ConstPropMix" else 2e-7      # explicit chemistry of the radical-seeded air: stay inside its time scale
        for _ in range(2):
            ex.time_step(scheme, dt, True, chem, 1.0)
        torch.cuda.synchronize()
        # the CFL reduction: every rank must end up with the same (global) spectral radius
        sr = torch.tensor([ex.max_spectral_radius()], dtype=torch.float64, device="cuda")
        allsr = [torch.zeros_like(sr) for _ in range(world)]
        dist.all_gather(allsr, sr)
        if rank == 0:
            same_sr = all(float(a.item()) == float(sr.item()) for a in allsr)
            print(f"{cfg['type']:12s} {variant or 'prod':6s} max spectral radius {float(sr.item()):.6e} identical on all ranks={same_sr}", flush=True)
            ok = ok and same_sr
        got = fb.download(["Conserved", "temperature", "rho", "mu"])
        if rank == 0:
            # the whole box as one periodic tile; the stretched sizes need not be equal
            fl = s.new_fluid(gN[0], gN[1], gN[2], halo=0, origin=(0, 0, 0))
            f = synthetic.taylor_green(gN, cfg, species_wave=1.0 if cfg["type"] != "ConstPropMix" else 0.0)
            for name, arr in f.items():
                if name in fl.t:
                    fl.t[name].copy_(torch.from_numpy(np.ascontiguousarray(arr)))
            s.UpdatePropertiesFromPrimitive(fl); s.UpdateConservedFromPrimitive(fl)
            for _ in range(2):
                s.UpdatePrimitiveFromConserved(fl); s.UpdatePropertiesFromPrimitive(fl)
            for _ in range(2):
                s.TimeStep(fl, scheme, dt, True, chem, 1.0)
            torch.cuda.synchronize()
            ref = fl.download(["Conserved", "temperature", "rho", "mu"])
            for k in got:
                r = ref[k][:n, :n, :n]
                same = np.array_equal(got[k], r)
                err = float(np.max(np.abs(got[k] - r))/np.max(np.abs(r)))
                print(f"{cfg['type']:12s} {variant or 'prod':6s} {k:12s} identical={same} max rel diff={err:.2e}", flush=True)
                ok = ok and (same if variant == "strict" else err <= 1e-12)
        dist.barrier()
if rank == 0:
    print("MULTI-GPU TIME STEP PARITY:", "OK" if ok else "FAILED", f"({world} ranks, {n}^3 per rank)")
dist.destroy_process_group()
