"""One RK3 time step of the 512^3 AirMix box (the bench workload) - the program profiled by tools/measure_round.sh."""
import sys, torch
sys.path[:0] = ['.', 'tests']
import bench, htr_solver_b200 as htr
from htr_solver_b200 import synthetic
cfg = synthetic.AIRMIX_3DPERIODIC
n = int(sys.argv[1]) if len(sys.argv) > 1 else 512
s = htr.Solver(cfg)
fl = s.new_fluid(n, n, n)
bench.fill_block(s, fl, cfg, n, (0, 0, 0), (n, n, n), 0)
s.TimeStep(fl, "Hybrid", 1e-7, True, True, 1.0)
torch.cuda.synchronize()
print("finite", bool(torch.isfinite(fl.t["Conserved"]).all().item()))
