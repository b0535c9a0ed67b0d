"""One UpdateDerivatives pass (htr_rhs) on an n^3 AirMix Taylor-Green box: the command profiled under ncu.
    python tools/exp_rhs_once.py [n] [reps] [opt=value ...]"""
import sys, torch, numpy as np
sys.path[:0] = ['.', 'tests']
import htr_solver_b200 as htr
from htr_solver_b200 import synthetic
import bench
n = int(sys.argv[1]) if len(sys.argv) > 1 else 256
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 2
cfg = synthetic.AIRMIX_3DPERIODIC
s = htr.Solver(cfg)
for kv in sys.argv[3:]:
    k, v = kv.split("="); s.set_option(k, int(v))
fl = s.new_fluid(n, n, n)
bench.fill_block(s, fl, cfg, n, (0, 0, 0), (n, n, n), 0)
s.UpdateDucrosSensor(fl, 1.0)
for d in range(3):
    s.UpdateShockSensor(fl, d)
ts = []
for _ in range(reps):
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record(); s.UpdateDerivatives(fl, "Hybrid", True, True, (1.0, 0.0), 1e-3); b.record(); torch.cuda.synchronize()
    ts.append(a.elapsed_time(b))
print(f"htr_rhs {n}^3 AirMix Hybrid+diffusion+chem: {ts} ms; checksum {fl.t['Conserved_t'].abs().sum().item():.10e}")
