import sys, torch, numpy as np
sys.path[:0] = ['.', 'tests']
import bench, htr_solver_b200 as htr
from htr_solver_b200 import synthetic
cfg = synthetic.AIRMIX_3DPERIODIC
s = htr.Solver(cfg, variant=(sys.argv[1] if len(sys.argv) > 1 else "")); n = 256
fl = s.new_fluid(n, n, n)
bench.fill_block(s, fl, cfg, n, (0, 0, 0), (n, n, n), 0)
s.UpdateDucrosSensor(fl, 1.0)
for d in range(3): s.UpdateShockSensor(fl, d)
def t(fn, reps=5):
    ts = []
    for _ in range(reps):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record(); torch.cuda.synchronize(); ts.append(a.elapsed_time(b))
    return np.mean(ts[1:])
print("K1 chem   ", t(lambda: s.UpdateDerivativesParts(1, fl, "Hybrid", True, True, (1., 0.), 1e-3)))
print("K1 nochem ", t(lambda: s.UpdateDerivativesParts(1, fl, "Hybrid", True, False, (1., 0.), 1e-3)))
print("chem alone", t(lambda: s.AddChemistrySources(fl)))
