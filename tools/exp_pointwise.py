"""Times of the cell-local kernels at 256^3 AirMix (RK glue, primitive / property recovery, sensors)."""
import sys, torch, numpy as np
sys.path[:0] = ['.', 'tests']
import bench, htr_solver_b200 as htr
from htr_solver_b200 import synthetic
cfg = synthetic.AIRMIX_3DPERIODIC
s = htr.Solver(cfg, variant=(sys.argv[1] if len(sys.argv) > 1 else "")); n = 256
fl = s.new_fluid(n, n, n)
bench.fill_block(s, fl, cfg, n, (0, 0, 0), (n, n, n), 0)
fl.t["Conserved_old"].copy_(fl.t["Conserved"]); fl.t["Conserved_t"].zero_()
def t(fn, reps=5):
    ts = []
    for _ in range(reps):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record(); torch.cuda.synchronize(); ts.append(a.elapsed_time(b))
    return np.mean(ts[1:])
print("update_vars        ", t(lambda: s.UpdateVars(fl, 2, 1e-7)))
print("primitive          ", t(lambda: s.UpdatePrimitiveFromConserved(fl)))
print("properties         ", t(lambda: s.UpdatePropertiesFromPrimitive(fl)))
print("rk_stage (fused)   ", t(lambda: s.UpdateVarsAndRecover(fl, 2, 1e-7)))
print("ducros             ", t(lambda: s.UpdateDucrosSensor(fl, 1.0)))
for d in range(3):
    print("shock sensor", d, "    ", t(lambda: s.UpdateShockSensor(fl, d)))
print("init temporaries   ", t(lambda: s.InitializeTemporaries(fl)))
