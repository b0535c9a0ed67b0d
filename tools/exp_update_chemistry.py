"""UpdateChemistry (implicit Rosenbrock integration, src/prometeo_chem.cc:36-71) on a hot reacting AirMix box: time of
one launch with CUDA events; run under ncu for the fp64-pipe numbers."""
import sys, torch, numpy as np
sys.path[:0] = ['.', 'tests']
import htr_solver_b200 as htr
from htr_solver_b200 import synthetic
cfg = synthetic.AIRMIX_3DPERIODIC
n = int(sys.argv[1]) if len(sys.argv) > 1 else 96
s = htr.Solver(cfg, variant=(sys.argv[2] if len(sys.argv) > 2 else ''))
fl = s.new_fluid(n, n, n)
f = synthetic.taylor_green((n, n, n), cfg, species_wave=1.0, front=0.2, p0=1.0, T0=6.0, U=0.5)
for name, arr in f.items():
    if name in fl.t:
        fl.t[name].copy_(torch.from_numpy(np.ascontiguousarray(arr)))
s.UpdatePropertiesFromPrimitive(fl); s.UpdateConservedFromPrimitive(fl)
fl.t["Conserved_t"].zero_()
keep = {k: fl.t[k].clone() for k in ("Conserved", "temperature")}
dt = 2e-3
ts = []
for _ in range(3):
    for k, v in keep.items():
        fl.t[k].copy_(v)
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record(); s.UpdateChemistry(fl, dt); b.record(); torch.cuda.synchronize()
    ts.append(a.elapsed_time(b))
fails = s.L.htr_chemistry_failures()
d = (fl.t["Conserved"] - keep["Conserved"]).abs().amax(dim=(0, 1, 2)) / keep["Conserved"].abs().amax(dim=(0, 1, 2))
print(f"UpdateChemistry {n}^3 AirMix dt={dt}: {np.mean(ts[1:]):.3f} ms = {n**3/np.mean(ts[1:])/1e3:.1f} M cells/s, failures {fails}, "
      f"finite {bool(torch.isfinite(fl.t['Conserved']).all())}, max relative change of the partial densities {d[:5].max().item():.2e}")
