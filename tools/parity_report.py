#!/usr/bin/env python3
"""GPU-box diagnostic: prints the parity error of every kernel vs the oracle for both library
variants (production = FMA contraction on; strict = -fmad=false + libdevice pow).
Usage: python tools/parity_report.py [n]   (writes nothing; redirect to profiles/ if wanted)"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests")]
import numpy as np
import htr_solver_b200 as htr
from htr_solver_b200 import synthetic
from parity import oracle_for, host_state, to_device, rel_err

CP, AIR = synthetic.CONSTPROP_VISCOUS, synthetic.AIRMIX_3DPERIODIC
CASES = [("constprop", CP, {}), ("constprop-front", CP, {"stretch": 0.2, "front": 0.3}),
         ("air", AIR, {"species_wave": 1.0}),
         ("air-hot-front", AIR, {"species_wave": 1.0, "front": 0.3, "T0": 6.0, "p0": 1.0, "U": 0.5})]


def comp_err(a, b):
    num = np.abs(a - b).reshape(-1, a.shape[-1]).max(axis=0)
    den = np.abs(b).reshape(-1, a.shape[-1]).max(axis=0)
    return num / np.maximum(den, 1e-300)


def main():
    shape = (24, 20, 16)
    for name, cfg, kw in CASES:
        o = oracle_for(cfg)
        f = host_state(o, shape, cfg, **kw)
        o.ducros_sensor(f, 1.0)
        for d in range(3):
            o.shock_sensor(f, d)
        print(f"== {name}: smooth fraction X/Y/Z = {f['shockSensorX'].mean():.2f}/{f['shockSensorY'].mean():.2f}/{f['shockSensorZ'].mean():.2f}")
        for var in ("", "strict"):
            s = htr.Solver(cfg, variant=var)
            fl = to_device(s, f)
            for scheme in ["SkewSymmetric", "TENO", "TENOA", "TENOLAD", "Hybrid"]:
                for d in range(3):
                    o.init_mix(cfg)
                    f["Conserved_t"][...] = 0.0; fl.t["Conserved_t"].zero_()
                    o.euler_flux(f, scheme, d, 0.75, 0.25, 1e-3)
                    s.UpdateUsingEulerFlux(fl, scheme, d, (0.75, 0.25), 1e-3)
                    got = fl.download(["Conserved_t"])["Conserved_t"]
                    ce = comp_err(got, f["Conserved_t"])
                    nbad = int((np.abs(got - f["Conserved_t"]).max(axis=-1) > 1e-9 * np.abs(f["Conserved_t"]).max()).sum())
                    print(f"{name:16s} {var or 'prod':6s} euler {scheme:13s} d={d} max={ce.max():.2e} per-comp={np.array2string(ce, precision=1)} cells>1e-9:{nbad}")
            for d in range(3):
                f["Conserved_t"][...] = 0.0; fl.t["Conserved_t"].zero_()
                o.diffusion_flux(f, d); s.UpdateUsingDiffusionFlux(fl, d)
                got = fl.download(["Conserved_t"])["Conserved_t"]
                ce = comp_err(got, f["Conserved_t"])
                print(f"{name:16s} {var or 'prod':6s} diffusion d={d} max={ce.max():.2e} per-comp={np.array2string(ce, precision=1)}")
            s.UpdateDucrosSensor(fl, 1.0)
            print(f"{name:16s} {var or 'prod':6s} ducros abs err {np.abs(fl.download(['DucrosSensor'])['DucrosSensor'] - f['DucrosSensor']).max():.2e}")
            fl.upload(f, ["DucrosSensor"])
            for d, nm in enumerate(["shockSensorX", "shockSensorY", "shockSensorZ"]):
                s.UpdateShockSensor(fl, d)
                print(f"{name:16s} {var or 'prod':6s} shock sensor d={d} flips {int((fl.download([nm])[nm] != f[nm]).sum())}")
            fl.upload(f)
            s.UpdatePropertiesFromPrimitive(fl)
            g = fl.download(["rho", "mu", "lam", "Di", "SoS", "MassFracs"])
            print(f"{name:16s} {var or 'prod':6s} properties " + " ".join(f"{k}={rel_err(v, f[k]):.1e}" for k, v in g.items()))
            fl.upload(f); s.UpdateConservedFromPrimitive(fl)
            print(f"{name:16s} {var or 'prod':6s} conserved {rel_err(fl.download(['Conserved'])['Conserved'], f['Conserved']):.1e}")
            f2 = {k: v.copy() for k, v in f.items()}
            f2["temperature"] *= 1.03
            fl.upload(f2); o.update_primitive_from_conserved(f2); s.UpdatePrimitiveFromConserved(fl)
            g = fl.download(["temperature", "pressure", "velocity", "MolarFracs"])
            print(f"{name:16s} {var or 'prod':6s} primitive " + " ".join(f"{k}={rel_err(v, f2[k]):.1e}" for k, v in g.items()))
            fl.upload(f)
            ref = o.max_spectral_radius(f); got = s.CalculateMaxSpectralRadius(fl)
            print(f"{name:16s} {var or 'prod':6s} cfl rel err {abs(got-ref)/ref:.1e}")
            if cfg is AIR:
                f["Conserved_t"][...] = 0.0; fl.t["Conserved_t"].zero_()
                o.add_chemistry_sources(f); s.AddChemistrySources(fl)
                got = fl.download(["Conserved_t"])["Conserved_t"]
                print(f"{name:16s} {var or 'prod':6s} chem source max|w|={np.abs(f['Conserved_t']).max():.2e} err={comp_err(got[..., :5], f['Conserved_t'][..., :5])}")


if __name__ == "__main__":
    main()
