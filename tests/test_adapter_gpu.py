"""GPU: the Legion adapter (htr-solver_b200/csrc/legion_adapter.cc) EXECUTED outside Legion.

oracle/adapter_driver.cc builds, for every task, the PhysicalRegions the reference's region requirements produce (region
order and per-region field lists of src/prometeo_{rhs,sensor,chem,variables,cfl}.cc) over this test's device arrays, with
SPARSE ghost regions, fills the task's Args and calls T::gpu_base_impl.  The stub legion.h aborts on a field that is not
part of the region it is asked from, on ptr() outside an instance and on the Rect of a sparse domain -- so this checks the
region order / field ids / stride unpacking of the one file that has to be right for the drop-in claim.  Every task must
reproduce the direct C-ABI call bit for bit.  (Legion itself is not available: this is the stub, not a Legion run.)"""
import ctypes as C
import os
import numpy as np
import pytest
import torch
import htr_solver_b200 as htr
from htr_solver_b200 import synthetic
from parity import oracle_for, host_state

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CP, AIR = synthetic.CONSTPROP_VISCOUS, synthetic.AIRMIX_3DPERIODIC


def adapter(mix):
    path = os.path.join(ROOT, "oracle", "_ref", f"libhtr_adapter_{mix}.so")
    if not os.path.exists(path):
        pytest.skip("oracle/_ref/libhtr_adapter_*.so not built (make -C oracle adapter)")
    htr.load_library(mix)            # the adapter links the product library: have it resident first
    A = C.CDLL(path)
    A.htr_adapter_run.argtypes = [C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_char_p,
                                  C.c_double, C.c_double, C.c_double, C.c_void_p, C.c_void_p]
    return A


@pytest.mark.parametrize("cfg", [CP, AIR], ids=["constprop", "air"])
def test_every_adapter_task_matches_the_direct_call(cfg):
    from htr_solver_b200 import halo
    mix = cfg["type"]
    A = adapter(mix)
    o = oracle_for(cfg)
    n = (24, 20, 16)
    f = host_state(o, n, cfg, species_wave=1.0 if mix != "ConstPropMix" else 0.0, front=0.3)
    o.ducros_sensor(f, 1.0)
    for d in range(3):
        o.shock_sensor(f, d)
    f["shockSensorX"][:, :, 5::7] = 0; f["shockSensorZ"][3::5, :, :] = 0
    s = htr.Solver(cfg)
    assert A.htr_adapter_sizeof_mix() == len(s.mix_bytes)
    # a block with resident ghosts at a non-zero origin (what a Legion tile instance looks like); unbounded bounds
    fl = s.new_fluid(*n, halo=halo.HALO, origin=(7, -3, 40))
    fl.upload(f)
    halo.exchange([t for k, t in fl.t.items()], n, 0, (1, 1, 1), None)
    tile = fl.tile()
    o3, h = fl.origin, halo.HALO
    alloc = htr.Rect.make((o3[0] - h, o3[1] - h, o3[2] - h), (o3[0] + n[0] - 1 + h, o3[1] + n[1] - 1 + h, o3[2] + n[2] - 1 + h))
    nowrap = htr.Rect.make((-2 ** 40,) * 3, (2 ** 40,) * 3)
    side = torch.cuda.Stream()        # a non-default task stream: the adapter must launch on it
    res = torch.zeros(1, dtype=torch.float64, device="cuda")
    rk, dt = (0.75, 0.25), 1e-3

    def via_adapter(task, d=0, s0=0.0, s1=0.0, s2=0.0):
        torch.cuda.synchronize()
        rc = A.htr_adapter_run(task, d, C.byref(fl.struct()), C.byref(tile), C.byref(alloc), C.byref(nowrap), s.mix_bytes,
                               s0, s1, s2, C.c_void_p(side.cuda_stream), C.c_void_p(res.data_ptr()))
        assert rc == 0
        side.synchronize()

    def compare(out_names, direct, adapted, pre=None):
        keep = {k: fl.t[k].clone() for k in out_names}
        got = []
        for fn in (direct, adapted):
            for k in out_names:
                fl.t[k].copy_(keep[k])
            if pre:
                pre()
            fn()
            torch.cuda.synchronize()
            got.append({k: fl.t[k].clone() for k in out_names})
        for k in out_names:
            fl.t[k].copy_(keep[k])
        for k in out_names:
            assert torch.equal(got[0][k], got[1][k]), k
        return got[0]

    zero_ct = lambda: fl.t["Conserved_t"].zero_()
    schemes = ["Hybrid", "TENO", "TENOA", "TENOLAD", "SkewSymmetric"]
    for task, scheme in enumerate(schemes):
        for d in range(3):
            r = compare(["Conserved_t"], lambda: s.UpdateUsingEulerFlux(fl, scheme, d, rk, dt, tile=tile, bounds=nowrap),
                        lambda: via_adapter(task, d, rk[0], rk[1], dt) if scheme != "SkewSymmetric" else via_adapter(task, d), pre=zero_ct)
            assert float(r["Conserved_t"].abs().max()) > 0
    for d in range(3):
        compare(["Conserved_t"], lambda: s.UpdateUsingDiffusionFlux(fl, d, tile=tile, bounds=nowrap), lambda: via_adapter(5, d), pre=zero_ct)
    compare(["DucrosSensor"], lambda: s.UpdateDucrosSensor(fl, 1.0, tile=tile, bounds=nowrap), lambda: via_adapter(6, 0, 1.0))
    for d, k in enumerate(["shockSensorX", "shockSensorY", "shockSensorZ"]):
        compare([k], lambda: s.UpdateShockSensor(fl, d, tile=tile, bounds=nowrap), lambda: via_adapter(7, d))
    compare(["Conserved_t"], lambda: s.AddChemistrySources(fl, tile=tile), lambda: via_adapter(8), pre=zero_ct)
    compare(["MassFracs", "rho", "mu", "lam", "Di", "SoS"], lambda: s.UpdatePropertiesFromPrimitive(fl, tile=tile), lambda: via_adapter(10))
    compare(["temperature", "pressure", "MolarFracs", "velocity"], lambda: s.UpdatePrimitiveFromConserved(fl, tile=tile), lambda: via_adapter(12))
    # UpdateConservedFromPrimitive over a union of rectangles: exactly those cells change
    before = fl.interior("Conserved").clone()
    fl.interior("Conserved").fill_(-1.0)
    via_adapter(11)
    after = fl.interior("Conserved").clone()
    pieces = torch.zeros(after.shape[:3], dtype=torch.bool, device="cuda")
    pieces[0] = True; pieces[-1] = True; pieces[1:-1, 0, :] = True
    assert torch.equal(after[pieces], before[pieces]) if mix == "ConstPropMix" else torch.allclose(after[pieces], before[pieces], rtol=1e-13, atol=0)
    assert bool((after[~pieces] == -1.0).all())
    fl.interior("Conserved").copy_(before)
    # CFL through the DeferredValue
    direct = s.CalculateMaxSpectralRadius(fl, tile=tile)
    via_adapter(13)
    assert float(res.item()) == direct and direct > 0
    if mix != "ConstPropMix":
        fl.t["Conserved_t"].zero_()
        compare(["Conserved", "Conserved_t_old", "temperature"], lambda: s.UpdateChemistry(fl, 1e-4, tile=tile), lambda: via_adapter(9, 0, 1e-4))
