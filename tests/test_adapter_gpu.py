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


@pytest.mark.parametrize("cfg", [CP, AIR], ids=["constprop", "air"])
def test_boundary_adapter_tasks_match_the_direct_calls(cfg):
    """The NSCBC right-hand-side tasks and the boundary-cell primitive updates (src/prometeo_rhs.cu:625-1560,
    src/prometeo_bc.cu:120-410) through T::gpu_base_impl over stub regions -- the written region is the plane of boundary
    cells, the interior neighbours come with the sparse Fluid / ghost regions -- against the direct C-ABI calls, bit for bit."""
    from htr_solver_b200 import halo
    from htr_solver_b200.bc import BCFields, BoundaryTasks, shock_params
    mix = cfg["type"]
    A = adapter(mix)
    A.htr_adapter_run_bc.argtypes = [C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_char_p,
                                     C.c_double, C.c_double, C.c_double, C.c_void_p, C.c_void_p]
    o = oracle_for(cfg)
    n = (20, 16, 12)
    f = host_state(o, n, cfg, species_wave=1.0 if mix != "ConstPropMix" else 0.0, front=0.3)
    c = f["SoS"]
    f["velocity"][..., 0] = c*np.linspace(-1.6, 1.6, n[1])[None, :, None]
    f["velocity"][..., 1] = c*np.linspace(-1.6, 1.6, n[2])[:, None, None]
    f["velocity"][..., 2] = c*np.linspace(-1.6, 1.6, n[0])[None, None, :]
    o.update_conserved_from_primitive(f)
    s = htr.Solver(cfg)
    fl = s.new_fluid(*n, halo=halo.HALO, origin=(7, -3, 40))
    fl.upload(f)
    halo.exchange([t for k, t in fl.t.items()], n, 0, (1, 1, 1), None)
    bcf = BCFields(fl)
    rng = np.random.default_rng(5)
    X = f["MolarFracs"]*(1.0 + 0.3*rng.random(f["MolarFracs"].shape))
    bcf.upload({"dudtBoundary": rng.standard_normal(f["velocity"].shape), "dTdtBoundary": rng.standard_normal(c.shape),
                "velocity_profile": f["velocity"]*1.1, "temperature_profile": f["temperature"]*1.05,
                "MolarFracs_profile": X/X.sum(axis=-1, keepdims=True)})
    bt = BoundaryTasks(s)
    o3, h = fl.origin, halo.HALO
    alloc = htr.Rect.make((o3[0] - h, o3[1] - h, o3[2] - h), (o3[0] + n[0] - 1 + h, o3[1] + n[1] - 1 + h, o3[2] + n[2] - 1 + h))
    nowrap = htr.Rect.make((-2 ** 40,) * 3, (2 ** 40,) * 3)
    side_stream = torch.cuda.Stream()
    P0 = float(f["pressure"].mean()); T0 = float(f["temperature"].mean()); c0 = float(c.mean())
    q = shock_params(o3[0] + n[0]//2, np.full(o.nspec, 1.0/o.nspec), P0, T0, (2*c0, -0.3*c0, 0.0), 2.5*P0, 1.4*T0, (1.5*c0, -0.6*c0, 0.0))

    def plane(d, side):
        lo = list(o3); hi = [o3[i] + n[i] - 1 for i in range(3)]
        lo[d] = hi[d] = (o3[d] if side == 0 else o3[d] + n[d] - 1)
        return htr.Rect.make(lo, hi)

    def via_adapter(task, d, side, bc, MaxMach=0.0, L=0.0, P=0.0):
        torch.cuda.synchronize()
        rc = A.htr_adapter_run_bc(task, d, side, C.byref(fl.struct()), C.byref(bcf.struct()), C.byref(bc), C.byref(alloc), C.byref(nowrap),
                                  s.mix_bytes, MaxMach, L, P, q.ctypes.data, C.c_void_p(side_stream.cuda_stream))
        assert rc == 0
        side_stream.synchronize()

    def compare(out_names, direct, adapted):
        keep = {k: fl.t[k].clone() for k in out_names}
        got = []
        for fn in (direct, adapted):
            for k in out_names:
                fl.t[k].copy_(keep[k])
            fn()
            torch.cuda.synchronize()
            got.append({k: fl.t[k].clone() for k in out_names})
        for k in out_names:
            fl.t[k].copy_(keep[k])
            assert torch.equal(got[0][k], got[1][k]), k
        assert any(not torch.equal(got[0][k], keep[k]) for k in out_names)      # the task wrote something
    fl.t["Conserved_t"].fill_(3.0)
    prim = ["pressure", "temperature", "MolarFracs", "velocity"]
    for d in range(3):
        for side in (0, 1):
            bc = plane(d, side)
            compare(["Conserved_t"], lambda: bt.UpdateUsingFluxNSCBCInflow(fl, bcf, d, side, bc), lambda: via_adapter(0, d, side, bc))
            for MaxMach in (0.5, 1.2):
                compare(["Conserved_t"], lambda: bt.UpdateUsingFluxNSCBCOutflow(fl, d, side, bc, nowrap, MaxMach, 0.7, P0),
                        lambda: via_adapter(1, d, side, bc, MaxMach, 0.7, P0))
                compare(["Conserved_t"], lambda: bt.UpdateUsingFluxNSCBCFarField(fl, bcf, d, side, bc, MaxMach, 0.7, P0),
                        lambda: via_adapter(2, d, side, bc, MaxMach, 0.7, P0))
        bc = plane(d, 0)
        compare(prim, lambda: bt.SetNSCBC_InflowBC(fl, bcf, d, bc, 1.3*P0), lambda: via_adapter(4, d, 0, bc, P=1.3*P0))
    bc = plane(1, 1)
    compare(["Conserved_t"], lambda: bt.UpdateUsingFluxIncomingShock(fl, bc, 0.5, 0.7, q), lambda: via_adapter(3, 1, 1, bc, 0.5, 0.7))
    fl.t["temperature"].mul_(1.02)
    compare(prim, lambda: bt.SetNSCBC_OutflowBC(fl, bc), lambda: via_adapter(5, 1, 1, bc))
    compare(prim, lambda: bt.SetIncomingShockBC(fl, bc, q), lambda: via_adapter(6, 1, 1, bc))
