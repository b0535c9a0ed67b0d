"""GPU: the NSCBC right-hand-side tasks (src/prometeo_rhs.inl:1285-1930) and the boundary-cell primitive updates
(src/prometeo_bc.cc:80-264) through the C ABI against the reference's own CPU tasks (oracle/ref_bc_driver.cc).

The box is bounded by collocated boundary nodes in all three directions (L_C Lp1_C Std ... Rm2_C Rm1_C R_C), every task runs
on the boundary plane of its (direction, side), and the normal velocity sweeps -1.6 .. 1.6 speeds of sound ACROSS each
plane, so one call exercises every branch of a task: supersonic / subsonic inflow, subsonic / supersonic outflow, back flow.
Strict build: the same IEEE operations in the same order -> agreement to rounding of the few libm calls; production build:
FMA contraction only."""
import numpy as np
import pytest
import htr_solver_b200 as htr
from htr_solver_b200 import synthetic
from htr_solver_b200.bc import BCFields, BoundaryTasks, shock_params, MINUS, PLUS
from htr_solver_b200.core import Rect
from oracle_lib import alloc_bc_fields
from parity import oracle_for, host_state, to_device, rel_err, TOL_FLUX, TOL_CHEM

pytestmark = pytest.mark.gpu
CP = synthetic.CONSTPROP_VISCOUS
AIR = synthetic.AIRMIX_3DPERIODIC
L_C, Lp1_C, Rm2_C, Rm1_C, R_C = 8, 9, 10, 11, 12
N = (22, 18, 14)      # nx, ny, nz


def bounded_state(cfg, seed=0):
    o = oracle_for(cfg)
    kw = {"stretch": 0.2, "front": 0.3}
    if cfg is AIR:
        kw["species_wave"] = 1.0
    f = host_state(o, N, cfg, **kw)
    for name, ax in (("nType_x", 2), ("nType_y", 1), ("nType_z", 0)):
        t = f[name]
        t[...] = 0
        idx = [slice(None)]*3
        for k, v in ((0, L_C), (1, Lp1_C), (-3, Rm2_C), (-2, Rm1_C), (-1, R_C)):
            idx[ax] = k
            t[tuple(idx)] = v
            idx[ax] = slice(None)
    # normal velocity of every boundary plane sweeps -1.6 .. 1.6 c along a transverse index
    nz, ny, nx = f["rho"].shape
    c = f["SoS"]
    g = lambda n: np.linspace(-1.6, 1.6, n)
    f["velocity"][..., 0] = c*g(ny)[None, :, None]
    f["velocity"][..., 1] = c*g(nz)[:, None, None]
    f["velocity"][..., 2] = c*g(nx)[None, None, :]
    o.update_conserved_from_primitive(f)
    rng = np.random.default_rng(seed)
    b = alloc_bc_fields(nx, ny, nz, o.nspec)
    b["dudtBoundary"][...] = rng.standard_normal(b["dudtBoundary"].shape)*np.abs(c).mean()
    b["dTdtBoundary"][...] = rng.standard_normal(b["dTdtBoundary"].shape)*f["temperature"].mean()
    b["velocity_profile"][...] = f["velocity"]*(1.0 + 0.2*rng.standard_normal(f["velocity"].shape))
    b["temperature_profile"][...] = f["temperature"]*(1.0 + 0.1*rng.random(f["temperature"].shape))
    X = f["MolarFracs"]*(1.0 + 0.3*rng.random(f["MolarFracs"].shape))
    b["MolarFracs_profile"][...] = X/X.sum(axis=-1, keepdims=True)
    return o, f, b


def plane(d, side):
    lo = [0, 0, 0]; hi = [N[0] - 1, N[1] - 1, N[2] - 1]
    lo[d] = hi[d] = (0 if side == MINUS else N[d] - 1)
    return lo, hi


def on_plane(a, d, side):
    idx = [slice(None)]*3
    idx[2 - d] = 0 if side == MINUS else -1
    return a[tuple(idx)]


def tol_for(cfg, variant):
    if variant == "strict":
        return 1e-13
    return TOL_CHEM if cfg is AIR else TOL_FLUX


@pytest.mark.parametrize("variant", ["", "strict"], ids=["prod", "strict"])
@pytest.mark.parametrize("cfg", [CP, AIR], ids=["constprop", "air"])
def test_nscbc_rhs_tasks(cfg, variant):
    o, f, b = bounded_state(cfg)
    s = htr.Solver(cfg, variant=variant)
    fl = to_device(s, f)
    bcf = BCFields(fl); bcf.upload(b)
    bt = BoundaryTasks(s)
    full = fl.tile()
    tol = tol_for(cfg, variant)
    P0 = float(f["pressure"].mean())
    Lsc = 0.7
    for d in range(3):
        for side in (MINUS, PLUS):
            lo, hi = plane(d, side)
            bc = Rect.make(lo, hi)
            for task in ("inflow", "outflow-lowMach", "outflow-highMach", "farfield-lowMach", "farfield-highMach"):
                MaxMach = 0.5 if task.endswith("lowMach") else 1.2
                f["Conserved_t"][...] = 3.0
                fl.t["Conserved_t"].fill_(3.0)
                if task == "inflow":
                    o.nscbc_inflow(f, b, d, side, lo, hi)
                    bt.UpdateUsingFluxNSCBCInflow(fl, bcf, d, side, bc)
                elif task.startswith("outflow"):
                    o.nscbc_outflow(f, d, side, lo, hi, MaxMach, Lsc, P0)
                    bt.UpdateUsingFluxNSCBCOutflow(fl, d, side, bc, full, MaxMach, Lsc, P0)
                else:
                    o.nscbc_farfield(f, b, d, side, lo, hi, MaxMach, Lsc, P0)
                    bt.UpdateUsingFluxNSCBCFarField(fl, bcf, d, side, bc, MaxMach, Lsc, P0)
                got = fl.download(["Conserved_t"])["Conserved_t"]
                ref = f["Conserved_t"]
                assert np.isfinite(ref).all()
                changed = np.abs(on_plane(ref, d, side) - 3.0).max()
                assert changed > 0, (d, side, task)                 # the task did something
                e = rel_err(on_plane(got, d, side) - 3.0, on_plane(ref, d, side) - 3.0)
                assert e <= tol, (d, side, task, e)
                # nothing outside the boundary plane was touched
                mask = np.ones(ref.shape[:3], dtype=bool)
                idx = [slice(None)]*3; idx[2 - d] = 0 if side == MINUS else -1
                mask[tuple(idx)] = False
                assert (got[mask] == 3.0).all(), (d, side, task)


@pytest.mark.parametrize("variant", ["", "strict"], ids=["prod", "strict"])
@pytest.mark.parametrize("cfg", [CP, AIR], ids=["constprop", "air"])
def test_incoming_shock_tasks(cfg, variant):
    o, f, b = bounded_state(cfg, seed=1)
    s = htr.Solver(cfg, variant=variant)
    fl = to_device(s, f)
    bt = BoundaryTasks(s)
    tol = tol_for(cfg, variant)
    lo, hi = plane(1, PLUS)
    bc = Rect.make(lo, hi)
    X = np.linspace(1.0, 2.0, o.nspec); X /= X.sum()
    c = float(f["SoS"].mean()); T = float(f["temperature"].mean()); P = float(f["pressure"].mean())
    q = shock_params(N[0]//2, X, P, T, (2.0*c, -0.3*c, 0.0), 2.5*P, 1.4*T, (1.5*c, -0.6*c, 0.0))
    for MaxMach in (0.5, 1.2):
        f["Conserved_t"][...] = 3.0
        fl.t["Conserved_t"].fill_(3.0)
        o.incoming_shock(f, lo, hi, MaxMach, 0.7, q)
        bt.UpdateUsingFluxIncomingShock(fl, bc, MaxMach, 0.7, q)
        got = fl.download(["Conserved_t"])["Conserved_t"]
        e = rel_err(on_plane(got, 1, PLUS) - 3.0, on_plane(f["Conserved_t"], 1, PLUS) - 3.0)
        assert e <= tol, (MaxMach, e)
    # SetIncomingShockBC: primitives of the plane (outflow treatment + the three injected cells)
    names = ["pressure", "temperature", "MolarFracs", "velocity"]
    f["temperature"] *= 1.02                         # Newton start away from the solution, same on both sides
    fl.upload(f, ["temperature"])
    o.set_incoming_shock_bc(f, lo, hi, q)
    bt.SetIncomingShockBC(fl, bc, q)
    got = fl.download(names)
    for n in names:
        assert rel_err(on_plane(got[n], 1, PLUS), on_plane(f[n], 1, PLUS)) <= tol, n
    inj = on_plane(got["pressure"], 1, PLUS)[:, N[0]//2 - 1:N[0]//2 + 2]
    assert np.allclose(inj, [0.75*P + 0.25*2.5*P, 0.25*P + 0.75*2.5*P, 2.5*P], rtol=1e-14)


@pytest.mark.parametrize("variant", ["", "strict"], ids=["prod", "strict"])
@pytest.mark.parametrize("cfg", [CP, AIR], ids=["constprop", "air"])
def test_set_nscbc_bc_tasks(cfg, variant):
    o, f, b = bounded_state(cfg, seed=2)
    s = htr.Solver(cfg, variant=variant)
    fl = to_device(s, f)
    bcf = BCFields(fl); bcf.upload(b)
    bt = BoundaryTasks(s)
    tol = tol_for(cfg, variant)
    names = ["pressure", "temperature", "MolarFracs", "velocity"]
    Pbc = 1.3*float(f["pressure"].mean())
    for d in range(3):
        lo, hi = plane(d, MINUS)
        o.set_nscbc_inflow_bc(f, b, d, lo, hi, Pbc)
        bt.SetNSCBC_InflowBC(fl, bcf, d, Rect.make(lo, hi), Pbc)
        got = fl.download(names)
        sup = np.abs(on_plane(b["velocity_profile"], d, MINUS)[..., d]) >= on_plane(f["SoS"], d, MINUS)
        assert sup.any() and (~sup).any()            # both the imposed and the computed pressure occur
        assert (on_plane(got["pressure"], d, MINUS)[sup] == Pbc).all()
        for n in names:
            assert rel_err(on_plane(got[n], d, MINUS), on_plane(f[n], d, MINUS)) <= tol, (d, n)
        lo, hi = plane(d, PLUS)
        f["temperature"][...] = f["temperature"]*1.03
        fl.upload(f, ["temperature"])
        o.set_nscbc_outflow_bc(f, lo, hi)
        bt.SetNSCBC_OutflowBC(fl, Rect.make(lo, hi))
        got = fl.download(names)
        for n in names:
            assert rel_err(on_plane(got[n], d, PLUS), on_plane(f[n], d, PLUS)) <= max(tol, 1e-12), (d, n)
