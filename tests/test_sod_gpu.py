"""GPU: BASELINE config C1, testcases/SodProblem/100.json -- the 1-D Sod shock tube on 100 x 1 x 1 cells, ConstPropMix
(R = 1, gamma = 1.4, inviscid), Hybrid scheme (vorticityScale 1e-6), Dirichlet boundaries in x (staggered boundary nodes
L_S Lp1_S Lp2_S Std ... Rm3_S Rm2_S Rm1_S R_S, one boundary cell per side), 200 steps of dt = 1e-3 to t = 0.2.

The Dirichlet ghost fill of a constant profile leaves the two boundary cells at their initial state, so the case runs with
the hot-path kernels alone: sensors on the whole box, per stage UpdateDerivatives + UpdateVars/primitive/property recovery
on the 100 interior cells.  Checked (i) against the reference's CPU task chain on the same inputs, step for step, and
(ii) against the reference's exact solution (tests/golden/sod_exact.txt = testcases/SodProblem/exact.txt)."""
import os
import numpy as np
import pytest
import htr_solver_b200 as htr
from htr_solver_b200.core import RK_C
from oracle_lib import alloc_fields
from parity import oracle_for, rel_err, TOL_FLUX

pytestmark = pytest.mark.gpu
SOD = {"type": "ConstPropMix", "gasConstant": 1.0, "gamma": 1.4, "prandtl": 0.71, "viscosityModel": {"type": "Constant", "Visc": 0.0}}
Std, L_S, Lp1_S, Lp2_S, Rm3_S, Rm2_S, Rm1_S, R_S = range(8)
NAMES = ["Conserved", "temperature", "pressure", "velocity", "MolarFracs", "rho", "mu", "lam", "Di", "SoS", "MassFracs"]


def sod_state(o, nx=100):
    """Host fields of the (nx + 2) x 1 x 1 box: one staggered boundary cell per side (src/prometeo_grid.rg:296-316: its
    centre is the boundary node), metrics from the reference's InitializeMetricTask, state of src/prometeo_init.rg:235-255."""
    n = nx + 2
    f = alloc_fields(n, 1, 1, 1)
    t = np.zeros(n, dtype=np.int32)
    t[:3] = [L_S, Lp1_S, Lp2_S]; t[-4:] = [Rm3_S, Rm2_S, Rm1_S, R_S]
    f["nType_x"][0, 0, :] = t
    nodes = np.arange(nx + 1) / nx
    xc = np.concatenate([[nodes[0]], 0.5*(nodes[1:] + nodes[:-1]), [nodes[-1]]])
    cc = np.zeros((1, 1, n, 3)); cc[0, 0, :, 0] = xc; cc[..., 1] = 0.5; cc[..., 2] = 0.5
    o.initialize_metric(f, cc, [1.0, 1.0, 1.0])
    left = xc < 0.5
    f["pressure"][0, 0, :] = np.where(left, 1.0, 0.1)
    f["temperature"][0, 0, :] = np.where(left, 1.0, 0.8)
    f["MolarFracs"][...] = 1.0
    o.update_properties_from_primitive(f)
    o.update_conserved_from_primitive(f)
    f["Conserved_old"][...] = f["Conserved"]
    return f, xc


def gpu_step(s, fl, dt, full, inner):
    s.UpdateDucrosSensor(fl, 1.0e-6, tile=full, bounds=full)
    s.UpdateShockSensor(fl, 0, tile=full, bounds=full)
    for d in (1, 2):
        s.UpdateShockSensor(fl, d, tile=full, bounds=full)
    s.InitializeTemporaries(fl, tile=full)
    for stage in (1, 2, 3):
        c1, c2, c3 = RK_C[stage - 1]
        s.UpdateDerivatives(fl, "Hybrid", True, False, (c1, c2), dt*c3, tile=inner, bounds=full)
        s.UpdateVarsAndRecover(fl, stage, dt, False, tile=inner)


def test_sod_shock_tube():
    o = oracle_for(SOD)
    f, xc = sod_state(o)
    n = f["rho"].shape[2]
    s = htr.Solver(SOD)
    fl = s.new_fluid(n, 1, 1)
    fl.upload(f)
    full = htr.Rect.make((0, 0, 0), (n - 1, 0, 0))
    inner = htr.Rect.make((1, 0, 0), (n - 2, 0, 0))
    dt = 1.0e-3
    bc = {k: f[k][:, :, [0, -1]].copy() for k in NAMES}
    # (i) five steps against the reference's task chain (boundary cells re-imposed after every stage: Dirichlet, constant)
    for step in range(5):
        o.ducros_sensor(f, 1.0e-6)
        for d in range(3):
            o.shock_sensor(f, d)
        f["Conserved_old"] = f["Conserved"].copy()
        for stage in (1, 2, 3):
            c1, c2, c3 = RK_C[stage - 1]
            o.rhs(f, "Hybrid", diffusion=True, chemistry=False, RK0=c1, RK1=c2, dt=dt*c3)
            f["Conserved"] = np.ascontiguousarray((c1*f["Conserved_old"] + c2*f["Conserved"]) + (c3*f["Conserved_t"])*dt)
            o.update_primitive_from_conserved(f)
            o.update_properties_from_primitive(f)
            for k in NAMES:
                f[k][:, :, [0, -1]] = bc[k]
        gpu_step(s, fl, dt, full, inner)
        got = fl.download(NAMES)
        for k in ("Conserved", "pressure", "temperature", "velocity", "rho"):
            assert rel_err(got[k], f[k]) <= 1e-11, (step, k, rel_err(got[k], f[k]))
    # (ii) to t = 0.2 and against the exact solution
    for step in range(5, 200):
        gpu_step(s, fl, dt, full, inner)
    got = fl.download(["rho", "velocity", "pressure"])
    ex = np.loadtxt(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "sod_exact.txt"))
    xe = 0.5*(ex[:, 0] + 1.0)
    xi = xc[1:-1]
    err = {}
    for name, col, val in (("rho", 1, got["rho"][0, 0, 1:-1]), ("u", 2, got["velocity"][0, 0, 1:-1, 0]), ("p", 3, got["pressure"][0, 0, 1:-1])):
        exact = np.interp(xi, xe, ex[:, col])
        err[name] = float(np.sqrt(np.mean((val - exact)**2)))
        assert np.isfinite(val).all()
    print("Sod 100 cells, t = 0.2: L2 error vs exact solution", {k: f"{v:.3e}" for k, v in err.items()})
    # a 6th-order hybrid scheme on 100 cells resolves the three waves within a few cells: O(1e-2) L2 errors
    assert err["rho"] < 1.5e-2 and err["u"] < 4.0e-2 and err["p"] < 1.2e-2, err      # the reference chain itself: 1.29e-2, 3.59e-2, 8.8e-3
    # boundary cells untouched, plateau values of the exact solution recovered between the waves
    for k in ("rho", "pressure"):
        assert np.array_equal(got[k][:, :, [0, -1]], bc[k])
    mid = (xi > 0.55) & (xi < 0.65)        # between the expansion tail and the contact
    assert abs(got["pressure"][0, 0, 1:-1][mid].mean() - 0.30313) < 2e-3
    assert abs(got["velocity"][0, 0, 1:-1, 0][mid].mean() - 0.92745) < 5e-3
