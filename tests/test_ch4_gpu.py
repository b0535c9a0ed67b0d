"""GPU parity for the CH4 one-step mixture (src/Mixtures/CH41StMix.hpp: 4 species, one irreversible reaction with
specified forward orders, FWD_ORDERS) against the reference's CPU tasks: variables, properties, chemistry sources,
the full right-hand side through every htr_rhs path, and two RK3 time steps."""
import numpy as np
import pytest
import htr_solver_b200 as htr
from htr_solver_b200 import synthetic
from htr_solver_b200.core import RK_C
from parity import oracle_for, host_state, to_device, rel_err, euler_scale, scaled_err, TOL_CHEM

pytestmark = pytest.mark.gpu
CH4 = synthetic.CH41ST_3DPERIODIC
HOT = {"species_wave": 1.0, "front": 0.2, "T0": 5.0, "p0": 1.0, "U": 0.3}     # ~1500 K: the reaction is active


def _sensors(o, f):
    o.ducros_sensor(f, 1.0)
    for d in range(3):
        o.shock_sensor(f, d)


def test_ch4_variables_properties_sources():
    o = oracle_for(CH4)
    f = host_state(o, (20, 16, 12), CH4, **HOT)
    s = htr.Solver(CH4)
    fl = to_device(s, f)
    # properties from primitives
    for k in ("rho", "mu", "lam", "Di", "SoS", "MassFracs"):
        fl.t[k].zero_()
    s.UpdatePropertiesFromPrimitive(fl)
    got = fl.download(["rho", "mu", "lam", "Di", "SoS", "MassFracs"])
    for k, v in got.items():
        assert rel_err(v, f[k]) <= TOL_CHEM, k
    # primitives from conserved, from a perturbed temperature guess
    fl.t["temperature"].mul_(1.03)
    ref = {k: f[k].copy() for k in ("temperature", "pressure", "MolarFracs", "velocity")}
    f["temperature"] *= 1.03
    o.update_primitive_from_conserved(f)
    s.UpdatePrimitiveFromConserved(fl)
    got = fl.download(list(ref))
    for k in ref:
        assert rel_err(got[k], f[k]) <= TOL_CHEM, k
    for k in ref:
        f[k][...] = ref[k]
    fl.upload(ref)
    # chemistry sources
    f["Conserved_t"][...] = 0.0
    fl.t["Conserved_t"].zero_()
    o.add_chemistry_sources(f)
    s.AddChemistrySources(fl)
    w = f["Conserved_t"][..., :4]
    assert np.abs(w).max() > 1e-6, "the reaction must be active in this state"
    assert rel_err(fl.download(["Conserved_t"])["Conserved_t"][..., :4], w) <= TOL_CHEM


@pytest.mark.parametrize("variant", ["", "strict"], ids=["prod", "strict"])
def test_ch4_full_rhs_all_paths(variant):
    o = oracle_for(CH4)
    f = host_state(o, (45, 23, 19), CH4, **HOT)
    _sensors(o, f)
    f["shockSensorX"][:, :, 7::11] = 0; f["shockSensorZ"][3::8, :, :] = 0
    s = htr.Solver(CH4, variant=variant)
    fl = to_device(s, f)
    out = {}
    try:
        for path, (fast, march) in {1: (1, 3), 2: (1, 0), 3: (1, 1), 0: (0, 0)}.items():
            s.set_option("fast", fast); s.set_option("march", march)
            fl.t["Conserved_t"].fill_(-3.0)
            s.UpdateDerivatives(fl, "Hybrid", diffusion=True, chemistry=True, RK_coeffs=(0.75, 0.25), deltaTime=1e-3)
            out[path] = fl.download(["Conserved_t"])["Conserved_t"]
    finally:
        s.set_option("fast", 1); s.set_option("march", 1)
    o.rhs(f, "Hybrid", diffusion=True, chemistry=True, RK0=0.75, RK1=0.25, dt=1e-3)
    scale = np.maximum.reduce([euler_scale(f, d) for d in range(3)])
    ref = f["Conserved_t"].copy()
    f["Conserved_t"][...] = 0.0
    o.add_chemistry_sources(f)
    scale = scale + np.abs(f["Conserved_t"]).reshape(-1, scale.size).max(axis=0)
    for k in (1, 2, 3):
        if variant == "strict":
            assert np.array_equal(out[k], out[0]), k
        assert scaled_err(out[k], ref, scale) <= TOL_CHEM, k


def test_ch4_time_step():
    o = oracle_for(CH4)
    f = host_state(o, (24, 20, 16), CH4, **HOT)
    s = htr.Solver(CH4)
    fl = to_device(s, f)
    dt = 0.4/o.max_spectral_radius(f)
    f["Conserved_t"][...] = 0.0
    o.add_chemistry_sources(f)
    w = np.abs(f["Conserved_t"][..., :4]).max(axis=-1)
    dt = min(dt, 0.005*float((f["rho"]/np.maximum(w, 1e-300)).min()))
    names = ["Conserved", "temperature", "pressure", "velocity", "MolarFracs", "rho", "mu", "lam", "Di", "SoS", "MassFracs"]
    for step in range(2):
        _sensors(o, f)
        f["Conserved_old"] = f["Conserved"].copy()
        for stage in (1, 2, 3):
            c1, c2, c3 = RK_C[stage - 1]
            o.rhs(f, "Hybrid", diffusion=True, chemistry=True, RK0=c1, RK1=c2, dt=dt*c3)
            f["Conserved"] = np.ascontiguousarray((c1*f["Conserved_old"] + c2*f["Conserved"]) + (c3*f["Conserved_t"])*dt)
            o.update_primitive_from_conserved(f)
            o.update_properties_from_primitive(f)
        s.TimeStep(fl, "Hybrid", dt, diffusion=True, chemistry=True, vorticityScale=1.0)
        got = fl.download(names)
        for k in names:
            assert rel_err(got[k], f[k]) <= TOL_CHEM, (step, k)
