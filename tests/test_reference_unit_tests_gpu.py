"""GPU: the reference's own unit tests (unitTests/{variablesTest,chemTest,cflTest}) replayed through the C ABI with the
reference's golden values and tolerances -- the same checks a maintainer would run after swapping in the B200 objects."""
import numpy as np
import pytest
import torch
import htr_solver_b200 as htr
from test_oracle_goldens_cpu import _unit_test_air_cfg

pytestmark = pytest.mark.gpu


def _fill(fl, name, value):
    t = fl.t[name]
    t.copy_(torch.as_tensor(np.broadcast_to(np.asarray(value, dtype=np.float64), t.shape).copy()))


def test_variablestest_gpu():
    """unitTests/variablesTest/variablesTest.rg:56-72,119-200: 1e-8."""
    cfg, q = _unit_test_air_cfg()
    s = htr.Solver(cfg)
    fl = s.new_fluid(16, 16, 16)
    P, T = 101325.0, 5000.0
    Xi = np.array([0.4, 0.2, 0.15, 0.15, 0.1]); v = np.array([1.0, 2.0, 3.0])
    _fill(fl, "pressure", P / 101325.0); _fill(fl, "temperature", T / 300.0)
    _fill(fl, "MolarFracs", Xi); _fill(fl, "velocity", v / q["uRef"])
    s.UpdatePropertiesFromPrimitive(fl)
    d = fl.download(["rho", "mu", "lam", "Di", "SoS"])
    tol = 1e-8
    assert np.all(np.abs(d["rho"] * q["rhoRef"] / 6.2899525132668e-02 - 1.0) < tol)
    assert np.all(np.abs(d["mu"] * q["muRef"] / 1.2424432854120e-04 - 1.0) < tol)
    assert np.all(np.abs(d["lam"] * q["lamRef"] / 2.4040406505225e-01 - 1.0) < tol)
    eDi = np.array([2.5146942638910e-03, 2.4389378958326e-03, 2.4551032686824e-03, 3.6168376636972e-03, 3.9165741891924e-03])
    assert np.all(np.abs(d["Di"] * q["DiRef"] / eDi - 1.0) < tol)
    assert np.all(np.abs(d["SoS"] * q["uRef"] / 1.4518745895533e+03 - 1.0) < tol)
    s.UpdateConservedFromPrimitive(fl)
    eC = np.array([2.3342371515176e-02, 1.3331617683677e-02, 9.3760512904744e-03, 4.3766946590956e-03, 3.3329044209192e-03,
                   1.8268107194243e-04, 3.6536214388485e-04, 5.4804321582728e-04, 5.3385045774457e+00])
    assert np.all(np.abs(fl.download(["Conserved"])["Conserved"] / eC - 1.0) < tol)
    for k in ("pressure", "MolarFracs", "velocity"):
        fl.t[k].zero_()
    _fill(fl, "temperature", 500.0 / 300.0)
    s.UpdatePrimitiveFromConserved(fl)
    d = fl.download(["pressure", "temperature", "MolarFracs", "velocity"])
    assert np.all(np.abs(d["pressure"] * 101325.0 / P - 1.0) < tol)
    assert np.all(np.abs(d["temperature"] * 300.0 / T - 1.0) < tol)
    assert np.all(np.abs(d["MolarFracs"] / Xi - 1.0) < tol)
    assert np.all(np.abs(d["velocity"] * q["uRef"] / v - 1.0) < tol)


def test_chemtest_gpu():
    """unitTests/chemTest/chemTest.rg:45-49,118-147: explicit sources and one implicit step, 1e-5."""
    from parity import oracle_for
    cfg, q = _unit_test_air_cfg()
    s = htr.Solver(cfg)
    o = oracle_for(cfg)
    P, T = 101325.0, 5000.0
    Yi = np.array([0.78, 0.22, 1.0e-60, 1.0e-60, 1.0e-60])
    r = o.mix_eval(P / 101325.0, T / 300.0, Yi)      # only used to build the initial state, like InitializeCell
    fl = s.new_fluid(2, 2, 2)
    _fill(fl, "pressure", P / 101325.0); _fill(fl, "temperature", T / 300.0)
    _fill(fl, "MassFracs", Yi); _fill(fl, "MolarFracs", r["Xi"]); _fill(fl, "rho", r["rho"])
    C = np.zeros(9); C[:5] = r["rho"] * Yi; C[8] = r["rho"] * r["e"]
    _fill(fl, "Conserved", C)
    s.AddChemistrySources(fl)
    Eprod = np.array([-1.658328e-01, -1.445940e+03, 1.332770e-54, 1.658328e-01, 1.445940e+03, 0.0, 0.0, 0.0, 0.0])
    got = fl.download(["Conserved_t"])["Conserved_t"] * q["wiRef"]
    err = np.where(Eprod != 0, (got - Eprod) / np.where(Eprod != 0, Eprod, 1.0), got - Eprod)
    assert np.all(np.abs(err) < 1e-5), err.reshape(-1, 9)[0]
    dt = 1e-6 * np.sqrt(101325.0 / q["rhoRef"])
    s.UpdateChemistry(fl, dt)                         # Conserved_t still holds the sources, as in the reference test
    Cres = np.array([5.460212e-02, 1.296359e-02, 2.994386e-04, 1.502734e-05, 2.321006e-03, 0.0, 0.0, 0.0, 3.141423e+05])
    got = fl.download(["Conserved"])["Conserved"] * q["rhoRef"]
    got[..., 8] *= q["eRef"]
    err = np.where(Cres != 0, (got - Cres) / np.where(Cres != 0, Cres, 1.0), got - Cres)
    assert np.all(np.abs(err) < 1e-5), err.reshape(-1, 9)[0]


def test_cfltest_gpu():
    """unitTests/cflTest/cflTest.rg:83-113: 11, 40, 11.42857, 40 (1e-3)."""
    s = htr.Solver({"type": "ConstPropMix", "gasConstant": 1.0, "gamma": 1.4, "prandtl": 0.71,
                    "viscosityModel": {"type": "Constant", "Visc": 0.0}})
    fl = s.new_fluid(2, 2, 2)
    for k in ("dcsi_d", "deta_d", "dzet_d", "rho", "MassFracs", "MolarFracs"):
        fl.t[k].fill_(1.0)
    fl.t["velocity"].fill_(1.0); fl.t["SoS"].fill_(10.0)
    assert abs(s.CalculateMaxSpectralRadius(fl) / 11.0 - 1.0) < 1e-3
    fl.t["velocity"].zero_(); fl.t["mu"].fill_(10.0)
    assert abs(s.CalculateMaxSpectralRadius(fl) / 40.0 - 1.0) < 1e-3
    fl.t["mu"].zero_(); fl.t["lam"].fill_(10.0)
    assert abs(s.CalculateMaxSpectralRadius(fl) / 11.42857 - 1.0) < 1e-3
    fl.t["lam"].zero_(); fl.t["Di"].fill_(10.0)
    assert abs(s.CalculateMaxSpectralRadius(fl) / 40.0 - 1.0) < 1e-3
