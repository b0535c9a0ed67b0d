"""CPU tests that PIN the oracle (oracle/_ref: the reference's own headers compiled in place behind a stub legion.h)
on every known-answer vector the reference's unit tests hold for the hot path:

  unitTests/mixTest/mixTest.rg:130-169        AirMix thermodynamics, transport, kinetics at 750 K (printed to 7 digits)
  unitTests/variablesTest/variablesTest.rg    rho, mu, lam, D_i, SoS, Conserved at 5000 K (14 digits), and the inverse
  unitTests/chemTest/chemTest.rg:45-49        production rates and the state after one implicit step
  unitTests/cflTest/cflTest.rg:83-113         the four branches of the spectral radius
  unitTests/mathUtilsTest/CPPmathUtilsTest.cc LU solve (1e-12) and the Rosenbrock integrator (1e-6)

The tolerances are the reference's own, or tighter where the golden is printed with enough digits.
"""
import numpy as np
import pytest
from oracle_lib import Oracle, alloc_fields

R = 8.3144598


def have_oracle(mix):
    try:
        Oracle(mix)
        return True
    except FileNotFoundError:
        return False


needs_air = pytest.mark.skipif(not have_oracle("AirMix"), reason="oracle/_ref not built")
needs_cp = pytest.mark.skipif(not have_oracle("ConstPropMix"), reason="oracle/_ref not built")

# unitTests/mixTest/AirMix.json: "XiRef": N2 0.78, O2 0.22
MIXTEST_AIR = {"type": "AirMix", "LRef": 1.0, "PRef": 101325.0, "TRef": 300.0,
               "XiRef": {"Species": [{"Name": "N2", "MolarFrac": 0.78}, {"Name": "O2", "MolarFrac": 0.22}]}}


def refs(PRef, TRef, MixWRef, LRef=1.0):
    rhoRef = PRef * MixWRef / (R * TRef)
    return dict(rhoRef=rhoRef, eRef=PRef / rhoRef, uRef=np.sqrt(PRef / rhoRef), CpRef=R / MixWRef,
                muRef=np.sqrt(PRef * rhoRef) * LRef, lamRef=np.sqrt(PRef * rhoRef) * LRef * R / MixWRef,
                DiRef=np.sqrt(PRef / rhoRef) * LRef, wiRef=np.sqrt(PRef * rhoRef) / LRef)


@needs_air
def test_mixtest_airmix_all_goldens():
    """unitTests/mixTest/mixTest.rg:130-169 (expected values) and :756-890 (checks, 1e-3).  The goldens are printed with
    7 significant digits; the oracle reproduces every one of them to that precision."""
    o = Oracle("AirMix"); o.init_mix(MIXTEST_AIR)
    P, T = 101325.0, 750.0
    MixWRef = 2.889018800000000e-02
    q = refs(101325.0, 300.0, MixWRef)
    r = o.mix_eval(P / 101325.0, T / 300.0, [0.2] * 5)
    W = np.array([o.lib.htr_oracle_species_W(i) for i in range(5)])
    assert np.allclose(W, [2 * 14.0067e-3, 2 * 15.9994e-3, 30.00610e-03, 14.0067e-03, 15.9994e-03], rtol=1e-12)
    EMixW = 2.135186e-02
    Erho = P * EMixW / (R * T)
    tol = 1e-6          # 7 printed digits
    def chk(got, exp, what):
        assert abs(got / exp - 1.0) < tol, (what, got, exp)
    chk(1.0 / np.sum(0.2 / W), EMixW, "MixW")
    for i, e in enumerate([1.524403e-01, 1.334541e-01, 1.423168e-01, 3.048806e-01, 2.669082e-01]):
        chk(r["Xi"][i], e, f"Xi[{i}]")
    chk(r["rho"] * q["rhoRef"], Erho, "rho")
    chk(r["cp"] * q["CpRef"], 1.206133e+03, "cp")
    chk(r["h"] * q["eRef"], 1.100436e+07, "h")
    chk(r["e"] * q["eRef"], 1.100436e+07 - P / Erho, "e")
    chk(r["mu"] * q["muRef"], 3.723077e-05, "mu")
    chk(r["lam"] * q["lamRef"], 6.736994e-02, "lam")
    chk(r["gamma"], 1.476782e+00, "gamma")
    chk(r["SoS"] * q["uRef"], 6.567317e+02, "SoS")
    for i, e in enumerate([1.192983e-04, 1.174019e-04, 1.162986e-04, 1.820668e-04, 1.882347e-04]):
        chk(r["Di"][i] * q["DiRef"], e, f"Di[{i}]")
    for i, e in enumerate([3.413124e+07, 2.664566e+04, -3.578332e+07, -1.742776e+07, 1.905320e+07]):
        chk(r["w"][i] * q["wiRef"], e, f"w[{i}]")
    chk(r["dpde"] * q["rhoRef"], Erho * (1.476782e+00 - 1.0), "dpde")
    for i, e in enumerate([5.206698e+06, 5.184248e+06, 3.742883e+06, -1.064452e+07, -2.029053e+06]):
        chk(r["dpdrhoi"][i] * q["eRef"], e, f"dpdrhoi[{i}]")
    for i, e in enumerate([4.812927e+05, 4.425045e+05, 3.505703e+06, 3.441706e+07, 1.617526e+07]):
        chk(o.lib.htr_oracle_species_enthalpy(i, T / 300.0) * q["eRef"], e, f"hsp[{i}]")
    # Newton inversion e -> T from a 5 % perturbed guess (mixTest.rg: GetTFromInternalEnergy check)
    chk(r["T_from_e"], T / 300.0, "T(e)")
    chk(r["MixW_Xi"], r["MixW"], "MixW(Xi)")


def _unit_test_air_cfg():
    """The mixture of variablesTest.rg / chemTest.rg: reference state from Y_O2 = 0.22, Y_N2 = 0.78."""
    MixWRef = 1.0 / (0.78 / 28.0134e-3 + 0.22 / (2 * 15.999e-3))
    cfg = {"type": "AirMix", "LRef": 1.0, "PRef": 101325.0, "TRef": 300.0,
           "XiRef": {"Species": [{"Name": "O2", "MolarFrac": MixWRef * 0.22 / (2 * 15.999e-3)},
                                 {"Name": "N2", "MolarFrac": MixWRef * 0.78 / 28.0134e-3}]}}
    return cfg, refs(101325.0, 300.0, MixWRef)


@needs_air
def test_variablestest_goldens():
    """unitTests/variablesTest/variablesTest.rg:56-72 (14-digit goldens), checks :119-200 with the reference's 1e-8:
    UpdatePropertiesFromPrimitive, UpdateConservedFromPrimitive and back through UpdatePrimitiveFromConserved."""
    cfg, q = _unit_test_air_cfg()
    o = Oracle("AirMix"); o.init_mix(cfg)
    P, T = 101325.0, 5000.0
    Xi = np.array([0.4, 0.2, 0.15, 0.15, 0.1]); v = np.array([1.0, 2.0, 3.0])
    f = alloc_fields(3, 2, 2, 5)
    f["pressure"][...] = P / 101325.0
    f["temperature"][...] = T / 300.0
    f["MolarFracs"][...] = Xi
    f["velocity"][...] = v / q["uRef"]
    o.update_properties_from_primitive(f)
    tol = 1e-8
    assert np.all(np.abs(f["rho"] * q["rhoRef"] / 6.2899525132668e-02 - 1.0) < tol)
    assert np.all(np.abs(f["mu"] * q["muRef"] / 1.2424432854120e-04 - 1.0) < tol)
    assert np.all(np.abs(f["lam"] * q["lamRef"] / 2.4040406505225e-01 - 1.0) < tol)
    eDi = np.array([2.5146942638910e-03, 2.4389378958326e-03, 2.4551032686824e-03, 3.6168376636972e-03, 3.9165741891924e-03])
    assert np.all(np.abs(f["Di"] * q["DiRef"] / eDi - 1.0) < tol)
    assert np.all(np.abs(f["SoS"] * q["uRef"] / 1.4518745895533e+03 - 1.0) < tol)
    o.update_conserved_from_primitive(f)
    eC = np.array([2.3342371515176e-02, 1.3331617683677e-02, 9.3760512904744e-03, 4.3766946590956e-03, 3.3329044209192e-03,
                   1.8268107194243e-04, 3.6536214388485e-04, 5.4804321582728e-04, 5.3385045774457e+00])
    assert np.all(np.abs(f["Conserved"] / eC - 1.0) < tol)
    # and back, from the test's restart temperature Tres = 500 K as the Newton guess
    f["pressure"][...] = 0.0; f["MolarFracs"][...] = 0.0; f["velocity"][...] = 0.0
    f["temperature"][...] = 500.0 / 300.0
    o.update_primitive_from_conserved(f)
    assert np.all(np.abs(f["pressure"] * 101325.0 / P - 1.0) < tol)
    assert np.all(np.abs(f["temperature"] * 300.0 / T - 1.0) < tol)
    assert np.all(np.abs(f["MolarFracs"] / Xi - 1.0) < tol)
    assert np.all(np.abs(f["velocity"] * q["uRef"] / v - 1.0) < tol)


@needs_air
def test_chemtest_goldens():
    """unitTests/chemTest/chemTest.rg:45-49: AddChemistrySources (Eprod) and UpdateChemistry over 1e-6 s (Cres), both to
    the reference's 1e-5."""
    cfg, q = _unit_test_air_cfg()
    o = Oracle("AirMix"); o.init_mix(cfg)
    P, T = 101325.0, 5000.0
    Yi = np.array([0.78, 0.22, 1.0e-60, 1.0e-60, 1.0e-60])
    r = o.mix_eval(P / 101325.0, T / 300.0, Yi)
    f = alloc_fields(2, 2, 2, 5)
    f["pressure"][...] = P / 101325.0
    f["temperature"][...] = T / 300.0
    f["MassFracs"][...] = Yi
    f["MolarFracs"][...] = r["Xi"]
    f["rho"][...] = r["rho"]
    f["Conserved"][..., :5] = r["rho"] * Yi
    f["Conserved"][..., 8] = r["rho"] * r["e"]
    o.add_chemistry_sources(f)
    Eprod = np.array([-1.658328e-01, -1.445940e+03, 1.332770e-54, 1.658328e-01, 1.445940e+03, 0.0, 0.0, 0.0, 0.0])
    got = f["Conserved_t"] * q["wiRef"]
    err = np.where(Eprod != 0, (got - Eprod) / np.where(Eprod != 0, Eprod, 1.0), got - Eprod)
    assert np.all(np.abs(err) < 1e-5), err.reshape(-1, 9)[0]
    # implicit step.  The reference test does NOT clear Conserved_t in between (chemTest.rg:118-135): the production
    # rates just added act as the frozen explicit part of the split right-hand side (Conserved_t_old = Conserved_t).
    dt = 1e-6 * np.sqrt(101325.0 / q["rhoRef"]) / 1.0
    o.update_chemistry(f, dt)
    Cres = np.array([5.460212e-02, 1.296359e-02, 2.994386e-04, 1.502734e-05, 2.321006e-03, 0.0, 0.0, 0.0, 3.141423e+05])
    got = f["Conserved"] * q["rhoRef"]
    got[..., 8] *= q["eRef"]
    err = np.where(Cres != 0, (got - Cres) / np.where(Cres != 0, Cres, 1.0), got - Cres)
    assert np.all(np.abs(err) < 1e-5), err.reshape(-1, 9)[0]


@needs_cp
def test_cfltest_goldens():
    """unitTests/cflTest/cflTest.rg:83-113: acoustic 11, momentum diffusion 40, energy diffusion 11.42857,
    species diffusion 40 (1e-3)."""
    o = Oracle("ConstPropMix")
    o.init_mix({"type": "ConstPropMix", "gasConstant": 1.0, "gamma": 1.4, "prandtl": 0.71,
                "viscosityModel": {"type": "Constant", "Visc": 0.0}})
    f = alloc_fields(2, 2, 2, 1)
    for k in ("dcsi_d", "deta_d", "dzet_d", "rho"):
        f[k][...] = 1.0
    f["MassFracs"][...] = 1.0; f["MolarFracs"][...] = 1.0
    f["velocity"][...] = 1.0; f["SoS"][...] = 10.0
    assert abs(o.max_spectral_radius(f) / 11.0 - 1.0) < 1e-3
    f["velocity"][...] = 0.0
    f["mu"][...] = 10.0
    assert abs(o.max_spectral_radius(f) / 40.0 - 1.0) < 1e-3
    f["mu"][...] = 0.0
    f["lam"][...] = 10.0
    assert abs(o.max_spectral_radius(f) / 11.42857 - 1.0) < 1e-3
    f["lam"][...] = 0.0
    f["Di"][...] = 10.0
    assert abs(o.max_spectral_radius(f) / 40.0 - 1.0) < 1e-3


@needs_air
def test_mathutils_goldens():
    """unitTests/mathUtilsTest/CPPmathUtilsTest.cc:6-55: LUdec<3> (1e-12) and Rosenbrock<3> on the stiff 3-equation
    kinetics problem (1e-6)."""
    o = Oracle("AirMix")
    A = np.array([[4.0, 3.0, 0.0], [3.0, 4.0, -1.0], [0.0, -1.0, 4.0]])
    b = np.array([24.0, 30.0, -24.0]); x = np.zeros(3)
    o.lib.htr_oracle_lu3_solve(A.ctypes.data, b.ctypes.data, x.ctypes.data)
    assert np.all(np.abs(x - np.array([3.0, 4.0, -5.0])) < 1e-12)
    y = np.array([1.0, 1.0, 0.0])
    o.lib.htr_oracle_rosenbrock3(y.ctypes.data, 1.0e-3, 25.0)
    assert np.all(np.abs(1.0 - y / np.array([7.818640e-01, 1.218133e+00, -2.655799e-06])) < 1e-6)


@pytest.mark.skipif(not have_oracle("CH41StMix"), reason="oracle/_ref not built")
def test_mixtest_ch41stmix_goldens():
    """unitTests/mixTest/mixTest.rg:173-215 (CH41StMix: T = 750 K, P = 1 atm, Y_i = 0.25; reference state pure O2):
    thermodynamics, transport and the one-step kinetics with specified forward orders, to the 7 printed digits."""
    cfg = {"type": "CH41StMix", "LRef": 1.0, "PRef": 101325.0, "TRef": 300.0,
           "XiRef": {"Species": [{"Name": "O2", "MolarFrac": 1.0}]}}
    o = Oracle("CH41StMix"); o.init_mix(cfg)
    P, T = 101325.0, 750.0
    q = refs(101325.0, 300.0, 3.1998800000000e-02)
    r = o.mix_eval(1.0, T / 300.0, [0.25] * 4)
    tol = 1e-6
    def chk(got, exp, what):
        assert abs(got / exp - 1.0) < tol, (what, got, exp)
    EMixW = 2.328035e-02
    Erho = P * EMixW / (R * T)
    W = np.array([o.lib.htr_oracle_species_W(i) for i in range(4)])
    assert np.allclose(W, [16.04206e-03, 2 * 15.9994e-3, 44.00950e-03, 18.01508e-03], rtol=1e-7)
    chk(1.0 / np.sum(0.25 / W), EMixW, "MixW")
    for i, e in enumerate([3.628018e-01, 1.818846e-01, 1.322462e-01, 3.230675e-01]):
        chk(r["Xi"][i], e, f"Xi[{i}]")
    chk(r["rho"] * q["rhoRef"], Erho, "rho")
    chk(r["cp"] * q["CpRef"], 2.032494e+03, "cp")
    chk(r["h"] * q["eRef"], -5.965912e+06, "h")
    chk(r["mu"] * q["muRef"], 3.108639e-05, "mu")
    chk(r["lam"] * q["lamRef"], 8.767210e-02, "lam")
    chk(r["gamma"], 1.213177e+00, "gamma")
    chk(r["SoS"] * q["uRef"], 5.700526e+02, "SoS")
    for i, e in enumerate([1.343289e-04, 1.007327e-04, 7.760974e-05, 1.350278e-04]):
        chk(r["Di"][i] * q["DiRef"], e, f"Di[{i}]")
    for i, e in enumerate([-2.658098e+00, -1.060412e+01, 7.292178e+00, 5.970037e+00]):
        chk(r["w"][i] * q["wiRef"], e, f"w[{i}]")
    for i, e in enumerate([-1.550405e+05, -1.186804e+06, 6.509754e+05, 1.762304e+06]):
        chk(r["dpdrhoi"][i] * q["eRef"], e, f"dpdrhoi[{i}]")
    for i, e in enumerate([-3.294307e+06, 4.425045e+05, -8.481094e+06, -1.253075e+07]):
        chk(o.lib.htr_oracle_species_enthalpy(i, T / 300.0) * q["eRef"], e, f"hsp[{i}]")


@pytest.mark.skipif(not have_oracle("CH4_30SpMix"), reason="oracle/_ref not built")
def test_mixtest_ch4_30spmix_goldens():
    """unitTests/mixTest/mixTest.rg:215-296 (CH4_30SpMix: 30 species, 184 reactions with third bodies and fall-off,
    T = 1500 K, P = 1 atm, a lean partially burnt methane/air composition; reference state pure O2).  Oracle only: the
    product library of this mixture is not built yet, this pins the reference build it will be tested against."""
    cfg = {"type": "CH4_30SpMix", "LRef": 1.0, "PRef": 101325.0, "TRef": 300.0,
           "XiRef": {"Species": [{"Name": "O2", "MolarFrac": 1.0}]}}
    o = Oracle("CH4_30SpMix"); o.init_mix(cfg)
    assert o.nspec == 30
    P, T = 101325.0, 1500.0
    q = refs(101325.0, 300.0, 3.1998800000000e-02)
    Y = [5.44971e-04, 1e-60, 1e-60, 2.15391e-01, 1e-60, 5.62779e-03, 2.33245e-05, 1e-60, 1e-60, 1e-60,
         1e-60, 1e-60, 1e-60, 5.27455e-02, 1.49287e-03, 5.98395e-04, 1e-60, 6.19628e-05, 1e-60, 1e-60,
         1.23090e-05, 1e-60, 1e-60, 1.22500e-05, 1e-60, 3.12909e-05, 1e-60, 1e-60, 1e-60, 7.23443e-01]
    r = o.mix_eval(1.0, T / 300.0, Y)
    tol = 1e-5      # the goldens are printed with 7 digits, the mass fractions with 6
    def chk(got, exp, what, t=tol):
        assert abs(got / exp - 1.0) < t, (what, got, exp)
    EMixW = 2.739850e-02
    EWi = [2.015680e-03, 1.007840e-03, 1.599940e-02, 3.199880e-02, 1.700724e-02, 1.801508e-02, 3.300664e-02, 3.401448e-02,
           1.201070e-02, 1.301854e-02, 1.402638e-02, 1.402638e-02, 1.503422e-02, 1.604206e-02, 2.801010e-02, 4.400950e-02,
           2.901794e-02, 3.002578e-02, 3.103362e-02, 3.103362e-02, 3.204146e-02, 2.603708e-02, 2.704492e-02, 2.805276e-02,
           2.906060e-02, 3.006844e-02, 4.102864e-02, 4.203648e-02, 4.304432e-02, 2.801340e-02]
    for i, e in enumerate(EWi):
        chk(o.lib.htr_oracle_species_W(i), e, f"W[{i}]", 1e-6)
    chk(r["MixW"], EMixW, "MixW")
    EXi = [7.407618e-03, 2.718537e-59, 1.712470e-60, 1.844254e-01, 1.610990e-60, 8.559107e-03, 1.936145e-05, 8.054952e-61,
           2.281174e-60, 2.104575e-60, 1.953355e-60, 1.953355e-60, 1.822409e-60, 9.008491e-02, 1.460273e-03, 3.725361e-04,
           9.441918e-61, 5.654101e-05, 8.828651e-61, 8.828651e-61, 1.052537e-05, 1.052288e-60, 1.013074e-60, 1.196430e-05,
           9.428057e-61, 2.851241e-05, 6.677896e-61, 6.517791e-61, 6.365184e-61, 7.075633e-01]
    for i, e in enumerate(EXi):
        chk(r["Xi"][i], e, f"Xi[{i}]")
    chk(r["rho"] * q["rhoRef"], P * EMixW / (R * T), "rho")
    chk(r["cp"] * q["CpRef"], 1.4689114e+03, "cp")
    chk(r["h"] * q["eRef"], 1.2239974e+06, "h")
    chk(r["mu"] * q["muRef"], 5.4245594e-05, "mu")
    chk(r["lam"] * q["lamRef"], 1.1047398e-01, "lam")
    chk(r["gamma"], 1.2603842e+00, "gamma")
    chk(r["SoS"] * q["uRef"], 7.5744434e+02, "SoS")
    Edif = [1.159092e-03, 1.919946e-03, 4.970050e-04, 3.137183e-04, 4.876624e-04, 4.248935e-04, 3.228673e-04, 3.206944e-04,
            4.671199e-04, 5.315370e-04, 3.640265e-04, 3.640265e-04, 3.559047e-04, 3.673578e-04, 3.204063e-04, 2.611356e-04,
            2.816701e-04, 2.793719e-04, 2.738767e-04, 2.738767e-04, 2.732409e-04, 2.720047e-04, 2.694003e-04, 2.693776e-04,
            2.473667e-04, 2.453524e-04, 4.018125e-04, 2.370506e-04, 2.359528e-04, 3.184521e-04]
    for i, e in enumerate(Edif):
        chk(r["Di"][i] * q["DiRef"], e, f"Di[{i}]")
    Eprod = [-2.619665e-04, 8.087741e-04, 7.719548e-05, -3.614274e-03, 2.444384e-03, -7.054158e-04, -4.116491e-02, 4.263025e-02,
             -1.594798e-53, -3.279458e-53, 1.364118e-53, 1.622365e-05, 3.088778e-02, -2.795221e-02, -2.962706e-03, 4.701254e-03,
             3.466765e-03, -3.706480e-03, 9.079769e-05, 9.819966e-08, -8.255164e-05, 2.713259e-05, 4.723412e-08, -3.162481e-05,
             1.908141e-05, -4.687640e-03, -8.071670e-55, 1.316211e-54, 2.042675e-56, 0.000000e+00]
    for i, e in enumerate(Eprod):      # fall-off, third-body and plain reactions all contribute
        if e == 0.0:
            assert r["w"][i] * q["wiRef"] == 0.0, i
        else:
            chk(r["w"][i] * q["wiRef"], e, f"w[{i}]", 2e-4)      # printed inputs (6 digits) through stiff rate expressions
    Edpdrho = [3.310111e+06, -4.697869e+07, -3.284257e+06, 3.610331e+05, -4.207291e+04, 3.871137e+06, 1.469175e+05, 1.159114e+06,
               -1.457004e+07, -1.130521e+07, -6.921959e+06, -7.602934e+06, -2.450997e+06, 1.092006e+06, 1.427804e+06, 2.520530e+06,
               -1.247804e+05, 1.072735e+06, 9.971174e+04, -1.386793e+05, 1.530256e+06, -2.250267e+06, -2.952978e+06, -6.743695e+05,
               -1.372558e+06, 3.044764e+05, -1.047489e+06, 2.804470e+05, -2.391786e+05, 4.043327e+05]
    for i, e in enumerate(Edpdrho):
        chk(r["dpdrhoi"][i] * q["eRef"], e, f"dpdrhoi[{i}]", 1e-4)
    Ehsp = [1.800606e+07, 2.410888e+08, 1.715512e+07, 1.268862e+06, 4.479985e+06, -1.074720e+07, 2.033562e+06, -1.907952e+06,
            6.175101e+07, 4.882339e+07, 3.165640e+07, 3.427167e+07, 1.419724e+07, 3.381412e+05, -2.559395e+06, -7.539517e+06,
            3.328419e+06, -1.340443e+06, 2.331136e+06, 3.246672e+06, -3.224028e+06, 1.172948e+07, 1.434182e+07, 5.510683e+06,
            8.117429e+06, 1.607184e+06, 6.263048e+06, 1.127859e+06, 3.089845e+06, 1.370973e+06]
    for i, e in enumerate(Ehsp):
        chk(o.lib.htr_oracle_species_enthalpy(i, T / 300.0) * q["eRef"], e, f"hsp[{i}]")


@needs_cp
def test_metrics_uniform_grid_and_tanh_fixture():
    """InitializeMetricTask through the oracle: on a uniform periodic grid all nine metrics are 1/dx to round-off
    (SURVEY appendix A); the committed Tanh fixture (tests/golden/tanh_y_metrics.npz, read by bench.py's air512_tanh
    workload) is what the reference's code gives for those nodes."""
    import os
    from oracle_lib import tanh_y_metrics
    from htr_solver_b200 import synthetic
    o = Oracle("ConstPropMix")
    n, w = (12, 10, 8), 2 * np.pi
    f = alloc_fields(*n, 1)
    cc = np.zeros(n[::-1] + (3,))
    for d, nn in enumerate(n):
        c = synthetic.cell_centers(synthetic.grid_nodes(nn, 0.0, w))
        shape = [1, 1, 1]; shape[2 - d] = nn
        cc[..., d] = c.reshape(shape)
    o.initialize_metric(f, cc, [w, w, w])
    for d, nn in enumerate(n):
        for kind in "eds":
            m = f[("dcsi_", "deta_", "dzet_")[d] + kind]
            assert np.allclose(m, nn / w, rtol=1e-13), (d, kind)
    fix = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "tanh_y_metrics.npz"))
    m = tanh_y_metrics(o, 40, float(fix["stretching"]), float(fix["width"]))
    for k, v in m.items():
        assert np.array_equal(v, fix[f"ny40_{k}"]), k
    # Tanh: small cells at both ends, symmetric, sizes sum to the width
    dy = 1.0 / m["deta_e"]
    assert abs(dy.sum() - float(fix["width"])) < 1e-12 and dy[0] < 0.3 * dy[20] and abs(dy[3] / dy[-4] - 1) < 1e-10
