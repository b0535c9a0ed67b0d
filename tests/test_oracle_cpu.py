"""CPU tests: the oracle pinned on the reference's own golden vectors; the product's mixture
block byte-identical with the reference's struct Mix; C-ABI exports."""
import ctypes as C
import os, re
import numpy as np
import pytest
from oracle_lib import Oracle
import htr_solver_b200 as htr
from htr_solver_b200 import synthetic

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def have_oracle(mix):
    try:
        Oracle(mix)
        return True
    except FileNotFoundError:
        return False


needs_air = pytest.mark.skipif(not have_oracle("AirMix"), reason="oracle/_ref not built")

# unitTests/mixTest/AirMix.json
MIXTEST_AIR = {"type": "AirMix", "LRef": 1.0, "PRef": 101325.0, "TRef": 300.0,
               "XiRef": {"Species": [{"Name": "N2", "MolarFrac": 0.79}, {"Name": "O2", "MolarFrac": 0.21}]}}


@needs_air
def test_mixtest_airmix_goldens():
    """unitTests/mixTest/mixTest.rg:130-169, T=750 K, P=1 atm, Yi=0.2, tolerance 1e-3 (:764-890)."""
    o = Oracle("AirMix"); o.init_mix(MIXTEST_AIR)
    r = o.mix_eval(101325.0 / 101325.0, 750.0 / 300.0, [0.2] * 5)
    m = C.create_string_buffer(o.lib.htr_oracle_sizeof_mix()); o.lib.htr_oracle_mix_bytes(m)
    # dimensional goldens printed by the reference test; compare the scale-free ones
    assert abs(r["gamma"] - 1.476782) / 1.476782 < 1e-3
    # MixW golden is dimensional [kg/mol]
    W = np.array([o.lib.htr_oracle_species_W(i) for i in range(5)])
    assert abs(1.0 / np.sum(0.2 / W) - 2.135186e-02) / 2.135186e-02 < 1e-3


@pytest.mark.parametrize("mix,cfg", [("AirMix", synthetic.AIRMIX_3DPERIODIC), ("AirMix", MIXTEST_AIR), ("CH41StMix", synthetic.CH41ST_3DPERIODIC),
                                     ("ConstPropMix", synthetic.CONSTPROP_VISCOUS), ("ConstPropMix", synthetic.CONSTPROP_3DPERIODIC)])
def test_mixture_block_is_byte_identical(mix, cfg):
    if not have_oracle(mix):
        pytest.skip("oracle/_ref not built")
    o = Oracle(mix); o.init_mix(cfg)
    n = o.lib.htr_oracle_sizeof_mix()
    buf = (C.c_uint8 * n)(); o.lib.htr_oracle_mix_bytes(buf)
    L = htr.load_library(mix)
    assert L.htr_sizeof_mix() == n
    assert htr.mixture_bytes(L, cfg) == bytes(buf)


def test_cabi_exports_every_declared_symbol():
    hdr = open(os.path.join(ROOT, "include", "htr_b200.h")).read()
    names = sorted(set(re.findall(r"\b(htr_[a-z_0-9]+)\s*\(", hdr)))
    assert len(names) >= 20
    for mix in ("ConstPropMix", "AirMix", "CH41StMix"):
        L = htr.load_library(mix)
        for n in names:
            assert hasattr(L, n), (mix, n)


def test_wrong_mixture_constructor_is_rejected():
    L = htr.load_library("ConstPropMix")
    buf = (C.c_uint8 * 4096)()
    fr = np.array([1.0])
    assert L.htr_mix_init_multicomp(buf, 1.0, 1e5, 300.0, 1, b"N2".ljust(10, b"\0"), fr.ctypes.data) == -3
    L = htr.load_library("AirMix")
    assert L.htr_mix_init_constprop(buf, 1.0, 1.4, 0.71, 0, 0.0, 0.0, 0.0) == -3
    fr = np.array([0.5])
    assert L.htr_mix_init_multicomp(buf, 1.0, 1e5, 300.0, 1, b"XX".ljust(10, b"\0"), fr.ctypes.data) == -1   # unknown species
    assert L.htr_mix_init_multicomp(buf, 1.0, 1e5, 300.0, 1, b"N2".ljust(10, b"\0"), fr.ctypes.data) == -1   # fractions do not sum to 1


def test_no_device_means_error_not_fallback():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(htr.HTRError):
        htr.Fluid(4, 4, 4, 1)


@needs_air
def test_nscbc_oracle_leaves_a_uniform_state_at_rest_alone():
    """The reference has no unit test for its NSCBC tasks, so the BC part of the oracle (oracle/ref_bc_driver.cc) is pinned on
    what the LODI relations must give: on a uniform gas at rest at the target pressure every wave amplitude vanishes, so the
    outflow, far-field and inflow tasks add nothing; a pressure excess at a subsonic outflow enters only through the relaxation
    term K (p - PInf) with K = sigma (1 - MaxMach^2) c / L (src/prometeo_rhs.inl:1464-1470)."""
    from oracle_lib import alloc_fields, alloc_bc_fields
    from parity import oracle_for
    cfg = synthetic.AIRMIX_3DPERIODIC
    o = oracle_for(cfg)
    nx, ny, nz = 6, 5, 4
    f = alloc_fields(nx, ny, nz, o.nspec)
    prim = synthetic.taylor_green((nx, ny, nz), cfg, U=0.0)
    for k, v in prim.items():
        f[k][...] = v
    f["velocity"][...] = 0.0
    f["pressure"][...] = 1.0
    f["temperature"][...] = 1.0
    f["nType_x"][:, :, 0] = 8; f["nType_x"][:, :, -1] = 12        # L_C / R_C
    o.update_properties_from_primitive(f)
    o.update_conserved_from_primitive(f)
    b = alloc_bc_fields(nx, ny, nz, o.nspec)
    b["velocity_profile"][...] = 0.0
    b["temperature_profile"][...] = 1.0
    b["MolarFracs_profile"][...] = f["MolarFracs"]
    for side, x in ((0, 0), (1, nx - 1)):
        lo, hi = [x, 0, 0], [x, ny - 1, nz - 1]
        f["Conserved_t"][...] = 0.0
        o.nscbc_outflow(f, 0, side, lo, hi, 0.3, 2.0, 1.0)
        o.nscbc_farfield(f, b, 0, side, lo, hi, 0.3, 2.0, 1.0)
        o.nscbc_inflow(f, b, 0, side, lo, hi)
        assert np.abs(f["Conserved_t"]).max() == 0.0
        # 1 % pressure excess at the outflow: d1 = K dp / (2 c^2) is the only non-zero LODI term besides dM_n
        o.nscbc_outflow(f, 0, side, lo, hi, 0.3, 2.0, 1.0/1.01)
        c = f["SoS"][0, 0, x]; rho = f["rho"][0, 0, x]
        K = 0.25*(1.0 - 0.3**2)*c/2.0
        d1 = 0.5*K*(1.0 - 1.0/1.01)/c**2
        Y = f["MassFracs"][0, 0, x]
        assert np.allclose(f["Conserved_t"][0, 0, x, :o.nspec], -d1*Y, rtol=1e-12, atol=0)
        dMn = (1 if side == 0 else -1)*K*(1.0 - 1.0/1.01)/(2*rho*c)
        assert np.isclose(f["Conserved_t"][0, 0, x, o.nspec], -rho*dMn, rtol=1e-12, atol=0)
