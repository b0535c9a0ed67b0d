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
