"""GPU: htr_rhs_no_alloc (include/htr_b200.h) -- htr_rhs with its scratch in the library's static device array, for callers
that must not allocate.  No counterpart in the reference, so the check is
against the existing entry point htr_rhs through the C ABI: same inputs, two calls in a row on DIFFERENT states and tile
sizes (the second call finds the first call's masks, flags and list in the array), within the tolerances of the parity tests;
then the size limit.  Collected last on purpose."""
import ctypes as C
import numpy as np
import pytest
import htr_solver_b200 as htr
from htr_solver_b200 import synthetic
from htr_solver_b200.core import HTR_ETOOBIG, Rect
from parity import oracle_for, host_state, to_device, euler_scale, scaled_err, TOL_FLUX, TOL_CHEM

pytestmark = pytest.mark.gpu
CP = synthetic.CONSTPROP_VISCOUS
AIR = synthetic.AIRMIX_3DPERIODIC


def _state(cfg, n, **kw):
    o = oracle_for(cfg)
    f = host_state(o, n, cfg, **kw)
    o.ducros_sensor(f, 1.0)
    for d in range(3):
        o.shock_sensor(f, d)
    return o, f


def _both(s, fl, scheme, chem):
    """Conserved_t of the existing entry point and of the no-allocation one on the same device state."""
    out = []
    for call in (s.UpdateDerivatives, s.UpdateDerivativesNoAlloc):
        fl.t["Conserved_t"].fill_(-7.0)
        call(fl, scheme, diffusion=True, chemistry=chem, RK_coeffs=(0.75, 0.25), deltaTime=1e-3)
        out.append(fl.download(["Conserved_t"])["Conserved_t"])
    return out


@pytest.mark.parametrize("cfg,chem,tol", [(AIR, True, TOL_CHEM), (CP, False, TOL_FLUX)], ids=["air", "constprop"])
def test_no_alloc_rhs_equals_htr_rhs_on_two_calls_in_a_row(cfg, chem, tol):
    s = htr.Solver(cfg)
    assert s.L.htr_rhs_no_alloc_max_cells() == 128**3
    # call 1: a box with a front (deferred cells: masks, flags and the list are all used); call 2: a smaller, different
    # box and scheme, whose scratch region still holds call 1's bytes
    cases = [((70, 37, 41), "Hybrid", {"front": 0.3, "stretch": 0.1}), ((45, 23, 19), "SkewSymmetric", {"stretch": 0.2}),
             ((70, 37, 41), "Hybrid", {"front": 0.3, "stretch": 0.1})]
    for n, scheme, kw in cases:
        if chem:
            kw = dict(kw, species_wave=1.0)
        o, f = _state(cfg, n, **kw)
        fl = to_device(s, f)
        ref, got = _both(s, fl, scheme, chem)
        assert np.isfinite(got).all()
        scale = np.maximum.reduce([euler_scale(f, d) for d in range(3)])
        if chem:
            f["Conserved_t"][...] = 0.0
            o.add_chemistry_sources(f)
            scale = scale + np.abs(f["Conserved_t"]).reshape(-1, scale.size).max(axis=0)
        e = scaled_err(got, ref, scale)
        assert e <= tol, (n, scheme, e)
        # and the existing entry point is what the other parity tests pin on the oracle
        o.rhs(f, scheme, diffusion=True, chemistry=chem, RK0=0.75, RK1=0.25, dt=1e-3)
        assert scaled_err(got, f["Conserved_t"], scale) <= tol, (n, scheme)


def test_no_alloc_rhs_rejects_a_tile_beyond_the_static_scratch():
    s = htr.Solver(CP)
    o, f = _state(CP, (16, 8, 8))
    fl = to_device(s, f)
    fl.t["Conserved_t"].fill_(3.0)
    big = Rect.make((0, 0, 0), (128, 127, 127))        # 129 x 128 x 128 cells: one plane more than the limit
    rc = s.UpdateDerivativesNoAlloc(fl, "Hybrid", tile=big, bounds=big)
    assert rc == HTR_ETOOBIG
    assert float(fl.t["Conserved_t"].min()) == 3.0 and float(fl.t["Conserved_t"].max()) == 3.0     # nothing was launched
    assert s.UpdateDerivativesNoAlloc(fl, "Hybrid") == 0
