"""Lane-cooperative TENO kernels (csrc/htr_teno_coop.cuh: eight lanes per face) against the one-thread-per-face kernels
(the default; htr_set_option "coop" = 1 selects the cooperative form, which measured slower) and against the reference's CPU tasks.  Strict build: the two GPU forms are the
same IEEE operations per value, so they must agree bit for bit (NaNs of the outermost boundary nodes included)."""
import numpy as np
import pytest
import htr_solver_b200 as htr
from htr_solver_b200 import synthetic
from parity import oracle_for, host_state, to_device, euler_scale, scaled_err, TOL_FLUX, TOL_CHEM
from test_parity_gpu import _sensors, _bounded_node_types

pytestmark = pytest.mark.gpu

CP = synthetic.CONSTPROP_VISCOUS
AIR = synthetic.AIRMIX_3DPERIODIC


def _same(a, b):
    return np.array_equal(a, b, equal_nan=True)


@pytest.mark.parametrize("cfg,kw", [(CP, {"stretch": 0.2, "front": 0.3}),
                                    (AIR, {"species_wave": 1.0, "front": 0.3, "T0": 6.0, "p0": 1.0, "U": 0.5})], ids=["constprop", "air-hot-front"])
@pytest.mark.parametrize("bounded", [False, True], ids=["periodic", "bounded"])
@pytest.mark.parametrize("variant", ["", "strict"], ids=["prod", "strict"])
def test_coop_teno_flux_tasks(cfg, kw, bounded, variant):
    """UpdateUsing{TENO,TENOA,TENOLAD}EulerFlux in all three directions on a box whose sizes are no multiple of the run
    length (8), periodic and with all 13 node types: cooperative vs per-face kernels, and both vs the oracle."""
    o = oracle_for(cfg)
    n = (45, 26, 13)
    f = host_state(o, n, cfg, **kw)
    if bounded:
        _bounded_node_types(f)
    _sensors(o, f)
    s = htr.Solver(cfg, variant=variant)
    fl = to_device(s, f)
    rk = (0.75, 0.25); dt = 1e-3
    inner = (slice(None), slice(1, -1), slice(1, -1)) if bounded else (slice(None),)*3
    try:
        for scheme in ("TENO", "TENOA", "TENOLAD"):
            for d in range(3):
                out = {}
                for coop in (1, 0):
                    s.set_option("coop", coop)
                    fl.t["Conserved_t"].fill_(0.25)
                    s.UpdateUsingEulerFlux(fl, scheme, d, rk, dt)
                    out[coop] = fl.download(["Conserved_t"])["Conserved_t"]
                f["Conserved_t"][...] = 0.25
                o.euler_flux(f, scheme, d, rk[0], rk[1], dt)
                e = scaled_err(out[1][inner], f["Conserved_t"][inner], euler_scale(f, d))
                assert e <= (1e-14 if variant else TOL_FLUX), (scheme, d, e)
                if variant == "strict":
                    assert _same(out[1], out[0]), (scheme, d)
                else:
                    assert scaled_err(out[1][inner], out[0][inner], euler_scale(f, d)) <= TOL_FLUX, (scheme, d)
    finally:
        s.set_option("coop", 0)       # library default


@pytest.mark.parametrize("cfg,chem,kw", [(CP, False, {"front": 0.3, "stretch": 0.2}),
                                         (AIR, True, {"species_wave": 1.0, "front": 0.3, "stretch": 0.1})], ids=["constprop", "air"])
@pytest.mark.parametrize("variant", ["", "strict"], ids=["prod", "strict"])
def test_coop_deferred_cells_of_the_hybrid_rhs(cfg, chem, kw, variant):
    """htr_rhs (Hybrid) with a mixed sensor field: the deferred cells go through the cooperative kernel as runs of one
    cell from the compacted list; same result as with the per-face listed kernel, and as the oracle's task sequence."""
    o = oracle_for(cfg)
    f = host_state(o, (45, 23, 19), cfg, **kw)
    _sensors(o, f)
    f["shockSensorX"][:, :, 7::11] = 0; f["shockSensorY"][:, 5::9, :] = 0; f["shockSensorZ"][3::8, :, :] = 0
    s = htr.Solver(cfg, variant=variant)
    fl = to_device(s, f)
    out = {}
    try:
        for coop in (1, 0):
            s.set_option("coop", coop)
            fl.t["Conserved_t"].fill_(-3.0)
            s.UpdateDerivatives(fl, "Hybrid", diffusion=True, chemistry=chem, RK_coeffs=(0.75, 0.25), deltaTime=1e-3)
            out[coop] = fl.download(["Conserved_t"])["Conserved_t"]
    finally:
        s.set_option("coop", 0)       # library default
    o.rhs(f, "Hybrid", diffusion=True, chemistry=chem, RK0=0.75, RK1=0.25, dt=1e-3)
    scale = np.maximum.reduce([euler_scale(f, d) for d in range(3)])
    if chem:
        ref = f["Conserved_t"].copy()
        f["Conserved_t"][...] = 0.0
        o.add_chemistry_sources(f)
        scale = scale + np.abs(f["Conserved_t"]).reshape(-1, scale.size).max(axis=0)
        f["Conserved_t"][...] = ref
    if variant == "strict":
        assert _same(out[1], out[0])
    else:
        assert scaled_err(out[1], out[0], scale) <= TOL_FLUX
    assert scaled_err(out[1], f["Conserved_t"], scale) <= (TOL_CHEM if chem else TOL_FLUX)
