"""TEST INFRASTRUCTURE: shared helpers for the GPU-vs-oracle parity tests."""
import numpy as np
import htr_solver_b200 as htr
from htr_solver_b200 import synthetic
from oracle_lib import Oracle, alloc_fields

# Tolerances stated by the north star (BASELINE.json): relative, fp64.
TOL_FLUX = 1e-12      # flux-only
TOL_CHEM = 1e-10      # with chemistry / transcendental-heavy property evaluation

_ORACLES = {}


def oracle_for(cfg):
    key = (cfg["type"], repr(sorted(cfg.items(), key=str)))
    o = _ORACLES.get(cfg["type"])
    if o is None:
        o = _ORACLES[cfg["type"]] = Oracle(cfg["type"])
    o.init_mix(cfg)      # the oracle keeps one global mixture: re-initialise for every use
    return o


def host_state(o, n, cfg, **kw):
    """Full host field set for an n^3 Taylor-Green box: primitives from synthetic.py, then the
    oracle's UpdatePropertiesFromPrimitive + UpdateConservedFromPrimitive."""
    nx, ny, nz = (n, n, n) if np.isscalar(n) else n
    f = alloc_fields(nx, ny, nz, o.nspec)
    prim = synthetic.taylor_green(n, cfg, **kw)
    for k, v in prim.items():
        f[k][...] = v
    o.update_properties_from_primitive(f)
    o.update_conserved_from_primitive(f)
    f["Conserved_old"][...] = f["Conserved"]
    return f


def to_device(solver, f):
    nz, ny, nx = f["rho"].shape
    fl = solver.new_fluid(nx, ny, nz)
    fl.upload(f)
    return fl


def rel_err(a, b):
    """max |a-b| / max |b| per trailing component (L-inf relative error of a field component).
    Point-wise relative error is meaningless for a flux divergence that crosses zero."""
    a = np.asarray(a, dtype=np.float64); b = np.asarray(b, dtype=np.float64)
    if a.ndim == 3:
        a = a[..., None]; b = b[..., None]
    num = np.abs(a - b).reshape(-1, a.shape[-1]).max(axis=0)
    den = np.abs(b).reshape(-1, a.shape[-1]).max(axis=0)
    return float(np.max(num / np.maximum(den, 1e-300)))


def solver_tol_ok(a, b):
    """The reference's own solver-test rule (solverTests/testAll.py:23-36):
    abs(d) <= max(1e-10*abs(ref), 1e-12)."""
    a = np.asarray(a); b = np.asarray(b)
    return bool(np.all(np.abs(a - b) <= np.maximum(1e-10 * np.abs(b), 1e-12)))


def euler_scale(f, d):
    """Magnitude of the operands of the Euler flux difference, per conserved component:
    m_e * (|u_n| + c) * |U_i|  (+ m_e * p for momentum, (|u_n|+c) * (rhoE + p) for energy).
    Conserved_t[i] = m*(F+ - F-) is a difference of two numbers of this size, so its absolute
    rounding floor is eps * scale even where the divergence itself vanishes (e.g. the z-flux of
    the Taylor-Green box, w = 0).  The parity error of a flux RHS is reported relative to it."""
    m = f[("dcsi_e", "deta_e", "dzet_e")[d]].max()
    lam = (np.abs(f["velocity"][..., d]) + f["SoS"]).max()
    U = np.abs(f["Conserved"]).reshape(-1, f["Conserved"].shape[-1]).max(axis=0)
    n = f["Conserved"].shape[-1]
    s = m * lam * U
    p = f["pressure"].max()
    s[n - 4:n - 1] = np.maximum(s[n - 4:n - 1], m * p)
    s[n - 1] = m * lam * (U[n - 1] + p)
    return s


def scaled_err(got, ref, scale):
    num = np.abs(got - ref).reshape(-1, got.shape[-1]).max(axis=0)
    return float(np.max(num / scale))
