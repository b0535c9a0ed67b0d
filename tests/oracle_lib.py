"""TEST INFRASTRUCTURE ONLY: ctypes loader for the CPU oracles (oracle/htr_oracle_api.h).

Preference order: oracle/_ref/libhtr_ref_<Mix>.so (the reference's own CPU
numerics, "reference") then oracle/libhtr_port_<Mix>.so (our restatement, "port").
Never imported by the product package.
"""
import ctypes as C
import os
import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

SCHEMES = {"Hybrid": 0, "TENO": 1, "TENOA": 2, "TENOLAD": 3, "SkewSymmetric": 4}

# (name, components: 0 scalar, 'S' nSpec, 'E' nEq, 3 vec3, dtype)
FIELD_SPECS = [
    ("nType_x", 0, np.int32), ("nType_y", 0, np.int32), ("nType_z", 0, np.int32),
    ("dcsi_e", 0, np.float64), ("deta_e", 0, np.float64), ("dzet_e", 0, np.float64),
    ("dcsi_d", 0, np.float64), ("deta_d", 0, np.float64), ("dzet_d", 0, np.float64),
    ("dcsi_s", 0, np.float64), ("deta_s", 0, np.float64), ("dzet_s", 0, np.float64),
    ("rho", 0, np.float64), ("mu", 0, np.float64), ("lam", 0, np.float64),
    ("Di", "S", np.float64), ("SoS", 0, np.float64),
    ("pressure", 0, np.float64), ("temperature", 0, np.float64),
    ("MassFracs", "S", np.float64), ("MolarFracs", "S", np.float64), ("velocity", 3, np.float64),
    ("Conserved", "E", np.float64), ("Conserved_old", "E", np.float64),
    ("Conserved_t", "E", np.float64), ("Conserved_t_old", "E", np.float64),
    ("shockSensorX", 0, np.uint8), ("shockSensorY", 0, np.uint8), ("shockSensorZ", 0, np.uint8),
    ("DucrosSensor", 0, np.float64),
]


class OracleFields(C.Structure):
    _fields_ = [("nx", C.c_int64), ("ny", C.c_int64), ("nz", C.c_int64)] + \
               [(n, C.c_void_p) for n, _, _ in FIELD_SPECS]


# boundary fields of the NSCBC tasks (oracle/htr_oracle_api.h: htr_oracle_bc_fields)
BC_FIELD_SPECS = [("dudtBoundary", 3), ("dTdtBoundary", 0), ("velocity_profile", 3), ("temperature_profile", 0),
                  ("MolarFracs_profile", "S")]


class OracleBCFields(C.Structure):
    _fields_ = [(n, C.c_void_p) for n, _ in BC_FIELD_SPECS]


def alloc_bc_fields(nx, ny, nz, nspec):
    return {name: np.zeros((nz, ny, nx) if c == 0 else (nz, ny, nx, ncomp(c, nspec))) for name, c in BC_FIELD_SPECS}


def ncomp(c, nspec):
    return {0: 1, 3: 3, "S": nspec, "E": nspec + 4}[c]


def alloc_fields(nx, ny, nz, nspec):
    """dict of zero C-contiguous arrays [z][y][x](comp) for one tile."""
    out = {}
    for name, c, dt in FIELD_SPECS:
        shape = (nz, ny, nx) if c == 0 else (nz, ny, nx, ncomp(c, nspec))
        out[name] = np.zeros(shape, dtype=dt)
    return out


def oracle_paths(mix, omp=False):
    sfx = "_omp" if omp else ""
    return [os.path.join(ROOT, "oracle", "_ref", f"libhtr_ref_{mix}{sfx}.so"),
            os.path.join(ROOT, "oracle", f"libhtr_port_{mix}{sfx}.so")]


class Oracle:
    def __init__(self, mix, omp=False, kind=None):
        paths = oracle_paths(mix, omp)
        if kind == "reference":
            paths = paths[:1]
        elif kind == "port":
            paths = paths[1:]
        path = next((p for p in paths if os.path.exists(p)), None)
        if path is None:
            raise FileNotFoundError(f"no oracle library for {mix}: tried {paths}; run `make -C oracle`")
        self.path = path
        L = self.lib = C.CDLL(path)
        L.htr_oracle_kind.restype = C.c_char_p
        L.htr_oracle_mixture.restype = C.c_char_p
        L.htr_oracle_max_spectral_radius.restype = C.c_double
        L.htr_oracle_species_enthalpy.restype = C.c_double
        L.htr_oracle_species_enthalpy.argtypes = [C.c_int, C.c_double]
        L.htr_oracle_species_W.restype = C.c_double
        L.htr_oracle_species_W.argtypes = [C.c_int]
        L.htr_oracle_init_constprop.argtypes = [C.c_double] * 3 + [C.c_int] + [C.c_double] * 3
        L.htr_oracle_init_multicomp.argtypes = [C.c_double] * 3 + [C.c_int, C.c_char_p, C.c_void_p]
        L.htr_oracle_mix_eval.argtypes = [C.c_double, C.c_double, C.c_void_p, C.c_void_p]
        L.htr_oracle_euler_flux.argtypes = [C.c_int, C.c_int, C.c_void_p, C.c_double, C.c_double, C.c_double]
        L.htr_oracle_diffusion_flux.argtypes = [C.c_int, C.c_void_p]
        L.htr_oracle_ducros_sensor.argtypes = [C.c_void_p, C.c_double]
        L.htr_oracle_shock_sensor.argtypes = [C.c_int, C.c_void_p]
        L.htr_oracle_update_chemistry.argtypes = [C.c_void_p, C.c_double]
        for fn in ("add_chemistry_sources", "update_primitive_from_conserved",
                   "update_properties_from_primitive", "update_conserved_from_primitive",
                   "max_spectral_radius"):
            getattr(L, "htr_oracle_" + fn).argtypes = [C.c_void_p]
        L.htr_oracle_lu3_solve.argtypes = [C.c_void_p] * 3
        L.htr_oracle_rosenbrock3.argtypes = [C.c_void_p, C.c_double, C.c_double]
        self.kind = L.htr_oracle_kind().decode()
        self.mixture = L.htr_oracle_mixture().decode()
        self.nspec = L.htr_oracle_nspec()
        self.neq = L.htr_oracle_neq()
        self.threads = L.htr_oracle_max_threads()
        if omp and hasattr(L, "htr_oracle_set_threads"):
            # all host cores this process may use, whatever OMP_NUM_THREADS it inherited (torchrun exports 1)
            try:
                ncpu = len(os.sched_getaffinity(0))
            except AttributeError:
                ncpu = os.cpu_count() or 1
            L.htr_oracle_set_threads(C.c_int(ncpu))
            self.threads = L.htr_oracle_max_threads()
        assert self.mixture == mix

    # ---- mixture ----
    def init_mix(self, cfg):
        if cfg["type"] == "ConstPropMix":
            vm = cfg.get("viscosityModel", {"type": "Constant", "Visc": 0.0})
            t = {"Constant": 0, "PowerLaw": 1, "Sutherland": 2}[vm["type"]]
            p = {0: [vm.get("Visc", 0.0), 0, 0],
                 1: [vm.get("ViscRef", 0.0), vm.get("TempRef", 0.0), 0],
                 2: [vm.get("ViscRef", 0.0), vm.get("TempRef", 0.0), vm.get("SRef", 0.0)]}[t]
            rc = self.lib.htr_oracle_init_constprop(cfg["gasConstant"], cfg["gamma"], cfg["prandtl"], t, *map(float, p))
        else:
            sp = cfg["XiRef"]["Species"]
            names = b"".join(s["Name"].encode().ljust(10, b"\0") for s in sp)
            fr = np.array([s["MolarFrac"] for s in sp], dtype=np.float64)
            rc = self.lib.htr_oracle_init_multicomp(cfg["LRef"], cfg["PRef"], cfg["TRef"], len(sp), names, fr.ctypes.data)
        assert rc == 0, "oracle mixture init failed (wrong mixture type for this library?)"

    def mix_eval(self, P, T, Yi):
        Yi = np.ascontiguousarray(Yi, dtype=np.float64)
        out = np.zeros(16 + 4 * self.nspec)
        self.lib.htr_oracle_mix_eval(P, T, Yi.ctypes.data, out.ctypes.data)
        n = self.nspec
        keys = ["MixW", "rho", "cp", "h", "e", "mu", "lam", "gamma", "SoS", "dpde", "MixW_Xi", "T_from_e"]
        d = {k: out[i] for i, k in enumerate(keys)}
        d.update(Xi=out[16:16 + n], Di=out[16 + n:16 + 2 * n], w=out[16 + 2 * n:16 + 3 * n], dpdrhoi=out[16 + 3 * n:16 + 4 * n])
        return d

    # ---- tasks ----
    def _struct(self, f):
        s = OracleFields()
        nz, ny, nx = f["rho"].shape
        s.nx, s.ny, s.nz = nx, ny, nz
        for name, c, dt in FIELD_SPECS:
            a = f[name]
            assert a.flags["C_CONTIGUOUS"] and a.dtype == dt, name
            assert a.shape[:3] == (nz, ny, nx), name
            setattr(s, name, a.ctypes.data)
        return s

    def euler_flux(self, f, scheme, d, RK0=1.0, RK1=0.0, dt=0.0):
        self.lib.htr_oracle_euler_flux(SCHEMES[scheme], d, C.byref(self._struct(f)), RK0, RK1, dt)

    def diffusion_flux(self, f, d):
        self.lib.htr_oracle_diffusion_flux(d, C.byref(self._struct(f)))

    def ducros_sensor(self, f, vorticityScale):
        self.lib.htr_oracle_ducros_sensor(C.byref(self._struct(f)), vorticityScale)

    def shock_sensor(self, f, d):
        self.lib.htr_oracle_shock_sensor(d, C.byref(self._struct(f)))

    def add_chemistry_sources(self, f):
        self.lib.htr_oracle_add_chemistry_sources(C.byref(self._struct(f)))

    def update_chemistry(self, f, dt):
        self.lib.htr_oracle_update_chemistry(C.byref(self._struct(f)), dt)

    def initialize_metric(self, f, centers, width):
        """InitializeMetricTask (src/prometeo_metric.cc:36-190): fills the nine metric fields of `f` from the cell-centre
        coordinates `centers` [z][y][x][3] of a box of physical size `width`, node types taken from f."""
        centers = np.ascontiguousarray(centers, dtype=np.float64)
        w = np.ascontiguousarray(width, dtype=np.float64)
        self.lib.htr_oracle_initialize_metric.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
        self.lib.htr_oracle_initialize_metric(C.byref(self._struct(f)), centers.ctypes.data, w.ctypes.data)

    def update_chemistry_counted(self, f, dt):
        """UpdateChemistry + the number of rhs() evaluations of every cell (signature of its step sequence)."""
        nz, ny, nx = f["rho"].shape
        calls = np.zeros((nz, ny, nx), dtype=np.int32)
        self.lib.htr_oracle_update_chemistry_counted.argtypes = [C.c_void_p, C.c_double, C.c_void_p]
        self.lib.htr_oracle_update_chemistry_counted(C.byref(self._struct(f)), dt, calls.ctypes.data)
        return calls

    def update_primitive_from_conserved(self, f):
        self.lib.htr_oracle_update_primitive_from_conserved(C.byref(self._struct(f)))

    def update_properties_from_primitive(self, f):
        self.lib.htr_oracle_update_properties_from_primitive(C.byref(self._struct(f)))

    def update_conserved_from_primitive(self, f):
        self.lib.htr_oracle_update_conserved_from_primitive(C.byref(self._struct(f)))

    def max_spectral_radius(self, f):
        return self.lib.htr_oracle_max_spectral_radius(C.byref(self._struct(f)))

    # ---- non-periodic boundaries (oracle/ref_bc_driver.cc) ----
    @staticmethod
    def _bc_struct(nspec, bcf):
        s = OracleBCFields()
        for name, c in BC_FIELD_SPECS:
            a = bcf[name]
            assert a.flags["C_CONTIGUOUS"] and a.dtype == np.float64, name
            setattr(s, name, a.ctypes.data)
        return s

    @staticmethod
    def _box(lo, hi):
        return (C.c_int64 * 3)(*lo), (C.c_int64 * 3)(*hi)

    def nscbc_inflow(self, f, bcf, d, side, lo, hi):
        L = self.lib.htr_oracle_nscbc_inflow
        L.argtypes = [C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        l, h = self._box(lo, hi)
        L(d, side, C.byref(self._struct(f)), C.byref(self._bc_struct(self.nspec, bcf)), l, h)

    def nscbc_outflow(self, f, d, side, lo, hi, MaxMach, LengthScale, PInf):
        L = self.lib.htr_oracle_nscbc_outflow
        L.argtypes = [C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_double, C.c_double, C.c_double]
        l, h = self._box(lo, hi)
        L(d, side, C.byref(self._struct(f)), l, h, MaxMach, LengthScale, PInf)

    def nscbc_farfield(self, f, bcf, d, side, lo, hi, MaxMach, LengthScale, PInf):
        L = self.lib.htr_oracle_nscbc_farfield
        L.argtypes = [C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_double, C.c_double, C.c_double]
        l, h = self._box(lo, hi)
        L(d, side, C.byref(self._struct(f)), C.byref(self._bc_struct(self.nspec, bcf)), l, h, MaxMach, LengthScale, PInf)

    def incoming_shock(self, f, lo, hi, MaxMach, LengthScale, params):
        L = self.lib.htr_oracle_incoming_shock
        L.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_double, C.c_double, C.c_void_p]
        l, h = self._box(lo, hi)
        q = np.ascontiguousarray(params, dtype=np.float64)
        L(C.byref(self._struct(f)), l, h, MaxMach, LengthScale, q.ctypes.data)

    def set_nscbc_inflow_bc(self, f, bcf, d, lo, hi, Pbc):
        L = self.lib.htr_oracle_set_nscbc_inflow_bc
        L.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_double]
        l, h = self._box(lo, hi)
        L(d, C.byref(self._struct(f)), C.byref(self._bc_struct(self.nspec, bcf)), l, h, Pbc)

    def set_nscbc_outflow_bc(self, f, lo, hi):
        L = self.lib.htr_oracle_set_nscbc_outflow_bc
        L.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
        l, h = self._box(lo, hi)
        L(C.byref(self._struct(f)), l, h)

    def set_incoming_shock_bc(self, f, lo, hi, params):
        L = self.lib.htr_oracle_set_incoming_shock_bc
        L.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        l, h = self._box(lo, hi)
        q = np.ascontiguousarray(params, dtype=np.float64)
        L(C.byref(self._struct(f)), l, h, q.ctypes.data)

    def rhs(self, f, scheme, diffusion=True, chemistry=False, RK0=1.0, RK1=0.0, dt=0.0):
        """One UpdateDerivatives pass in the reference's task order
        (src/prometeo.rg:502-578): 0, +chem, +Euler z,y,x, +Diffusion z,y,x."""
        f["Conserved_t"][...] = 0.0
        if chemistry:
            self.add_chemistry_sources(f)
        for d in (2, 1, 0):
            self.euler_flux(f, scheme, d, RK0, RK1, dt)
        if diffusion:
            for d in (2, 1, 0):
                self.diffusion_flux(f, d)


def tanh_y_metrics(o, ny, stretching=1.3843992415368547, width=2 * np.pi):
    """Cell centres and y metrics of a periodic direction whose nodes follow the reference's Tanh map
    (src/prometeo_grid.rg:123-230), metrics by the reference's InitializeMetricTask through the oracle `o`."""
    import sys
    sys.path.insert(0, ROOT)
    from htr_solver_b200 import synthetic
    nodes = synthetic.grid_nodes(ny, 0.0, width, "Tanh", stretching)
    yc = synthetic.cell_centers(nodes)
    f = alloc_fields(1, ny, 1, o.nspec)              # a line: the metrics are separable
    cc = np.zeros((1, ny, 1, 3))
    cc[0, :, 0, 1] = yc
    cc[..., 0] = 0.5 * width; cc[..., 2] = 0.5 * width
    o.initialize_metric(f, cc, [width, width, width])
    return dict(centers=yc, deta_e=f["deta_e"][0, :, 0].copy(), deta_d=f["deta_d"][0, :, 0].copy(), deta_s=f["deta_s"][0, :, 0].copy())
