"""Host-side mirror of the reference's task interface for the RHS hot path.

Thin ctypes layer over the C ABI in include/htr_b200.h (libhtr_b200_<Mix>.so, hand-written
sm_100a kernels).  PyTorch is used only for device memory and streams.  Names follow the
reference's tasks (src/prometeo_{rhs,sensor,chem,variables,cfl}.rg): UpdateUsingEulerFlux,
UpdateUsingDiffusionFlux, UpdateDucrosSensor, UpdateShockSensor, AddChemistrySources,
UpdateChemistry, UpdatePrimitiveFromConserved, UpdatePropertiesFromPrimitive,
UpdateConservedFromPrimitive, CalculateMaxSpectralRadius.

There is NO CPU fallback: a missing library or a missing GPU raises.
"""
import ctypes as C
import os
import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
HTR_ETOOBIG = -4        # include/htr_b200.h
SCHEMES = {"Hybrid": 0, "TENO": 1, "TENOA": 2, "TENOLAD": 3, "SkewSymmetric": 4}

# (name, components: 0 scalar, 'S' nSpec, 'E' nEq, 3 vec3, numpy dtype) -- order of htr_fluid
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
FIELD_NAMES = [n for n, _, _ in FIELD_SPECS]


class Acc(C.Structure):
    _fields_ = [("base", C.c_void_p), ("stride", C.c_int64 * 3)]


class Rect(C.Structure):
    _fields_ = [("lo", C.c_int64 * 3), ("hi", C.c_int64 * 3)]

    @staticmethod
    def make(lo, hi):
        r = Rect()
        r.lo[:] = list(lo)
        r.hi[:] = list(hi)
        return r


class FluidStruct(C.Structure):
    _fields_ = [(n, Acc) for n in FIELD_NAMES]


def ncomp(c, nspec):
    return {0: 1, 3: 3, "S": nspec, "E": nspec + 4}[c]


def rhs_launches(scheme, diffusion, chemistry, fast=True, march=True, listed=True):
    """Kernels one htr_rhs call launches (csrc/htr_kernels.cu: htr_rhs).  `march` = the plane-march form ("march" & 2);
    `listed` = deferred cells from the compacted list (the default): one compaction per mask + one listed pass per direction,
    against one block-scan pass per direction."""
    fastpath = fast and scheme in ("Hybrid", "SkewSymmetric")
    if fastpath and march:
        # Euler-z march + its deferred pass; plane march + deferred Euler y, x (+ deferred diffusion z, y, x) on one mask
        return 2 + 3 + (3 if diffusion else 0) + ((1 + 1) if listed else 0)
    if fastpath:
        # Euler: z(+chem), xy, three deferred passes over two masks; diffusion: z, xy, three deferred passes over two masks
        return 5 + (5 if diffusion else 0) + ((2 + (2 if diffusion else 0)) if listed else 0)
    return 1 + (1 if chemistry else 0) + 3 + (3 if diffusion else 0)


class HTRError(RuntimeError):
    pass


_LIBS = {}


def lib_path(mix, variant=""):
    return os.path.join(HERE, f"libhtr_b200_{mix}{'_' + variant if variant else ''}.so")


def load_library(mix, variant=""):
    """dlopen libhtr_b200_<Mix>[_strict].so; raises if it has not been built (no fallback).
    variant "strict" = built without FMA contraction and with libdevice pow (parity debugging)."""
    key = (mix, variant)
    if key in _LIBS:
        return _LIBS[key]
    path = lib_path(mix, variant)
    if not os.path.exists(path):
        raise HTRError(f"{path} not found: build it with `python htr-solver_b200/build.py` "
                       "(there is no CPU or PyTorch fallback)")
    L = C.CDLL(path)
    vp, i, d = C.c_void_p, C.c_int, C.c_double
    L.htr_mixture_name.restype = C.c_char_p
    L.htr_build_info.restype = C.c_char_p
    L.htr_mix_init_constprop.argtypes = [vp, d, d, d, i, d, d, d]
    L.htr_mix_init_multicomp.argtypes = [vp, d, d, d, i, C.c_char_p, vp]
    L.htr_set_mixture.argtypes = [vp, C.c_size_t]
    L.htr_set_option.argtypes = [C.c_char_p, i]
    L.htr_euler_flux.argtypes = [i, i, vp, vp, vp, d, d, d, vp]
    L.htr_diffusion_flux.argtypes = [i, vp, vp, vp, vp]
    L.htr_ducros_sensor.argtypes = [vp, vp, vp, d, vp]
    L.htr_shock_sensor.argtypes = [i, vp, vp, vp, vp]
    L.htr_add_chemistry_sources.argtypes = [vp, vp, vp]
    L.htr_update_chemistry.argtypes = [vp, vp, d, vp]
    L.htr_update_primitive_from_conserved.argtypes = [vp, vp, vp]
    L.htr_update_properties_from_primitive.argtypes = [vp, vp, vp]
    L.htr_update_ghost_primitive_from_conserved.argtypes = [vp, vp, vp]
    L.htr_update_conserved_from_primitive.argtypes = [vp, vp, vp]
    L.htr_max_spectral_radius.argtypes = [vp, vp, vp, vp]
    L.htr_rk_initialize_temporaries.argtypes = [vp, vp, vp]
    L.htr_rk_initialize_time_derivatives.argtypes = [vp, vp, i, vp]
    L.htr_rk_update_vars.argtypes = [i, d, i, vp, vp, vp]
    L.htr_rk_stage.argtypes = [i, d, i, vp, vp, vp]
    L.htr_time_step.argtypes = [i, i, i, i, d, vp, vp, vp, d, vp]
    L.htr_rhs.argtypes = [i, i, i, vp, vp, vp, d, d, d, vp]
    L.htr_rhs_parts.argtypes = [i, i, i, i, vp, vp, vp, d, d, d, vp]
    L.htr_rhs_no_alloc.argtypes = [i, i, i, vp, vp, vp, d, d, d, vp]
    L.htr_rhs_no_alloc_max_cells.restype = C.c_int64
    L.htr_rhs_no_alloc_max_cells.argtypes = []
    L.htr_host_rhs.argtypes = [i, i, i, i, d, vp, vp, vp, vp, vp, vp, vp, d, d, d, vp]
    L.htr_probe_fp64.argtypes = [d, vp, vp]
    L.htr_chemistry_failures_on.argtypes = [vp]
    assert L.htr_mixture_name().decode() == mix
    _LIBS[key] = L
    return L


def mixture_bytes(L, cfg):
    """Builds the mixture parameter block from a reference-style config dict
    (Flow.mixture of the reference's JSON input)."""
    n = L.htr_sizeof_mix()
    buf = (C.c_uint8 * n)()
    if cfg["type"] == "ConstPropMix":
        vm = cfg.get("viscosityModel", {"type": "Constant", "Visc": 0.0})
        t = {"Constant": 0, "PowerLaw": 1, "Sutherland": 2}[vm["type"]]
        p = {0: [vm.get("Visc", 0.0), 0.0, 0.0],
             1: [vm.get("ViscRef", 0.0), vm.get("TempRef", 0.0), 0.0],
             2: [vm.get("ViscRef", 0.0), vm.get("TempRef", 0.0), vm.get("SRef", 0.0)]}[t]
        rc = L.htr_mix_init_constprop(buf, cfg["gasConstant"], cfg["gamma"], cfg["prandtl"], t, *map(float, p))
    else:
        sp = cfg["XiRef"]["Species"]
        names = b"".join(s["Name"].encode().ljust(10, b"\0") for s in sp)
        fr = np.array([s["MolarFrac"] for s in sp], dtype=np.float64)
        rc = L.htr_mix_init_multicomp(buf, cfg["LRef"], cfg["PRef"], cfg["TRef"], len(sp), names, fr.ctypes.data)
    if rc != 0:
        raise HTRError(f"mixture construction failed (rc={rc}) for {cfg['type']} in {L.htr_mixture_name().decode()} library")
    return bytes(buf)


class Fluid:
    """Device-resident fields of one tile: interior [nz][ny][nx](comp) plus `halo` ghost cells
    per side in every direction, one allocation per field (x fastest), addressed in global
    coordinates with origin `origin` -- the layout of a single Legion instance per tile
    (src/prometeo_mapper.cc:1142-1204)."""

    def __init__(self, nx, ny, nz, nspec, halo=0, origin=(0, 0, 0), device="cuda"):
        import torch
        if not torch.cuda.is_available():
            raise HTRError("htr_solver_b200 needs a CUDA device (no CPU fallback)")
        self.torch = torch
        self.nx, self.ny, self.nz, self.nspec, self.halo = nx, ny, nz, nspec, halo
        self.origin = tuple(origin)
        self.device = torch.device(device)
        self.t = {}
        hx, hy, hz = self.halos()
        for name, c, dt in FIELD_SPECS:
            shape = (nz + 2 * hz, ny + 2 * hy, nx + 2 * hx) + (() if c == 0 else (ncomp(c, nspec),))
            tdt = {np.int32: torch.int32, np.float64: torch.float64, np.uint8: torch.uint8}[dt]
            self.t[name] = torch.zeros(shape, dtype=tdt, device=self.device)
        self._struct = None

    def halos(self):
        h = self.halo
        return (h, h, h) if isinstance(h, int) else tuple(h)

    # ---- geometry ----
    def tile(self):
        o = self.origin
        return Rect.make(o, (o[0] + self.nx - 1, o[1] + self.ny - 1, o[2] + self.nz - 1))

    def struct(self):
        if self._struct is None:
            s = FluidStruct()
            hx, hy, hz = self.halos()
            for name, c, dt in FIELD_SPECS:
                t = self.t[name]
                es = t.element_size()
                st = t.stride()
                a = Acc()
                sx, sy, sz = st[2] * es, st[1] * es, st[0] * es
                # base such that global point (origin - halo) is element 0
                gx, gy, gz = self.origin[0] - hx, self.origin[1] - hy, self.origin[2] - hz
                a.base = t.data_ptr() - (gx * sx + gy * sy + gz * sz)
                a.stride[:] = [sx, sy, sz]
                setattr(s, name, a)
            self._struct = s
        return self._struct

    # ---- host <-> device ----
    def interior(self, name):
        hx, hy, hz = self.halos()
        t = self.t[name]
        return t[hz:hz + self.nz, hy:hy + self.ny, hx:hx + self.nx]

    def upload(self, host, names=None):
        for name in (names or host.keys()):
            if name in self.t:
                self.interior(name).copy_(self.torch.from_numpy(np.ascontiguousarray(host[name])))

    def download(self, names=None):
        return {n: self.interior(n).cpu().numpy() for n in (names or FIELD_NAMES)}

    def save_npz(self, path, names=None):
        """Interior of the fields as one .npz file: the input of tools/compare_fields.py (the numpy restatement of the
        reference's scripts/compare_hdf.py and of its solver-test tolerance rule; HDF5 is absent on this image)."""
        import numpy as np
        np.savez(path, **self.download(names))


# Runge-Kutta coefficients RK_C[STAGE] = (old state, current state, time derivative), src/prometeo_const.rg:54-58
RK_C = ((1.0, 0.0, 1.0), (3.0/4.0, 1.0/4.0, 1.0/4.0), (1.0/3.0, 2.0/3.0, 2.0/3.0))


class Solver:
    """The reference's leaf tasks for one mixture, on the current CUDA device."""

    def __init__(self, mix_cfg, variant=""):
        import torch
        self.torch = torch
        self.mix = mix_cfg["type"]
        self.L = load_library(self.mix, variant)
        self.nspec = self.L.htr_nspec()
        self.neq = self.L.htr_neq()
        self.mix_bytes = mixture_bytes(self.L, mix_cfg)
        self._check(self.L.htr_set_mixture(self.mix_bytes, len(self.mix_bytes)), "htr_set_mixture")
        self._cfl = torch.zeros(1, dtype=torch.float64, device="cuda")

    def _check(self, rc, what):
        if rc != 0:
            raise HTRError(f"{what} failed with code {rc}")

    def set_option(self, name, value):
        """Library switch (htr_set_option): "fast" 1/0 = staged fast path of htr_rhs on/off."""
        self._check(self.L.htr_set_option(name.encode(), int(value)), "htr_set_option")

    def _stream(self):
        return C.c_void_p(self.torch.cuda.current_stream().cuda_stream)

    def new_fluid(self, nx, ny, nz, halo=0, origin=(0, 0, 0)):
        return Fluid(nx, ny, nz, self.nspec, halo, origin)

    @staticmethod
    def _rects(fl, tile, bounds):
        t = tile if tile is not None else fl.tile()
        b = bounds if bounds is not None else fl.tile()
        return C.byref(t), C.byref(b)

    # src/prometeo_rhs.rg:103-139 / src/prometeo_rhs.cc:37-376
    def UpdateUsingEulerFlux(self, fl, scheme, d, RK_coeffs=(1.0, 0.0), deltaTime=0.0, tile=None, bounds=None):
        t, b = self._rects(fl, tile, bounds)
        self._check(self.L.htr_euler_flux(SCHEMES[scheme], d, C.byref(fl.struct()), t, b,
                                          RK_coeffs[0], RK_coeffs[1], deltaTime, self._stream()), "htr_euler_flux")

    def UpdateUsingDiffusionFlux(self, fl, d, tile=None, bounds=None):
        t, b = self._rects(fl, tile, bounds)
        self._check(self.L.htr_diffusion_flux(d, C.byref(fl.struct()), t, b, self._stream()), "htr_diffusion_flux")

    def UpdateDucrosSensor(self, fl, vorticityScale, tile=None, bounds=None):
        t, b = self._rects(fl, tile, bounds)
        self._check(self.L.htr_ducros_sensor(C.byref(fl.struct()), t, b, vorticityScale, self._stream()), "htr_ducros_sensor")

    def UpdateShockSensor(self, fl, d, tile=None, bounds=None):
        t, b = self._rects(fl, tile, bounds)
        self._check(self.L.htr_shock_sensor(d, C.byref(fl.struct()), t, b, self._stream()), "htr_shock_sensor")

    def AddChemistrySources(self, fl, tile=None):
        t, _ = self._rects(fl, tile, None)
        self._check(self.L.htr_add_chemistry_sources(C.byref(fl.struct()), t, self._stream()), "htr_add_chemistry_sources")

    def UpdateChemistry(self, fl, Integrator_deltaTime, tile=None, rhs_counts=None):
        """rhs_counts: optional int32 CUDA tensor (one entry per cell of the tile) that receives the number of
        right-hand-side evaluations of every cell (diagnostic of the adaptive step sequence)."""
        t, _ = self._rects(fl, tile, None)
        if rhs_counts is not None:
            self.L.htr_chemistry_rhs_counts.argtypes = [C.c_void_p]
            self._check(self.L.htr_chemistry_rhs_counts(C.c_void_p(rhs_counts.data_ptr())), "htr_chemistry_rhs_counts")
        try:
            self._check(self.L.htr_update_chemistry(C.byref(fl.struct()), t, Integrator_deltaTime, self._stream()), "htr_update_chemistry")
        finally:
            if rhs_counts is not None:
                self.torch.cuda.current_stream().synchronize()
                self.L.htr_chemistry_rhs_counts(C.c_void_p(0))
        nfail = self.L.htr_chemistry_failures_on(self._stream())
        if nfail:
            raise HTRError(f"implicit chemistry did not converge in {nfail} cells")

    def UpdatePrimitiveFromConserved(self, fl, tile=None):
        t, _ = self._rects(fl, tile, None)
        self._check(self.L.htr_update_primitive_from_conserved(C.byref(fl.struct()), t, self._stream()), "htr_update_primitive_from_conserved")

    def UpdatePropertiesFromPrimitive(self, fl, tile=None):
        t, _ = self._rects(fl, tile, None)
        self._check(self.L.htr_update_properties_from_primitive(C.byref(fl.struct()), t, self._stream()), "htr_update_properties_from_primitive")

    def UpdateGhostPrimitiveFromConserved(self, fl, tile=None):
        """Primitives of ghost copies from the exchanged (converged) temperature, no Newton iteration."""
        t, _ = self._rects(fl, tile, None)
        self._check(self.L.htr_update_ghost_primitive_from_conserved(C.byref(fl.struct()), t, self._stream()),
                    "htr_update_ghost_primitive_from_conserved")

    def UpdateConservedFromPrimitive(self, fl, tile=None):
        t, _ = self._rects(fl, tile, None)
        self._check(self.L.htr_update_conserved_from_primitive(C.byref(fl.struct()), t, self._stream()), "htr_update_conserved_from_primitive")

    # src/prometeo_rk.rg:46-111
    def InitializeTemporaries(self, fl, tile=None):
        t, _ = self._rects(fl, tile, None)
        self._check(self.L.htr_rk_initialize_temporaries(C.byref(fl.struct()), t, self._stream()), "htr_rk_initialize_temporaries")

    def InitializeTimeDerivatives(self, fl, implicitChemistry=False, tile=None):
        t, _ = self._rects(fl, tile, None)
        self._check(self.L.htr_rk_initialize_time_derivatives(C.byref(fl.struct()), t, int(implicitChemistry), self._stream()),
                    "htr_rk_initialize_time_derivatives")

    def UpdateVars(self, fl, stage, Integrator_deltaTime, implicitChemistry=False, tile=None):
        t, _ = self._rects(fl, tile, None)
        self._check(self.L.htr_rk_update_vars(int(stage), float(Integrator_deltaTime), int(implicitChemistry),
                                              C.byref(fl.struct()), t, self._stream()), "htr_rk_update_vars")

    def UpdateVarsAndRecover(self, fl, stage, Integrator_deltaTime, implicitChemistry=False, tile=None):
        """UpdateVars(STAGE) + UpdatePrimitiveFromConserved + UpdatePropertiesFromPrimitive in one pass
        (src/prometeo.rg:1427, :452, :470)."""
        t, _ = self._rects(fl, tile, None)
        self._check(self.L.htr_rk_stage(int(stage), float(Integrator_deltaTime), int(implicitChemistry),
                                        C.byref(fl.struct()), t, self._stream()), "htr_rk_stage")

    def TimeStep(self, fl, scheme, Integrator_deltaTime, diffusion=True, chemistry=False, vorticityScale=1.0,
                 tile=None, bounds=None, implicitChemistry=False):
        """One RK3 step of a block without boundary nodes (MainLoopBody, src/prometeo.rg:1348-1470)."""
        t, b = self._rects(fl, tile, bounds)
        self._check(self.L.htr_time_step(SCHEMES[scheme], int(diffusion), int(chemistry), int(implicitChemistry), float(vorticityScale),
                                         C.byref(fl.struct()), t, b, float(Integrator_deltaTime), self._stream()), "htr_time_step")

    def CalculateMaxSpectralRadius(self, fl, tile=None, sync=True):
        t, _ = self._rects(fl, tile, None)
        self._check(self.L.htr_max_spectral_radius(C.byref(fl.struct()), t, C.c_void_p(self._cfl.data_ptr()), self._stream()),
                    "htr_max_spectral_radius")
        return float(self._cfl.item()) if sync else self._cfl

    # src/prometeo.rg:481-608 (UpdateDerivatives), hot-path part
    def UpdateDerivatives(self, fl, scheme, diffusion=True, chemistry=False, RK_coeffs=(1.0, 0.0), deltaTime=0.0,
                          tile=None, bounds=None):
        t, b = self._rects(fl, tile, bounds)
        self._check(self.L.htr_rhs(SCHEMES[scheme], int(diffusion), int(chemistry), C.byref(fl.struct()), t, b,
                                   RK_coeffs[0], RK_coeffs[1], deltaTime, self._stream()), "htr_rhs")

    def UpdateDerivativesNoAlloc(self, fl, scheme, diffusion=True, chemistry=False, RK_coeffs=(1.0, 0.0), deltaTime=0.0,
                                 tile=None, bounds=None):
        """UpdateDerivatives through htr_rhs_no_alloc: scratch in the library's static device array (tiles of at most
        htr_rhs_no_alloc_max_cells() cells), no allocation.  Returns the C return code instead of raising on HTR_ETOOBIG."""
        t, b = self._rects(fl, tile, bounds)
        rc = self.L.htr_rhs_no_alloc(SCHEMES[scheme], int(diffusion), int(chemistry), C.byref(fl.struct()), t, b,
                                     RK_coeffs[0], RK_coeffs[1], deltaTime, self._stream())
        if rc != HTR_ETOOBIG:
            self._check(rc, "htr_rhs_no_alloc")
        return rc

    def UpdateDerivativesParts(self, parts, fl, scheme, diffusion=True, chemistry=False, RK_coeffs=(1.0, 0.0), deltaTime=0.0,
                               tile=None, bounds=None):
        """Some of the staged kernels of UpdateDerivatives (htr_rhs_parts; timing only)."""
        t, b = self._rects(fl, tile, bounds)
        self._check(self.L.htr_rhs_parts(int(parts), SCHEMES[scheme], int(diffusion), int(chemistry), C.byref(fl.struct()), t, b,
                                         RK_coeffs[0], RK_coeffs[1], deltaTime, self._stream()), "htr_rhs_parts")

    # host-buffer stage: torch CPU (pinned) tensors in, Conserved_t out; see htr_host_rhs in include/htr_b200.h
    def HostRHS(self, fl, scheme, h_Conserved, h_temperature, h_Conserved_t, h_Conserved_old=None, diffusion=True,
                chemistry=False, update_sensors=True, vorticityScale=1.0, RK_coeffs=(1.0, 0.0), deltaTime=0.0):
        t, b = self._rects(fl, None, None)
        old = C.c_void_p(h_Conserved_old.data_ptr()) if h_Conserved_old is not None else C.c_void_p(0)
        self._check(self.L.htr_host_rhs(SCHEMES[scheme], int(diffusion), int(chemistry), int(update_sensors), vorticityScale,
                                        C.byref(fl.struct()), t, b, C.c_void_p(h_Conserved.data_ptr()), old,
                                        C.c_void_p(h_temperature.data_ptr()), C.c_void_p(h_Conserved_t.data_ptr()),
                                        RK_coeffs[0], RK_coeffs[1], deltaTime, self._stream()), "htr_host_rhs")
