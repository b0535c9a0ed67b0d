"""Non-periodic boundaries: host-side mirror of the reference's NSCBC right-hand-side tasks and boundary-cell primitive
updates (src/prometeo_rhs.hpp:479-875, src/prometeo_bc.hpp:113-215), thin ctypes callers of the C ABI
(include/htr_b200.h: htr_nscbc_*, htr_set_*_bc).  Task names follow the reference's."""
import ctypes as C
import numpy as np
from .core import Acc, HTRError, ncomp

# boundary fields of struct Fluid_columns read by these tasks (src/prometeo_types.h:49-113); same order as htr_bc_fields
BC_FIELD_SPECS = [("dudtBoundary", 3), ("dTdtBoundary", 0), ("velocity_profile", 3), ("temperature_profile", 0),
                  ("MolarFracs_profile", "S")]
MINUS, PLUS = 0, 1      # HTR_SIDE_MINUS / HTR_SIDE_PLUS


class BCStruct(C.Structure):
    _fields_ = [(n, Acc) for n, _ in BC_FIELD_SPECS]


class BCFields:
    """Device-resident boundary fields laid out like the fields of a `Fluid` (same box, halo and origin)."""

    def __init__(self, fl):
        torch = fl.torch
        self.fl = fl
        hx, hy, hz = fl.halos()
        self.t = {}
        for name, c in BC_FIELD_SPECS:
            shape = (fl.nz + 2*hz, fl.ny + 2*hy, fl.nx + 2*hx) + (() if c == 0 else (ncomp(c, fl.nspec),))
            self.t[name] = torch.zeros(shape, dtype=torch.float64, device=fl.device)
        self._struct = None

    def struct(self):
        if self._struct is None:
            s = BCStruct()
            fl = self.fl
            hx, hy, hz = fl.halos()
            for name, _ in BC_FIELD_SPECS:
                t = self.t[name]
                es = t.element_size()
                st = t.stride()
                a = Acc()
                sx, sy, sz = st[2]*es, st[1]*es, st[0]*es
                gx, gy, gz = fl.origin[0] - hx, fl.origin[1] - hy, fl.origin[2] - hz
                a.base = t.data_ptr() - (gx*sx + gy*sy + gz*sz)
                a.stride[:] = [sx, sy, sz]
                setattr(s, name, a)
            self._struct = s
        return self._struct

    def upload(self, host):
        fl = self.fl
        hx, hy, hz = fl.halos()
        for name, _ in BC_FIELD_SPECS:
            if name in host:
                self.t[name][hz:hz + fl.nz, hy:hy + fl.ny, hx:hx + fl.nx].copy_(fl.torch.from_numpy(np.ascontiguousarray(host[name])))


def shock_params(iShock, MolarFracs, pressure0, temperature0, velocity0, pressure1, temperature1, velocity1):
    """struct IncomingShockParams (src/prometeo_bc_types.h:46-59) flattened the way the C ABI takes it."""
    return np.ascontiguousarray([float(iShock)] + list(MolarFracs) + [pressure0, temperature0] + list(velocity0) +
                                [pressure1, temperature1] + list(velocity1), dtype=np.float64)


_D = C.c_double
_VP = C.c_void_p
_I = C.c_int


def _bind(L):
    if getattr(L, "_htr_bc_bound", False):
        return
    L.htr_nscbc_inflow.argtypes = [_I, _I, _VP, _VP, _VP, _VP]
    L.htr_nscbc_outflow.argtypes = [_I, _I, _VP, _VP, _VP, _D, _D, _D, _VP]
    L.htr_nscbc_farfield.argtypes = [_I, _I, _VP, _VP, _VP, _D, _D, _D, _VP]
    L.htr_nscbc_incoming_shock.argtypes = [_VP, _VP, _D, _D, _VP, _VP]
    L.htr_set_nscbc_inflow_bc.argtypes = [_I, _VP, _VP, _VP, _D, _VP]
    L.htr_set_nscbc_outflow_bc.argtypes = [_VP, _VP, _VP]
    L.htr_set_incoming_shock_bc.argtypes = [_VP, _VP, _VP, _VP]
    L._htr_bc_bound = True


class BoundaryTasks:
    """The boundary tasks of one Solver.  `bc` = Rect of the boundary cells the task runs on (the reference's p_xNeg[0][c]
    etc., src/prometeo_rhs.rg:670-760), `bounds` = Rect of the whole grid."""

    def __init__(self, solver):
        self.s = solver
        self.L = solver.L
        _bind(self.L)

    def _check(self, rc, what):
        if rc != 0:
            raise HTRError(f"{what} failed with code {rc}")

    # src/prometeo_rhs.hpp:479-582
    def UpdateUsingFluxNSCBCInflow(self, fl, bcf, dir, side, bc):
        self._check(self.L.htr_nscbc_inflow(dir, side, C.byref(fl.struct()), C.byref(bcf.struct()), C.byref(bc), self.s._stream()),
                    "htr_nscbc_inflow")

    # src/prometeo_rhs.hpp:584-720
    def UpdateUsingFluxNSCBCOutflow(self, fl, dir, side, bc, bounds, MaxMach, LengthScale, PInf):
        self._check(self.L.htr_nscbc_outflow(dir, side, C.byref(fl.struct()), C.byref(bc), C.byref(bounds), MaxMach, LengthScale, PInf,
                                             self.s._stream()), "htr_nscbc_outflow")

    # src/prometeo_rhs.hpp:722-838
    def UpdateUsingFluxNSCBCFarField(self, fl, bcf, dir, side, bc, MaxMach, LengthScale, PInf):
        self._check(self.L.htr_nscbc_farfield(dir, side, C.byref(fl.struct()), C.byref(bcf.struct()), C.byref(bc), MaxMach, LengthScale,
                                              PInf, self.s._stream()), "htr_nscbc_farfield")

    # src/prometeo_rhs.hpp:840-875
    def UpdateUsingFluxIncomingShock(self, fl, bc, MaxMach, LengthScale, params):
        q = np.ascontiguousarray(params, dtype=np.float64)
        self._check(self.L.htr_nscbc_incoming_shock(C.byref(fl.struct()), C.byref(bc), MaxMach, LengthScale, q.ctypes.data,
                                                    self.s._stream()), "htr_nscbc_incoming_shock")

    # src/prometeo_bc.hpp:113-152
    def SetNSCBC_InflowBC(self, fl, bcf, dir, bc, Pbc):
        self._check(self.L.htr_set_nscbc_inflow_bc(dir, C.byref(fl.struct()), C.byref(bcf.struct()), C.byref(bc), Pbc, self.s._stream()),
                    "htr_set_nscbc_inflow_bc")

    # src/prometeo_bc.hpp:160-183
    def SetNSCBC_OutflowBC(self, fl, bc):
        self._check(self.L.htr_set_nscbc_outflow_bc(C.byref(fl.struct()), C.byref(bc), self.s._stream()), "htr_set_nscbc_outflow_bc")

    # src/prometeo_bc.hpp:190-215
    def SetIncomingShockBC(self, fl, bc, params):
        q = np.ascontiguousarray(params, dtype=np.float64)
        self._check(self.L.htr_set_incoming_shock_bc(C.byref(fl.struct()), C.byref(bc), q.ctypes.data, self.s._stream()),
                    "htr_set_incoming_shock_bc")
