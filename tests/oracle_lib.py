"""ctypes bindings for the CHECKERS (test infrastructure, never product):

* ``oracle/liboracle.so``     -- plain-C restatement of the reference path (oracle/fs_oracle.c,
                                 oracle/pfmg_oracle.c)
* ``oracle/_ref/libfsref.so`` -- the UNMODIFIED reference (headers + vendored HYPRE) behind
                                 oracle/ref_harness.cpp; present where oracle/Makefile built it.

All arrays are numpy float64 in the reference's storage (src/Container.hpp:105-111): ghost
inclusive, j fastest.  U-like (NX+3, NY+2), V-like (NX+2, NY+3), cell (NX+2, NY+2).
"""
from __future__ import annotations

import ctypes as C
import os
from dataclasses import dataclass, field

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_SO = os.path.join(ROOT, "oracle", "liboracle.so")
REF_SO = os.path.join(ROOT, "oracle", "_ref", "libfsref.so")

DIRICHLET, NEUMANN, PERIODIC, SYMMETRY = 0, 1, 2, 3
LEFT, RIGHT, BOTTOM, TOP = 0, 1, 2, 3
PSD_NONE, PSD_LEFT, PSD_RIGHT, PSD_BOTTOM, PSD_TOP = 0, 1, 2, 3, 4

DP = C.POINTER(C.c_double)


class CFs(C.Structure):
    _fields_ = [("nx", C.c_int64), ("ny", C.c_int64), ("dx", C.c_double), ("dy", C.c_double),
                ("rho_gas", C.c_double), ("rho_liquid", C.c_double), ("visc_gas", C.c_double),
                ("visc_liquid", C.c_double), ("sigma", C.c_double)] + [
        (n, DP) for n in ("visc", "p", "p_jump_u", "p_jump_v", "rho_u_old", "rho_v_old", "U_old",
                          "V_old", "rho_u", "rho_v", "U", "V")]


class CBconds(C.Structure):
    _fields_ = [("kind", C.c_int * 4), ("u_val", C.c_double * 4), ("v_val", C.c_double * 4),
                ("u_prof", DP * 4), ("v_prof", DP * 4), ("clipped", C.c_int * 4)]


class CSolverOpts(C.Structure):
    _fields_ = [("tol", C.c_double), ("max_iter", C.c_int), ("max_levels", C.c_int),
                ("relax_type", C.c_int), ("skip_relax", C.c_int), ("snap_zero_diag", C.c_int)]


class CStepOpts(C.Structure):
    _fields_ = [("num_subiter", C.c_int), ("outflow_fix", C.c_int), ("psd", C.c_int),
                ("rebuild_operator_each_subiter", C.c_int), ("density_update", C.c_int),
                ("capillary", C.c_int), ("vf", DP), ("curv", DP), ("ilen", DP), ("total_flow", C.c_double),
                ("ib_mode", C.c_int), ("ib_u", DP), ("ib_v", DP)]


def _step_opts(num_subiter, outflow_fix, psd, rebuild, density, capillary, total_flow=0.0, ib=None):
    """outflow_fix: 0 none, 1 LaminarChannel, 2 CustomDirichletChannel (dy-weighted), 3 PeriodicChannel
    (both ghost columns forced to total_flow)"""
    o = CStepOpts(num_subiter, outflow_fix, psd, rebuild, int(density), int(capillary is not None))
    o.total_flow = total_flow
    if ib is not None:  # (mode, ib_corr_u_stag, ib_corr_v_stag): 1 implicit Euler, 2 semi-analytical
        o._keep_ib = [np.ascontiguousarray(a, dtype=np.float64) for a in ib[1:]]
        o.ib_mode = int(ib[0])
        o.ib_u, o.ib_v = (a.ctypes.data_as(DP) for a in o._keep_ib)
    if capillary is not None:
        o._keep = [np.ascontiguousarray(a, dtype=np.float64) for a in capillary]
        o.vf, o.curv, o.ilen = (a.ctypes.data_as(DP) for a in o._keep)
    return o


FIELDS_U = ("p_jump_u", "rho_u_old", "U_old", "rho_u", "U")
FIELDS_V = ("p_jump_v", "rho_v_old", "V_old", "rho_v", "V")
FIELDS_C = ("visc", "p")
ALL_FIELDS = FIELDS_C + FIELDS_U + FIELDS_V


def shape_of(name: str, nx: int, ny: int):
    if name in FIELDS_U or name in ("dmomU",):
        return (nx + 3, ny + 2)
    if name in FIELDS_V or name in ("dmomV",):
        return (nx + 2, ny + 3)
    return (nx + 2, ny + 2)


@dataclass
class State:
    """Host copy of the reference's ``FS`` (src/FS.hpp:28-52) as numpy arrays."""
    nx: int
    ny: int
    dx: float
    dy: float
    rho_gas: float = 1.0
    rho_liquid: float = 1.0
    visc_gas: float = 1e-3
    visc_liquid: float = 1e-3
    sigma: float = 0.0
    a: dict = field(default_factory=dict)

    def __post_init__(self):
        for n in ALL_FIELDS:
            if n not in self.a:
                self.a[n] = np.zeros(shape_of(n, self.nx, self.ny))
            else:
                self.a[n] = np.ascontiguousarray(self.a[n], dtype=np.float64)
                assert self.a[n].shape == shape_of(n, self.nx, self.ny), n

    def copy(self) -> "State":
        return State(self.nx, self.ny, self.dx, self.dy, self.rho_gas, self.rho_liquid, self.visc_gas,
                     self.visc_liquid, self.sigma, {k: v.copy() for k, v in self.a.items()})

    def cstruct(self) -> CFs:
        s = CFs(self.nx, self.ny, self.dx, self.dy, self.rho_gas, self.rho_liquid, self.visc_gas,
                self.visc_liquid, self.sigma)
        for n in ALL_FIELDS:
            setattr(s, n, self.a[n].ctypes.data_as(DP))
        return s

    def __getattr__(self, n):
        a = self.__dict__.get("a", {})
        if n in a:
            return a[n]
        raise AttributeError(n)


@dataclass
class BConds:
    """Mirror of ``FlowBConds`` (src/BoundaryConditions.hpp:207-212); order left,right,bottom,top."""
    kind: tuple = (NEUMANN,) * 4
    u_val: tuple = (0.0,) * 4
    v_val: tuple = (0.0,) * 4
    clipped: tuple = (0,) * 4
    u_prof: tuple = (None,) * 4
    v_prof: tuple = (None,) * 4

    def cstruct(self) -> CBconds:
        b = CBconds()
        self._keep = []
        for s in range(4):
            b.kind[s] = self.kind[s]
            b.u_val[s] = self.u_val[s]
            b.v_val[s] = self.v_val[s]
            b.clipped[s] = self.clipped[s]
            for nm, prof in (("u_prof", self.u_prof[s]), ("v_prof", self.v_prof[s])):
                if prof is not None:
                    arr = np.ascontiguousarray(prof, dtype=np.float64)
                    self._keep.append(arr)
                    getattr(b, nm)[s] = arr.ctypes.data_as(DP)
        return b


def _ptr(a):
    return a.ctypes.data_as(DP)


class Oracle:
    """oracle/liboracle.so"""

    def __init__(self, path: str = ORACLE_SO):
        self.lib = L = C.CDLL(path)
        L.fso_adjust_dt.restype = C.c_double
        L.fso_adjust_dt.argtypes = [C.POINTER(CFs), C.c_double, C.c_double]
        L.fso_outflow_fix.restype = C.c_double
        L.fso_flow_control.restype = C.c_double
        L.fso_flow_control.argtypes = [C.POINTER(CFs), C.c_int, C.c_double, DP]
        L.fso_abs_max.restype = C.c_double
        L.fso_integrate.restype = C.c_double
        L.fso_integrate.argtypes = [DP, C.c_int64, C.c_int64, C.c_double, C.c_double, C.c_int, C.c_int]
        L.fso_sample.argtypes = [C.c_int, DP, DP, C.c_int64, C.c_int64, DP, DP, C.c_int64, DP, DP, DP, DP]
        L.fso_grad_centered.argtypes = [DP, C.c_int64, C.c_int64, C.c_double, C.c_double, DP, DP]
        L.fso_abs_max.argtypes = [DP, C.c_int64, C.c_int64, C.c_int]
        L.fso_calc_mid_time.argtypes = [DP, DP, C.c_int64]
        L.fso_update_velocity.argtypes = [C.POINTER(CFs), DP, DP, C.c_double]
        L.fso_update_density.argtypes = [C.POINTER(CFs), DP, DP, C.c_double]
        L.fso_fill_pressure_operator.argtypes = [C.POINTER(CFs), C.c_int, DP]
        L.fso_fill_pressure_rhs.argtypes = [C.POINTER(CFs), DP, C.c_double, C.c_int, DP]
        L.fso_correct.argtypes = [C.POINTER(CFs), DP, C.c_double]
        L.fso_apply_neumann_bconds.argtypes = [DP, C.c_int64, C.c_int64]
        L.fso_calc_pressure_jump.argtypes = [C.POINTER(CFs), DP, DP, DP]
        L.fso_interpolate_UV.argtypes = [C.POINTER(CFs), DP, DP]
        L.fso_conserved.argtypes = [C.POINTER(CFs), DP, DP, DP]
        L.fso_correct_velocity_ib_implicit_euler.argtypes = [C.POINTER(CFs), DP, DP, C.c_double]
        L.fso_update_velocity_ib_semi_analytical.argtypes = [C.POINTER(CFs), DP, DP, C.c_double, DP, DP]
        L.fso_capillary_divergence.argtypes = [C.POINTER(CFs), DP, DP, DP, C.c_double, DP]
        L.fso_solver_create.restype = C.c_void_p
        L.fso_solver_create.argtypes = [C.c_int64, C.c_int64, C.POINTER(CSolverOpts)]
        L.fso_solver_destroy.argtypes = [C.c_void_p]
        L.fso_solver_set_operator.argtypes = [C.c_void_p, DP]
        L.fso_solver_solve.restype = C.c_int
        L.fso_solver_solve.argtypes = [C.c_void_p, DP, DP, DP]
        L.fso_solver_num_levels.argtypes = [C.c_void_p]
        L.fso_step.restype = C.c_int
        L.fso_step.argtypes = [C.POINTER(CFs), C.POINTER(CBconds), C.c_void_p, C.POINTER(CStepOpts),
                               C.c_double, DP]

    # each method mutates / returns numpy arrays; `st` is a State
    def adjust_dt(self, st, cfl, dt_max):
        return self.lib.fso_adjust_dt(C.byref(st.cstruct()), cfl, dt_max)

    def save_old_state(self, st):
        self.lib.fso_save_old_state(C.byref(st.cstruct()))

    def calc_mid_time(self, st):
        self.lib.fso_calc_mid_time(_ptr(st.U), _ptr(st.U_old), st.U.size)
        self.lib.fso_calc_mid_time(_ptr(st.V), _ptr(st.V_old), st.V.size)

    def calc_dmomdt(self, st):
        dU = np.zeros(shape_of("dmomU", st.nx, st.ny))
        dV = np.zeros(shape_of("dmomV", st.nx, st.ny))
        self.lib.fso_calc_dmomdt(C.byref(st.cstruct()), _ptr(dU), _ptr(dV))
        return dU, dV

    def calc_drhodt(self, st):
        dU = np.zeros(shape_of("dmomU", st.nx, st.ny))
        dV = np.zeros(shape_of("dmomV", st.nx, st.ny))
        self.lib.fso_calc_drhodt(C.byref(st.cstruct()), _ptr(dU), _ptr(dV))
        return dU, dV

    def update_velocity(self, st, dU, dV, dt):
        self.lib.fso_update_velocity(C.byref(st.cstruct()), _ptr(dU), _ptr(dV), dt)

    def update_density(self, st, dru, drv, dt):
        self.lib.fso_update_density(C.byref(st.cstruct()), _ptr(dru), _ptr(drv), dt)

    def calc_pressure_jump(self, st, vf, curv, ilen):
        self.lib.fso_calc_pressure_jump(C.byref(st.cstruct()), _ptr(vf), _ptr(curv), _ptr(ilen))

    def capillary_divergence(self, st, vf, curv, ilen, dt, div):
        self.lib.fso_capillary_divergence(C.byref(st.cstruct()), _ptr(vf), _ptr(curv), _ptr(ilen), dt, _ptr(div))

    def interpolate_UV(self, st):
        ui, vi = np.zeros((st.nx + 2, st.ny + 2)), np.zeros((st.nx + 2, st.ny + 2))
        self.lib.fso_interpolate_UV(C.byref(st.cstruct()), _ptr(ui), _ptr(vi))
        return ui, vi

    def conserved(self, st):
        a, b, c = C.c_double(0), C.c_double(0), C.c_double(0)
        self.lib.fso_conserved(C.byref(st.cstruct()), C.byref(a), C.byref(b), C.byref(c))
        return a.value, b.value, c.value

    def ib_implicit_euler(self, st, ibu, ibv, dt):
        self.lib.fso_correct_velocity_ib_implicit_euler(C.byref(st.cstruct()), _ptr(ibu), _ptr(ibv), dt)

    def ib_semi_analytical(self, st, dU, dV, dt, ibu, ibv):
        self.lib.fso_update_velocity_ib_semi_analytical(C.byref(st.cstruct()), _ptr(dU), _ptr(dV), dt,
                                                        _ptr(ibu), _ptr(ibv))

    def apply_velocity_bconds(self, st, bc):
        self.lib.fso_apply_velocity_bconds(C.byref(st.cstruct()), C.byref(bc.cstruct()))

    def outflow_fix(self, st):
        return self.lib.fso_outflow_fix(C.byref(st.cstruct()))

    def flow_control(self, st, mode, total_flow=0.0):
        """-> (inflow, outflow) before the correction"""
        io = np.zeros(2)
        self.lib.fso_flow_control(C.byref(st.cstruct()), mode, total_flow, _ptr(io))
        return io[0], io[1]

    def calc_divergence(self, st):
        div = np.zeros((st.nx + 2, st.ny + 2))
        self.lib.fso_calc_divergence(C.byref(st.cstruct()), _ptr(div))
        return div

    def fill_pressure_operator(self, st, psd=PSD_NONE):
        op = np.zeros((st.ny + 2, st.nx + 2, 5))  # Layout::F: [j][i][5]
        self.lib.fso_fill_pressure_operator(C.byref(st.cstruct()), psd, _ptr(op))
        return op

    def fill_pressure_rhs(self, st, div, dt, psd=PSD_NONE):
        rhs = np.zeros((st.ny + 2, st.nx + 2))  # Layout::F
        self.lib.fso_fill_pressure_rhs(C.byref(st.cstruct()), _ptr(div), dt, psd, _ptr(rhs))
        return rhs

    def shift_pressure_to_zero(self, st, dp):
        self.lib.fso_shift_pressure_to_zero(C.byref(st.cstruct()), _ptr(dp))

    def correct(self, st, dp, dt):
        self.lib.fso_correct(C.byref(st.cstruct()), _ptr(dp), dt)

    def abs_max(self, a, include_ghost=False):
        """abs_max<INCLUDE_GHOST = false>(field): interior entries unless include_ghost"""
        a = np.ascontiguousarray(a)
        return self.lib.fso_abs_max(_ptr(a), a.shape[0] - 2, a.shape[1] - 2, int(include_ghost))

    def calc_rho_visc_const(self, st):
        self.lib.fso_calc_rho_visc_const(C.byref(st.cstruct()))

    def calc_rho_visc_vof(self, st, vf):
        self.lib.fso_calc_rho_visc_vof(C.byref(st.cstruct()), _ptr(vf))

    def integrate(self, a, dx, dy, include_ghost=False, l1=False):
        """integrate<INCLUDE_GHOST> / L1_norm<INCLUDE_GHOST>(dx, dy, field), src/Operators.hpp:51-76"""
        a = np.ascontiguousarray(a)
        return self.lib.fso_integrate(_ptr(a), a.shape[0] - 2, a.shape[1] - 2, dx, dy, int(include_ghost), int(l1))

    def sample(self, mode, fa, fb, xm, ym, px, py):
        """mode 0: eval_flow_field_at(xm, ym, fa = Ui, fb = Vi, x, y) -> (u, v); mode 1: bilinear_interpolate(xm, ym, fa, x, y)
        -> values (src/Operators.hpp:85-161); xm, ym incl. ghost entries"""
        fa = np.ascontiguousarray(fa)
        px, py = np.ascontiguousarray(px, dtype=np.float64), np.ascontiguousarray(py, dtype=np.float64)
        xm, ym = np.ascontiguousarray(xm, dtype=np.float64), np.ascontiguousarray(ym, dtype=np.float64)
        oa, ob = np.zeros_like(px), np.zeros_like(px)
        if fb is not None:
            fb = np.ascontiguousarray(fb)
        self.lib.fso_sample(mode, _ptr(fa), _ptr(fb) if fb is not None else None, fa.shape[0] - 2, fa.shape[1] - 2,
                            _ptr(xm), _ptr(ym), px.size, _ptr(px), _ptr(py), _ptr(oa), _ptr(ob) if fb is not None else None)
        return (oa, ob) if fb is not None else oa

    def grad_centered(self, f, dx, dy):
        """calc_grad_of_centered_points, src/Operators.hpp:164-201 -> (dfdx, dfdy)"""
        f = np.ascontiguousarray(f)
        gx, gy = np.zeros_like(f), np.zeros_like(f)
        self.lib.fso_grad_centered(_ptr(f), f.shape[0] - 2, f.shape[1] - 2, dx, dy, _ptr(gx), _ptr(gy))
        return gx, gy

    def apply_neumann_bconds(self, f):
        self.lib.fso_apply_neumann_bconds(_ptr(f), f.shape[0] - 2, f.shape[1] - 2)

    def solver(self, nx, ny, tol=1e-6, max_iter=50, max_levels=16, relax_type=1, skip_relax=1, snap_zero_diag=0):
        return OracleSolver(self, nx, ny, CSolverOpts(tol, max_iter, max_levels, relax_type, skip_relax, snap_zero_diag))

    def step(self, st, bc, solver, dt, num_subiter, outflow_fix=0, psd=PSD_NONE, rebuild=0, density=0,
             capillary=None, total_flow=0.0, ib=None):
        """capillary = (vf, curv, interface_length) cell arrays, or None; ib = (mode, ib_corr_u, ib_corr_v) or None"""
        res = C.c_double(0)
        o = _step_opts(num_subiter, outflow_fix, psd, rebuild, density, capillary, total_flow, ib)
        it = self.lib.fso_step(C.byref(st.cstruct()), C.byref(bc.cstruct()), solver.h, C.byref(o), dt,
                               C.byref(res))
        return it, res.value


class OracleSolver:
    def __init__(self, o: Oracle, nx, ny, opts: CSolverOpts):
        self.o, self.nx, self.ny = o, nx, ny
        self.h = C.c_void_p(o.lib.fso_solver_create(nx, ny, C.byref(opts)))

    def set_operator(self, op):
        op = np.ascontiguousarray(op)
        self.o.lib.fso_solver_set_operator(self.h, _ptr(op))

    def solve(self, rhs):
        sol = np.zeros((self.nx + 2, self.ny + 2))
        res = C.c_double(0)
        rhs = np.ascontiguousarray(rhs)
        it = self.o.lib.fso_solver_solve(self.h, _ptr(rhs), _ptr(sol), C.byref(res))
        return sol, res.value, it

    @property
    def num_levels(self):
        return self.o.lib.fso_solver_num_levels(self.h)

    def __del__(self):
        try:
            self.o.lib.fso_solver_destroy(self.h)
        except Exception:
            pass


def have_ref() -> bool:
    return os.path.exists(REF_SO)


def ref_variant(max_levels):
    """the reference rebuilt with -DPS_PFMG_MAX_LEVELS=<n> (its own knob,
    src/LinearSolver_StructHypre.hpp:221-229): oracle/_ref/libfsref_ml<n>.so, or None"""
    so = os.path.join(ROOT, "oracle", "_ref", f"libfsref_ml{max_levels}.so")
    if not os.path.exists(so):
        return None
    return type(f"RefML{max_levels}", (Ref,), {"_lib": None, "SO": so})


class Ref:
    """oracle/_ref/libfsref.so: one session = one reference ``FS`` + solver of a fixed size."""

    _lib = None
    SO = REF_SO

    @classmethod
    def lib(cls):
        if cls._lib is None:
            L = C.CDLL(cls.SO)
            L.fsref_create.restype = C.c_void_p
            L.fsref_create.argtypes = [C.c_int64, C.c_int64]
            L.fsref_supported.argtypes = [C.c_int64, C.c_int64]
            for n in ("destroy", "save_old_state", "calc_mid_time", "apply_velocity_bconds",
                      "solver_set_operator"):
                getattr(L, "fsref_" + n).argtypes = [C.c_void_p]
            L.fsref_load.argtypes = [C.c_void_p, C.POINTER(CFs)]
            L.fsref_store.argtypes = [C.c_void_p, C.POINTER(CFs)]
            L.fsref_set_bconds.argtypes = [C.c_void_p, C.POINTER(CBconds)]
            L.fsref_adjust_dt.restype = C.c_double
            L.fsref_adjust_dt.argtypes = [C.c_void_p, C.c_double, C.c_double]
            L.fsref_calc_dmomdt.argtypes = [C.c_void_p, DP, DP]
            L.fsref_calc_drhodt.argtypes = [C.c_void_p, DP, DP]
            L.fsref_update_velocity.argtypes = [C.c_void_p, DP, DP, C.c_double]
            L.fsref_calc_divergence.argtypes = [C.c_void_p, DP]
            L.fsref_fill_pressure_operator.argtypes = [C.c_void_p, C.c_int, DP]
            L.fsref_fill_pressure_rhs.argtypes = [C.c_void_p, DP, C.c_double, C.c_int, DP]
            L.fsref_shift_pressure_to_zero.argtypes = [C.c_void_p, DP]
            L.fsref_calc_rho_visc_vof.argtypes = [C.c_void_p, DP]
            L.fsref_solver_create.argtypes = [C.c_void_p, C.c_double, C.c_int, C.c_int]
            L.fsref_solver_solve.restype = C.c_int
            L.fsref_solver_solve.argtypes = [C.c_void_p, DP, C.c_double, DP, DP]
            L.fsref_step.restype = C.c_int
            L.fsref_step.argtypes = [C.c_void_p, C.POINTER(CStepOpts), C.c_double, DP, DP]
            L.fsref_diagnostics.argtypes = [C.c_void_p, DP, DP, DP]
            L.fsref_update_density.argtypes = [C.c_void_p, DP, DP, C.c_double]
            L.fsref_calc_pressure_jump.argtypes = [C.c_void_p, DP, DP, DP]
            L.fsref_interpolate_uv.argtypes = [C.c_void_p, DP, DP]
            L.fsref_conserved.argtypes = [C.c_void_p, DP, DP, DP]
            L.fsref_flow_control.argtypes = [C.c_void_p, C.c_int, C.c_double, DP]
            L.fsref_to_npy.argtypes = [C.c_void_p, C.c_char_p]
            L.fsref_ib_channel_correction.argtypes = [C.c_void_p, C.c_double, C.c_double, DP, DP]
            L.fsref_vf_drop.argtypes = [C.c_void_p, DP]
            L.fsref_sample.argtypes = [C.c_void_p, C.c_int, C.c_int, DP, DP, DP, DP]
            L.fsref_monitor_log.argtypes = [C.c_char_p, C.c_int, C.POINTER(C.c_char_p), C.c_int, C.POINTER(C.c_char_p),
                                            C.c_int, DP, C.POINTER(C.c_int64)]
            L.fsref_should_save.argtypes = [C.c_double] * 4
            L.fsref_integrate_cell.argtypes = [C.c_void_p, DP, C.c_int, C.c_int]
            L.fsref_integrate_cell.restype = C.c_double
            L.fsref_grad_centered.argtypes = [C.c_void_p, DP, DP, DP]
            L.fsref_write_vtk.argtypes = [C.c_void_p, C.c_char_p, C.c_double]
            L.fsref_flow_control.restype = C.c_double
            L.fsref_ib_implicit_euler.argtypes = [C.c_void_p, DP, DP, C.c_double]
            L.fsref_ib_semi_analytical.argtypes = [C.c_void_p, DP, DP, C.c_double, DP, DP]
            L.fsref_set_threads.argtypes = [C.c_int]
            cls._lib = L
        return cls._lib

    @classmethod
    def monitor_log(cls, filename, fnames, inames, frows, irows):
        """the reference's Monitor (src/Monitor.hpp) writing len(frows) rows; column order: all Float columns but the
        last, the Index columns, the last Float column"""
        fr = np.ascontiguousarray(frows, dtype=np.float64)
        ir = np.ascontiguousarray(irows, dtype=np.int64)
        fn = (C.c_char_p * len(fnames))(*[n.encode() for n in fnames])
        inn = (C.c_char_p * len(inames))(*[n.encode() for n in inames])
        cls.lib().fsref_monitor_log(os.fsencode(filename), len(fnames), fn, len(inames), inn, fr.shape[0], _ptr(fr),
                                    ir.ctypes.data_as(C.POINTER(C.c_int64)))

    @classmethod
    def should_save(cls, t, dt, dt_write, t_end):
        return bool(cls.lib().fsref_should_save(t, dt, dt_write, t_end))

    def __init__(self, nx, ny):
        L = self.lib()
        if not L.fsref_supported(nx, ny):
            raise ValueError(f"reference harness has no instantiation for {nx}x{ny}")
        self.nx, self.ny = nx, ny
        self.h = C.c_void_p(L.fsref_create(nx, ny))

    def __del__(self):
        try:
            self.lib().fsref_destroy(self.h)
        except Exception:
            pass

    def load(self, st):
        self.lib().fsref_load(self.h, C.byref(st.cstruct()))

    def store(self, st):
        self.lib().fsref_store(self.h, C.byref(st.cstruct()))

    def set_bconds(self, bc):
        self.lib().fsref_set_bconds(self.h, C.byref(bc.cstruct()))

    def adjust_dt(self, cfl, dt_max):
        return self.lib().fsref_adjust_dt(self.h, cfl, dt_max)

    def save_old_state(self):
        self.lib().fsref_save_old_state(self.h)

    def calc_mid_time(self):
        self.lib().fsref_calc_mid_time(self.h)

    def calc_dmomdt(self):
        dU = np.zeros(shape_of("dmomU", self.nx, self.ny))
        dV = np.zeros(shape_of("dmomV", self.nx, self.ny))
        self.lib().fsref_calc_dmomdt(self.h, _ptr(dU), _ptr(dV))
        return dU, dV

    def calc_drhodt(self):
        dU = np.zeros(shape_of("dmomU", self.nx, self.ny))
        dV = np.zeros(shape_of("dmomV", self.nx, self.ny))
        self.lib().fsref_calc_drhodt(self.h, _ptr(dU), _ptr(dV))
        return dU, dV

    def update_velocity(self, dU, dV, dt):
        self.lib().fsref_update_velocity(self.h, _ptr(dU), _ptr(dV), dt)

    def apply_velocity_bconds(self):
        self.lib().fsref_apply_velocity_bconds(self.h)

    def calc_divergence(self):
        div = np.zeros((self.nx + 2, self.ny + 2))
        self.lib().fsref_calc_divergence(self.h, _ptr(div))
        return div

    def fill_pressure_operator(self, psd=PSD_NONE):
        op = np.zeros((self.ny + 2, self.nx + 2, 5))
        self.lib().fsref_fill_pressure_operator(self.h, psd, _ptr(op))
        return op

    def fill_pressure_rhs(self, div, dt, psd=PSD_NONE):
        rhs = np.zeros((self.ny + 2, self.nx + 2))
        self.lib().fsref_fill_pressure_rhs(self.h, _ptr(div), dt, psd, _ptr(rhs))
        return rhs

    def shift_pressure_to_zero(self, dp):
        self.lib().fsref_shift_pressure_to_zero(self.h, _ptr(dp))

    def calc_rho_visc_vof(self, vf):
        self.lib().fsref_calc_rho_visc_vof(self.h, _ptr(vf))

    def solver_create(self, tol=1e-6, max_iter=50, psd=PSD_NONE):
        self.lib().fsref_solver_create(self.h, tol, max_iter, psd)

    def solver_set_operator(self):
        self.lib().fsref_solver_set_operator(self.h)

    def solver_solve(self, div, dt):
        dp = np.zeros((self.nx + 2, self.ny + 2))
        res = C.c_double(0)
        it = self.lib().fsref_solver_solve(self.h, _ptr(div), dt, _ptr(dp), C.byref(res))
        return dp, res.value, it

    def update_density(self, dru, drv, dt):
        self.lib().fsref_update_density(self.h, _ptr(dru), _ptr(drv), dt)

    def calc_pressure_jump(self, vf, curv, ilen):
        self.lib().fsref_calc_pressure_jump(self.h, _ptr(vf), _ptr(curv), _ptr(ilen))

    def interpolate_UV(self):
        ui, vi = np.zeros((self.nx + 2, self.ny + 2)), np.zeros((self.nx + 2, self.ny + 2))
        self.lib().fsref_interpolate_uv(self.h, _ptr(ui), _ptr(vi))
        return ui, vi

    def conserved(self):
        a, b, c = C.c_double(0), C.c_double(0), C.c_double(0)
        self.lib().fsref_conserved(self.h, C.byref(a), C.byref(b), C.byref(c))
        return a.value, b.value, c.value

    def ib_implicit_euler(self, ibu, ibv, dt):
        self.lib().fsref_ib_implicit_euler(self.h, _ptr(ibu), _ptr(ibv), dt)

    def ib_semi_analytical(self, dU, dV, dt, ibu, ibv):
        self.lib().fsref_ib_semi_analytical(self.h, _ptr(dU), _ptr(dV), dt, _ptr(ibu), _ptr(ibv))

    def integrate(self, a, include_ghost=False, l1=False):
        a = np.ascontiguousarray(a)
        return self.lib().fsref_integrate_cell(self.h, _ptr(a), int(include_ghost), int(l1))

    def sample(self, mode, px, py):
        """mode 0: eval_flow_field_at on Ui, Vi of the loaded state -> (u, v); mode 1: bilinear_interpolate of fs.p"""
        px, py = np.ascontiguousarray(px, dtype=np.float64), np.ascontiguousarray(py, dtype=np.float64)
        a, b = np.zeros_like(px), np.zeros_like(px)
        self.lib().fsref_sample(self.h, mode, px.size, _ptr(px), _ptr(py), _ptr(a), _ptr(b))
        return (a, b) if mode == 0 else a

    def grad_centered(self, f):
        f = np.ascontiguousarray(f)
        gx, gy = np.zeros_like(f), np.zeros_like(f)
        self.lib().fsref_grad_centered(self.h, _ptr(f), _ptr(gx), _ptr(gy))
        return gx, gy

    def ib_channel_correction(self, y0, y1):
        """calc_ib_correction_shape (src/IB.hpp:45-107) for solid walls below y0 and above y1 -> (ib_corr_u, ib_corr_v)"""
        ibu, ibv = np.zeros((self.nx + 3, self.ny + 2)), np.zeros((self.nx + 2, self.ny + 3))
        self.lib().fsref_ib_channel_correction(self.h, y0, y1, _ptr(ibu), _ptr(ibv))
        return ibu, ibv

    def vf_drop(self):
        """the initial volume fraction of test/StationaryDrop.cpp:47-52,209-212 on this session's grid"""
        vf = np.zeros((self.nx + 2, self.ny + 2))
        self.lib().fsref_vf_drop(self.h, _ptr(vf))
        return vf

    def to_npy(self, directory):
        """to_npy(directory, fs), src/IO.hpp:231-269 -- REMOVES and re-creates `directory`"""
        assert self.lib().fsref_to_npy(self.h, os.fsencode(directory)) == 0

    def write_vtk(self, directory, t):
        """the drivers' DataWriter (pressure, divergence, velocity) -> directory/state_NNNNNN.vtk"""
        assert self.lib().fsref_write_vtk(self.h, os.fsencode(directory), t) == 0

    def flow_control(self, mode, total_flow=0.0):
        io = np.zeros(2)
        self.lib().fsref_flow_control(self.h, mode, total_flow, _ptr(io))
        return io[0], io[1]

    def step(self, dt, num_subiter, outflow_fix=0, psd=PSD_NONE, rebuild=0, phases=None, density=0,
             capillary=None, total_flow=0.0, ib=None):
        res = C.c_double(0)
        o = _step_opts(num_subiter, outflow_fix, psd, rebuild, density, capillary, total_flow, ib)
        ph = _ptr(phases) if phases is not None else None
        it = self.lib().fsref_step(self.h, C.byref(o), dt, C.byref(res), ph)
        return it, res.value

    def diagnostics(self):
        a, b, c = C.c_double(0), C.c_double(0), C.c_double(0)
        self.lib().fsref_diagnostics(self.h, C.byref(a), C.byref(b), C.byref(c))
        return a.value, b.value, c.value
