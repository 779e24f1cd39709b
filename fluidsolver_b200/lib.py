"""ctypes binding of libfsgpu.so: the reference's call surface (src/FS.hpp, src/Operators.hpp,
src/BoundaryConditions.hpp, src/LinearSolver_StructHypre.hpp) over the C ABI of include/fsgpu.h.

Method names and argument meaning follow the reference functions they stand for; host arrays are
numpy float64 in the reference's storage (ghost inclusive, j fastest).  The product path never
touches oracle/ and has no CPU fallback: if libfsgpu.so is missing or no CUDA device is present
the calls raise.
"""
from __future__ import annotations

import ctypes as C
import os
import weakref

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
# FSGPU_LIB: load another build of the same library (kernel-variant A/B runs); default in-tree
LIB_PATH = os.environ.get("FSGPU_LIB") or os.path.join(HERE, "libfsgpu.so")

FIELDS = ("VISC", "P", "P_JUMP_U", "P_JUMP_V", "RHO_U_OLD", "RHO_V_OLD", "U_OLD", "V_OLD", "RHO_U",
          "RHO_V", "U", "V", "DMOM_U", "DMOM_V", "DIV", "DELTA_P", "VF", "RHS",
          # extension fields, device memory allocated on first use
          "DRHO_U", "DRHO_V", "CURV", "INTERFACE_LENGTH", "PJ_OLD_U", "PJ_OLD_V", "UI", "VI", "RHOI",
          "IB_CORR_U", "IB_CORR_V", "GRAD_X", "GRAD_Y", "GRAD_XX", "GRAD_YY")
FIELD_ID = {n: i for i, n in enumerate(FIELDS)}
U_LIKE = {"P_JUMP_U", "RHO_U_OLD", "U_OLD", "RHO_U", "U", "DMOM_U", "DRHO_U", "PJ_OLD_U", "IB_CORR_U"}
V_LIKE = {"P_JUMP_V", "RHO_V_OLD", "V_OLD", "RHO_V", "V", "DMOM_V", "DRHO_V", "PJ_OLD_V", "IB_CORR_V"}

DIRICHLET, NEUMANN, PERIODIC, SYMMETRY = 0, 1, 2, 3
LEFT, RIGHT, BOTTOM, TOP = 0, 1, 2, 3
PSD_NONE, PSD_LEFT, PSD_RIGHT, PSD_BOTTOM, PSD_TOP = 0, 1, 2, 3, 4
RELAX_JACOBI = 1
STEP_OUTFLOW_FIX, STEP_REBUILD_OPERATOR, STEP_DENSITY, STEP_CAPILLARY = 1, 2, 4, 8
FLOW_OUTFLOW_FIX, FLOW_OUTFLOW_FIX_DY, FLOW_FORCE_TOTAL = 1, 2, 3

DP = C.POINTER(C.c_double)

# every symbol include/fsgpu.h declares (tests check the library exports all of them)
SYMBOLS = (
    "fsgpu_last_error", "fsgpu_launch_count", "fsgpu_device_count", "fsgpu_prof_enable",
    "fsgpu_prof_read", "fsgpu_create", "fsgpu_destroy",
    "fsgpu_set_props", "fsgpu_upload", "fsgpu_download", "fsgpu_fill", "fsgpu_sync",
    "fsgpu_device_field", "fsgpu_stream", "fsgpu_set_bconds", "fsgpu_adjust_dt",
    "fsgpu_save_old_state", "fsgpu_save_old_velocity", "fsgpu_save_old_density",
    "fsgpu_calc_mid_time", "fsgpu_calc_dmomdt", "fsgpu_update_velocity", "fsgpu_momentum_update",
    "fsgpu_apply_velocity_bconds", "fsgpu_outflow_fix", "fsgpu_flow_control", "fsgpu_set_flow_control",
    "fsgpu_field_max", "fsgpu_field_min", "fsgpu_count_nonfinite", "fsgpu_integrate", "fsgpu_l1_norm", "fsgpu_calc_grad_of_centered_points",
    "fsgpu_eval_flow_field_at", "fsgpu_bilinear_interpolate",
    "fsgpu_to_npy", "fsgpu_to_npy_state", "fsgpu_npy_write_host", "fsgpu_write_vtk", "fsgpu_calc_divergence",
    "fsgpu_shift_pressure_to_zero", "fsgpu_correct_velocity", "fsgpu_abs_max", "fsgpu_abs_max_all",
    "fsgpu_calc_rho_visc_const", "fsgpu_calc_rho_visc_vof", "fsgpu_apply_neumann_bconds",
    "fsgpu_ps_create", "fsgpu_ps_destroy", "fsgpu_ps_set_operator", "fsgpu_ps_set_rhs",
    "fsgpu_ps_solve", "fsgpu_ps_num_levels", "fsgpu_substep", "fsgpu_step",
    "fsgpu_comm_id", "fsgpu_create_slab", "fsgpu_slab_range", "fsgpu_slab_coarse_range",
    "fsgpu_ps_num_distributed_levels", "fsgpu_halo_is_direct",
    "fsgpu_calc_drhodt", "fsgpu_update_density", "fsgpu_density_step", "fsgpu_calc_pressure_jump",
    "fsgpu_capillary_divergence", "fsgpu_correct_velocity_ib_implicit_euler",
    "fsgpu_update_velocity_ib_semi_analytical", "fsgpu_interpolate_cell_centres",
    "fsgpu_calc_conserved_quantities", "fsgpu_monitor",
    "fsgpu_project", "fsgpu_ps_set_rhs_from_velocity", "fsgpu_selftest_division",
)
TRANSPORT_NCCL, TRANSPORT_LOCAL = 0, 1
COMM_ID_BYTES = 128


class PsOpts(C.Structure):
    _fields_ = [("tol", C.c_double), ("max_iter", C.c_int), ("psd", C.c_int), ("relax", C.c_int),
                ("max_levels", C.c_int), ("skip_relax", C.c_int)]


class FsgpuError(RuntimeError):
    pass


_lib = None


def load():
    """dlopen libfsgpu.so (built by fluidsolver_b200.build); raises if it is missing."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise FsgpuError(f"{LIB_PATH} not built: run `python -m fluidsolver_b200.build` "
                         "(there is no CPU fallback)")
    L = C.CDLL(LIB_PATH)
    vp, i64, dbl, i32 = C.c_void_p, C.c_int64, C.c_double, C.c_int
    L.fsgpu_last_error.restype = C.c_char_p
    L.fsgpu_launch_count.restype = i64
    L.fsgpu_device_count.argtypes = [C.POINTER(i32)]
    L.fsgpu_create.argtypes = [i64, i64, dbl, dbl, i32, C.POINTER(vp)]
    L.fsgpu_comm_id.argtypes = [C.c_char_p]
    L.fsgpu_create_slab.argtypes = [i64, i64, dbl, dbl, i32, i32, i32, C.c_char_p, i32, C.POINTER(vp)]
    L.fsgpu_slab_range.argtypes = [i64, i32, i32, C.POINTER(i64), C.POINTER(i64)]
    L.fsgpu_slab_coarse_range.argtypes = [i64, i64, i64, C.POINTER(i64), C.POINTER(i64)]
    L.fsgpu_ps_num_distributed_levels.argtypes = [vp, C.POINTER(i32)]
    L.fsgpu_set_props.argtypes = [vp, dbl, dbl, dbl, dbl, dbl]
    L.fsgpu_upload.argtypes = [vp, i32, DP]
    L.fsgpu_download.argtypes = [vp, i32, DP]
    L.fsgpu_fill.argtypes = [vp, i32, dbl]
    L.fsgpu_device_field.argtypes = [vp, i32, C.POINTER(DP), C.POINTER(i64), C.POINTER(i64),
                                     C.POINTER(i64)]
    L.fsgpu_stream.argtypes = [vp, C.POINTER(vp)]
    L.fsgpu_set_bconds.argtypes = [vp, i32, i32, dbl, dbl, DP, DP, i32]
    L.fsgpu_adjust_dt.argtypes = [vp, dbl, dbl, DP]
    for n in ("destroy", "sync", "save_old_state", "save_old_velocity", "save_old_density",
              "calc_mid_time", "calc_dmomdt", "apply_velocity_bconds", "calc_divergence",
              "shift_pressure_to_zero", "calc_rho_visc_const", "calc_rho_visc_vof", "calc_drhodt",
              "calc_pressure_jump", "interpolate_cell_centres"):
        getattr(L, "fsgpu_" + n).argtypes = [vp]
    for n in ("update_velocity", "momentum_update", "correct_velocity", "project", "update_density", "density_step",
              "capillary_divergence", "correct_velocity_ib_implicit_euler",
              "update_velocity_ib_semi_analytical"):
        getattr(L, "fsgpu_" + n).argtypes = [vp, dbl]
    L.fsgpu_calc_conserved_quantities.argtypes = [vp, DP, DP, DP]
    L.fsgpu_monitor.argtypes = [vp, DP, DP, DP]
    L.fsgpu_outflow_fix.argtypes = [vp, DP]
    L.fsgpu_flow_control.argtypes = [vp, C.c_int, C.c_double, DP, DP]
    L.fsgpu_set_flow_control.argtypes = [vp, C.c_int, C.c_double]
    L.fsgpu_to_npy.argtypes = [vp, i32, C.c_char_p]
    L.fsgpu_to_npy_state.argtypes = [vp, C.c_char_p, DP, DP]
    L.fsgpu_npy_write_host.argtypes = [C.c_char_p, DP, i64, i64]
    L.fsgpu_write_vtk.argtypes = [vp, C.c_char_p, dbl, DP, DP, i32, C.POINTER(i32), C.POINTER(C.c_char_p), i32,
                                  C.POINTER(i32), C.POINTER(C.c_char_p)]
    L.fsgpu_abs_max.argtypes = [vp, i32, DP]
    L.fsgpu_integrate.argtypes = [vp, i32, i32, DP]
    L.fsgpu_field_max.argtypes = [vp, i32, i32, DP]
    L.fsgpu_count_nonfinite.argtypes = [vp, i32, C.POINTER(i64), C.POINTER(i64)]
    L.fsgpu_field_min.argtypes = [vp, i32, i32, DP]
    L.fsgpu_l1_norm.argtypes = [vp, i32, i32, DP]
    L.fsgpu_calc_grad_of_centered_points.argtypes = [vp, i32, i32, i32]
    L.fsgpu_eval_flow_field_at.argtypes = [vp, DP, DP, i64, DP, DP, DP, DP]
    L.fsgpu_bilinear_interpolate.argtypes = [vp, i32, DP, DP, i64, DP, DP, DP]
    L.fsgpu_abs_max_all.argtypes = [vp, i32, DP]
    L.fsgpu_apply_neumann_bconds.argtypes = [vp, i32]
    L.fsgpu_ps_create.argtypes = [vp, C.POINTER(PsOpts), C.POINTER(vp)]
    L.fsgpu_ps_destroy.argtypes = [vp]
    L.fsgpu_ps_set_operator.argtypes = [vp]
    L.fsgpu_ps_set_rhs.argtypes = [vp, dbl]
    L.fsgpu_ps_set_rhs_from_velocity.argtypes = [vp, dbl]
    L.fsgpu_ps_solve.argtypes = [vp, DP, C.POINTER(i32)]
    L.fsgpu_ps_num_levels.argtypes = [vp, C.POINTER(i32)]
    L.fsgpu_substep.argtypes = [vp, vp, dbl, i32, DP, C.POINTER(i32)]
    L.fsgpu_step.argtypes = [vp, vp, dbl, i32, i32, DP, C.POINTER(i32)]
    _lib = L
    return L


def _chk(rc):
    if rc != 0:
        raise FsgpuError(load().fsgpu_last_error().decode())


def launch_count() -> int:
    return int(load().fsgpu_launch_count())


PROF_CLASSES = ("momentum", "bconds", "div_rhs", "shift_correct", "matvec_dot", "xr_update", "dot_rz",
                "p_update", "vcycle_down_l0", "vcycle_up_l0", "vcycle_down_coarse", "vcycle_up_coarse",
                "coarsest_solve", "operator_setup", "other")


def prof_enable(classes=()):
    """enable device timing for the named kernel classes (empty = off)"""
    mask = 0
    for c in classes:
        mask |= 1 << PROF_CLASSES.index(c)
    L = load()
    L.fsgpu_prof_enable.argtypes = [C.c_uint]
    _chk(L.fsgpu_prof_enable(mask))


def prof_read():
    """-> {class: (total_ms, count)} for classes with records; clears the records"""
    L = load()
    n = len(PROF_CLASSES)
    ms, cnt = (C.c_double * n)(), (C.c_int64 * n)()
    L.fsgpu_prof_read.argtypes = [C.POINTER(C.c_double), C.POINTER(C.c_int64), C.c_int]
    _chk(L.fsgpu_prof_read(ms, cnt, n))
    return {PROF_CLASSES[i]: (ms[i], cnt[i]) for i in range(n) if cnt[i]}


def selftest_division(y, n=1 << 24, seed=1):
    """-> (fast, mismatches): the kernels' division-free quotient x / y against the FP64 division on n random x"""
    L = load()
    fast, bad = C.c_int(), C.c_int64()
    L.fsgpu_selftest_division.argtypes = [C.c_double, C.c_int64, C.c_uint64, C.POINTER(C.c_int), C.POINTER(C.c_int64)]
    _chk(L.fsgpu_selftest_division(y, n, seed, C.byref(fast), C.byref(bad)))
    return fast.value, bad.value


def device_count() -> int:
    n = C.c_int(0)
    _chk(load().fsgpu_device_count(C.byref(n)))
    return n.value


def comm_id() -> bytes:
    """ncclGetUniqueId (rank 0 calls it and broadcasts the 128 bytes to the other ranks)"""
    buf = C.create_string_buffer(COMM_ID_BYTES)
    _chk(load().fsgpu_comm_id(buf))
    return buf.raw


def slab_range(nx: int, rank: int, nranks: int):
    """ghost-inclusive cell rows [a_lo, a_hi) of `rank` (host-only partition rule)"""
    lo, hi = C.c_int64(), C.c_int64()
    _chk(load().fsgpu_slab_range(nx, rank, nranks, C.byref(lo), C.byref(hi)))
    return lo.value, hi.value


def slab_coarse_range(lo: int, hi: int, foff: int):
    a, b = C.c_int64(), C.c_int64()
    _chk(load().fsgpu_slab_coarse_range(lo, hi, foff, C.byref(a), C.byref(b)))
    return a.value, b.value


def slab_rows(name: str, nx: int, ny: int, rank: int, nranks: int):
    """-> (r0, r1): rows of the GLOBAL host array of field `name` that fsgpu_upload reads on `rank`
    (its owned rows + the halo rows that exist); fsgpu_download writes the owned sub-range"""
    m = field_shape(name, nx, ny)[0]
    lo, hi = slab_range(nx, rank, nranks)
    return max(lo - 2, 0), min(hi + 2, m)


def field_shape(name: str, nx: int, ny: int):
    if name in U_LIKE:
        return (nx + 3, ny + 2)
    if name in V_LIKE:
        return (nx + 2, ny + 3)
    return (nx + 2, ny + 2)


def npy_write_host(filename, arr):
    """the host half of FS.to_npy (src/IO.hpp:25-95 + payload) on a 1-D or 2-D float64 array"""
    a = np.ascontiguousarray(arr, dtype=np.float64)
    assert a.ndim in (1, 2)
    _chk(load().fsgpu_npy_write_host(os.fsencode(filename), a.ctypes.data_as(DP), a.shape[0],
                                     a.shape[1] if a.ndim == 2 else 0))


class FS:
    """The reference's ``FS<Float,NX,NY,1>`` (src/FS.hpp:28-52) resident on one B200, or -- with
    rank / nranks / comm_id -- one slab of it (include/fsgpu.h "slab decomposition"): every method
    is then collective and host arrays stay GLOBAL arrays."""

    def __init__(self, nx, ny, dx, dy, rho_gas=1.0, rho_liquid=1.0, visc_gas=1e-3, visc_liquid=1e-3,
                 sigma=0.0, device=0, rank=0, nranks=1, comm_id=None, transport=TRANSPORT_NCCL):
        self.L = load()
        self.nx, self.ny, self.dx, self.dy = int(nx), int(ny), float(dx), float(dy)
        self.h = C.c_void_p()
        self._solvers = []  # weak refs: solvers hold device state of this context
        self.rank, self.nranks = int(rank), int(nranks)
        if self.nranks > 1:
            assert comm_id is not None and len(comm_id) == COMM_ID_BYTES
            _chk(self.L.fsgpu_create_slab(self.nx, self.ny, self.dx, self.dy, device, self.rank,
                                          self.nranks, comm_id, transport, C.byref(self.h)))
        else:
            _chk(self.L.fsgpu_create(self.nx, self.ny, self.dx, self.dy, device, C.byref(self.h)))
        self.set_props(rho_gas, rho_liquid, visc_gas, visc_liquid, sigma)

    def close(self):
        if getattr(self, "h", None) and self.h.value:
            for w in self._solvers:  # a solver must go before its context (any GC order)
                ps = w()
                if ps is not None:
                    ps.close()
            self._solvers = []
            self.L.fsgpu_destroy(self.h)
            self.h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def set_props(self, rho_gas, rho_liquid, visc_gas, visc_liquid, sigma=0.0):
        self.rho_gas, self.rho_liquid = rho_gas, rho_liquid
        self.visc_gas, self.visc_liquid, self.sigma = visc_gas, visc_liquid, sigma
        _chk(self.L.fsgpu_set_props(self.h, rho_gas, rho_liquid, visc_gas, visc_liquid, sigma))

    # ---- Field2D::get_data() ------------------------------------------------------------------
    def upload(self, name, arr):
        a = np.ascontiguousarray(arr, dtype=np.float64)
        assert a.shape == field_shape(name, self.nx, self.ny), (name, a.shape)
        _chk(self.L.fsgpu_upload(self.h, FIELD_ID[name], a.ctypes.data_as(DP)))

    def download(self, name, out=None):
        a = out if out is not None else np.empty(field_shape(name, self.nx, self.ny))
        _chk(self.L.fsgpu_download(self.h, FIELD_ID[name], a.ctypes.data_as(DP)))
        return a

    # slab decomposition: host buffers that hold only this rank's rows of the global array
    def local_rows(self, name):
        """-> (r0, r1): the rows of the global array fsgpu_upload reads on this rank (owned rows
        + halo rows); fsgpu_download writes a sub-range of them"""
        return slab_rows(name, self.nx, self.ny, self.rank, self.nranks)

    def _shifted(self, name, rows, r0):
        n = field_shape(name, self.nx, self.ny)[1]
        assert rows.dtype == np.float64 and rows.flags.c_contiguous and rows.shape[1] == n
        assert (r0, r0 + rows.shape[0]) == self.local_rows(name)
        return C.cast(C.c_void_p(rows.ctypes.data - r0 * n * 8), DP)  # base of the virtual global array

    def upload_rows(self, name, rows, r0):
        _chk(self.L.fsgpu_upload(self.h, FIELD_ID[name], self._shifted(name, rows, r0)))

    def download_rows(self, name, rows, r0):
        _chk(self.L.fsgpu_download(self.h, FIELD_ID[name], self._shifted(name, rows, r0)))

    def halo_is_direct(self) -> bool:
        """True if FSGPU_HALO=p2p took: halos / scalar reductions go through peer-mapped mailboxes"""
        on = C.c_int(0)
        self.L.fsgpu_halo_is_direct.argtypes = [C.c_void_p, C.POINTER(C.c_int)]
        _chk(self.L.fsgpu_halo_is_direct(self.h, C.byref(on)))
        return bool(on.value)

    def fill(self, name, value):
        _chk(self.L.fsgpu_fill(self.h, FIELD_ID[name], value))

    def sync(self):
        _chk(self.L.fsgpu_sync(self.h))

    def stream(self) -> int:
        s = C.c_void_p()
        _chk(self.L.fsgpu_stream(self.h, C.byref(s)))
        return s.value or 0

    def device_field(self, name):
        p, pitch, r0, c0 = DP(), C.c_int64(), C.c_int64(), C.c_int64()
        _chk(self.L.fsgpu_device_field(self.h, FIELD_ID[name], C.byref(p), C.byref(pitch),
                                       C.byref(r0), C.byref(c0)))
        return C.cast(p, C.c_void_p).value, pitch.value, r0.value, c0.value

    # ---- FlowBConds ---------------------------------------------------------------------------
    def set_bconds(self, side, kind, u=0.0, v=0.0, u_prof=None, v_prof=None, clipped=False):
        up = vp = None
        if u_prof is not None:
            u_prof = np.ascontiguousarray(u_prof, dtype=np.float64)
            up = u_prof.ctypes.data_as(DP)
        if v_prof is not None:
            v_prof = np.ascontiguousarray(v_prof, dtype=np.float64)
            vp = v_prof.ctypes.data_as(DP)
        _chk(self.L.fsgpu_set_bconds(self.h, side, kind, u, v, up, vp, int(clipped)))

    # ---- the reference's free functions -------------------------------------------------------
    def adjust_dt(self, cfl_max, dt_max):
        dt = C.c_double()
        _chk(self.L.fsgpu_adjust_dt(self.h, cfl_max, dt_max, C.byref(dt)))
        return dt.value

    def save_old_state(self):
        _chk(self.L.fsgpu_save_old_state(self.h))

    def save_old_velocity(self):
        _chk(self.L.fsgpu_save_old_velocity(self.h))

    def save_old_density(self):
        _chk(self.L.fsgpu_save_old_density(self.h))

    def calc_mid_time(self):
        _chk(self.L.fsgpu_calc_mid_time(self.h))

    def calc_dmomdt(self):
        _chk(self.L.fsgpu_calc_dmomdt(self.h))

    def update_velocity(self, dt):
        _chk(self.L.fsgpu_update_velocity(self.h, dt))

    def momentum_update(self, dt):
        _chk(self.L.fsgpu_momentum_update(self.h, dt))

    def apply_velocity_bconds(self):
        _chk(self.L.fsgpu_apply_velocity_bconds(self.h))

    def outflow_fix(self, want_error=True):
        e = C.c_double()
        _chk(self.L.fsgpu_outflow_fix(self.h, C.byref(e) if want_error else None))
        return e.value

    def flow_control(self, mode, total_flow=0.0):
        """the channel drivers' ghost-column corrections (FLOW_OUTFLOW_FIX, FLOW_OUTFLOW_FIX_DY, FLOW_FORCE_TOTAL);
        returns (inflow, outflow) before the correction"""
        i, o = C.c_double(), C.c_double()
        _chk(self.L.fsgpu_flow_control(self.h, mode, total_flow, C.byref(i), C.byref(o)))
        return i.value, o.value

    def set_flow_control(self, mode, total_flow=0.0):
        """what STEP_OUTFLOW_FIX applies inside substep()/step()"""
        _chk(self.L.fsgpu_set_flow_control(self.h, mode, total_flow))

    def calc_divergence(self):
        _chk(self.L.fsgpu_calc_divergence(self.h))

    def shift_pressure_to_zero(self):
        _chk(self.L.fsgpu_shift_pressure_to_zero(self.h))

    def correct_velocity(self, dt):
        _chk(self.L.fsgpu_correct_velocity(self.h, dt))

    def project(self, dt):
        """shift_pressure_to_zero + p += delta_p + velocity correction in one pass (the kernel fsgpu_substep runs)"""
        _chk(self.L.fsgpu_project(self.h, dt))

    def abs_max(self, name, include_ghost=False):
        """abs_max<INCLUDE_GHOST = false>(field), src/Utility.hpp:44-88: interior entries unless include_ghost"""
        o = C.c_double()
        fn = self.L.fsgpu_abs_max_all if include_ghost else self.L.fsgpu_abs_max
        _chk(fn(self.h, FIELD_ID[name], C.byref(o)))
        return o.value

    def count_nonfinite(self, name):
        """(number of NaN, number of +-Inf) entries of the ghost-inclusive array: has_nan / has_inf, src/Container.hpp:191-204"""
        a, b = C.c_int64(), C.c_int64()
        _chk(self.L.fsgpu_count_nonfinite(self.h, FIELD_ID[name], C.byref(a), C.byref(b)))
        return a.value, b.value

    def max(self, name, include_ghost=False):
        """max<INCLUDE_GHOST>(field), src/Utility.hpp:91-101"""
        o = C.c_double()
        _chk(self.L.fsgpu_field_max(self.h, FIELD_ID[name], int(include_ghost), C.byref(o)))
        return o.value

    def min(self, name, include_ghost=False):
        """min<INCLUDE_GHOST>(field), src/Utility.hpp:103-115"""
        o = C.c_double()
        _chk(self.L.fsgpu_field_min(self.h, FIELD_ID[name], int(include_ghost), C.byref(o)))
        return o.value

    def integrate(self, name, include_ghost=False):
        """integrate<INCLUDE_GHOST>(dx, dy, field), src/Operators.hpp:51-62"""
        o = C.c_double()
        _chk(self.L.fsgpu_integrate(self.h, FIELD_ID[name], int(include_ghost), C.byref(o)))
        return o.value

    def l1_norm(self, name, include_ghost=False):
        """L1_norm<INCLUDE_GHOST>(dx, dy, field), src/Operators.hpp:65-76"""
        o = C.c_double()
        _chk(self.L.fsgpu_l1_norm(self.h, FIELD_ID[name], int(include_ghost), C.byref(o)))
        return o.value

    def calc_grad_of_centered_points(self, f, dfdx="GRAD_X", dfdy="GRAD_Y"):
        """src/Operators.hpp:164-201 on cell fields"""
        _chk(self.L.fsgpu_calc_grad_of_centered_points(self.h, FIELD_ID[f], FIELD_ID[dfdx], FIELD_ID[dfdy]))

    def _centres(self, x, y):
        x, y = self._coords(x, y)
        assert x.size == self.nx + 3 and y.size == self.ny + 3
        return np.ascontiguousarray((x[:-1] + x[1:]) / 2), np.ascontiguousarray((y[:-1] + y[1:]) / 2)  # init_grid, FS.hpp:71-72

    def eval_flow_field_at(self, x, y, px, py):
        """eval_flow_field_at(xm, ym, Ui, Vi, x, y), src/Operators.hpp:118-161, at the points (px, py) -> (u, v);
        x, y: fs.x / fs.y incl. ghosts (grid_coords())"""
        x, y = self._centres(x, y)
        px, py = self._coords(px, py)
        u, v = np.zeros_like(px), np.zeros_like(px)
        _chk(self.L.fsgpu_eval_flow_field_at(self.h, x.ctypes.data_as(DP), y.ctypes.data_as(DP), px.size,
                                             px.ctypes.data_as(DP), py.ctypes.data_as(DP), u.ctypes.data_as(DP),
                                             v.ctypes.data_as(DP)))
        return u, v

    def bilinear_interpolate(self, name, x, y, px, py):
        """bilinear_interpolate(xm, ym, field, x, y), src/Operators.hpp:85-115, on a cell field"""
        x, y = self._centres(x, y)
        px, py = self._coords(px, py)
        out = np.zeros_like(px)
        _chk(self.L.fsgpu_bilinear_interpolate(self.h, FIELD_ID[name], x.ctypes.data_as(DP), y.ctypes.data_as(DP),
                                               px.size, px.ctypes.data_as(DP), py.ctypes.data_as(DP),
                                               out.ctypes.data_as(DP)))
        return out

    def calc_rho_visc_const(self):
        _chk(self.L.fsgpu_calc_rho_visc_const(self.h))

    def calc_rho_visc_vof(self):
        _chk(self.L.fsgpu_calc_rho_visc_vof(self.h))

    def apply_neumann_bconds(self, name):
        _chk(self.L.fsgpu_apply_neumann_bconds(self.h, FIELD_ID[name]))

    # ---- rows either side of the single-phase path ---------------------------------
    def calc_drhodt(self):
        _chk(self.L.fsgpu_calc_drhodt(self.h))

    def update_density(self, dt):
        _chk(self.L.fsgpu_update_density(self.h, dt))

    def density_step(self, dt):
        _chk(self.L.fsgpu_density_step(self.h, dt))

    def calc_pressure_jump(self):
        _chk(self.L.fsgpu_calc_pressure_jump(self.h))

    def capillary_divergence(self, dt):
        _chk(self.L.fsgpu_capillary_divergence(self.h, dt))

    def correct_velocity_ib_implicit_euler(self, dt):
        _chk(self.L.fsgpu_correct_velocity_ib_implicit_euler(self.h, dt))

    def update_velocity_ib_semi_analytical(self, dt):
        _chk(self.L.fsgpu_update_velocity_ib_semi_analytical(self.h, dt))

    def interpolate_cell_centres(self):
        _chk(self.L.fsgpu_interpolate_cell_centres(self.h))

    def calc_conserved_quantities(self):
        m, mx, my = C.c_double(), C.c_double(), C.c_double()
        _chk(self.L.fsgpu_calc_conserved_quantities(self.h, C.byref(m), C.byref(mx), C.byref(my)))
        return m.value, mx.value, my.value

    def monitor(self):
        """-> (abs_max(U), abs_max(V), abs_max(div)) over the interior entries of each field, as the
        drivers' monitor row; div (all cells) is stored in field DIV"""
        a, b, d = C.c_double(), C.c_double(), C.c_double()
        _chk(self.L.fsgpu_monitor(self.h, C.byref(a), C.byref(b), C.byref(d)))
        return a.value, b.value, d.value

    # ---- fused path ---------------------------------------------------------------------------
    # ---- file output from the device buffers (src/IO.hpp, src/VTKWriter.hpp) -----------------------
    @staticmethod
    def _coords(x, y):
        x = np.ascontiguousarray(x, dtype=np.float64)
        y = np.ascontiguousarray(y, dtype=np.float64)
        return x, y

    def grid_coords(self, x_min=0.0, y_min=0.0):
        """fs.x / fs.y of init_grid (src/FS.hpp:55-73) including their ghost entries"""
        return (x_min + np.arange(-1, self.nx + 2, dtype=np.float64) * self.dx,
                y_min + np.arange(-1, self.ny + 2, dtype=np.float64) * self.dy)

    def to_npy(self, name, filename):
        _chk(self.L.fsgpu_to_npy(self.h, FIELD_ID[name], os.fsencode(filename)))

    def to_npy_state(self, directory, x=None, y=None):
        """to_npy(directory, fs), src/IO.hpp:231-269 (x, y: fs.x / fs.y incl. ghosts, or None to skip them)"""
        if x is None or y is None:
            _chk(self.L.fsgpu_to_npy_state(self.h, os.fsencode(directory), None, None))
            return
        x, y = self._coords(x, y)
        assert x.size == self.nx + 3 and y.size == self.ny + 3
        _chk(self.L.fsgpu_to_npy_state(self.h, os.fsencode(directory), x.ctypes.data_as(DP), y.ctypes.data_as(DP)))

    def write_vtk(self, filename, t, x, y, scalars=(), vectors=()):
        """VTKWriter::write, src/VTKWriter.hpp:127-152. scalars: [(vtk name, field)], vectors: [(vtk name,
        x field, y field)] with cell fields ("P", "DIV", "UI", "VI", ...)"""
        x, y = self._coords(x, y)
        assert x.size == self.nx + 3 and y.size == self.ny + 3
        ns, nv = len(scalars), len(vectors)
        sf = (C.c_int32 * max(ns, 1))(*[FIELD_ID[f] for _, f in scalars])
        sn = (C.c_char_p * max(ns, 1))(*[n.encode() for n, _ in scalars])
        vf = (C.c_int32 * max(2 * nv, 1))(*[FIELD_ID[f] for _, fx, fy in vectors for f in (fx, fy)])
        vn = (C.c_char_p * max(nv, 1))(*[n.encode() for n, _, _ in vectors])
        _chk(self.L.fsgpu_write_vtk(self.h, os.fsencode(filename), t, x.ctypes.data_as(DP), y.ctypes.data_as(DP),
                                    ns, sf, sn, nv, vf, vn))

    def substep(self, ps, dt, flags=0):
        res, it = C.c_double(), C.c_int()
        _chk(self.L.fsgpu_substep(self.h, ps.h, dt, flags, C.byref(res), C.byref(it)))
        return res.value, it.value

    def step(self, ps, dt, num_subiter, flags=0):
        res, it = C.c_double(), C.c_int()
        _chk(self.L.fsgpu_step(self.h, ps.h, dt, num_subiter, flags, C.byref(res), C.byref(it)))
        return res.value, it.value


class LinearSolver:
    """``LinearSolver_StructHypre<Float,NX,NY,1>`` (src/LinearSolver_StructHypre.hpp:28-192)."""

    def __init__(self, fs: FS, tol=1e-6, max_iter=50, psd=PSD_NONE, relax=0, max_levels=0,
                 skip_relax=-1):
        self.fs, self.L = fs, fs.L
        self.h = C.c_void_p()
        o = PsOpts(tol, max_iter, psd, relax, max_levels, skip_relax)
        _chk(self.L.fsgpu_ps_create(fs.h, C.byref(o), C.byref(self.h)))
        fs._solvers.append(weakref.ref(self))

    def close(self):
        if getattr(self, "h", None) and self.h.value:
            self.L.fsgpu_ps_destroy(self.h)
            self.h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def set_pressure_operator(self):
        _chk(self.L.fsgpu_ps_set_operator(self.h))

    def set_pressure_rhs(self, dt):
        _chk(self.L.fsgpu_ps_set_rhs(self.h, dt))

    def set_pressure_rhs_from_velocity(self, dt):
        """calc_divergence + set_pressure_rhs in one pass (the kernel fsgpu_substep runs)"""
        _chk(self.L.fsgpu_ps_set_rhs_from_velocity(self.h, dt))

    def solve(self):
        """-> (residual, num_iter); the solution is field DELTA_P"""
        res, it = C.c_double(), C.c_int()
        _chk(self.L.fsgpu_ps_solve(self.h, C.byref(res), C.byref(it)))
        return res.value, it.value

    @property
    def num_distributed_levels(self):
        n = C.c_int()
        _chk(self.L.fsgpu_ps_num_distributed_levels(self.h, C.byref(n)))
        return n.value

    @property
    def num_levels(self):
        n = C.c_int()
        _chk(self.L.fsgpu_ps_num_levels(self.h, C.byref(n)))
        return n.value
