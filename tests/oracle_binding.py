"""ctypes binding of the CPU oracle (oracle/_build/liboracle.so).  TEST INFRASTRUCTURE: only tests/,
__graft_entry__.smoke() and bench.py's CPU-baseline legs load this."""
import ctypes as C
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")

c_double_p = C.POINTER(C.c_double)
c_int_p = C.POINTER(C.c_int)
_VP = C.c_void_p


class SolverStats(C.Structure):
    _fields_ = [("iterations", C.c_int32), ("initial_residual", C.c_double), ("final_residual", C.c_double),
                ("tolerance", C.c_double)]


_libs = {}


def lib(fast=False):
    """fast=True: the -O3 -march=native build, compiled on this machine on first use (timed CPU baseline)."""
    name = "liboracle_fast.so" if fast else "liboracle.so"
    if name not in _libs:
        path = os.path.join(ORACLE_DIR, "_build", name)
        src = os.path.join(ORACLE_DIR, "oracle.cpp")
        stale = os.path.exists(path) and os.path.getmtime(path) < os.path.getmtime(src)
        if not os.path.exists(path) or stale or fast:
            env = dict(os.environ)
            env.pop("CXX", None)
            if fast and os.path.exists(path):
                os.remove(path)  # -march=native code must be built where it runs
            subprocess.run(["make", "fast" if fast else "all"], cwd=ORACLE_DIR, env=env, check=True,
                           stdout=subprocess.DEVNULL)
        h = C.CDLL(path)
        h.oracle_last_error.restype = C.c_char_p
        h.oracle_create.argtypes = [C.c_void_p, C.POINTER(_VP)]
        h.oracle_destroy.argtypes = [_VP]
        h.oracle_destroy.restype = None
        for f in ("oracle_set_field", "oracle_get_field", "oracle_get_operator_values"):
            getattr(h, f).argtypes = [_VP, C.c_int, C.c_int, c_double_p]
        for f in ("oracle_interpolate_velocity", "oracle_assemble_advection_term", "oracle_projection_step",
                  "oracle_update_pressure"):
            getattr(h, f).argtypes = [_VP]
        h.oracle_diffusion_step.argtypes = [_VP, C.c_int]
        h.oracle_advance.argtypes = [_VP, C.c_int, C.c_int, C.c_int, C.c_int]
        h.oracle_solver_stats.argtypes = [_VP, C.c_int, C.POINTER(SolverStats)]
        h.oracle_section_seconds.argtypes = [_VP, C.c_char_p, c_double_p]
        h.oracle_spmv.argtypes = [C.c_int, C.c_int, c_int_p, c_int_p, c_double_p, c_double_p, c_double_p, C.c_int, C.c_int]
        h.oracle_condense.argtypes = [C.c_int, c_int_p, c_int_p, c_double_p, c_double_p, C.c_int, c_int_p, c_double_p]
        h.oracle_ilu0_factor.argtypes = [C.c_int, c_int_p, c_int_p, c_double_p, C.c_double, c_double_p]
        h.oracle_ilu0_apply.argtypes = [C.c_int, c_int_p, c_int_p, c_double_p, c_double_p, c_double_p]
        h.oracle_gmres.argtypes = [C.c_int, c_int_p, c_int_p, c_double_p, c_double_p, c_double_p, c_double_p, C.c_double,
                                   C.c_int, C.c_int, C.POINTER(SolverStats)]
        h.oracle_cg.argtypes = [C.c_int, c_int_p, c_int_p, c_double_p, c_double_p, c_double_p, c_double_p, C.c_double,
                                C.c_int, C.POINTER(SolverStats)]
        _libs[name] = h
    return _libs[name]


def _d(a):
    return a.ctypes.data_as(c_double_p)


def _i(a):
    return a.ctypes.data_as(c_int_p)


def _f64(a):
    return np.ascontiguousarray(a, dtype=np.float64)


def _i32(a):
    return np.ascontiguousarray(a, dtype=np.int32)


class OracleError(RuntimeError):
    pass


def _check(h, rc):
    if rc != 0:
        raise OracleError(h.oracle_last_error().decode())


class Oracle:
    """CPU restatement of the time loop on the same problem description the device context takes."""

    def __init__(self, problem, fast=False, threads=1):
        self.h = lib(fast)
        self.h.oracle_set_threads(threads)
        p = problem.contents if hasattr(problem, "contents") else problem
        self.dim, self.nv, self.np_ = p.dim, p.n_vel, p.n_pres
        handle = _VP()
        _check(self.h, self.h.oracle_create(C.cast(C.pointer(p), C.c_void_p), C.byref(handle)))
        self.handle = handle

    def close(self):
        if self.handle is not None:
            self.h.oracle_destroy(self.handle)
        self.handle = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _len(self, field):
        return self.nv if field in (0, 1, 2, 3) else self.np_

    def set_field(self, field, values, component=0):
        values = _f64(values)
        assert values.shape == (self._len(field),)
        _check(self.h, self.h.oracle_set_field(self.handle, field, component, _d(values)))

    def get_field(self, field, component=0):
        out = np.empty(self._len(field))
        _check(self.h, self.h.oracle_get_field(self.handle, field, component, _d(out)))
        return out

    def get_operator_values(self, op, nnz, component=0):
        out = np.empty(nnz)
        _check(self.h, self.h.oracle_get_operator_values(self.handle, op, component, _d(out)))
        return out

    def interpolate_velocity(self):
        _check(self.h, self.h.oracle_interpolate_velocity(self.handle))

    def assemble_advection_term(self):
        _check(self.h, self.h.oracle_assemble_advection_term(self.handle))

    def diffusion_step(self, reinit_prec):
        _check(self.h, self.h.oracle_diffusion_step(self.handle, 1 if reinit_prec else 0))

    def projection_step(self):
        _check(self.h, self.h.oracle_projection_step(self.handle))

    def update_pressure(self):
        _check(self.h, self.h.oracle_update_pressure(self.handle))

    def advance(self, first_n, n_steps, start_n=2, start_offset_n=2):
        _check(self.h, self.h.oracle_advance(self.handle, first_n, n_steps, start_n, start_offset_n))

    def solver_stats(self, component):
        st = SolverStats()
        self.h.oracle_solver_stats(self.handle, component, C.byref(st))
        return st

    def section_seconds(self, name):
        s = C.c_double()
        self.h.oracle_section_seconds(self.handle, name.encode(), C.byref(s))
        return s.value


def spmv(n_rows, n_cols, rowptr, col, val, x, y=None, transpose=False):
    h = lib()
    rowptr, col, val, x = _i32(rowptr), _i32(col), _f64(val), _f64(x)
    add = y is not None
    out = _f64(y).copy() if add else np.zeros(n_cols if transpose else n_rows)
    _check(h, h.oracle_spmv(n_rows, n_cols, _i(rowptr), _i(col), _d(val), _d(x), _d(out), int(transpose), int(add)))
    return out


def condense(n, rowptr, col, val, rhs, dofs, values=None):
    h = lib()
    rowptr, col, dofs = _i32(rowptr), _i32(col), _i32(dofs)
    val, rhs = _f64(val).copy(), _f64(rhs).copy()
    vp = _d(_f64(values)) if values is not None else None
    _check(h, h.oracle_condense(n, _i(rowptr), _i(col), _d(val), _d(rhs), len(dofs), _i(dofs), vp))
    return val, rhs


def ilu0_factor(n, rowptr, col, val, strengthen=0.0):
    h = lib()
    rowptr, col, val = _i32(rowptr), _i32(col), _f64(val)
    lu = np.empty_like(val)
    _check(h, h.oracle_ilu0_factor(n, _i(rowptr), _i(col), _d(val), strengthen, _d(lu)))
    return lu


def ilu0_apply(n, rowptr, col, lu, src):
    h = lib()
    rowptr, col, lu, src = _i32(rowptr), _i32(col), _f64(lu), _f64(src)
    dst = np.empty_like(src)
    _check(h, h.oracle_ilu0_apply(n, _i(rowptr), _i(col), _d(lu), _d(src), _d(dst)))
    return dst


def gmres(n, rowptr, col, val, lu, x0, b, tol, max_its=1000, krylov=30):
    h = lib()
    rowptr, col, val, lu, b = _i32(rowptr), _i32(col), _f64(val), _f64(lu), _f64(b)
    x = _f64(x0).copy()
    st = SolverStats()
    rc = h.oracle_gmres(n, _i(rowptr), _i(col), _d(val), _d(lu), _d(x), _d(b), tol, max_its, krylov, C.byref(st))
    if rc not in (0, 3):
        _check(h, rc)
    return x, st, rc == 0


def cg(n, rowptr, col, val, lu, x0, b, tol, max_its=1000):
    h = lib()
    rowptr, col, val, b = _i32(rowptr), _i32(col), _f64(val), _f64(b)
    x = _f64(x0).copy()
    st = SolverStats()
    lup = _d(_f64(lu)) if lu is not None else None
    rc = h.oracle_cg(n, _i(rowptr), _i(col), _d(val), lup, _d(x), _d(b), tol, max_its, C.byref(st))
    if rc not in (0, 3):
        _check(h, rc)
    return x, st, rc == 0
