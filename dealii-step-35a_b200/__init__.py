"""ctypes binding of libstep35a_b200.so (include/step35a_b200.h, include/step35a_host.h).

The shared library holds the sm_100a kernels, their C ABI and the C++ host side (mesh input, DoF
numbering, one-off operators, the NavierStokesProjection driver).  This module only marshals numpy
arrays across that C ABI; it contains no numerics and no fallback: if the library is missing, loading
fails, and on a machine without a CUDA device every compute entry point raises S35Error.

The directory name has a hyphen; load it with ``__graft_entry__.load_package()``.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libstep35a_b200.so")
HOST_LIB_PATH = os.path.join(os.path.dirname(os.path.abspath(__file__)), "libstep35a_host.so")

c_double_p = C.POINTER(C.c_double)
c_int32_p = C.POINTER(C.c_int32)

STATUS_NAMES = {0: "OK", 1: "INVALID", 2: "CUDA", 3: "NO_CONVERGENCE", 4: "ZERO_PIVOT", 5: "UNSUPPORTED", 6: "IO"}

# s35_field
U_N, U_N_MINUS_1, U_STAR, FORCE, PRES_N, PRES_N_MINUS_1, PHI_N, PHI_N_MINUS_1, PRES_TMP = range(9)
# s35_operator
VEL_ADVECTION, VEL_IT_MATRIX, VEL_ILU = range(3)
# s35_bench_kernel
(BENCH_SPMV_VEL, BENCH_SPMV_PRES, BENCH_SPMV_GRAD, BENCH_SPMV_GRAD_T, BENCH_ASSEMBLE, BENCH_ILU_APPLY_VEL,
 BENCH_ILU_APPLY_PRES, BENCH_DOT, BENCH_ADD_AND_DOT, BENCH_ILU_FACTOR_VEL, BENCH_FP64_FMA) = range(11)

SECTIONS = ("interpolate_velocity", "setup_diffusion_rhs", "solve_diffusion_component", "assemble_advection_term",
            "projection_step_setup", "projection_step_solve", "update_pressure")


class S35Error(RuntimeError):
    def __init__(self, status, message):
        super().__init__("%s: %s" % (STATUS_NAMES.get(status, status), message))
        self.status = status


class SolverStats(C.Structure):
    _fields_ = [("iterations", C.c_int32), ("initial_residual", C.c_double), ("final_residual", C.c_double),
                ("tolerance", C.c_double)]


class Halo(C.Structure):
    _fields_ = [("n_neighbors", C.c_int32), ("neighbor_rank", c_int32_p), ("send_ptr", c_int32_p), ("send_idx", c_int32_p),
                ("recv_ptr", c_int32_p)]


class Dist(C.Structure):
    _fields_ = [("rank", C.c_int32), ("n_ranks", C.c_int32), ("n_vel_owned", C.c_int32), ("n_pres_owned", C.c_int32),
                ("n_vel_global", C.c_int64), ("n_pres_global", C.c_int64), ("vel_halo", Halo), ("pres_halo", Halo),
                ("vel_global_index", c_int32_p), ("pres_global_index", c_int32_p),
                ("gradT_rowptr", c_int32_p), ("gradT_col", c_int32_p), ("gradT_val", c_double_p * 3)]


class Problem(C.Structure):
    _fields_ = [
        ("dim", C.c_int32), ("n_cells", C.c_int32), ("n_vel", C.c_int32), ("n_pres", C.c_int32),
        ("cell_vertices", c_double_p), ("vel_cell_dofs", c_int32_p), ("cell_color", c_int32_p),
        ("velc_rowptr", c_int32_p), ("velc_col", c_int32_p),
        ("velm_rowptr", c_int32_p), ("velm_col", c_int32_p), ("velm_val", c_double_p),
        ("grad_rowptr", c_int32_p), ("grad_col", c_int32_p), ("grad_val", c_double_p * 3),
        ("plap_rowptr", c_int32_p), ("plap_col", c_int32_p), ("plap_val", c_double_p),
        ("pmass_rowptr", c_int32_p), ("pmass_col", c_int32_p), ("pmass_val", c_double_p),
        ("n_wall", C.c_int32), ("wall_dofs", c_int32_p),
        ("n_flow", C.c_int32 * 3), ("flow_dofs", c_int32_p * 3), ("flow_vals", c_double_p * 3),
        ("n_proj", C.c_int32), ("proj_dofs", c_int32_p),
        ("dt", C.c_double), ("reynolds", C.c_double), ("vel_eps", C.c_double), ("vel_diag_strength", C.c_double),
        ("vel_max_iterations", C.c_int32), ("vel_krylov_size", C.c_int32), ("vel_off_diagonals", C.c_int32),
        ("vel_update_prec", C.c_int32), ("rotational", C.c_int32), ("ilu_blocks", C.c_int32),
        ("dist", C.POINTER(Dist)),
    ]


class Params(C.Structure):
    _fields_ = [
        ("rotational", C.c_int32),
        ("initial_time", C.c_double), ("final_time", C.c_double), ("reynolds", C.c_double),
        ("dt", C.c_double),
        ("n_global_refines", C.c_int32), ("pressure_degree", C.c_int32),
        ("vel_max_iterations", C.c_int32), ("vel_krylov_size", C.c_int32), ("vel_off_diagonals", C.c_int32),
        ("vel_update_prec", C.c_int32),
        ("vel_eps", C.c_double), ("vel_diag_strength", C.c_double),
        ("verbose", C.c_int32), ("output_interval", C.c_int32),
        ("load_checkpoint_file_name", C.c_char * 512), ("save_checkpoint_file_name", C.c_char * 512),
    ]


class DiscInfo(C.Structure):
    _fields_ = [
        ("dim", C.c_int32), ("n_cells", C.c_int32), ("n_vertices", C.c_int32), ("n_vel", C.c_int32),
        ("n_pres", C.c_int32), ("n_colors", C.c_int32),
        ("nnz_vel_constrained", C.c_int64), ("nnz_vel_unconstrained", C.c_int64), ("nnz_grad", C.c_int64),
        ("nnz_proj", C.c_int64), ("nnz_pres", C.c_int64),
        ("width", C.c_double), ("height", C.c_double), ("min_cell_diameter", C.c_double), ("setup_seconds", C.c_double),
    ]


_lib = None

# every symbol the two headers declare: (name, restype, argtypes)
_VP = C.c_void_p
_SIGNATURES = [
    ("s35_last_error", C.c_char_p, []),
    ("s35_version", C.c_char_p, []),
    ("s35_device_count", C.c_int, []),
    ("s35_set_device", C.c_int, [C.c_int]),
    ("s35_matrix_create", C.c_int, [C.c_int32, C.c_int32, c_int32_p, c_int32_p, c_double_p, C.POINTER(_VP)]),
    ("s35_matrix_destroy", None, [_VP]),
    ("s35_matrix_set_values", C.c_int, [_VP, c_double_p]),
    ("s35_matrix_get_values", C.c_int, [_VP, c_double_p]),
    ("s35_matrix_nnz", C.c_int64, [_VP]),
    ("s35_spmv", C.c_int, [_VP, c_double_p, c_double_p, C.c_int, C.c_int]),
    ("s35_condense", C.c_int, [_VP, c_double_p, C.c_int32, c_int32_p, c_double_p]),
    ("s35_ilu_create", C.c_int, [_VP, C.POINTER(_VP)]),
    ("s35_ilu_destroy", None, [_VP]),
    ("s35_ilu_factor", C.c_int, [_VP, _VP, C.c_double]),
    ("s35_ilu_apply", C.c_int, [_VP, c_double_p, c_double_p]),
    ("s35_ilu_get_values", C.c_int, [_VP, c_double_p]),
    ("s35_ilu_levels", C.c_int, [_VP, c_int32_p, c_int32_p]),
    ("s35_gmres", C.c_int, [_VP, _VP, c_double_p, c_double_p, C.c_double, C.c_int32, C.c_int32, C.POINTER(SolverStats)]),
    ("s35_cg", C.c_int, [_VP, _VP, c_double_p, c_double_p, C.c_double, C.c_int32, C.POINTER(SolverStats)]),
    ("s35_vec_lincomb", C.c_int, [C.c_int64, C.c_double, c_double_p, C.c_double, c_double_p, c_double_p]),
    ("s35_vec_dot", C.c_int, [C.c_int64, c_double_p, c_double_p, c_double_p]),
    ("s35_vec_add_and_dot", C.c_int, [C.c_int64, C.c_double, c_double_p, c_double_p, c_double_p, c_double_p]),
    ("s35_vec_dot_device_noalloc", C.c_int, [C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    ("s35_comm_unique_id", C.c_int, [C.c_char_p]),
    ("s35_comm_init", C.c_int, [C.c_int32, C.c_int32, C.c_char_p]),
    ("s35_comm_finalize", None, []),
    ("s35_comm_thread_group", C.c_int, [C.c_int32]),
    ("s35_comm_thread_bind", C.c_int, [C.c_int32]),
    ("s35_ctx_create", C.c_int, [C.POINTER(Problem), C.POINTER(_VP)]),
    ("s35_ctx_destroy", None, [_VP]),
    ("s35_ctx_set_field", C.c_int, [_VP, C.c_int, C.c_int32, c_double_p]),
    ("s35_ctx_get_field", C.c_int, [_VP, C.c_int, C.c_int32, c_double_p]),
    ("s35_ctx_get_operator_values", C.c_int, [_VP, C.c_int, C.c_int32, c_double_p]),
    ("s35_interpolate_velocity", C.c_int, [_VP]),
    ("s35_assemble_advection_term", C.c_int, [_VP]),
    ("s35_diffusion_step", C.c_int, [_VP, C.c_int]),
    ("s35_projection_step", C.c_int, [_VP]),
    ("s35_update_pressure", C.c_int, [_VP]),
    ("s35_advance", C.c_int, [_VP, C.c_int32, C.c_int32, C.c_int32, C.c_int32, c_double_p]),
    ("s35_synchronize", C.c_int, [_VP]),
    ("s35_ctx_solver_stats", C.c_int, [_VP, C.c_int32, C.POINTER(SolverStats)]),
    ("s35_ctx_section_ms", C.c_int, [_VP, C.c_char_p, c_double_p, C.POINTER(C.c_int64)]),
    ("s35_ctx_bench_kernel", C.c_int, [_VP, C.c_int, C.c_int32, C.c_int32, C.c_int, c_double_p, c_double_p,
                                       C.POINTER(C.c_int64)]),
    ("s35_launch_count", C.c_int64, []),
    # host side
    ("s35_params_defaults", None, [C.POINTER(Params)]),
    ("s35_params_read", C.c_int, [C.c_char_p, C.POINTER(Params)]),
    ("s35_params_parse", C.c_int, [C.c_char_p, C.POINTER(Params)]),
    ("s35_disc_create", C.c_int, [C.c_char_p, C.c_int32, C.POINTER(Params), C.POINTER(_VP)]),
    ("s35_disc_destroy", None, [_VP]),
    ("s35_disc_problem", C.POINTER(Problem), [_VP]),
    ("s35_disc_get_info", C.c_int, [_VP, C.POINTER(DiscInfo)]),
    ("s35_disc_support_points", c_double_p, [_VP, C.c_int32]),
    ("s35_disc_initial_pressure", c_double_p, [_VP]),
    ("s35_disc_mesh", C.c_int, [_VP, C.POINTER(c_double_p), C.POINTER(c_int32_p), C.POINTER(c_int32_p),
                                C.POINTER(c_int32_p), C.POINTER(C.c_int64)]),
    ("s35_disc_pres_cell_dofs", c_int32_p, [_VP]),
    ("s35_part_create", C.c_int, [_VP, C.c_int32, C.c_int32, C.POINTER(_VP)]),
    ("s35_part_destroy", None, [_VP]),
    ("s35_part_problem", C.POINTER(Problem), [_VP]),
    ("s35_part_initial_pressure", c_double_p, [_VP]),
    ("s35_disc_partition_cells", C.c_int, [_VP, C.c_int32, c_int32_p]),
    ("s35_mesh_convert", C.c_int, [C.c_char_p, C.c_int32, C.c_char_p]),
    ("s35_mesh_h5_save", C.c_int, [C.c_char_p, C.c_char_p, c_double_p, C.c_int64]),
    ("s35_mesh_h5_load", C.c_int, [C.c_char_p, C.c_char_p, c_double_p, C.c_int64, C.POINTER(C.c_int64)]),
    ("s35_sim_create", C.c_int, [C.POINTER(Params), C.c_char_p, C.c_int32, C.POINTER(_VP)]),
    ("s35_sim_destroy", None, [_VP]),
    ("s35_sim_run", C.c_int, [_VP, C.c_int32, C.c_int32]),
    ("s35_sim_ctx", _VP, [_VP]),
    ("s35_sim_disc", _VP, [_VP]),
    ("s35_sim_n_steps", C.c_int32, [_VP]),
    ("s35_sim_drag_lift", C.c_int, [_VP, c_double_p]),
    ("s35_disc_drag_lift", C.c_int, [_VP, C.c_double, c_double_p, c_double_p, c_double_p]),
]
DECLARED_SYMBOLS = [s[0] for s in _SIGNATURES]
# what libstep35a_host.so (set-up, partition, drag / lift: no device code, links no CUDA library) exports
HOST_ONLY_SYMBOLS = ["s35_last_error"] + [s[0] for s in _SIGNATURES
                                          if s[0].startswith(("s35_params_", "s35_disc_", "s35_part_", "s35_mesh_"))]
_host_lib = None


def lib():
    """The loaded shared library.  Raises if it has not been built (run __graft_entry__.build())."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError("libstep35a_b200.so is not built (%s); run `python -c 'import __graft_entry__ as g; "
                              "g.build()'` -- there is no Python or CPU fallback" % LIB_PATH)
        handle = C.CDLL(LIB_PATH)
        for name, restype, argtypes in _SIGNATURES:
            fn = getattr(handle, name)
            fn.restype = restype
            fn.argtypes = argtypes
        _lib = handle
    return _lib


def host_lib():
    """The host-only library: discretisation, partition and drag / lift.  Everything that only sets a problem up goes
    through it, so that a process that never touches a GPU (the CPU reference arm of bench.py, the oracle tests) maps
    no CUDA library."""
    global _host_lib
    if _host_lib is None:
        if not os.path.exists(HOST_LIB_PATH):
            raise ImportError("libstep35a_host.so is not built (%s); run `python -c 'import __graft_entry__ as g; "
                              "g.build()'`" % HOST_LIB_PATH)
        handle = C.CDLL(HOST_LIB_PATH)
        for name, restype, argtypes in _SIGNATURES:
            if name in HOST_ONLY_SYMBOLS:
                fn = getattr(handle, name)
                fn.restype = restype
                fn.argtypes = argtypes
        _host_lib = handle
    return _host_lib


def check(status):
    if status != 0:
        raise S35Error(status, lib().s35_last_error().decode())


def check_host(status):
    if status != 0:
        raise S35Error(status, host_lib().s35_last_error().decode())


def dptr(a):
    return a.ctypes.data_as(c_double_p)


def iptr(a):
    return a.ctypes.data_as(c_int32_p)


def f64(a):
    return np.ascontiguousarray(a, dtype=np.float64)


def i32(a):
    return np.ascontiguousarray(a, dtype=np.int32)


def device_count():
    return lib().s35_device_count()


def launch_count():
    return lib().s35_launch_count()


def default_params(**overrides):
    p = Params()
    host_lib().s35_params_defaults(C.byref(p))
    for k, v in overrides.items():
        setattr(p, k, v)
    return p


def read_params(prm_file, **overrides):
    p = Params()
    check_host(host_lib().s35_params_read(prm_file.encode(), C.byref(p)))
    for k, v in overrides.items():
        setattr(p, k, v)
    return p


def parse_params(text):
    p = Params()
    check_host(host_lib().s35_params_parse(text.encode(), C.byref(p)))
    return p


def _view(ptr, shape, dtype):
    n = int(np.prod(shape))
    if n == 0:
        return np.zeros(shape, dtype=dtype)
    return np.ctypeslib.as_array(ptr, shape=(n,)).reshape(shape)


class Disc:
    """The discretisation: mesh, DoFs, constraints, patterns and one-off operators (host arrays)."""

    def __init__(self, mesh_file, dim, params, _handle=None, _owner=None):
        self._owner = _owner
        if _handle is not None:
            self.handle = _handle
            self._owns = False
        else:
            h = _VP()
            check_host(host_lib().s35_disc_create(mesh_file.encode(), dim, C.byref(params), C.byref(h)))
            self.handle = h
            self._owns = True
        self.info = DiscInfo()
        check_host(host_lib().s35_disc_get_info(self.handle, C.byref(self.info)))
        self.problem = host_lib().s35_disc_problem(self.handle)  # POINTER(Problem)
        p = self.problem.contents
        self.dim, self.n_cells, self.nv, self.np_ = p.dim, p.n_cells, p.n_vel, p.n_pres
        dim = self.dim
        self.vel_support = _view(host_lib().s35_disc_support_points(self.handle, 0), (self.nv, dim), np.float64)
        self.pres_support = _view(host_lib().s35_disc_support_points(self.handle, 1), (self.np_, dim), np.float64)
        self.initial_pressure = _view(host_lib().s35_disc_initial_pressure(self.handle), (self.np_,), np.float64)

    def close(self):
        if self._owns and self.handle is not None:
            host_lib().s35_disc_destroy(self.handle)
        self.handle = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def csr(self, which):
        """(rowptr, col, values-or-list) of one of 'velc', 'velm', 'grad', 'plap', 'pmass' as numpy views."""
        p = self.problem.contents
        n_rows = {"velc": self.nv, "velm": self.nv, "grad": self.nv, "plap": self.np_, "pmass": self.np_}[which]
        rowptr = _view(getattr(p, which + "_rowptr"), (n_rows + 1,), np.int32)
        nnz = int(rowptr[-1])
        col = _view(getattr(p, which + "_col"), (nnz,), np.int32)
        if which == "velc":
            val = None
        elif which == "grad":
            val = [_view(p.grad_val[d], (nnz,), np.float64) for d in range(self.dim)]
        else:
            val = _view(getattr(p, which + "_val"), (nnz,), np.float64)
        return rowptr, col, val

    def constraints(self):
        p = self.problem.contents
        wall = _view(p.wall_dofs, (p.n_wall,), np.int32)
        flow = [(_view(p.flow_dofs[d], (p.n_flow[d],), np.int32), _view(p.flow_vals[d], (p.n_flow[d],), np.float64))
                for d in range(self.dim)]
        proj = _view(p.proj_dofs, (p.n_proj,), np.int32)
        return wall, flow, proj

    def cell_tables(self):
        p = self.problem.contents
        vpc, dpc = 1 << self.dim, 3 ** self.dim
        verts = _view(p.cell_vertices, (self.n_cells, vpc, self.dim), np.float64)
        dofs = _view(p.vel_cell_dofs, (self.n_cells, dpc), np.int32)
        color = _view(p.cell_color, (self.n_cells,), np.int32)
        pdofs = _view(host_lib().s35_disc_pres_cell_dofs(self.handle), (self.n_cells, vpc), np.int32)
        return verts, dofs, color, pdofs

    def mesh(self):
        v, c, b, ids, nb = c_double_p(), c_int32_p(), c_int32_p(), c_int32_p(), C.c_int64()
        check_host(host_lib().s35_disc_mesh(self.handle, C.byref(v), C.byref(c), C.byref(b), C.byref(ids), C.byref(nb)))
        vpc, vpf = 1 << self.dim, 1 << (self.dim - 1)
        return (_view(v, (self.info.n_vertices, self.dim), np.float64), _view(c, (self.n_cells, vpc), np.int32),
                _view(b, (nb.value, vpf), np.int32), _view(ids, (nb.value,), np.int32))

    def drag_lift(self, reynolds, u, p):
        out = np.zeros(2)
        u = f64(u)
        p = f64(p)
        check_host(host_lib().s35_disc_drag_lift(self.handle, reynolds, dptr(u), dptr(p), dptr(out)))
        return out


class Part:
    """One rank's part of a mesh-partitioned problem (s35_part): local problem, ownership and halo plans."""

    def __init__(self, disc, n_ranks, rank):
        h = _VP()
        check_host(host_lib().s35_part_create(disc.handle, n_ranks, rank, C.byref(h)))
        self.handle = h
        self.disc = disc
        self.problem = host_lib().s35_part_problem(h)
        p = self.problem.contents
        d = p.dist.contents
        self.rank, self.n_ranks, self.dim = d.rank, d.n_ranks, p.dim
        self.nv, self.np_ = p.n_vel, p.n_pres
        self.nv_owned, self.np_owned = d.n_vel_owned, d.n_pres_owned
        self.vel_global = _view(d.vel_global_index, (self.nv,), np.int32)
        self.pres_global = _view(d.pres_global_index, (self.np_,), np.int32)
        self.initial_pressure = _view(host_lib().s35_part_initial_pressure(h), (self.np_,), np.float64)

    def halo(self, which):
        d = self.problem.contents.dist.contents
        hl = d.vel_halo if which == "vel" else d.pres_halo
        nb = _view(hl.neighbor_rank, (hl.n_neighbors,), np.int32)
        sp_ = _view(hl.send_ptr, (hl.n_neighbors + 1,), np.int32)
        rp = _view(hl.recv_ptr, (hl.n_neighbors + 1,), np.int32)
        si = _view(hl.send_idx, (int(sp_[-1]) if hl.n_neighbors else 0,), np.int32)
        return nb, sp_, si, rp

    def csr(self, which):
        """local rows (owned DoFs) of 'velc', 'velm', 'grad', 'gradT', 'plap', 'pmass': (rowptr, col, values)"""
        p = self.problem.contents
        d = p.dist.contents
        if which == "gradT":
            rowptr = _view(d.gradT_rowptr, (self.np_owned + 1,), np.int32)
            nnz = int(rowptr[-1])
            return rowptr, _view(d.gradT_col, (nnz,), np.int32), [_view(d.gradT_val[k], (nnz,), np.float64) for k in range(self.dim)]
        n_rows = self.nv_owned if which in ("velc", "velm", "grad") else self.np_owned
        rowptr = _view(getattr(p, which + "_rowptr"), (n_rows + 1,), np.int32)
        nnz = int(rowptr[-1])
        col = _view(getattr(p, which + "_col"), (nnz,), np.int32)
        if which == "velc":
            val = None
        elif which == "grad":
            val = [_view(p.grad_val[k], (nnz,), np.float64) for k in range(self.dim)]
        else:
            val = _view(getattr(p, which + "_val"), (nnz,), np.float64)
        return rowptr, col, val

    def close(self):
        if self.handle is not None:
            host_lib().s35_part_destroy(self.handle)
        self.handle = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def comm_unique_id():
    buf = C.create_string_buffer(128)
    check(lib().s35_comm_unique_id(buf))
    return buf.raw


def comm_init(rank, n_ranks, unique_id):
    check(lib().s35_comm_init(rank, n_ranks, unique_id))


def comm_finalize():
    lib().s35_comm_finalize()


def comm_thread_group(n_ranks):
    """ranks as threads of this process sharing one GPU (single-GPU tests of the partitioned path)"""
    check(lib().s35_comm_thread_group(n_ranks))


def comm_thread_bind(rank):
    check(lib().s35_comm_thread_bind(rank))


def with_ilu_blocks(problem, k):
    """Copy of a problem description with the gated block-ILU alternative switched on (s35_problem.ilu_blocks)."""
    src = problem.contents if hasattr(problem, "contents") else problem
    p = Problem.from_buffer_copy(src)
    p.ilu_blocks = int(k)
    p._keepalive = problem
    return p


class Ctx:
    """Device context of the time loop (s35_ctx)."""

    def __init__(self, problem, _handle=None, _owner=None):
        self._owner = _owner
        self._owns = _handle is None
        if _handle is None:
            h = _VP()
            check(lib().s35_ctx_create(problem, C.byref(h)))
            self.handle = h
        else:
            self.handle = _handle
        p = problem.contents if hasattr(problem, "contents") else problem
        self.dim, self.nv, self.np_ = p.dim, p.n_vel, p.n_pres
        self.nnz_vel = None

    def close(self):
        if self._owns and self.handle is not None:
            lib().s35_ctx_destroy(self.handle)
        self.handle = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _len(self, field):
        return self.nv if field in (U_N, U_N_MINUS_1, U_STAR, FORCE) else self.np_

    def set_field(self, field, values, component=0):
        values = f64(values)
        assert values.shape == (self._len(field),)
        check(lib().s35_ctx_set_field(self.handle, field, component, dptr(values)))

    def get_field(self, field, component=0):
        out = np.empty(self._len(field))
        check(lib().s35_ctx_get_field(self.handle, field, component, dptr(out)))
        return out

    def get_operator_values(self, op, nnz, component=0):
        out = np.empty(nnz)
        check(lib().s35_ctx_get_operator_values(self.handle, op, component, dptr(out)))
        return out

    def interpolate_velocity(self):
        check(lib().s35_interpolate_velocity(self.handle))

    def assemble_advection_term(self):
        check(lib().s35_assemble_advection_term(self.handle))

    def diffusion_step(self, reinit_prec):
        check(lib().s35_diffusion_step(self.handle, 1 if reinit_prec else 0))

    def projection_step(self):
        check(lib().s35_projection_step(self.handle))

    def update_pressure(self):
        check(lib().s35_update_pressure(self.handle))

    def advance(self, first_n, n_steps, start_n=2, start_offset_n=2):
        ms = C.c_double()
        check(lib().s35_advance(self.handle, first_n, n_steps, start_n, start_offset_n, C.byref(ms)))
        return ms.value

    def synchronize(self):
        check(lib().s35_synchronize(self.handle))

    def solver_stats(self, component):
        st = SolverStats()
        check(lib().s35_ctx_solver_stats(self.handle, component, C.byref(st)))
        return st

    def section_ms(self, name):
        ms, calls = C.c_double(), C.c_int64()
        check(lib().s35_ctx_section_ms(self.handle, name.encode(), C.byref(ms), C.byref(calls)))
        return ms.value, calls.value

    def bench_kernel(self, kernel, warmup=3, reps=10, flush_l2=True):
        ms, nbytes, launches = C.c_double(), C.c_double(), C.c_int64()
        check(lib().s35_ctx_bench_kernel(self.handle, kernel, warmup, reps, 1 if flush_l2 else 0, C.byref(ms),
                                         C.byref(nbytes), C.byref(launches)))
        return ms.value, nbytes.value, launches.value


class Matrix:
    """Device-resident CSR matrix (s35_matrix)."""

    def __init__(self, n_rows, n_cols, rowptr, col, val=None):
        self.n_rows, self.n_cols = n_rows, n_cols
        rowptr, col = i32(rowptr), i32(col)
        self.nnz = int(rowptr[-1])
        h = _VP()
        vp = dptr(f64(val)) if val is not None else None
        check(lib().s35_matrix_create(n_rows, n_cols, iptr(rowptr), iptr(col), vp, C.byref(h)))
        self.handle = h

    def close(self):
        if self.handle is not None:
            lib().s35_matrix_destroy(self.handle)
        self.handle = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def set_values(self, val):
        val = f64(val)
        check(lib().s35_matrix_set_values(self.handle, dptr(val)))

    def get_values(self):
        out = np.empty(self.nnz)
        check(lib().s35_matrix_get_values(self.handle, dptr(out)))
        return out

    def spmv(self, x, y=None, transpose=False):
        x = f64(x)
        add = y is not None
        out = f64(y).copy() if add else np.empty(self.n_cols if transpose else self.n_rows)
        check(lib().s35_spmv(self.handle, dptr(x), dptr(out), 1 if transpose else 0, 1 if add else 0))
        return out

    def condense(self, rhs, dofs, values=None):
        rhs = f64(rhs).copy()
        dofs = i32(dofs)
        vp = dptr(f64(values)) if values is not None else None
        check(lib().s35_condense(self.handle, dptr(rhs), len(dofs), iptr(dofs), vp))
        return rhs


class Ilu:
    """ILU(0) preconditioner on the pattern of a Matrix (s35_ilu)."""

    def __init__(self, matrix):
        self.matrix = matrix
        h = _VP()
        check(lib().s35_ilu_create(matrix.handle, C.byref(h)))
        self.handle = h

    def close(self):
        if self.handle is not None:
            lib().s35_ilu_destroy(self.handle)
        self.handle = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def factor(self, matrix=None, strengthen_diagonal=0.0):
        check(lib().s35_ilu_factor(self.handle, (matrix or self.matrix).handle, strengthen_diagonal))

    def apply(self, src):
        src = f64(src)
        dst = np.empty_like(src)
        check(lib().s35_ilu_apply(self.handle, dptr(src), dptr(dst)))
        return dst

    def values(self):
        out = np.empty(self.matrix.nnz)
        check(lib().s35_ilu_get_values(self.handle, dptr(out)))
        return out

    def levels(self):
        f, b = C.c_int32(), C.c_int32()
        check(lib().s35_ilu_levels(self.handle, C.byref(f), C.byref(b)))
        return f.value, b.value


def gmres(matrix, ilu, x0, b, tol, max_iterations=1000, krylov_size=30):
    x = f64(x0).copy()
    b = f64(b)
    st = SolverStats()
    status = lib().s35_gmres(matrix.handle, ilu.handle, dptr(x), dptr(b), tol, max_iterations, krylov_size, C.byref(st))
    if status not in (0, 3):
        check(status)
    return x, st, status == 0


def cg(matrix, ilu, x0, b, tol, max_iterations=1000):
    x = f64(x0).copy()
    b = f64(b)
    st = SolverStats()
    status = lib().s35_cg(matrix.handle, ilu.handle if ilu is not None else None, dptr(x), dptr(b), tol, max_iterations,
                          C.byref(st))
    if status not in (0, 3):
        check(status)
    return x, st, status == 0


def vec_lincomb(a, x, b=0.0, y=None):
    x = f64(x)
    z = np.empty_like(x)
    yp = dptr(f64(y)) if y is not None else None
    check(lib().s35_vec_lincomb(x.size, a, dptr(x), b, yp, dptr(z)))
    return z


def vec_dot(x, y):
    x, y = f64(x), f64(y)
    r = C.c_double()
    check(lib().s35_vec_dot(x.size, dptr(x), dptr(y), C.byref(r)))
    return r.value


def vec_add_and_dot(a, x, y, w):
    """Returns (y + a x, (y + a x) . w)."""
    alias = w is y
    x = f64(x)
    y = f64(y).copy()
    w = y if alias else f64(w)
    r = C.c_double()
    check(lib().s35_vec_add_and_dot(x.size, a, dptr(x), dptr(y), dptr(w), C.byref(r)))
    return y, r.value


def h5_save(file_name, dataset_name, values):
    """H5::save of the reference's h5.h: one 1-D double dataset, added to the file if it exists"""
    v = f64(values)
    check_host(host_lib().s35_mesh_h5_save(file_name.encode(), dataset_name.encode(), dptr(v), v.size))


def h5_load(file_name, dataset_name):
    n = C.c_int64()
    check_host(host_lib().s35_mesh_h5_load(file_name.encode(), dataset_name.encode(), None, 0, C.byref(n)))
    out = np.zeros(n.value)
    check_host(host_lib().s35_mesh_h5_load(file_name.encode(), dataset_name.encode(), dptr(out), n.value, C.byref(n)))
    return out


def mesh_convert(in_file, dim, out_file):
    check_host(host_lib().s35_mesh_convert(in_file.encode(), dim, out_file.encode()))


class Simulation:
    """NavierStokesProjection<dim> (host driver + device context)."""

    def __init__(self, params, mesh_file, dim):
        h = _VP()
        check(lib().s35_sim_create(C.byref(params), mesh_file.encode() if mesh_file else None, dim, C.byref(h)))
        self.handle = h
        self.disc = Disc(None, dim, None, _handle=_VP(lib().s35_sim_disc(h)), _owner=self)
        self.ctx = Ctx(self.disc.problem, _handle=_VP(lib().s35_sim_ctx(h)), _owner=self)
        self.n_steps = lib().s35_sim_n_steps(h)

    def run(self, verbose=False, output_interval=999999):
        check(lib().s35_sim_run(self.handle, 1 if verbose else 0, output_interval))

    def drag_lift(self):
        out = np.zeros(2)
        check(lib().s35_sim_drag_lift(self.handle, dptr(out)))
        return out

    def close(self):
        if self.handle is not None:
            self.ctx.handle = None
            self.disc.handle = None
            lib().s35_sim_destroy(self.handle)
        self.handle = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
