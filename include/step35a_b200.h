/*
 * step35a_b200 -- C ABI of the B200 (sm_100a) device path for the per-time-step
 * hot path of step-35a's NavierStokesProjection.
 *
 * The reference has no FFI or plugin layer: its seam is the set of member
 * functions of NavierStokesProjection<dim> (step-35a.cc:529-543, 626, 681)
 * working on deal.II objects held by value (step-35a.cc:474-520).  Every entry
 * point below names the reference call it stands in for.  All pointers are
 * plain host pointers unless a comment says "device"; sizes are element
 * counts; indices are 32-bit like deal.II's types::global_dof_index.
 *
 * CSR convention: `rowptr` has n_rows+1 entries, columns are ASCENDING within
 * each row, and square matrices store their diagonal entry (deal.II keeps the
 * diagonal first in each row; INTEGRATION.md shows the two-line conversion).
 *
 * Error convention: every function returns an s35_status.  The reference
 * throws C++ exceptions (AssertThrow at step-35a.cc:246,731; deal.II's
 * SolverControl::NoConvergence) that main() catches (step-35a.cc:1730-1752);
 * the C++ host wrapper turns non-zero codes back into std::runtime_error.
 * s35_last_error() returns a message for the calling thread's last failure.
 *
 * There is no CPU fallback: if no CUDA device is usable every compute entry
 * point returns S35_ERR_CUDA.
 */
#ifndef STEP35A_B200_H
#define STEP35A_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef enum s35_status
{
  S35_OK = 0,
  S35_ERR_INVALID = 1,        /* bad argument (null pointer, size mismatch, unsorted row ...) */
  S35_ERR_CUDA = 2,           /* CUDA runtime error, or no device */
  S35_ERR_NO_CONVERGENCE = 3, /* SolverControl::NoConvergence */
  S35_ERR_ZERO_PIVOT = 4,     /* SparseILU: zero pivot */
  S35_ERR_UNSUPPORTED = 5,    /* parameter combination outside this build (e.g. off_diagonals > 0 on a partition) */
  S35_ERR_IO = 6
} s35_status;

const char *s35_last_error(void);
/* library version string, and the number of visible CUDA devices (0 on a CPU-only machine) */
const char *s35_version(void);
int s35_device_count(void);
/* select the CUDA device for subsequently created objects of the calling thread (default 0) */
s35_status s35_set_device(int device);

/* iteration record of the last solve (replaces the deallog output enabled at step-35a.cc:1724) */
typedef struct s35_solver_stats
{
  int32_t iterations;
  double initial_residual;
  double final_residual;
  double tolerance;
} s35_solver_stats;

/* ------------------------------------------------------------------------
 * Device-resident sparse matrix (deal.II SparseMatrix<double> + SparsityPattern,
 * step-35a.cc:474-500).
 * ---------------------------------------------------------------------- */
typedef struct s35_matrix s35_matrix;

/* uploads the pattern and (if val != NULL) the values; otherwise values are zero */
s35_status s35_matrix_create(int32_t n_rows, int32_t n_cols, const int32_t *rowptr, const int32_t *col,
                             const double *val, s35_matrix **out);
void s35_matrix_destroy(s35_matrix *m);
s35_status s35_matrix_set_values(s35_matrix *m, const double *val);
s35_status s35_matrix_get_values(const s35_matrix *m, double *val);
int64_t s35_matrix_nnz(const s35_matrix *m);

/* SparseMatrix::vmult / vmult_add / Tvmult / Tvmult_add (step-35a.cc:1287,1289,1462,1464):
 * y = [y +] A x   or   y = [y +] A^T x.  x and y are host vectors; the copies are part of the call. */
s35_status s35_spmv(const s35_matrix *A, const double *x, double *y, int transpose, int add);

/* ConstraintMatrix::condense(SparseMatrix&, Vector&) for constraints without dependencies
 * (x_c = g_c), step-35a.cc:1311 and Appendix A.4 of SURVEY.md: rhs lifted, constrained rows and
 * columns cleared, the mean |diagonal| written on constrained diagonals. */
s35_status s35_condense(s35_matrix *A, double *rhs, int32_t n_constrained, const int32_t *dofs, const double *values);

/* ------------------------------------------------------------------------
 * ILU(0) preconditioner (deal.II SparseILU<double>, step-35a.cc:1070-1074,1312-1319).
 * ---------------------------------------------------------------------- */
typedef struct s35_ilu s35_ilu;

/* symbolic set-up on the pattern of A (use_previous_sparsity=true: done once).  The preconditioner keeps a
 * reference to A's pattern and works on A's stream: destroy it BEFORE A (SparseILU has the same rule for its
 * SparsityPattern). */
s35_status s35_ilu_create(const s35_matrix *A, s35_ilu **out);
void s35_ilu_destroy(s35_ilu *p);
/* SparseILU::initialize(A, AdditionalData(strengthen_diagonal, 0)): numeric factorisation,
 * inverse pivots stored on the diagonal */
s35_status s35_ilu_factor(s35_ilu *p, const s35_matrix *A, double strengthen_diagonal);
/* SparseILU::vmult(dst, src): dst = (LU)^-1 src, host vectors */
s35_status s35_ilu_apply(s35_ilu *p, const double *src, double *dst);
/* factor values in the layout of A's pattern (unit-L strictly lower, U upper, 1/pivot on the diagonal) */
s35_status s35_ilu_get_values(const s35_ilu *p, double *val);
/* number of dependency levels of the forward / backward substitution */
s35_status s35_ilu_levels(const s35_ilu *p, int32_t *forward_levels, int32_t *backward_levels);

/* ------------------------------------------------------------------------
 * Krylov solvers with the ILU(0) preconditioner, host vectors in and out.
 * ---------------------------------------------------------------------- */
/* SolverGMRES<>::solve with AdditionalData(max_n_tmp_vectors = krylov_size), left preconditioning,
 * stopping on the preconditioned residual (step-35a.cc:1334-1337).  x: initial guess in, solution out. */
s35_status s35_gmres(const s35_matrix *A, s35_ilu *prec, double *x, const double *b, double tolerance,
                     int32_t max_iterations, int32_t krylov_size, s35_solver_stats *stats);
/* SolverCG<>::solve (step-35a.cc:1496-1499).  prec may be NULL (PreconditionIdentity). */
s35_status s35_cg(const s35_matrix *A, s35_ilu *prec, double *x, const double *b, double tolerance,
                  int32_t max_iterations, s35_solver_stats *stats);

/* ------------------------------------------------------------------------
 * Fused vector kernels (deal.II Vector<double>::equ/add/sadd/operator*, add_and_dot, l2_norm;
 * step-35a.cc:1251-1252,1272-1273,1285-1286,1506,1539-1540 and the Krylov internals).
 * Host vectors; used by the parity tests and the bandwidth benchmark.
 * ---------------------------------------------------------------------- */
/* z = a x + b y (y may be NULL when b == 0) */
s35_status s35_vec_lincomb(int64_t n, double a, const double *x, double b, const double *y, double *z);
s35_status s35_vec_dot(int64_t n, const double *x, const double *y, double *result);
/* y += a x; result = y . w   (Vector::add_and_dot) */
s35_status s35_vec_add_and_dot(int64_t n, double a, const double *x, double *y, const double *w, double *result);

/* result = x . y without allocating anything.  This entry point has no counterpart in the reference; it exists
 * for callers that cannot call cudaMalloc: programs that reserve all of their device memory up front and run
 * against that fixed budget, callers confined to a memory pool of their own, and code that must not trigger
 * the device-wide stall an allocation causes inside its solver loop.
 *   - x_device, y_device (n doubles each) and result_device (one double) are DEVICE pointers owned by the
 *     caller; `stream` is a cudaStream_t (NULL: the default stream).  The call only enqueues work: read
 *     *result_device after synchronising that stream.
 *   - The reduction scratch is a fixed array inside the library image (8.25 KB on every GPU that loads the
 *     library, for the life of the process), so n is limited to 2 097 152 elements; larger n returns
 *     S35_ERR_INVALID with a message naming the limit.  No cudaMalloc, cudaFree or allocating library call
 *     is made.  The result is bit-identical to s35_vec_dot's for the same vectors.
 *   - The first call on each device looks up the device address of that array; make that first call before
 *     starting any stream capture.
 *   - All calls share the one array: two calls that can run on the GPU at the same time (different streams
 *     without an ordering between them) race and give wrong results.  Calls on one stream are safe.
 *   - The looked-up address is remembered per device for the life of the process: after a cudaDeviceReset() it is
 *     stale, and this entry point must not be used again in that process.
 *   - n == 0 stores 0.0 in *result_device and reads neither x nor y (they may be NULL). */
s35_status s35_vec_dot_device_noalloc(int64_t n, const double *x_device, const double *y_device, double *result_device,
                                      void *stream);

/* ------------------------------------------------------------------------
 * The time-step context: everything NavierStokesProjection<dim> keeps between steps.
 * ---------------------------------------------------------------------- */
typedef struct s35_problem
{
  int32_t dim;                /* 2 or 3 */
  int32_t n_cells;
  int32_t n_vel;              /* velocity DoFs per component (Q2) */
  int32_t n_pres;             /* pressure DoFs (Q1) */

  /* geometry and cell -> DoF tables for assemble_advection_term (step-35a.cc:1366-1374).
   * Vertices and local DoFs in lexicographic order (x fastest). */
  const double *cell_vertices;   /* [n_cells][2^dim][dim] */
  const int32_t *vel_cell_dofs;  /* [n_cells][3^dim] */
  const int32_t *cell_color;     /* [n_cells], cells of one colour share no DoF; NULL: computed by the library */

  /* constrained_velocity_sparsity_pattern: vel_it_matrix[d], vel_Advection (step-35a.cc:972-984) */
  const int32_t *velc_rowptr, *velc_col;
  /* vel_Mass on unconstrained_velocity_sparsity_pattern (step-35a.cc:965-970,986-989) */
  const int32_t *velm_rowptr, *velm_col;
  const double *velm_val;
  /* pres_Diff[d], n_vel x n_pres, one shared pattern (step-35a.cc:1084-1091,1112) */
  const int32_t *grad_rowptr, *grad_col;
  const double *grad_val[3];
  /* pres_Laplace on projection_sparsity_pattern, pres_Mass on pressure_sparsity_pattern (step-35a.cc:996-1058) */
  const int32_t *plap_rowptr, *plap_col;
  const double *plap_val;
  const int32_t *pmass_rowptr, *pmass_col;
  const double *pmass_val;

  /* constraints without dependencies: lists of (dof, value), dofs ascending (step-35a.cc:903-958) */
  int32_t n_wall;             /* velocity_wall_constraints (homogeneous) */
  const int32_t *wall_dofs;
  int32_t n_flow[3];          /* velocity_flow_constraints[d] */
  const int32_t *flow_dofs[3];
  const double *flow_vals[3];
  int32_t n_proj;             /* projection_constraints (homogeneous) */
  const int32_t *proj_dofs;

  /* RunTimeParameters::Data_Storage fields used on the path (step-35a.cc:135-151) */
  double dt, reynolds;
  double vel_eps, vel_diag_strength;
  int32_t vel_max_iterations, vel_krylov_size, vel_off_diagonals, vel_update_prec;
  /* vel_off_diagonals = k > 0 (SparseILU::AdditionalData::extra_off_diagonals, step-35a.cc:1070-1074, 1312-1319): both
   * ILU factors live on the matrix pattern plus the entries (r, r - k .. r + k).  One GPU only; not with ilu_blocks.
   * s35_ctx_get_operator_values(S35_VEL_ILU) then returns the factor's entries at the matrix's own positions. */
  int32_t rotational;         /* METHOD_ROTATIONAL = 1, METHOD_STANDARD = 0 */

  /* GATED ALTERNATIVE, not the reference's preconditioner.  0 or 1: ILU(0) of the whole matrix, exactly what
   * SparseILU computes (the default).  k > 1: the rows are cut into k contiguous index ranges (slabs of the
   * Cuthill-McKee numbering) and couplings between different ranges are dropped from the FACTORS only; the
   * matrices are untouched.  The k blocks are independent, so the dependency chains of the substitutions get
   * about k times shorter and k times wider -- the single-GPU analogue of what the mesh partition does across
   * GPUs.  Iteration counts rise; fields agree with the default to the solver tolerance. */
  int32_t ilu_blocks;

  /* NULL: the whole problem on one GPU.  Otherwise this is ONE rank's part of a mesh-partitioned problem
   * (new capability, the reference is single-process): n_vel / n_pres count the rank's local DoFs, owned
   * ones first, then ghosts grouped by owning rank; every matrix has rows for the owned DoFs only and
   * columns in local numbering; n_cells counts the cells that touch an owned velocity DoF; constraint lists
   * hold local indices of owned and ghost DoFs. */
  const struct s35_dist *dist;
} s35_problem;

/* ghost-value exchange of one DoF kind: neighbour q receives the owned entries send_idx[send_ptr[q] ..
 * send_ptr[q+1]) and delivers the ghost entries [n_owned + recv_ptr[q], n_owned + recv_ptr[q+1]) */
typedef struct s35_halo
{
  int32_t n_neighbors;
  const int32_t *neighbor_rank;
  const int32_t *send_ptr, *send_idx;
  const int32_t *recv_ptr;
} s35_halo;

typedef struct s35_dist
{
  int32_t rank, n_ranks;
  int32_t n_vel_owned, n_pres_owned;
  int64_t n_vel_global, n_pres_global;
  s35_halo vel_halo, pres_halo;
  const int32_t *vel_global_index;  /* [n_vel]  local -> global DoF index */
  const int32_t *pres_global_index; /* [n_pres] */
  /* rows of pres_Diff[d]^T for the owned pressure DoFs, columns = local velocity DoFs (owned and ghost): a rank
   * cannot form them from its own rows of pres_Diff[d], which only cover its owned velocity DoFs */
  const int32_t *gradT_rowptr, *gradT_col;
  const double *gradT_val[3];
} s35_dist;

typedef struct s35_ctx s35_ctx;

/* Joins the calling process (one per GPU) to the communicator the distributed contexts of this process will
 * use: NCCL over NVLink, `unique_id` = the 128 bytes of an ncclUniqueId created by rank 0 (s35_comm_unique_id)
 * and handed to the other ranks by the launcher (bench.py broadcasts it with torch.distributed). */
s35_status s35_comm_unique_id(unsigned char id[128]);
s35_status s35_comm_init(int32_t rank, int32_t n_ranks, const unsigned char unique_id[128]);
void s35_comm_finalize(void);
/* The same partitioned path with the ranks as THREADS of one process that share one GPU (NCCL refuses two
 * ranks on one device): s35_comm_thread_group(n) once, then every rank's thread calls s35_comm_thread_bind(rank)
 * before creating its context, and all rank threads make the same sequence of context calls.  For testing the
 * multi-rank code on a single GPU; not a performance path. */
s35_status s35_comm_thread_group(int32_t n_ranks);
s35_status s35_comm_thread_bind(int32_t rank);

/* copies everything it needs to the device; the caller keeps ownership of the host arrays */
s35_status s35_ctx_create(const s35_problem *problem, s35_ctx **out);
void s35_ctx_destroy(s35_ctx *ctx);

/* state vectors: the checkpoint set of step-35a.cc:865-875 plus the work vectors the tests look at */
typedef enum s35_field
{
  S35_U_N = 0,         /* component = 0..dim-1 */
  S35_U_N_MINUS_1 = 1, /* component = 0..dim-1 */
  S35_U_STAR = 2,      /* component = 0..dim-1 */
  S35_FORCE = 3,       /* component = 0..dim-1 */
  S35_PRES_N = 4,
  S35_PRES_N_MINUS_1 = 5,
  S35_PHI_N = 6,
  S35_PHI_N_MINUS_1 = 7,
  S35_PRES_TMP = 8
} s35_field;
s35_status s35_ctx_set_field(s35_ctx *ctx, s35_field field, int32_t component, const double *values);
s35_status s35_ctx_get_field(s35_ctx *ctx, s35_field field, int32_t component, double *values);

/* matrices the tests look at */
typedef enum s35_operator
{
  S35_VEL_ADVECTION = 0, /* vel_Advection */
  S35_VEL_IT_MATRIX = 1, /* vel_it_matrix[component] after condense */
  S35_VEL_ILU = 2        /* prec_velocity[component] factor values */
} s35_operator;
s35_status s35_ctx_get_operator_values(s35_ctx *ctx, s35_operator op, int32_t component, double *values);

/* one call per reference member function */
s35_status s35_interpolate_velocity(s35_ctx *ctx);                  /* step-35a.cc:1244-1254 */
s35_status s35_assemble_advection_term(s35_ctx *ctx);               /* step-35a.cc:1346-1453 */
s35_status s35_diffusion_step(s35_ctx *ctx, int reinit_prec);       /* step-35a.cc:1268-1338 */
s35_status s35_projection_step(s35_ctx *ctx);                       /* step-35a.cc:1456-1508, #else branch */
s35_status s35_update_pressure(s35_ctx *ctx);                       /* step-35a.cc:1511-1545, #else branch */
/* the body of run()'s loop for steps n = first_n .. first_n + n_steps - 1 (step-35a.cc:1190-1239):
 * start_n is `step_start_n`, start_offset_n is `step_start_offset_n` of the reference.
 * Device time of the whole call in milliseconds is returned in elapsed_ms if not NULL. */
s35_status s35_advance(s35_ctx *ctx, int32_t first_n, int32_t n_steps, int32_t start_n, int32_t start_offset_n,
                       double *elapsed_ms);
s35_status s35_synchronize(s35_ctx *ctx);

/* iteration counts of the last diffusion_step (component 0..dim-1) and projection_step (component = -1),
 * and of the pressure-mass solve in update_pressure (component = -2) */
s35_status s35_ctx_solver_stats(s35_ctx *ctx, int32_t component, s35_solver_stats *stats);

/* accumulated device time per reference timer section (TimerOutput names, step-35a.cc:1248-1515);
 * section = one of "interpolate_velocity", "setup_diffusion_rhs", "solve_diffusion_component",
 * "assemble_advection_term", "projection_step_setup", "projection_step_solve", "update_pressure" */
s35_status s35_ctx_section_ms(s35_ctx *ctx, const char *section, double *ms, int64_t *calls);

/* ------------------------------------------------------------------------
 * Kernel benchmarks on device-resident data (inputs already in HBM): average device time of one
 * launch in milliseconds over `reps` launches after `warmup` ones, CUDA events on the launch stream.
 * `flush_l2` != 0 writes a buffer larger than L2 between launches (outside the events).
 * ---------------------------------------------------------------------- */
typedef enum s35_bench_kernel
{
  S35_BENCH_SPMV_VEL = 0,      /* y = vel_it_matrix[0] x */
  S35_BENCH_SPMV_PRES = 1,     /* y = pres_Laplace x */
  S35_BENCH_SPMV_GRAD = 2,     /* y = pres_Diff[0] x */
  S35_BENCH_SPMV_GRAD_T = 3,   /* y = sum_d pres_Diff[d]^T x_d */
  S35_BENCH_ASSEMBLE = 4,      /* assemble_advection_term */
  S35_BENCH_ILU_APPLY_VEL = 5, /* prec_velocity[0].vmult */
  S35_BENCH_ILU_APPLY_PRES = 6,
  S35_BENCH_DOT = 7,
  S35_BENCH_ADD_AND_DOT = 8,
  S35_BENCH_ILU_FACTOR_VEL = 9,
  S35_BENCH_FP64_FMA = 10      /* FP64 FMA rate of the device; *algorithmic_bytes returns FLOPS per launch */
} s35_bench_kernel;
s35_status s35_ctx_bench_kernel(s35_ctx *ctx, s35_bench_kernel kernel, int32_t warmup, int32_t reps, int flush_l2,
                                double *avg_ms, double *algorithmic_bytes, int64_t *launches);

/* number of kernels this library has launched in the calling process so far */
int64_t s35_launch_count(void);

#ifdef __cplusplus
}
#endif
#endif /* STEP35A_B200_H */
