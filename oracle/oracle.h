/*
 * oracle -- CPU restatement of the per-time-step path of step-35a's NavierStokesProjection.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing under oracle/ is part of the product: only tests/,
 * __graft_entry__.smoke() and bench.py's CPU-baseline legs may build, load or call it.
 *
 * PARITY UNPINNED: the reference ships no tests, golden vectors or fixtures, and its arithmetic lives in
 * deal.II (8.5-9.0, un-vendored, not installable here), so this restatement cannot be checked against
 * reference outputs.  It follows step-35a.cc:1167-1545 for control flow and coefficients and SURVEY.md
 * Appendix A (recalled deal.II semantics: SolverGMRES, SolverCG, SparseILU, ConstraintMatrix) for the
 * library calls; it is cross-checked against an independent numpy/scipy restatement (oracle/oracle_np.py)
 * and against analytic identities (tests/).
 *
 * Serial and deterministic: one thread, fixed summation orders (cells in index order, matrix rows with
 * the diagonal first and then ascending columns like deal.II's SparsityPattern).  oracle_set_threads(n>1)
 * switches on OpenMP for the timed CPU baseline in the way the reference uses TBB (cells of the
 * assembly, the dim velocity components, matrix rows in mat-vecs); results stay bit-identical except for
 * the mat-vec row partition, which does not change them either.
 *
 * The problem description is the same plain struct the device library takes (include/step35a_b200.h).
 */
#ifndef S35_ORACLE_H
#define S35_ORACLE_H

#include "../include/step35a_b200.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef struct oracle_ctx oracle_ctx;

void oracle_set_threads(int n);
int oracle_get_threads(void);
const char *oracle_last_error(void);

int oracle_create(const s35_problem *problem, oracle_ctx **out);
void oracle_destroy(oracle_ctx *ctx);
int oracle_set_field(oracle_ctx *ctx, int field, int component, const double *values);
int oracle_get_field(oracle_ctx *ctx, int field, int component, double *values);
int oracle_get_operator_values(oracle_ctx *ctx, int op, int component, double *values);

int oracle_interpolate_velocity(oracle_ctx *ctx);          /* step-35a.cc:1244-1254 */
int oracle_assemble_advection_term(oracle_ctx *ctx);       /* step-35a.cc:1346-1453 */
int oracle_diffusion_step(oracle_ctx *ctx, int reinit_prec); /* step-35a.cc:1268-1338 */
int oracle_projection_step(oracle_ctx *ctx);               /* step-35a.cc:1456-1508 */
int oracle_update_pressure(oracle_ctx *ctx);               /* step-35a.cc:1511-1545 */
int oracle_advance(oracle_ctx *ctx, int first_n, int n_steps, int start_n, int start_offset_n); /* :1190-1239 */
int oracle_solver_stats(oracle_ctx *ctx, int component, s35_solver_stats *stats);
/* wall-clock seconds accumulated per reference timer section */
int oracle_section_seconds(oracle_ctx *ctx, const char *section, double *seconds);

/* stand-alone pieces, CSR with ascending columns */
int oracle_spmv(int n_rows, int n_cols, const int *rowptr, const int *col, const double *val, const double *x, double *y,
                int transpose, int add);
int oracle_condense(int n, const int *rowptr, const int *col, double *val, double *rhs, int n_constrained, const int *dofs,
                    const double *values);
int oracle_ilu0_factor(int n, const int *rowptr, const int *col, const double *val, double strengthen, double *lu);
int oracle_ilu0_apply(int n, const int *rowptr, const int *col, const double *lu, const double *src, double *dst);
int oracle_gmres(int n, const int *rowptr, const int *col, const double *val, const double *lu, double *x, const double *b,
                 double tol, int max_its, int krylov_size, s35_solver_stats *stats);
int oracle_cg(int n, const int *rowptr, const int *col, const double *val, const double *lu, double *x, const double *b,
              double tol, int max_its, s35_solver_stats *stats);

#ifdef __cplusplus
}
#endif
#endif
