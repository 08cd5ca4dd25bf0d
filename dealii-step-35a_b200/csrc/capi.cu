// extern "C" layer of include/step35a_b200.h: argument checks, exception -> status translation, and the
// host <-> device copies of the entry points that take host vectors.
#include "ctx.cuh"

#include <cstring>
#include <functional>

using namespace s35;

struct s35_matrix
{
  DevMatrix m;
  cudaStream_t stream = nullptr;
  ReduceWs ws;
};
struct s35_ilu
{
  DevIlu p;
  const s35_matrix *owner = nullptr;
};
struct s35_ctx
{
  Ctx *c = nullptr;
};

namespace {
thread_local std::string g_last_error;
}
namespace s35 {
void set_last_error(const std::string &msg) { g_last_error = msg; }
} // namespace s35

namespace {

template <typename F>
s35_status guarded(F &&f)
{
  try
    {
      f();
      return S35_OK;
    }
  catch (const Error &e)
    {
      g_last_error = e.what();
      return e.code;
    }
  catch (const std::bad_alloc &)
    {
      g_last_error = "out of host memory";
      return S35_ERR_INVALID;
    }
  catch (const std::exception &e)
    {
      g_last_error = e.what();
      return S35_ERR_INVALID;
    }
  catch (...)
    {
      g_last_error = "unknown failure";   // nothing may unwind through the C boundary
      return S35_ERR_INVALID;
    }
}

void require_device()
{
  int n = 0;
  const cudaError_t e = cudaGetDeviceCount(&n);
  if (e != cudaSuccess || n == 0)
    {
      cudaGetLastError();
      throw Error(S35_ERR_CUDA, "no CUDA device available: this library has no CPU fallback");
    }
}

// temporary device vector filled from / copied back to host memory
struct Staged
{
  DevBuf<double> d;
  Staged(const double *host, int64_t n, cudaStream_t s) : d((size_t)n)
  {
    if (host && n)
      S35_CUDA(cudaMemcpyAsync(d.ptr, host, n * sizeof(double), cudaMemcpyHostToDevice, s));
  }
  void back(double *host, cudaStream_t s)
  {
    if (d.n)
      {
        S35_CUDA(cudaMemcpyAsync(host, d.ptr, d.n * sizeof(double), cudaMemcpyDeviceToHost, s));
        S35_CUDA(cudaStreamSynchronize(s));
      }
  }
};

} // namespace

namespace {
// eight independent FMA chains per thread: enough to cover the 8-cycle latency at 8 warps per scheduler
__global__ void k_fp64_fma(double *out, int iters, int n_out)
{
  double a0 = threadIdx.x * 1e-9, a1 = a0 + 1., a2 = a0 + 2., a3 = a0 + 3., a4 = a0 + 4., a5 = a0 + 5., a6 = a0 + 6., a7 = a0 + 7.;
  const double m = 1.0 - 1e-12, b = 1e-12;
  for (int k = 0; k < iters; ++k)
    {
      a0 = fma(a0, m, b);
      a1 = fma(a1, m, b);
      a2 = fma(a2, m, b);
      a3 = fma(a3, m, b);
      a4 = fma(a4, m, b);
      a5 = fma(a5, m, b);
      a6 = fma(a6, m, b);
      a7 = fma(a7, m, b);
    }
  const double sum = ((a0 + a1) + (a2 + a3)) + ((a4 + a5) + (a6 + a7));
  if (sum == 123.456) // never: keeps the chains alive
    out[(blockIdx.x * blockDim.x + threadIdx.x) % n_out] = sum;
}
} // namespace

extern "C" {

const char *s35_last_error(void) { return g_last_error.c_str(); }
const char *s35_version(void) { return "step35a_b200 0.1 (sm_100a)"; }

int s35_device_count(void)
{
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess)
    {
      cudaGetLastError();
      return 0;
    }
  return n;
}

s35_status s35_set_device(int device)
{
  return guarded([&] {
    require_device();
    S35_CUDA(cudaSetDevice(device));
  });
}

int64_t s35_launch_count(void) { return g_launch_count.load(); }

/* ---------------- matrices ---------------- */

s35_status s35_matrix_create(int32_t n_rows, int32_t n_cols, const int32_t *rowptr, const int32_t *col, const double *val,
                             s35_matrix **out)
{
  return guarded([&] {
    S35_REQUIRE(out != nullptr, "matrix_create: null output pointer");
    *out = nullptr;
    require_device();
    std::unique_ptr<s35_matrix> m(new s35_matrix);
    S35_CUDA(cudaStreamCreateWithFlags(&m->stream, cudaStreamNonBlocking));
    m->m.create(n_rows, n_cols, rowptr, col, val, m->stream);
    m->ws.init(m->stream);
    S35_CUDA(cudaStreamSynchronize(m->stream));
    *out = m.release();
  });
}

void s35_matrix_destroy(s35_matrix *m)
{
  if (!m)
    return;
  if (m->stream)
    cudaStreamDestroy(m->stream);
  delete m;
}

s35_status s35_matrix_set_values(s35_matrix *m, const double *val)
{
  return guarded([&] {
    S35_REQUIRE(m && val, "matrix_set_values: null argument");
    m->m.set_values(val, m->stream);
  });
}

s35_status s35_matrix_get_values(const s35_matrix *m, double *val)
{
  return guarded([&] {
    S35_REQUIRE(m && val, "matrix_get_values: null argument");
    m->m.val.download(val, m->stream);
  });
}

int64_t s35_matrix_nnz(const s35_matrix *m) { return m ? m->m.nnz : 0; }

s35_status s35_spmv(const s35_matrix *A, const double *x, double *y, int transpose, int add)
{
  return guarded([&] {
    S35_REQUIRE(A && x && y, "spmv: null argument");
    cudaStream_t s = A->stream;
    const DevMatrix &M = transpose ? A->m.transpose(s) : A->m;
    Staged dx(x, M.n_cols, s), dy(add ? y : nullptr, M.n_rows, s);
    spmv(M, dx.d, dy.d, add != 0, s);
    dy.back(y, s);
  });
}

s35_status s35_condense(s35_matrix *A, double *rhs, int32_t n_constrained, const int32_t *dofs, const double *values)
{
  return guarded([&] {
    S35_REQUIRE(A && rhs && (n_constrained == 0 || dofs), "condense: null argument");
    S35_REQUIRE(A->m.n_rows == A->m.n_cols, "condense: matrix must be square");
    cudaStream_t s = A->stream;
    CondensePlan plan;
    plan.create(A->m, n_constrained, dofs, values, s);
    Staged drhs(rhs, A->m.n_rows, s);
    DevBuf<double> avg(1);
    diag_abs_mean(A->m, A->m.val, avg, A->ws, s);
    plan.condense(A->m, A->m.val, drhs.d, avg, s);
    A->m.values_changed();
    drhs.back(rhs, s);
  });
}

/* ---------------- ILU(0) ---------------- */

s35_status s35_ilu_create(const s35_matrix *A, s35_ilu **out)
{
  return guarded([&] {
    S35_REQUIRE(A && out, "ilu_create: null argument");
    *out = nullptr;
    std::unique_ptr<s35_ilu> p(new s35_ilu);
    p->owner = A;
    p->p.create(A->m, A->stream);
    S35_CUDA(cudaStreamSynchronize(A->stream));
    *out = p.release();
  });
}

void s35_ilu_destroy(s35_ilu *p) { delete p; }

s35_status s35_ilu_factor(s35_ilu *p, const s35_matrix *A, double strengthen_diagonal)
{
  return guarded([&] {
    S35_REQUIRE(p && A, "ilu_factor: null argument");
    S35_REQUIRE(A->m.nnz == p->p.pattern->nnz && A->m.n_rows == p->p.n, "ilu_factor: matrix does not match the pattern");
    p->p.factor(A->m.val, strengthen_diagonal, p->owner->stream);
    p->p.check_pivots(p->owner->stream);
  });
}

s35_status s35_ilu_apply(s35_ilu *p, const double *src, double *dst)
{
  return guarded([&] {
    S35_REQUIRE(p && src && dst, "ilu_apply: null argument");
    cudaStream_t s = p->owner->stream;
    Staged dsrc(src, p->p.n, s), ddst(nullptr, p->p.n, s);
    p->p.apply(ddst.d, dsrc.d, s);
    ddst.back(dst, s);
  });
}

s35_status s35_ilu_get_values(const s35_ilu *p, double *val)
{
  return guarded([&] {
    S35_REQUIRE(p && val, "ilu_get_values: null argument");
    p->p.lu.download(val, p->owner->stream);
  });
}

s35_status s35_ilu_levels(const s35_ilu *p, int32_t *forward_levels, int32_t *backward_levels)
{
  return guarded([&] {
    S35_REQUIRE(p, "ilu_levels: null argument");
    if (forward_levels)
      *forward_levels = p->p.levels_fwd;
    if (backward_levels)
      *backward_levels = p->p.levels_bwd;
  });
}

/* ---------------- Krylov solvers ---------------- */

s35_status s35_gmres(const s35_matrix *A, s35_ilu *prec, double *x, const double *b, double tolerance, int32_t max_iterations,
                     int32_t krylov_size, s35_solver_stats *stats)
{
  return guarded([&] {
    S35_REQUIRE(A && prec && x && b, "gmres: null argument");
    S35_REQUIRE(A->m.n_rows == A->m.n_cols && prec->p.n == A->m.n_rows, "gmres: size mismatch");
    cudaStream_t s = A->stream;
    const int32_t n = A->m.n_rows;
    GmresSolver solver;
    solver.init(n, n, krylov_size, s);
    Staged dx(x, n, s), db(b, n, s);
    s35_solver_stats st = {};
    const bool ok = solver.solve(A->m, A->m.val, prec->p, dx.d, db.d, tolerance, max_iterations, st);
    dx.back(x, s);
    if (stats)
      *stats = st;
    if (!ok)
      throw Error(S35_ERR_NO_CONVERGENCE, "GMRES: no convergence after " + std::to_string(st.iterations) + " iterations");
  });
}

s35_status s35_cg(const s35_matrix *A, s35_ilu *prec, double *x, const double *b, double tolerance, int32_t max_iterations,
                  s35_solver_stats *stats)
{
  return guarded([&] {
    S35_REQUIRE(A && x && b, "cg: null argument");
    S35_REQUIRE(A->m.n_rows == A->m.n_cols && (!prec || prec->p.n == A->m.n_rows), "cg: size mismatch");
    cudaStream_t s = A->stream;
    const int32_t n = A->m.n_rows;
    CgSolver solver;
    solver.init(n, n, s);
    Staged dx(x, n, s), db(b, n, s);
    s35_solver_stats st = {};
    const bool ok = solver.solve(A->m, A->m.val, prec ? &prec->p : nullptr, dx.d, db.d, tolerance, max_iterations, st);
    dx.back(x, s);
    if (stats)
      *stats = st;
    if (!ok)
      throw Error(S35_ERR_NO_CONVERGENCE, "CG: no convergence after " + std::to_string(st.iterations) + " iterations");
  });
}

/* ---------------- vector kernels ---------------- */

s35_status s35_vec_lincomb(int64_t n, double a, const double *x, double b, const double *y, double *z)
{
  return guarded([&] {
    S35_REQUIRE(n >= 0 && x && z && (y || b == 0.0), "vec_lincomb: null argument");
    require_device();
    Staged dx(x, n, 0), dy(y, y ? n : 0, 0), dz(nullptr, n, 0);
    vec_lincomb(dz.d, DScal(a), dx.d, DScal(b), y ? dy.d.ptr : nullptr, n, 0);
    dz.back(z, 0);
  });
}

s35_status s35_vec_dot(int64_t n, const double *x, const double *y, double *result)
{
  return guarded([&] {
    S35_REQUIRE(n >= 0 && x && y && result, "vec_dot: null argument");
    require_device();
    ReduceWs ws;
    ws.init(0);
    Staged dx(x, n, 0), dy(y, n, 0), dr(nullptr, 1, 0);
    vec_dot(dr.d, dx.d, dy.d, n, kSum, ws, 0);
    dr.back(result, 0);
  });
}

s35_status s35_vec_add_and_dot(int64_t n, double a, const double *x, double *y, const double *w, double *result)
{
  return guarded([&] {
    S35_REQUIRE(n >= 0 && x && y && w && result, "vec_add_and_dot: null argument");
    require_device();
    ReduceWs ws;
    ws.init(0);
    Staged dx(x, n, 0), dy(y, n, 0), dw(w == y ? nullptr : w, w == y ? 0 : n, 0), dr(nullptr, 1, 0);
    vec_add_and_dot(dr.d, dy.d, DScal(a), dx.d, w == y ? dy.d.ptr : dw.d.ptr, n, kSum, ws, 0);
    dy.back(y, 0);
    dr.back(result, 0);
  });
}

/* ---------------- communicator of the partitioned path ---------------- */

s35_status s35_comm_unique_id(unsigned char id[128])
{
  return guarded([&] {
    S35_REQUIRE(id != nullptr, "comm_unique_id: null argument");
    require_device();
    comm_unique_id(id);
  });
}

s35_status s35_comm_init(int32_t rank, int32_t n_ranks, const unsigned char unique_id[128])
{
  return guarded([&] {
    S35_REQUIRE(unique_id != nullptr, "comm_init: null argument");
    require_device();
    comm_init(rank, n_ranks, unique_id);
  });
}

void s35_comm_finalize(void) { comm_finalize(); }

s35_status s35_comm_thread_group(int32_t n_ranks)
{
  return guarded([&] { comm_thread_group(n_ranks); });
}

s35_status s35_comm_thread_bind(int32_t rank)
{
  return guarded([&] {
    require_device();
    comm_thread_bind(rank);
  });
}

/* ---------------- time-step context ---------------- */

s35_status s35_ctx_create(const s35_problem *problem, s35_ctx **out)
{
  return guarded([&] {
    S35_REQUIRE(problem && out, "ctx_create: null argument");
    *out = nullptr;
    require_device();
    std::unique_ptr<s35_ctx> h(new s35_ctx);
    h->c = new Ctx(*problem);
    *out = h.release();
  });
}

void s35_ctx_destroy(s35_ctx *ctx)
{
  if (!ctx)
    return;
  delete ctx->c;
  delete ctx;
}

s35_status s35_ctx_set_field(s35_ctx *ctx, s35_field field, int32_t component, const double *values)
{
  return guarded([&] {
    S35_REQUIRE(ctx && values, "ctx_set_field: null argument");
    int64_t len = 0;
    double *d = ctx->c->field(field, component, len);
    S35_CUDA(cudaMemcpyAsync(d, values, len * sizeof(double), cudaMemcpyHostToDevice, ctx->c->stream));
    S35_CUDA(cudaStreamSynchronize(ctx->c->stream));
  });
}

s35_status s35_ctx_get_field(s35_ctx *ctx, s35_field field, int32_t component, double *values)
{
  return guarded([&] {
    S35_REQUIRE(ctx && values, "ctx_get_field: null argument");
    int64_t len = 0;
    double *d = ctx->c->field(field, component, len);
    S35_CUDA(cudaMemcpyAsync(values, d, len * sizeof(double), cudaMemcpyDeviceToHost, ctx->c->stream));
    S35_CUDA(cudaStreamSynchronize(ctx->c->stream));
  });
}

s35_status s35_ctx_get_operator_values(s35_ctx *ctx, s35_operator op, int32_t component, double *values)
{
  return guarded([&] {
    S35_REQUIRE(ctx && values, "ctx_get_operator_values: null argument");
    Ctx &c = *ctx->c;
    const double *src = nullptr;
    switch (op)
      {
      case S35_VEL_ADVECTION: src = c.vel_adv.val; break;
      case S35_VEL_IT_MATRIX:
        S35_REQUIRE(component >= 0 && component < c.dim, "component out of range");
        src = c.it_val[component];
        break;
      case S35_VEL_ILU:
        S35_REQUIRE(component >= 0 && component < c.dim, "component out of range");
        // with ilu_blocks > 1 or on a partition the factor lives on a filtered pattern (fewer entries, other positions):
        // there is no value array on the caller's pattern to hand out
        if (c.off_diagonals > 0)
          { // the factor's entries at the matrix's own positions; the added side diagonals have no place in the caller's array
            std::vector<double> wide(c.vel_blk.nnz);
            S35_CUDA(cudaMemcpyAsync(wide.data(), c.prec_vel[component].lu, wide.size() * sizeof(double), cudaMemcpyDeviceToHost,
                                     c.stream));
            S35_CUDA(cudaStreamSynchronize(c.stream));
            for (size_t i = 0; i < wide.size(); ++i)
              if (c.vel_wide_src[i] >= 0)
                values[c.vel_wide_src[i]] = wide[i];
            return;
          }
        if (c.use_blocks)
          throw Error(S35_ERR_UNSUPPORTED, "the ILU factor of a block / partitioned preconditioner is not on the matrix pattern");
        src = c.prec_vel[component].lu;
        break;
      default: throw Error(S35_ERR_INVALID, "unknown operator id");
      }
    S35_CUDA(cudaMemcpyAsync(values, src, c.vel_adv.nnz * sizeof(double), cudaMemcpyDeviceToHost, c.stream));
    S35_CUDA(cudaStreamSynchronize(c.stream));
  });
}

#define S35_CTX_CALL(expr)                               \
  return guarded([&] {                                   \
    S35_REQUIRE(ctx != nullptr, "null context");         \
    Ctx &c = *ctx->c;                                    \
    S35_CUDA(cudaSetDevice(c.device));                   \
    expr;                                                \
    c.resolve_sections();                                \
  })

s35_status s35_interpolate_velocity(s35_ctx *ctx) { S35_CTX_CALL(c.interpolate_velocity()); }
s35_status s35_assemble_advection_term(s35_ctx *ctx) { S35_CTX_CALL(c.assemble_advection_term()); }
s35_status s35_diffusion_step(s35_ctx *ctx, int reinit_prec) { S35_CTX_CALL(c.diffusion_step(reinit_prec != 0)); }
s35_status s35_projection_step(s35_ctx *ctx) { S35_CTX_CALL(c.projection_step()); }
s35_status s35_update_pressure(s35_ctx *ctx) { S35_CTX_CALL(c.update_pressure()); }

s35_status s35_advance(s35_ctx *ctx, int32_t first_n, int32_t n_steps, int32_t start_n, int32_t start_offset_n,
                       double *elapsed_ms)
{
  S35_CTX_CALL(c.advance(first_n, n_steps, start_n, start_offset_n, elapsed_ms));
}

s35_status s35_synchronize(s35_ctx *ctx)
{
  return guarded([&] {
    S35_REQUIRE(ctx != nullptr, "null context");
    S35_CUDA(cudaStreamSynchronize(ctx->c->stream));
  });
}

s35_status s35_ctx_solver_stats(s35_ctx *ctx, int32_t component, s35_solver_stats *stats)
{
  return guarded([&] {
    S35_REQUIRE(ctx && stats, "ctx_solver_stats: null argument");
    Ctx &c = *ctx->c;
    if (component == -1)
      *stats = c.stats_proj;
    else if (component == -2)
      *stats = c.stats_mass;
    else
      {
        S35_REQUIRE(component >= 0 && component < c.dim, "component out of range");
        *stats = c.stats_vel[component];
      }
  });
}

s35_status s35_ctx_section_ms(s35_ctx *ctx, const char *section, double *ms, int64_t *calls)
{
  return guarded([&] {
    S35_REQUIRE(ctx && section, "ctx_section_ms: null argument");
    auto it = ctx->c->sections.find(section);
    if (ms)
      *ms = it == ctx->c->sections.end() ? 0.0 : it->second.ms;
    if (calls)
      *calls = it == ctx->c->sections.end() ? 0 : it->second.calls;
  });
}

s35_status s35_ctx_bench_kernel(s35_ctx *ctx, s35_bench_kernel kernel, int32_t warmup, int32_t reps, int flush,
                                double *avg_ms, double *algorithmic_bytes, int64_t *launches)
{
  return guarded([&] {
    S35_REQUIRE(ctx && avg_ms && reps > 0 && warmup >= 0, "ctx_bench_kernel: bad argument");
    Ctx &c = *ctx->c;
    cudaStream_t s = c.stream;
    double bytes = 0.0;
    // scratch operands: GMRES work vectors (velocity sized) and CG work vectors (pressure sized)
    double *xv = c.gmres[0].tmp.ptr, *yv = c.gmres[0].tmp.ptr + c.gmres[0].stride, *zv = c.gmres[0].tmp.ptr + 2 * c.gmres[0].stride;
    double *xp = c.cg.g, *yp = c.cg.d;
    vec_set(xv, 1.0, c.nv, s);
    vec_set(yv, 0.5, c.nv, s);
    vec_set(zv, 0.25, c.nv, s);
    vec_set(xp, 1.0, c.np, s);
    const double *us[3] = {c.u_star[0], c.u_star[1], c.u_star[2]};
    const double *gvals[3] = {c.gradT_val[0], c.gradT_val[1], c.gradT_val[2]};
    const double *gx[3] = {xv, yv, zv};
    // what prec_vel[0] factorises: the iteration matrix's values, gathered onto the block pattern when the factor lives
    // there (ilu_blocks > 1, or a partition) -- as Ctx::solve_component does
    auto factor_input = [&]() -> const double * {
      vec_copy(c.it_val[0], c.vel_adv.val, c.vel_adv.nnz, s);
      if (!c.use_blocks)
        return c.it_val[0];
      gather_values(c.blk_val[0], c.it_val[0], c.vel_blk_src, c.vel_blk.nnz, s);
      return c.blk_val[0];
    };
    if ((kernel == S35_BENCH_ILU_APPLY_VEL) && !c.prec_vel[0].factored)
      c.prec_vel[0].factor(factor_input(), c.diag_strength, s);
    std::function<void()> run;
    switch (kernel)
      {
      case S35_BENCH_SPMV_VEL:
        bytes = c.vel_adv.spmv_bytes();
        run = [&] { spmv_values(c.vel_adv, c.it_val[0], xv, yv, false, s); };
        break;
      case S35_BENCH_SPMV_PRES:
        bytes = c.plap.spmv_bytes();
        run = [&] { spmv(c.plap, xp, yp, false, s); };
        break;
      case S35_BENCH_SPMV_GRAD:
        bytes = c.grad.spmv_bytes();
        run = [&] { spmv_values(c.grad, c.grad_val[0], xp, yv, false, s); };
        break;
      case S35_BENCH_SPMV_GRAD_T:
        bytes = (4.0 + 8.0 * c.dim) * c.grad.nnz + 4.0 * (c.np + 1) + 8.0 * c.np + 8.0 * c.dim * c.nv;
        run = [&] { spmv_multi(*c.gradT, c.dim, gvals, gx, yp, false, s); };
        break;
      case S35_BENCH_ASSEMBLE:
        bytes = c.assembler.algorithmic_bytes(c.vel_adv.nnz, c.n_vertices, c.nv);
        run = [&] { c.assembler.assemble(c.vel_adv.val, c.vel_adv.nnz, us, c.dt, c.reynolds, s); };
        break;
      case S35_BENCH_ILU_APPLY_VEL:
        bytes = c.prec_vel[0].apply_bytes();
        run = [&] { c.prec_vel[0].apply(yv, xv, s); };
        break;
      case S35_BENCH_ILU_APPLY_PRES:
        bytes = c.prec_pres.apply_bytes();
        run = [&] { c.prec_pres.apply(yp, xp, s); };
        break;
      case S35_BENCH_DOT:
        bytes = 16.0 * c.nv;
        run = [&] { vec_dot(c.scal.ptr + 6, xv, yv, c.nv, kSum, c.ws, s); };
        break;
      case S35_BENCH_ADD_AND_DOT:
        bytes = 32.0 * c.nv;
        run = [&] { vec_add_and_dot(c.scal.ptr + 6, yv, DScal(1e-9), xv, zv, c.nv, kSum, c.ws, s); };
        break;
      case S35_BENCH_ILU_FACTOR_VEL:
        bytes = 20.0 * (c.use_blocks ? c.vel_blk.nnz : c.vel_adv.nnz);
        {
          const double *input = factor_input();
          run = [&c, input, s] { c.prec_vel[0].factor(input, c.diag_strength, s); };
        }
        break;
      case S35_BENCH_FP64_FMA:
        {
          // the FP64 FMA rate this GPU sustains: the roof of the assembly ("measure FP64 peak")
          const int blocks = sm_count() * 8, iters = 4096;
          bytes = 2.0 * 8.0 * iters * (double)blocks * 256.0; // flops, not bytes
          run = [&c, blocks, iters, yv, s] { S35_LAUNCH(k_fp64_fma, blocks, 256, 0, s, yv, iters, (int)c.nv); };
        }
        break;
      default: throw Error(S35_ERR_INVALID, "unknown bench kernel");
      }
    for (int i = 0; i < warmup; ++i)
      run();
    cudaEvent_t e0, e1;
    S35_CUDA(cudaEventCreate(&e0));
    S35_CUDA(cudaEventCreate(&e1));
    double total = 0.0;
    const int64_t launches_before = g_launch_count.load();
    for (int i = 0; i < reps; ++i)
      {
        if (flush)
          flush_l2(s);
        S35_CUDA(cudaEventRecord(e0, s));
        run();
        S35_CUDA(cudaEventRecord(e1, s));
        S35_CUDA(cudaEventSynchronize(e1));
        float ms = 0.f;
        S35_CUDA(cudaEventElapsedTime(&ms, e0, e1));
        total += ms;
      }
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    *avg_ms = total / reps;
    if (algorithmic_bytes)
      *algorithmic_bytes = bytes;
    if (launches)
      *launches = (g_launch_count.load() - launches_before) / reps;
    if (kernel == S35_BENCH_ASSEMBLE)
      c.vel_adv.values_changed();
  });
}

} // extern "C"
