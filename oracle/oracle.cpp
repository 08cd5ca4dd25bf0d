// CPU restatement of step-35a's per-time-step path.  TEST INFRASTRUCTURE -- see oracle.h.
// Citations: "ref:N" = /root/reference/step-35a.cc line N; "A.k" = SURVEY.md Appendix A.k.
#include "oracle.h"

#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstring>
#include <limits>
#include <map>
#include <stdexcept>
#include <string>
#include <vector>

#ifdef _OPENMP
#include <omp.h>

#include <mutex>
#include <thread>
#endif

namespace {

int g_threads = 1;
thread_local std::string g_error;

typedef std::vector<double> Vec;

struct Csr
{
  int n_rows = 0, n_cols = 0;
  std::vector<int> rowptr, col, diag;
  void set(int nr, int nc, const int *rp, const int *c)
  {
    n_rows = nr;
    n_cols = nc;
    rowptr.assign(rp, rp + nr + 1);
    col.assign(c, c + rp[nr]);
    diag.assign(nr, -1);
    if (nr == nc)
      for (int r = 0; r < nr; ++r)
        for (int k = rowptr[r]; k < rowptr[r + 1]; ++k)
          if (col[k] == r)
            diag[r] = k;
  }
  int nnz() const { return (int)col.size(); }
};

// SparseMatrix::vmult / vmult_add: row sums in deal.II's storage order, i.e. the diagonal first (square
// matrices), then ascending columns [A.3]
// Threads a matrix-vector product may use.  deal.II's SparseMatrix::vmult splits its rows over the task
// scheduler's workers wherever it is called from; here: the whole team at top level, an equal share inside the
// per-component tasks of the diffusion step, and what the caller set for a side task.
int g_share_in_tasks = 1;
thread_local int t_team_override = 0;
int team_size()
{
  if (t_team_override > 0)
    return t_team_override;
  return omp_get_level() > 0 ? g_share_in_tasks : g_threads;
}

void vmult(const Csr &P, const double *val, const double *x, double *y, bool add)
{
  const int team = team_size();
#pragma omp parallel for schedule(static) num_threads(team) if (team > 1)
  for (int r = 0; r < P.n_rows; ++r)
    {
      double s = 0.0;
      const int d = P.n_rows == P.n_cols ? P.diag[r] : -1;
      if (d >= 0)
        s += val[d] * x[P.col[d]];
      for (int k = P.rowptr[r]; k < P.rowptr[r + 1]; ++k)
        if (k != d)
          s += val[k] * x[P.col[k]];
      y[r] = add ? y[r] + s : s;
    }
}

// SparseMatrix::Tvmult / Tvmult_add: scatter over rows in ascending order
void Tvmult(const Csr &P, const double *val, const double *x, double *y, bool add)
{
  if (!add)
    std::fill(y, y + P.n_cols, 0.0);
  for (int r = 0; r < P.n_rows; ++r)
    for (int k = P.rowptr[r]; k < P.rowptr[r + 1]; ++k)
      y[P.col[k]] += val[k] * x[r];
}

double dot(const double *a, const double *b, int n)
{
  double s = 0.0;
  for (int i = 0; i < n; ++i)
    s += a[i] * b[i];
  return s;
}
double l2_norm(const double *a, int n) { return std::sqrt(dot(a, a, n)); }
// Vector::add_and_dot: this += a * V; return this . W
double add_and_dot(double *self, double a, const double *V, const double *W, int n)
{
  double s = 0.0;
  for (int i = 0; i < n; ++i)
    {
      self[i] += a * V[i];
      s += self[i] * W[i];
    }
  return s;
}

// SparseILU<double>::initialize [A.3]: IKJ ILU(0) in place, inverse pivot on the diagonal.
// blocks > 1 is the gated alternative of s35_problem.ilu_blocks (no counterpart in the reference): couplings
// between different index ranges are zeroed before factoring.  They stay exactly zero through the elimination
// and add exact zeros in the substitutions, so this equals ILU(0) of the block-diagonal part entry for entry.
void ilu0_factor(const Csr &P, const double *a, double strengthen, double *lu, int blocks = 1)
{
  const int n = P.n_rows;
  std::copy(a, a + P.nnz(), lu);
  if (blocks > 1)
    for (int r = 0; r < n; ++r)
      for (int k = P.rowptr[r]; k < P.rowptr[r + 1]; ++k)
        if ((long long)P.col[k] * blocks / n != (long long)r * blocks / n)
          lu[k] = 0.0;
  if (strengthen > 0.0)
    for (int r = 0; r < n; ++r)
      {
        double rowsum = 0.0;
        for (int k = P.rowptr[r]; k < P.rowptr[r + 1]; ++k)
          if (k != P.diag[r])
            rowsum += std::fabs(lu[k]);
        lu[P.diag[r]] += strengthen * rowsum;
      }
  std::vector<int> iw(n, -1);
  for (int k = 0; k < n; ++k)
    {
      const int j1 = P.rowptr[k], j2 = P.rowptr[k + 1];
      for (int j = j1; j < j2; ++j)
        iw[P.col[j]] = j;
      for (int j = j1; j < j2 && P.col[j] < k; ++j)
        {
          const int jrow = P.col[j];
          const double t1 = lu[j] * lu[P.diag[jrow]];
          lu[j] = t1;
          for (int jj = P.diag[jrow] + 1; jj < P.rowptr[jrow + 1]; ++jj)
            {
              const int jw = iw[P.col[jj]];
              if (jw != -1)
                lu[jw] -= t1 * lu[jj];
            }
        }
      if (lu[P.diag[k]] == 0.0)
        throw std::runtime_error("SparseILU: zero pivot");
      lu[P.diag[k]] = 1. / lu[P.diag[k]];
      for (int j = j1; j < j2; ++j)
        iw[P.col[j]] = -1;
    }
}

// SparseILU<double>::vmult [A.3]
void ilu0_apply(const Csr &P, const double *lu, const double *src, double *dst)
{
  const int n = P.n_rows;
  std::copy(src, src + n, dst);
  for (int r = 0; r < n; ++r)
    {
      double v = dst[r];
      for (int k = P.rowptr[r]; k < P.diag[r]; ++k)
        v -= lu[k] * dst[P.col[k]];
      dst[r] = v;
    }
  for (int r = n - 1; r >= 0; --r)
    {
      double v = dst[r];
      for (int k = P.diag[r] + 1; k < P.rowptr[r + 1]; ++k)
        v -= lu[k] * dst[P.col[k]];
      dst[r] = v * lu[P.diag[r]];
    }
}

// SparseLUDecomposition with AdditionalData::extra_off_diagonals = k [A.3]: the factor's pattern is the matrix's pattern
// plus the k sub- and super-diagonals next to the main diagonal (entries (r, r - k .. r + k) inside the matrix).
// source[i]: where entry i of the new pattern sits in the matrix's value array, -1 for an added entry (value 0).
void augment_pattern(const Csr &P, int k, Csr &out, std::vector<int> &source)
{
  const int n = P.n_rows;
  std::vector<int> rp(n + 1, 0), cl;
  source.clear();
  for (int r = 0; r < n; ++r)
    {
      int pos = P.rowptr[r];
      const int end = P.rowptr[r + 1];
      auto flush_below = [&](int c_limit) { // matrix entries with column < c_limit
        while (pos < end && P.col[pos] < c_limit)
          {
            cl.push_back(P.col[pos]);
            source.push_back(pos);
            ++pos;
          }
      };
      for (int c = std::max(0, r - k); c <= std::min(n - 1, r + k); ++c)
        {
          flush_below(c);
          if (pos < end && P.col[pos] == c)
            {
              cl.push_back(c);
              source.push_back(pos);
              ++pos;
            }
          else
            {
              cl.push_back(c);
              source.push_back(-1);
            }
        }
      flush_below(n);
      rp[r + 1] = (int)cl.size();
    }
  out.set(n, n, rp.data(), cl.data());
}

// the matrix's values on the factor's pattern
void gather_on_pattern(const std::vector<int> &source, const double *a, std::vector<double> &out)
{
  out.resize(source.size());
  for (size_t i = 0; i < source.size(); ++i)
    out[i] = source[i] >= 0 ? a[source[i]] : 0.0;
}

// SolverControl::check
enum State
{
  ITERATE,
  SUCCESS,
  FAILURE
};
State check(int step, double value, double tol, int max_steps)
{
  if (value <= tol)
    return SUCCESS;
  if (step >= max_steps || value != value)
    return FAILURE;
  return ITERATE;
}

// SolverGMRES<Vector<double>>::solve, left preconditioning, default residual [A.1]
bool gmres(const Csr &P, const double *A, const double *lu, double *x, const double *b, double tol, int max_its,
           int n_tmp_vectors, s35_solver_stats &stats, const Csr *factor_pattern = nullptr)
{
  const int n = P.n_rows;
  if (n_tmp_vectors < 3)
    throw std::runtime_error("GMRES: Krylov_size must be at least 3");
  std::vector<Vec> tmp(n_tmp_vectors, Vec(n, 0.0));
  std::vector<Vec> H(n_tmp_vectors, Vec(n_tmp_vectors - 1, 0.0));
  Vec gamma(n_tmp_vectors), ci(n_tmp_vectors - 1), si(n_tmp_vectors - 1), h(n_tmp_vectors - 1);
  int accumulated = 0;
  unsigned dim = 0;
  State state = ITERATE;
  double last_res = -std::numeric_limits<double>::max();
  bool re_orthogonalize = false;
  Vec &v = tmp[0], &p = tmp[n_tmp_vectors - 1];
  stats.tolerance = tol;
  stats.initial_residual = 0.0;
  do
    {
      std::fill(h.begin(), h.end(), 0.0);
      vmult(P, A, x, p.data(), false);
      for (int i = 0; i < n; ++i)
        p[i] = -1. * p[i] + 1. * b[i]; // p.sadd(-1., 1., b)
      ilu0_apply(factor_pattern ? *factor_pattern : P, lu, p.data(), v.data());
      double rho = l2_norm(v.data(), n);
      if (accumulated == 0)
        stats.initial_residual = rho;
      last_res = rho;
      state = check(accumulated, rho, tol, max_its);
      if (state != ITERATE)
        break;
      gamma[0] = rho;
      for (int i = 0; i < n; ++i)
        v[i] *= 1. / rho;
      for (int inner = 0; inner < n_tmp_vectors - 2 && state == ITERATE; ++inner)
        {
          ++accumulated;
          Vec &vv = tmp[inner + 1];
          vmult(P, A, tmp[inner].data(), p.data(), false);
          ilu0_apply(factor_pattern ? *factor_pattern : P, lu, p.data(), vv.data());
          dim = inner + 1;
          // modified_gram_schmidt
          double norm_vv_start = 0;
          const bool consider = (re_orthogonalize == false) && (inner % 5 == 4);
          if (consider)
            norm_vv_start = l2_norm(vv.data(), n);
          h[0] = dot(vv.data(), tmp[0].data(), n);
          for (unsigned i = 1; i < dim; ++i)
            h[i] = add_and_dot(vv.data(), -h[i - 1], tmp[i - 1].data(), tmp[i].data(), n);
          double norm_vv = std::sqrt(add_and_dot(vv.data(), -h[dim - 1], tmp[dim - 1].data(), vv.data(), n));
          bool done = false;
          if (consider)
            {
              if (norm_vv > 10. * norm_vv_start * std::sqrt(std::numeric_limits<double>::epsilon()))
                done = true;
              else
                re_orthogonalize = true;
            }
          if (!done && re_orthogonalize)
            {
              double htmp = dot(vv.data(), tmp[0].data(), n);
              h[0] += htmp;
              for (unsigned i = 1; i < dim; ++i)
                {
                  htmp = add_and_dot(vv.data(), -htmp, tmp[i - 1].data(), tmp[i].data(), n);
                  h[i] += htmp;
                }
              norm_vv = std::sqrt(add_and_dot(vv.data(), -htmp, tmp[dim - 1].data(), vv.data(), n));
            }
          const double s = norm_vv;
          h[inner + 1] = s;
          if (std::isfinite(1. / s))
            for (int i = 0; i < n; ++i)
              vv[i] *= 1. / s;
          // givens_rotation
          for (int i = 0; i < inner; ++i)
            {
              const double sg = si[i], cg = ci[i], dummy = h[i];
              h[i] = cg * dummy + sg * h[i + 1];
              h[i + 1] = -sg * dummy + cg * h[i + 1];
            }
          const double r = 1. / std::sqrt(h[inner] * h[inner] + h[inner + 1] * h[inner + 1]);
          si[inner] = h[inner + 1] * r;
          ci[inner] = h[inner] * r;
          h[inner] = ci[inner] * h[inner] + si[inner] * h[inner + 1];
          gamma[inner + 1] = -si[inner] * gamma[inner];
          gamma[inner] *= ci[inner];
          for (unsigned i = 0; i < dim; ++i)
            H[i][inner] = h[i];
          rho = std::fabs(gamma[dim]);
          last_res = rho;
          state = check(accumulated, rho, tol, max_its);
        }
      Vec y(dim);
      for (int i = (int)dim - 1; i >= 0; --i)
        {
          double s = gamma[i];
          for (unsigned j = i + 1; j < dim; ++j)
            s -= H[i][j] * y[j];
          y[i] = s / H[i][i];
        }
      for (unsigned i = 0; i < dim; ++i)
        for (int k = 0; k < n; ++k)
          x[k] += y[i] * tmp[i][k];
    }
  while (state == ITERATE);
  stats.iterations = accumulated;
  stats.final_residual = last_res;
  return state == SUCCESS;
}

// SolverCG<Vector<double>>::solve [A.2]; lu == nullptr: PreconditionIdentity, inv_diag: Jacobi
bool cg(const Csr &P, const double *A, const double *lu, double *x, const double *b, double tol, int max_its,
        s35_solver_stats &stats, const double *inv_diag = nullptr, const Csr *factor_pattern = nullptr)
{
  const int n = P.n_rows;
  Vec g(n), d(n), h(n);
  auto precondition = [&](double *dst, const double *src) {
    if (lu)
      ilu0_apply(factor_pattern ? *factor_pattern : P, lu, src, dst);
    else if (inv_diag)
      for (int i = 0; i < n; ++i)
        dst[i] = inv_diag[i] * src[i];
    else
      std::copy(src, src + n, dst);
  };
  stats.tolerance = tol;
  int it = 0;
  bool all_zero = true;
  for (int i = 0; i < n && all_zero; ++i)
    all_zero = x[i] == 0.0;
  if (!all_zero)
    {
      vmult(P, A, x, g.data(), false);
      for (int i = 0; i < n; ++i)
        g[i] += -1. * b[i];
    }
  else
    for (int i = 0; i < n; ++i)
      g[i] = -1. * b[i];
  double res = l2_norm(g.data(), n);
  stats.initial_residual = res;
  State state = check(0, res, tol, max_its);
  if (state == ITERATE)
    {
      precondition(h.data(), g.data());
      for (int i = 0; i < n; ++i)
        d[i] = -1. * h[i];
      double gh = dot(g.data(), h.data(), n);
      while (state == ITERATE)
        {
          ++it;
          vmult(P, A, d.data(), h.data(), false);
          double alpha = dot(d.data(), h.data(), n);
          alpha = gh / alpha;
          for (int i = 0; i < n; ++i)
            x[i] += alpha * d[i];
          res = std::sqrt(add_and_dot(g.data(), alpha, h.data(), g.data(), n));
          state = check(it, res, tol, max_its);
          if (state != ITERATE)
            break;
          precondition(h.data(), g.data());
          double beta = gh;
          gh = dot(g.data(), h.data(), n);
          beta = gh / beta;
          for (int i = 0; i < n; ++i)
            d[i] = beta * d[i] - 1. * h[i]; // d.sadd(beta, -1., h)
        }
    }
  stats.iterations = it;
  stats.final_residual = res;
  return state == SUCCESS;
}

// ConstraintMatrix::condense(SparseMatrix&, Vector&) for constraints x_c = g_c [A.4]
void condense(const Csr &P, double *val, double *rhs, const std::vector<int> &dofs, const std::vector<double> &g)
{
  const int n = P.n_rows;
  std::vector<int> distribute(n, -1);
  for (size_t i = 0; i < dofs.size(); ++i)
    distribute[dofs[i]] = (int)i;
  double average_diagonal = 0;
  for (int i = 0; i < n; ++i)
    average_diagonal += std::fabs(val[P.diag[i]]);
  average_diagonal /= n;
  for (int row = 0; row < n; ++row)
    {
      if (distribute[row] == -1)
        {
          // deal.II walks the row in storage order: the diagonal (never a constrained column here), then ascending
          for (int k = P.rowptr[row]; k < P.rowptr[row + 1]; ++k)
            {
              const int column = P.col[k];
              if (distribute[column] != -1)
                {
                  rhs[row] -= val[k] * g[distribute[column]];
                  val[k] = 0.;
                }
            }
        }
      else
        {
          for (int k = P.rowptr[row]; k < P.rowptr[row + 1]; ++k)
            {
              const int column = P.col[k];
              if (distribute[column] == -1)
                val[k] = 0.;
              else
                val[k] = (row == column ? average_diagonal : 0.);
            }
          rhs[row] = 0.;
        }
    }
}

// ---------------------------------------------------------------------------------------------
// FEValues stand-in: Q2 shape functions (lexicographic), 3-point Gauss rule, d-linear mapping [A.7]
// ---------------------------------------------------------------------------------------------
void lagrange2(int i, double x, double &v, double &d)
{
  // nodes 0, 1/2, 1
  switch (i)
    {
    case 0:
      v = 2. * (x - .5) * (x - 1.);
      d = 4. * x - 3.;
      break;
    case 1:
      v = -4. * x * (x - 1.);
      d = -8. * x + 4.;
      break;
    default:
      v = 2. * x * (x - .5);
      d = 4. * x - 1.;
      break;
    }
}

struct FE
{
  int dim, dpc, nqp, nv;
  Vec w, value, grad, dN; // value[q][i], grad[q][i][d] (reference), dN[q][v][d] (Q1 mapping)
  explicit FE(int dim_) : dim(dim_)
  {
    dpc = dim == 3 ? 27 : 9;
    nqp = dpc;
    nv = 1 << dim;
    const double a = 0.5 * std::sqrt(0.6);
    const double xq[3] = {0.5 - a, 0.5, 0.5 + a}, wq[3] = {5. / 18., 8. / 18., 5. / 18.};
    w.resize(nqp);
    value.resize(nqp * dpc);
    grad.resize(nqp * dpc * dim);
    dN.resize(nqp * nv * dim);
    for (int q = 0; q < nqp; ++q)
      {
        const int qi[3] = {q % 3, (q / 3) % 3, q / 9};
        double x[3] = {0, 0, 0};
        w[q] = 1.;
        for (int d = 0; d < dim; ++d)
          {
            x[d] = xq[qi[d]];
            w[q] *= wq[qi[d]];
          }
        for (int i = 0; i < dpc; ++i)
          {
            const int ii[3] = {i % 3, (i / 3) % 3, i / 9};
            double v1[3], d1[3];
            for (int d = 0; d < dim; ++d)
              lagrange2(ii[d], x[d], v1[d], d1[d]);
            double v = 1.;
            for (int d = 0; d < dim; ++d)
              v *= v1[d];
            value[q * dpc + i] = v;
            for (int d = 0; d < dim; ++d)
              {
                double g = 1.;
                for (int e = 0; e < dim; ++e)
                  g *= e == d ? d1[e] : v1[e];
                grad[(q * dpc + i) * dim + d] = g;
              }
          }
        for (int v = 0; v < nv; ++v)
          for (int d = 0; d < dim; ++d)
            {
              double g = 1.;
              for (int e = 0; e < dim; ++e)
                {
                  const int bit = (v >> e) & 1;
                  g *= e == d ? (bit ? 1. : -1.) : (bit ? x[e] : 1. - x[e]);
                }
              dN[(q * nv + v) * dim + d] = g;
            }
      }
  }
};

} // namespace

struct oracle_ctx
{
  int dim = 0, n_cells = 0, nv = 0, np = 0;
  double dt = 0, Re = 0, eps = 0, diag_strength = 0;
  int ilu_blocks = 1;
  // off_diagonals > 0 (ref:1070-1074, 1312-1319): the factors live on patterns widened by that many side diagonals
  int off_diagonals = 0;
  Csr velc_ilu, plap_ilu;
  std::vector<int> velc_ilu_src, plap_ilu_src;
  int max_its = 0, krylov = 0, update_prec = 0;
  bool rotational = true;
  FE *fe = nullptr;
  Vec cell_vertices;
  std::vector<int> cell_dofs;
  Csr velc, velm, grad, plap, pmass;
  Vec vel_mass, pres_diff[3], pres_laplace, pres_mass, pmass_inv_diag;
  Vec vel_advection, vel_it[3], prec_vel[3], prec_pres;
  std::vector<int> wall, flow_dofs[3], proj;
  std::vector<char> wall_mask;
  Vec flow_vals[3];
  Vec u_n[3], u_nm1[3], u_star[3], force[3], v_tmp[3];
  Vec pres_n, pres_nm1, phi_n, phi_nm1, pres_tmp;
  s35_solver_stats stats_vel[3], stats_proj, stats_mass;
  std::map<std::string, double> seconds;
  // envelope (skyline) Cholesky factor of pres_Mass, standing in for SparseDirectUMFPACK when it fits
  bool have_cholesky = false;
  std::vector<long> sky_ptr;
  Vec sky;
  ~oracle_ctx() { delete fe; }
};

namespace {

struct Timer
{
  oracle_ctx &c;
  const char *name;
  std::chrono::steady_clock::time_point t0;
  Timer(oracle_ctx &c_, const char *n) : c(c_), name(n), t0(std::chrono::steady_clock::now()) {}
  ~Timer()
  {
    static std::mutex guard; // sections are timed from concurrent tasks
    const double dt = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    std::lock_guard<std::mutex> lock(guard);
    c.seconds[name] += dt;
  }
};

// assemble_one_cell_of_advection (ref:1366-1424), written like the reference: FEValues quantities first,
// then the (q, i, j) triple loop with the bilinear form spelled out term by term
void cell_advection(const oracle_ctx &c, int cell, double *A /* dpc x dpc */, Vec &scratch)
{
  const FE &fe = *c.fe;
  const int dim = c.dim, dpc = fe.dpc, nqp = fe.nqp, nvx = fe.nv;
  // scratch layout: shape_grad[q][i][dim], JxW[q], u_star_local[q][dim], u_star_tmp[q]
  double *shape_grad = scratch.data();
  double *JxW = shape_grad + nqp * dpc * dim;
  double *u_star_local = JxW + nqp;
  double *u_star_tmp = u_star_local + nqp * dim;
  const double *X = &c.cell_vertices[(size_t)cell * nvx * dim];
  const int *ldi = &c.cell_dofs[(size_t)cell * dpc];
  for (int q = 0; q < nqp; ++q)
    {
      double J[3][3] = {{0, 0, 0}, {0, 0, 0}, {0, 0, 0}};
      for (int v = 0; v < nvx; ++v)
        for (int k = 0; k < dim; ++k)
          for (int d = 0; d < dim; ++d)
            J[k][d] += X[v * dim + k] * fe.dN[(q * nvx + v) * dim + d];
      double inv[3][3], det;
      if (dim == 2)
        {
          det = J[0][0] * J[1][1] - J[0][1] * J[1][0];
          inv[0][0] = J[1][1] / det;
          inv[0][1] = -J[0][1] / det;
          inv[1][0] = -J[1][0] / det;
          inv[1][1] = J[0][0] / det;
        }
      else
        {
          det = J[0][0] * (J[1][1] * J[2][2] - J[1][2] * J[2][1]) - J[0][1] * (J[1][0] * J[2][2] - J[1][2] * J[2][0]) +
                J[0][2] * (J[1][0] * J[2][1] - J[1][1] * J[2][0]);
          inv[0][0] = (J[1][1] * J[2][2] - J[1][2] * J[2][1]) / det;
          inv[0][1] = (J[0][2] * J[2][1] - J[0][1] * J[2][2]) / det;
          inv[0][2] = (J[0][1] * J[1][2] - J[0][2] * J[1][1]) / det;
          inv[1][0] = (J[1][2] * J[2][0] - J[1][0] * J[2][2]) / det;
          inv[1][1] = (J[0][0] * J[2][2] - J[0][2] * J[2][0]) / det;
          inv[1][2] = (J[0][2] * J[1][0] - J[0][0] * J[1][2]) / det;
          inv[2][0] = (J[1][0] * J[2][1] - J[1][1] * J[2][0]) / det;
          inv[2][1] = (J[0][1] * J[2][0] - J[0][0] * J[2][1]) / det;
          inv[2][2] = (J[0][0] * J[1][1] - J[0][1] * J[1][0]) / det;
        }
      JxW[q] = det * fe.w[q];
      for (int i = 0; i < dpc; ++i)
        for (int k = 0; k < dim; ++k)
          {
            double s = 0.;
            for (int d = 0; d < dim; ++d)
              s += inv[d][k] * fe.grad[(q * dpc + i) * dim + d];
            shape_grad[(q * dpc + i) * dim + k] = s;
          }
    }
  // get_function_values (ref:1376-1381)
  for (int d = 0; d < dim; ++d)
    for (int q = 0; q < nqp; ++q)
      {
        double s = 0.;
        for (int i = 0; i < dpc; ++i)
          s += c.u_star[d][ldi[i]] * fe.value[q * dpc + i];
        u_star_local[q * dim + d] = s;
      }
  // divergence from get_function_gradients (ref:1383-1392)
  for (int d = 0; d < dim; ++d)
    for (int q = 0; q < nqp; ++q)
      {
        double s = 0.;
        for (int i = 0; i < dpc; ++i)
          s += c.u_star[d][ldi[i]] * shape_grad[(q * dpc + i) * dim + d];
        if (d == 0)
          u_star_tmp[q] = 0.;
        u_star_tmp[q] += s;
      }
  const double time_step_term = 1.5 / c.dt;
  const double inverse_re = 1.0 / c.Re;
  for (int q = 0; q < nqp; ++q)
    {
      u_star_tmp[q] *= 0.5;
      u_star_tmp[q] += time_step_term;
    }
  std::fill(A, A + dpc * dpc, 0.);
  for (int q = 0; q < nqp; ++q)
    for (int i = 0; i < dpc; ++i)
      for (int j = 0; j < dpc; ++j)
        {
          double gg = 0., ug = 0.;
          for (int k = 0; k < dim; ++k)
            {
              gg += shape_grad[(q * dpc + i) * dim + k] * shape_grad[(q * dpc + j) * dim + k];
              ug += u_star_local[q * dim + k] * shape_grad[(q * dpc + j) * dim + k];
            }
          A[i * dpc + j] +=
            (inverse_re * gg + fe.value[q * dpc + i] * (ug + u_star_tmp[q] * fe.value[q * dpc + j])) * JxW[q];
        }
}

// copy_advection_local_to_global (ref:1427-1453) = ConstraintMatrix::distribute_local_to_global with the
// homogeneous wall constraints [A.5]
void scatter_advection(oracle_ctx &c, int cell, const double *A)
{
  const int dpc = c.fe->dpc;
  const int *ldi = &c.cell_dofs[(size_t)cell * dpc];
  double average_diagonal = 0.;
  for (int i = 0; i < dpc; ++i)
    average_diagonal += std::fabs(A[i * dpc + i]);
  average_diagonal /= dpc;
  auto slot = [&](int r, int col) {
    const int *b = &c.velc.col[c.velc.rowptr[r]], *e = &c.velc.col[0] + c.velc.rowptr[r + 1];
    const int *p = std::lower_bound(b, e, col);
    if (p == e || *p != col)
      throw std::runtime_error("oracle: sparsity pattern lacks an entry");
    return (int)(p - &c.velc.col[0]);
  };
  for (int i = 0; i < dpc; ++i)
    {
      if (c.wall_mask[ldi[i]])
        {
          const double a = std::fabs(A[i * dpc + i]);
          c.vel_advection[slot(ldi[i], ldi[i])] += (a != 0. ? a : average_diagonal);
          continue;
        }
      for (int j = 0; j < dpc; ++j)
        if (!c.wall_mask[ldi[j]])
          c.vel_advection[slot(ldi[i], ldi[j])] += A[i * dpc + j];
    }
}

void build_mass_cholesky(oracle_ctx &c)
{
  // envelope of the lower triangle
  const int n = c.np;
  c.sky_ptr.assign(n + 1, 0);
  std::vector<int> first(n);
  long total = 0;
  for (int r = 0; r < n; ++r)
    {
      first[r] = std::min(r, c.pmass.col[c.pmass.rowptr[r]]);
      total += r - first[r] + 1;
      c.sky_ptr[r + 1] = total;
    }
  if (total > 60000000L)
    return; // too large: update_pressure falls back to CG run to 1e-15
  c.sky.assign(total, 0.);
  auto L = [&](int r, int col) -> double & { return c.sky[c.sky_ptr[r] + (col - first[r])]; };
  for (int r = 0; r < n; ++r)
    for (int k = c.pmass.rowptr[r]; k < c.pmass.rowptr[r + 1]; ++k)
      if (c.pmass.col[k] <= r)
        L(r, c.pmass.col[k]) = c.pres_mass[k];
  for (int r = 0; r < n; ++r)
    {
      for (int j = first[r]; j <= r; ++j)
        {
          double s = L(r, j);
          const int lo = std::max(first[r], first[j]);
          for (int k = lo; k < j; ++k)
            s -= L(r, k) * L(j, k);
          if (j < r)
            L(r, j) = s / L(j, j);
          else
            {
              if (!(s > 0.))
                throw std::runtime_error("oracle: pres_Mass is not positive definite");
              L(r, r) = std::sqrt(s);
            }
        }
    }
  c.have_cholesky = true;
  // keep `first` implicit: first[r] = r - (row length - 1)
}

void mass_solve(oracle_ctx &c, Vec &x /* in: rhs, out: solution */)
{
  const int n = c.np;
  if (c.have_cholesky)
    {
      auto first = [&](int r) { return r - (int)(c.sky_ptr[r + 1] - c.sky_ptr[r] - 1); };
      for (int r = 0; r < n; ++r)
        {
          double s = x[r];
          const int f = first(r);
          for (int k = f; k < r; ++k)
            s -= c.sky[c.sky_ptr[r] + (k - f)] * x[k];
          x[r] = s / c.sky[c.sky_ptr[r] + (r - f)];
        }
      for (int r = n - 1; r >= 0; --r)
        {
          const int f = first(r);
          x[r] /= c.sky[c.sky_ptr[r] + (r - f)];
          for (int k = f; k < r; ++k)
            x[k] -= c.sky[c.sky_ptr[r] + (k - f)] * x[r];
        }
      c.stats_mass = s35_solver_stats{0, 0., 0., 0.};
    }
  else
    {
      Vec b = x;
      std::fill(x.begin(), x.end(), 0.);
      cg(c.pmass, c.pres_mass.data(), nullptr, x.data(), b.data(), 1e-15 * l2_norm(b.data(), n), 500, c.stats_mass,
         c.pmass_inv_diag.data());
    }
}

template <typename F>
int guarded(F &&f)
{
  try
    {
      f();
      return 0;
    }
  catch (const std::exception &e)
    {
      g_error = e.what();
      return 1;
    }
}

Vec *field_of(oracle_ctx &c, int field, int comp)
{
  switch (field)
    {
    case S35_U_N: return &c.u_n[comp];
    case S35_U_N_MINUS_1: return &c.u_nm1[comp];
    case S35_U_STAR: return &c.u_star[comp];
    case S35_FORCE: return &c.force[comp];
    case S35_PRES_N: return &c.pres_n;
    case S35_PRES_N_MINUS_1: return &c.pres_nm1;
    case S35_PHI_N: return &c.phi_n;
    case S35_PHI_N_MINUS_1: return &c.phi_nm1;
    case S35_PRES_TMP: return &c.pres_tmp;
    }
  throw std::runtime_error("oracle: unknown field");
}

} // namespace

extern "C" {

void oracle_set_threads(int n) { g_threads = n < 1 ? 1 : n; }
int oracle_get_threads(void) { return g_threads; }
const char *oracle_last_error(void) { return g_error.c_str(); }

int oracle_create(const s35_problem *p, oracle_ctx **out)
{
  return guarded([&] {
    if (p->vel_off_diagonals < 0 || (p->vel_off_diagonals > 0 && p->ilu_blocks > 1))
      throw std::runtime_error("oracle: off_diagonals < 0, or off_diagonals together with ilu_blocks");
    oracle_ctx *c = new oracle_ctx;
    c->dim = p->dim;
    c->n_cells = p->n_cells;
    c->nv = p->n_vel;
    c->np = p->n_pres;
    c->dt = p->dt;
    c->Re = p->reynolds;
    c->eps = p->vel_eps;
    c->diag_strength = p->vel_diag_strength;
    c->ilu_blocks = std::max(1, (int)p->ilu_blocks);
    c->off_diagonals = p->vel_off_diagonals;
    c->max_its = p->vel_max_iterations;
    c->krylov = p->vel_krylov_size;
    c->update_prec = p->vel_update_prec;
    c->rotational = p->rotational != 0;
    c->fe = new FE(p->dim);
    const int dpc = c->fe->dpc, nvx = c->fe->nv, dim = p->dim;
    c->cell_vertices.assign(p->cell_vertices, p->cell_vertices + (size_t)p->n_cells * nvx * dim);
    c->cell_dofs.assign(p->vel_cell_dofs, p->vel_cell_dofs + (size_t)p->n_cells * dpc);
    c->velc.set(c->nv, c->nv, p->velc_rowptr, p->velc_col);
    c->velm.set(c->nv, c->nv, p->velm_rowptr, p->velm_col);
    c->grad.set(c->nv, c->np, p->grad_rowptr, p->grad_col);
    c->plap.set(c->np, c->np, p->plap_rowptr, p->plap_col);
    if (c->off_diagonals > 0)
      {
        augment_pattern(c->velc, c->off_diagonals, c->velc_ilu, c->velc_ilu_src);
        augment_pattern(c->plap, c->off_diagonals, c->plap_ilu, c->plap_ilu_src);
      }
    c->pmass.set(c->np, c->np, p->pmass_rowptr, p->pmass_col);
    c->vel_mass.assign(p->velm_val, p->velm_val + c->velm.nnz());
    c->pres_laplace.assign(p->plap_val, p->plap_val + c->plap.nnz());
    c->pres_mass.assign(p->pmass_val, p->pmass_val + c->pmass.nnz());
    c->pmass_inv_diag.resize(c->np);
    for (int r = 0; r < c->np; ++r)
      c->pmass_inv_diag[r] = 1. / c->pres_mass[c->pmass.diag[r]];
    c->wall.assign(p->wall_dofs, p->wall_dofs + p->n_wall);
    c->wall_mask.assign(c->nv, 0);
    for (int i : c->wall)
      c->wall_mask[i] = 1;
    c->proj.assign(p->proj_dofs, p->proj_dofs + p->n_proj);
    c->vel_advection.assign(c->velc.nnz(), 0.);
    for (int d = 0; d < dim; ++d)
      {
        c->pres_diff[d].assign(p->grad_val[d], p->grad_val[d] + c->grad.nnz());
        c->flow_dofs[d].assign(p->flow_dofs[d], p->flow_dofs[d] + p->n_flow[d]);
        c->flow_vals[d].assign(p->flow_vals[d], p->flow_vals[d] + p->n_flow[d]);
        c->vel_it[d].assign(c->velc.nnz(), 0.);
        c->prec_vel[d].assign(c->off_diagonals > 0 ? c->velc_ilu.nnz() : c->velc.nnz(), 0.);
        for (Vec *v : {&c->u_n[d], &c->u_nm1[d], &c->u_star[d], &c->force[d], &c->v_tmp[d]})
          v->assign(c->nv, 0.);
      }
    for (Vec *v : {&c->pres_n, &c->pres_nm1, &c->phi_n, &c->phi_nm1, &c->pres_tmp})
      v->assign(c->np, 0.);
    // pressure preconditioners are built once (ref:1070-1075)
    if (c->off_diagonals > 0)
      {
        Vec widened;
        gather_on_pattern(c->plap_ilu_src, c->pres_laplace.data(), widened);
        c->prec_pres.resize(c->plap_ilu.nnz());
        ilu0_factor(c->plap_ilu, widened.data(), c->diag_strength, c->prec_pres.data());
      }
    else
      {
        c->prec_pres.resize(c->plap.nnz());
        ilu0_factor(c->plap, c->pres_laplace.data(), c->diag_strength, c->prec_pres.data(), c->ilu_blocks);
      }
    build_mass_cholesky(*c);
    std::memset(c->stats_vel, 0, sizeof(c->stats_vel));
    c->stats_proj = c->stats_mass = s35_solver_stats{0, 0., 0., 0.};
    *out = c;
  });
}

void oracle_destroy(oracle_ctx *c) { delete c; }

int oracle_set_field(oracle_ctx *c, int field, int comp, const double *values)
{
  return guarded([&] {
    Vec *v = field_of(*c, field, comp);
    std::copy(values, values + v->size(), v->begin());
  });
}
int oracle_get_field(oracle_ctx *c, int field, int comp, double *values)
{
  return guarded([&] {
    Vec *v = field_of(*c, field, comp);
    std::copy(v->begin(), v->end(), values);
  });
}
int oracle_get_operator_values(oracle_ctx *c, int op, int comp, double *values)
{
  return guarded([&] {
    const Vec *v = op == S35_VEL_ADVECTION ? &c->vel_advection : op == S35_VEL_IT_MATRIX ? &c->vel_it[comp] : &c->prec_vel[comp];
    if (op == S35_VEL_ILU && c->off_diagonals > 0)
      { // the factor's entries at the matrix's own positions; the added side diagonals have no place in the caller's array
        for (size_t i = 0; i < c->velc_ilu_src.size(); ++i)
          if (c->velc_ilu_src[i] >= 0)
            values[c->velc_ilu_src[i]] = (*v)[i];
        return;
      }
    std::copy(v->begin(), v->end(), values);
  });
}

int oracle_interpolate_velocity(oracle_ctx *c)
{
  return guarded([&] {
    Timer t(*c, "interpolate_velocity");
    for (int d = 0; d < c->dim; ++d)
      for (int i = 0; i < c->nv; ++i)
        {
          c->u_star[d][i] = 2.0 * c->u_n[d][i];     // equ(2.0, u_n)
          c->u_star[d][i] += -1.0 * c->u_nm1[d][i]; // add(-1.0, u_n_minus_1)
        }
  });
}

int oracle_assemble_advection_term(oracle_ctx *c)
{
  return guarded([&] {
    Timer t(*c, "assemble_advection_term");
    std::fill(c->vel_advection.begin(), c->vel_advection.end(), 0.); // ref:1351
    const int dpc = c->fe->dpc, nqp = c->fe->nqp, dim = c->dim;
    const size_t scratch_len = (size_t)nqp * dpc * dim + nqp + (size_t)nqp * dim + nqp;
    if (g_threads <= 1)
      {
        Vec A(dpc * dpc), scratch(scratch_len);
        for (int cell = 0; cell < c->n_cells; ++cell)
          {
            cell_advection(*c, cell, A.data(), scratch);
            scatter_advection(*c, cell, A.data());
          }
      }
    else
      {
        // WorkStream: workers in parallel, copier serial and in cell order (ref:1357-1362)
        const int batch = 2048;
        Vec A((size_t)batch * dpc * dpc);
        for (int begin = 0; begin < c->n_cells; begin += batch)
          {
            const int end = std::min(c->n_cells, begin + batch);
#pragma omp parallel num_threads(g_threads)
            {
              Vec scratch(scratch_len);
#pragma omp for schedule(dynamic, 16)
              for (int cell = begin; cell < end; ++cell)
                cell_advection(*c, cell, &A[(size_t)(cell - begin) * dpc * dpc], scratch);
            }
            for (int cell = begin; cell < end; ++cell)
              scatter_advection(*c, cell, &A[(size_t)(cell - begin) * dpc * dpc]);
          }
      }
  });
}

int oracle_diffusion_step(oracle_ctx *c, int reinit_prec)
{
  return guarded([&] {
    const int dim = c->dim, nv = c->nv, np = c->np;
    const double dt = c->dt;
    for (int i = 0; i < np; ++i)
      {
        c->pres_tmp[i] = -1.0 * c->pres_n[i];                                        // equ(-1.0, pres_n)
        c->pres_tmp[i] += -4.0 / 3.0 * c->phi_n[i] + 1.0 / 3.0 * c->phi_nm1[i];      // add(-4/3, phi_n, 1/3, phi_n_minus_1)
      }
    {
      Timer t(*c, "setup_diffusion_rhs");
#pragma omp parallel for num_threads(std::min(g_threads, dim)) if (g_threads > 1)
      for (int d = 0; d < dim; ++d)
        {
          for (int i = 0; i < nv; ++i)
            {
              c->force[d][i] = 0.;
              c->v_tmp[d][i] = 2.0 / dt * c->u_n[d][i];
              c->v_tmp[d][i] += -0.5 / dt * c->u_nm1[d][i];
            }
          vmult(c->velm, c->vel_mass.data(), c->v_tmp[d].data(), c->force[d].data(), true);
          vmult(c->grad, c->pres_diff[d].data(), c->pres_tmp.data(), c->force[d].data(), true);
          c->u_nm1[d] = c->u_n[d];
          c->vel_it[d] = c->vel_advection; // copy_from
        }
    }
    {
      Timer t(*c, "solve_diffusion_component");
      std::string failure;
      // one task per component as in the reference (ref:1277-1325); the products inside share the other threads
      g_share_in_tasks = std::max(1, g_threads / dim);
      omp_set_max_active_levels(2);
#pragma omp parallel for num_threads(std::min(g_threads, dim)) if (g_threads > 1)
      for (int d = 0; d < dim; ++d)
        {
          try
            {
              condense(c->velc, c->vel_it[d].data(), c->force[d].data(), c->flow_dofs[d], c->flow_vals[d]);
              if (reinit_prec && c->off_diagonals > 0)
                {
                  Vec widened;
                  gather_on_pattern(c->velc_ilu_src, c->vel_it[d].data(), widened);
                  ilu0_factor(c->velc_ilu, widened.data(), c->diag_strength, c->prec_vel[d].data());
                }
              else if (reinit_prec)
                ilu0_factor(c->velc, c->vel_it[d].data(), c->diag_strength, c->prec_vel[d].data(), c->ilu_blocks);
              // diffusion_component_solve (ref:1330-1338)
              const double tol = c->eps * l2_norm(c->force[d].data(), nv);
              const bool ok = gmres(c->velc, c->vel_it[d].data(), c->prec_vel[d].data(), c->u_n[d].data(), c->force[d].data(),
                                    tol, c->max_its, c->krylov, c->stats_vel[d], c->off_diagonals > 0 ? &c->velc_ilu : nullptr);
              if (!ok)
                throw std::runtime_error("GMRES: no convergence");
              for (size_t i = 0; i < c->flow_dofs[d].size(); ++i)
                c->u_n[d][c->flow_dofs[d][i]] = c->flow_vals[d][i];
              for (int i : c->wall)
                c->u_n[d][i] = 0.;
            }
          catch (const std::exception &e)
            {
#pragma omp critical
              failure = e.what();
            }
        }
      if (!failure.empty())
        throw std::runtime_error(failure);
    }
  });
}

int oracle_projection_step(oracle_ctx *c)
{
  return guarded([&] {
    const int np = c->np;
    {
      Timer t(*c, "projection_step_setup");
      Tvmult(c->grad, c->pres_diff[0].data(), c->u_n[0].data(), c->pres_tmp.data(), false);
      for (int d = 1; d < c->dim; ++d)
        Tvmult(c->grad, c->pres_diff[d].data(), c->u_n[d].data(), c->pres_tmp.data(), true);
      c->phi_nm1 = c->phi_n;
    }
    {
      Timer t(*c, "projection_step_solve");
      const double tol = c->eps * l2_norm(c->pres_tmp.data(), np); // before condense (ref:1496-1498)
      for (int i : c->proj)
        c->pres_tmp[i] = 0.;
      const bool ok = cg(c->plap, c->pres_laplace.data(), c->prec_pres.data(), c->phi_n.data(), c->pres_tmp.data(), tol,
                         c->max_its, c->stats_proj, nullptr, c->off_diagonals > 0 ? &c->plap_ilu : nullptr);
      if (!ok)
        throw std::runtime_error("CG: no convergence");
      for (int i : c->proj)
        c->phi_n[i] = 0.;
    }
    {
      Timer t(*c, "projection_step_setup");
      for (int i = 0; i < np; ++i)
        c->phi_n[i] *= 1.5 / c->dt;
    }
  });
}

int oracle_update_pressure(oracle_ctx *c)
{
  return guarded([&] {
    Timer t(*c, "update_pressure");
    const int np = c->np;
    c->pres_nm1 = c->pres_n;
    if (!c->rotational)
      for (int i = 0; i < np; ++i)
        c->pres_n[i] += c->phi_n[i];
    else
      {
        c->pres_n = c->pres_tmp;
        mass_solve(*c, c->pres_n); // prec_pres_Mass.solve(pres_n), UMFPACK in the reference (ref:1537)
        for (int i = 0; i < np; ++i)
          {
            c->pres_n[i] = 1.0 / c->Re * c->pres_n[i] + 1.0 * c->pres_nm1[i]; // sadd(1/Re, 1, pres_n_minus_1)
            c->pres_n[i] += 1.0 * c->phi_n[i];
          }
      }
  });
}

int oracle_advance(oracle_ctx *c, int first_n, int n_steps, int start_n, int start_offset_n)
{
  for (int n = first_n; n < first_n + n_steps; ++n)
    {
      int rc = 0;
      if (n == start_offset_n)
        {
          rc = rc || oracle_interpolate_velocity(c);
          rc = rc || oracle_assemble_advection_term(c);
        }
      rc = rc || oracle_diffusion_step(c, (n % c->update_prec == 0) || (n == start_n));
      rc = rc || oracle_interpolate_velocity(c);
      if (rc)
        return rc;
      // Two tasks as in the reference (ref:1219-1235): the assembly next to projection + pressure update.  They
      // touch disjoint data, so the results do not depend on running them together; with one thread they run
      // one after the other.
      if (g_threads > 1)
        {
          int rc_side = 0;
          std::string side_error; // the message lives in the task's own thread
          std::thread side([&] {
            t_team_override = std::min(2, g_threads);
            rc_side = oracle_projection_step(c);
            rc_side = rc_side ? rc_side : oracle_update_pressure(c);
            if (rc_side)
              side_error = g_error;
          });
          rc = oracle_assemble_advection_term(c);
          side.join();
          if (!rc && rc_side)
            {
              g_error = side_error;
              rc = rc_side;
            }
        }
      else
        {
          rc = rc || oracle_assemble_advection_term(c);
          rc = rc || oracle_projection_step(c);
          rc = rc || oracle_update_pressure(c);
        }
      if (rc)
        return rc;
    }
  return 0;
}

int oracle_solver_stats(oracle_ctx *c, int component, s35_solver_stats *stats)
{
  *stats = component == -1 ? c->stats_proj : component == -2 ? c->stats_mass : c->stats_vel[component];
  return 0;
}

int oracle_section_seconds(oracle_ctx *c, const char *section, double *seconds)
{
  auto it = c->seconds.find(section);
  *seconds = it == c->seconds.end() ? 0. : it->second;
  return 0;
}

int oracle_spmv(int n_rows, int n_cols, const int *rowptr, const int *col, const double *val, const double *x, double *y,
                int transpose, int add)
{
  return guarded([&] {
    Csr P;
    P.set(n_rows, n_cols, rowptr, col);
    if (transpose)
      Tvmult(P, val, x, y, add != 0);
    else
      vmult(P, val, x, y, add != 0);
  });
}

int oracle_condense(int n, const int *rowptr, const int *col, double *val, double *rhs, int n_constrained, const int *dofs,
                    const double *values)
{
  return guarded([&] {
    Csr P;
    P.set(n, n, rowptr, col);
    std::vector<int> d(dofs, dofs + n_constrained);
    std::vector<double> g(n_constrained, 0.);
    if (values)
      g.assign(values, values + n_constrained);
    condense(P, val, rhs, d, g);
  });
}

int oracle_ilu0_factor(int n, const int *rowptr, const int *col, const double *val, double strengthen, double *lu)
{
  return guarded([&] {
    Csr P;
    P.set(n, n, rowptr, col);
    ilu0_factor(P, val, strengthen, lu);
  });
}

int oracle_ilu0_apply(int n, const int *rowptr, const int *col, const double *lu, const double *src, double *dst)
{
  return guarded([&] {
    Csr P;
    P.set(n, n, rowptr, col);
    ilu0_apply(P, lu, src, dst);
  });
}

int oracle_gmres(int n, const int *rowptr, const int *col, const double *val, const double *lu, double *x, const double *b,
                 double tol, int max_its, int krylov_size, s35_solver_stats *stats)
{
  int converged = 1;
  const int rc = guarded([&] {
    Csr P;
    P.set(n, n, rowptr, col);
    s35_solver_stats st = {0, 0., 0., 0.};
    converged = gmres(P, val, lu, x, b, tol, max_its, krylov_size, st) ? 1 : 0;
    if (stats)
      *stats = st;
  });
  return rc ? rc : (converged ? 0 : 3);
}

int oracle_cg(int n, const int *rowptr, const int *col, const double *val, const double *lu, double *x, const double *b,
              double tol, int max_its, s35_solver_stats *stats)
{
  int converged = 1;
  const int rc = guarded([&] {
    Csr P;
    P.set(n, n, rowptr, col);
    s35_solver_stats st = {0, 0., 0., 0.};
    converged = cg(P, val, lu, x, b, tol, max_its, st) ? 1 : 0;
    if (stats)
      *stats = st;
  });
  return rc ? rc : (converged ? 0 : 3);
}

} // extern "C"
