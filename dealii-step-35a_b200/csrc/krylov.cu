#include "krylov.cuh"

#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <limits>

namespace s35 {

void GmresSolver::init(int32_t n_owned, int32_t n_local_, int krylov_size, cudaStream_t s)
{
  S35_REQUIRE(krylov_size >= 3, "GMRES: Krylov_size must be at least 3 (two of the vectors are work space)");
  S35_REQUIRE(krylov_size <= 34, "GMRES: Krylov_size above 34 is not supported by this build");
  n = n_owned;
  n_local = n_local_;
  n_tmp = krylov_size;
  stride = ((int64_t)n_local + 1) & ~(int64_t)1;
  tmp.alloc((size_t)stride * n_tmp);
  tmp.zero(s);
  h_dev.alloc(n_tmp + 4);
  h_host.alloc(n_tmp + 4);
  ws.init(s);
  stream = s;
}

bool GmresSolver::solve(const DevMatrix &P, const double *values, DevIlu &prec, double *x, const double *b, double tol,
                        int max_iterations, s35_solver_stats &stats)
{
  cudaStream_t s = stream;
  auto vec = [&](int i) { return tmp.ptr + (size_t)i * stride; };
  double *const v = vec(0), *const p = vec(n_tmp - 1);
  double *const hd = h_dev.ptr;
  double *const hh = h_host.ptr;
  const int i_norm = n_tmp, i_start = n_tmp + 1; // slots for ||vv|| and ||vv||_start
  auto matvec = [&](double *in, double *out) {
    dist.refresh_ghosts(in, s);
    spmv_values(P, values, in, out, false, s);
  };
  auto dot = [&](double *result, const double *a, const double *c, bool take_sqrt) {
    vec_dot(result, a, c, n, dist.local_post(take_sqrt), ws, s);
    dist.finish_sums(result, 1, take_sqrt, s);
  };
  auto add_and_dot = [&](double *result, double *y, DScal a, const double *xx, const double *w, bool take_sqrt) {
    vec_add_and_dot(result, y, a, xx, w, n, dist.local_post(take_sqrt), ws, s);
    dist.finish_sums(result, 1, take_sqrt, s);
  };

  std::vector<double> H((size_t)n_tmp * (n_tmp - 1), 0.0); // H(i,j) = H[i*(n_tmp-1)+j]
  std::vector<double> gamma(n_tmp), ci(n_tmp - 1), si(n_tmp - 1), h(n_tmp - 1);
  auto Hm = [&](int i, int j) -> double & { return H[(size_t)i * (n_tmp - 1) + j]; };

  int accumulated = 0;
  unsigned dim = 0;
  SolverState state = SolverState::iterate;
  double last_res = -std::numeric_limits<double>::max();
  bool reorthogonalize = false;
  stats.initial_residual = 0.0;
  stats.tolerance = tol;

  // modified Gram-Schmidt pass over the first `dim` basis vectors; coefficients at hd[0..dim), norm at hd[i_norm]
  auto gram_schmidt_pass = [&](double *vv) {
    dot(hd + 0, vv, vec(0), false);
    for (unsigned i = 1; i < dim; ++i)
      add_and_dot(hd + i, vv, DScal(hd + i - 1, -1.0), vec(i - 1), vec(i), false);
    add_and_dot(hd + i_norm, vv, DScal(hd + dim - 1, -1.0), vec(dim - 1), vv, true);
    S35_CUDA(cudaMemcpyAsync(hh, hd, (n_tmp + 2) * sizeof(double), cudaMemcpyDeviceToHost, s));
    S35_CUDA(cudaStreamSynchronize(s));
  };

  do
    {
      std::fill(h.begin(), h.end(), 0.0);
      // p = b - A x ; v = P^-1 p
      matvec(x, p);
      vec_lincomb(p, DScal(-1.0), p, DScal(1.0), b, n, s);
      prec.apply(v, p, s);
      dot(hd + i_norm, v, v, true);
      S35_CUDA(cudaMemcpyAsync(hh + i_norm, hd + i_norm, sizeof(double), cudaMemcpyDeviceToHost, s));
      S35_CUDA(cudaStreamSynchronize(s));
      double rho = hh[i_norm];
      if (accumulated == 0)
        stats.initial_residual = rho;
      last_res = rho;
      state = solver_check(accumulated, rho, tol, max_iterations);
      if (state != SolverState::iterate)
        break;
      gamma[0] = rho;
      vec_lincomb(v, DScal(1.0 / rho), v, DScal(0.0), nullptr, n, s);

      for (int inner = 0; inner < n_tmp - 2 && state == SolverState::iterate; ++inner)
        {
          ++accumulated;
          double *vv = vec(inner + 1);
          matvec(vec(inner), p);
          prec.apply(vv, p, s);
          dim = inner + 1;

          // SolverGMRES::modified_gram_schmidt
          const bool consider_reorthogonalize = !reorthogonalize && (inner % 5 == 4);
          if (consider_reorthogonalize)
            dot(hd + i_start, vv, vv, true);
          gram_schmidt_pass(vv);
          for (unsigned i = 0; i < dim; ++i)
            h[i] = hh[i];
          double norm_vv = hh[i_norm];
          if (consider_reorthogonalize)
            {
              const double norm_vv_start = hh[i_start];
              if (!(norm_vv > 10. * norm_vv_start * std::sqrt(std::numeric_limits<double>::epsilon())))
                reorthogonalize = true;
            }
          if (reorthogonalize)
            {
              // second pass, entered once loss of orthogonality has been detected; corrections add to h
              gram_schmidt_pass(vv);
              for (unsigned i = 0; i < dim; ++i)
                h[i] += hh[i];
              norm_vv = hh[i_norm];
            }
          const double sn = norm_vv;
          h[inner + 1] = sn;
          // s = 0 is a lucky breakdown; do not divide by zero
          if (std::isfinite(1. / sn))
            vec_lincomb(vv, DScal(1. / sn), vv, DScal(0.0), nullptr, n, s);

          // SolverGMRES::givens_rotation
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
            Hm(i, inner) = h[i];
          rho = std::fabs(gamma[dim]);
          last_res = rho;
          state = solver_check(accumulated, rho, tol, max_iterations);
        }
      // back substitution H1 y = gamma (FullMatrix::backward), then x += sum y_i v_i
      std::vector<double> y(dim);
      for (int i = (int)dim - 1; i >= 0; --i)
        {
          double sum = gamma[i];
          for (unsigned j = i + 1; j < dim; ++j)
            sum -= Hm(i, j) * y[j];
          y[i] = sum / Hm(i, i);
        }
      vec_multi_axpy(x, vec(0), stride, y.data(), (int)dim, n, s);
    }
  while (state == SolverState::iterate);

  stats.iterations = accumulated;
  stats.final_residual = last_res;
  return state == SolverState::success;
}

namespace {

// before the loop graph: tolerance and iteration limit where the device-side check reads them
__global__ void k_cg_arm(double *sc, double tol, double max_iterations)
{
  sc[6] = tol;
  sc[7] = max_iterations;
  sc[8] = 0.0; // iterations done
  sc[9] = 0.0; // SolverControl state: 0 iterate, 1 success, 2 failure
}

// SolverControl::check (solver_check above) after the residual update of an iteration, and the WHILE condition
__global__ void k_cg_check(double *sc, cudaGraphConditionalHandle loop)
{
  const double it = sc[8] + 1.0, res = sc[3];
  sc[8] = it;
  unsigned int go_on = 1;
  if (res <= sc[6])
    {
      sc[9] = 1.0;
      go_on = 0;
    }
  else if (it >= sc[7] || res != res)
    {
      sc[9] = 2.0;
      go_on = 0;
    }
  cudaGraphSetConditional(loop, go_on);
}

bool cg_graph_enabled()
{
  const char *t = std::getenv("S35_CG_GRAPH");
  return !(t && std::atoi(t) == 0);
}

} // namespace

CgSolver::~CgSolver()
{
  if (res_copied)
    cudaEventDestroy(res_copied);
  for (LoopGraph &l : loops)
    {
      if (l.exec)
        cudaGraphExecDestroy(l.exec);
      if (l.graph)
        cudaGraphDestroy(l.graph);
    }
}

void CgSolver::init(int32_t n_owned, int32_t n_local_, cudaStream_t s)
{
  n = n_owned;
  n_local = n_local_;
  g.alloc(n_local);
  d.alloc(n_local);
  h.alloc(n_local);
  g.zero(s);
  d.zero(s);
  h.zero(s);
  scal.alloc(16);
  scal.zero(s);
  res_host.alloc(4);
  ws.init(s);
  stream = s;
}

bool CgSolver::solve(const DevMatrix &P, const double *values, DevIlu *prec, double *x, const double *b, double tol,
                     int max_iterations, s35_solver_stats &stats, const double *inv_diag)
{
  cudaStream_t s = stream;
  double *const sc = scal.ptr;
  enum
  {
    DH = 0,
    ALPHA = 1,
    GH = 2,
    RES = 3,
    GHNEW = 4,
    BETA = 5
  };
  auto read_res = [&]() {
    S35_CUDA(cudaMemcpyAsync(res_host.ptr, sc + RES, sizeof(double), cudaMemcpyDeviceToHost, s));
    S35_CUDA(cudaStreamSynchronize(s));
    return res_host.ptr[0];
  };
  auto precondition = [&](double *dst, const double *src) {
    if (prec)
      prec->apply(dst, src, s);
    else if (inv_diag)
      vec_multiply(dst, inv_diag, src, n, s);
    else
      vec_copy(dst, src, n, s);
  };
  auto matvec = [&](double *in, double *out) {
    dist.refresh_ghosts(in, s);
    spmv_values(P, values, in, out, false, s);
  };
  auto dot = [&](double *result, const double *a, const double *c) {
    vec_dot(result, a, c, n, kSum, ws, s);
    dist.finish_sums(result, 1, false, s);
  };
  stats.tolerance = tol;
  int it = 0;
  // g = A x - b (for x == 0 this is exactly -b, the reference's short cut)
  matvec(x, g);
  vec_add_and_dot(sc + RES, g, DScal(-1.0), b, g, n, dist.local_post(true), ws, s);
  dist.finish_sums(sc + RES, 1, true, s);
  double res = read_res();
  stats.initial_residual = res;
  SolverState state = solver_check(0, res, tol, max_iterations);
  if (state == SolverState::iterate)
    {
      precondition(h, g);
      vec_lincomb(d, DScal(-1.0), h, DScal(0.0), nullptr, n, s);
      dot(sc + GH, g, h);
      if (!dist.comm && cg_graph_enabled())
        {
          // the whole loop as one graph launch (krylov.cuh); one graph per combination of operands
          const void *key[6] = {P.rowptr.ptr, values, prec, x, b, inv_diag};
          LoopGraph *loop = nullptr;
          for (LoopGraph &l : loops)
            if (std::equal(key, key + 6, l.key))
              loop = &l;
          if (!loop)
            {
              if (loops.size() >= 8) // a caller that keeps changing its operands: forget the oldest graph
                {
                  cudaGraphExecDestroy(loops.front().exec);
                  cudaGraphDestroy(loops.front().graph);
                  loops.erase(loops.begin());
                }
              loops.emplace_back();
              loop = &loops.back();
              std::copy(key, key + 6, loop->key);
              S35_CUDA(cudaGraphCreate(&loop->graph, 0));
              cudaGraphConditionalHandle handle;
              S35_CUDA(cudaGraphConditionalHandleCreate(&handle, loop->graph, 1, cudaGraphCondAssignDefault));
              cudaGraphNodeParams params = {};
              params.type = cudaGraphNodeTypeConditional;
              params.conditional.handle = handle;
              params.conditional.type = cudaGraphCondTypeWhile;
              params.conditional.size = 1;
              cudaGraphNode_t node;
              S35_CUDA(cudaGraphAddNode(&node, loop->graph, nullptr, 0, &params));
              cudaGraph_t body = params.conditional.phGraph_out[0];
              const int64_t launches_before = g_launch_count.load();
              S35_CUDA(cudaStreamBeginCaptureToGraph(s, body, nullptr, nullptr, 0, cudaStreamCaptureModeRelaxed));
              try
                {
                  // one iteration; after the check the rest only touches h, d and the scalars, so running it once more
                  // after the stopping iteration (the loop ends at the END of a body) leaves x and g as the reference has them
                  matvec(d, h);
                  dot(sc + DH, d, h);
                  scalar_divide(sc + ALPHA, sc + GH, sc + DH, nullptr, s);
                  vec_cg_update(sc + RES, x, g, DScal(sc + ALPHA), d, h, n, kSqrt, ws, s);
                  S35_LAUNCH(k_cg_check, 1, 1, 0, s, sc, handle);
                  precondition(h, g);
                  dot(sc + GHNEW, g, h);
                  scalar_divide(sc + BETA, sc + GHNEW, sc + GH, sc + GH, s);
                  vec_lincomb(d, DScal(sc + BETA), d, DScal(-1.0), h, n, s);
                }
              catch (...)
                {
                  cudaGraph_t dummy = nullptr;
                  cudaStreamEndCapture(s, &dummy);
                  throw;
                }
              S35_CUDA(cudaStreamEndCapture(s, nullptr));
              loop->launches_per_iteration = g_launch_count.load() - launches_before;
              g_launch_count.fetch_sub(loop->launches_per_iteration); // recorded, not run
              S35_CUDA(cudaGraphInstantiate(&loop->exec, loop->graph, 0));
            }
          S35_LAUNCH(k_cg_arm, 1, 1, 0, s, sc, tol, (double)max_iterations);
          S35_CUDA(cudaGraphLaunch(loop->exec, s));
          S35_CUDA(cudaMemcpyAsync(res_host.ptr, sc + RES, sizeof(double), cudaMemcpyDeviceToHost, s));
          S35_CUDA(cudaMemcpyAsync(res_host.ptr + 1, sc + 8, 2 * sizeof(double), cudaMemcpyDeviceToHost, s));
          S35_CUDA(cudaStreamSynchronize(s));
          res = res_host.ptr[0];
          it = (int)res_host.ptr[1];
          state = res_host.ptr[2] == 1.0 ? SolverState::success : SolverState::failure;
          g_launch_count.fetch_add((int64_t)it * loop->launches_per_iteration);
        }
      while (state == SolverState::iterate)
        {
          ++it;
          matvec(d, h);
          dot(sc + DH, d, h);
          scalar_divide(sc + ALPHA, sc + GH, sc + DH, nullptr, s);
          vec_cg_update(sc + RES, x, g, DScal(sc + ALPHA), d, h, n, dist.local_post(true), ws, s);
          dist.finish_sums(sc + RES, 1, true, s);
          // The rest of the iteration touches only h, gh, beta and d -- not x, not g -- so it is launched before the
          // host has seen the residual: the wait for the copy and the next launches overlap with the preconditioner
          // instead of leaving the GPU idle.  If this was the last iteration, one application was for nothing.
          if (!res_copied)
            S35_CUDA(cudaEventCreateWithFlags(&res_copied, cudaEventDisableTiming));
          S35_CUDA(cudaMemcpyAsync(res_host.ptr, sc + RES, sizeof(double), cudaMemcpyDeviceToHost, s));
          S35_CUDA(cudaEventRecord(res_copied, s));
          precondition(h, g);
          dot(sc + GHNEW, g, h);
          // beta = gh_new / gh_old ; gh = gh_new
          scalar_divide(sc + BETA, sc + GHNEW, sc + GH, sc + GH, s);
          vec_lincomb(d, DScal(sc + BETA), d, DScal(-1.0), h, n, s);
          S35_CUDA(cudaEventSynchronize(res_copied));
          res = res_host.ptr[0];
          state = solver_check(it, res, tol, max_iterations);
        }
    }
  stats.iterations = it;
  stats.final_residual = res;
  return state == SolverState::success;
}

} // namespace s35
