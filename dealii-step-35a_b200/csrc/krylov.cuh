// Device Krylov solvers that follow deal.II's SolverGMRES / SolverCG step by
// step (call sites step-35a.cc:1334-1337 and
// 1496-1499), so that they stop at the same iterate as the reference.
// Vectors live on the device; the Hessenberg / Givens bookkeeping of GMRES
// (a handful of scalars per iteration) is done on the host exactly as
// deal.II does it, after one read-back per iteration.
//
// On a mesh-partitioned problem (`dist` set) a vector holds the rank's owned entries followed by its ghost
// entries: vector operations and dot products run over the owned part, every dot product is completed by a
// sum over the ranks, and the ghost entries of a mat-vec's input are refreshed right before it.
#pragma once
#include "comm.cuh"
#include "ilu.cuh"
#include "vecops.cuh"

#include <vector>

namespace s35 {

// SolverControl::check
enum class SolverState
{
  iterate,
  success,
  failure
};
inline SolverState solver_check(int step, double value, double tol, int max_steps)
{
  if (value <= tol)
    return SolverState::success;
  if (step >= max_steps || value != value)
    return SolverState::failure;
  return SolverState::iterate;
}

// how a solver reaches the other ranks; both null on a single GPU
struct DistLink
{
  Comm *comm = nullptr;
  HaloDev *halo = nullptr;
  // x's ghost entries <- owners' values
  void refresh_ghosts(double *x, cudaStream_t s) const
  {
    if (comm && halo)
      comm->exchange(*halo, x, s);
  }
  // completes `count` local sums at `result` (device) and applies the square root if asked
  void finish_sums(double *result, int count, bool take_sqrt, cudaStream_t s) const
  {
    if (comm)
      {
        comm->allreduce_sum(result, count, s);
        if (take_sqrt)
          for (int i = 0; i < count; ++i)
            scalar_sqrt(result + i, s);
      }
  }
  // the reduction kernels apply the square root themselves only when no cross-rank sum follows
  ReducePost local_post(bool take_sqrt) const { return (take_sqrt && !comm) ? kSqrt : kSum; }
};

struct GmresSolver
{
  int32_t n = 0;          // owned entries: length of vector operations
  int32_t n_local = 0;    // owned + ghost entries: length of a stored vector
  int n_tmp = 0;          // AdditionalData::max_n_tmp_vectors (= Krylov_size)
  int64_t stride = 0;     // distance between basis vectors
  DevBuf<double> tmp;     // n_tmp vectors
  DevBuf<double> h_dev;   // MGS coefficients + norms
  PinnedBuf<double> h_host;
  ReduceWs ws;
  cudaStream_t stream = 0;
  DistLink dist;

  void init(int32_t n_owned, int32_t n_local, int krylov_size, cudaStream_t s);
  // returns false on SolverControl failure (iteration limit / NaN); x holds the last iterate either way
  bool solve(const DevMatrix &pattern, const double *values, DevIlu &prec, double *x, const double *b, double tol,
             int max_iterations, s35_solver_stats &stats);
};

struct CgSolver
{
  int32_t n = 0, n_local = 0;
  DevBuf<double> g, d, h;
  DevBuf<double> scal;    // [0] d.h  [1] alpha  [2] gh  [3] res  [4] gh_new  [5] beta  [6] tol  [7] max its  [8] it  [9] state
  PinnedBuf<double> res_host;
  cudaEvent_t res_copied = nullptr; // host-driven loop: the residual of the iteration has reached res_host
  ReduceWs ws;
  cudaStream_t stream = 0;
  DistLink dist;

  // Single GPU: the iteration runs as ONE graph launch -- a WHILE node (CUDA conditional graph node) whose body is one
  // CG iteration and whose condition a one-thread kernel sets from SolverControl::check on the device -- so that no
  // iteration waits for the host.  The iterates and the stopping iteration are those of the host-driven loop.  One
  // graph per (matrix, preconditioner, x, b) combination is kept.  S35_CG_GRAPH=0 switches back to the host loop.
  struct LoopGraph
  {
    const void *key[6] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
    cudaGraph_t graph = nullptr;
    cudaGraphExec_t exec = nullptr;
    int64_t launches_per_iteration = 0;
  };
  std::vector<LoopGraph> loops;
  ~CgSolver();

  void init(int32_t n_owned, int32_t n_local, cudaStream_t s);
  // prec may be null; then inv_diag (Jacobi, h = inv_diag .* g) is used if given, else PreconditionIdentity
  bool solve(const DevMatrix &pattern, const double *values, DevIlu *prec, double *x, const double *b, double tol,
             int max_iterations, s35_solver_stats &stats, const double *inv_diag = nullptr);
};

} // namespace s35
