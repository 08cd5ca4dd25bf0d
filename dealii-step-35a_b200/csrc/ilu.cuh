// ILU(0) on the device: deal.II SparseILU<double>::initialize (Saad's IKJ
// variant, in place, inverse pivots on the diagonal) and ::vmult (forward
// substitution with the unit lower factor, backward substitution with the
// upper factor) -- step-35a.cc:1070-1074, 1312-1319 and the preconditioner
// calls inside SolverGMRES / SolverCG.  SURVEY.md Appendix A.3.
//
// The factors are exactly those of the reference's elimination order; only the
// schedule differs.  Both phases are "sync-free": rows are handed out in a
// topological order (dependency level, then row index) through a ticket
// counter, and a row waits for the rows it depends on.
//  * factorisation: one warp per row, a ready flag per row;
//  * substitution: one warp per row; the solution vector itself is the flag --
//    it is pre-filled with a NaN bit pattern no arithmetic produces, and a
//    consumer polls the entry it needs (through L2) until it changes.
#pragma once
#include "sparse.cuh"

namespace s35 {

struct DevIlu
{
  const DevMatrix *pattern = nullptr; // not owned; must outlive this object
  int32_t n = 0;
  DevBuf<double> lu;                  // factor values on the pattern of A
  DevBuf<int32_t> order_fwd, order_bwd;
  DevBuf<double> y;                   // forward-substitution result, kept armed with the sentinel
  DevBuf<int32_t> ready;              // factorisation: row finished
  DevBuf<int32_t> tickets;            // [0] forward, [1] backward, [2] factor, [3] zero-pivot flag
  int32_t levels_fwd = 0, levels_bwd = 0;
  int per_lane = 2;                   // matrix entries a lane holds in registers per chunk (one warp per row)
  int blocks_per_sm = 4;              // resident blocks per SM of the persistent sweep kernels
  int backoff_ns = 0;                 // longest pause between polling rounds of a waiting row
  bool factored = false;

  // Wavefront mode (for factors whose dependency levels are narrow): a sweep runs in ONE block that walks the levels
  // in order with the recent solution values in shared memory (wavefront.h).  Index 0: forward sweep, 1: backward.
  struct Wave
  {
    bool on = false;
    int32_t n_tiles[2] = {0, 0}, n_levels[2] = {0, 0};
    int32_t groups = 0, group_rows = 0, window = 0, ring_bytes = 0;
    size_t smem_bytes = 0;
    int64_t n_values[2] = {0, 0}, stream_bytes[2] = {0, 0};
    DevBuf<unsigned char> blob[2];
    DevBuf<int32_t> directory[2], rhs_dir[2], value_dst[2], value_src[2];
    DevBuf<int32_t> rhs_slot;        // per row: its place in the forward sweep's rhs layout
    DevBuf<double> by_position[2], rhs[2]; // rhs: the sweeps' right-hand sides in rhs layout
    DevBuf<long long> trace; // S35_WAVE_TRACE: clocks of the forward sweep's tiles
    int trace_calls = 0;
    double far_fraction[2] = {0., 0.}, fresh_fraction[2] = {0., 0.};
  } wave;

  void create(const DevMatrix &A, cudaStream_t s);
  // lu = values (+ strengthen * sum|offdiag| on the diagonal), then factorise
  void factor(const double *values, double strengthen_diagonal, cudaStream_t s);
  // throws S35_ERR_ZERO_PIVOT if the last factorisation met a zero pivot (synchronises the stream)
  void check_pivots(cudaStream_t s);
  // dst = (LU)^-1 src; dst must not alias src
  void apply(double *dst, const double *src, cudaStream_t s);
  double apply_bytes() const { return 12.0 * pattern->nnz + 4.0 * (n + 1.0) + 4.0 * n + 32.0 * n; }
};

} // namespace s35
