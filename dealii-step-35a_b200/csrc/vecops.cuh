// Fused vector kernels (deal.II Vector<double>::equ / add / sadd / operator* /
// add_and_dot / l2_norm; used at step-35a.cc:1251-1252,1272-1273,1285-1286,
// 1506,1539-1540 and inside SolverGMRES / SolverCG).
//
// All operands are device pointers.  Scalar coefficients can come from device
// memory (DScal) so that Krylov recurrences run without a host round trip.
// Reductions are single-pass and deterministic: every block writes one
// partial, the last block to finish (ticket counter) adds the partials in a
// fixed order.
#pragma once
#include "common.cuh"

namespace s35 {

// value = mul * (ptr ? *ptr : 1); `recip` uses 1 / *ptr instead
struct DScal
{
  const double *ptr;
  double mul;
  int recip;
  DScal(double v) : ptr(nullptr), mul(v), recip(0) {}
  DScal(const double *p, double m = 1.0, int r = 0) : ptr(p), mul(m), recip(r) {}
};

// scratch for single-pass reductions; one per concurrently used stream
struct ReduceWs
{
  static constexpr int kMaxBlocks = 1024;
  DevBuf<double> partials;     // kMaxBlocks * kMaxResults
  DevBuf<unsigned int> ticket; // zero between launches
  static constexpr int kMaxResults = 32;
  void init(cudaStream_t s)
  {
    partials.alloc((size_t)kMaxBlocks * kMaxResults);
    ticket.alloc(1);
    ticket.zero(s);
  }
};

enum ReducePost
{
  kSum = 0,
  kSqrt = 1
};

// z = a x + b y (y may be null if b == 0; z may alias x or y)
void vec_lincomb(double *z, DScal a, const double *x, DScal b, const double *y, int64_t n, cudaStream_t s);
// z = a x + b y + c w
void vec_lincomb3(double *z, DScal a, const double *x, DScal b, const double *y, DScal c, const double *w, int64_t n,
                  cudaStream_t s);
// z = x .* y
void vec_multiply(double *z, const double *x, const double *y, int64_t n, cudaStream_t s);
void vec_copy(double *dst, const double *src, int64_t n, cudaStream_t s);
void vec_set(double *dst, double value, int64_t n, cudaStream_t s);
// *result = x . y   (or sqrt of it)
void vec_dot(double *result, const double *x, const double *y, int64_t n, ReducePost post, ReduceWs &ws, cudaStream_t s);
// y += a x; *result = y . w (w may alias y)   -- Vector::add_and_dot
void vec_add_and_dot(double *result, double *y, DScal a, const double *x, const double *w, int64_t n, ReducePost post,
                     ReduceWs &ws, cudaStream_t s);
// x += alpha d; g += alpha h; *result = g . g (or its square root)   -- the CG update pair (SolverCG)
void vec_cg_update(double *result, double *x, double *g, DScal alpha, const double *d, const double *h, int64_t n,
                   ReducePost post, ReduceWs &ws, cudaStream_t s);
// *result = (sum_i |values[idx[i]]|) / n   (mean |diagonal| for ConstraintMatrix::condense)
void vec_gather_abs_mean(double *result, const double *values, const int32_t *idx, int32_t n, ReduceWs &ws, cudaStream_t s);
// x += sum_i coef[i] v_i, i < m <= 32 (GMRES solution update); basis vectors are v0 + i * stride
void vec_multi_axpy(double *x, const double *v0, int64_t stride, const double *coef_host, int m, int64_t n, cudaStream_t s);
// scatter: x[idx[i]] = val ? val[i] : 0   (ConstraintMatrix::distribute, step-35a.cc:1322-1323,1500)
void vec_scatter_set(double *x, const int32_t *idx, const double *val, int32_t m, cudaStream_t s);
// out[0] = num[0] / den[0]; optionally also store num into save (CG's alpha / beta bookkeeping)
void scalar_divide(double *out, const double *num, const double *den, double *save_num, cudaStream_t s);

// bytes written by the L2 flush helper must exceed the 126 MB L2
void flush_l2(cudaStream_t s);

} // namespace s35
