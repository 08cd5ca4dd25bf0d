// Device-resident CSR matrix and SpMV kernels (deal.II SparseMatrix<double>:
// vmult / vmult_add / Tvmult / Tvmult_add at step-35a.cc:1287,1289,1462,1464 and
// the mat-vecs inside SolverGMRES / SolverCG).
#pragma once
#include "common.cuh"

#include <memory>

namespace s35 {

struct DevMatrix
{
  int32_t n_rows = 0, n_cols = 0;
  int64_t nnz = 0;
  DevBuf<int32_t> rowptr, col;
  DevBuf<double> val;
  DevBuf<int32_t> diag; // position of the diagonal entry of each row (square matrices only)
  int lanes = 8;        // threads cooperating on one row in SpMV
  // host copy of the pattern (symbolic phases: transpose, ILU analysis, condense lists)
  std::vector<int32_t> h_rowptr, h_col;

  // explicit transpose, built on first use of Tvmult
  struct Transposed
  {
    std::unique_ptr<DevMatrix> T;
    DevBuf<int32_t> perm; // T.val[k] = val[perm[k]]
    bool values_current = false;
  };
  mutable std::unique_ptr<Transposed> transposed;

  void create(int32_t n_rows, int32_t n_cols, const int32_t *rowptr, const int32_t *col, const double *val,
              cudaStream_t s);
  void set_values(const double *host_val, cudaStream_t s);
  void values_changed() const
  {
    if (transposed)
      transposed->values_current = false;
  }
  const DevMatrix &transpose(cudaStream_t s) const;
  double spmv_bytes() const { return 12.0 * nnz + 4.0 * (n_rows + 1) + 8.0 * n_rows + 8.0 * n_cols; }
};

// y = [y +] A x, all device pointers
void spmv(const DevMatrix &A, const double *x, double *y, bool add, cudaStream_t s);
// same pattern, explicit value array (vel_it_matrix[d] share vel_Advection's pattern)
void spmv_values(const DevMatrix &pattern, const double *val, const double *x, double *y, bool add, cudaStream_t s);
// y = [y +] sum_{d<m} A_d x_d for m <= 3 value arrays on one pattern (projection rhs, step-35a.cc:1462-1464)
void spmv_multi(const DevMatrix &pattern, int m, const double *const *val, const double *const *x, double *y, bool add,
                cudaStream_t s);
// dst[k] = src[perm[k]]
void gather_values(double *dst, const double *src, const int32_t *perm, int64_t n, cudaStream_t s);

} // namespace s35
