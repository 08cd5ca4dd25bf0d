// Device kernels for the parts of diffusion_step / assemble_advection_term
// that touch matrix entries:
//  * assemble_advection_term + assemble_one_cell_of_advection +
//    copy_advection_local_to_global (step-35a.cc:1346-1453): element-coloured,
//    atomic-free quadrature assembly straight into the device CSR;
//  * vel_it_matrix[d].copy_from(vel_Advection) + ConstraintMatrix::condense
//    (step-35a.cc:1297,1311).
#pragma once
#include "sparse.cuh"
#include "vecops.cuh"

namespace s35 {

struct AdvectionAssembler
{
  int dim = 0;
  int32_t n_cells = 0;
  int dpc = 0;
  int n_colors = 0;
  std::vector<int32_t> color_offset; // cells of colour c: cell_list[color_offset[c] .. color_offset[c+1])
  DevBuf<int32_t> cell_list;
  DevBuf<double> cell_vertices; // [n_cells][2^dim][dim]
  DevBuf<int32_t> cell_dofs;    // [n_cells][dpc]
  // CSR position of local entry (i,j) of each cell: >= 0 add; -1 dropped (wall-constrained row or column);
  // <= -2: wall-constrained diagonal, add |a_ii| at position -(s+2)   
  DevBuf<int32_t> slots;        // [n_cells][dpc][dpc]
  DevBuf<double> tables;        // phi, reference gradients, Q1 gradients, weights
  size_t smem_bytes = 0;

  void create(int dim, int32_t n_cells, const double *cell_vertices, const int32_t *cell_dofs, const int32_t *cell_color,
              const DevMatrix &pattern, const std::vector<char> &wall_mask, cudaStream_t s);
  // values = 0; values += sum over cells of the local matrices
  void assemble(double *values, int64_t nnz, const double *const *u_star, double dt, double reynolds, cudaStream_t s);
  double algorithmic_bytes(int64_t nnz, int64_t n_vertices, int32_t n_vel) const
  {
    return 8.0 * nnz + 8.0 * dim * n_vertices + 4.0 * dpc * (double)n_cells + 8.0 * dim * n_vel;
  }
  double flops() const;
};

// Constraint sets without dependencies (x_c = g_c), prepared for condense on a fixed pattern.
struct CondensePlan
{
  int32_t n_constrained = 0;
  DevBuf<int32_t> dofs;     // constrained DoFs (ascending; on a partition: owned and ghost ones)
  int32_t n_row_dofs = 0;
  DevBuf<int32_t> row_dofs; // those of them that are rows of this rank's matrix
  DevBuf<double> values;    // their inhomogeneities
  // unconstrained rows that couple to a constrained column
  int32_t n_lift_rows = 0;
  DevBuf<int32_t> lift_row, lift_ptr; // rows and CSR-style offsets into lift_pos / lift_val
  DevBuf<int32_t> lift_pos;           // matrix positions of the entries (r, c), ascending within a row
  DevBuf<double> lift_val;            // g_c for each of those entries

  void create(const DevMatrix &pattern, int32_t n, const int32_t *dofs, const double *values, cudaStream_t s);
  // *avg_diag must hold mean |a_ii| of the uncondensed matrix (device scalar)
  void condense(const DevMatrix &pattern, double *values, double *rhs, const double *avg_diag, cudaStream_t s) const;
};

// *result = (sum_i |a_ii|) / n
void diag_abs_mean(const DevMatrix &pattern, const double *values, double *result, ReduceWs &ws, cudaStream_t s);

} // namespace s35
