#include "assembly.cuh"

#include "../host/fe.h"

#include <algorithm>

namespace s35 {

namespace {

constexpr int kAsmWarps = 8;

template <int DIM>
struct AsmDims
{
  static constexpr int N1 = 3;                           // Q2: 3 nodes and 3 Gauss points per direction
  static constexpr int DPC = DIM == 3 ? 27 : 9;          // dofs per cell
  static constexpr int NQ = DPC;                         // quadrature points per cell
  static constexpr int NV = 1 << DIM;                    // vertices per cell
  static constexpr int TW = DIM + 1;                     // per (q, j): scaled gradient (DIM) and advection/mass weight
  static constexpr int QS = (DPC * TW + 2 + 1) & ~1;     // padded stride between quadrature points (even)
  // block-wide tables
  static constexpr int OFF_PHI = 0;                      // [NQ][DPC]
  static constexpr int OFF_GREF = OFF_PHI + NQ * DPC;    // [NQ][DPC][DIM]
  static constexpr int OFF_DN = OFF_GREF + NQ * DPC * DIM; // [NQ][NV][DIM]
  static constexpr int OFF_W = OFF_DN + NQ * NV * DIM;   // [NQ]
  static constexpr int TABLE = (OFF_W + NQ + 1) & ~1;
  // per-warp scratch
  static constexpr int WARP_T = NQ * QS;                 // T[q][j][TW]
  static constexpr int WARP_U = (DIM * DPC + 1) & ~1;    // u*[d][j]
  static constexpr int WARP = WARP_T + WARP_U;
  static constexpr size_t SMEM = sizeof(double) * (TABLE + kAsmWarps * WARP);
};

// One warp per cell, cells of one colour per launch.
//   phase A (lane = quadrature point q): Jacobian of the d-linear map, physical gradients G_j(q), u*(q), div u*(q),
//            T[q][j] = { JxW/Re * G_j(q),  JxW * (u*(q).G_j(q) + (0.5 div u*(q) + 1.5/dt) phi_j(q)) }
//   phase B (lane = local row i): A(i,j) = sum_q  G_i(q) . T[q][j][0..DIM) + phi_i(q) * T[q][j][DIM]
//   scatter: row i adds its 3^DIM entries at precomputed CSR positions; cells of a colour share no DoF, so plain
//            read-modify-write is race free and the summation order is fixed by the colour order.
template <int DIM>
__global__ void __launch_bounds__(kAsmWarps * 32)
k_assemble_advection(int32_t n_cells_in_color, const int32_t *__restrict__ cell_list, const double *__restrict__ cell_vertices,
                     const int32_t *__restrict__ cell_dofs, const int32_t *__restrict__ slots,
                     const double *__restrict__ u0, const double *__restrict__ u1, const double *__restrict__ u2,
                     const double *__restrict__ tables, double inv_re, double time_step_term, double *values)
{
  using D = AsmDims<DIM>;
  extern __shared__ __align__(16) double smem[];
  for (int i = threadIdx.x; i < D::TABLE; i += blockDim.x)
    smem[i] = tables[i];
  __syncthreads();
  const double *s_phi = smem + D::OFF_PHI, *s_gref = smem + D::OFF_GREF, *s_dn = smem + D::OFF_DN, *s_w = smem + D::OFF_W;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int ci = blockIdx.x * kAsmWarps + warp;
  if (ci >= n_cells_in_color)
    return;
  const int32_t cell = cell_list[ci];
  double *T = smem + D::TABLE + warp * D::WARP;
  double *us = T + D::WARP_T;

  const int li = lane < D::DPC ? lane : D::DPC - 1; // clamped: idle lanes shadow the last one and never store
  const bool active = lane < D::DPC;
  const int32_t dof = cell_dofs[(size_t)cell * D::DPC + li];
  if (active)
    {
      us[0 * D::DPC + li] = u0[dof];
      us[1 * D::DPC + li] = u1[dof];
      if (DIM == 3)
        us[2 * D::DPC + li] = u2[dof];
    }
  const double xv = lane < D::NV * DIM ? cell_vertices[(size_t)cell * D::NV * DIM + lane] : 0.0;
  __syncwarp();

  // ---- phase A: lane = quadrature point
  double scale; // Re / JxW(q)
  {
    const int q = li;
    double J[DIM][DIM];
#pragma unroll
    for (int k = 0; k < DIM; ++k)
#pragma unroll
      for (int d = 0; d < DIM; ++d)
        J[k][d] = 0.0;
#pragma unroll
    for (int v = 0; v < D::NV; ++v)
#pragma unroll
      for (int k = 0; k < DIM; ++k)
        {
          const double xk = __shfl_sync(0xffffffffu, xv, v * DIM + k);
#pragma unroll
          for (int d = 0; d < DIM; ++d)
            J[k][d] += xk * s_dn[(q * D::NV + v) * DIM + d];
        }
    double iJT[DIM][DIM], det;
    if (DIM == 2)
      {
        det = J[0][0] * J[1][1] - J[0][1] * J[1][0];
        const double id = 1.0 / det;
        // iJT[k][d] = inv(J)[d][k]
        iJT[0][0] = J[1][1] * id;
        iJT[1][0] = -J[0][1] * id;
        iJT[0][1] = -J[1][0] * id;
        iJT[1][1] = J[0][0] * id;
      }
    else
      {
        const double c00 = J[1][1] * J[2 % DIM][2 % DIM] - J[1][2 % DIM] * J[2 % DIM][1];
        const double c01 = J[1][2 % DIM] * J[2 % DIM][0] - J[1][0] * J[2 % DIM][2 % DIM];
        const double c02 = J[1][0] * J[2 % DIM][1] - J[1][1] * J[2 % DIM][0];
        det = J[0][0] * c00 + J[0][1] * c01 + J[0][2 % DIM] * c02;
        const double id = 1.0 / det;
        // inv[d][k]; iJT[k][d] = inv[d][k]
        iJT[0][0] = c00 * id;
        iJT[0][1] = c01 * id;
        iJT[0][2 % DIM] = c02 * id;
        iJT[1][0] = (J[0][2 % DIM] * J[2 % DIM][1] - J[0][1] * J[2 % DIM][2 % DIM]) * id;
        iJT[1][1] = (J[0][0] * J[2 % DIM][2 % DIM] - J[0][2 % DIM] * J[2 % DIM][0]) * id;
        iJT[1][2 % DIM] = (J[0][1] * J[2 % DIM][0] - J[0][0] * J[2 % DIM][1]) * id;
        iJT[2 % DIM][0] = (J[0][1] * J[1][2 % DIM] - J[0][2 % DIM] * J[1][1]) * id;
        iJT[2 % DIM][1] = (J[0][2 % DIM] * J[1][0] - J[0][0] * J[1][2 % DIM]) * id;
        iJT[2 % DIM][2 % DIM] = (J[0][0] * J[1][1] - J[0][1] * J[1][0]) * id;
      }
    const double jxw = det * s_w[q];
    double uq[DIM];
#pragma unroll
    for (int d = 0; d < DIM; ++d)
      uq[d] = 0.0;
    for (int j = 0; j < D::DPC; ++j)
      {
        const double ph = s_phi[q * D::DPC + j];
#pragma unroll
        for (int d = 0; d < DIM; ++d)
          uq[d] += us[d * D::DPC + j] * ph;
      }
    double div = 0.0;
    double *Tq = T + q * D::QS;
    for (int j = 0; j < D::DPC; ++j)
      {
        double g[DIM];
#pragma unroll
        for (int k = 0; k < DIM; ++k)
          {
            double sum = 0.0;
#pragma unroll
            for (int d = 0; d < DIM; ++d)
              sum += iJT[k][d] * s_gref[(q * D::DPC + j) * DIM + d];
            g[k] = sum;
            div += us[k * D::DPC + j] * sum;
          }
        if (active)
#pragma unroll
          for (int k = 0; k < DIM; ++k)
            Tq[j * D::TW + k] = g[k];
      }
    const double cq = 0.5 * div + time_step_term;
    const double sc = jxw * inv_re;
    if (active)
      for (int j = 0; j < D::DPC; ++j)
        {
          double adv = cq * s_phi[q * D::DPC + j];
#pragma unroll
          for (int k = 0; k < DIM; ++k)
            {
              const double g = Tq[j * D::TW + k];
              adv += uq[k] * g;
              Tq[j * D::TW + k] = g * sc;
            }
          Tq[j * D::TW + DIM] = jxw * adv;
        }
    scale = 1.0 / sc;
  }
  __syncwarp();

  // ---- phase B: lane = local row
  double acc[D::DPC];
#pragma unroll
  for (int j = 0; j < D::DPC; ++j)
    acc[j] = 0.0;
  for (int q = 0; q < D::NQ; ++q)
    {
      const double sq = __shfl_sync(0xffffffffu, scale, q);
      const double *Tq = T + q * D::QS;
      double gi[DIM];
#pragma unroll
      for (int k = 0; k < DIM; ++k)
        gi[k] = Tq[li * D::TW + k] * sq;
      const double ph = s_phi[q * D::DPC + li];
      if (DIM == 3)
        {
          const double2 *Tq2 = reinterpret_cast<const double2 *>(Tq);
#pragma unroll
          for (int j = 0; j < D::DPC; ++j)
            {
              const double2 a = Tq2[j * 2], b = Tq2[j * 2 + 1];
              acc[j] += gi[0] * a.x + gi[1] * a.y + gi[2 % DIM] * b.x + ph * b.y;
            }
        }
      else
        {
#pragma unroll
          for (int j = 0; j < D::DPC; ++j)
            acc[j] += gi[0] * Tq[j * D::TW] + gi[1] * Tq[j * D::TW + 1] + ph * Tq[j * D::TW + 2];
        }
    }

  // ---- scatter (ConstraintMatrix::distribute_local_to_global with the wall constraints)
  double diag = 0.0;
#pragma unroll
  for (int j = 0; j < D::DPC; ++j)
    if (j == li)
      diag = acc[j];
  double diag_sum = active ? fabs(diag) : 0.0;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1)
    diag_sum += __shfl_xor_sync(0xffffffffu, diag_sum, o);
  if (!active)
    return;
  const int32_t *srow = slots + ((size_t)cell * D::DPC + li) * D::DPC;
#pragma unroll
  for (int j = 0; j < D::DPC; ++j)
    {
      const int32_t sl = srow[j];
      if (sl >= 0)
        values[sl] += acc[j];
      else if (sl <= -2)
        {
          const double a = fabs(acc[j]);
          values[-(sl + 2)] += (a != 0.0 ? a : diag_sum / D::DPC);
        }
    }
}

// unconstrained rows: rhs_r -= a_rc g_c ; a_rc = 0 for the listed entries (ascending column order)
__global__ void k_condense_lift(int32_t n_rows, const int32_t *__restrict__ lift_row, const int32_t *__restrict__ lift_ptr,
                                const int32_t *__restrict__ lift_pos, const double *__restrict__ lift_val, double *values,
                                double *rhs)
{
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n_rows)
    return;
  const int r = lift_row[i];
  double b = rhs[r];
  for (int k = lift_ptr[i]; k < lift_ptr[i + 1]; ++k)
    {
      const int pos = lift_pos[k];
      b -= values[pos] * lift_val[k];
      values[pos] = 0.0;
    }
  rhs[r] = b;
}

// constrained rows: everything zero except the diagonal, which gets the mean |diagonal|; rhs_c = 0
__global__ void k_condense_rows(int32_t n_constrained, const int32_t *__restrict__ dofs, const int32_t *__restrict__ rowptr,
                                const int32_t *__restrict__ diag, double *values, double *rhs, const double *avg_diag)
{
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n_constrained)
    return;
  const int r = dofs[i];
  const int d = diag[r];
  const double avg = *avg_diag;
  for (int k = rowptr[r]; k < rowptr[r + 1]; ++k)
    values[k] = (k == d) ? avg : 0.0;
  rhs[r] = 0.0;
}

} // namespace

void AdvectionAssembler::create(int dim_, int32_t n_cells_, const double *h_vertices, const int32_t *h_cell_dofs,
                                const int32_t *h_color, const DevMatrix &P, const std::vector<char> &wall_mask, cudaStream_t s)
{
  S35_REQUIRE(dim_ == 2 || dim_ == 3, "assembler: dim must be 2 or 3");
  dim = dim_;
  n_cells = n_cells_;
  dpc = dim == 3 ? 27 : 9;
  const int nv = 1 << dim;
  // cells grouped by colour
  int max_color = -1;
  for (int32_t c = 0; c < n_cells; ++c)
    {
      S35_REQUIRE(h_color[c] >= 0 && h_color[c] < 64, "assembler: cell colour out of range");
      max_color = std::max(max_color, (int)h_color[c]);
    }
  n_colors = max_color + 1;
  color_offset.assign(n_colors + 1, 0);
  for (int32_t c = 0; c < n_cells; ++c)
    color_offset[h_color[c] + 1]++;
  for (int c = 0; c < n_colors; ++c)
    color_offset[c + 1] += color_offset[c];
  std::vector<int32_t> list(n_cells), fill(color_offset.begin(), color_offset.end() - 1);
  for (int32_t c = 0; c < n_cells; ++c)
    list[fill[h_color[c]]++] = c;
  // verify the colouring: no DoF may appear twice within one colour
  {
    std::vector<int32_t> last_color(P.n_cols, -1);
    for (int col = 0; col < n_colors; ++col)
      for (int32_t k = color_offset[col]; k < color_offset[col + 1]; ++k)
        for (int i = 0; i < dpc; ++i)
          {
            const int32_t d = h_cell_dofs[(size_t)list[k] * dpc + i];
            S35_REQUIRE(d >= 0 && d < P.n_cols, "assembler: cell DoF index out of range");
            S35_REQUIRE(last_color[d] != col, "assembler: two cells of one colour share a DoF");
            last_color[d] = col;
          }
  }
  cell_list.upload(list, s);
  cell_vertices.upload(h_vertices, (size_t)n_cells * nv * dim, s);
  cell_dofs.upload(h_cell_dofs, (size_t)n_cells * dpc, s);

  // CSR positions of the local entries
  std::vector<int32_t> h_slots((size_t)n_cells * dpc * dpc);
  bool missing = false;
#pragma omp parallel for schedule(static) reduction(|| : missing)
  for (int32_t c = 0; c < n_cells; ++c)
    {
      const int32_t *cd = h_cell_dofs + (size_t)c * dpc;
      int32_t *sl = h_slots.data() + (size_t)c * dpc * dpc;
      for (int i = 0; i < dpc; ++i)
        {
          const int32_t r = cd[i];
          if (r >= P.n_rows)
            {
              // a ghost row: its owner assembles it
              for (int j = 0; j < dpc; ++j)
                sl[i * dpc + j] = -1;
              continue;
            }
          const int32_t *b = P.h_col.data() + P.h_rowptr[r], *e = P.h_col.data() + P.h_rowptr[r + 1];
          for (int j = 0; j < dpc; ++j)
            {
              const int32_t cc = cd[j];
              int32_t v = -1;
              if (wall_mask[r] || wall_mask[cc])
                {
                  if (i == j)
                    {
                      const int32_t *p = std::lower_bound(b, e, cc);
                      if (p == e || *p != cc)
                        missing = true;
                      else
                        v = -((int32_t)(p - P.h_col.data()) + 2);
                    }
                }
              else
                {
                  const int32_t *p = std::lower_bound(b, e, cc);
                  if (p == e || *p != cc)
                    missing = true;
                  else
                    v = (int32_t)(p - P.h_col.data());
                }
              sl[i * dpc + j] = v;
            }
        }
    }
  S35_REQUIRE(!missing, "assembler: the sparsity pattern lacks an entry the cells couple");
  slots.upload(h_slots, s);

  // reference-element tables: Q2 at the 3-point Gauss rule, Q1 gradients for the mapping
  const ShapeTable q2(dim, 2, 3), q1(dim, 1, 3);
  std::vector<double> t;
  t.insert(t.end(), q2.value.begin(), q2.value.end());
  t.insert(t.end(), q2.grad.begin(), q2.grad.end());
  t.insert(t.end(), q1.grad.begin(), q1.grad.end());
  t.insert(t.end(), q2.weight.begin(), q2.weight.end());
  const size_t table_len = dim == 3 ? AsmDims<3>::TABLE : AsmDims<2>::TABLE;
  t.resize(table_len, 0.0);
  tables.upload(t, s);
  smem_bytes = dim == 3 ? AsmDims<3>::SMEM : AsmDims<2>::SMEM;
  if (dim == 3)
    S35_CUDA(cudaFuncSetAttribute(k_assemble_advection<3>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_bytes));
  else
    S35_CUDA(cudaFuncSetAttribute(k_assemble_advection<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_bytes));
}

void AdvectionAssembler::assemble(double *values, int64_t nnz, const double *const *u_star, double dt, double reynolds,
                                  cudaStream_t s)
{
  S35_CUDA(cudaMemsetAsync(values, 0, (size_t)nnz * sizeof(double), s)); // vel_Advection = 0
  const double inv_re = 1.0 / reynolds, ts = 1.5 / dt;
  for (int c = 0; c < n_colors; ++c)
    {
      const int32_t count = color_offset[c + 1] - color_offset[c];
      if (count == 0)
        continue;
      const int grid = (count + kAsmWarps - 1) / kAsmWarps;
      if (dim == 3)
        S35_LAUNCH(k_assemble_advection<3>, grid, kAsmWarps * 32, smem_bytes, s, count, cell_list.ptr + color_offset[c],
                   cell_vertices.ptr, cell_dofs.ptr, slots.ptr, u_star[0], u_star[1], u_star[2], tables.ptr, inv_re, ts, values);
      else
        S35_LAUNCH(k_assemble_advection<2>, grid, kAsmWarps * 32, smem_bytes, s, count, cell_list.ptr + color_offset[c],
                   cell_vertices.ptr, cell_dofs.ptr, slots.ptr, u_star[0], u_star[1], u_star[1], tables.ptr, inv_re, ts, values);
    }
}

double AdvectionAssembler::flops() const
{
  // phase B: 2 * (dim + 1) flops per (q, i, j)
  return (double)n_cells * dpc * dpc * dpc * 2.0 * (dim + 1);
}

void CondensePlan::create(const DevMatrix &P, int32_t n, const int32_t *h_dofs, const double *h_values, cudaStream_t s)
{
  n_constrained = n;
  std::vector<double> g(P.n_cols, 0.0);
  std::vector<char> mask(P.n_cols, 0);
  for (int32_t i = 0; i < n; ++i)
    {
      S35_REQUIRE(h_dofs[i] >= 0 && h_dofs[i] < P.n_cols, "condense: constrained DoF out of range");
      mask[h_dofs[i]] = 1;
      g[h_dofs[i]] = h_values ? h_values[i] : 0.0;
    }
  dofs.upload(h_dofs, n, s);
  std::vector<double> vals(n);
  for (int32_t i = 0; i < n; ++i)
    vals[i] = g[h_dofs[i]];
  values.upload(vals, s);
  // constrained rows this rank owns (ghost DoFs only appear as columns)
  std::vector<int32_t> own_rows;
  for (int32_t i = 0; i < n; ++i)
    if (h_dofs[i] < P.n_rows)
      own_rows.push_back(h_dofs[i]);
  n_row_dofs = (int32_t)own_rows.size();
  row_dofs.upload(own_rows, s);
  std::vector<int32_t> rows, ptr(1, 0), pos;
  std::vector<double> lval;
  for (int32_t r = 0; r < P.n_rows; ++r)
    {
      if (mask[r])
        continue;
      bool any = false;
      for (int32_t k = P.h_rowptr[r]; k < P.h_rowptr[r + 1]; ++k)
        if (mask[P.h_col[k]])
          {
            pos.push_back(k);
            lval.push_back(g[P.h_col[k]]);
            any = true;
          }
      if (any)
        {
          rows.push_back(r);
          ptr.push_back((int32_t)pos.size());
        }
    }
  n_lift_rows = (int32_t)rows.size();
  lift_row.upload(rows, s);
  lift_ptr.upload(ptr, s);
  lift_pos.upload(pos, s);
  lift_val.upload(lval, s);
}

void CondensePlan::condense(const DevMatrix &P, double *vals, double *rhs, const double *avg_diag, cudaStream_t s) const
{
  if (n_lift_rows)
    S35_LAUNCH(k_condense_lift, (n_lift_rows + 127) / 128, 128, 0, s, n_lift_rows, lift_row.ptr, lift_ptr.ptr, lift_pos.ptr,
               lift_val.ptr, vals, rhs);
  if (n_row_dofs)
    S35_LAUNCH(k_condense_rows, (n_row_dofs + 127) / 128, 128, 0, s, n_row_dofs, row_dofs.ptr, P.rowptr.ptr, P.diag.ptr, vals,
               rhs, avg_diag);
}

void diag_abs_mean(const DevMatrix &P, const double *values, double *result, ReduceWs &ws, cudaStream_t s)
{
  vec_gather_abs_mean(result, values, P.diag.ptr, P.n_rows, ws, s);
}

} // namespace s35
