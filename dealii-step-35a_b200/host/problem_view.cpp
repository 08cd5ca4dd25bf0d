// ProblemView: the flat-array view of a Discretization that crosses the C ABI (no device code: part of the host-only
// library too, so that set-up and the CPU oracle can run without loading any CUDA library).
#include "navier_stokes_projection.h"

namespace Step35 {

ProblemView::ProblemView(const s35::Discretization &D, const RunTimeParameters::Data_Storage &data)
{
  const int dim = D.dim, vpc = 1 << dim;
  const int64_t n_cells = D.mesh.n_cells();
  cell_vertices.resize((size_t)n_cells * vpc * dim);
  for (int64_t c = 0; c < n_cells; ++c)
    for (int v = 0; v < vpc; ++v)
      for (int k = 0; k < dim; ++k)
        cell_vertices[((size_t)c * vpc + v) * dim + k] = D.mesh.vertices[(size_t)D.mesh.cells[c * vpc + v] * dim + k];
  s35_problem &p = problem;
  p = s35_problem{};
  p.dim = dim;
  p.n_cells = (int32_t)n_cells;
  p.n_vel = D.vel.n_dofs;
  p.n_pres = D.pres.n_dofs;
  p.cell_vertices = cell_vertices.data();
  p.vel_cell_dofs = D.vel.cell_dofs.data();
  p.cell_color = D.cell_color.data();
  p.velc_rowptr = D.vel_pattern_constrained.rowptr.data();
  p.velc_col = D.vel_pattern_constrained.col.data();
  p.velm_rowptr = D.vel_pattern_unconstrained.rowptr.data();
  p.velm_col = D.vel_pattern_unconstrained.col.data();
  p.velm_val = D.vel_mass.data();
  p.grad_rowptr = D.grad_pattern.rowptr.data();
  p.grad_col = D.grad_pattern.col.data();
  for (int d = 0; d < dim; ++d)
    p.grad_val[d] = D.pres_diff[d].data();
  p.plap_rowptr = D.proj_pattern.rowptr.data();
  p.plap_col = D.proj_pattern.col.data();
  p.plap_val = D.pres_laplace.data();
  p.pmass_rowptr = D.pres_pattern.rowptr.data();
  p.pmass_col = D.pres_pattern.col.data();
  p.pmass_val = D.pres_mass.data();
  p.n_wall = (int32_t)D.wall_dofs.size();
  p.wall_dofs = D.wall_dofs.data();
  for (int d = 0; d < dim; ++d)
    {
      p.n_flow[d] = (int32_t)D.flow_dofs[d].size();
      p.flow_dofs[d] = D.flow_dofs[d].data();
      p.flow_vals[d] = D.flow_vals[d].data();
    }
  p.n_proj = (int32_t)D.proj_dofs.size();
  p.proj_dofs = D.proj_dofs.data();
  p.dt = data.dt;
  p.reynolds = data.Reynolds;
  p.vel_eps = data.vel_eps;
  p.vel_diag_strength = data.vel_diag_strength;
  p.vel_max_iterations = (int32_t)data.vel_max_iterations;
  p.vel_krylov_size = (int32_t)data.vel_Krylov_size;
  p.vel_off_diagonals = (int32_t)data.vel_off_diagonals;
  p.vel_update_prec = (int32_t)data.vel_update_prec;
  p.rotational = data.form == RunTimeParameters::METHOD_ROTATIONAL ? 1 : 0;
}

} // namespace Step35
