// Mesh partitioning for the multi-GPU path (new capability; the reference is a single process and
// replicates everything, step-35a.cc:1016,1471,1526).
//
// Cells are split among the ranks by recursive coordinate bisection of their centroids.  A DoF is owned
// by the lowest rank among the cells it belongs to.  Rank k works on its owned DoFs plus one layer of
// ghosts: the DoFs of every cell that contains an owned velocity DoF ("halo cells").  With that layer
//   * every matrix row of an owned DoF is complete (all its columns are local),
//   * the advection assembly of owned rows needs no communication (halo cells are recomputed locally),
//   * the only data-path communication is the ghost-value exchange before a mat-vec and the sum
//     reductions of the Krylov dot products.
// Local numbering: owned DoFs first, in ascending global (Cuthill-McKee) order, then the ghosts grouped by
// owning rank, ascending global index inside a group -- so a neighbour's ghost values arrive as one
// contiguous segment.
#pragma once
#include "../../include/step35a_b200.h"
#include "discretization.h"

#include <memory>
#include <vector>

namespace s35 {

struct HaloPlan
{
  std::vector<int32_t> neighbor_rank, send_ptr, send_idx, recv_ptr;
  s35_halo view() const
  {
    s35_halo h;
    h.n_neighbors = (int32_t)neighbor_rank.size();
    h.neighbor_rank = neighbor_rank.data();
    h.send_ptr = send_ptr.data();
    h.send_idx = send_idx.data();
    h.recv_ptr = recv_ptr.data();
    return h;
  }
};

// rank of every cell: recursive coordinate bisection along the longest extent, equal cell counts
std::vector<int32_t> partition_cells(const Mesh &mesh, int n_ranks);

struct LocalProblem
{
  int rank = 0, n_ranks = 1;
  int32_t n_vel_owned = 0, n_pres_owned = 0;
  std::vector<int32_t> vel_global, pres_global; // local -> global
  HaloPlan vel_halo, pres_halo;
  std::vector<int32_t> cells;                   // global ids of the halo cells
  std::vector<double> cell_vertices;
  std::vector<int32_t> vel_cell_dofs, pres_cell_dofs, cell_color;
  Csr velc, velm, grad, gradT, plap, pmass;     // rows: owned DoFs; columns: local numbering, ascending
  std::vector<double> velm_val, grad_val[3], gradT_val[3], plap_val, pmass_val;
  std::vector<int32_t> wall, flow_dofs[3], proj;
  std::vector<double> flow_vals[3];
  std::vector<double> initial_pressure;         // local
  s35_dist dist;
  s35_problem problem;
};

std::unique_ptr<LocalProblem> extract_local_problem(const Discretization &D, const std::vector<int32_t> &cell_rank, int rank,
                                                    int n_ranks, const s35_problem &global);

} // namespace s35
