// Host scaffolding: DoF numbering, sparsity patterns, constraints and the
// one-off operators of step-35a, i.e. what the reference obtains from deal.II
// in create_triangulation_and_dofs / initialize_* (step-35a.cc:738-1164).
// Everything here runs once at set-up; its outputs are the inputs of the
// per-time-step device path.
#pragma once
#include "fe.h"
#include "mesh.h"

#include <cstdint>
#include <memory>
#include <unordered_map>
#include <vector>

namespace s35 {

// CSR pattern, columns ascending within each row.
struct Csr
{
  int32_t n_rows = 0, n_cols = 0;
  std::vector<int32_t> rowptr; // n_rows + 1
  std::vector<int32_t> col;    // nnz
  int64_t nnz() const { return (int64_t)col.size(); }
  // position of (r, c) or -1
  int32_t find(int32_t r, int32_t c) const;
};

struct DofMap
{
  int dim = 0, degree = 0, dpc = 0;
  int32_t n_dofs = 0;
  std::vector<int32_t> cell_dofs; // n_cells * dpc, lexicographic local order
  std::vector<double> support;    // n_dofs * dim support points
  // DoF of the mesh entity spanned by a sorted vertex tuple (vertex, edge, quad)
  struct Key
  {
    int32_t v[4];
    bool operator==(const Key &o) const { return v[0] == o.v[0] && v[1] == o.v[1] && v[2] == o.v[2] && v[3] == o.v[3]; }
  };
  struct KeyHash
  {
    size_t operator()(const Key &k) const;
  };
  std::unordered_map<Key, int32_t, KeyHash> entity_dof;
};

// DoFHandler::distribute_dofs for FE_Q(degree), degree 1 or 2 (step-35a.cc:764,766).
DofMap distribute_dofs(const Mesh &mesh, int degree);

// DoFRenumbering::boost::Cuthill_McKee (step-35a.cc:765,767).  deal.II writes
// boost's ordering through a reverse iterator, i.e. the default is a reversed
// Cuthill-McKee numbering; `reversed` selects that.  Start vertex: a
// pseudo-peripheral node (George-Liu); neighbours are visited by increasing
// degree, ties by index.  Returns new_index[old_index].
std::vector<int32_t> cuthill_mckee(const Csr &graph, bool reversed = true);
void renumber_dofs(DofMap &dofs, const std::vector<int32_t> &new_index);

// DoFTools::make_sparsity_pattern.  With `constrained` given (square case) the
// pattern is the keep_constrained_dofs=false one: constrained rows keep only
// their diagonal and no row couples to a constrained column
// (step-35a.cc:972-980, 996-1004).
Csr make_sparsity_pattern(const DofMap &rows, const DofMap &cols, int64_t n_cells,
                          const std::vector<char> *constrained = nullptr);

// DoFs with support on boundary faces of the given indicator, ascending
// (VectorTools::interpolate_boundary_values, step-35a.cc:917-957).
std::vector<int32_t> boundary_dofs(const Mesh &mesh, const DofMap &dofs, int boundary_id);

// Greedy colouring of cells so that cells of one colour share no vertex (and
// hence no DoF).  Returns the colour of each cell; n_colors is set.
std::vector<int32_t> color_cells(const Mesh &mesh, int &n_colors);

// Geometry of the d-linear (Q1) mapping at the points of a tensor Gauss rule
// (FEValues::reinit without a mapping argument, step-35a.cc:1374).
struct CellGeometry
{
  int dim, nqp;
  ShapeTable q1;
  std::vector<double> JxW;   // [nqp]
  std::vector<double> invJT; // [nqp][dim][dim]; physical grad = invJT * reference grad
  CellGeometry(int dim, int n_q_1d);
  void reinit(const Mesh &mesh, int64_t cell);
};

// Everything the per-step path consumes, as flat host arrays.
struct Discretization
{
  int dim = 0;
  Mesh mesh;
  DofMap vel, pres; // Q2 / Q1
  double width = 0, height = 0;

  std::vector<int32_t> cell_color;
  int n_colors = 0;

  // constraints (step-35a.cc:903-958)
  std::vector<int32_t> wall_dofs;            // ids 1 and 4, value 0
  std::vector<int32_t> flow_dofs[3];         // id 2 (+ id 3 for d >= 1)
  std::vector<double> flow_vals[3];
  std::vector<int32_t> proj_dofs;            // pressure DoFs on id 3, value 0

  // patterns and one-off operators (step-35a.cc:960-1164)
  Csr vel_pattern_constrained;   // vel_it_matrix[d], vel_Advection
  Csr vel_pattern_unconstrained; // vel_Mass
  Csr grad_pattern;              // pres_Diff[d] (Nv x Np)
  Csr proj_pattern;              // pres_Laplace
  Csr pres_pattern;              // pres_Mass
  std::vector<double> vel_mass, pres_laplace, pres_mass;
  std::vector<double> pres_diff[3];

  // initial conditions (step-35a.cc:834-854)
  std::vector<double> pres_n0, pres_nm1_0;
};

struct DiscretizationOptions
{
  int n_refines = 0;
  int pressure_degree = 1;
  double dt = 5e-4;         // only used for the initial pressure at t0 + dt (which has no time dependence)
  bool renumber = true;     // Cuthill-McKee
};

// Inflow profile EquationData::Velocity::value (step-35a.cc:364-387), component 0.
double inflow_velocity(int dim, const double *p, double width, double height);

std::unique_ptr<Discretization> build_discretization(const Mesh &coarse, const DiscretizationOptions &opt);

} // namespace s35
