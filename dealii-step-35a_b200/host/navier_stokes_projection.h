// Host mirror of the reference's NavierStokesProjection<dim> (step-35a.cc:439-686):
// same constructor and run() signature, same protected step methods, same
// stdout lines and timer section names.  The method bodies are calls through
// the C ABI of include/step35a_b200.h; all state vectors and matrices of the
// time loop live on the device.
#pragma once
#include "../../include/step35a_host.h"
#include "discretization.h"
#include "parameters.h"

#include <memory>
#include <string>

namespace Step35 {

// flat-array view of a Discretization in the layout of the C ABI
struct ProblemView
{
  s35_problem problem;
  std::vector<double> cell_vertices; // [n_cells][2^dim][dim]
  explicit ProblemView(const s35::Discretization &D, const RunTimeParameters::Data_Storage &data);
};

template <int dim>
class NavierStokesProjection
{
public:
  // mesh_file: the reference hard-wires "cylinderextruded.msh" (step-35a.cc:748); empty selects that
  // name for dim == 3 and "nsbench2.inp" for dim == 2.
  NavierStokesProjection(const RunTimeParameters::Data_Storage &data, const std::string &mesh_file = "");
  ~NavierStokesProjection();

  void run(const bool verbose = false, const unsigned int n_plots = 10);

  // access for tests, benchmarks and the C wrapper
  s35_ctx *context() { return ctx; }
  const s35::Discretization &discretization() const { return *disc; }
  const ProblemView &problem_view() const { return *view; }
  unsigned int n_time_steps() const;
  double setup_seconds = 0.0;
  std::string output_directory = ".";

protected:
  RunTimeParameters::MethodFormulation type;
  const unsigned int deg;
  const double dt;
  const double t_0, T, Re;
  std::string load_checkpoint_file_name;
  std::string save_checkpoint_file_name;

  std::unique_ptr<s35::Discretization> disc;
  std::unique_ptr<ProblemView> view;
  s35_ctx *ctx = nullptr;

  double height = 0, width = 0;
  bool write_mesh = true;                  // step-35a.cc: mesh.h5 is written with the first output only
  std::vector<std::string> xdmf_entries;   // one grid per output step (step-35a.cc:1652-1656)

  void create_triangulation_and_dofs(const unsigned int n_refines, const std::string &mesh_file);
  void initialize();
  void load_checkpoint();
  void save_checkpoint();
  void interpolate_velocity();
  void diffusion_step(const bool reinit_prec);
  void projection_step();
  void update_pressure();
  void assemble_advection_term();
  void output_results(const unsigned int step, const double time);

private:
  unsigned int vel_max_its, vel_Krylov_size, vel_off_diagonals, vel_update_prec;
  double vel_eps, vel_diag_strength;
  RunTimeParameters::Data_Storage parameters;
  void print_timer_summary() const;
  static void check(s35_status st);
};

} // namespace Step35
