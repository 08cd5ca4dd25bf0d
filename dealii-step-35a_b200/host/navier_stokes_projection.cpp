#include "navier_stokes_projection.h"

#include "h5lite.h"

#include <chrono>
#include <cstdlib>
#include <cstdio>
#include <iomanip>
#include <iostream>
#include <sstream>
#include <stdexcept>

namespace Step35 {

template <int dim>
void NavierStokesProjection<dim>::check(s35_status st)
{
  if (st != S35_OK)
    throw std::runtime_error(s35_last_error());
}

template <int dim>
NavierStokesProjection<dim>::NavierStokesProjection(const RunTimeParameters::Data_Storage &data, const std::string &mesh_file)
  : type(data.form), deg(data.pressure_degree), dt(data.dt), t_0(data.initial_time), T(data.final_time), Re(data.Reynolds),
    load_checkpoint_file_name(data.load_checkpoint_file_name), save_checkpoint_file_name(data.save_checkpoint_file_name),
    vel_max_its(data.vel_max_iterations), vel_Krylov_size(data.vel_Krylov_size), vel_off_diagonals(data.vel_off_diagonals),
    vel_update_prec(data.vel_update_prec), vel_eps(data.vel_eps), vel_diag_strength(data.vel_diag_strength), parameters(data)
{
  if (deg < 1)
    std::cout << " WARNING: The chosen pair of finite element spaces is not stable." << std::endl
              << " The obtained results will be nonsense" << std::endl;
  // AssertThrow at step-35a.cc:731
  if ((dt <= 0.) || (dt > .5 * T))
    {
      std::ostringstream msg;
      msg << " The time step " << dt << " is out of range." << std::endl << " The permitted range is (0," << .5 * T << "]";
      throw std::runtime_error(msg.str());
    }
  create_triangulation_and_dofs(data.n_global_refines, mesh_file);
  initialize();
}

template <int dim>
NavierStokesProjection<dim>::~NavierStokesProjection()
{
  s35_ctx_destroy(ctx);
}

template <int dim>
void NavierStokesProjection<dim>::create_triangulation_and_dofs(const unsigned int n_refines, const std::string &mesh_file)
{
  const auto t0 = std::chrono::steady_clock::now();
  const std::string filename = !mesh_file.empty() ? mesh_file : (dim == 3 ? "cylinderextruded.msh" : "nsbench2.inp");
  s35::Mesh coarse = s35::read_mesh(filename, dim);
  std::cout << "Number of refines = " << n_refines << '\n';
  s35::DiscretizationOptions opt;
  opt.n_refines = (int)n_refines;
  opt.pressure_degree = (int)deg;
  opt.dt = dt;
  disc = s35::build_discretization(coarse, opt);
  std::cout << "minimum cell diameter: " << s35::minimal_cell_diameter(disc->mesh) << '\n';
  std::cout << "Number of active cells: " << disc->mesh.n_cells() << std::endl;
  width = disc->width;
  height = disc->height;
  std::cout << "width: " << width << std::endl;
  std::cout << "height: " << height << std::endl;

  view.reset(new ProblemView(*disc, parameters));
  // The parameter file is the reference's and has no entry for the gated block-ILU alternative
  // (s35_problem.ilu_blocks); the program takes it from the environment and says so.
  if (const char *blocks = std::getenv("S35_ILU_BLOCKS"))
    if (std::atoi(blocks) > 1)
      {
        view->problem.ilu_blocks = std::atoi(blocks);
        std::cout << "GATED ALTERNATIVE: ILU(0) of " << view->problem.ilu_blocks
                  << " index-range blocks instead of the reference's SparseILU" << std::endl;
      }
  check(s35_ctx_create(&view->problem, &ctx));

  std::cout << "dim (X_h) = " << ((long)disc->vel.n_dofs * dim) << std::endl
            << "dim (M_h) = " << disc->pres.n_dofs << std::endl
            << "Re        = " << Re << std::endl
            << std::endl;
  setup_seconds = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
}

// step-35a.cc:834-854: p = 25 - x at both time levels, phi = 0, u = 0
template <int dim>
void NavierStokesProjection<dim>::initialize()
{
  std::vector<double> zero_v(disc->vel.n_dofs, 0.0), zero_p(disc->pres.n_dofs, 0.0);
  check(s35_ctx_set_field(ctx, S35_PRES_N_MINUS_1, 0, disc->pres_nm1_0.data()));
  check(s35_ctx_set_field(ctx, S35_PRES_N, 0, disc->pres_n0.data()));
  check(s35_ctx_set_field(ctx, S35_PHI_N, 0, zero_p.data()));
  check(s35_ctx_set_field(ctx, S35_PHI_N_MINUS_1, 0, zero_p.data()));
  for (int d = 0; d < dim; ++d)
    {
      check(s35_ctx_set_field(ctx, S35_U_N_MINUS_1, d, zero_v.data()));
      check(s35_ctx_set_field(ctx, S35_U_N, d, zero_v.data()));
    }
}

// The checkpoint set of the reference (step-35a.cc:856-900): dataset name -> state vector.
struct CheckpointEntry
{
  std::string dataset;
  s35_field field;
  int component;
};

static std::vector<CheckpointEntry> checkpoint_entries(int dim)
{
  std::vector<CheckpointEntry> e;
  for (int d = 0; d < dim; ++d)
    {
      e.push_back({"u_n" + std::to_string(d), S35_U_N, d});
      e.push_back({"u_n_minus_1" + std::to_string(d), S35_U_N_MINUS_1, d});
    }
  e.push_back({"phi_n", S35_PHI_N, 0});
  e.push_back({"phi_n_minus_1", S35_PHI_N_MINUS_1, 0});
  e.push_back({"pres_n", S35_PRES_N, 0});
  e.push_back({"pres_n_minus_1", S35_PRES_N_MINUS_1, 0});
  return e;
}

template <int dim>
void NavierStokesProjection<dim>::load_checkpoint()
{
  if (load_checkpoint_file_name.empty())
    return;
  std::vector<double> values;
  for (const CheckpointEntry &e : checkpoint_entries(dim))
    {
      H5::load(load_checkpoint_file_name, e.dataset, values);
      const size_t expected = (e.field == S35_U_N || e.field == S35_U_N_MINUS_1) ? disc->vel.n_dofs : disc->pres.n_dofs;
      if (values.size() != expected)
        throw std::runtime_error("checkpoint dataset " + e.dataset + " has the wrong length");
      check(s35_ctx_set_field(ctx, e.field, e.component, values.data()));
    }
}

template <int dim>
void NavierStokesProjection<dim>::save_checkpoint()
{
  if (save_checkpoint_file_name.empty())
    return;
  const std::string file = output_directory + "/" + save_checkpoint_file_name;
  std::cout << "outfile name is '" << save_checkpoint_file_name << "'" << std::endl;
  std::remove(file.c_str());
  std::vector<double> values;
  for (const CheckpointEntry &e : checkpoint_entries(dim))
    {
      values.resize((e.field == S35_U_N || e.field == S35_U_N_MINUS_1) ? disc->vel.n_dofs : disc->pres.n_dofs);
      check(s35_ctx_get_field(ctx, e.field, e.component, values.data()));
      H5::save(file, e.dataset, values);
    }
}

namespace {

// Which step numbers run() visits.  Fresh start: n = 2 .. (T - t0)/dt, and the advection matrix is assembled
// once before step 2.  Restart from a checkpoint: n = t0/dt .. t0/dt + (T - t0)/dt, and -- a quirk of the
// reference kept here -- the matrix is NOT assembled before the first diffusion step, because the prologue is
// tied to n == 0 (step-35a.cc:1174-1191, 1204-1208).
struct StepRange
{
  unsigned int first, last, prologue_at;
};

StepRange step_range(double t_0, double T, double dt, bool restart)
{
  const unsigned int n_steps = static_cast<unsigned int>((T - t_0) / dt);
  if (!restart)
    return StepRange{2u, n_steps, 2u};
  const unsigned int first = static_cast<unsigned int>(t_0 / dt);
  return StepRange{first, first + n_steps, 0u};
}

} // namespace

template <int dim>
unsigned int NavierStokesProjection<dim>::n_time_steps() const
{
  const StepRange r = step_range(t_0, T, dt, !load_checkpoint_file_name.empty());
  return r.last >= r.first ? r.last - r.first + 1 : 0;
}

template <int dim>
void NavierStokesProjection<dim>::run(const bool verbose, const unsigned int output_interval)
{
  auto say = [&](const char *msg) {
    if (verbose)
      std::cout << msg << std::endl;
  };
  const bool restart = !load_checkpoint_file_name.empty();
  const StepRange steps = step_range(t_0, T, dt, restart);
  if (restart)
    load_checkpoint();
  else
    output_results(1, t_0);

  for (unsigned int n = steps.first; n <= steps.last; ++n)
    {
      const double current_time = t_0 + (n - steps.first + steps.prologue_at) * dt;
      if (n % output_interval == 0)
        {
          say("Plotting Solution");
          output_results(n, current_time);
        }
      std::cout << "Step = " << n << " Time = " << current_time << std::endl;
      say("  Interpolating the velocity ");
      if (n == steps.prologue_at)
        {
          interpolate_velocity();
          assemble_advection_term();
        }
      say("  Diffusion Step");
      const bool refresh_preconditioner = (n % vel_update_prec == 0);
      if (refresh_preconditioner)
        say("    With reinitialization of the preconditioner");
      diffusion_step(refresh_preconditioner || n == steps.first);
      interpolate_velocity();
      // two concurrent tasks in the reference (step-35a.cc:1219-1235); they touch disjoint data
      assemble_advection_term();
      say("  Projection Step");
      projection_step();
      say("  Updating the Pressure");
      update_pressure();
      if (verbose)
        {
          // stands in for deal.II's deallog output of the solvers (step-35a.cc:1724)
          s35_solver_stats st;
          for (int d = 0; d < dim; ++d)
            {
              check(s35_ctx_solver_stats(ctx, d, &st));
              std::cout << "    GMRES[" << d << "]: " << st.iterations << " iterations, residual " << st.final_residual << std::endl;
            }
          check(s35_ctx_solver_stats(ctx, -1, &st));
          std::cout << "    CG: " << st.iterations << " iterations, residual " << st.final_residual << std::endl;
        }
    }
  save_checkpoint();
  print_timer_summary();
}

template <int dim>
void NavierStokesProjection<dim>::interpolate_velocity()
{
  check(s35_interpolate_velocity(ctx));
}
template <int dim>
void NavierStokesProjection<dim>::diffusion_step(const bool reinit_prec)
{
  check(s35_diffusion_step(ctx, reinit_prec ? 1 : 0));
}
template <int dim>
void NavierStokesProjection<dim>::projection_step()
{
  check(s35_projection_step(ctx));
}
template <int dim>
void NavierStokesProjection<dim>::update_pressure()
{
  check(s35_update_pressure(ctx));
}
template <int dim>
void NavierStokesProjection<dim>::assemble_advection_term()
{
  check(s35_assemble_advection_term(ctx));
}

// The reference writes mesh.h5 / solution-%09d.h5 / solution.xdmf through deal.II's DataOut and a
// snapshot-%09d.h5 of u_n (step-35a.cc:1548-1661).  Here the fields go into solution-%09d.h5 and
// snapshot-%09d.h5 in the container of h5lite.h, datasets "v0".."v{dim-1}", "p" and "a0".."a{dim-1}".
template <int dim>
void NavierStokesProjection<dim>::output_results(const unsigned int step, const double time)
{
  (void)time;
  char name[64];
  std::snprintf(name, sizeof(name), "solution-%09u.h5", step);
  const std::string sol = output_directory + "/" + name;
  std::snprintf(name, sizeof(name), "snapshot-%09u.h5", step);
  const std::string snap = output_directory + "/" + name;
  std::remove(sol.c_str());
  std::remove(snap.c_str());
  std::vector<double> v(disc->vel.n_dofs), p(disc->pres.n_dofs);
  for (int d = 0; d < dim; ++d)
    {
      check(s35_ctx_get_field(ctx, S35_U_N, d, v.data()));
      H5::save(sol, "v" + std::to_string(d), v);
      H5::save(snap, "a" + std::to_string(d), v);
    }
  check(s35_ctx_get_field(ctx, S35_PRES_N, 0, p.data()));
  H5::save(sol, "p", p);
}

// TimerOutput summary (step-35a.cc:508,714): device time per section
template <int dim>
void NavierStokesProjection<dim>::print_timer_summary() const
{
  static const char *names[] = {"assemble_advection_term", "interpolate_velocity",     "projection_step_setup", "projection_step_solve",
                                "setup_diffusion_rhs",     "solve_diffusion_component", "update_pressure"};
  double total = 0.0;
  double ms[7];
  int64_t calls[7];
  for (int i = 0; i < 7; ++i)
    {
      s35_ctx_section_ms(ctx, names[i], &ms[i], &calls[i]);
      total += ms[i];
    }
  std::cout << "\n+---------------------------------------------+------------+------------+\n"
            << "| Total device time elapsed in sections       |" << std::setw(10) << std::fixed << std::setprecision(3)
            << total * 1e-3 << "s |            |\n"
            << "| Section                         | no. calls  |  dev time  | % of total |\n"
            << "+---------------------------------+------------+------------+------------+\n";
  for (int i = 0; i < 7; ++i)
    std::cout << "| " << std::left << std::setw(32) << names[i] << "|" << std::right << std::setw(11) << calls[i] << " |"
              << std::setw(10) << std::setprecision(3) << ms[i] * 1e-3 << "s |" << std::setw(10) << std::setprecision(1)
              << (total > 0 ? 100.0 * ms[i] / total : 0.0) << "% |\n";
  std::cout << "+---------------------------------+------------+------------+------------+\n" << std::endl;
  std::cout.unsetf(std::ios::fixed);
  std::cout << std::setprecision(6);
}

template class NavierStokesProjection<2>;
template class NavierStokesProjection<3>;

} // namespace Step35
