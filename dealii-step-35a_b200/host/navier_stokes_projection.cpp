#include "navier_stokes_projection.h"

#include "h5.h"

#include <chrono>
#include <cstdlib>
#include <cstdio>
#include <fstream>
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
  // the file name as given, like the reference (step-35a.cc:891) and like load_checkpoint(): a run restarted with the
  // same name finds the file
  const std::string &file = save_checkpoint_file_name;
  std::cout << "outfile name is '" << save_checkpoint_file_name << "'" << std::endl;
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

// step-35a.cc:1548-1661.  The reference hands the fields to deal.II's DataOut, which writes mesh.h5 (once),
// solution-%09d.h5 and solution.xdmf, and saves u_n as snapshot-%09d.h5 with H5::save_block_vector.  Here the same
// files are written by hand as real HDF5 (h5min.h): mesh.h5 holds /nodes (vertices x 3) and /cells (cells x 2^dim,
// vertex indices in XDMF order), solution-*.h5 the vertex values /v0 ... /v{dim-1} and /p, solution.xdmf one grid per
// output step that points into them, snapshot-*.h5 the blocks /a0 ... /a{dim-1} of u_n.
template <int dim>
void NavierStokesProjection<dim>::output_results(const unsigned int step, const double time)
{
  char name[64];
  std::snprintf(name, sizeof(name), "solution-%09u.h5", step);
  const std::string sol_name = name, sol = output_directory + "/" + sol_name;
  std::snprintf(name, sizeof(name), "snapshot-%09u.h5", step);
  const std::string snap = output_directory + "/" + name;
  const int vpc = 1 << dim, dpc = dim == 2 ? 9 : 27;
  const size_t n_vertices = disc->mesh.vertices.size() / dim, n_cells = (size_t)disc->mesh.n_cells();
  if (write_mesh)
    {
      write_mesh = false;
      h5min::File mesh;
      h5min::Dataset &nodes = mesh["nodes"], &cells = mesh["cells"];
      nodes.dims = {n_vertices, 3};
      nodes.data.assign(3 * n_vertices, 0.0);
      for (size_t v = 0; v < n_vertices; ++v)
        for (int k = 0; k < dim; ++k)
          nodes.data[3 * v + k] = disc->mesh.vertices[v * dim + k];
      // lexicographic vertex order -> the order XDMF wants for Quadrilateral / Hexahedron
      const int xdmf_order[8] = {0, 1, 3, 2, 4, 5, 7, 6};
      cells.dims = {n_cells, (uint64_t)vpc};
      cells.data.resize(n_cells * vpc);
      for (size_t c = 0; c < n_cells; ++c)
        for (int v = 0; v < vpc; ++v)
          cells.data[c * vpc + v] = (double)disc->mesh.cells[c * vpc + xdmf_order[v]];
      h5min::write_file(output_directory + "/mesh.h5", mesh);
      std::ofstream(output_directory + "/finite_element.txt") << "FE_Q<" << dim << ">(" << deg + 1 << ")" << std::endl;
    }
  // vertex values: a Q2 cell holds its vertices at the corners of the 3^dim lexicographic lattice, a Q1 cell in order
  std::vector<std::vector<double>> blocks(dim, std::vector<double>(disc->vel.n_dofs));
  std::vector<double> p(disc->pres.n_dofs);
  for (int d = 0; d < dim; ++d)
    check(s35_ctx_get_field(ctx, S35_U_N, d, blocks[d].data()));
  check(s35_ctx_get_field(ctx, S35_PRES_N, 0, p.data()));
  h5min::File solution;
  for (int d = 0; d <= dim; ++d)
    {
      h5min::Dataset &set = solution[d < dim ? "v" + std::to_string(d) : std::string("p")];
      set.dims = {n_vertices};
      set.data.assign(n_vertices, 0.0);
      for (size_t c = 0; c < n_cells; ++c)
        for (int v = 0; v < vpc; ++v)
          {
            const int corner = (v & 1) * 2 + ((v >> 1) & 1) * 6 + ((v >> 2) & 1) * 18;
            const size_t vertex = (size_t)disc->mesh.cells[c * vpc + v];
            set.data[vertex] = d < dim ? blocks[d][disc->vel.cell_dofs[c * dpc + corner]] : p[disc->pres.cell_dofs[c * vpc + v]];
          }
    }
  h5min::write_file(sol, solution);
  H5::save_block_vector(snap, blocks);
  // solution.xdmf: rewritten with one more grid at every output step, like DataOut::write_xdmf_file
  {
    std::ostringstream e;
    e << "      <Grid Name=\"mesh\" GridType=\"Uniform\">\n        <Time Value=\"" << time << "\"/>\n"
      << "        <Geometry GeometryType=\"XYZ\">\n          <DataItem Dimensions=\"" << n_vertices
      << " 3\" NumberType=\"Float\" Precision=\"8\" Format=\"HDF\">mesh.h5:/nodes</DataItem>\n        </Geometry>\n"
      << "        <Topology TopologyType=\"" << (dim == 2 ? "Quadrilateral" : "Hexahedron") << "\" NumberOfElements=\"" << n_cells
      << "\">\n          <DataItem Dimensions=\"" << n_cells << " " << vpc
      << "\" NumberType=\"Float\" Precision=\"8\" Format=\"HDF\">mesh.h5:/cells</DataItem>\n        </Topology>\n";
    for (int d = 0; d <= dim; ++d)
      {
        const std::string field = d < dim ? "v" + std::to_string(d) : std::string("p");
        e << "        <Attribute Name=\"" << field << "\" AttributeType=\"Scalar\" Center=\"Node\">\n          <DataItem Dimensions=\""
          << n_vertices << "\" NumberType=\"Float\" Precision=\"8\" Format=\"HDF\">" << sol_name << ":/" << field
          << "</DataItem>\n        </Attribute>\n";
      }
    e << "      </Grid>\n";
    xdmf_entries.push_back(e.str());
    std::ofstream x(output_directory + "/solution.xdmf");
    x << "<?xml version=\"1.0\" ?>\n<!DOCTYPE Xdmf SYSTEM \"Xdmf.dtd\" []>\n<Xdmf Version=\"2.0\">\n  <Domain>\n"
      << "    <Grid Name=\"CellTime\" GridType=\"Collection\" CollectionType=\"Temporal\">\n";
    for (const std::string &g : xdmf_entries)
      x << g;
    x << "    </Grid>\n  </Domain>\n</Xdmf>\n";
  }
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
