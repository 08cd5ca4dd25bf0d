// extern "C" layer of include/step35a_host.h
#include "../../include/step35a_host.h"
#include "navier_stokes_projection.h"
#include "partition.h"
#include "h5.h"

#include <chrono>
#include <cstring>
#include <memory>
#include <stdexcept>

// S35_HOST_ONLY: the build of this file that goes into libstep35a_host.so (set-up, partition, drag / lift; no device
// code, no simulation driver).  It keeps its own error string.
namespace s35 {
#ifdef S35_HOST_ONLY
namespace {
thread_local std::string g_host_last_error;
}
void set_last_error(const std::string &msg) { g_host_last_error = msg; }
#else
void set_last_error(const std::string &msg); // csrc/capi.cu
#endif
void drag_lift(const Discretization &D, double reynolds, const double *u, const double *p, double *out);
}

using Step35::RunTimeParameters::Data_Storage;

struct s35_disc
{
  const s35::Discretization *D = nullptr;
  const Step35::ProblemView *view = nullptr;
  std::unique_ptr<s35::Discretization> own_D;       // empty when the handle is a view into a simulation
  std::unique_ptr<Step35::ProblemView> own_view;
  double setup_seconds = 0.0;
};

struct s35_part
{
  std::unique_ptr<s35::LocalProblem> L;
};

#ifndef S35_HOST_ONLY
struct s35_sim
{
  int dim = 0;
  std::unique_ptr<Step35::NavierStokesProjection<2>> sim2;
  std::unique_ptr<Step35::NavierStokesProjection<3>> sim3;
  s35_disc disc_view; // non-owning view for s35_sim_disc
};
#endif

namespace {

template <typename F>
s35_status guarded(F &&f, s35_status fallback = S35_ERR_INVALID)
{
  try
    {
      f();
      return S35_OK;
    }
  catch (const std::exception &e)
    {
      s35::set_last_error(e.what());
      return fallback;
    }
}

void to_struct(const Data_Storage &d, s35_params *p)
{
  std::memset(p, 0, sizeof(*p));
  p->rotational = d.form == Step35::RunTimeParameters::METHOD_ROTATIONAL ? 1 : 0;
  p->initial_time = d.initial_time;
  p->final_time = d.final_time;
  p->reynolds = d.Reynolds;
  p->dt = d.dt;
  p->n_global_refines = (int32_t)d.n_global_refines;
  p->pressure_degree = (int32_t)d.pressure_degree;
  p->vel_max_iterations = (int32_t)d.vel_max_iterations;
  p->vel_krylov_size = (int32_t)d.vel_Krylov_size;
  p->vel_off_diagonals = (int32_t)d.vel_off_diagonals;
  p->vel_update_prec = (int32_t)d.vel_update_prec;
  p->vel_eps = d.vel_eps;
  p->vel_diag_strength = d.vel_diag_strength;
  p->verbose = d.verbose ? 1 : 0;
  p->output_interval = (int32_t)d.output_interval;
  std::strncpy(p->load_checkpoint_file_name, d.load_checkpoint_file_name.c_str(), sizeof(p->load_checkpoint_file_name) - 1);
  std::strncpy(p->save_checkpoint_file_name, d.save_checkpoint_file_name.c_str(), sizeof(p->save_checkpoint_file_name) - 1);
}

Data_Storage from_struct(const s35_params &p)
{
  Data_Storage d;
  d.form = p.rotational ? Step35::RunTimeParameters::METHOD_ROTATIONAL : Step35::RunTimeParameters::METHOD_STANDARD;
  d.initial_time = p.initial_time;
  d.final_time = p.final_time;
  d.Reynolds = p.reynolds;
  d.dt = p.dt;
  d.n_global_refines = (unsigned)p.n_global_refines;
  d.pressure_degree = (unsigned)p.pressure_degree;
  d.vel_max_iterations = (unsigned)p.vel_max_iterations;
  d.vel_Krylov_size = (unsigned)p.vel_krylov_size;
  d.vel_off_diagonals = (unsigned)p.vel_off_diagonals;
  d.vel_update_prec = (unsigned)p.vel_update_prec;
  d.vel_eps = p.vel_eps;
  d.vel_diag_strength = p.vel_diag_strength;
  d.verbose = p.verbose != 0;
  d.output_interval = (unsigned)p.output_interval;
  d.load_checkpoint_file_name = p.load_checkpoint_file_name;
  d.save_checkpoint_file_name = p.save_checkpoint_file_name;
  return d;
}

} // namespace

extern "C" {

void s35_params_defaults(s35_params *p)
{
  Data_Storage d;
  to_struct(d, p);
}

s35_status s35_params_read(const char *prm_file, s35_params *p)
{
  return guarded(
    [&] {
      if (!prm_file || !p)
        throw std::runtime_error("params_read: null argument");
      Data_Storage d;
      d.read_data(prm_file);
      to_struct(d, p);
    },
    S35_ERR_IO);
}

s35_status s35_params_parse(const char *prm_text, s35_params *p)
{
  return guarded([&] {
    if (!prm_text || !p)
      throw std::runtime_error("params_parse: null argument");
    Data_Storage d;
    d.read_data_from_string(prm_text);
    to_struct(d, p);
  });
}

s35_status s35_disc_create(const char *mesh_file, int32_t dim, const s35_params *params, s35_disc **out)
{
  return guarded([&] {
    if (!mesh_file || !params || !out)
      throw std::runtime_error("disc_create: null argument");
    *out = nullptr;
    if (dim != 2 && dim != 3)
      throw std::runtime_error("disc_create: dim must be 2 or 3");
    const auto t0 = std::chrono::steady_clock::now();
    std::unique_ptr<s35_disc> d(new s35_disc);
    s35::Mesh coarse = s35::read_mesh(mesh_file, dim);
    s35::DiscretizationOptions opt;
    opt.n_refines = params->n_global_refines;
    opt.pressure_degree = params->pressure_degree;
    opt.dt = params->dt;
    d->own_D = s35::build_discretization(coarse, opt);
    d->own_view.reset(new Step35::ProblemView(*d->own_D, from_struct(*params)));
    d->D = d->own_D.get();
    d->view = d->own_view.get();
    d->setup_seconds = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    *out = d.release();
  });
}

void s35_disc_destroy(s35_disc *d) { delete d; }

const s35_problem *s35_disc_problem(const s35_disc *d) { return d ? &d->view->problem : nullptr; }

static const s35::Discretization &disc_of(const s35_disc *d)
{
  if (!d || !d->D)
    throw std::runtime_error("null discretisation handle");
  return *d->D;
}

s35_status s35_disc_get_info(const s35_disc *d, s35_disc_info *info)
{
  return guarded([&] {
    if (!info)
      throw std::runtime_error("disc_get_info: null argument");
    const s35::Discretization &D = disc_of(d);
    info->dim = D.dim;
    info->n_cells = (int32_t)D.mesh.n_cells();
    info->n_vertices = (int32_t)D.mesh.n_vertices();
    info->n_vel = D.vel.n_dofs;
    info->n_pres = D.pres.n_dofs;
    info->n_colors = D.n_colors;
    info->nnz_vel_constrained = D.vel_pattern_constrained.nnz();
    info->nnz_vel_unconstrained = D.vel_pattern_unconstrained.nnz();
    info->nnz_grad = D.grad_pattern.nnz();
    info->nnz_proj = D.proj_pattern.nnz();
    info->nnz_pres = D.pres_pattern.nnz();
    info->width = D.width;
    info->height = D.height;
    info->min_cell_diameter = s35::minimal_cell_diameter(D.mesh);
    info->setup_seconds = d->setup_seconds;
  });
}

const double *s35_disc_support_points(const s35_disc *d, int32_t which)
{
  if (!d || !d->D)
    return nullptr;
  return which == 0 ? d->D->vel.support.data() : d->D->pres.support.data();
}

const double *s35_disc_initial_pressure(const s35_disc *d) { return d && d->D ? d->D->pres_n0.data() : nullptr; }

s35_status s35_disc_mesh(const s35_disc *d, const double **vertices, const int32_t **cells, const int32_t **bfaces,
                         const int32_t **bface_ids, int64_t *n_bfaces)
{
  return guarded([&] {
    const s35::Discretization &D = disc_of(d);
    if (vertices)
      *vertices = D.mesh.vertices.data();
    if (cells)
      *cells = D.mesh.cells.data();
    if (bfaces)
      *bfaces = D.mesh.bfaces.data();
    if (bface_ids)
      *bface_ids = D.mesh.bface_ids.data();
    if (n_bfaces)
      *n_bfaces = D.mesh.n_bfaces();
  });
}

const int32_t *s35_disc_pres_cell_dofs(const s35_disc *d) { return d && d->D ? d->D->pres.cell_dofs.data() : nullptr; }

s35_status s35_part_create(const s35_disc *d, int32_t n_ranks, int32_t rank, s35_part **out)
{
  return guarded([&] {
    if (!out)
      throw std::runtime_error("part_create: null argument");
    *out = nullptr;
    const s35::Discretization &D = disc_of(d);
    if (n_ranks < 1 || rank < 0 || rank >= n_ranks)
      throw std::runtime_error("part_create: bad rank / n_ranks");
    std::unique_ptr<s35_part> p(new s35_part);
    p->L = s35::extract_local_problem(D, s35::partition_cells(D.mesh, n_ranks), rank, n_ranks, d->view->problem);
    *out = p.release();
  });
}

void s35_part_destroy(s35_part *p) { delete p; }

const s35_problem *s35_part_problem(const s35_part *p) { return p ? &p->L->problem : nullptr; }

const double *s35_part_initial_pressure(const s35_part *p) { return p ? p->L->initial_pressure.data() : nullptr; }

s35_status s35_disc_partition_cells(const s35_disc *d, int32_t n_ranks, int32_t *cell_rank)
{
  return guarded([&] {
    if (!cell_rank)
      throw std::runtime_error("partition_cells: null argument");
    const s35::Discretization &D = disc_of(d);
    const std::vector<int32_t> r = s35::partition_cells(D.mesh, n_ranks);
    std::copy(r.begin(), r.end(), cell_rank);
  });
}

s35_status s35_mesh_convert(const char *in_file, int32_t dim, const char *out_file)
{
  return guarded(
    [&] {
      if (!in_file || !out_file)
        throw std::runtime_error("mesh_convert: null argument");
      const s35::Mesh m = s35::read_mesh(in_file, dim);
      const std::string out(out_file);
      auto ends_with = [&](const char *suffix) {
        const size_t n = std::strlen(suffix);
        return out.size() >= n && out.compare(out.size() - n, n, suffix) == 0;
      };
      if (ends_with(".s35m"))
        s35::write_s35m(m, out);
      else if (ends_with(".msh"))
        s35::write_msh(m, out);
      else if (ends_with(".inp"))
        s35::write_ucd(m, out);
      else
        throw std::runtime_error("mesh_convert: unknown output extension");
    },
    S35_ERR_IO);
}

s35_status s35_mesh_h5_save(const char *file_name, const char *dataset_name, const double *values, int64_t n)
{
  return guarded([&] {
    if (!file_name || !dataset_name || n < 0 || (n > 0 && !values))
      throw std::invalid_argument("h5 save: null argument or negative length");
    H5::save(file_name, dataset_name, std::vector<double>(values, values + n));
  });
}

s35_status s35_mesh_h5_load(const char *file_name, const char *dataset_name, double *values, int64_t capacity, int64_t *n)
{
  return guarded([&] {
    if (!file_name || !dataset_name || !n || capacity < 0 || (capacity > 0 && !values))
      throw std::invalid_argument("h5 load: null argument or negative capacity");
    std::vector<double> v;
    H5::load(file_name, dataset_name, v);
    *n = (int64_t)v.size();
    std::copy(v.begin(), v.begin() + std::min<int64_t>(capacity, *n), values);
  });
}

#ifndef S35_HOST_ONLY
s35_status s35_sim_create(const s35_params *params, const char *mesh_file, int32_t dim, s35_sim **out)
{
  return guarded([&] {
    if (!out)
      throw std::runtime_error("sim_create: null argument");
    *out = nullptr;
    Data_Storage data = params ? from_struct(*params) : Data_Storage();
    std::unique_ptr<s35_sim> s(new s35_sim);
    s->dim = dim;
    const std::string mesh = mesh_file ? mesh_file : "";
    if (dim == 2)
      s->sim2.reset(new Step35::NavierStokesProjection<2>(data, mesh));
    else if (dim == 3)
      s->sim3.reset(new Step35::NavierStokesProjection<3>(data, mesh));
    else
      throw std::runtime_error("sim_create: dim must be 2 or 3");
    s->disc_view.D = dim == 2 ? &s->sim2->discretization() : &s->sim3->discretization();
    s->disc_view.view = dim == 2 ? &s->sim2->problem_view() : &s->sim3->problem_view();
    s->disc_view.setup_seconds = dim == 2 ? s->sim2->setup_seconds : s->sim3->setup_seconds;
    *out = s.release();
  });
}

const s35_disc *s35_sim_disc(const s35_sim *sim) { return sim ? &sim->disc_view : nullptr; }

void s35_sim_destroy(s35_sim *sim) { delete sim; }

s35_status s35_sim_run(s35_sim *sim, int32_t verbose, int32_t output_interval)
{
  return guarded([&] {
    if (!sim)
      throw std::runtime_error("sim_run: null argument");
    if (sim->dim == 2)
      sim->sim2->run(verbose != 0, (unsigned)output_interval);
    else
      sim->sim3->run(verbose != 0, (unsigned)output_interval);
  });
}

s35_ctx *s35_sim_ctx(s35_sim *sim)
{
  if (!sim)
    return nullptr;
  return sim->dim == 2 ? sim->sim2->context() : sim->sim3->context();
}

int32_t s35_sim_n_steps(const s35_sim *sim)
{
  if (!sim)
    return 0;
  return (int32_t)(sim->dim == 2 ? sim->sim2->n_time_steps() : sim->sim3->n_time_steps());
}

#endif

s35_status s35_disc_drag_lift(const s35_disc *d, double reynolds, const double *u, const double *p, double *out)
{
  return guarded([&] {
    if (!u || !p || !out)
      throw std::runtime_error("drag_lift: null argument");
    s35::drag_lift(disc_of(d), reynolds, u, p, out);
  });
}

#ifndef S35_HOST_ONLY
s35_status s35_sim_drag_lift(s35_sim *sim, double *out)
{
  return guarded([&] {
    if (!sim || !out)
      throw std::runtime_error("sim_drag_lift: null argument");
    const s35::Discretization &D = sim->dim == 2 ? sim->sim2->discretization() : sim->sim3->discretization();
    const s35_problem &pr = sim->dim == 2 ? sim->sim2->problem_view().problem : sim->sim3->problem_view().problem;
    s35_ctx *ctx = s35_sim_ctx(sim);
    std::vector<double> u((size_t)D.dim * D.vel.n_dofs), p(D.pres.n_dofs);
    for (int d = 0; d < D.dim; ++d)
      if (s35_ctx_get_field(ctx, S35_U_N, d, u.data() + (size_t)d * D.vel.n_dofs) != S35_OK)
        throw std::runtime_error(s35_last_error());
    if (s35_ctx_get_field(ctx, S35_PRES_N, 0, p.data()) != S35_OK)
      throw std::runtime_error(s35_last_error());
    s35::drag_lift(D, pr.reynolds, u.data(), p.data(), out);
  });
}
#else
const char *s35_last_error(void) { return s35::g_host_last_error.c_str(); }
#endif

} // extern "C"
