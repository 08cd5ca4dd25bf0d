/*
 * step35a_host -- C ABI of the host side of step35a_b200: the set-up the reference does with deal.II
 * (mesh input, refinement, DoF numbering, constraints, the one-off operators; step-35a.cc:738-1164) and
 * the NavierStokesProjection / RunTimeParameters driver (step-35a.cc:121-293, 439-735, 1167-1241).
 *
 * None of this is on the per-time-step path; it exists so that the device path of step35a_b200.h can be
 * fed and driven without deal.II, which is not available on the target machines.  Python tests and
 * bench.py bind these functions with ctypes; the C++ program `step-35a` links the same objects.
 */
#ifndef STEP35A_HOST_H
#define STEP35A_HOST_H

#include "step35a_b200.h"

#ifdef __cplusplus
extern "C" {
#endif

/* RunTimeParameters::Data_Storage (step-35a.cc:129-154) as a plain struct */
typedef struct s35_params
{
  int32_t rotational;           /* Method_Form: 1 rotational, 0 standard */
  double initial_time, final_time, reynolds;
  double dt;
  int32_t n_global_refines, pressure_degree;
  int32_t vel_max_iterations, vel_krylov_size, vel_off_diagonals, vel_update_prec;
  double vel_eps, vel_diag_strength;
  int32_t verbose;
  int32_t output_interval;
  char load_checkpoint_file_name[512];
  char save_checkpoint_file_name[512];
} s35_params;

/* declared defaults of the reference (step-35a.cc:157-234) */
void s35_params_defaults(s35_params *p);
/* Data_Storage::read_data (step-35a.cc:243-292) */
s35_status s35_params_read(const char *prm_file, s35_params *p);
s35_status s35_params_parse(const char *prm_text, s35_params *p);

/* ---- discretisation: everything create_triangulation_and_dofs + initialize_* build ---- */
typedef struct s35_disc s35_disc;

/* mesh_file: .inp (UCD), .msh (gmsh 2.2) or .s35m; dim 2 or 3 */
s35_status s35_disc_create(const char *mesh_file, int32_t dim, const s35_params *params, s35_disc **out);
void s35_disc_destroy(s35_disc *d);
/* the problem description to hand to s35_ctx_create; pointers stay valid until s35_disc_destroy */
const s35_problem *s35_disc_problem(const s35_disc *d);

typedef struct s35_disc_info
{
  int32_t dim, n_cells, n_vertices, n_vel, n_pres, n_colors;
  int64_t nnz_vel_constrained, nnz_vel_unconstrained, nnz_grad, nnz_proj, nnz_pres;
  double width, height;         /* max y / max z over the vertices (step-35a.cc:789-798) */
  double min_cell_diameter;
  double setup_seconds;
} s35_disc_info;
s35_status s35_disc_get_info(const s35_disc *d, s35_disc_info *info);
/* support points: [n][dim]; which = 0 velocity, 1 pressure */
const double *s35_disc_support_points(const s35_disc *d, int32_t which);
/* initial pressure 25 - x at the pressure support points (step-35a.cc:838-841) */
const double *s35_disc_initial_pressure(const s35_disc *d);
/* mesh arrays: vertices [n_vertices][dim], cells [n_cells][2^dim] (lexicographic),
 * boundary faces [n_bfaces][2^(dim-1)] and their ids */
s35_status s35_disc_mesh(const s35_disc *d, const double **vertices, const int32_t **cells, const int32_t **bfaces,
                         const int32_t **bface_ids, int64_t *n_bfaces);
/* pressure cell -> DoF table [n_cells][2^dim] (lexicographic) */
const int32_t *s35_disc_pres_cell_dofs(const s35_disc *d);

/* ---- one rank's part of the mesh-partitioned problem (multi-GPU path; new capability, the reference is a
 * single process).  Cells are split by recursive coordinate bisection; a DoF belongs to the lowest rank among
 * its cells; the local problem holds the owned DoFs and one layer of ghosts. ---- */
typedef struct s35_part s35_part;
s35_status s35_part_create(const s35_disc *d, int32_t n_ranks, int32_t rank, s35_part **out);
void s35_part_destroy(s35_part *p);
/* the local problem (problem->dist is set); pointers stay valid until s35_part_destroy */
const s35_problem *s35_part_problem(const s35_part *p);
/* initial pressure at the local pressure DoFs */
const double *s35_part_initial_pressure(const s35_part *p);
/* rank of every cell of the global mesh under the partition into n_ranks parts: [n_cells] */
s35_status s35_disc_partition_cells(const s35_disc *d, int32_t n_ranks, int32_t *cell_rank);

/* mesh file conversion (tools/pack_meshes.py and the reader round-trip tests) */
s35_status s35_mesh_convert(const char *in_file, int32_t dim, const char *out_file);

/* ---- h5.h: H5::save / H5::load (reference h5.h:66-121) ----
 * One 1-D dataset of doubles in a real HDF5 file (written by hand, host/h5min.h: libhdf5 is not needed).  save adds
 * the dataset to an existing file or creates the file; load reads at most `capacity` values into `values` and reports
 * the dataset's length in *n (call with capacity 0 to ask for the length). */
s35_status s35_mesh_h5_save(const char *file_name, const char *dataset_name, const double *values, int64_t n);
s35_status s35_mesh_h5_load(const char *file_name, const char *dataset_name, double *values, int64_t capacity, int64_t *n);

/* ---- the driver: NavierStokesProjection<dim> ---- */
typedef struct s35_sim s35_sim;

/* constructor (step-35a.cc:689-735): builds the discretisation, the device context, initial conditions.
 * prm_file may be NULL (declared defaults).  mesh_file may be NULL: "cylinderextruded.msh" for dim 3 like
 * the reference (step-35a.cc:748), "nsbench2.inp" for dim 2 (step-35a.cc:747). */
s35_status s35_sim_create(const s35_params *params, const char *mesh_file, int32_t dim, s35_sim **out);
void s35_sim_destroy(s35_sim *sim);
/* run(verbose, output_interval) (step-35a.cc:1167-1241); output_dir NULL = current directory */
s35_status s35_sim_run(s35_sim *sim, int32_t verbose, int32_t output_interval);
s35_ctx *s35_sim_ctx(s35_sim *sim);
const s35_disc *s35_sim_disc(const s35_sim *sim);
/* number of time steps run() would do (step-35a.cc:1174,1190-1191) */
int32_t s35_sim_n_steps(const s35_sim *sim);
/* drag and lift coefficients on the obstacle from the current fields (definition);
 * out[0] = c_drag, out[1] = c_lift */
s35_status s35_sim_drag_lift(s35_sim *sim, double *out);

/* drag/lift functional on host arrays (the same code evaluates oracle fields in the tests).
 * u: [dim][n_vel], p: [n_pres] */
s35_status s35_disc_drag_lift(const s35_disc *d, double reynolds, const double *u, const double *p, double *out);

#ifdef __cplusplus
}
#endif
#endif /* STEP35A_HOST_H */
