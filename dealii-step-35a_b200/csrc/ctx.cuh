// Device context of one NavierStokesProjection object: every matrix, vector,
// constraint list and preconditioner the time loop touches
// (step-35a.cc:474-520), resident in HBM between steps.
#pragma once
#include "assembly.cuh"
#include "krylov.cuh"

#include <map>
#include <string>

namespace s35 {

struct SectionTimer
{
  double ms = 0.0;
  int64_t calls = 0;
};

struct Ctx
{
  int dim = 0;
  int32_t n_cells = 0;
  int32_t nv = 0, np = 0;         // local DoFs: owned, then ghosts (all DoFs on a single GPU)
  int32_t nv_own = 0, np_own = 0; // owned DoFs = matrix rows on this rank
  int64_t nv_global = 0, np_global = 0;
  int64_t n_vertices = 0;
  double dt = 0, reynolds = 0, eps = 0, diag_strength = 0;
  int max_its = 0, krylov = 0, update_prec = 0;
  bool rotational = true;
  int device = 0;
  cudaStream_t stream = nullptr;

  // matrices
  DevMatrix vel_adv;              // constrained pattern; values = vel_Advection
  DevBuf<double> it_val[3];       // vel_it_matrix[d] values on the same pattern
  DevMatrix vel_mass;             // unconstrained pattern
  DevMatrix grad;                 // pres_Diff pattern; values of component 0
  DevBuf<double> grad_val[3];     // pres_Diff[d] values
  const DevMatrix *gradT = nullptr; // transposed pattern (owned by grad, or gradT_own on a partition)
  DevMatrix gradT_own;
  // partition: communicator, ghost exchange plans, diagonal blocks for the ILU(0) factors
  Comm *comm = nullptr;
  HaloDev vel_halo, pres_halo;
  DistLink vel_link, pres_link;
  HaloDev comp_halo[3];           // per velocity component: its own send buffer and communicator
  DistLink comp_link[3];
  bool use_blocks = false;        // the ILU(0) factors live on a pattern of their own (partition, ilu_blocks, off_diagonals)
  int32_t ilu_blocks = 1;
  int32_t off_diagonals = 0;      // > 0: the filtered pattern is the matrix's widened by that many side diagonals
  std::vector<int32_t> vel_wide_src; // host copy of vel_blk_src in that mode (s35_ctx_get_operator_values)
  DevMatrix vel_blk, plap_blk;
  DevBuf<int32_t> vel_blk_src, plap_blk_src;
  DevBuf<double> blk_val[3];
  DevBuf<double> gradT_val[3];
  DevMatrix plap, pmass;
  DevBuf<double> pmass_inv_diag;

  // constraints
  CondensePlan flow[3];
  DevBuf<int32_t> wall_dofs, proj_dofs;
  int32_t n_wall = 0, n_proj = 0;

  // preconditioners and solvers
  DevIlu prec_vel[3], prec_pres;
  GmresSolver gmres[3];           // one per velocity component
  ReduceWs ws_comp[3];
  cudaStream_t comp_stream[3] = {nullptr, nullptr, nullptr};
  cudaEvent_t fork_event = nullptr, join_event[3] = {nullptr, nullptr, nullptr};
  // the assembly task of step-35a.cc:1219-1235 on a stream of its own, next to projection_step + update_pressure
  cudaStream_t asm_stream = nullptr;
  cudaEvent_t asm_fork = nullptr, asm_done = nullptr;
  bool asm_in_flight = false;
  bool concurrent_components = true;
  CgSolver cg;
  AdvectionAssembler assembler;
  ReduceWs ws;
  DevBuf<double> scal; // [0] mean |diag|, [1..3] ||force_d||, [4] ||pres_tmp||
  PinnedBuf<double> scal_host;

  // vectors
  DevBuf<double> u_n[3], u_nm1[3], u_star[3], force[3], v_tmp[3];
  DevBuf<double> pres_n, pres_nm1, phi_n, phi_nm1, pres_tmp, p_work;

  s35_solver_stats stats_vel[3] = {}, stats_proj = {}, stats_mass = {};
  std::map<std::string, SectionTimer> sections;
  bool timers_enabled = true;

  // open timer sections of the current call: (name, start, stop)
  struct Pending
  {
    std::string name;
    cudaEvent_t a, b;
  };
  std::vector<Pending> pending;
  std::vector<cudaEvent_t> event_pool;

  explicit Ctx(const s35_problem &p);
  ~Ctx();
  void build(const s35_problem &p);   // the constructor's body
  void release_handles();              // streams and events (device buffers release themselves)
  Ctx(const Ctx &) = delete;
  Ctx &operator=(const Ctx &) = delete;

  void interpolate_velocity();
  void assemble_advection_term(bool overlap = false);
  void join_assembly();
  void diffusion_step(bool reinit_prec);
  void solve_component(int d, bool reinit_prec, cudaStream_t s);
  void projection_step();
  void update_pressure();
  void advance(int first_n, int n_steps, int start_n, int start_offset_n, double *elapsed_ms);

  double *field(s35_field f, int component, int64_t &len);

  size_t begin_section(const char *name, cudaStream_t on = nullptr);
  void end_section(size_t id, cudaStream_t on = nullptr);
  void resolve_sections();
};

} // namespace s35
