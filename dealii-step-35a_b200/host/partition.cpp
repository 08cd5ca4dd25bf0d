#include "partition.h"

#include <algorithm>
#include <numeric>
#include <stdexcept>

namespace s35 {

namespace {

void bisect(const Mesh &mesh, const std::vector<double> &centroid, std::vector<int32_t> &ids, size_t begin, size_t end, int rank0,
            int n_ranks, std::vector<int32_t> &cell_rank)
{
  if (n_ranks == 1)
    {
      for (size_t i = begin; i < end; ++i)
        cell_rank[ids[i]] = rank0;
      return;
    }
  const int dim = mesh.dim;
  // longest extent of this set of cells
  double lo[3] = {1e300, 1e300, 1e300}, hi[3] = {-1e300, -1e300, -1e300};
  for (size_t i = begin; i < end; ++i)
    for (int k = 0; k < dim; ++k)
      {
        lo[k] = std::min(lo[k], centroid[(size_t)ids[i] * dim + k]);
        hi[k] = std::max(hi[k], centroid[(size_t)ids[i] * dim + k]);
      }
  int axis = 0;
  for (int k = 1; k < dim; ++k)
    if (hi[k] - lo[k] > hi[axis] - lo[axis])
      axis = k;
  const int left_ranks = n_ranks / 2;
  const size_t mid = begin + (end - begin) * left_ranks / n_ranks;
  std::nth_element(ids.begin() + begin, ids.begin() + mid, ids.begin() + end, [&](int32_t a, int32_t b) {
    const double xa = centroid[(size_t)a * dim + axis], xb = centroid[(size_t)b * dim + axis];
    return xa != xb ? xa < xb : a < b;
  });
  bisect(mesh, centroid, ids, begin, mid, rank0, left_ranks, cell_rank);
  bisect(mesh, centroid, ids, mid, end, rank0 + left_ranks, n_ranks - left_ranks, cell_rank);
}

// owner of every DoF: the lowest rank among the cells that contain it
std::vector<int32_t> dof_owner(const DofMap &dofs, const std::vector<int32_t> &cell_rank)
{
  std::vector<int32_t> owner(dofs.n_dofs, 0x7fffffff);
  const size_t n_cells = cell_rank.size();
  for (size_t c = 0; c < n_cells; ++c)
    for (int i = 0; i < dofs.dpc; ++i)
      {
        int32_t &o = owner[dofs.cell_dofs[c * dofs.dpc + i]];
        o = std::min(o, cell_rank[c]);
      }
  return owner;
}

// rows of the owned DoFs of a global matrix, columns renumbered to local indices and sorted
void extract_rows(const Csr &G, const std::vector<const std::vector<double> *> &gval, const std::vector<int32_t> &row_global,
                  int32_t n_rows_owned, const std::vector<int32_t> &col_g2l, int32_t n_cols_local, Csr &L,
                  std::vector<std::vector<double> *> lval)
{
  L.n_rows = n_rows_owned;
  L.n_cols = n_cols_local;
  L.rowptr.assign(n_rows_owned + 1, 0);
  for (int32_t r = 0; r < n_rows_owned; ++r)
    L.rowptr[r + 1] = L.rowptr[r] + (G.rowptr[row_global[r] + 1] - G.rowptr[row_global[r]]);
  L.col.resize(L.rowptr[n_rows_owned]);
  for (auto *v : lval)
    v->resize(L.col.size());
  bool missing = false;
#pragma omp parallel
  {
    std::vector<std::pair<int32_t, int32_t>> row; // (local column, global position)
#pragma omp for schedule(static)
    for (int32_t r = 0; r < n_rows_owned; ++r)
      {
        const int32_t g = row_global[r];
        row.clear();
        for (int32_t k = G.rowptr[g]; k < G.rowptr[g + 1]; ++k)
          {
            const int32_t lc = col_g2l[G.col[k]];
            if (lc < 0)
              missing = true;
            row.emplace_back(lc, k);
          }
        std::sort(row.begin(), row.end());
        for (size_t i = 0; i < row.size(); ++i)
          {
            L.col[L.rowptr[r] + i] = row[i].first;
            for (size_t m = 0; m < gval.size(); ++m)
              (*lval[m])[L.rowptr[r] + i] = (*gval[m])[row[i].second];
          }
      }
  }
  if (missing)
    throw std::runtime_error("partition: a column of an owned row is not local");
}

} // namespace

std::vector<int32_t> partition_cells(const Mesh &mesh, int n_ranks)
{
  if (n_ranks < 1)
    throw std::runtime_error("partition_cells: n_ranks must be positive");
  const int dim = mesh.dim, vpc = mesh.verts_per_cell();
  const int64_t n_cells = mesh.n_cells();
  std::vector<double> centroid((size_t)n_cells * dim, 0.0);
  for (int64_t c = 0; c < n_cells; ++c)
    for (int v = 0; v < vpc; ++v)
      for (int k = 0; k < dim; ++k)
        centroid[c * dim + k] += mesh.vertices[(size_t)mesh.cells[c * vpc + v] * dim + k] / vpc;
  std::vector<int32_t> ids(n_cells), cell_rank(n_cells, 0);
  std::iota(ids.begin(), ids.end(), 0);
  bisect(mesh, centroid, ids, 0, ids.size(), 0, n_ranks, cell_rank);
  return cell_rank;
}

std::unique_ptr<LocalProblem> extract_local_problem(const Discretization &D, const std::vector<int32_t> &cell_rank, int rank,
                                                    int n_ranks, const s35_problem &global)
{
  const int dim = D.dim, vpc = 1 << dim, vdpc = D.vel.dpc, pdpc = D.pres.dpc;
  const int64_t n_cells = D.mesh.n_cells();
  if ((int64_t)cell_rank.size() != n_cells)
    throw std::runtime_error("partition: cell_rank has the wrong length");
  std::unique_ptr<LocalProblem> lp(new LocalProblem);
  LocalProblem &L = *lp;
  L.rank = rank;
  L.n_ranks = n_ranks;
  const std::vector<int32_t> vel_owner = dof_owner(D.vel, cell_rank), pres_owner = dof_owner(D.pres, cell_rank);

  // ranks that touch a cell = owners of its velocity DoFs; halo cells of `rank` = cells it touches
  auto touching_ranks = [&](int64_t c, int32_t *out) {
    int n = 0;
    for (int i = 0; i < vdpc; ++i)
      {
        const int32_t o = vel_owner[D.vel.cell_dofs[c * vdpc + i]];
        bool seen = false;
        for (int k = 0; k < n; ++k)
          seen = seen || out[k] == o;
        if (!seen)
          out[n++] = o;
      }
    return n;
  };
  std::vector<char> vel_local(D.vel.n_dofs, 0), pres_local(D.pres.n_dofs, 0);
  // (neighbour rank, global index) of owned DoFs that are ghosts elsewhere
  std::vector<std::pair<int32_t, int32_t>> vel_send, pres_send;
  int32_t touch[27];
  for (int64_t c = 0; c < n_cells; ++c)
    {
      const int nt = touching_ranks(c, touch);
      bool mine = false;
      for (int k = 0; k < nt; ++k)
        mine = mine || touch[k] == rank;
      if (!mine)
        continue;
      L.cells.push_back((int32_t)c);
      for (int i = 0; i < vdpc; ++i)
        {
          const int32_t g = D.vel.cell_dofs[c * vdpc + i];
          vel_local[g] = 1;
          if (vel_owner[g] == rank)
            for (int k = 0; k < nt; ++k)
              if (touch[k] != rank)
                vel_send.emplace_back(touch[k], g);
        }
      for (int i = 0; i < pdpc; ++i)
        {
          const int32_t g = D.pres.cell_dofs[c * pdpc + i];
          pres_local[g] = 1;
          if (pres_owner[g] == rank)
            for (int k = 0; k < nt; ++k)
              if (touch[k] != rank)
                pres_send.emplace_back(touch[k], g);
        }
    }

  // local numbering and halo plans
  auto number = [&](const std::vector<char> &is_local, const std::vector<int32_t> &owner, std::vector<int32_t> &l2g,
                    std::vector<int32_t> &g2l, int32_t &n_owned, std::vector<std::pair<int32_t, int32_t>> &send, HaloPlan &halo) {
    const int32_t n_global = (int32_t)owner.size();
    g2l.assign(n_global, -1);
    std::vector<std::pair<int32_t, int32_t>> ghosts; // (owner, global)
    for (int32_t g = 0; g < n_global; ++g)
      if (is_local[g])
        {
          if (owner[g] == rank)
            l2g.push_back(g);
          else
            ghosts.emplace_back(owner[g], g);
        }
    n_owned = (int32_t)l2g.size();
    std::sort(ghosts.begin(), ghosts.end());
    for (auto &gh : ghosts)
      l2g.push_back(gh.second);
    for (size_t l = 0; l < l2g.size(); ++l)
      g2l[l2g[l]] = (int32_t)l;
    std::sort(send.begin(), send.end());
    send.erase(std::unique(send.begin(), send.end()), send.end());
    // neighbours: every rank we receive from or send to (the relation is symmetric by construction)
    std::vector<int32_t> nb;
    for (auto &gh : ghosts)
      nb.push_back(gh.first);
    for (auto &s : send)
      nb.push_back(s.first);
    std::sort(nb.begin(), nb.end());
    nb.erase(std::unique(nb.begin(), nb.end()), nb.end());
    halo.neighbor_rank = nb;
    halo.send_ptr.assign(nb.size() + 1, 0);
    halo.recv_ptr.assign(nb.size() + 1, 0);
    for (size_t q = 0; q < nb.size(); ++q)
      {
        int32_t ns = 0, nr = 0;
        for (auto &s : send)
          ns += s.first == nb[q];
        for (auto &gh : ghosts)
          nr += gh.first == nb[q];
        halo.send_ptr[q + 1] = halo.send_ptr[q] + ns;
        halo.recv_ptr[q + 1] = halo.recv_ptr[q] + nr;
      }
    for (auto &s : send)
      halo.send_idx.push_back(g2l[s.second]);
  };
  std::vector<int32_t> vg2l, pg2l;
  number(vel_local, vel_owner, L.vel_global, vg2l, L.n_vel_owned, vel_send, L.vel_halo);
  number(pres_local, pres_owner, L.pres_global, pg2l, L.n_pres_owned, pres_send, L.pres_halo);
  const int32_t nv = (int32_t)L.vel_global.size(), np = (int32_t)L.pres_global.size();

  // cells
  const int32_t nc = (int32_t)L.cells.size();
  L.cell_vertices.resize((size_t)nc * vpc * dim);
  L.vel_cell_dofs.resize((size_t)nc * vdpc);
  L.pres_cell_dofs.resize((size_t)nc * pdpc);
  L.cell_color.resize(nc);
  for (int32_t lc = 0; lc < nc; ++lc)
    {
      const int64_t c = L.cells[lc];
      for (int v = 0; v < vpc; ++v)
        for (int k = 0; k < dim; ++k)
          L.cell_vertices[((size_t)lc * vpc + v) * dim + k] = D.mesh.vertices[(size_t)D.mesh.cells[c * vpc + v] * dim + k];
      for (int i = 0; i < vdpc; ++i)
        L.vel_cell_dofs[(size_t)lc * vdpc + i] = vg2l[D.vel.cell_dofs[c * vdpc + i]];
      for (int i = 0; i < pdpc; ++i)
        L.pres_cell_dofs[(size_t)lc * pdpc + i] = pg2l[D.pres.cell_dofs[c * pdpc + i]];
      L.cell_color[lc] = D.cell_color[c];
    }

  // matrices: rows of the owned DoFs
  extract_rows(D.vel_pattern_constrained, {}, L.vel_global, L.n_vel_owned, vg2l, nv, L.velc, {});
  extract_rows(D.vel_pattern_unconstrained, {&D.vel_mass}, L.vel_global, L.n_vel_owned, vg2l, nv, L.velm, {&L.velm_val});
  {
    std::vector<const std::vector<double> *> gv;
    std::vector<std::vector<double> *> lv;
    for (int d = 0; d < dim; ++d)
      {
        gv.push_back(&D.pres_diff[d]);
        lv.push_back(&L.grad_val[d]);
      }
    extract_rows(D.grad_pattern, gv, L.vel_global, L.n_vel_owned, pg2l, np, L.grad, lv);
  }
  {
    // global transpose of the gradient operator, then the rows of the owned pressure DoFs
    const Csr &G = D.grad_pattern;
    Csr T;
    T.n_rows = G.n_cols;
    T.n_cols = G.n_rows;
    T.rowptr.assign(T.n_rows + 1, 0);
    for (int32_t c2 : G.col)
      T.rowptr[c2 + 1]++;
    for (int32_t r = 0; r < T.n_rows; ++r)
      T.rowptr[r + 1] += T.rowptr[r];
    T.col.resize(G.col.size());
    std::vector<double> tval[3];
    for (int d = 0; d < dim; ++d)
      tval[d].resize(G.col.size());
    std::vector<int32_t> fill(T.rowptr.begin(), T.rowptr.end() - 1);
    for (int32_t r = 0; r < G.n_rows; ++r)
      for (int32_t k = G.rowptr[r]; k < G.rowptr[r + 1]; ++k)
        {
          const int32_t pos = fill[G.col[k]]++;
          T.col[pos] = r;
          for (int d = 0; d < dim; ++d)
            tval[d][pos] = D.pres_diff[d][k];
        }
    std::vector<const std::vector<double> *> gv;
    std::vector<std::vector<double> *> lv;
    for (int d = 0; d < dim; ++d)
      {
        gv.push_back(&tval[d]);
        lv.push_back(&L.gradT_val[d]);
      }
    extract_rows(T, gv, L.pres_global, L.n_pres_owned, vg2l, nv, L.gradT, lv);
  }
  extract_rows(D.proj_pattern, {&D.pres_laplace}, L.pres_global, L.n_pres_owned, pg2l, np, L.plap, {&L.plap_val});
  extract_rows(D.pres_pattern, {&D.pres_mass}, L.pres_global, L.n_pres_owned, pg2l, np, L.pmass, {&L.pmass_val});

  // constraints, local indices of owned and ghost DoFs
  auto localise = [&](const std::vector<int32_t> &global, const std::vector<double> *gvals, const std::vector<int32_t> &g2l,
                      std::vector<int32_t> &out, std::vector<double> *ovals) {
    std::vector<std::pair<int32_t, double>> items;
    for (size_t i = 0; i < global.size(); ++i)
      if (g2l[global[i]] >= 0)
        items.emplace_back(g2l[global[i]], gvals ? (*gvals)[i] : 0.0);
    std::sort(items.begin(), items.end());
    for (auto &it : items)
      {
        out.push_back(it.first);
        if (ovals)
          ovals->push_back(it.second);
      }
  };
  localise(D.wall_dofs, nullptr, vg2l, L.wall, nullptr);
  for (int d = 0; d < dim; ++d)
    localise(D.flow_dofs[d], &D.flow_vals[d], vg2l, L.flow_dofs[d], &L.flow_vals[d]);
  localise(D.proj_dofs, nullptr, pg2l, L.proj, nullptr);
  L.initial_pressure.resize(np);
  for (int32_t l = 0; l < np; ++l)
    L.initial_pressure[l] = D.pres_n0[L.pres_global[l]];

  // C views
  s35_dist &dist = L.dist;
  dist.rank = rank;
  dist.n_ranks = n_ranks;
  dist.n_vel_owned = L.n_vel_owned;
  dist.n_pres_owned = L.n_pres_owned;
  dist.n_vel_global = D.vel.n_dofs;
  dist.n_pres_global = D.pres.n_dofs;
  dist.vel_halo = L.vel_halo.view();
  dist.pres_halo = L.pres_halo.view();
  dist.vel_global_index = L.vel_global.data();
  dist.pres_global_index = L.pres_global.data();
  dist.gradT_rowptr = L.gradT.rowptr.data();
  dist.gradT_col = L.gradT.col.data();
  for (int d = 0; d < 3; ++d)
    dist.gradT_val[d] = d < dim ? L.gradT_val[d].data() : nullptr;
  s35_problem &p = L.problem;
  p = s35_problem{};
  p.dim = dim;
  p.n_cells = nc;
  p.n_vel = nv;
  p.n_pres = np;
  p.cell_vertices = L.cell_vertices.data();
  p.vel_cell_dofs = L.vel_cell_dofs.data();
  p.cell_color = L.cell_color.data();
  p.velc_rowptr = L.velc.rowptr.data();
  p.velc_col = L.velc.col.data();
  p.velm_rowptr = L.velm.rowptr.data();
  p.velm_col = L.velm.col.data();
  p.velm_val = L.velm_val.data();
  p.grad_rowptr = L.grad.rowptr.data();
  p.grad_col = L.grad.col.data();
  for (int d = 0; d < dim; ++d)
    p.grad_val[d] = L.grad_val[d].data();
  p.plap_rowptr = L.plap.rowptr.data();
  p.plap_col = L.plap.col.data();
  p.plap_val = L.plap_val.data();
  p.pmass_rowptr = L.pmass.rowptr.data();
  p.pmass_col = L.pmass.col.data();
  p.pmass_val = L.pmass_val.data();
  p.n_wall = (int32_t)L.wall.size();
  p.wall_dofs = L.wall.data();
  for (int d = 0; d < dim; ++d)
    {
      p.n_flow[d] = (int32_t)L.flow_dofs[d].size();
      p.flow_dofs[d] = L.flow_dofs[d].data();
      p.flow_vals[d] = L.flow_vals[d].data();
    }
  p.n_proj = (int32_t)L.proj.size();
  p.proj_dofs = L.proj.data();
  p.dt = global.dt;
  p.reynolds = global.reynolds;
  p.vel_eps = global.vel_eps;
  p.vel_diag_strength = global.vel_diag_strength;
  p.vel_max_iterations = global.vel_max_iterations;
  p.vel_krylov_size = global.vel_krylov_size;
  p.vel_off_diagonals = global.vel_off_diagonals;
  p.vel_update_prec = global.vel_update_prec;
  p.rotational = global.rotational;
  p.dist = &L.dist;
  return lp;
}

} // namespace s35
