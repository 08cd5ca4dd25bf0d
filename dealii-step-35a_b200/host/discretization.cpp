#include "discretization.h"

#include <algorithm>
#include <cmath>
#include <cstring>
#include <numeric>
#include <stdexcept>

namespace s35 {

int32_t Csr::find(int32_t r, int32_t c) const
{
  const int32_t *b = col.data() + rowptr[r], *e = col.data() + rowptr[r + 1];
  const int32_t *p = std::lower_bound(b, e, c);
  return (p != e && *p == c) ? (int32_t)(p - col.data()) : -1;
}

size_t DofMap::KeyHash::operator()(const Key &k) const
{
  uint64_t h = 0x9E3779B97F4A7C15ull;
  for (int i = 0; i < 4; ++i)
    {
      h ^= (uint64_t)(uint32_t)k.v[i] + 0x9E3779B97F4A7C15ull + (h << 6) + (h >> 2);
      h *= 0xff51afd7ed558ccdull;
    }
  return (size_t)(h ^ (h >> 32));
}

namespace {

// corner vertices (of an entity with 2^edim lexicographic vertices `ev`) that
// span the sub-entity on which the node with multi-index idx in {0..degree}^edim
// lives; degree 1 or 2.
int node_corners(const int32_t *ev, int edim, int degree, const int *idx, int32_t *corners)
{
  int n = 0;
  for (int c = 0; c < (1 << edim); ++c)
    {
      bool inside = true;
      for (int d = 0; d < edim; ++d)
        {
          const bool interior = idx[d] != 0 && idx[d] != degree;
          if (!interior && ((c >> d) & 1) != (idx[d] == degree ? 1 : 0))
            inside = false;
        }
      if (inside)
        corners[n++] = ev[c];
    }
  return n;
}

DofMap::Key make_key(const int32_t *corners, int n)
{
  DofMap::Key k{{-1, -1, -1, -1}};
  for (int i = 0; i < n; ++i)
    k.v[i] = corners[i];
  std::sort(k.v, k.v + n);
  return k;
}

} // namespace

DofMap distribute_dofs(const Mesh &mesh, int degree)
{
  if (degree != 1 && degree != 2)
    throw std::runtime_error("distribute_dofs: only Q1 and Q2 are implemented (Q2/Q1 Taylor-Hood)");
  const int dim = mesh.dim, vpc = mesh.verts_per_cell();
  DofMap dofs;
  dofs.dim = dim;
  dofs.degree = degree;
  const int n1 = degree + 1;
  dofs.dpc = ipow(n1, dim);
  const int64_t n_cells = mesh.n_cells();
  dofs.cell_dofs.resize((size_t)n_cells * dofs.dpc);
  dofs.support = mesh.vertices; // vertex DoFs first, numbered like the vertices
  int32_t next = (int32_t)mesh.n_vertices();
  dofs.entity_dof.reserve((size_t)mesh.n_vertices() * (degree == 2 ? 8 : 1));
  for (int32_t v = 0; v < next; ++v)
    dofs.entity_dof.emplace(make_key(&v, 1), v);

  auto new_dof_at_mean = [&](const int32_t *corners, int n) {
    for (int k = 0; k < dim; ++k)
      {
        double s = 0.0;
        for (int i = 0; i < n; ++i)
          s += mesh.vertices[(size_t)corners[i] * dim + k];
        dofs.support.push_back(s / n);
      }
    return next++;
  };

  for (int64_t c = 0; c < n_cells; ++c)
    {
      const int32_t *cv = &mesh.cells[c * vpc];
      for (int i = 0; i < dofs.dpc; ++i)
        {
          const int idx[3] = {i % n1, (i / n1) % n1, i / (n1 * n1)};
          int32_t corners[8];
          const int n = node_corners(cv, dim, degree, idx, corners);
          int32_t dof;
          if (n == 1)
            dof = corners[0];
          else if (n == vpc)
            dof = new_dof_at_mean(corners, n);
          else
            {
              const DofMap::Key key = make_key(corners, n);
              auto it = dofs.entity_dof.find(key);
              if (it == dofs.entity_dof.end())
                {
                  dof = new_dof_at_mean(corners, n);
                  dofs.entity_dof.emplace(key, dof);
                }
              else
                dof = it->second;
            }
          dofs.cell_dofs[(size_t)c * dofs.dpc + i] = dof;
        }
    }
  dofs.n_dofs = next;
  return dofs;
}

void renumber_dofs(DofMap &dofs, const std::vector<int32_t> &new_index)
{
  for (auto &d : dofs.cell_dofs)
    d = new_index[d];
  for (auto &kv : dofs.entity_dof)
    kv.second = new_index[kv.second];
  std::vector<double> support(dofs.support.size());
  for (int32_t i = 0; i < dofs.n_dofs; ++i)
    for (int k = 0; k < dofs.dim; ++k)
      support[(size_t)new_index[i] * dofs.dim + k] = dofs.support[(size_t)i * dofs.dim + k];
  dofs.support.swap(support);
}

std::vector<int32_t> cuthill_mckee(const Csr &g, bool reversed)
{
  const int32_t n = g.n_rows;
  std::vector<int32_t> degree(n);
  for (int32_t i = 0; i < n; ++i)
    {
      int32_t d = 0;
      for (int32_t k = g.rowptr[i]; k < g.rowptr[i + 1]; ++k)
        d += g.col[k] != i;
      degree[i] = d;
    }
  std::vector<int32_t> order;
  order.reserve(n);
  std::vector<int32_t> level(n, -1);
  std::vector<char> visited(n, 0);
  std::vector<int32_t> queue, nbrs;

  // breadth-first level structure rooted at `root` over unvisited nodes;
  // returns the eccentricity and a minimum-degree node of the last level
  auto bfs_levels = [&](int32_t root, int32_t &last_min_degree_node) {
    queue.clear();
    queue.push_back(root);
    level[root] = 0;
    size_t head = 0;
    while (head < queue.size())
      {
        const int32_t u = queue[head++];
        for (int32_t k = g.rowptr[u]; k < g.rowptr[u + 1]; ++k)
          {
            const int32_t v = g.col[k];
            if (!visited[v] && level[v] < 0)
              {
                level[v] = level[u] + 1;
                queue.push_back(v);
              }
          }
      }
    const int32_t ecc = level[queue.back()];
    last_min_degree_node = queue.back();
    for (size_t i = queue.size(); i-- > 0 && level[queue[i]] == ecc;)
      if (degree[queue[i]] < degree[last_min_degree_node] ||
          (degree[queue[i]] == degree[last_min_degree_node] && queue[i] < last_min_degree_node))
        last_min_degree_node = queue[i];
    for (int32_t u : queue)
      level[u] = -1;
    return ecc;
  };

  int32_t scan = 0;
  while ((int32_t)order.size() < n)
    {
      // lowest-degree unvisited node of the next component
      while (visited[scan])
        ++scan;
      int32_t root = scan;
      {
        // restrict the search for the minimum degree to this component
        int32_t dummy;
        bfs_levels(root, dummy);
        for (int32_t u : queue)
          if (degree[u] < degree[root] || (degree[u] == degree[root] && u < root))
            root = u;
      }
      // George-Liu pseudo-peripheral node
      int32_t candidate;
      int32_t ecc = bfs_levels(root, candidate);
      for (;;)
        {
          int32_t next_candidate;
          const int32_t e2 = bfs_levels(candidate, next_candidate);
          if (e2 > ecc)
            {
              root = candidate;
              ecc = e2;
              candidate = next_candidate;
            }
          else
            break;
        }
      // Cuthill-McKee sweep
      size_t head = order.size();
      order.push_back(root);
      visited[root] = 1;
      while (head < order.size())
        {
          const int32_t u = order[head++];
          nbrs.clear();
          for (int32_t k = g.rowptr[u]; k < g.rowptr[u + 1]; ++k)
            {
              const int32_t v = g.col[k];
              if (!visited[v])
                {
                  visited[v] = 1;
                  nbrs.push_back(v);
                }
            }
          std::sort(nbrs.begin(), nbrs.end(), [&](int32_t a, int32_t b) {
            return degree[a] != degree[b] ? degree[a] < degree[b] : a < b;
          });
          order.insert(order.end(), nbrs.begin(), nbrs.end());
        }
    }
  std::vector<int32_t> new_index(n);
  for (int32_t pos = 0; pos < n; ++pos)
    new_index[order[pos]] = reversed ? n - 1 - pos : pos;
  return new_index;
}

Csr make_sparsity_pattern(const DofMap &rows, const DofMap &cols, int64_t n_cells, const std::vector<char> *constrained)
{
  Csr p;
  p.n_rows = rows.n_dofs;
  p.n_cols = cols.n_dofs;
  // row dof -> cells
  std::vector<int32_t> cptr(rows.n_dofs + 1, 0);
  for (size_t k = 0; k < rows.cell_dofs.size(); ++k)
    cptr[rows.cell_dofs[k] + 1]++;
  for (int32_t i = 0; i < rows.n_dofs; ++i)
    cptr[i + 1] += cptr[i];
  std::vector<int32_t> clist(rows.cell_dofs.size());
  {
    std::vector<int32_t> fill(cptr.begin(), cptr.end() - 1);
    for (int64_t c = 0; c < n_cells; ++c)
      for (int i = 0; i < rows.dpc; ++i)
        clist[fill[rows.cell_dofs[(size_t)c * rows.dpc + i]]++] = (int32_t)c;
  }
  p.rowptr.assign(rows.n_dofs + 1, 0);
  auto gather_row = [&](int32_t r, std::vector<int32_t> &buf) {
    buf.clear();
    if (constrained && (*constrained)[r])
      {
        buf.push_back(r);
        return;
      }
    for (int32_t k = cptr[r]; k < cptr[r + 1]; ++k)
      {
        const int32_t *cd = &cols.cell_dofs[(size_t)clist[k] * cols.dpc];
        buf.insert(buf.end(), cd, cd + cols.dpc);
      }
    std::sort(buf.begin(), buf.end());
    buf.erase(std::unique(buf.begin(), buf.end()), buf.end());
    if (constrained)
      buf.erase(std::remove_if(buf.begin(), buf.end(), [&](int32_t c) { return (*constrained)[c] != 0; }), buf.end());
  };
#pragma omp parallel
  {
    std::vector<int32_t> buf;
#pragma omp for schedule(static)
    for (int32_t r = 0; r < rows.n_dofs; ++r)
      {
        gather_row(r, buf);
        p.rowptr[r + 1] = (int32_t)buf.size();
      }
  }
  int64_t total = 0;
  for (int32_t r = 0; r < rows.n_dofs; ++r)
    {
      total += p.rowptr[r + 1];
      if (total > 0x7fffffffLL)
        throw std::runtime_error("sparsity pattern exceeds 2^31 entries (32-bit CSR offsets)");
      p.rowptr[r + 1] = (int32_t)total;
    }
  p.col.resize(total);
#pragma omp parallel
  {
    std::vector<int32_t> buf;
#pragma omp for schedule(static)
    for (int32_t r = 0; r < rows.n_dofs; ++r)
      {
        gather_row(r, buf);
        std::copy(buf.begin(), buf.end(), p.col.begin() + p.rowptr[r]);
      }
  }
  return p;
}

std::vector<int32_t> boundary_dofs(const Mesh &mesh, const DofMap &dofs, int boundary_id)
{
  const int fdim = mesh.dim - 1, vpf = mesh.verts_per_face(), n1 = dofs.degree + 1;
  const int npf = ipow(n1, fdim);
  std::vector<int32_t> out;
  for (int64_t f = 0; f < mesh.n_bfaces(); ++f)
    {
      if (mesh.bface_ids[f] != boundary_id)
        continue;
      const int32_t *fv = &mesh.bfaces[f * vpf];
      for (int i = 0; i < npf; ++i)
        {
          const int idx[3] = {i % n1, i / n1, 0};
          int32_t corners[4];
          const int n = node_corners(fv, fdim, dofs.degree, idx, corners);
          auto it = dofs.entity_dof.find(make_key(corners, n));
          if (it == dofs.entity_dof.end())
            throw std::runtime_error("boundary face does not belong to any cell");
          out.push_back(it->second);
        }
    }
  std::sort(out.begin(), out.end());
  out.erase(std::unique(out.begin(), out.end()), out.end());
  return out;
}

std::vector<int32_t> color_cells(const Mesh &mesh, int &n_colors)
{
  const int vpc = mesh.verts_per_cell();
  const int64_t n_cells = mesh.n_cells(), nv = mesh.n_vertices();
  std::vector<int32_t> vptr(nv + 1, 0);
  for (int32_t v : mesh.cells)
    vptr[v + 1]++;
  for (int64_t i = 0; i < nv; ++i)
    vptr[i + 1] += vptr[i];
  std::vector<int32_t> vcells(mesh.cells.size());
  {
    std::vector<int32_t> fill(vptr.begin(), vptr.end() - 1);
    for (int64_t c = 0; c < n_cells; ++c)
      for (int i = 0; i < vpc; ++i)
        vcells[fill[mesh.cells[c * vpc + i]]++] = (int32_t)c;
  }
  std::vector<int32_t> color(n_cells, -1);
  n_colors = 0;
  uint64_t used;
  for (int64_t c = 0; c < n_cells; ++c)
    {
      used = 0;
      for (int i = 0; i < vpc; ++i)
        {
          const int32_t v = mesh.cells[c * vpc + i];
          for (int32_t k = vptr[v]; k < vptr[v + 1]; ++k)
            if (color[vcells[k]] >= 0)
              used |= 1ull << color[vcells[k]];
        }
      int col = 0;
      while (used & (1ull << col))
        ++col;
      if (col >= 63)
        throw std::runtime_error("color_cells: more than 63 colours needed");
      color[c] = col;
      n_colors = std::max(n_colors, col + 1);
    }
  return color;
}

CellGeometry::CellGeometry(int dim_, int n_q_1d) : dim(dim_), nqp(ipow(n_q_1d, dim_)), q1(dim_, 1, n_q_1d)
{
  JxW.resize(nqp);
  invJT.resize((size_t)nqp * dim * dim);
}

void CellGeometry::reinit(const Mesh &mesh, int64_t cell)
{
  const int vpc = 1 << dim;
  const int32_t *cv = &mesh.cells[cell * vpc];
  for (int q = 0; q < nqp; ++q)
    {
      double J[3][3] = {{0, 0, 0}, {0, 0, 0}, {0, 0, 0}};
      for (int v = 0; v < vpc; ++v)
        for (int k = 0; k < dim; ++k)
          for (int d = 0; d < dim; ++d)
            J[k][d] += mesh.vertices[(size_t)cv[v] * dim + k] * q1.grad[((size_t)q * vpc + v) * dim + d];
      double det, inv[3][3];
      if (dim == 2)
        {
          det = J[0][0] * J[1][1] - J[0][1] * J[1][0];
          inv[0][0] = J[1][1] / det;
          inv[0][1] = -J[0][1] / det;
          inv[1][0] = -J[1][0] / det;
          inv[1][1] = J[0][0] / det;
        }
      else
        {
          const double c00 = J[1][1] * J[2][2] - J[1][2] * J[2][1];
          const double c01 = J[1][2] * J[2][0] - J[1][0] * J[2][2];
          const double c02 = J[1][0] * J[2][1] - J[1][1] * J[2][0];
          det = J[0][0] * c00 + J[0][1] * c01 + J[0][2] * c02;
          inv[0][0] = c00 / det;
          inv[1][0] = c01 / det;
          inv[2][0] = c02 / det;
          inv[0][1] = (J[0][2] * J[2][1] - J[0][1] * J[2][2]) / det;
          inv[1][1] = (J[0][0] * J[2][2] - J[0][2] * J[2][0]) / det;
          inv[2][1] = (J[0][1] * J[2][0] - J[0][0] * J[2][1]) / det;
          inv[0][2] = (J[0][1] * J[1][2] - J[0][2] * J[1][1]) / det;
          inv[1][2] = (J[0][2] * J[1][0] - J[0][0] * J[1][2]) / det;
          inv[2][2] = (J[0][0] * J[1][1] - J[0][1] * J[1][0]) / det;
        }
      JxW[q] = det * q1.weight[q];
      // physical gradient_k = sum_d inv[d][k] * reference gradient_d  (J^{-T})
      for (int k = 0; k < dim; ++k)
        for (int d = 0; d < dim; ++d)
          invJT[((size_t)q * dim + k) * dim + d] = inv[d][k];
    }
}

double inflow_velocity(int dim, const double *p, double width, double height)
{
  const double Um = 1.5;
  if (dim == 2)
    return 4.0 * Um * p[1] * (width - p[1]) / (width * width);
  // the reference divides by height^2 only (step-35a.cc:377-378)
  return 4.0 * Um * p[1] * (width - p[1]) * p[2] * (height - p[2]) / (height * height);
}

namespace {

std::vector<std::vector<int32_t>> cells_by_color(const std::vector<int32_t> &color, int n_colors)
{
  std::vector<std::vector<int32_t>> lists(n_colors);
  for (size_t c = 0; c < color.size(); ++c)
    lists[color[c]].push_back((int32_t)c);
  return lists;
}

void physical_gradients(const CellGeometry &geo, const ShapeTable &fe, int q, double *g /* [dpc][dim] */)
{
  const int dim = geo.dim;
  for (int i = 0; i < fe.dpc; ++i)
    for (int k = 0; k < dim; ++k)
      {
        double s = 0.0;
        for (int d = 0; d < dim; ++d)
          s += geo.invJT[((size_t)q * dim + k) * dim + d] * fe.grad[((size_t)q * fe.dpc + i) * dim + d];
        g[i * dim + k] = s;
      }
}

} // namespace

std::unique_ptr<Discretization> build_discretization(const Mesh &coarse, const DiscretizationOptions &opt)
{
  if (opt.pressure_degree != 1)
    throw std::runtime_error("pressure_fe_degree must be 1: this build implements the Q2/Q1 pair only");
  std::unique_ptr<Discretization> dp(new Discretization);
  Discretization &D = *dp;
  D.dim = coarse.dim;
  D.mesh = coarse;
  for (int r = 0; r < opt.n_refines; ++r)
    D.mesh = refine_uniform(D.mesh);
  const int dim = D.dim;
  const int64_t n_cells = D.mesh.n_cells();

  D.vel = distribute_dofs(D.mesh, opt.pressure_degree + 1);
  D.pres = distribute_dofs(D.mesh, opt.pressure_degree);
  if (opt.renumber)
    {
      renumber_dofs(D.vel, cuthill_mckee(make_sparsity_pattern(D.vel, D.vel, n_cells)));
      renumber_dofs(D.pres, cuthill_mckee(make_sparsity_pattern(D.pres, D.pres, n_cells)));
    }

  // the reference takes max y / max z over all vertices (step-35a.cc:789-798)
  for (int64_t v = 0; v < D.mesh.n_vertices(); ++v)
    {
      D.width = std::max(D.width, D.mesh.vertices[v * dim + 1]);
      if (dim == 3)
        D.height = std::max(D.height, D.mesh.vertices[v * dim + 2]);
    }

  // constraints
  {
    const std::vector<int32_t> inflow = boundary_dofs(D.mesh, D.vel, 2);
    const std::vector<int32_t> outflow = boundary_dofs(D.mesh, D.vel, 3);
    for (int d = 0; d < dim; ++d)
      {
        std::vector<std::pair<int32_t, double>> lines;
        for (int32_t i : inflow)
          lines.emplace_back(i, d == 0 ? inflow_velocity(dim, &D.vel.support[(size_t)i * dim], D.width, D.height) : 0.0);
        if (d != 0)
          for (int32_t i : outflow)
            if (!std::binary_search(inflow.begin(), inflow.end(), i))
              lines.emplace_back(i, 0.0);
        std::sort(lines.begin(), lines.end());
        for (auto &l : lines)
          {
            D.flow_dofs[d].push_back(l.first);
            D.flow_vals[d].push_back(l.second);
          }
      }
    D.wall_dofs = boundary_dofs(D.mesh, D.vel, 1);
    const std::vector<int32_t> obstacle = boundary_dofs(D.mesh, D.vel, 4);
    D.wall_dofs.insert(D.wall_dofs.end(), obstacle.begin(), obstacle.end());
    std::sort(D.wall_dofs.begin(), D.wall_dofs.end());
    D.wall_dofs.erase(std::unique(D.wall_dofs.begin(), D.wall_dofs.end()), D.wall_dofs.end());
    D.proj_dofs = boundary_dofs(D.mesh, D.pres, 3);
  }

  // patterns
  std::vector<char> wall_mask(D.vel.n_dofs, 0), proj_mask(D.pres.n_dofs, 0);
  for (int32_t i : D.wall_dofs)
    wall_mask[i] = 1;
  for (int32_t i : D.proj_dofs)
    proj_mask[i] = 1;
  D.vel_pattern_unconstrained = make_sparsity_pattern(D.vel, D.vel, n_cells);
  D.vel_pattern_constrained = make_sparsity_pattern(D.vel, D.vel, n_cells, &wall_mask);
  D.grad_pattern = make_sparsity_pattern(D.vel, D.pres, n_cells);
  D.proj_pattern = make_sparsity_pattern(D.pres, D.pres, n_cells, &proj_mask);
  D.pres_pattern = make_sparsity_pattern(D.pres, D.pres, n_cells);

  D.cell_color = color_cells(D.mesh, D.n_colors);
  const auto color_lists = cells_by_color(D.cell_color, D.n_colors);

  // one-off operators.  Cells of one colour share no DoF, so they can be
  // scattered concurrently; colours run in sequence, which fixes the order of
  // the floating-point sums.
  D.vel_mass.assign(D.vel_pattern_unconstrained.nnz(), 0.0);
  D.pres_laplace.assign(D.proj_pattern.nnz(), 0.0);
  D.pres_mass.assign(D.pres_pattern.nnz(), 0.0);
  for (int d = 0; d < dim; ++d)
    D.pres_diff[d].assign(D.grad_pattern.nnz(), 0.0);

  const int nq_v = opt.pressure_degree + 2, nq_p = opt.pressure_degree + 1;
  const ShapeTable fe_v(dim, D.vel.degree, nq_v);   // velocity at quadrature_velocity
  const ShapeTable fe_pv(dim, D.pres.degree, nq_v); // pressure at quadrature_velocity (gradient operator)
  const ShapeTable fe_p(dim, D.pres.degree, nq_p);  // pressure at quadrature_pressure
  const int vdpc = fe_v.dpc, pdpc = fe_p.dpc;

  for (const auto &cells : color_lists)
    {
#pragma omp parallel
      {
        CellGeometry geo_v(dim, nq_v), geo_p(dim, nq_p);
        std::vector<double> mass(vdpc * vdpc), grad(vdpc * pdpc * dim), lap(pdpc * pdpc), pmass(pdpc * pdpc);
        std::vector<double> gv(vdpc * dim), gp(pdpc * dim);
#pragma omp for schedule(static)
        for (size_t ci = 0; ci < cells.size(); ++ci)
          {
            const int64_t c = cells[ci];
            const int32_t *vd = &D.vel.cell_dofs[(size_t)c * vdpc];
            const int32_t *pd = &D.pres.cell_dofs[(size_t)c * pdpc];
            geo_v.reinit(D.mesh, c);
            geo_p.reinit(D.mesh, c);
            std::fill(mass.begin(), mass.end(), 0.0);
            std::fill(grad.begin(), grad.end(), 0.0);
            std::fill(lap.begin(), lap.end(), 0.0);
            std::fill(pmass.begin(), pmass.end(), 0.0);
            // vel_Mass (MatrixCreator::create_mass_matrix, step-35a.cc:986-989) and
            // pres_Diff[d](i,j) = -JxW * d_d phi^v_i * phi^p_j (step-35a.cc:1143-1150)
            for (int q = 0; q < fe_v.nqp; ++q)
              {
                physical_gradients(geo_v, fe_v, q, gv.data());
                const double jxw = geo_v.JxW[q];
                const double *phi = &fe_v.value[(size_t)q * vdpc];
                const double *psi = &fe_pv.value[(size_t)q * pdpc];
                for (int i = 0; i < vdpc; ++i)
                  {
                    for (int j = 0; j < vdpc; ++j)
                      mass[i * vdpc + j] += phi[i] * phi[j] * jxw;
                    for (int j = 0; j < pdpc; ++j)
                      for (int d = 0; d < dim; ++d)
                        grad[(d * vdpc + i) * pdpc + j] += -jxw * gv[i * dim + d] * psi[j];
                  }
              }
            // pres_Laplace / pres_Mass (step-35a.cc:1035-1058)
            for (int q = 0; q < fe_p.nqp; ++q)
              {
                physical_gradients(geo_p, fe_p, q, gp.data());
                const double jxw = geo_p.JxW[q];
                const double *psi = &fe_p.value[(size_t)q * pdpc];
                for (int i = 0; i < pdpc; ++i)
                  for (int j = 0; j < pdpc; ++j)
                    {
                      double gg = 0.0;
                      for (int d = 0; d < dim; ++d)
                        gg += gp[i * dim + d] * gp[j * dim + d];
                      lap[i * pdpc + j] += gg * jxw;
                      pmass[i * pdpc + j] += psi[i] * psi[j] * jxw;
                    }
              }
            for (int i = 0; i < vdpc; ++i)
              {
                for (int j = 0; j < vdpc; ++j)
                  D.vel_mass[D.vel_pattern_unconstrained.find(vd[i], vd[j])] += mass[i * vdpc + j];
                for (int j = 0; j < pdpc; ++j)
                  {
                    const int32_t slot = D.grad_pattern.find(vd[i], pd[j]);
                    for (int d = 0; d < dim; ++d)
                      D.pres_diff[d][slot] += grad[(d * vdpc + i) * pdpc + j];
                  }
              }
            // ConstraintMatrix::distribute_local_to_global with homogeneous
            // Dirichlet constraints: constrained rows/columns are dropped and
            // |A_loc(i,i)| goes to the diagonal (mean |diag| if that is zero).
            double avg_diag = 0.0;
            for (int i = 0; i < pdpc; ++i)
              avg_diag += std::fabs(lap[i * pdpc + i]);
            avg_diag /= pdpc;
            for (int i = 0; i < pdpc; ++i)
              {
                if (proj_mask[pd[i]])
                  {
                    const double a = std::fabs(lap[i * pdpc + i]);
                    D.pres_laplace[D.proj_pattern.find(pd[i], pd[i])] += (a != 0.0 ? a : avg_diag);
                  }
                else
                  for (int j = 0; j < pdpc; ++j)
                    if (!proj_mask[pd[j]])
                      D.pres_laplace[D.proj_pattern.find(pd[i], pd[j])] += lap[i * pdpc + j];
                for (int j = 0; j < pdpc; ++j)
                  D.pres_mass[D.pres_pattern.find(pd[i], pd[j])] += pmass[i * pdpc + j];
              }
          }
      }
    }

  // initial conditions: p = 25 - x at t0 and t0 + dt (no time dependence)
  D.pres_n0.resize(D.pres.n_dofs);
  for (int32_t i = 0; i < D.pres.n_dofs; ++i)
    D.pres_n0[i] = 25.0 - D.pres.support[(size_t)i * dim];
  D.pres_nm1_0 = D.pres_n0;
  (void)opt.dt;
  return dp;
}

} // namespace s35
