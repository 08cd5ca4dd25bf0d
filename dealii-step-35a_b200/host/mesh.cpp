#include "mesh.h"

#include <algorithm>
#include <array>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <fstream>
#include <map>
#include <sstream>
#include <stdexcept>
#include <unordered_map>

namespace s35 {
namespace {

// Sub-entity key: sorted vertex ids, padded with -1 (edge: 2 ids, quad: 4 ids).
struct EntityKey
{
  int32_t v[4];
  bool operator==(const EntityKey &o) const { return std::memcmp(v, o.v, sizeof(v)) == 0; }
};
struct EntityKeyHash
{
  size_t operator()(const EntityKey &k) const
  {
    uint64_t h = 0x9E3779B97F4A7C15ull;
    for (int i = 0; i < 4; ++i)
      {
        h ^= (uint64_t)(uint32_t)k.v[i] + 0x9E3779B97F4A7C15ull + (h << 6) + (h >> 2);
        h *= 0xff51afd7ed558ccdull;
      }
    return (size_t)(h ^ (h >> 32));
  }
};

EntityKey make_key(const int32_t *ids, int n)
{
  EntityKey k{{-1, -1, -1, -1}};
  for (int i = 0; i < n; ++i)
    k.v[i] = ids[i];
  std::sort(k.v, k.v + n);
  return k;
}

// gmsh / UCD order vertices counter-clockwise (bottom face first, then top);
// lexicographic order swaps the last two of every group of four.
const int kCcwToLex2[4] = {0, 1, 3, 2};
const int kCcwToLex3[8] = {0, 1, 3, 2, 4, 5, 7, 6};

struct RawElement
{
  int type; // 1 line, 3 quad, 5 hex (gmsh codes)
  int tag;
  std::vector<int64_t> nodes;
};

Mesh build_mesh(int dim, const std::map<int64_t, std::array<double, 3>> &nodes,
                const std::vector<RawElement> &elements)
{
  const int cell_type = dim == 3 ? 5 : 3;
  const int face_type = dim == 3 ? 3 : 1;
  Mesh mesh;
  mesh.dim = dim;
  // keep only vertices used by cells, in file order (GridIn drops unused ones)
  std::map<int64_t, int32_t> used;
  for (const auto &e : elements)
    if (e.type == cell_type)
      for (int64_t n : e.nodes)
        used.emplace(n, 0);
  int32_t next = 0;
  for (auto &kv : used)
    {
      auto it = nodes.find(kv.first);
      if (it == nodes.end())
        throw std::runtime_error("mesh: cell refers to unknown vertex id");
      kv.second = next++;
      for (int d = 0; d < dim; ++d)
        mesh.vertices.push_back(it->second[d]);
    }
  for (const auto &e : elements)
    {
      if (e.type == cell_type)
        {
          const int n = 1 << dim;
          if ((int)e.nodes.size() != n)
            throw std::runtime_error("mesh: wrong vertex count in cell");
          for (int i = 0; i < n; ++i)
            {
              const int src = dim == 3 ? kCcwToLex3[i] : kCcwToLex2[i];
              mesh.cells.push_back(used.at(e.nodes[src]));
            }
        }
      else if (e.type == face_type)
        {
          const int n = 1 << (dim - 1);
          if ((int)e.nodes.size() != n)
            throw std::runtime_error("mesh: wrong vertex count in boundary face");
          bool ok = true;
          for (int i = 0; i < n; ++i)
            ok = ok && used.count(e.nodes[dim == 3 ? kCcwToLex2[i] : i]);
          if (!ok)
            continue;
          for (int i = 0; i < n; ++i)
            mesh.bfaces.push_back(used.at(e.nodes[dim == 3 ? kCcwToLex2[i] : i]));
          mesh.bface_ids.push_back(e.tag);
        }
    }
  if (mesh.cells.empty())
    throw std::runtime_error("mesh: no cells of the requested dimension found");
  check_orientation(mesh);
  return mesh;
}

double det_at_vertex(const Mesh &m, const int32_t *c, int corner)
{
  const int dim = m.dim;
  double J[3][3];
  for (int d = 0; d < dim; ++d)
    {
      const int other = corner ^ (1 << d);
      const double sign = (corner >> d) & 1 ? -1.0 : 1.0;
      for (int k = 0; k < dim; ++k)
        J[k][d] = sign * (m.vertices[(size_t)c[other] * dim + k] - m.vertices[(size_t)c[corner] * dim + k]);
    }
  if (dim == 2)
    return J[0][0] * J[1][1] - J[0][1] * J[1][0];
  return J[0][0] * (J[1][1] * J[2][2] - J[1][2] * J[2][1]) - J[0][1] * (J[1][0] * J[2][2] - J[1][2] * J[2][0]) +
         J[0][2] * (J[1][0] * J[2][1] - J[1][1] * J[2][0]);
}

} // namespace

void check_orientation(const Mesh &mesh)
{
  const int vpc = mesh.verts_per_cell();
  for (int64_t c = 0; c < mesh.n_cells(); ++c)
    for (int v = 0; v < vpc; ++v)
      if (!(det_at_vertex(mesh, &mesh.cells[c * vpc], v) > 0.0))
        {
          std::ostringstream msg;
          msg << "mesh: cell " << c << " has a non-positive Jacobian at vertex " << v;
          throw std::runtime_error(msg.str());
        }
}

Mesh read_ucd(const std::string &filename, int dim)
{
  std::ifstream in(filename);
  if (!in)
    throw std::runtime_error("cannot open file " + filename);
  std::string line;
  // skip comment lines
  do
    {
      if (!std::getline(in, line))
        throw std::runtime_error("ucd: empty file " + filename);
    }
  while (!line.empty() && line[0] == '#');
  std::istringstream header(line);
  long n_vertices = 0, n_elements = 0;
  header >> n_vertices >> n_elements;
  std::map<int64_t, std::array<double, 3>> nodes;
  for (long i = 0; i < n_vertices; ++i)
    {
      int64_t id;
      std::array<double, 3> x{{0, 0, 0}};
      in >> id >> x[0] >> x[1] >> x[2];
      nodes[id] = x;
    }
  std::vector<RawElement> elements;
  for (long i = 0; i < n_elements; ++i)
    {
      int64_t id;
      int material;
      std::string type;
      in >> id >> material >> type;
      if (!in)
        throw std::runtime_error("ucd: truncated element list in " + filename);
      RawElement e;
      e.tag = material;
      int n = 0;
      if (type == "line")
        e.type = 1, n = 2;
      else if (type == "quad")
        e.type = 3, n = 4;
      else if (type == "hex")
        e.type = 5, n = 8;
      else
        throw std::runtime_error("ucd: unknown element type " + type);
      e.nodes.resize(n);
      for (int k = 0; k < n; ++k)
        in >> e.nodes[k];
      elements.push_back(std::move(e));
    }
  return build_mesh(dim, nodes, elements);
}

Mesh read_msh(const std::string &filename, int dim)
{
  std::ifstream in(filename);
  if (!in)
    throw std::runtime_error("cannot open file " + filename);
  std::string tok;
  std::map<int64_t, std::array<double, 3>> nodes;
  std::vector<RawElement> elements;
  while (in >> tok)
    {
      if (tok == "$MeshFormat")
        {
          double version;
          int file_type, data_size;
          in >> version >> file_type >> data_size;
          if (version < 2.0 || version >= 3.0 || file_type != 0)
            throw std::runtime_error("msh: only gmsh 2.x ASCII files are supported");
        }
      else if (tok == "$Nodes")
        {
          long n;
          in >> n;
          for (long i = 0; i < n; ++i)
            {
              int64_t id;
              std::array<double, 3> x;
              in >> id >> x[0] >> x[1] >> x[2];
              nodes[id] = x;
            }
        }
      else if (tok == "$Elements")
        {
          long n;
          in >> n;
          for (long i = 0; i < n; ++i)
            {
              int64_t id;
              int type, ntags;
              in >> id >> type >> ntags;
              RawElement e;
              e.type = type;
              e.tag = 0;
              for (int t = 0; t < ntags; ++t)
                {
                  int tag;
                  in >> tag;
                  if (t == 0)
                    e.tag = tag;
                }
              int nn = 0;
              switch (type)
                {
                case 1: nn = 2; break;
                case 3: nn = 4; break;
                case 5: nn = 8; break;
                case 15: nn = 1; break;
                default: throw std::runtime_error("msh: unsupported element type");
                }
              e.nodes.resize(nn);
              for (int k = 0; k < nn; ++k)
                in >> e.nodes[k];
              if (!in)
                throw std::runtime_error("msh: truncated element list in " + filename);
              elements.push_back(std::move(e));
            }
        }
    }
  return build_mesh(dim, nodes, elements);
}

static const char kMagic[8] = {'S', '3', '5', 'M', 'E', 'S', 'H', '1'};

void write_s35m(const Mesh &mesh, const std::string &filename)
{
  std::ofstream out(filename, std::ios::binary);
  if (!out)
    throw std::runtime_error("cannot open file " + filename);
  int64_t hdr[4] = {mesh.dim, mesh.n_vertices(), mesh.n_cells(), mesh.n_bfaces()};
  out.write(kMagic, 8);
  out.write((const char *)hdr, sizeof(hdr));
  out.write((const char *)mesh.vertices.data(), mesh.vertices.size() * sizeof(double));
  out.write((const char *)mesh.cells.data(), mesh.cells.size() * sizeof(int32_t));
  out.write((const char *)mesh.bfaces.data(), mesh.bfaces.size() * sizeof(int32_t));
  out.write((const char *)mesh.bface_ids.data(), mesh.bface_ids.size() * sizeof(int32_t));
}

Mesh read_s35m(const std::string &filename)
{
  std::ifstream in(filename, std::ios::binary);
  if (!in)
    throw std::runtime_error("cannot open file " + filename);
  char magic[8];
  int64_t hdr[4];
  in.read(magic, 8);
  in.read((char *)hdr, sizeof(hdr));
  if (!in || std::memcmp(magic, kMagic, 8) != 0)
    throw std::runtime_error("not an s35m mesh file: " + filename);
  Mesh mesh;
  mesh.dim = (int)hdr[0];
  if (mesh.dim != 2 && mesh.dim != 3)
    throw std::runtime_error("s35m: bad dimension");
  mesh.vertices.resize(hdr[1] * mesh.dim);
  mesh.cells.resize(hdr[2] * mesh.verts_per_cell());
  mesh.bfaces.resize(hdr[3] * mesh.verts_per_face());
  mesh.bface_ids.resize(hdr[3]);
  in.read((char *)mesh.vertices.data(), mesh.vertices.size() * sizeof(double));
  in.read((char *)mesh.cells.data(), mesh.cells.size() * sizeof(int32_t));
  in.read((char *)mesh.bfaces.data(), mesh.bfaces.size() * sizeof(int32_t));
  in.read((char *)mesh.bface_ids.data(), mesh.bface_ids.size() * sizeof(int32_t));
  if (!in)
    throw std::runtime_error("s35m: truncated file " + filename);
  check_orientation(mesh);
  return mesh;
}

void write_ucd(const Mesh &mesh, const std::string &filename)
{
  std::ofstream out(filename);
  if (!out)
    throw std::runtime_error("cannot open file " + filename);
  out.precision(17);
  out << mesh.n_vertices() << " " << mesh.n_cells() + mesh.n_bfaces() << " 0 0 0\n";
  for (int64_t v = 0; v < mesh.n_vertices(); ++v)
    {
      out << v;
      for (int d = 0; d < 3; ++d)
        out << " " << (d < mesh.dim ? mesh.vertices[v * mesh.dim + d] : 0.0);
      out << "\n";
    }
  const int vpc = mesh.verts_per_cell(), vpf = mesh.verts_per_face();
  int64_t id = 1;
  for (int64_t c = 0; c < mesh.n_cells(); ++c, ++id)
    {
      out << id << " 1 " << (mesh.dim == 3 ? "hex" : "quad");
      for (int i = 0; i < vpc; ++i)
        out << " " << mesh.cells[c * vpc + (mesh.dim == 3 ? kCcwToLex3[i] : kCcwToLex2[i])];
      out << "\n";
    }
  for (int64_t f = 0; f < mesh.n_bfaces(); ++f, ++id)
    {
      out << id << " " << mesh.bface_ids[f] << " " << (mesh.dim == 3 ? "quad" : "line");
      for (int i = 0; i < vpf; ++i)
        out << " " << mesh.bfaces[f * vpf + (mesh.dim == 3 ? kCcwToLex2[i] : i)];
      out << "\n";
    }
}

void write_msh(const Mesh &mesh, const std::string &filename)
{
  std::ofstream out(filename);
  if (!out)
    throw std::runtime_error("cannot open file " + filename);
  out.precision(17);
  out << "$MeshFormat\n2.2 0 8\n$EndMeshFormat\n$Nodes\n" << mesh.n_vertices() << "\n";
  for (int64_t v = 0; v < mesh.n_vertices(); ++v)
    {
      out << v + 1;
      for (int d = 0; d < 3; ++d)
        out << " " << (d < mesh.dim ? mesh.vertices[v * mesh.dim + d] : 0.0);
      out << "\n";
    }
  out << "$EndNodes\n$Elements\n" << mesh.n_cells() + mesh.n_bfaces() << "\n";
  const int vpc = mesh.verts_per_cell(), vpf = mesh.verts_per_face();
  int64_t id = 1;
  for (int64_t f = 0; f < mesh.n_bfaces(); ++f, ++id)
    {
      out << id << " " << (mesh.dim == 3 ? 3 : 1) << " 2 " << mesh.bface_ids[f] << " 0";
      for (int i = 0; i < vpf; ++i)
        out << " " << mesh.bfaces[f * vpf + (mesh.dim == 3 ? kCcwToLex2[i] : i)] + 1;
      out << "\n";
    }
  for (int64_t c = 0; c < mesh.n_cells(); ++c, ++id)
    {
      out << id << " " << (mesh.dim == 3 ? 5 : 3) << " 2 0 0";
      for (int i = 0; i < vpc; ++i)
        out << " " << mesh.cells[c * vpc + (mesh.dim == 3 ? kCcwToLex3[i] : kCcwToLex2[i])] + 1;
      out << "\n";
    }
  out << "$EndElements\n";
}

Mesh read_mesh(const std::string &filename, int dim)
{
  auto ends_with = [&](const char *suffix) {
    const size_t n = std::strlen(suffix);
    return filename.size() >= n && filename.compare(filename.size() - n, n, suffix) == 0;
  };
  if (ends_with(".inp"))
    return read_ucd(filename, dim);
  if (ends_with(".msh"))
    return read_msh(filename, dim);
  if (ends_with(".s35m"))
    {
      Mesh m = read_s35m(filename);
      if (m.dim != dim)
        throw std::runtime_error("mesh dimension does not match the requested one: " + filename);
      return m;
    }
  throw std::runtime_error("unknown mesh file extension: " + filename);
}

Mesh refine_uniform(const Mesh &coarse)
{
  const int dim = coarse.dim;
  const int vpc = coarse.verts_per_cell();
  const int vpf = coarse.verts_per_face();
  Mesh fine;
  fine.dim = dim;
  fine.vertices = coarse.vertices;
  fine.cells.reserve(coarse.cells.size() * vpc);
  fine.bfaces.reserve(coarse.bfaces.size() * vpf);
  fine.bface_ids.reserve(coarse.bface_ids.size() * vpf);

  std::unordered_map<EntityKey, int32_t, EntityKeyHash> entity_vertex;
  entity_vertex.reserve((size_t)coarse.n_cells() * (dim == 3 ? 7 : 3));

  // vertex for the node with multi-index idx[d] in {0,1,2} of an entity whose
  // corner vertices are ev[0 .. 2^edim) (lexicographic)
  auto node_vertex = [&](const int32_t *ev, int edim, const int *idx) -> int32_t {
    int32_t corners[8];
    int n = 0;
    for (int c = 0; c < (1 << edim); ++c)
      {
        bool inside = true;
        for (int d = 0; d < edim; ++d)
          if (idx[d] != 1 && ((c >> d) & 1) != idx[d] / 2)
            inside = false;
        if (inside)
          corners[n++] = ev[c];
      }
    if (n == 1)
      return corners[0];
    auto add_mean = [&]() {
      const int32_t id = (int32_t)(fine.vertices.size() / dim);
      for (int k = 0; k < dim; ++k)
        {
          double s = 0.0;
          for (int i = 0; i < n; ++i)
            s += coarse.vertices[(size_t)corners[i] * dim + k];
          fine.vertices.push_back(s / n);
        }
      return id;
    };
    if (n == 8)
      return add_mean(); // cell interior: unique, no lookup
    const EntityKey key = make_key(corners, n);
    auto it = entity_vertex.find(key);
    if (it != entity_vertex.end())
      return it->second;
    const int32_t id = add_mean();
    entity_vertex.emplace(key, id);
    return id;
  };

  int32_t grid[27];
  for (int64_t c = 0; c < coarse.n_cells(); ++c)
    {
      const int32_t *cv = &coarse.cells[c * vpc];
      const int nn = dim == 3 ? 27 : 9;
      for (int n = 0; n < nn; ++n)
        {
          const int idx[3] = {n % 3, (n / 3) % 3, n / 9};
          grid[n] = node_vertex(cv, dim, idx);
        }
      for (int child = 0; child < vpc; ++child)
        for (int v = 0; v < vpc; ++v)
          {
            const int ix = (child & 1) + (v & 1);
            const int iy = ((child >> 1) & 1) + ((v >> 1) & 1);
            const int iz = ((child >> 2) & 1) + ((v >> 2) & 1);
            fine.cells.push_back(grid[ix + 3 * iy + 9 * iz]);
          }
    }
  for (int64_t f = 0; f < coarse.n_bfaces(); ++f)
    {
      const int32_t *fv = &coarse.bfaces[f * vpf];
      const int nn = dim == 3 ? 9 : 3;
      for (int n = 0; n < nn; ++n)
        {
          const int idx[3] = {n % 3, n / 3, 0};
          grid[n] = node_vertex(fv, dim - 1, idx);
        }
      for (int child = 0; child < vpf; ++child)
        {
          for (int v = 0; v < vpf; ++v)
            {
              const int ix = (child & 1) + (v & 1);
              const int iy = ((child >> 1) & 1) + ((v >> 1) & 1);
              fine.bfaces.push_back(grid[ix + 3 * iy]);
            }
          fine.bface_ids.push_back(coarse.bface_ids[f]);
        }
    }
  return fine;
}

double minimal_cell_diameter(const Mesh &mesh)
{
  const int dim = mesh.dim, vpc = mesh.verts_per_cell();
  double best = 1e300;
  for (int64_t c = 0; c < mesh.n_cells(); ++c)
    {
      // deal.II's cell diameter: the longest of the space diagonals
      double diam = 0.0;
      for (int a = 0; a < vpc / 2; ++a)
        {
          const int b = (vpc - 1) ^ a;
          double s = 0.0;
          for (int k = 0; k < dim; ++k)
            {
              const double d = mesh.vertices[(size_t)mesh.cells[c * vpc + a] * dim + k] -
                               mesh.vertices[(size_t)mesh.cells[c * vpc + b] * dim + k];
              s += d * d;
            }
          diam = std::max(diam, s);
        }
      best = std::min(best, diam);
    }
  return std::sqrt(best);
}

} // namespace s35
