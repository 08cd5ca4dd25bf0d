// Host scaffolding: hexahedral / quadrilateral meshes for the step-35a hot path.
//
// Replaces what the reference gets from deal.II's GridIn + Triangulation
// (reference step-35a.cc:744-756: read_msh / read_ucd, refine_global) with a
// flat-array mesh that can be handed to the device as is.  No manifold is
// attached in the reference, so refinement is straight-sided (edge midpoints,
// face and cell centroids).
//
// Cell vertices are stored in lexicographic order (bit0 = x, bit1 = y,
// bit2 = z of the reference cell), which is what the d-linear mapping and the
// tensor-product shape tables in fe.h expect.
#pragma once
#include <cstdint>
#include <string>
#include <vector>

namespace s35 {

struct Mesh
{
  int dim = 0;
  std::vector<double> vertices;   // n_vertices * dim
  std::vector<int32_t> cells;     // n_cells * 2^dim, lexicographic
  std::vector<int32_t> bfaces;    // n_bfaces * 2^(dim-1) vertex ids (lexicographic on the face)
  std::vector<int32_t> bface_ids; // n_bfaces boundary indicators

  int verts_per_cell() const { return 1 << dim; }
  int verts_per_face() const { return 1 << (dim - 1); }
  int64_t n_vertices() const { return (int64_t)vertices.size() / dim; }
  int64_t n_cells() const { return (int64_t)cells.size() / verts_per_cell(); }
  int64_t n_bfaces() const { return (int64_t)bface_ids.size(); }
};

// AVS-UCD reader (nsbench2.inp; reference step-35a.cc:747,752 `read_ucd`).
// dim = 2: `quad` lines are cells, `line` lines carry the boundary id in their
// material column.  dim = 3: `hex` cells, `quad` boundary faces.
Mesh read_ucd(const std::string &filename, int dim);

// gmsh 2.2 ASCII reader (cylinder3d.msh, cylinderextruded.msh; reference
// step-35a.cc:748,751 `read_msh`).  Element type 5 = hex, 3 = quad, 1 = line;
// the first tag is the physical id = boundary indicator for faces.
Mesh read_msh(const std::string &filename, int dim);

// Compact binary container used to ship meshes with the repository.
Mesh read_s35m(const std::string &filename);
void write_s35m(const Mesh &mesh, const std::string &filename);
// Writers for the two text formats (used for reader round-trip tests).
void write_ucd(const Mesh &mesh, const std::string &filename);
void write_msh(const Mesh &mesh, const std::string &filename);

// Dispatch on the file extension (.inp, .msh, .s35m).
Mesh read_mesh(const std::string &filename, int dim);

// One step of uniform refinement (Triangulation::refine_global(1) with a flat
// manifold): every cell is split into 2^dim children, boundary faces inherit
// the boundary id of their parent.
Mesh refine_uniform(const Mesh &coarse);

// Smallest cell diameter (longest space diagonal of a cell), as printed by the
// reference at step-35a.cc:757-759.
double minimal_cell_diameter(const Mesh &mesh);

// Throws std::runtime_error if a cell has a non-positive Jacobian at a vertex.
void check_orientation(const Mesh &mesh);

} // namespace s35
