// Drag and lift coefficients of the obstacle.  The reference computes no force
// functional at all; the north star asks for one, so it is
// defined here once and evaluated by the same code on device results and on
// oracle results:
//
//   F = int_S sigma n dS,  sigma = -p I + (1/Re)(grad u + grad u^T),
//   n = unit normal on the obstacle surface S pointing into the fluid,
//   c_drag = 2 F_x / (Ubar^2 A),  c_lift = 2 F_y / (Ubar^2 A),
//
// S: boundary faces with indicator 4 (nsbench2.inp, cylinderextruded.msh); if the mesh has none
// (cylinder3d.msh marks its square obstacle with the wall indicator 1) the indicator-1 faces whose
// vertices all lie in the box [2.25,3.25] x [2,3].  Ubar is the mean of the inflow profile the reference
// actually prescribes (step-35a.cc:373, 377-378): 2 Um / 3 in 2D and 4 Um w^2 / 36 in 3D, Um = 1.5.
// A is the frontal area of S: its y-extent, times its z-extent in 3D.
// Quadrature: 3-point Gauss rule per face direction; gradients from the Q2 shape functions of the
// adjacent cell under the d-linear mapping.
#include "discretization.h"

#include <algorithm>
#include <array>
#include <cmath>
#include <map>
#include <stdexcept>

namespace s35 {

void drag_lift(const Discretization &D, double reynolds, const double *u, const double *p, double *out)
{
  const int dim = D.dim, vpc = 1 << dim, vpf = 1 << (dim - 1);
  const Mesh &m = D.mesh;
  // obstacle faces
  bool has4 = false;
  for (int32_t id : m.bface_ids)
    has4 = has4 || id == 4;
  std::map<std::array<int32_t, 4>, int> obstacle; // sorted vertex tuple -> 1
  double ymin = 1e300, ymax = -1e300, zmin = 1e300, zmax = -1e300;
  for (int64_t f = 0; f < m.n_bfaces(); ++f)
    {
      bool take = false;
      if (has4)
        take = m.bface_ids[f] == 4;
      else if (m.bface_ids[f] == 1)
        {
          take = true;
          for (int v = 0; v < vpf; ++v)
            {
              const double *x = &m.vertices[(size_t)m.bfaces[f * vpf + v] * dim];
              if (x[0] < 2.25 - 1e-9 || x[0] > 3.25 + 1e-9 || x[1] < 2.0 - 1e-9 || x[1] > 3.0 + 1e-9)
                take = false;
            }
        }
      if (!take)
        continue;
      std::array<int32_t, 4> key{{-1, -1, -1, -1}};
      for (int v = 0; v < vpf; ++v)
        {
          key[v] = m.bfaces[f * vpf + v];
          const double *x = &m.vertices[(size_t)key[v] * dim];
          ymin = std::min(ymin, x[1]);
          ymax = std::max(ymax, x[1]);
          if (dim == 3)
            {
              zmin = std::min(zmin, x[2]);
              zmax = std::max(zmax, x[2]);
            }
        }
      std::sort(key.begin(), key.begin() + vpf);
      obstacle[key] = 1;
    }
  if (obstacle.empty())
    throw std::runtime_error("drag_lift: the mesh has no obstacle faces");

  const int nq1 = 3;
  std::vector<double> xq, wq;
  gauss_rule(nq1, xq, wq);
  const int vdpc = D.vel.dpc, pdpc = D.pres.dpc;
  const int32_t nv = D.vel.n_dofs;
  double F[3] = {0, 0, 0};

  for (int64_t c = 0; c < m.n_cells(); ++c)
    for (int face = 0; face < 2 * dim; ++face)
      {
        const int dir = face / 2, side = face % 2;
        std::array<int32_t, 4> key{{-1, -1, -1, -1}};
        int nk = 0;
        for (int v = 0; v < vpc; ++v)
          if (((v >> dir) & 1) == side)
            key[nk++] = m.cells[c * vpc + v];
        std::sort(key.begin(), key.begin() + vpf);
        if (!obstacle.count(key))
          continue;
        const int nfq = dim == 3 ? nq1 * nq1 : nq1;
        for (int q = 0; q < nfq; ++q)
          {
            double xi[3] = {0, 0, 0};
            double w = 1.0;
            int t = q;
            for (int d = 0; d < dim; ++d)
              {
                if (d == dir)
                  xi[d] = side;
                else
                  {
                    xi[d] = xq[t % nq1];
                    w *= wq[t % nq1];
                    t /= nq1;
                  }
              }
            // Jacobian of the d-linear map at xi
            double J[3][3] = {{0, 0, 0}, {0, 0, 0}, {0, 0, 0}};
            for (int v = 0; v < vpc; ++v)
              for (int d = 0; d < dim; ++d)
                {
                  double g = 1.0;
                  for (int e = 0; e < dim; ++e)
                    {
                      const int bit = (v >> e) & 1;
                      g *= (e == d) ? (bit ? 1.0 : -1.0) : (bit ? xi[e] : 1.0 - xi[e]);
                    }
                  for (int k = 0; k < dim; ++k)
                    J[k][d] += m.vertices[(size_t)m.cells[c * vpc + v] * dim + k] * g;
                }
            double inv[3][3], det;
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
            // outward normal of the cell through this face: J^{-T} e_dir * (+-1), not normalised;
            // |det J| * |J^{-T} n_ref| is the surface element
            double nrm[3] = {0, 0, 0}, len = 0.0;
            for (int k = 0; k < dim; ++k)
              {
                nrm[k] = inv[dir][k] * (side ? 1.0 : -1.0);
                len += nrm[k] * nrm[k];
              }
            len = std::sqrt(len);
            const double dS = std::fabs(det) * len * w;
            double nb[3]; // unit normal pointing from the body into the fluid = minus the cell's outward normal
            for (int k = 0; k < dim; ++k)
              nb[k] = -nrm[k] / len;
            // velocity gradient and pressure at xi
            double gu[3][3] = {{0, 0, 0}, {0, 0, 0}, {0, 0, 0}};
            for (int i = 0; i < vdpc; ++i)
              {
                const int ii[3] = {i % 3, (i / 3) % 3, i / 9};
                double gref[3];
                for (int d = 0; d < dim; ++d)
                  {
                    double g = 1.0;
                    for (int e = 0; e < dim; ++e)
                      {
                        double val, der;
                        lagrange_1d(2, ii[e], xi[e], val, der);
                        g *= (e == d) ? der : val;
                      }
                    gref[d] = g;
                  }
                const int32_t dof = D.vel.cell_dofs[(size_t)c * vdpc + i];
                for (int k = 0; k < dim; ++k)
                  {
                    double gk = 0.0;
                    for (int d = 0; d < dim; ++d)
                      gk += inv[d][k] * gref[d];
                    for (int comp = 0; comp < dim; ++comp)
                      gu[comp][k] += u[(size_t)comp * nv + dof] * gk;
                  }
              }
            double pq = 0.0;
            for (int j = 0; j < pdpc; ++j)
              {
                double v = 1.0;
                for (int e = 0; e < dim; ++e)
                  v *= ((j >> e) & 1) ? xi[e] : 1.0 - xi[e];
                pq += p[D.pres.cell_dofs[(size_t)c * pdpc + j]] * v;
              }
            for (int k = 0; k < dim; ++k)
              {
                double tk = -pq * nb[k];
                for (int l = 0; l < dim; ++l)
                  tk += (gu[k][l] + gu[l][k]) * nb[l] / reynolds;
                F[k] += tk * dS;
              }
          }
      }
  const double Um = 1.5;
  const double ubar = dim == 2 ? 2.0 * Um / 3.0 : 4.0 * Um * D.width * D.width / 36.0;
  const double area = (ymax - ymin) * (dim == 3 ? (zmax - zmin) : 1.0);
  out[0] = 2.0 * F[0] / (ubar * ubar * area);
  out[1] = 2.0 * F[1] / (ubar * ubar * area);
}

} // namespace s35
