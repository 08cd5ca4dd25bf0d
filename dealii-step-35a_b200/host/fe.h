// Host scaffolding: tensor-product Lagrange elements Q1/Q2 and Gauss quadrature
// on the unit cell, replacing deal.II's FE_Q / QGauss / FEValues tables
// (reference step-35a.cc:702-708: fe_velocity(deg+1), fe_pressure(deg),
// quadrature_pressure(deg+1), quadrature_velocity(deg+2)).
//
// Local DoFs and quadrature points are numbered lexicographically (x fastest).
// Nodes are equidistant, which for degree <= 2 coincides with the
// Gauss-Lobatto nodes FE_Q uses.
#pragma once
#include <cmath>
#include <stdexcept>
#include <vector>

namespace s35 {

inline int ipow(int b, int e)
{
  int r = 1;
  while (e-- > 0)
    r *= b;
  return r;
}

// n-point Gauss-Legendre rule on [0,1]
inline void gauss_rule(int n, std::vector<double> &x, std::vector<double> &w)
{
  x.resize(n);
  w.resize(n);
  switch (n)
    {
    case 1:
      x = {0.5};
      w = {1.0};
      break;
    case 2:
      {
        const double a = 0.5 / std::sqrt(3.0);
        x = {0.5 - a, 0.5 + a};
        w = {0.5, 0.5};
        break;
      }
    case 3:
      {
        const double a = 0.5 * std::sqrt(0.6);
        x = {0.5 - a, 0.5, 0.5 + a};
        w = {5.0 / 18.0, 8.0 / 18.0, 5.0 / 18.0};
        break;
      }
    case 4:
      {
        const double a = 0.5 * std::sqrt(3.0 / 7.0 - 2.0 / 7.0 * std::sqrt(1.2));
        const double b = 0.5 * std::sqrt(3.0 / 7.0 + 2.0 / 7.0 * std::sqrt(1.2));
        const double wa = (18.0 + std::sqrt(30.0)) / 72.0, wb = (18.0 - std::sqrt(30.0)) / 72.0;
        x = {0.5 - b, 0.5 - a, 0.5 + a, 0.5 + b};
        w = {wb, wa, wa, wb};
        break;
      }
    default:
      throw std::runtime_error("gauss_rule: only 1..4 points are tabulated");
    }
}

// value and derivative of the i-th 1D Lagrange polynomial of the given degree
// (equidistant nodes on [0,1]) at x
inline void lagrange_1d(int degree, int i, double x, double &value, double &deriv)
{
  value = 1.0;
  deriv = 0.0;
  const double h = 1.0 / degree;
  for (int k = 0; k <= degree; ++k)
    if (k != i)
      value *= (x - k * h) / ((i - k) * h);
  for (int m = 0; m <= degree; ++m)
    if (m != i)
      {
        double t = 1.0 / ((i - m) * h);
        for (int k = 0; k <= degree; ++k)
          if (k != i && k != m)
            t *= (x - k * h) / ((i - k) * h);
        deriv += t;
      }
}

// Shape values and reference gradients of Q_degree at the points of the
// n_q_1d-point tensor Gauss rule.
struct ShapeTable
{
  int dim = 0, degree = 0, n_q_1d = 0;
  int dpc = 0; // dofs per cell
  int nqp = 0; // quadrature points per cell
  std::vector<double> weight; // [nqp]
  std::vector<double> qpoint; // [nqp][dim]
  std::vector<double> value;  // [nqp][dpc]
  std::vector<double> grad;   // [nqp][dpc][dim] (reference gradients)
  // 1D factors, for sum-factorised kernels
  std::vector<double> value_1d; // [n_q_1d][degree+1]
  std::vector<double> deriv_1d; // [n_q_1d][degree+1]
  std::vector<double> weight_1d;

  ShapeTable() = default;
  ShapeTable(int dim_, int degree_, int n_q_1d_) : dim(dim_), degree(degree_), n_q_1d(n_q_1d_)
  {
    const int n1 = degree + 1;
    dpc = ipow(n1, dim);
    nqp = ipow(n_q_1d, dim);
    std::vector<double> xq;
    gauss_rule(n_q_1d, xq, weight_1d);
    value_1d.resize(n_q_1d * n1);
    deriv_1d.resize(n_q_1d * n1);
    for (int q = 0; q < n_q_1d; ++q)
      for (int i = 0; i < n1; ++i)
        lagrange_1d(degree, i, xq[q], value_1d[q * n1 + i], deriv_1d[q * n1 + i]);
    weight.resize(nqp);
    qpoint.resize(nqp * dim);
    value.resize((size_t)nqp * dpc);
    grad.resize((size_t)nqp * dpc * dim);
    for (int q = 0; q < nqp; ++q)
      {
        int qi[3] = {q % n_q_1d, (q / n_q_1d) % n_q_1d, q / (n_q_1d * n_q_1d)};
        double w = 1.0;
        for (int d = 0; d < dim; ++d)
          {
            w *= weight_1d[qi[d]];
            qpoint[q * dim + d] = xq[qi[d]];
          }
        weight[q] = w;
        for (int i = 0; i < dpc; ++i)
          {
            int ii[3] = {i % n1, (i / n1) % n1, i / (n1 * n1)};
            double v = 1.0;
            for (int d = 0; d < dim; ++d)
              v *= value_1d[qi[d] * n1 + ii[d]];
            value[(size_t)q * dpc + i] = v;
            for (int d = 0; d < dim; ++d)
              {
                double g = 1.0;
                for (int e = 0; e < dim; ++e)
                  g *= (e == d ? deriv_1d : value_1d)[qi[e] * n1 + ii[e]];
                grad[((size_t)q * dpc + i) * dim + d] = g;
              }
          }
      }
  }
};

} // namespace s35
