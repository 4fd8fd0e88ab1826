// ORACLE (test infrastructure, not product code).
//
// 1-D ingredients of the FE_Q(p) / QGauss(p+1) sum-factorisation kernels.
// deal.II (>= 9.5.0, not vendored under /root/reference) supplies these to the
// reference at common/laplace_operator.h:535 (FEEvaluation<dim,p,p+1,1,Number>)
// and common/multigrid_solver.h:160 (QGauss<1>(fe_degree+1)); this file restates
// the published definitions:
//   * FE_Q(p) support points      = Gauss-Lobatto points on [0,1]
//   * quadrature                  = Gauss-Legendre with p+1 points on [0,1]
//   * S[q][i] = phi_i(x_q), G[q][i] = phi_i'(x_q)
//   * D[q][r] = l_r'(x_q)  (collocation derivative on the Gauss points, G = D S)
//   * P1[f][i] = phi_i(xf_f), xf = the 2p+1 fine support points of the two
//     children of the unit interval (MGTransferMatrixFree's 1-D embedding)
// All evaluated in long double and rounded once.
#pragma once
#include <cmath>
#include <vector>

namespace orc
{
  struct Basis1D
  {
    int                 p = 0, n = 0;
    std::vector<double> gll;    // n support points
    std::vector<double> xq, wq; // n Gauss points / weights
    std::vector<double> S, G, D; // n*n, row = q
    std::vector<double> P1;      // (2p+1)*n
  };

  inline long double
  legendre(const int k, const long double x, long double *deriv = nullptr)
  {
    long double p0 = 1, p1 = x;
    if (k == 0)
      {
        if (deriv)
          *deriv = 0;
        return 1;
      }
    for (int j = 2; j <= k; ++j)
      {
        const long double p2 = ((2 * j - 1) * x * p1 - (j - 1) * p0) / j;
        p0                   = p1;
        p1                   = p2;
      }
    if (deriv)
      *deriv = k * (x * p1 - p0) / (x * x - 1);
    return p1;
  }

  inline long double
  lagrange(const std::vector<long double> &nodes, const int i, const long double x)
  {
    long double v = 1;
    for (int l = 0; l < (int)nodes.size(); ++l)
      if (l != i)
        v *= (x - nodes[l]) / (nodes[i] - nodes[l]);
    return v;
  }

  inline long double
  lagrange_deriv(const std::vector<long double> &nodes, const int i, const long double x)
  {
    long double s = 0;
    for (int m = 0; m < (int)nodes.size(); ++m)
      if (m != i)
        {
          long double v = 1 / (nodes[i] - nodes[m]);
          for (int l = 0; l < (int)nodes.size(); ++l)
            if (l != i && l != m)
              v *= (x - nodes[l]) / (nodes[i] - nodes[l]);
          s += v;
        }
    return s;
  }

  inline Basis1D
  make_basis(const int p)
  {
    Basis1D b;
    b.p = p;
    b.n = p + 1;
    const int         n  = p + 1;
    const long double pi = 3.141592653589793238462643383279502884L;
    // Gauss-Lobatto points on [-1,1]: +-1 and the roots of P_p'
    std::vector<long double> gl(n), gq(n), gw(n);
    gl[0] = -1;
    gl[p] = 1;
    for (int i = 1; i < p; ++i)
      {
        long double x = -std::cos(pi * i / p);
        for (int it = 0; it < 100; ++it)
          {
            // Newton on f = P_p'(x); f' from Legendre ODE: (1-x^2) P'' = 2x P' - p(p+1) P
            long double       dP;
            const long double P   = legendre(p, x, &dP);
            const long double ddP = (2 * x * dP - p * (p + 1) * P) / (1 - x * x);
            const long double dx  = dP / ddP;
            x -= dx;
            if (std::fabs((double)dx) < 1e-19)
              break;
          }
        gl[i] = x;
      }
    for (int i = 0; i < n; ++i)
      {
        long double x = -std::cos(pi * (i + 0.75L) / (n + 0.5L));
        long double dP;
        for (int it = 0; it < 100; ++it)
          {
            const long double P  = legendre(n, x, &dP);
            const long double dx = P / dP;
            x -= dx;
            if (std::fabs((double)dx) < 1e-19)
              break;
          }
        legendre(n, x, &dP);
        gq[i] = x;
        gw[i] = 2 / ((1 - x * x) * dP * dP);
      }
    // symmetrise and map to [0,1]
    std::vector<long double> xl(n), xg(n);
    for (int i = 0; i < n; ++i)
      {
        xl[i] = 0.5L * (1 + 0.5L * (gl[i] - gl[n - 1 - i]));
        xg[i] = 0.5L * (1 + 0.5L * (gq[i] - gq[n - 1 - i]));
      }
    b.gll.resize(n);
    b.xq.resize(n);
    b.wq.resize(n);
    for (int i = 0; i < n; ++i)
      {
        b.gll[i] = (double)xl[i];
        b.xq[i]  = (double)xg[i];
        b.wq[i]  = (double)(0.25L * (gw[i] + gw[n - 1 - i]));
      }
    b.S.resize(n * n);
    b.G.resize(n * n);
    b.D.resize(n * n);
    for (int q = 0; q < n; ++q)
      for (int i = 0; i < n; ++i)
        {
          b.S[q * n + i] = (double)lagrange(xl, i, xg[q]);
          b.G[q * n + i] = (double)lagrange_deriv(xl, i, xg[q]);
          b.D[q * n + i] = (double)lagrange_deriv(xg, i, xg[q]);
        }
    b.P1.resize((2 * p + 1) * n);
    for (int f = 0; f <= 2 * p; ++f)
      {
        const long double xf = f <= p ? 0.5L * xl[f] : 0.5L + 0.5L * xl[f - p];
        for (int i = 0; i < n; ++i)
          b.P1[f * n + i] = (double)lagrange(xl, i, xf);
      }
    return b;
  }
} // namespace orc
