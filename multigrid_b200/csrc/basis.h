// Host-side 1-D ingredients for FE_Q(p) with QGauss(p+1) on the unit interval.
// Replaces what deal.II's FE_Q / QGauss / ShapeInfo hand to FEEvaluation<dim,p,p+1,1,Number>
// (reference call sites: common/laplace_operator.h:535, common/multigrid_solver.h:160).
#pragma once
#include <cmath>
#include <vector>

namespace mgb
{
  struct Basis
  {
    int                 p, n;
    std::vector<double> nodes;   // Gauss-Lobatto support points on [0,1]
    std::vector<double> xq, wq;  // Gauss-Legendre points / weights on [0,1]
    std::vector<double> S;       // S[q*n+i]  = phi_i(x_q)
    std::vector<double> G;       // G[q*n+i]  = phi_i'(x_q)
    std::vector<double> D;       // D[q*n+r]  = l_r'(x_q), collocation derivative (G = D S)
    std::vector<double> M1;      // sum_q S[q][i] w_q S[q][j]      (reference-cell 1-D mass)
    std::vector<double> K1;      // sum_q G[q][i] w_q G[q][j]      (reference-cell 1-D stiffness)
    std::vector<double> P1;      // P1[f*n+i] = phi_i(child support point f), f = 0..2p
  };

  namespace detail
  {
    typedef long double ld;

    // Legendre polynomial P_k and derivative by the three-term recurrence
    inline void
    leg(const int k, const ld x, ld &P, ld &dP)
    {
      ld a = 1, b = x, da = 0, db = 1;
      if (k == 0)
        {
          P  = 1;
          dP = 0;
          return;
        }
      for (int j = 2; j <= k; ++j)
        {
          const ld c  = ((2 * j - 1) * x * b - (j - 1) * a) / j;
          const ld dc = ((2 * j - 1) * (b + x * db) - (j - 1) * da) / j;
          a           = b;
          b           = c;
          da          = db;
          db          = dc;
        }
      P  = b;
      dP = db;
    }

    // roots of f by bracketing + bisection/Newton safeguard on a fine scan (robust, set-up only)
    template <typename F>
    inline std::vector<ld>
    roots_in(F f, const ld lo, const ld hi, const int expected)
    {
      std::vector<ld> r;
      const int       scan = 4000 * (expected + 1);
      ld              x0 = lo, f0 = f(x0);
      for (int s = 1; s <= scan; ++s)
        {
          const ld x1 = lo + (hi - lo) * s / scan, f1 = f(x1);
          if ((f0 < 0 && f1 >= 0) || (f0 > 0 && f1 <= 0))
            {
              ld a = x0, b = x1, fa = f0;
              for (int it = 0; it < 200; ++it)
                {
                  const ld m = 0.5L * (a + b), fm = f(m);
                  if ((fa < 0 && fm < 0) || (fa > 0 && fm > 0))
                    {
                      a  = m;
                      fa = fm;
                    }
                  else
                    b = m;
                }
              r.push_back(0.5L * (a + b));
            }
          x0 = x1;
          f0 = f1;
        }
      return r;
    }

    inline ld
    lag(const std::vector<ld> &x, const int i, const ld t)
    {
      ld v = 1;
      for (size_t l = 0; l < x.size(); ++l)
        if ((int)l != i)
          v *= (t - x[l]) / (x[i] - x[l]);
      return v;
    }

    inline ld
    dlag(const std::vector<ld> &x, const int i, const ld t)
    {
      ld s = 0;
      for (size_t m = 0; m < x.size(); ++m)
        if ((int)m != i)
          {
            ld v = 1 / (x[i] - x[m]);
            for (size_t l = 0; l < x.size(); ++l)
              if ((int)l != i && l != m)
                v *= (t - x[l]) / (x[i] - x[l]);
            s += v;
          }
      return s;
    }
  } // namespace detail

  inline Basis
  make_basis(const int p)
  {
    using namespace detail;
    Basis b;
    b.p         = p;
    b.n         = p + 1;
    const int n = p + 1;
    // Gauss points: roots of P_n on (-1,1); Lobatto interior points: roots of P_p'
    std::vector<ld> g = roots_in(
      [&](ld x) {
        ld P, dP;
        leg(n, x, P, dP);
        return P;
      },
      -1.L,
      1.L,
      n);
    std::vector<ld> l(1, -1.L);
    if (p > 1)
      {
        std::vector<ld> in = roots_in(
          [&](ld x) {
            ld P, dP;
            leg(p, x, P, dP);
            return dP;
          },
          -1.L + 1e-12L,
          1.L - 1e-12L,
          p - 1);
        l.insert(l.end(), in.begin(), in.end());
      }
    l.push_back(1.L);
    std::vector<ld> xl(n), xg(n), wg(n);
    for (int i = 0; i < n; ++i)
      {
        // enforce exact symmetry about 1/2
        xl[i] = 0.5L + 0.25L * (l[i] - l[n - 1 - i]);
        xg[i] = 0.5L + 0.25L * (g[i] - g[n - 1 - i]);
      }
    for (int i = 0; i < n; ++i)
      {
        const ld x = 2 * xg[i] - 1;
        ld       P, dP;
        leg(n, x, P, dP);
        wg[i] = 1 / ((1 - x * x) * dP * dP); // = (2/((1-x^2) dP^2)) / 2 for [0,1]
      }
    b.nodes.resize(n);
    b.xq.resize(n);
    b.wq.resize(n);
    for (int i = 0; i < n; ++i)
      {
        b.nodes[i] = (double)xl[i];
        b.xq[i]    = (double)xg[i];
        b.wq[i]    = (double)(0.5L * (wg[i] + wg[n - 1 - i]));
      }
    std::vector<ld> Sl(n * n), Gl(n * n);
    b.S.resize(n * n);
    b.G.resize(n * n);
    b.D.resize(n * n);
    b.M1.resize(n * n);
    b.K1.resize(n * n);
    for (int q = 0; q < n; ++q)
      for (int i = 0; i < n; ++i)
        {
          Sl[q * n + i]  = lag(xl, i, xg[q]);
          Gl[q * n + i]  = dlag(xl, i, xg[q]);
          b.S[q * n + i] = (double)Sl[q * n + i];
          b.G[q * n + i] = (double)Gl[q * n + i];
          b.D[q * n + i] = (double)dlag(xg, i, xg[q]);
        }
    for (int i = 0; i < n; ++i)
      for (int j = 0; j < n; ++j)
        {
          ld m = 0, k = 0;
          for (int q = 0; q < n; ++q)
            {
              const ld w = 0.5L * (wg[q] + wg[n - 1 - q]);
              m += Sl[q * n + i] * w * Sl[q * n + j];
              k += Gl[q * n + i] * w * Gl[q * n + j];
            }
          b.M1[i * n + j] = (double)m;
          b.K1[i * n + j] = (double)k;
        }
    b.P1.resize((2 * p + 1) * n);
    for (int f = 0; f <= 2 * p; ++f)
      {
        const ld xf = f <= p ? 0.5L * xl[f] : 0.5L + 0.5L * xl[f - p];
        for (int i = 0; i < n; ++i)
          b.P1[f * n + i] = (double)lag(xl, i, xf);
      }
    return b;
  }
} // namespace mgb
