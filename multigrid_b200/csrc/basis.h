// Host-side 1-D ingredients for FE_Q(p) with QGauss(p+1) on the unit interval.
// Replaces what deal.II's FE_Q / QGauss / ShapeInfo hand to FEEvaluation<dim,p,p+1,1,Number>
// (reference call sites: common/laplace_operator.h:535, common/multigrid_solver.h:160).
#pragma once
#include <cmath>
#include <utility>
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
    std::vector<double> R1;      // R1[i*(2p+1)+f]: value at coarse support point i of the fine function on the two children
                                 // (FE_Q restriction matrices = interpolation; the child containing the point contributes)
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
    b.R1.assign(n * (2 * p + 1), 0.);
    for (int i = 0; i < n; ++i)
      {
        const int child = xl[i] <= 0.5L ? 0 : 1;
        for (int m = 0; m < n; ++m)
          b.R1[i * (2 * p + 1) + child * p + m] = (double)lag(xl, m, 2 * xl[i] - child);
      }
    return b;
  }
  // ---- 1-D bases of the DG operator (laplace_operator_dg.h:346-348, matvec_dg/program.cc:44-49):
  //      MGB_DG_BASIS_GAUSS_LOBATTO  FE_DGQ(p), Lagrange polynomials in the Gauss-Lobatto points  (reference type 1)
  //      MGB_DG_BASIS_HERMITE        FE_DGQHermite(p), "Hermite-like" basis: only phi_0, phi_1 (phi_{p-1}, phi_p) have
  //                                  a value or a first derivative at x = 0 (x = 1)                (reference type 0)
  //      MGB_DG_BASIS_GAUSS          FE_DGQArbitraryNodes(QGauss(p+1)), collocated with the quadrature (reference type 2)
  //      Every function is kept in product form  w prod_k (x - r_k).
  enum
  {
    DG_BASIS_GAUSS_LOBATTO = 0,
    DG_BASIS_HERMITE       = 1,
    DG_BASIS_GAUSS         = 2
  };

  struct DGBasis
  {
    int                 p, n, kind;
    std::vector<double> xq, wq;   // Gauss points / weights on [0,1]
    std::vector<double> S, G, D;  // S[q n + i] = phi_i(x_q), G = phi_i'(x_q), D = collocation derivative
    std::vector<double> f0, f1;   // phi_i(0), phi_i(1)
    std::vector<double> e0, e1;   // phi_i'(0), phi_i'(1)
    std::vector<double> B;        // B[j n + i] = phi_i(Gauss-Lobatto point j): nodal values from coefficients
    std::vector<double> Binv;     // coefficients from the values in the Gauss-Lobatto points
    // h-embedding of the 1-D basis into the two children of a cell (the DG level transfer of MGTransferMatrixFree):
    // E[o][jf n + ic] = coefficient jf of the child o = 0 / 1 of the parent's basis function ic, i.e. the child's
    // representation of phi_ic((o + xi) / 2)
    std::vector<double> E[2];
  };

  namespace detail
  {
    struct ProductPoly
    {
      std::vector<ld> r;
      ld              w = 1;
      ld
      value(const ld x) const
      {
        ld v = w;
        for (const ld rk : r)
          v *= x - rk;
        return v;
      }
      ld
      derivative(const ld x) const
      {
        ld s = 0;
        for (size_t m = 0; m < r.size(); ++m)
          {
            ld v = w;
            for (size_t k = 0; k < r.size(); ++k)
              if (k != m)
                v *= x - r[k];
            s += v;
          }
        return s;
      }
    };

    inline std::vector<ProductPoly>
    lagrange_basis(const std::vector<ld> &x)
    {
      std::vector<ProductPoly> b(x.size());
      for (size_t i = 0; i < x.size(); ++i)
        for (size_t l = 0; l < x.size(); ++l)
          if (l != i)
            {
              b[i].r.push_back(x[l]);
              b[i].w /= x[i] - x[l];
            }
      return b;
    }

    // roots on [0,1] of the Jacobi polynomial P_m^(4,4), proportional to the Gegenbauer polynomial C_m^(9/2)
    inline std::vector<ld>
    jacobi44_roots(const int m)
    {
      if (m <= 0)
        return std::vector<ld>();
      const ld        lam = 4.5L;
      std::vector<ld> r   = roots_in(
        [&](ld x) {
          ld a = 1, b = 2 * lam * x;
          for (int k = 2; k <= m; ++k)
            {
              const ld c = (2 * x * (k + lam - 1) * b - (k + 2 * lam - 2) * a) / k;
              a          = b;
              b          = c;
            }
          return b;
        },
        -1.L + 1e-12L,
        1.L - 1e-12L,
        m);
      std::vector<ld> out(m);
      for (int i = 0; i < m; ++i)
        out[i] = 0.5L + 0.25L * (r[i] - r[m - 1 - i]);
      return out;
    }

    // Polynomials::HermiteLikeInterpolation of deal.II (the 1-D basis of FE_DGQHermite), restated: for p >= 4
    //   phi_0 = w0 (x - x*) prod_j (x - r_j) (x - 1)^2,  phi_0(0) = 1,  x* such that int phi_0 phi_1 = 0
    //   phi_1 = w1  x       prod_j (x - r_j) (x - 1)^2,  phi_0'(0) + phi_1'(0) = 0
    //   phi_i = wi  x^2 prod_{j != i-2} (x - r_j) (x - 1)^2,  phi_i(r_{i-2}) = 1,   i = 2 .. p-2
    //   phi_{p-1}(x) = phi_1(1 - x),  phi_p(x) = phi_0(1 - x)
    // with r_j the p-3 roots of the Jacobi polynomial P^(4,4)_{p-3} on [0,1]; p = 3 is the same construction without
    // interior functions (x* = 2/7), p = 2: (x-1)^2, -2 x (x-1), x^2; p = 1: 1-x, x.
    inline std::vector<ProductPoly>
    hermite_like_basis(const int p, const std::vector<ld> &xg, const std::vector<ld> &wg)
    {
      std::vector<ProductPoly> b(p + 1);
      if (p == 1)
        {
          b[0].r = {1.L};
          b[0].w = -1;
          b[1].r = {0.L};
          b[1].w = 1;
          return b;
        }
      if (p == 2)
        {
          b[0].r = {1.L, 1.L};
          b[0].w = 1;
          b[1].r = {0.L, 1.L};
          b[1].w = -2;
          b[2].r = {0.L, 0.L};
          b[2].w = 1;
          return b;
        }
      const std::vector<ld> jr = jacobi44_roots(p - 3);
      // x* = int x^2 c / int x c with c = prod (x - r_j)^2 (x - 1)^4; x^2 c has degree 2 p, which the p + 1 Gauss
      // points integrate exactly
      ld i1 = 0, i2 = 0;
      for (size_t q = 0; q < xg.size(); ++q)
        {
          const ld x = xg[q];
          ld       c = (x - 1) * (x - 1) * (x - 1) * (x - 1);
          for (const ld rj : jr)
            c *= (x - rj) * (x - rj);
          i1 += wg[q] * x * c;
          i2 += wg[q] * x * x * c;
        }
      const ld xs = i2 / i1;
      b[0].r.push_back(xs);
      b[1].r.push_back(0.L);
      for (int k = 0; k < 2; ++k)
        {
          for (const ld rj : jr)
            b[k].r.push_back(rj);
          b[k].r.push_back(1.L);
          b[k].r.push_back(1.L);
        }
      b[0].w = 1 / b[0].value(0.L);
      b[1].w = -b[0].derivative(0.L) / b[1].derivative(0.L);
      for (int i = 2; i <= p - 2; ++i)
        {
          b[i].r = {0.L, 0.L};
          for (int j = 0; j < p - 3; ++j)
            if (j != i - 2)
              b[i].r.push_back(jr[j]);
          b[i].r.push_back(1.L);
          b[i].r.push_back(1.L);
          b[i].w = 1 / b[i].value(jr[i - 2]);
        }
      for (int k = 0; k < 2; ++k)
        { // mirror images: phi(1 - x) = w (-1)^p prod (x - (1 - r))
          ProductPoly m;
          m.w = b[k].w * ((p % 2) ? -1 : 1);
          for (const ld rk : b[k].r)
            m.r.push_back(1 - rk);
          b[p - k] = m;
        }
      return b;
    }
  } // namespace detail

  inline DGBasis
  make_dg_basis(const int p, const int kind)
  {
    using namespace detail;
    const Basis q = make_basis(p);
    const int   n = p + 1;
    DGBasis     b;
    b.p    = p;
    b.n    = n;
    b.kind = kind;
    b.xq   = q.xq;
    b.wq   = q.wq;
    b.D    = q.D;
    std::vector<ld> xl(q.nodes.begin(), q.nodes.end()), xg(q.xq.begin(), q.xq.end()), wg(q.wq.begin(), q.wq.end());
    std::vector<ProductPoly> phi = kind == DG_BASIS_HERMITE ? hermite_like_basis(p, xg, wg) :
                                                              lagrange_basis(kind == DG_BASIS_GAUSS ? xg : xl);
    b.S.resize(n * n);
    b.G.resize(n * n);
    b.B.resize(n * n);
    b.f0.resize(n);
    b.f1.resize(n);
    b.e0.resize(n);
    b.e1.resize(n);
    const auto clean = [](const ld v) { return std::fabs((double)v) < 1e-15 ? 0. : (double)v; };
    for (int i = 0; i < n; ++i)
      {
        for (int k = 0; k < n; ++k)
          {
            b.S[k * n + i] = clean(phi[i].value(xg[k]));
            b.G[k * n + i] = clean(phi[i].derivative(xg[k]));
            b.B[k * n + i] = clean(phi[i].value(xl[k]));
          }
        b.f0[i] = clean(phi[i].value(0.L));
        b.f1[i] = clean(phi[i].value(1.L));
        b.e0[i] = clean(phi[i].derivative(0.L));
        b.e1[i] = clean(phi[i].derivative(1.L));
      }
    // Binv by Gauss-Jordan with partial pivoting (long double)
    std::vector<ld> A(n * 2 * n, 0);
    for (int i = 0; i < n; ++i)
      {
        for (int j = 0; j < n; ++j)
          A[i * 2 * n + j] = phi[j].value(xl[i]);
        A[i * 2 * n + n + i] = 1;
      }
    for (int c = 0; c < n; ++c)
      {
        int piv = c;
        for (int r = c + 1; r < n; ++r)
          if (std::fabs((double)A[r * 2 * n + c]) > std::fabs((double)A[piv * 2 * n + c]))
            piv = r;
        for (int j = 0; j < 2 * n; ++j)
          std::swap(A[c * 2 * n + j], A[piv * 2 * n + j]);
        const ld d = A[c * 2 * n + c];
        for (int j = 0; j < 2 * n; ++j)
          A[c * 2 * n + j] /= d;
        for (int r = 0; r < n; ++r)
          if (r != c)
            {
              const ld f = A[r * 2 * n + c];
              for (int j = 0; j < 2 * n; ++j)
                A[r * 2 * n + j] -= f * A[c * 2 * n + j];
            }
      }
    b.Binv.resize(n * n);
    for (int i = 0; i < n; ++i)
      for (int j = 0; j < n; ++j)
        b.Binv[i * n + j] = clean(A[i * 2 * n + n + j]);
    // embedding: values of the parent's functions in the Gauss-Lobatto points of the child, then to coefficients
    for (int o = 0; o < 2; ++o)
      {
        b.E[o].assign(n * n, 0.);
        for (int jf = 0; jf < n; ++jf)
          for (int ic = 0; ic < n; ++ic)
            {
              ld s = 0;
              for (int k = 0; k < n; ++k)
                s += A[jf * 2 * n + n + k] * phi[ic].value((o + xl[k]) / 2);
              b.E[o][jf * n + ic] = clean(s);
            }
      }
    return b;
  }
} // namespace mgb
