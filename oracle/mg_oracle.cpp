// ORACLE (test infrastructure, not product code; parity PINNED against README.md:135-159).
//
// CPU restatement of the matrix-free geometric-multigrid hot path of
// kronbichler/multigrid for FE_Q(p), p = 1..8, on (refined) box meshes. Only tests/,
// __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
// load this library. The GPU library never links or calls it.
//
// What is restated (file:line under /root/reference):
//   operator            common/laplace_operator.h:359-432 (merged coefficient),
//                       :438-523 (quadrature-point operation), :529-558 (cell),
//                       :575-601 vmult, :607-634 vmult_residual, :640-719 cg update,
//                       :747-800 compute_diagonal, :806-845 compute_residual
//   gather/scatter      common/vector_access_reduced.h:153-239, :407-495 (constrained
//                       entities read as 0 / skipped on write)
//   solver              common/multigrid_solver.h:100-292 (set-up), :298-343 (L2 error),
//                       :387-476 (FMG), :483-493 (solve_cg), :498-510 (vmult),
//                       :641-681 (v_cycle)
//   problem             poisson_cube/program.cc:79-144, :498-570
// What lives in deal.II (>= 9.5.0, not vendored) and is restated
// from its published algorithms: FEEvaluation sum factorisation, PreconditionChebyshev
// (first_kind), MGTransferMatrixFree, SolverCG incl. its Lanczos eigenvalue estimate.
//
// DoF numbering here: global lexicographic on the (N_d p + 1)^3 grid (x fastest).
#include <algorithm>
#include <array>
#include <chrono>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <memory>
#include <vector>
#ifdef _OPENMP
#include <omp.h>
#endif

#include "basis1d.h"

namespace orc
{
  using std::size_t;
  using std::vector;

  // ---------------------------------------------------------------- functions
  static const double PI = 3.14159265358979323846;

  // poisson_cube/program.cc:98-104
  static double
  sol_cube(const double *x)
  {
    return std::sin(PI * x[0] * 3.) * std::sin(PI * x[1] * 3.) * std::sin(PI * x[2] * 3.);
  }
  // poisson_cube/program.cc:140-144
  static double
  rhs_cube(const double *x)
  {
    return 3 * PI * 3. * PI * 3. * sol_cube(x);
  }
  static double
  coef_one(const double *)
  {
    return 1.;
  }
  // poisson_shell/program.cc:97-101
  static double
  sol_shell(const double *x)
  {
    return std::sin(2 * PI * (x[0] + x[1]));
  }
  // poisson_shell/program.cc:159-171
  static double
  coef_shell(const double *x)
  {
    double prod = 1.;
    for (int e = 0; e < 3; ++e)
      {
        const double c = std::cos(2 * PI * x[e] + 0.1 * e);
        prod *= c * c;
      }
    return 1.0 + 1.0e6 * prod;
  }
  // poisson_shell/program.cc:175-225
  static double
  rhs_shell(const double *x)
  {
    const double lap     = -2 * 2 * PI * 2 * PI * std::sin(2 * PI * (x[0] + x[1]));
    const double gsol    = 2 * PI * std::cos(2 * PI * (x[0] + x[1]));
    double       gdot    = 0;
    for (int d = 0; d < 2; ++d) // solution gradient is (g, g, 0)
      {
        double prod = 1.;
        for (int e = 0; e < 3; ++e)
          {
            const double c = std::cos(2 * PI * x[e] + 0.1 * e);
            if (e == d)
              prod *= -4 * PI * c * std::sin(2 * PI * x[e] + 0.1 * e);
            else
              prod *= c * c;
          }
        gdot += 1.0e6 * prod * gsol;
      }
    return -(lap * coef_shell(x) + gdot);
  }

  typedef double (*Fn)(const double *);

  // ---------------------------------------------------------------- geometry
  // Map from the unit box of mesh coordinates to physical space.  kind 0: the box [lo, hi] (axis-aligned cells,
  // subdivided_hyper_rectangle); kind 1: one of the six coarse cells of GridGenerator::hyper_shell(0, r0, r1, 6)
  // (poisson_shell/program.cc:426) -- a cube face of half-width tan(alpha) projected radially, radius r0..r1 --
  // and its uniform refinements; kind 2: affine, x = lo + A ((hi - lo) o xi).  Inside a cell the geometry is the
  // degree-p polynomial through the images of the Gauss-Lobatto support points, as MappingQGeneric(p)
  // (multigrid_solver.h:138-139) represents a manifold; deal.II's SphericalManifold places points by weighted
  // averages on the sphere, not by this analytic map, so against the reference's shell meshes this is PARITY UNPINNED.
  struct Geometry
  {
    int    kind = 0;
    double lo[3] = {0, 0, 0}, hi[3] = {1, 1, 1};
    double par[12] = {0};

    void
    point(const double xi[3], double x[3]) const
    {
      if (kind == 1)
        {
          const double half = std::tan(par[2]);
          const double v[3] = {half * (2 * xi[0] - 1), half * (2 * xi[1] - 1), 1.};
          const double len  = std::sqrt(v[0] * v[0] + v[1] * v[1] + v[2] * v[2]);
          const double r    = par[0] * (1 - xi[2]) + par[1] * xi[2];
          for (int d = 0; d < 3; ++d)
            x[d] = r * v[d] / len;
        }
      else
        {
          double u[3];
          for (int d = 0; d < 3; ++d)
            u[d] = (hi[d] - lo[d]) * xi[d];
          for (int d = 0; d < 3; ++d)
            x[d] = lo[d] + (kind == 2 ? par[3 * d] * u[0] + par[3 * d + 1] * u[1] + par[3 * d + 2] * u[2] : u[d]);
        }
    }
  };

  // inverse and determinant of a 3 x 3 matrix by Gauss-Jordan elimination with partial pivoting
  static double
  invert3(const double A[3][3], double inv[3][3])
  {
    double M[3][6];
    for (int i = 0; i < 3; ++i)
      for (int j = 0; j < 3; ++j)
        {
          M[i][j]     = A[i][j];
          M[i][3 + j] = i == j;
        }
    double det = 1;
    for (int c = 0; c < 3; ++c)
      {
        int piv = c;
        for (int r = c + 1; r < 3; ++r)
          if (std::fabs(M[r][c]) > std::fabs(M[piv][c]))
            piv = r;
        if (piv != c)
          {
            for (int j = 0; j < 6; ++j)
              std::swap(M[c][j], M[piv][j]);
            det = -det;
          }
        det *= M[c][c];
        const double d = 1. / M[c][c];
        for (int j = 0; j < 6; ++j)
          M[c][j] *= d;
        for (int r = 0; r < 3; ++r)
          if (r != c)
            {
              const double f = M[r][c];
              for (int j = 0; j < 6; ++j)
                M[r][j] -= f * M[c][j];
            }
      }
    for (int i = 0; i < 3; ++i)
      for (int j = 0; j < 3; ++j)
        inv[i][j] = M[i][3 + j];
    return det;
  }

  // ---------------------------------------------------------------- 1-D sweeps
  // out[.., o, ..] (+)= sum_i M[o][i] in[.., i, ..]   (transpose: M[i][o])
  template <typename T, int n, int dir, bool transpose, bool add>
  static inline void
  sweep(const T *__restrict M, const T *in, T *out)
  {
    constexpr int stride = dir == 0 ? 1 : (dir == 1 ? n : n * n);
    for (int a = 0; a < n; ++a)
      for (int b = 0; b < n; ++b)
        {
          const int base = dir == 0 ? (a * n + b) * n : (dir == 1 ? a * n * n + b : a * n + b);
          T         line[n];
          for (int i = 0; i < n; ++i)
            line[i] = in[base + i * stride];
          for (int o = 0; o < n; ++o)
            {
              T sum = 0;
              for (int i = 0; i < n; ++i)
                sum += (transpose ? M[i * n + o] : M[o * n + i]) * line[i];
              if (add)
                out[base + o * stride] += sum;
              else
                out[base + o * stride] = sum;
            }
        }
  }

  // ---------------------------------------------------------------- level operator
  template <typename T>
  struct LevelOp
  {
    const Basis1D *b = nullptr;
    int            p = 0, n = 0;
    int            nc[3], nd[3];
    size_t         ndofs = 0, ncells = 0;
    double         lo[3], h[3];
    vector<T>      S, D, wq;
    bool           variable = false;
    Geometry       geom;
    T              coef_affine[6];
    vector<T>      coef_q; // [cell][q][6]
    vector<T>      inv_diag;

    void
    init(const Basis1D &basis, const int ncell[3], const double lower[3], const double upper[3])
    {
      b = &basis;
      p = basis.p;
      n = basis.n;
      ndofs  = 1;
      ncells = 1;
      for (int d = 0; d < 3; ++d)
        {
          nc[d] = ncell[d];
          nd[d] = ncell[d] * p + 1;
          lo[d] = lower[d];
          h[d]  = (upper[d] - lower[d]) / ncell[d];
          ndofs *= nd[d];
          ncells *= nc[d];
        }
      S.resize(n * n);
      D.resize(n * n);
      wq.resize(n);
      for (int i = 0; i < n * n; ++i)
        {
          S[i] = (T)basis.S[i];
          D[i] = (T)basis.D[i];
        }
      for (int i = 0; i < n; ++i)
        wq[i] = (T)basis.wq[i];
    }

    size_t
    m() const
    {
      return ndofs;
    }

    bool
    on_boundary(const int i, const int j, const int k) const
    {
      return i == 0 || j == 0 || k == 0 || i == nd[0] - 1 || j == nd[1] - 1 || k == nd[2] - 1;
    }

    // control points of the polynomial geometry of cell (cx,cy,cz): X[l][d], l lexicographic over the support points
    void
    cell_nodes(const int cx, const int cy, const int cz, vector<double> &X) const
    {
      X.resize(3 * n * n * n);
      for (int k = 0, l = 0; k < n; ++k)
        for (int j = 0; j < n; ++j)
          for (int i = 0; i < n; ++i, ++l)
            {
              const double xi[3] = {(cx + b->gll[i]) / nc[0], (cy + b->gll[j]) / nc[1], (cz + b->gll[k]) / nc[2]};
              geom.point(xi, &X[3 * l]);
            }
    }

    // position and Jacobian d x_d / d xi_e of the cell geometry at quadrature point (q0,q1,q2)
    void
    cell_geometry(const vector<double> &X, const int q0, const int q1, const int q2, double x[3], double J[3][3]) const
    {
      for (int d = 0; d < 3; ++d)
        x[d] = J[d][0] = J[d][1] = J[d][2] = 0;
      for (int k = 0, l = 0; k < n; ++k)
        for (int j = 0; j < n; ++j)
          for (int i = 0; i < n; ++i, ++l)
            {
              const double v[3]  = {b->S[q0 * n + i], b->S[q1 * n + j], b->S[q2 * n + k]};
              const double dv[3] = {b->G[q0 * n + i], b->G[q1 * n + j], b->G[q2 * n + k]};
              for (int d = 0; d < 3; ++d)
                {
                  x[d] += X[3 * l + d] * v[0] * v[1] * v[2];
                  J[d][0] += X[3 * l + d] * dv[0] * v[1] * v[2];
                  J[d][1] += X[3 * l + d] * v[0] * dv[1] * v[2];
                  J[d][2] += X[3 * l + d] * v[0] * v[1] * dv[2];
                }
            }
    }

    // symmetric tensor scale * J^-1 J^-T in the storage order 00, 11, 22, 01, 02, 12 (laplace_operator.h:382-386)
    static void
    merged_tensor(const double J[3][3], const double scale_without_det, T *c)
    {
      double       inv[3][3];
      const double det = invert3(J, inv);
      const double f   = scale_without_det * std::fabs(det);
      const int    pe[6] = {0, 1, 2, 0, 0, 1}, pf[6] = {0, 1, 2, 1, 2, 2};
      for (int s = 0; s < 6; ++s)
        {
          double m = 0;
          for (int d = 0; d < 3; ++d)
            m += inv[pe[s]][d] * inv[pf[s]][d];
          c[s] = (T)(f * m);
        }
    }

    // laplace_operator.h:359-432
    void
    evaluate_coefficient(Fn coefficient, const bool is_constant)
    {
      const double detJ   = h[0] * h[1] * h[2];
      const int    nq     = n * n * n;
      variable            = !is_constant;
      if (geom.kind == 2 && is_constant)
        {
          // affine cells: one tensor per geometry class (:374-387), quadrature weight applied in the cell kernel
          double J[3][3];
          for (int d = 0; d < 3; ++d)
            for (int e = 0; e < 3; ++e)
              J[d][e] = geom.par[3 * d + e] * h[e];
          const double x0[3] = {0, 0, 0};
          merged_tensor(J, coefficient(x0), coef_affine);
          return;
        }
      if (geom.kind != 0)
        {
          // general cells: one tensor per quadrature point, JxW includes the weight (:388-430)
          variable = true;
          coef_q.resize(ncells * nq * 6);
#pragma omp parallel for collapse(2)
          for (int cz = 0; cz < nc[2]; ++cz)
            for (int cy = 0; cy < nc[1]; ++cy)
              for (int cx = 0; cx < nc[0]; ++cx)
                {
                  const size_t   cell = ((size_t)cz * nc[1] + cy) * nc[0] + cx;
                  vector<double> X;
                  cell_nodes(cx, cy, cz, X);
                  for (int q2 = 0; q2 < n; ++q2)
                    for (int q1 = 0; q1 < n; ++q1)
                      for (int q0 = 0; q0 < n; ++q0)
                        {
                          double x[3], J[3][3];
                          cell_geometry(X, q0, q1, q2, x, J);
                          merged_tensor(J, coefficient(x) * (b->wq[q0] * b->wq[q1] * b->wq[q2]),
                                        &coef_q[(cell * nq + (q2 * n + q1) * n + q0) * 6]);
                        }
                }
          return;
        }
      if (is_constant)
        {
          const double x0[3] = {0, 0, 0};
          const double c     = coefficient(x0);
          for (int d = 0; d < 3; ++d)
            coef_affine[d] = (T)(detJ / (h[d] * h[d]) * c);
          for (int d = 3; d < 6; ++d)
            coef_affine[d] = 0;
          return;
        }
      coef_q.resize(ncells * nq * 6);
#pragma omp parallel for collapse(2)
      for (int cz = 0; cz < nc[2]; ++cz)
        for (int cy = 0; cy < nc[1]; ++cy)
          for (int cx = 0; cx < nc[0]; ++cx)
            {
              const size_t cell = ((size_t)cz * nc[1] + cy) * nc[0] + cx;
              for (int q2 = 0; q2 < n; ++q2)
                for (int q1 = 0; q1 < n; ++q1)
                  for (int q0 = 0; q0 < n; ++q0)
                    {
                      const double x[3] = {lo[0] + (cx + b->xq[q0]) * h[0],
                                           lo[1] + (cy + b->xq[q1]) * h[1],
                                           lo[2] + (cz + b->xq[q2]) * h[2]};
                      const double f =
                        coefficient(x) * (b->wq[q0] * b->wq[q1] * b->wq[q2]) * detJ;
                      T *c = &coef_q[(cell * nq + (q2 * n + q1) * n + q0) * 6];
                      for (int d = 0; d < 3; ++d)
                        c[d] = (T)(f / (h[d] * h[d]));
                      c[3] = c[4] = c[5] = 0;
                    }
            }
    }

    // evaluate_coefficient(const Function &) with the function given by its samples c[cell][q] at the quadrature points
    void
    set_coefficient_values(const double *c)
    {
      evaluate_coefficient(coef_one, /*is_constant=*/false); // per-point tensors of the geometry alone
      const size_t nq = (size_t)n * n * n;
      variable        = true;
      if (coef_q.empty()) // (affine geometry with the constant-tensor shortcut cannot happen: is_constant is false)
        return;
      for (size_t cell = 0; cell < ncells; ++cell)
        for (size_t q = 0; q < nq; ++q)
          for (int e = 0; e < 6; ++e)
            coef_q[(cell * nq + q) * 6 + e] *= (T)c[cell * nq + q];
    }

    // out[q2][q1][q0] = sum_kji M2[q2][k] M1[q1][j] M0[q0][i] in[k][j][i]; plain loops (set-up code)
    static void
    tensor3(const int n, const double *M0, const double *M1, const double *M2, const double *in, double *out)
    {
      vector<double> a(n * n * n), c(n * n * n);
      for (int k = 0; k < n; ++k)
        for (int j = 0; j < n; ++j)
          for (int q0 = 0; q0 < n; ++q0)
            {
              double s = 0;
              for (int i = 0; i < n; ++i)
                s += M0[q0 * n + i] * in[(k * n + j) * n + i];
              a[(k * n + j) * n + q0] = s;
            }
      for (int k = 0; k < n; ++k)
        for (int q1 = 0; q1 < n; ++q1)
          for (int q0 = 0; q0 < n; ++q0)
            {
              double s = 0;
              for (int j = 0; j < n; ++j)
                s += M1[q1 * n + j] * a[(k * n + j) * n + q0];
              c[(k * n + q1) * n + q0] = s;
            }
      for (int q2 = 0; q2 < n; ++q2)
        for (int q1 = 0; q1 < n; ++q1)
          for (int q0 = 0; q0 < n; ++q0)
            {
              double s = 0;
              for (int k = 0; k < n; ++k)
                s += M2[q2 * n + k] * c[(k * n + q1) * n + q0];
              out[(q2 * n + q1) * n + q0] = s;
            }
    }

    // MinimalSurfaceOperator::evaluate_coefficient(first_time, solution) (minimal_surface/program.cc:117-163): the merged
    // coefficient JxW J^-1 T J^-T with the tensor T of the gradient g of `u` (all entries of u are read, boundary values
    // included: the FEEvaluation without Dirichlet constraints of :123):
    // mode 0 (first_time): T = I;  mode 1: T = (I - g g^T / (1 + |g|^2)) / sqrt(1 + |g|^2), the Newton linearisation;
    // mode 2: T = I / sqrt(1 + |g|^2), the flux of the nonlinear residual (:188-192) written as an operator on u
    void
    evaluate_solution_coefficient(const int mode, const T *u)
    {
      const int nq = n * n * n;
      variable     = true;
      coef_q.resize(ncells * nq * 6);
#pragma omp parallel for collapse(2)
      for (int cz = 0; cz < nc[2]; ++cz)
        for (int cy = 0; cy < nc[1]; ++cy)
          for (int cx = 0; cx < nc[0]; ++cx)
            {
              const size_t   cell = ((size_t)cz * nc[1] + cy) * nc[0] + cx;
              vector<double> X, ul(nq), gr[3] = {vector<double>(nq), vector<double>(nq), vector<double>(nq)};
              if (geom.kind != 0)
                cell_nodes(cx, cy, cz, X);
              for (int k = 0, l = 0; k < n; ++k)
                for (int j = 0; j < n; ++j)
                  for (int i = 0; i < n; ++i, ++l)
                    ul[l] = u[((size_t)(cz * p + k) * nd[1] + cy * p + j) * nd[0] + cx * p + i];
              const double *Sb = b->S.data(), *Gb = b->G.data();
              tensor3(n, Gb, Sb, Sb, ul.data(), gr[0].data());
              tensor3(n, Sb, Gb, Sb, ul.data(), gr[1].data());
              tensor3(n, Sb, Sb, Gb, ul.data(), gr[2].data());
              for (int q2 = 0, q = 0; q2 < n; ++q2)
                for (int q1 = 0; q1 < n; ++q1)
                  for (int q0 = 0; q0 < n; ++q0, ++q)
                    {
                      double x[3], J[3][3] = {{h[0], 0, 0}, {0, h[1], 0}, {0, 0, h[2]}}, inv[3][3];
                      if (geom.kind != 0)
                        cell_geometry(X, q0, q1, q2, x, J);
                      const double det = invert3(J, inv); // inv[e][d] = d xi_e / d x_d
                      double       g[3], Tn[3][3];
                      for (int d = 0; d < 3; ++d)
                        g[d] = inv[0][d] * gr[0][q] + inv[1][d] * gr[1][q] + inv[2][d] * gr[2][q];
                      const double g2 = g[0] * g[0] + g[1] * g[1] + g[2] * g[2];
                      for (int d = 0; d < 3; ++d)
                        for (int e = 0; e < 3; ++e)
                          {
                            const double id = d == e ? 1. : 0.;
                            Tn[d][e]        = mode == 0 ? id :
                                                          (mode == 1 ? (id - g[d] * g[e] / (1. + g2)) / std::sqrt(1. + g2) :
                                                                       id / std::sqrt(1. + g2));
                          }
                      const double f     = b->wq[q0] * b->wq[q1] * b->wq[q2] * std::fabs(det);
                      const int    pe[6] = {0, 1, 2, 0, 0, 1}, pf[6] = {0, 1, 2, 1, 2, 2};
                      for (int s = 0; s < 6; ++s)
                        {
                          double m = 0;
                          for (int d = 0; d < 3; ++d)
                            for (int e = 0; e < 3; ++e)
                              m += inv[pe[s]][d] * Tn[d][e] * inv[pf[s]][e];
                          coef_q[(cell * nq + q) * 6 + s] = (T)(f * m);
                        }
                    }
            }
    }

    // evaluate + quadrature-point operation + integrate on one cell
    // (laplace_operator.h:544-551 with :438-523 in between)
    template <int nn>
    inline void
    cell_kernel(const size_t cell, T *u /* n^3, overwritten with result */) const
    {
      constexpr int nq = nn * nn * nn;
      T             g0[nq], g1[nq], g2[nq];
      const T      *Sp = S.data(), *Dp = D.data();
      sweep<T, nn, 0, false, false>(Sp, u, u);
      sweep<T, nn, 1, false, false>(Sp, u, u);
      sweep<T, nn, 2, false, false>(Sp, u, u);
      sweep<T, nn, 0, false, false>(Dp, u, g0);
      sweep<T, nn, 1, false, false>(Dp, u, g1);
      sweep<T, nn, 2, false, false>(Dp, u, g2);
      if (!variable)
        {
          const T *c = coef_affine;
          for (int q2 = 0, q = 0; q2 < nn; ++q2)
            for (int q1 = 0; q1 < nn; ++q1)
              for (int q0 = 0; q0 < nn; ++q0, ++q)
                {
                  const T weight = wq[q0] * wq[q1] * wq[q2];
                  const T t0 = g0[q], t1 = g1[q], t2 = g2[q];
                  g0[q] = (c[0] * t0 + c[3] * t1 + c[4] * t2) * weight;
                  g1[q] = (c[3] * t0 + c[1] * t1 + c[5] * t2) * weight;
                  g2[q] = (c[4] * t0 + c[5] * t1 + c[2] * t2) * weight;
                }
        }
      else
        {
          const T *cq = &coef_q[cell * nq * 6];
          for (int q = 0; q < nq; ++q)
            {
              const T *c  = cq + 6 * q;
              const T  t0 = g0[q], t1 = g1[q], t2 = g2[q];
              g0[q] = (c[0] * t0 + c[3] * t1 + c[4] * t2);
              g1[q] = (c[3] * t0 + c[1] * t1 + c[5] * t2);
              g2[q] = (c[4] * t0 + c[5] * t1 + c[2] * t2);
            }
        }
      sweep<T, nn, 0, true, false>(Dp, g0, u);
      sweep<T, nn, 1, true, true>(Dp, g1, u);
      sweep<T, nn, 2, true, true>(Dp, g2, u);
      sweep<T, nn, 2, true, false>(Sp, u, u);
      sweep<T, nn, 1, true, false>(Sp, u, u);
      sweep<T, nn, 0, true, false>(Sp, u, u);
    }

    // dst += sum_cells R^T K_cell R src; Dirichlet entries read as zero (unless
    // read_plain) and never written.  Two-colour z-layer loop for thread safety.
    template <int nn>
    void
    apply_add_t(T *dst, const T *src, const bool read_plain) const
    {
      for (int colour = 0; colour < 2; ++colour)
        {
#pragma omp parallel for schedule(static)
          for (int cz = colour; cz < nc[2]; cz += 2)
            for (int cy = 0; cy < nc[1]; ++cy)
              for (int cx = 0; cx < nc[0]; ++cx)
                {
                  T            u[nn * nn * nn];
                  const size_t cell = ((size_t)cz * nc[1] + cy) * nc[0] + cx;
                  for (int k = 0, l = 0; k < nn; ++k)
                    for (int j = 0; j < nn; ++j)
                      for (int i = 0; i < nn; ++i, ++l)
                        {
                          const int gi = cx * p + i, gj = cy * p + j, gk = cz * p + k;
                          u[l] = (!read_plain && on_boundary(gi, gj, gk)) ?
                                   T(0) :
                                   src[((size_t)gk * nd[1] + gj) * nd[0] + gi];
                        }
                  cell_kernel<nn>(cell, u);
                  for (int k = 0, l = 0; k < nn; ++k)
                    for (int j = 0; j < nn; ++j)
                      for (int i = 0; i < nn; ++i, ++l)
                        {
                          const int gi = cx * p + i, gj = cy * p + j, gk = cz * p + k;
                          if (!on_boundary(gi, gj, gk))
                            dst[((size_t)gk * nd[1] + gj) * nd[0] + gi] += u[l];
                        }
                }
        }
    }

    void
    apply_add(T *dst, const T *src, const bool read_plain = false) const
    {
      switch (n)
        {
          case 2: apply_add_t<2>(dst, src, read_plain); break;
          case 3: apply_add_t<3>(dst, src, read_plain); break;
          case 4: apply_add_t<4>(dst, src, read_plain); break;
          case 5: apply_add_t<5>(dst, src, read_plain); break;
          case 6: apply_add_t<6>(dst, src, read_plain); break;
          case 7: apply_add_t<7>(dst, src, read_plain); break;
          case 8: apply_add_t<8>(dst, src, read_plain); break;
          case 9: apply_add_t<9>(dst, src, read_plain); break;
        }
    }

    template <typename F>
    void
    for_boundary(F &&f) const
    {
      for (int k = 0; k < nd[2]; ++k)
        for (int j = 0; j < nd[1]; ++j)
          {
            const bool edge = k == 0 || j == 0 || k == nd[2] - 1 || j == nd[1] - 1;
            if (edge)
              for (int i = 0; i < nd[0]; ++i)
                f(((size_t)k * nd[1] + j) * nd[0] + i);
            else
              {
                f(((size_t)k * nd[1] + j) * nd[0]);
                f(((size_t)k * nd[1] + j) * nd[0] + nd[0] - 1);
              }
          }
    }

    // laplace_operator.h:575-601
    void
    vmult(T *dst, const T *src) const
    {
#pragma omp parallel for
      for (size_t i = 0; i < ndofs; ++i)
        dst[i] = 0;
      apply_add(dst, src);
      for_boundary([&](const size_t i) { dst[i] = src[i]; });
    }

    // laplace_operator.h:607-634
    void
    vmult_residual(const T *rhs, const T *lhs, T *residual) const
    {
#pragma omp parallel for
      for (size_t i = 0; i < ndofs; ++i)
        residual[i] = 0;
      apply_add(residual, lhs);
#pragma omp parallel for
      for (size_t i = 0; i < ndofs; ++i)
        residual[i] = rhs[i] - residual[i];
      for_boundary([&](const size_t i) { residual[i] -= lhs[i]; });
    }

    // laplace_operator.h:640-719
    std::array<T, 4>
    vmult_with_cg_update(const T alpha, const T beta, const T *r, T *q, T *pp, T *x) const
    {
      if (alpha == T(0))
        {
#pragma omp parallel for
          for (size_t i = 0; i < ndofs; ++i)
            {
              pp[i] = q[i];
              q[i]  = 0;
            }
        }
      else
        {
#pragma omp parallel for
          for (size_t i = 0; i < ndofs; ++i)
            {
              x[i] += alpha * pp[i];
              pp[i] = beta * pp[i] + q[i];
              q[i]  = 0;
            }
        }
      apply_add(q, pp);
      // NOTE: the reference does not touch constrained rows here (q stays 0 there)
      double s0 = 0, s1 = 0, s2 = 0, s3 = 0;
#pragma omp parallel for reduction(+ : s0, s1, s2, s3)
      for (size_t i = 0; i < ndofs; ++i)
        {
          s0 += (double)q[i] * pp[i];
          s1 += (double)r[i] * r[i];
          s2 += (double)q[i] * r[i];
          s3 += (double)q[i] * q[i];
        }
      return {{(T)s0, (T)s1, (T)s2, (T)s3}};
    }

    // laplace_operator.h:747-800: apply the cell operator to every unit vector
    template <int nn>
    void
    compute_diagonal_t()
    {
      constexpr int nq = nn * nn * nn;
      inv_diag.assign(ndofs, T(0));
      for (int colour = 0; colour < 2; ++colour)
        {
#pragma omp parallel for schedule(static)
          for (int cz = colour; cz < nc[2]; cz += 2)
            for (int cy = 0; cy < nc[1]; ++cy)
              for (int cx = 0; cx < nc[0]; ++cx)
                {
                  const size_t cell = ((size_t)cz * nc[1] + cy) * nc[0] + cx;
                  T            u[nq], diag[nq];
                  for (int i = 0; i < nq; ++i)
                    {
                      for (int j = 0; j < nq; ++j)
                        u[j] = 0;
                      u[i] = 1;
                      cell_kernel<nn>(cell, u);
                      diag[i] = u[i];
                    }
                  for (int k = 0, l = 0; k < nn; ++k)
                    for (int j = 0; j < nn; ++j)
                      for (int i = 0; i < nn; ++i, ++l)
                        {
                          const int gi = cx * p + i, gj = cy * p + j, gk = cz * p + k;
                          if (!on_boundary(gi, gj, gk))
                            inv_diag[((size_t)gk * nd[1] + gj) * nd[0] + gi] += diag[l];
                        }
                }
        }
      for_boundary([&](const size_t i) { inv_diag[i] = 1; });
#pragma omp parallel for
      for (size_t i = 0; i < ndofs; ++i)
        inv_diag[i] = T(1) / inv_diag[i];
    }

    void
    compute_diagonal()
    {
      switch (n)
        {
          case 2: compute_diagonal_t<2>(); break;
          case 3: compute_diagonal_t<3>(); break;
          case 4: compute_diagonal_t<4>(); break;
          case 5: compute_diagonal_t<5>(); break;
          case 6: compute_diagonal_t<6>(); break;
          case 7: compute_diagonal_t<7>(); break;
          case 8: compute_diagonal_t<8>(); break;
          case 9: compute_diagonal_t<9>(); break;
        }
    }

    // laplace_operator.h:806-845: dst = int f phi - K u_bc (rows of Dirichlet DoFs dropped)
    void
    compute_residual(T *dst, const T *src, Fn rhs_function, const double *rhs_values = nullptr) const
    {
      for (size_t i = 0; i < ndofs; ++i)
        dst[i] = 0;
      vector<T> minus(ndofs);
      for (size_t i = 0; i < ndofs; ++i)
        minus[i] = -src[i];
      apply_add(dst, minus.data(), /*read_plain=*/true);
      // value part: S^T ( f(x_q) JxW )
      const double detJ = h[0] * h[1] * h[2];
      const int    nq   = n * n * n;
      for (int colour = 0; colour < 2; ++colour)
        {
#pragma omp parallel for schedule(static)
          for (int cz = colour; cz < nc[2]; cz += 2)
            for (int cy = 0; cy < nc[1]; ++cy)
              for (int cx = 0; cx < nc[0]; ++cx)
                {
                  vector<double> v(nq), t(nq), X;
                  if (geom.kind != 0)
                    cell_nodes(cx, cy, cz, X);
                  for (int q2 = 0, q = 0; q2 < n; ++q2)
                    for (int q1 = 0; q1 < n; ++q1)
                      for (int q0 = 0; q0 < n; ++q0, ++q)
                        {
                          if (geom.kind != 0)
                            {
                              double x[3], J[3][3], inv[3][3];
                              cell_geometry(X, q0, q1, q2, x, J);
                              const size_t cell = ((size_t)cz * nc[1] + cy) * nc[0] + cx;
                              const double f    = rhs_values ? rhs_values[cell * nq + q] : rhs_function(x);
                              v[q] = f * b->wq[q0] * b->wq[q1] * b->wq[q2] * std::fabs(invert3(J, inv));
                              continue;
                            }
                          const double x[3] = {lo[0] + (cx + b->xq[q0]) * h[0],
                                               lo[1] + (cy + b->xq[q1]) * h[1],
                                               lo[2] + (cz + b->xq[q2]) * h[2]};
                          const size_t cell = ((size_t)cz * nc[1] + cy) * nc[0] + cx;
                          const double f    = rhs_values ? rhs_values[cell * nq + q] : rhs_function(x);
                          v[q] = f * b->wq[q0] * b->wq[q1] * b->wq[q2] * detJ;
                        }
                  // S^T in x, y, z (plain loops; set-up only)
                  for (int dir = 0; dir < 3; ++dir)
                    {
                      const int stride = dir == 0 ? 1 : (dir == 1 ? n : n * n);
                      for (int a = 0; a < n; ++a)
                        for (int c = 0; c < n; ++c)
                          {
                            const int base =
                              dir == 0 ? (a * n + c) * n : (dir == 1 ? a * n * n + c : a * n + c);
                            for (int o = 0; o < n; ++o)
                              {
                                double s = 0;
                                for (int i = 0; i < n; ++i)
                                  s += b->S[i * n + o] * v[base + i * stride];
                                t[base + o * stride] = s;
                              }
                          }
                      v.swap(t);
                    }
                  for (int k = 0, l = 0; k < n; ++k)
                    for (int j = 0; j < n; ++j)
                      for (int i = 0; i < n; ++i, ++l)
                        {
                          const int gi = cx * p + i, gj = cy * p + j, gk = cz * p + k;
                          if (!on_boundary(gi, gj, gk))
                            dst[((size_t)gk * nd[1] + gj) * nd[0] + gi] += (T)v[l];
                        }
                }
        }
    }
  };

  // ---------------------------------------------------------------- vector helpers
  template <typename T>
  static double
  dot(const vector<T> &a, const vector<T> &b)
  {
    double s = 0;
#pragma omp parallel for reduction(+ : s)
    for (size_t i = 0; i < a.size(); ++i)
      s += (double)a[i] * (double)b[i];
    return s;
  }

  template <typename A, typename B>
  static void
  convert(vector<A> &dst, const vector<B> &src)
  {
#pragma omp parallel for
    for (size_t i = 0; i < src.size(); ++i)
      dst[i] = (A)src[i];
  }

  // extreme eigenvalues of a symmetric tridiagonal matrix by Sturm bisection
  static void
  tridiag_extreme_eigenvalues(const vector<double> &diag,
                              const vector<double> &off,
                              double               &emin,
                              double               &emax)
  {
    const int k = diag.size();
    if (k == 0)
      {
        emin = emax = 1;
        return;
      }
    double glo = 1e300, ghi = -1e300;
    for (int i = 0; i < k; ++i)
      {
        const double r = (i > 0 ? std::fabs(off[i - 1]) : 0.) + (i < k - 1 ? std::fabs(off[i]) : 0.);
        glo            = std::min(glo, diag[i] - r);
        ghi            = std::max(ghi, diag[i] + r);
      }
    auto count_below = [&](const double x) {
      int    cnt = 0;
      double d   = 1;
      for (int i = 0; i < k; ++i)
        {
          d = diag[i] - x - (i > 0 ? off[i - 1] * off[i - 1] / d : 0.);
          if (d == 0)
            d = 1e-300;
          if (d < 0)
            ++cnt;
        }
      return cnt;
    };
    auto kth = [&](const int idx) { // smallest x with count_below(x) > idx
      double a = glo, c = ghi;
      for (int it = 0; it < 200; ++it)
        {
          const double mid = 0.5 * (a + c);
          if (count_below(mid) > idx)
            c = mid;
          else
            a = mid;
        }
      return 0.5 * (a + c);
    };
    emin = kth(0);
    emax = kth(k - 1);
  }

  // ---------------------------------------------------------------- Chebyshev
  // deal.II PreconditionChebyshev, polynomial_type first_kind, preconditioner =
  // point Jacobi (multigrid_solver.h:271-288); SURVEY.md appendix A.5.
  template <typename T>
  struct Chebyshev
  {
    const LevelOp<T> *A = nullptr;
    double            smoothing_range = 20.;
    unsigned int      degree          = 1;
    unsigned int      eig_cg_n_iterations = 15;
    double            min_eig = 1, max_eig = 1; // raw CG/Lanczos estimates
    double            theta = 1, delta = 1;
    unsigned int      cg_iterations = 0;
    mutable vector<T> solution_old, temp1;

    static constexpr unsigned int invalid_degree = 0xffffffffu;

    // initial guess "global index mod 11, mean removed, constrained := 0"
    void
    initial_guess(vector<T> &v) const
    {
      const size_t N = A->ndofs;
      v.resize(N);
      double mean = 0;
      for (size_t i = 0; i < N; ++i)
        {
          v[i] = (T)(i % 11);
          mean += v[i];
        }
      mean /= N;
      for (size_t i = 0; i < N; ++i)
        v[i] -= (T)mean;
      A->for_boundary([&](const size_t i) { v[i] = 0; });
    }

    void
    initialize(const LevelOp<T> &op)
    {
      A              = &op;
      const size_t N = A->ndofs;
      solution_old.assign(N, 0);
      temp1.assign(N, 0);

      // Lanczos/CG eigenvalue estimate of D^-1 A
      vector<T> x(N, 0), g, h(N), d(N);
      initial_guess(g);
      const vector<T> &P = A->inv_diag;
      vector<double>   diag, off;
      // g = -b
      for (size_t i = 0; i < N; ++i)
        g[i] = -g[i];
      double res = std::sqrt(dot(g, g));
      cg_iterations = 0;
      if (res > 1e-10 && eig_cg_n_iterations > 0)
        {
          for (size_t i = 0; i < N; ++i)
            {
              h[i] = P[i] * g[i];
              d[i] = -h[i];
            }
          double gh = dot(g, h);
          double alpha_old = 0, beta_old = 0;
          for (unsigned int it = 1;; ++it)
            {
              A->vmult(h.data(), d.data());
              const double alpha = gh / dot(d, h);
              for (size_t i = 0; i < N; ++i)
                {
                  x[i] += (T)alpha * d[i];
                  g[i] += (T)alpha * h[i];
                }
              res = std::sqrt(dot(g, g));
              // Lanczos coefficients
              if (it == 1)
                diag.push_back(1. / alpha);
              else
                {
                  diag.push_back(1. / alpha + beta_old / alpha_old);
                  off.push_back(std::sqrt(beta_old) / alpha_old);
                }
              cg_iterations = it;
              if (it >= eig_cg_n_iterations || res <= 1e-10)
                break;
              for (size_t i = 0; i < N; ++i)
                h[i] = P[i] * g[i];
              const double gh_new = dot(g, h);
              const double beta   = gh_new / gh;
              gh                  = gh_new;
              for (size_t i = 0; i < N; ++i)
                d[i] = (T)beta * d[i] - h[i];
              alpha_old = alpha;
              beta_old  = beta;
            }
          tridiag_extreme_eigenvalues(diag, off, min_eig, max_eig);
        }
      else
        min_eig = max_eig = 1;

      const double max_est = 1.2 * max_eig;
      const double alpha   = smoothing_range > 1. ? max_est / smoothing_range :
                                                    std::min(0.9 * max_est, min_eig);
      if (degree == invalid_degree)
        {
          const double actual_range = max_est / alpha;
          const double sigma =
            (1. - std::sqrt(1. / actual_range)) / (1. + std::sqrt(1. / actual_range));
          const double eps = smoothing_range;
          degree           = 1 + static_cast<unsigned int>(
                         std::log(1. / eps + std::sqrt(1. / eps / eps - 1.)) / std::log(1. / sigma));
        }
      theta = (max_est + alpha) * 0.5;
      delta = (max_est - alpha) * 0.5;
    }

    // x_{k+1} = x_k + f1 (x_k - x_{k-1}) + f2 P^-1 (b - A x_k); temp1 holds A x_k
    void
    update(const unsigned int iteration_index,
           const double       f1,
           const double       f2,
           vector<T>         &x,
           const vector<T>   &b) const
    {
      const size_t N  = A->ndofs;
      const T     *P  = A->inv_diag.data();
      const T      F1 = (T)f1, F2 = (T)f2;
      if (iteration_index == 0)
        {
#pragma omp parallel for
          for (size_t i = 0; i < N; ++i)
            x[i] = F2 * P[i] * b[i];
        }
      else if (iteration_index == 1)
        {
#pragma omp parallel for
          for (size_t i = 0; i < N; ++i)
            {
              // x^{n+1} = x^n + f1 x^n + f2 P^-1 (b - A x^n)   (x^{n-1} = 0 implied)
              const T xi      = x[i];
              solution_old[i] = xi;
              x[i]            = (T(1) + F1) * xi + F2 * P[i] * (b[i] - temp1[i]);
            }
        }
      else
        {
#pragma omp parallel for
          for (size_t i = 0; i < N; ++i)
            {
              const T xi      = x[i];
              const T xo      = solution_old[i];
              solution_old[i] = xi;
              x[i]            = xi + F1 * (xi - xo) + F2 * P[i] * (b[i] - temp1[i]);
            }
        }
    }

    void
    loop(vector<T> &x, const vector<T> &b, const unsigned int index_offset) const
    {
      if (degree < 2)
        return;
      double       rhok  = delta / theta;
      const double sigma = theta / delta;
      for (unsigned int k = 0; k < degree - 1; ++k)
        {
          A->vmult(temp1.data(), x.data());
          const double rhokp = 1. / (2. * sigma - rhok);
          const double f1 = rhokp * rhok, f2 = 2. * rhokp / delta;
          rhok = rhokp;
          update(k + index_offset, f1, f2, x, b);
        }
    }

    void
    vmult(vector<T> &x, const vector<T> &b) const
    {
      update(0, 0., 1. / theta, x, b);
      loop(x, b, 1);
    }

    void
    step(vector<T> &x, const vector<T> &b) const
    {
      A->vmult(temp1.data(), x.data());
      update(1, 0., 1. / theta, x, b);
      loop(x, b, 2);
    }
  };

  // ---------------------------------------------------------------- transfer
  // MGTransferMatrixFree restated: P = nodal interpolation
  // coarse -> fine applied parent cell by parent cell with the 1-D embedding P1,
  // R = P^T with the fine entries weighted by 1/(number of parent patches sharing).
  template <typename T>
  struct Transfer
  {
    const Basis1D *b = nullptr;
    int            p, n, nf1;
    int            ncc[3], ndc[3], ndf[3];
    bool           constrained = false; // initialize_constraints(mg_constrained_dofs)
    vector<T>      P1;

    void
    init(const Basis1D &basis, const int ncoarse[3], const bool with_constraints)
    {
      b   = &basis;
      p   = basis.p;
      n   = basis.n;
      nf1 = 2 * p + 1;
      for (int d = 0; d < 3; ++d)
        {
          ncc[d] = ncoarse[d];
          ndc[d] = ncoarse[d] * p + 1;
          ndf[d] = 2 * ncoarse[d] * p + 1;
        }
      constrained = with_constraints;
      P1.resize(nf1 * n);
      for (int i = 0; i < nf1 * n; ++i)
        P1[i] = (T)basis.P1[i];
    }

    bool
    coarse_boundary(const int i, const int j, const int k) const
    {
      return i == 0 || j == 0 || k == 0 || i == ndc[0] - 1 || j == ndc[1] - 1 || k == ndc[2] - 1;
    }

    // fine (+)= P coarse
    void
    prolongate(vector<T> &fine, const vector<T> &coarse, const bool add) const
    {
      const int nf = nf1;
      vector<T> tmp;
      T        *out = fine.data();
      if (add)
        {
          tmp.assign(fine.size(), 0);
          out = tmp.data();
        }
#pragma omp parallel for collapse(2)
      for (int cz = 0; cz < ncc[2]; ++cz)
        for (int cy = 0; cy < ncc[1]; ++cy)
          for (int cx = 0; cx < ncc[0]; ++cx)
            {
              vector<T> u(n * n * n), a(nf * n * n), c(nf * nf * n), f(nf * nf * nf);
              for (int k = 0, l = 0; k < n; ++k)
                for (int j = 0; j < n; ++j)
                  for (int i = 0; i < n; ++i, ++l)
                    {
                      const int gi = cx * p + i, gj = cy * p + j, gk = cz * p + k;
                      u[l] = (constrained && coarse_boundary(gi, gj, gk)) ?
                               T(0) :
                               coarse[((size_t)gk * ndc[1] + gj) * ndc[0] + gi];
                    }
              // x
              for (int k = 0; k < n; ++k)
                for (int j = 0; j < n; ++j)
                  for (int o = 0; o < nf; ++o)
                    {
                      T s = 0;
                      for (int i = 0; i < n; ++i)
                        s += P1[o * n + i] * u[(k * n + j) * n + i];
                      a[(k * n + j) * nf + o] = s;
                    }
              // y
              for (int k = 0; k < n; ++k)
                for (int o = 0; o < nf; ++o)
                  for (int i = 0; i < nf; ++i)
                    {
                      T s = 0;
                      for (int j = 0; j < n; ++j)
                        s += P1[o * n + j] * a[(k * n + j) * nf + i];
                      c[(k * nf + o) * nf + i] = s;
                    }
              // z
              for (int o = 0; o < nf; ++o)
                for (int j = 0; j < nf; ++j)
                  for (int i = 0; i < nf; ++i)
                    {
                      T s = 0;
                      for (int k = 0; k < n; ++k)
                        s += P1[o * n + k] * c[(k * nf + j) * nf + i];
                      f[(o * nf + j) * nf + i] = s;
                    }
              // values on shared fine DoFs agree between parents: plain overwrite
              for (int k = 0; k < nf; ++k)
                for (int j = 0; j < nf; ++j)
                  for (int i = 0; i < nf; ++i)
                    {
                      const int gi = 2 * cx * p + i, gj = 2 * cy * p + j, gk = 2 * cz * p + k;
                      out[((size_t)gk * ndf[1] + gj) * ndf[0] + gi] = f[(k * nf + j) * nf + i];
                    }
            }
      if (add)
        {
#pragma omp parallel for
          for (size_t i = 0; i < fine.size(); ++i)
            fine[i] += tmp[i];
        }
    }

    // coarse += P^T fine
    void
    restrict_and_add(vector<T> &coarse, const vector<T> &fine) const
    {
      const int nf = nf1;
      for (int colour = 0; colour < 2; ++colour)
        {
#pragma omp parallel for schedule(static)
          for (int cz = colour; cz < ncc[2]; cz += 2)
            for (int cy = 0; cy < ncc[1]; ++cy)
              for (int cx = 0; cx < ncc[0]; ++cx)
                {
                  vector<T> u(n * n * n), a(nf * n * n), c(nf * nf * n), f(nf * nf * nf);
                  for (int k = 0; k < nf; ++k)
                    for (int j = 0; j < nf; ++j)
                      for (int i = 0; i < nf; ++i)
                        {
                          const int gi = 2 * cx * p + i, gj = 2 * cy * p + j,
                                    gk = 2 * cz * p + k;
                          T w = 1;
                          if ((i == 0 && gi > 0) || (i == nf - 1 && gi < ndf[0] - 1))
                            w *= T(0.5);
                          if ((j == 0 && gj > 0) || (j == nf - 1 && gj < ndf[1] - 1))
                            w *= T(0.5);
                          if ((k == 0 && gk > 0) || (k == nf - 1 && gk < ndf[2] - 1))
                            w *= T(0.5);
                          f[(k * nf + j) * nf + i] =
                            w * fine[((size_t)gk * ndf[1] + gj) * ndf[0] + gi];
                        }
                  // z^T
                  for (int k = 0; k < n; ++k)
                    for (int j = 0; j < nf; ++j)
                      for (int i = 0; i < nf; ++i)
                        {
                          T s = 0;
                          for (int o = 0; o < nf; ++o)
                            s += P1[o * n + k] * f[(o * nf + j) * nf + i];
                          c[(k * nf + j) * nf + i] = s;
                        }
                  // y^T
                  for (int k = 0; k < n; ++k)
                    for (int j = 0; j < n; ++j)
                      for (int i = 0; i < nf; ++i)
                        {
                          T s = 0;
                          for (int o = 0; o < nf; ++o)
                            s += P1[o * n + j] * c[(k * nf + o) * nf + i];
                          a[(k * n + j) * nf + i] = s;
                        }
                  // x^T
                  for (int k = 0; k < n; ++k)
                    for (int j = 0; j < n; ++j)
                      for (int i = 0; i < n; ++i)
                        {
                          T s = 0;
                          for (int o = 0; o < nf; ++o)
                            s += P1[o * n + i] * a[(k * n + j) * nf + o];
                          u[(k * n + j) * n + i] = s;
                        }
                  for (int k = 0, l = 0; k < n; ++k)
                    for (int j = 0; j < n; ++j)
                      for (int i = 0; i < n; ++i, ++l)
                        {
                          const int gi = cx * p + i, gj = cy * p + j, gk = cz * p + k;
                          if (!(constrained && coarse_boundary(gi, gj, gk)))
                            coarse[((size_t)gk * ndc[1] + gj) * ndc[0] + gi] += u[l];
                        }
                }
        }
    }
  };

  // ---------------------------------------------------------------- solver
  // MultigridSolver<3,p,float,double> (multigrid_solver.h:96-782)
  struct MultigridSolver
  {
    Basis1D      basis;
    int          p, nlevels;
    unsigned int degree_pre, degree_post, n_cycles;
    Fn           f_solution, f_rhs, f_coef;
    bool         const_coef;
    double       lo[3], hi[3];

    vector<LevelOp<float>>    matrix;
    vector<LevelOp<double>>   matrix_dp;
    vector<Chebyshev<float>>  smooth;
    vector<Transfer<float>>   transfer;               // [level] : level-1 -> level
    vector<Transfer<double>>  mg_transfer_no_boundary;
    vector<vector<double>>    solution, rhs, residual;
    vector<vector<float>>     defect, t, solution_update;
    vector<vector<std::pair<size_t, double>>> inhomogeneous_bc;
    vector<vector<double>>    exact_values; // caller-supplied exact solution at the quadrature points, per level
    double                    timings_setup[3] = {0, 0, 0};

    MultigridSolver(const int     degree,
                    const int     nsub[3],
                    const int     n_refine,
                    const double  lower[3],
                    const double  upper[3],
                    const unsigned int pre,
                    const unsigned int post,
                    const unsigned int cycles,
                    const int     problem,
                    const Geometry &geometry = Geometry())
      : p(degree)
      , nlevels(n_refine + 1)
      , degree_pre(pre)
      , degree_post(post)
      , n_cycles(cycles)
    {
      basis = make_basis(p);
      if (problem == 0)
        {
          f_solution = sol_cube;
          f_rhs      = rhs_cube;
          f_coef     = coef_one;
          const_coef = true;
        }
      else
        {
          f_solution = sol_shell;
          f_rhs      = rhs_shell;
          f_coef     = coef_shell;
          const_coef = false;
        }
      for (int d = 0; d < 3; ++d)
        {
          lo[d] = lower[d];
          hi[d] = upper[d];
        }
      matrix.resize(nlevels);
      matrix_dp.resize(nlevels);
      smooth.resize(nlevels);
      transfer.resize(nlevels);
      mg_transfer_no_boundary.resize(nlevels);
      solution.resize(nlevels);
      rhs.resize(nlevels);
      residual.resize(nlevels);
      defect.resize(nlevels);
      t.resize(nlevels);
      solution_update.resize(nlevels);
      inhomogeneous_bc.resize(nlevels);

      auto t0 = std::chrono::steady_clock::now();
      for (int l = 0; l < nlevels; ++l)
        {
          int nc[3];
          for (int d = 0; d < 3; ++d)
            nc[d] = nsub[d] << l;
          matrix[l].init(basis, nc, lo, hi);
          matrix[l].geom = geometry;
          matrix[l].evaluate_coefficient(f_coef, const_coef);
          matrix_dp[l].init(basis, nc, lo, hi);
          matrix_dp[l].geom = geometry;
          matrix_dp[l].evaluate_coefficient(f_coef, const_coef);
          const size_t N = matrix[l].ndofs;
          defect[l].assign(N, 0);
          t[l].assign(N, 0);
          solution_update[l].assign(N, 0);
          solution[l].assign(N, 0);
          rhs[l].assign(N, 0);
          residual[l].assign(N, 0);
          if (l > 0)
            {
              transfer[l].init(basis, matrix[l - 1].nc, true);
              mg_transfer_no_boundary[l].init(basis, matrix[l - 1].nc, false);
            }
        }
      timings_setup[0] = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
      t0               = std::chrono::steady_clock::now();
      // multigrid_solver.h:224-262
      for (int l = 0; l < nlevels; ++l)
        {
          const LevelOp<double> &op = matrix_dp[l];
          for (int k = 0; k < op.nd[2]; ++k)
            for (int j = 0; j < op.nd[1]; ++j)
              for (int i = 0; i < op.nd[0]; ++i)
                if (op.on_boundary(i, j, k))
                  {
                    double x[3] = {support_point(op, 0, i), support_point(op, 1, j), support_point(op, 2, k)};
                    if (op.geom.kind != 0)
                      {
                        // mapped support point = the geometry map at the node's mesh coordinates
                        const double xi[3] = {unit_coordinate(op, 0, i), unit_coordinate(op, 1, j),
                                              unit_coordinate(op, 2, k)};
                        op.geom.point(xi, x);
                      }
                    const double value = f_solution(x);
                    if (value != 0.0)
                      inhomogeneous_bc[l].emplace_back(((size_t)k * op.nd[1] + j) * op.nd[0] + i,
                                                       value);
                  }
          for (auto &i : inhomogeneous_bc[l])
            solution[l][i.first] = i.second;
          op.compute_residual(rhs[l].data(), solution[l].data(), f_rhs);
        }
      timings_setup[1] = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
      t0               = std::chrono::steady_clock::now();
      // multigrid_solver.h:268-289
      for (int l = 0; l < nlevels; ++l)
        {
          if (l > 0)
            {
              smooth[l].smoothing_range     = 20.;
              smooth[l].degree              = degree_pre;
              smooth[l].eig_cg_n_iterations = 15;
            }
          else
            {
              smooth[l].smoothing_range     = 1e-3;
              smooth[l].degree              = Chebyshev<float>::invalid_degree;
              smooth[l].eig_cg_n_iterations = matrix[0].m();
            }
          matrix[l].compute_diagonal();
          smooth[l].initialize(matrix[l]);
        }
      timings_setup[2] = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    }

    double
    support_point(const LevelOp<double> &op, const int d, const int i) const
    {
      const int cell = std::min(i / p, op.nc[d] - 1);
      return op.lo[d] + (cell + basis.gll[i - cell * p]) * op.h[d];
    }

    // ---- problem data supplied by samples (the Function<dim> arguments of multigrid_solver.h:100-106)
    void
    quadrature_points(const int level, double *out) const
    {
      const LevelOp<double> &op = matrix_dp[level];
      const int              n  = op.n;
      for (int cz = 0; cz < op.nc[2]; ++cz)
        for (int cy = 0; cy < op.nc[1]; ++cy)
          for (int cx = 0; cx < op.nc[0]; ++cx)
            {
              const size_t   cell = ((size_t)cz * op.nc[1] + cy) * op.nc[0] + cx;
              vector<double> X;
              if (op.geom.kind != 0)
                op.cell_nodes(cx, cy, cz, X);
              for (int q2 = 0, q = 0; q2 < n; ++q2)
                for (int q1 = 0; q1 < n; ++q1)
                  for (int q0 = 0; q0 < n; ++q0, ++q)
                    {
                      double x[3] = {op.lo[0] + (cx + basis.xq[q0]) * op.h[0], op.lo[1] + (cy + basis.xq[q1]) * op.h[1],
                                     op.lo[2] + (cz + basis.xq[q2]) * op.h[2]};
                      if (op.geom.kind != 0)
                        {
                          double J[3][3];
                          op.cell_geometry(X, q0, q1, q2, x, J);
                        }
                      for (int d = 0; d < 3; ++d)
                        out[(cell * n * n * n + q) * 3 + d] = x[d];
                    }
            }
    }

    void
    support_points(const int level, double *out) const
    {
      const LevelOp<double> &op = matrix_dp[level];
      for (int k = 0; k < op.nd[2]; ++k)
        for (int j = 0; j < op.nd[1]; ++j)
          for (int i = 0; i < op.nd[0]; ++i)
            {
              double x[3] = {support_point(op, 0, i), support_point(op, 1, j), support_point(op, 2, k)};
              if (op.geom.kind != 0)
                {
                  const double xi[3] = {unit_coordinate(op, 0, i), unit_coordinate(op, 1, j), unit_coordinate(op, 2, k)};
                  op.geom.point(xi, x);
                }
              for (int d = 0; d < 3; ++d)
                out[(((size_t)k * op.nd[1] + j) * op.nd[0] + i) * 3 + d] = x[d];
            }
    }

    void
    set_coefficient_values(const int level, const double *c)
    {
      matrix[level].set_coefficient_values(c);
      matrix_dp[level].set_coefficient_values(c);
    }

    void
    reinit_smoothers()
    {
      for (int l = 0; l < nlevels; ++l)
        {
          if (l == 0)
            {
              smooth[0].degree              = Chebyshev<float>::invalid_degree;
              smooth[0].eig_cg_n_iterations = matrix[0].m();
            }
          matrix[l].compute_diagonal();
          smooth[l].initialize(matrix[l]);
        }
    }

    void
    set_rhs_values(const int level, const double *f, const double *u_bc)
    {
      const LevelOp<double> &op = matrix_dp[level];
      inhomogeneous_bc[level].clear();
      std::fill(solution[level].begin(), solution[level].end(), 0.);
      if (u_bc)
        for (int k = 0; k < op.nd[2]; ++k)
          for (int j = 0; j < op.nd[1]; ++j)
            for (int i = 0; i < op.nd[0]; ++i)
              if (op.on_boundary(i, j, k))
                {
                  const size_t idx = ((size_t)k * op.nd[1] + j) * op.nd[0] + i;
                  if (u_bc[idx] != 0.0)
                    inhomogeneous_bc[level].emplace_back(idx, u_bc[idx]);
                }
      for (auto &i : inhomogeneous_bc[level])
        solution[level][i.first] = i.second;
      op.compute_residual(rhs[level].data(), solution[level].data(), f_rhs, f);
    }

    void
    set_exact_values(const int level, const double *u)
    {
      exact_values.resize(nlevels);
      const size_t nq = matrix_dp[level].ncells * (size_t)matrix_dp[level].n * matrix_dp[level].n * matrix_dp[level].n;
      exact_values[level].assign(u, u + nq);
    }

    double
    unit_coordinate(const LevelOp<double> &op, const int d, const int i) const
    {
      const int cell = std::min(i / p, op.nc[d] - 1);
      return (cell + basis.gll[i - cell * p]) / op.nc[d];
    }

    // multigrid_solver.h:298-343
    double
    compute_l2_error(const int level)
    {
      for (auto &i : inhomogeneous_bc[level])
        solution[level][i.first] = i.second;
      const LevelOp<double> &op   = matrix_dp[level];
      const int              n    = op.n;
      const double           detJ = op.h[0] * op.h[1] * op.h[2];
      double                 err = 0, vol = 0;
#pragma omp parallel for collapse(2) reduction(+ : err, vol)
      for (int cz = 0; cz < op.nc[2]; ++cz)
        for (int cy = 0; cy < op.nc[1]; ++cy)
          for (int cx = 0; cx < op.nc[0]; ++cx)
            {
              vector<double> u(n * n * n), w(n * n * n), X;
              if (op.geom.kind != 0)
                op.cell_nodes(cx, cy, cz, X);
              for (int k = 0, l = 0; k < n; ++k)
                for (int j = 0; j < n; ++j)
                  for (int i = 0; i < n; ++i, ++l)
                    u[l] = solution[level][((size_t)(cz * p + k) * op.nd[1] + cy * p + j) * op.nd[0] +
                                           cx * p + i];
              for (int dir = 0; dir < 3; ++dir)
                {
                  const int stride = dir == 0 ? 1 : (dir == 1 ? n : n * n);
                  for (int a = 0; a < n; ++a)
                    for (int c = 0; c < n; ++c)
                      {
                        const int base =
                          dir == 0 ? (a * n + c) * n : (dir == 1 ? a * n * n + c : a * n + c);
                        for (int o = 0; o < n; ++o)
                          {
                            double s = 0;
                            for (int i = 0; i < n; ++i)
                              s += basis.S[o * n + i] * u[base + i * stride];
                            w[base + o * stride] = s;
                          }
                      }
                  u.swap(w);
                }
              for (int q2 = 0, q = 0; q2 < n; ++q2)
                for (int q1 = 0; q1 < n; ++q1)
                  for (int q0 = 0; q0 < n; ++q0, ++q)
                    {
                      double x[3] = {op.lo[0] + (cx + basis.xq[q0]) * op.h[0],
                                     op.lo[1] + (cy + basis.xq[q1]) * op.h[1],
                                     op.lo[2] + (cz + basis.xq[q2]) * op.h[2]};
                      double JxW  = basis.wq[q0] * basis.wq[q1] * basis.wq[q2] * detJ;
                      if (op.geom.kind != 0)
                        {
                          double J[3][3], inv[3][3];
                          op.cell_geometry(X, q0, q1, q2, x, J);
                          JxW = basis.wq[q0] * basis.wq[q1] * basis.wq[q2] * std::fabs(invert3(J, inv));
                        }
                      const size_t cell = ((size_t)cz * op.nc[1] + cy) * op.nc[0] + cx;
                      const bool   have = (int)exact_values.size() > level && !exact_values[level].empty();
                      const double diff = u[q] - (have ? exact_values[level][cell * n * n * n + q] : f_solution(x));
                      err += diff * diff * JxW;
                      vol += JxW;
                    }
            }
      return std::sqrt(err / vol);
    }

    // multigrid_solver.h:641-681
    void
    v_cycle(const int level, const unsigned int my_n_cycles)
    {
      if (level == 0)
        {
          smooth[0].vmult(solution_update[0], defect[0]);
          return;
        }
      for (unsigned int c = 0; c < my_n_cycles; ++c)
        {
          if (c == 0)
            smooth[level].vmult(solution_update[level], defect[level]);
          else
            smooth[level].step(solution_update[level], defect[level]);
          matrix[level].vmult_residual(defect[level].data(),
                                       solution_update[level].data(),
                                       t[level].data());
          std::fill(defect[level - 1].begin(), defect[level - 1].end(), 0.f);
          transfer[level].restrict_and_add(defect[level - 1], t[level]);
          v_cycle(level - 1, 1);
          transfer[level].prolongate(solution_update[level], solution_update[level - 1], true);
          smooth[level].step(solution_update[level], defect[level]);
        }
    }

    // multigrid_solver.h:387-476; trace[level*4+{0,1,2,3}] = error start, residual
    // start, residual end, error end (only if do_analyze)
    double
    solve(const bool do_analyze, double *trace)
    {
      double reduction_rate = 1.;
      convert(defect[0], rhs[0]);
      smooth[0].vmult(t[0], defect[0]);
      smooth[0].step(t[0], defect[0]);
      convert(solution[0], t[0]);
      for (int level = 1; level < nlevels; ++level)
        {
          for (auto &i : inhomogeneous_bc[level - 1])
            solution[level - 1][i.first] = i.second;
          mg_transfer_no_boundary[level].prolongate(solution[level], solution[level - 1], false);
          double init_residual = 1.;
          if (do_analyze && trace)
            trace[level * 4 + 0] = compute_l2_error(level);
          for (auto &i : inhomogeneous_bc[level])
            solution[level][i.first] = 0;
          matrix_dp[level].vmult_residual(rhs[level].data(),
                                          solution[level].data(),
                                          residual[level].data());
          convert(defect[level], residual[level]);
          if (do_analyze)
            {
              init_residual = std::sqrt(dot(residual[level], residual[level]));
              if (trace)
                trace[level * 4 + 1] = init_residual;
            }
          v_cycle(level, n_cycles);
          for (size_t i = 0; i < solution[level].size(); ++i)
            solution[level][i] += solution_update[level][i];
          if (do_analyze)
            {
              for (auto &i : inhomogeneous_bc[level])
                solution[level][i.first] = 0;
              matrix_dp[level].vmult(residual[level].data(), solution[level].data());
              for (size_t i = 0; i < residual[level].size(); ++i)
                residual[level][i] = rhs[level][i] - residual[level][i];
              const double res_norm = std::sqrt(dot(residual[level], residual[level]));
              reduction_rate        = std::pow(res_norm / init_residual, 1. / n_cycles);
              if (trace)
                {
                  trace[level * 4 + 2] = res_norm;
                  trace[level * 4 + 3] = compute_l2_error(level);
                }
            }
        }
      return reduction_rate;
    }

    // multigrid_solver.h:498-510
    void
    vmult(vector<double> &dst, const vector<double> &src)
    {
      const int L = nlevels - 1;
      convert(defect[L], src);
      v_cycle(L, 1);
      convert(dst, solution_update[L]);
    }

    // multigrid_solver.h:483-493 with deal.II SolverCG + ReductionControl(1000,1e-16,1e-9)
    // 
    std::pair<unsigned int, double>
    solve_cg(vector<double> *history = nullptr, const unsigned int max_it = 1000, const double abs_tol = 1e-16,
             const double rel_tol = 1e-9)
    {
      const int              L = nlevels - 1;
      const LevelOp<double> &A = matrix_dp[L];
      const size_t           N = A.ndofs;
      vector<double>        &x = solution[L];
      std::fill(x.begin(), x.end(), 0.);
      vector<double> g(N), h(N), d(N);
      for (size_t i = 0; i < N; ++i)
        g[i] = -rhs[L][i];
      double       res  = std::sqrt(dot(g, g));
      const double res0 = res;
      if (history)
        history->push_back(res);
      unsigned int it = 0;
      if (res > abs_tol && res0 > 0)
        {
          vmult(h, g);
          for (size_t i = 0; i < N; ++i)
            d[i] = -h[i];
          double gh = dot(g, h);
          while (true)
            {
              ++it;
              A.vmult(h.data(), d.data());
              const double alpha = gh / dot(d, h);
#pragma omp parallel for
              for (size_t i = 0; i < N; ++i)
                {
                  x[i] += alpha * d[i];
                  g[i] += alpha * h[i];
                }
              res = std::sqrt(dot(g, g));
              if (history)
                history->push_back(res);
              if (res <= abs_tol || res <= rel_tol * res0 || it >= max_it)
                break;
              vmult(h, g);
              const double gh_new = dot(g, h);
              const double beta   = gh_new / gh;
              gh                  = gh_new;
#pragma omp parallel for
              for (size_t i = 0; i < N; ++i)
                d[i] = beta * d[i] - h[i];
            }
        }
      return std::make_pair(it, it > 0 ? std::pow(res / res0, 1. / it) : 0.);
    }

    // ------------------------------------------------------------------ minimal_surface/program.cc
    // Newton's method with line search for -div(grad u / sqrt(1 + |grad u|^2)) = 0 with Dirichlet data, the linear
    // systems solved by CG with the V-cycle on the linearised operator (LaplaceProblem::solve, :412-572).
    vector<vector<float>> coefficient_solutions;
    LevelOp<double>       residual_op;
    vector<double>        search_direction, tentative_solution, system_rhs;

    // interpolate_boundary_values (:396-408): the Dirichlet data in the boundary entries of `solution`, zero inside
    void
    ms_set_boundary_values(const double *u_bc)
    {
      const int              L  = nlevels - 1;
      const LevelOp<double> &op = matrix_dp[L];
      std::fill(solution[L].begin(), solution[L].end(), 0.);
      inhomogeneous_bc[L].clear();
      for (int k = 0; k < op.nd[2]; ++k)
        for (int j = 0; j < op.nd[1]; ++j)
          for (int i = 0; i < op.nd[0]; ++i)
            if (op.on_boundary(i, j, k))
              {
                const size_t idx = ((size_t)k * op.nd[1] + j) * op.nd[0] + i;
                if (u_bc)
                  solution[L][idx] = u_bc[idx];
                else
                  {
                    double x[3] = {support_point(op, 0, i), support_point(op, 1, j), support_point(op, 2, k)};
                    if (op.geom.kind != 0)
                      {
                        const double xi[3] = {unit_coordinate(op, 0, i), unit_coordinate(op, 1, j), unit_coordinate(op, 2, k)};
                        op.geom.point(xi, x);
                      }
                    solution[L][idx] = f_solution(x);
                  }
                inhomogeneous_bc[L].emplace_back(idx, solution[L][idx]);
              }
      residual_op      = matrix_dp[L];
      coefficient_solutions.resize(nlevels);
      smooth[0].smoothing_range = 1e-3; // :470-474
    }

    // the FE_Q restriction matrices of :433-447 (interpolation of the fine function in the support points of the coarse
    // cell, each child contributing the points inside it) as one 1-D matrix R[i][m] over the 2p+1 fine points of a
    // coarse cell, applied in the three directions
    void
    ms_restrict_solution(const int level, vector<float> &coarse, const vector<float> &fine) const
    {
      const LevelOp<float> &oc = matrix[level - 1], &of = matrix[level];
      const int             n = oc.n, nf = 2 * p + 1;
      vector<long double>   nodes(basis.gll.begin(), basis.gll.end());
      vector<double>        R(n * nf, 0.);
      for (int i = 0; i < n; ++i)
        {
          const int child = basis.gll[i] <= 0.5 ? 0 : 1;
          for (int m = 0; m < n; ++m)
            R[i * nf + child * p + m] = (double)lagrange(nodes, m, 2 * (long double)basis.gll[i] - child);
        }
      coarse.assign(oc.ndofs, 0.f);
#pragma omp parallel for collapse(2)
      for (int k = 0; k < oc.nd[2]; ++k)
        for (int j = 0; j < oc.nd[1]; ++j)
          for (int i = 0; i < oc.nd[0]; ++i)
            {
              const int c[3] = {std::min(i / p, oc.nc[0] - 1), std::min(j / p, oc.nc[1] - 1), std::min(k / p, oc.nc[2] - 1)};
              const int l[3] = {i - c[0] * p, j - c[1] * p, k - c[2] * p};
              double    sum  = 0;
              for (int mz = 0; mz < nf; ++mz)
                for (int my = 0; my < nf; ++my)
                  for (int mx = 0; mx < nf; ++mx)
                    {
                      const double w = R[l[0] * nf + mx] * R[l[1] * nf + my] * R[l[2] * nf + mz];
                      if (w != 0.)
                        sum += w * (double)fine[((size_t)(2 * c[2] * p + mz) * of.nd[1] + 2 * c[1] * p + my) * of.nd[0] +
                                                2 * c[0] * p + mx];
                    }
              coarse[((size_t)k * oc.nd[1] + j) * oc.nd[0] + i] = (float)sum;
            }
    }

    // MinimalSurfaceOperator::compute_residual (:169-198): dst_i = -(grad phi_i, grad u / sqrt(1 + |grad u|^2)) on the
    // rows without Dirichlet condition (first_time: the Laplace residual)
    void
    ms_residual(vector<double> &dst, const vector<double> &u, const bool first_time)
    {
      residual_op.evaluate_solution_coefficient(first_time ? 0 : 2, u.data());
      dst.assign(u.size(), 0.);
      residual_op.apply_add(dst.data(), u.data(), /*read_plain=*/true);
      for (double &v : dst)
        v = -v;
    }

    // LaplaceProblem::solve(first_time) (:412-572); info = {initial residual, final residual, linear iterations,
    // line search steps, step length}
    void
    ms_newton_step(const bool first_time, double *info)
    {
      const int L = nlevels - 1;
      coefficient_solutions[L].resize(solution[L].size());
      convert(coefficient_solutions[L], solution[L]);
      for (int l = L; l > 0; --l)
        ms_restrict_solution(l, coefficient_solutions[l - 1], coefficient_solutions[l]);
      matrix_dp[L].evaluate_solution_coefficient(first_time ? 0 : 1, solution[L].data());
      for (int l = 0; l < nlevels; ++l)
        {
          matrix[l].evaluate_solution_coefficient(first_time ? 0 : 1, coefficient_solutions[l].data());
          if (l == 0)
            {
              smooth[0].degree              = Chebyshev<float>::invalid_degree;
              smooth[0].eig_cg_n_iterations = matrix[0].m();
            }
          matrix[l].compute_diagonal();
          smooth[l].initialize(matrix[l]);
        }
      ms_residual(system_rhs, solution[L], first_time);
      const double initial_residual_norm = std::sqrt(dot(system_rhs, system_rhs));
      // ReductionControl(m(), 1e-13, 1e-4), zero start; the Newton update lives in solution[L] during the solve
      vector<double> current = solution[L];
      rhs[L]                 = system_rhs;
      const auto cg          = solve_cg(nullptr, (unsigned int)matrix_dp[L].m(), 1e-13, 1e-4);
      search_direction       = solution[L];
      solution[L]            = current;
      matrix_dp[L].for_boundary([&](const size_t i) { search_direction[i] = 0; }); // constraints.distribute
      double       final_residual_norm = initial_residual_norm, alpha = 1.;
      unsigned int n_steps             = 0;
      while (n_steps < 100)
        {
          tentative_solution = solution[L];
          for (size_t i = 0; i < tentative_solution.size(); ++i)
            tentative_solution[i] += alpha * search_direction[i];
          ms_residual(system_rhs, tentative_solution, false);
          final_residual_norm = std::sqrt(dot(system_rhs, system_rhs));
          if (final_residual_norm < initial_residual_norm)
            break;
          alpha /= 2.;
          ++n_steps;
        }
      solution[L] = tentative_solution;
      info[0]     = initial_residual_norm;
      info[1]     = final_residual_norm;
      info[2]     = cg.first;
      info[3]     = n_steps;
      info[4]     = alpha;
    }
  };
} // namespace orc

// ------------------------------------------------------------------ C interface
using namespace orc;

extern "C"
{
  void *
  orc_mg_create(int           degree,
                const int    *nsub,
                int           n_refine,
                const double *lower,
                const double *upper,
                unsigned int  pre,
                unsigned int  post,
                unsigned int  cycles,
                int           problem)
  {
    if (degree < 1 || degree > 8)
      return nullptr;
    return new MultigridSolver(degree, nsub, n_refine, lower, upper, pre, post, cycles, problem);
  }

  void
  orc_mg_quadrature_points(void *h, int level, double *out)
  {
    static_cast<MultigridSolver *>(h)->quadrature_points(level, out);
  }
  void
  orc_mg_support_points(void *h, int level, double *out)
  {
    static_cast<MultigridSolver *>(h)->support_points(level, out);
  }
  void
  orc_mg_set_coefficient_values(void *h, int level, const double *c)
  {
    static_cast<MultigridSolver *>(h)->set_coefficient_values(level, c);
  }
  void
  orc_mg_reinit_smoothers(void *h)
  {
    static_cast<MultigridSolver *>(h)->reinit_smoothers();
  }
  void
  orc_mg_set_rhs_values(void *h, int level, const double *f, const double *u_bc)
  {
    static_cast<MultigridSolver *>(h)->set_rhs_values(level, f, u_bc);
  }
  void
  orc_mg_set_exact_values(void *h, int level, const double *u)
  {
    static_cast<MultigridSolver *>(h)->set_exact_values(level, u);
  }

  // the same on a mapped mesh: geometry 0 box, 1 shell sector (params r0, r1, half-angle), 2 affine (params A row-major)
  void *
  orc_mg_create_mapped(int           degree,
                       const int    *nsub,
                       int           n_refine,
                       const double *lower,
                       const double *upper,
                       unsigned int  pre,
                       unsigned int  post,
                       unsigned int  cycles,
                       int           problem,
                       int           geometry,
                       const double *params)
  {
    if (degree < 1 || degree > 8 || geometry < 0 || geometry > 2)
      return nullptr;
    Geometry g;
    g.kind = geometry;
    for (int d = 0; d < 3; ++d)
      {
        g.lo[d] = lower[d];
        g.hi[d] = upper[d];
      }
    for (int i = 0; i < 12; ++i)
      g.par[i] = params[i];
    return new MultigridSolver(degree, nsub, n_refine, lower, upper, pre, post, cycles, problem, g);
  }

  void
  orc_mg_destroy(void *h)
  {
    delete static_cast<MultigridSolver *>(h);
  }

  void
  orc_mg_ms_set_boundary_values(void *h, const double *u_bc)
  {
    static_cast<MultigridSolver *>(h)->ms_set_boundary_values(u_bc);
  }

  void
  orc_mg_ms_newton_step(void *h, int first_time, double *info)
  {
    static_cast<MultigridSolver *>(h)->ms_newton_step(first_time != 0, info);
  }

  void
  orc_mg_ms_residual(void *h, const double *u, int first_time, double *dst)
  {
    MultigridSolver *m = static_cast<MultigridSolver *>(h);
    const size_t     N = m->matrix_dp[m->nlevels - 1].ndofs;
    vector<double>   in(u, u + N), out;
    m->ms_residual(out, in, first_time != 0);
    std::copy(out.begin(), out.end(), dst);
  }

  // coefficient_solutions[level] of the last Newton step, converted to double
  void
  orc_mg_ms_level_solution(void *h, int level, double *out)
  {
    MultigridSolver *m = static_cast<MultigridSolver *>(h);
    for (size_t i = 0; i < m->coefficient_solutions[level].size(); ++i)
      out[i] = m->coefficient_solutions[level][i];
  }

  int
  orc_mg_n_levels(void *h)
  {
    return static_cast<MultigridSolver *>(h)->nlevels;
  }

  long long
  orc_mg_n_dofs(void *h, int level)
  {
    return static_cast<MultigridSolver *>(h)->matrix[level].ndofs;
  }

  void
  orc_mg_dof_grid(void *h, int level, int *nd)
  {
    for (int d = 0; d < 3; ++d)
      nd[d] = static_cast<MultigridSolver *>(h)->matrix[level].nd[d];
  }

  double
  orc_mg_solve(void *h, int do_analyze, double *trace)
  {
    return static_cast<MultigridSolver *>(h)->solve(do_analyze != 0, trace);
  }

  double
  orc_mg_solve_cg(void *h, unsigned int *its, double *history, int history_len)
  {
    vector<double> hist;
    auto           r = static_cast<MultigridSolver *>(h)->solve_cg(&hist);
    *its             = r.first;
    for (int i = 0; i < history_len; ++i)
      history[i] = i < (int)hist.size() ? hist[i] : 0.;
    return r.second;
  }

  double
  orc_mg_l2_error(void *h, int level)
  {
    return static_cast<MultigridSolver *>(h)->compute_l2_error(level);
  }

  // precision: 0 = float operator (matrix), 1 = double operator (matrix_dp);
  // vectors cross the interface as double
  void
  orc_mg_vmult(void *h, int level, int precision, const double *src, double *dst)
  {
    MultigridSolver *s = static_cast<MultigridSolver *>(h);
    const size_t     N = s->matrix[level].ndofs;
    if (precision == 1)
      s->matrix_dp[level].vmult(dst, src);
    else
      {
        vector<float> a(src, src + N), b(N);
        s->matrix[level].vmult(b.data(), a.data());
        for (size_t i = 0; i < N; ++i)
          dst[i] = b[i];
      }
  }

  void
  orc_mg_vmult_residual(void *h, int level, int precision, const double *rhs, const double *lhs, double *res)
  {
    MultigridSolver *s = static_cast<MultigridSolver *>(h);
    const size_t     N = s->matrix[level].ndofs;
    if (precision == 1)
      s->matrix_dp[level].vmult_residual(rhs, lhs, res);
    else
      {
        vector<float> a(rhs, rhs + N), b(lhs, lhs + N), c(N);
        s->matrix[level].vmult_residual(a.data(), b.data(), c.data());
        for (size_t i = 0; i < N; ++i)
          res[i] = c[i];
      }
  }

  // fused CG update (double operator): in/out arrays r,q,p,x; sums[4]
  void
  orc_mg_vmult_with_cg_update(void *h, int level, double alpha, double beta, const double *r, double *q, double *p, double *x, double *sums)
  {
    MultigridSolver *s = static_cast<MultigridSolver *>(h);
    auto             a = s->matrix_dp[level].vmult_with_cg_update(alpha, beta, r, q, p, x);
    for (int i = 0; i < 4; ++i)
      sums[i] = a[i];
  }

  // which: 0 solution (with boundary values), 1 rhs, 2 inverse diagonal (float op),
  // 3 solution_update (float), 4 residual, 5 defect (float), 6 Chebyshev initial guess
  void
  orc_mg_get_vector(void *h, int which, int level, double *out)
  {
    MultigridSolver *s = static_cast<MultigridSolver *>(h);
    const size_t     N = s->matrix[level].ndofs;
    if (which == 0)
      {
        for (auto &i : s->inhomogeneous_bc[level])
          s->solution[level][i.first] = i.second;
        std::copy(s->solution[level].begin(), s->solution[level].end(), out);
      }
    else if (which == 1)
      std::copy(s->rhs[level].begin(), s->rhs[level].end(), out);
    else if (which == 2)
      for (size_t i = 0; i < N; ++i)
        out[i] = s->matrix[level].inv_diag[i];
    else if (which == 3)
      for (size_t i = 0; i < N; ++i)
        out[i] = s->solution_update[level][i];
    else if (which == 4)
      std::copy(s->residual[level].begin(), s->residual[level].end(), out);
    else if (which == 5)
      for (size_t i = 0; i < N; ++i)
        out[i] = s->defect[level][i];
    else if (which == 6)
      {
        vector<float> v;
        s->smooth[level].initial_guess(v);
        for (size_t i = 0; i < N; ++i)
          out[i] = v[i];
      }
  }

  // double-precision inverse diagonal computed afresh on matrix_dp (for parity tests)
  void
  orc_mg_inverse_diagonal_dp(void *h, int level, double *out)
  {
    MultigridSolver *s = static_cast<MultigridSolver *>(h);
    s->matrix_dp[level].compute_diagonal();
    std::copy(s->matrix_dp[level].inv_diag.begin(), s->matrix_dp[level].inv_diag.end(), out);
  }

  // info[0..5] = min_eig, max_eig (raw Lanczos), theta, delta, degree, cg iterations
  void
  orc_mg_chebyshev_info(void *h, int level, double *info)
  {
    const Chebyshev<float> &c = static_cast<MultigridSolver *>(h)->smooth[level];
    info[0]                   = c.min_eig;
    info[1]                   = c.max_eig;
    info[2]                   = c.theta;
    info[3]                   = c.delta;
    info[4]                   = c.degree;
    info[5]                   = c.cg_iterations;
  }

  // re-initialise the smoother of one level (multigrid_solver_dg.h:271-291 uses other degrees / ranges than
  // multigrid_solver.h:268-289): degree < 0 selects the automatic degree of the coarse solver
  void
  orc_mg_smoother_reinit(void *h, int level, double smoothing_range, int degree, int eig_cg_n_iterations)
  {
    MultigridSolver  *s = static_cast<MultigridSolver *>(h);
    Chebyshev<float> &c = s->smooth[level];
    c.smoothing_range     = smoothing_range;
    c.degree              = degree < 0 ? Chebyshev<float>::invalid_degree : (unsigned int)degree;
    c.eig_cg_n_iterations = eig_cg_n_iterations < 0 ? (unsigned int)s->matrix[level].m() : (unsigned int)eig_cg_n_iterations;
    c.initialize(s->matrix[level]);
  }

  // mode 0: x = vmult(b) (zero initial guess), mode 1: step(x, b)
  void
  orc_mg_chebyshev_apply(void *h, int level, int mode, double *x, const double *b)
  {
    MultigridSolver *s = static_cast<MultigridSolver *>(h);
    const size_t     N = s->matrix[level].ndofs;
    vector<float>    xf(x, x + N), bf(b, b + N);
    if (mode == 0)
      s->smooth[level].vmult(xf, bf);
    else
      s->smooth[level].step(xf, bf);
    for (size_t i = 0; i < N; ++i)
      x[i] = xf[i];
  }

  // mode 0: fine = P coarse (float, constrained); 1: fine += P coarse (float,
  // constrained); 2: coarse += P^T fine (float, constrained); 3: fine = P coarse
  // (double, no boundary constraints).  level = fine level.
  void
  orc_mg_transfer(void *h, int level, int mode, double *dst, const double *src)
  {
    MultigridSolver *s  = static_cast<MultigridSolver *>(h);
    const size_t     Nf = s->matrix[level].ndofs, Nc = s->matrix[level - 1].ndofs;
    if (mode == 3)
      {
        vector<double> f(Nf), c(src, src + Nc);
        s->mg_transfer_no_boundary[level].prolongate(f, c, false);
        std::copy(f.begin(), f.end(), dst);
      }
    else if (mode == 2)
      {
        vector<float> c(dst, dst + Nc), f(src, src + Nf);
        s->transfer[level].restrict_and_add(c, f);
        for (size_t i = 0; i < Nc; ++i)
          dst[i] = c[i];
      }
    else
      {
        vector<float> f(dst, dst + Nf), c(src, src + Nc);
        s->transfer[level].prolongate(f, c, mode == 1);
        for (size_t i = 0; i < Nf; ++i)
          dst[i] = f[i];
      }
  }

  // preconditioner application (one V-cycle in float): dst = M^-1 src on the finest level
  void
  orc_mg_vcycle(void *h, double *dst, const double *src)
  {
    MultigridSolver *s = static_cast<MultigridSolver *>(h);
    const size_t     N = s->matrix.back().ndofs;
    vector<double>   a(src, src + N), b(N);
    s->vmult(b, a);
    std::copy(b.begin(), b.end(), dst);
  }

  // seconds per apply, protocol of poisson_cube/program.cc:343-380 (one repeat)
  double
  orc_mg_time_matvec(void *h, int precision, int n_mv)
  {
    MultigridSolver *s = static_cast<MultigridSolver *>(h);
    const int        L = s->nlevels - 1;
    auto             t0 = std::chrono::steady_clock::now();
    for (int i = 0; i < n_mv; ++i)
      if (precision == 1)
        s->matrix_dp[L].vmult(s->residual[L].data(), s->solution[L].data());
      else
        s->matrix[L].vmult(s->solution_update[L].data(), s->defect[L].data());
    return std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count() / n_mv;
  }

  // threads the OpenMP runtime really uses (bench.py reports this, not the value it asked for)
  int
  orc_omp_max_threads()
  {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
  }

  void
  orc_omp_set_threads(int n)
  {
#ifdef _OPENMP
    if (n > 0)
      omp_set_num_threads(n);
#else
    (void)n;
#endif
  }

  void
  orc_mg_setup_times(void *h, double *out)
  {
    for (int i = 0; i < 3; ++i)
      out[i] = static_cast<MultigridSolver *>(h)->timings_setup[i];
  }

  // 1-D ingredients for cross-checks: S, G, D (n*n each), P1 ((2p+1)*n), gll, xq, wq (n each)
  void
  orc_basis(int p, double *S, double *G, double *D, double *P1, double *gll, double *xq, double *wq)
  {
    const Basis1D b = make_basis(p);
    const int     n = p + 1;
    std::copy(b.S.begin(), b.S.end(), S);
    std::copy(b.G.begin(), b.G.end(), G);
    std::copy(b.D.begin(), b.D.end(), D);
    std::copy(b.P1.begin(), b.P1.end(), P1);
    std::copy(b.gll.begin(), b.gll.end(), gll);
    std::copy(b.xq.begin(), b.xq.end(), xq);
    std::copy(b.wq.begin(), b.wq.end(), wq);
    (void)n;
  }
}
