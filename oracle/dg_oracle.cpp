// ORACLE (test infrastructure, not product code): symmetric interior-penalty DG Laplacian, FE_DGQ(p)
// (Lagrange basis in the Gauss-Lobatto points) with Gauss quadrature on p+1 points, on an affine
// (sheared) box mesh.  CPU restatement of the reference's in-repo DG specification
//   MFReference::LaplaceOperatorFaceBased          common/laplace_operator_dg_face.h:13-166
//     local_apply                (cell term  int grad v . grad u)                      :58-77
//     local_apply_face           (interior faces, sigma = (p+1)^2 * mean |n . J^-T|_n)  :79-129
//     local_apply_boundary_face  (mirror u+ = -u-, sigma_b = 2 sigma-)                  :131-160
// against which the reference itself verifies its fast operator LaplaceOperatorCompactCombine
// (matvec_dg/program.cc:206-207), on the mesh of matvec_dg/program.cc:55-77 (box, vertices mapped by
// the constant matrix I + 0.12 (d+1)(e+1)).  The loops are organised face by face like the reference's
// face-based operator (the CUDA kernel is organised cell by cell), plain quadrature sums without sum
// factorisation.  The reference has no golden vectors for this operator: PARITY UNPINNED against the
// reference itself; the oracle is pinned by properties instead (tests/test_dg_oracle.py: symmetry,
// a(u,v) = -int Laplace(u) v for global polynomials, comparison with a dense numpy assembly).
// DoF numbering: cell-wise contiguous (dof_indices_contiguous, laplace_operator_dg.h:1141-1142), cells
// lexicographic with x fastest, inside a cell (k n + j) n + i.
#include <cmath>
#include <cstring>
#include <vector>

#include "basis1d.h"

namespace
{
  struct DG
  {
    int    p, n, nc[3];
    double Jinv[3][3]; // d xi_d / d x_k
    double detJ;
    double gnorm[3];   // |grad xi_d|
    orc::Basis1D b;
    std::vector<double> S, G;   // [q][i] values / derivatives at the Gauss points
    std::vector<double> e0, e1; // phi_i'(0), phi_i'(1)
    std::vector<double> f0, f1; // phi_i(0), phi_i(1)
    long   ncells, ndofs;
  };

  void
  invert3(const double A[3][3], double B[3][3], double &det)
  {
    det = A[0][0] * (A[1][1] * A[2][2] - A[1][2] * A[2][1]) - A[0][1] * (A[1][0] * A[2][2] - A[1][2] * A[2][0]) +
          A[0][2] * (A[1][0] * A[2][1] - A[1][1] * A[2][0]);
    for (int i = 0; i < 3; ++i)
      for (int j = 0; j < 3; ++j)
        {
          const int i1 = (i + 1) % 3, i2 = (i + 2) % 3, j1 = (j + 1) % 3, j2 = (j + 2) % 3;
          B[j][i]      = (A[i1][j1] * A[i2][j2] - A[i1][j2] * A[i2][j1]) / det;
        }
  }

  // values and physical gradient of the cell function with nodal values u at reference point (x0,x1,x2)
  void
  eval_point(const DG &g, const double *u, const double *ph[3], const double *dph[3], double &val, double grad[3])
  {
    const int n = g.n;
    double    rg[3] = {0, 0, 0};
    val             = 0;
    for (int k = 0; k < n; ++k)
      for (int j = 0; j < n; ++j)
        for (int i = 0; i < n; ++i)
          {
            const double c = u[(k * n + j) * n + i];
            val += c * ph[0][i] * ph[1][j] * ph[2][k];
            rg[0] += c * dph[0][i] * ph[1][j] * ph[2][k];
            rg[1] += c * ph[0][i] * dph[1][j] * ph[2][k];
            rg[2] += c * ph[0][i] * ph[1][j] * dph[2][k];
          }
    for (int k = 0; k < 3; ++k)
      grad[k] = g.Jinv[0][k] * rg[0] + g.Jinv[1][k] * rg[1] + g.Jinv[2][k] * rg[2];
  }

  // dst_i += tv phi_i + tg . grad phi_i  (test with value and physical gradient at one point)
  void
  test_point(const DG &g, double *dst, const double *ph[3], const double *dph[3], const double tv, const double tg[3])
  {
    const int n = g.n;
    double    r[3];
    for (int d = 0; d < 3; ++d)
      r[d] = g.Jinv[d][0] * tg[0] + g.Jinv[d][1] * tg[1] + g.Jinv[d][2] * tg[2];
    for (int k = 0; k < n; ++k)
      for (int j = 0; j < n; ++j)
        for (int i = 0; i < n; ++i)
          dst[(k * n + j) * n + i] += tv * ph[0][i] * ph[1][j] * ph[2][k] + r[0] * dph[0][i] * ph[1][j] * ph[2][k] +
                                      r[1] * ph[0][i] * dph[1][j] * ph[2][k] + r[2] * ph[0][i] * ph[1][j] * dph[2][k];
  }
} // namespace

extern "C"
{
  // trafo: row-major 3x3 matrix applied to the vertices of the box [lo,hi] with ncell cells (matvec_dg/program.cc:66-76)
  void *
  dgo_create(int p, const int *ncell, const double *lo, const double *hi, const double *trafo)
  {
    if (p < 1 || p > 8)
      return nullptr;
    DG *g = new DG();
    g->p  = p;
    g->n  = p + 1;
    g->b  = orc::make_basis(p);
    double J[3][3];
    for (int d = 0; d < 3; ++d)
      {
        g->nc[d] = ncell[d];
        for (int k = 0; k < 3; ++k)
          J[k][d] = trafo[k * 3 + d] * (hi[d] - lo[d]) / ncell[d]; // d x_k / d xi_d
      }
    invert3(J, g->Jinv, g->detJ);
    g->detJ = std::fabs(g->detJ);
    for (int d = 0; d < 3; ++d)
      g->gnorm[d] = std::sqrt(g->Jinv[d][0] * g->Jinv[d][0] + g->Jinv[d][1] * g->Jinv[d][1] + g->Jinv[d][2] * g->Jinv[d][2]);
    g->S = g->b.S;
    g->G = g->b.G;
    std::vector<long double> nodes(g->b.gll.begin(), g->b.gll.end());
    g->e0.resize(g->n);
    g->e1.resize(g->n);
    g->f0.assign(g->n, 0.);
    g->f1.assign(g->n, 0.);
    g->f0[0]        = 1;
    g->f1[g->n - 1] = 1;
    for (int i = 0; i < g->n; ++i)
      {
        g->e0[i] = (double)orc::lagrange_deriv(nodes, i, 0.0L);
        g->e1[i] = (double)orc::lagrange_deriv(nodes, i, 1.0L);
      }
    g->ncells = (long)ncell[0] * ncell[1] * ncell[2];
    g->ndofs  = g->ncells * g->n * g->n * g->n;
    return g;
  }

  // Another 1-D basis of the same polynomial space (FE_DGQHermite, FE_DGQArbitraryNodes(QGauss): template parameter
  // `type` of LaplaceOperatorCompactCombine, laplace_operator_dg.h:346-350; matvec_dg/program.cc:44-49): the tables
  // S[q n + i] = phi_i(x_q), G = phi_i'(x_q), f0/f1 = phi_i(0/1), e0/e1 = phi_i'(0/1) come from oracle/orc.py
  void
  dgo_set_basis(void *h, const double *S, const double *G, const double *f0, const double *f1, const double *e0,
                const double *e1)
  {
    DG       *g = static_cast<DG *>(h);
    const int n = g->n;
    g->S.assign(S, S + n * n);
    g->G.assign(G, G + n * n);
    g->f0.assign(f0, f0 + n);
    g->f1.assign(f1, f1 + n);
    g->e0.assign(e0, e0 + n);
    g->e1.assign(e1, e1 + n);
  }

  void dgo_destroy(void *h) { delete static_cast<DG *>(h); }
  long long dgo_n_dofs(void *h) { return static_cast<DG *>(h)->ndofs; }

  // (p+1)^2 |grad xi_d|: the penalty of a face normal to direction d (get_penalty, laplace_operator_dg.h:789-793)
  void
  dgo_penalty(void *h, double *sigma)
  {
    const DG *g = static_cast<DG *>(h);
    for (int d = 0; d < 3; ++d)
      sigma[d] = (g->p + 1.0) * (g->p + 1.0) * g->gnorm[d];
  }

  void
  dgo_vmult(void *h, const double *src, double *dst)
  {
    const DG &g  = *static_cast<DG *>(h);
    const int n  = g.n, n3 = n * n * n;
    std::memset(dst, 0, sizeof(double) * g.ndofs);
    const std::vector<double> &v0 = g.f0, &v1 = g.f1;
    // ---- cells (laplace_operator_dg_face.h:58-77)
#pragma omp parallel for
    for (long c = 0; c < g.ncells; ++c)
      for (int q2 = 0; q2 < n; ++q2)
        for (int q1 = 0; q1 < n; ++q1)
          for (int q0 = 0; q0 < n; ++q0)
            {
              const double *ph[3]  = {&g.S[q0 * n], &g.S[q1 * n], &g.S[q2 * n]};
              const double *dph[3] = {&g.G[q0 * n], &g.G[q1 * n], &g.G[q2 * n]};
              double        val, grad[3];
              eval_point(g, src + c * n3, ph, dph, val, grad);
              const double JxW = g.detJ * g.b.wq[q0] * g.b.wq[q1] * g.b.wq[q2];
              for (int k = 0; k < 3; ++k)
                grad[k] *= JxW;
              test_point(g, dst + c * n3, ph, dph, 0., grad);
            }
    // ---- faces normal to direction d, visited from the cell on the lower side ("interior" side, normal n =
    //      grad xi_d / |grad xi_d|); the face on the low boundary is visited from its only cell with normal -n
    for (int d = 0; d < 3; ++d)
      {
        const int    a = (d + 1) % 3, bdir = (d + 2) % 3;
        const double sigma = (g.p + 1.0) * (g.p + 1.0) * g.gnorm[d]; // both sides equal on this mesh
        double       nrm[3];
        for (int k = 0; k < 3; ++k)
          nrm[k] = g.Jinv[d][k] / g.gnorm[d];
        // each visit writes to cell c and to its upper neighbour: two passes by the parity of the cell index
        // along d keep the threads apart
        for (int parity = 0; parity < 2; ++parity)
          {
#pragma omp parallel for
            for (long c = 0; c < g.ncells; ++c)
              {
                int cc[3] = {(int)(c % g.nc[0]), (int)((c / g.nc[0]) % g.nc[1]), (int)(c / ((long)g.nc[0] * g.nc[1]))};
                if ((cc[d] & 1) != parity)
                  continue;
                const long stride[3] = {1, g.nc[0], (long)g.nc[0] * g.nc[1]};
                for (int qb = 0; qb < n; ++qb)
                  for (int qa = 0; qa < n; ++qa)
                    {
                      const double  JxW = g.detJ * g.gnorm[d] * g.b.wq[qa] * g.b.wq[qb];
                      const double *ph_m[3], *dph_m[3], *ph_p[3], *dph_p[3];
                      ph_m[a] = ph_p[a] = &g.S[qa * n];
                      dph_m[a] = dph_p[a] = &g.G[qa * n];
                      ph_m[bdir] = ph_p[bdir] = &g.S[qb * n];
                      dph_m[bdir] = dph_p[bdir] = &g.G[qb * n];
                      // upper face of cell c: minus side at xi_d = 1, plus side (neighbour) at xi_d = 0
                      ph_m[d]  = v1.data();
                      dph_m[d] = g.e1.data();
                      ph_p[d]  = v0.data();
                      dph_p[d] = g.e0.data();
                      double um, gm[3];
                      eval_point(g, src + c * n3, ph_m, dph_m, um, gm);
                      const double dn_m = gm[0] * nrm[0] + gm[1] * nrm[1] + gm[2] * nrm[2];
                      if (cc[d] + 1 < g.nc[d])
                        {
                          // interior face (laplace_operator_dg_face.h:79-129)
                          const long cn = c + stride[d];
                          double     up, gp[3];
                          eval_point(g, src + cn * n3, ph_p, dph_p, up, gp);
                          const double dn_p    = gp[0] * nrm[0] + gp[1] * nrm[1] + gp[2] * nrm[2];
                          const double avg     = 0.5 * (um - up);
                          const double valgrad = avg * sigma * 2.0 - 0.5 * (dn_m + dn_p);
                          double       tg[3];
                          for (int k = 0; k < 3; ++k)
                            tg[k] = -avg * nrm[k] * JxW; // submit_normal_derivative(-average_value) on both sides
                          test_point(g, dst + c * n3, ph_m, dph_m, valgrad * JxW, tg);
                          test_point(g, dst + cn * n3, ph_p, dph_p, -valgrad * JxW, tg);
                        }
                      else
                        {
                          // boundary face with outward normal +n (:131-160)
                          const double valgrad = -dn_m + um * sigma * 2.0;
                          double       tg[3];
                          for (int k = 0; k < 3; ++k)
                            tg[k] = -um * nrm[k] * JxW;
                          test_point(g, dst + c * n3, ph_m, dph_m, valgrad * JxW, tg);
                        }
                      if (cc[d] == 0)
                        {
                          // boundary face on the low side, outward normal -n, own trace at xi_d = 0
                          double u0, g0[3];
                          eval_point(g, src + c * n3, ph_p, dph_p, u0, g0);
                          const double dn      = -(g0[0] * nrm[0] + g0[1] * nrm[1] + g0[2] * nrm[2]);
                          const double valgrad = -dn + u0 * sigma * 2.0;
                          double       tg[3];
                          for (int k = 0; k < 3; ++k)
                            tg[k] = u0 * nrm[k] * JxW; // -u * (-n)
                          test_point(g, dst + c * n3, ph_p, dph_p, valgrad * JxW, tg);
                        }
                    }
              }
          }
      }
  }
}
