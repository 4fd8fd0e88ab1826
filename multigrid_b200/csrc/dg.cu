// Host side and C ABI of the DG operator (include/mgb200.h, section "Discontinuous Galerkin").
//   DGOperator  <->  multigrid::LaplaceOperatorCompactCombine<3,p,Number,1>   common/laplace_operator_dg.h:360-2026
// There is no CPU fallback: every numerical step is a kernel launch on the operator's stream.
#include <cuda_runtime.h>

#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <functional>
#include <memory>
#include <stdexcept>
#include <string>
#include <vector>

#include "../../include/mgb200.h"
#include "basis.h"
#include "dg_kernels.cuh"
#include "level.h"

#include "functions.cuh"

struct mgb_solver_t;
const mgb::LevelDesc *mgb_internal_level_desc(mgb_solver_t *s, int level);
cudaStream_t          mgb_internal_stream(mgb_solver_t *s);
struct mgb_comm_t;
mgb_comm_t *mgb_internal_comm_create(int rank, int size, const unsigned char id[128], int device);
void        mgb_internal_comm_destroy(mgb_comm_t *c);
void        mgb_internal_comm_exchange(mgb_comm_t *c, cudaStream_t stream, const void *send_up, void *recv_dn, const void *send_dn,
                                       void *recv_up, size_t count, int is_double);
void        mgb_internal_comm_allreduce(mgb_comm_t *c, cudaStream_t stream, double *dev, int count);
mgb_comm_t *mgb_internal_comm_borrow(mgb_solver_t *s);

namespace mgb
{
#define MGB_DG_CHECK(call)                                                                      \
  do                                                                                            \
    {                                                                                           \
      cudaError_t err__ = (call);                                                               \
      if (err__ != cudaSuccess)                                                                 \
        throw std::runtime_error(std::string("CUDA error: ") + cudaGetErrorString(err__) +     \
                                 " at " + __FILE__ + ":" + std::to_string(__LINE__));           \
    }                                                                                           \
  while (0)

  // LocalBasisTransformer<dim, 0, degree, Number> (laplace_operator_dg.h:176-222): generalised eigenvectors of the
  // 1-D SIP Laplacian (cell term + both end faces, penalty N^2) with respect to the 1-D mass matrix, in the nodal
  // basis.  The reference calls LAPACK (compute_generalized_eigenvalues_symmetric); here: Cholesky of the mass
  // matrix + cyclic Jacobi rotations.  V[k n + j] = component j of eigenvector k, eigenvalues ascending.  Sign
  // and scaling of the vectors do not matter for V D^-1 V^T with D = diag(V^T A V).
  inline std::vector<double>
  jacobi_transformed_eigenvectors(const DGBasis &b)
  {
    const int                n = b.n;
    std::vector<double>      M(n * n), A(n * n);
    const double             penalty = (double)n * n; // penalty_factor = 1 (laplace_operator_dg.h:47,198)
    for (int i = 0; i < n; ++i)
      for (int j = 0; j < n; ++j)
        {
          double m = 0, l = 0;
          for (int q = 0; q < n; ++q)
            {
              m += b.S[q * n + i] * b.S[q * n + j] * b.wq[q];
              l += b.G[q * n + i] * b.G[q * n + j] * b.wq[q];
            }
          const double vi0 = b.f0[i], vj0 = b.f0[j], vi1 = b.f1[i], vj1 = b.f1[j];
          const double gi0 = b.e0[i], gj0 = b.e0[j], gi1 = b.e1[i], gj1 = b.e1[j];
          l += vi0 * vj0 * penalty + 0.5 * gi0 * vj0 + 0.5 * gj0 * vi0;
          l += vi1 * vj1 * penalty - 0.5 * gi1 * vj1 - 0.5 * gj1 * vi1;
          M[i * n + j] = m;
          A[i * n + j] = l;
        }
    // M = L L^T
    std::vector<double> L(n * n, 0.);
    for (int i = 0; i < n; ++i)
      for (int j = 0; j <= i; ++j)
        {
          double s = M[i * n + j];
          for (int k = 0; k < j; ++k)
            s -= L[i * n + k] * L[j * n + k];
          L[i * n + j] = i == j ? std::sqrt(s) : s / L[j * n + j];
        }
    // C = L^-1 A L^-T
    std::vector<double> X(n * n), C(n * n);
    for (int c = 0; c < n; ++c) // X = L^-1 A (column by column, forward substitution)
      for (int i = 0; i < n; ++i)
        {
          double s = A[i * n + c];
          for (int k = 0; k < i; ++k)
            s -= L[i * n + k] * X[k * n + c];
          X[i * n + c] = s / L[i * n + i];
        }
    for (int r = 0; r < n; ++r) // C = X L^-T: row r of C solves L C_r^T = X_r^T
      for (int i = 0; i < n; ++i)
        {
          double s = X[r * n + i];
          for (int k = 0; k < i; ++k)
            s -= L[i * n + k] * C[r * n + k];
          C[r * n + i] = s / L[i * n + i];
        }
    for (int i = 0; i < n; ++i)
      for (int j = 0; j < i; ++j)
        C[i * n + j] = C[j * n + i] = 0.5 * (C[i * n + j] + C[j * n + i]);
    // cyclic Jacobi: C W = W diag
    std::vector<double> W(n * n, 0.);
    for (int i = 0; i < n; ++i)
      W[i * n + i] = 1;
    for (int sweep = 0; sweep < 100; ++sweep)
      {
        double off = 0;
        for (int i = 0; i < n; ++i)
          for (int j = 0; j < i; ++j)
            off += C[i * n + j] * C[i * n + j];
        if (off < 1e-60)
          break;
        for (int pp = 0; pp < n; ++pp)
          for (int q = pp + 1; q < n; ++q)
            {
              if (C[pp * n + q] == 0.)
                continue;
              const double theta = (C[q * n + q] - C[pp * n + pp]) / (2. * C[pp * n + q]);
              const double t  = (theta >= 0 ? 1. : -1.) / (std::fabs(theta) + std::sqrt(theta * theta + 1.));
              const double cs = 1. / std::sqrt(t * t + 1.), sn = t * cs;
              for (int k = 0; k < n; ++k)
                {
                  const double ckp = C[k * n + pp], ckq = C[k * n + q];
                  C[k * n + pp]    = cs * ckp - sn * ckq;
                  C[k * n + q]     = sn * ckp + cs * ckq;
                }
              for (int k = 0; k < n; ++k)
                {
                  const double cpk = C[pp * n + k], cqk = C[q * n + k];
                  C[pp * n + k]    = cs * cpk - sn * cqk;
                  C[q * n + k]     = sn * cpk + cs * cqk;
                }
              for (int k = 0; k < n; ++k)
                {
                  const double wkp = W[k * n + pp], wkq = W[k * n + q];
                  W[k * n + pp]    = cs * wkp - sn * wkq;
                  W[k * n + q]     = sn * wkp + cs * wkq;
                }
            }
      }
    // eigenvectors of the pencil: v = L^-T w, sorted by eigenvalue
    std::vector<int> order(n);
    for (int i = 0; i < n; ++i)
      order[i] = i;
    std::sort(order.begin(), order.end(), [&](int x, int y) { return C[x * n + x] < C[y * n + y]; });
    std::vector<double> V(n * n);
    for (int k = 0; k < n; ++k)
      {
        const int col = order[k];
        for (int i = n - 1; i >= 0; --i) // back substitution with L^T
          {
            double s = W[i * n + col];
            for (int j = i + 1; j < n; ++j)
              s -= L[j * n + i] * V[k * n + j];
            V[k * n + i] = s / L[i * n + i];
          }
      }
    return V;
  }

  template <typename T, int n>
  static DGTables<T, n>
  make_tables(const DGBasis &b)
  {
    DGTables<T, n>           t;
    for (int i = 0; i < n * n; ++i)
      {
        t.S[i] = (T)b.S[i];
        t.G[i] = (T)b.G[i];
        t.D[i] = (T)b.D[i];
      }
    for (int i = 0; i < n; ++i)
      {
        t.e0[i] = (T)b.e0[i];
        t.e1[i] = (T)b.e1[i];
        t.f0[i] = (T)b.f0[i];
        t.f1[i] = (T)b.f1[i];
        t.w[i]  = (T)b.wq[i];
      }
    // traces of the collocation basis (Lagrange polynomials in the Gauss points): what FE_DGQArbitraryNodes(QGauss) has
    const DGBasis g = b.kind == DG_BASIS_GAUSS ? b : make_dg_basis(b.p, DG_BASIS_GAUSS);
    for (int i = 0; i < n; ++i)
      {
        t.qf0[i] = (T)g.f0[i];
        t.qf1[i] = (T)g.f1[i];
        t.qe0[i] = (T)g.e0[i];
        t.qe1[i] = (T)g.e1[i];
      }
    const std::vector<double> V = jacobi_transformed_eigenvectors(b);
    for (int i = 0; i < n * n; ++i)
      t.V[i] = (T)V[i];
    return t;
  }

  // diagonal entries of the 1-D matrices int phi phi, int phi' phi', int phi' phi and the end derivatives
  struct DGDiagTables
  {
    double M[9], K[9], C[9], e0[9], e1[9], f0[9], f1[9];
  };

  struct DGGeom
  {
    double gt[6], gs[3][3], sigma[3], area[3];
  };

  // a(phi_i, phi_i): cell term + the faces on which phi_i does not vanish (see dg_kernels.cuh header)
  template <typename T>
  __global__ void
  k_dg_inverse_diagonal(const int n, const int nc0, const int nc1, const int nc2, const int local_layers, const int cz_shift,
                        const DGGeom g, const DGDiagTables tb, T *__restrict__ out)
  {
    const size_t n3    = (size_t)n * n * n;
    const size_t total = (size_t)nc0 * nc1 * local_layers * n3;
    for (size_t idx = blockIdx.x * (size_t)blockDim.x + threadIdx.x; idx < total; idx += (size_t)gridDim.x * blockDim.x)
      {
        const size_t c = idx / n3;
        const int    l = (int)(idx - c * n3);
        const int    i[3]  = {l % n, (l / n) % n, l / (n * n)};
        const int    cc[3] = {(int)(c % nc0), (int)((c / nc0) % nc1), (int)(c / ((size_t)nc0 * nc1)) + cz_shift};
        const int    ncs[3] = {nc0, nc1, nc2};
        const double M0 = tb.M[i[0]], M1 = tb.M[i[1]], M2 = tb.M[i[2]];
        const double C0 = tb.C[i[0]], C1 = tb.C[i[1]], C2 = tb.C[i[2]];
        double       v = g.gt[0] * tb.K[i[0]] * M1 * M2 + g.gt[1] * M0 * tb.K[i[1]] * M2 + g.gt[2] * M0 * M1 * tb.K[i[2]] +
                   2. * (g.gt[3] * C0 * C1 * M2 + g.gt[4] * C0 * M1 * C2 + g.gt[5] * M0 * C1 * C2);
        for (int d = 0; d < 3; ++d)
          {
            const int    da = d == 0 ? 1 : 0, db = d == 2 ? 1 : 2;
            const double Ma = tb.M[i[da]], Mb = tb.M[i[db]], Ca = tb.C[i[da]], Cb = tb.C[i[db]];
            for (int s = 0; s < 2; ++s)
              {
                const double f = s ? tb.f1[i[d]] : tb.f0[i[d]]; // trace of the 1-D function: one node of the Gauss-Lobatto basis
                if (f == 0.)
                  continue;
                const bool   interior = s ? cc[d] + 1 < ncs[d] : cc[d] > 0;
                const double sgn = s ? 1. : -1., e = s ? tb.e1[i[d]] : tb.e0[i[d]];
                const double term = g.sigma[d] * f * f * Ma * Mb -
                                    sgn * f * (g.gs[d][d] * e * Ma * Mb + f * (g.gs[d][da] * Ca * Mb + g.gs[d][db] * Ma * Cb));
                v += (interior ? 1. : 2.) * g.area[d] * term;
              }
          }
        out[idx] = (T)(1. / v);
      }
  }

  template <typename F>
  __global__ void
  k_dg_foreach(const size_t n, F f)
  {
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
      f(i);
  }

  // sum_i f(i, acc) for NS sums (warp shuffles, one atomic per block and sum)
  template <int NS, typename F>
  __global__ void
  k_dg_reduce(const size_t off, const size_t n, F f, double *out)
  {
    double acc[NS];
    for (int s = 0; s < NS; ++s)
      acc[s] = 0;
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
      f(off + i, acc);
    __shared__ double red[NS][8];
    for (int s = 0; s < NS; ++s)
      {
        double v = acc[s];
        for (int o = 16; o > 0; o >>= 1)
          v += __shfl_down_sync(0xffffffffu, v, o);
        if ((threadIdx.x & 31) == 0)
          red[s][threadIdx.x >> 5] = v;
      }
    __syncthreads();
    if (threadIdx.x < NS)
      {
        double v = 0;
        for (int w = 0; w < (int)(blockDim.x >> 5); ++w)
          v += red[threadIdx.x][w];
        atomicAdd(&out[threadIdx.x], v);
      }
  }

  // four sums of vmult_with_cg_update: q.p, r.r, q.r, q.q (warp shuffles, one atomic per block and sum)
  // out[0..4] += column sums of partial[n_blocks][5] (the per-block sums of k_dg_apply in mode DG_SET_SUMS); out zeroed
  // by the caller, one atomicAdd per block and column
  __global__ void
  k_dg_sum_partials(const size_t n_blocks, const double *__restrict__ partial, double *out)
  {
    double       acc[5] = {0, 0, 0, 0, 0};
    const size_t total = n_blocks * 5, stride = (size_t)gridDim.x * blockDim.x * 5;
    for (size_t e = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) * 5; e < total; e += stride)
      for (int s = 0; s < 5; ++s)
        acc[s] += partial[e + s];
    __shared__ double red[5][32];
    for (int s = 0; s < 5; ++s)
      {
        double v = acc[s];
        for (int o = 16; o > 0; o >>= 1)
          v += __shfl_down_sync(0xffffffffu, v, o);
        if ((threadIdx.x & 31) == 0)
          red[s][threadIdx.x >> 5] = v;
      }
    __syncthreads();
    if (threadIdx.x < 5)
      {
        double v = 0;
        for (int w = 0; w < (int)(blockDim.x >> 5); ++w)
          v += red[threadIdx.x][w];
        atomicAdd(&out[threadIdx.x], v);
      }
  }

  template <typename T>
  __global__ void
  k_dg_cg_sums(const size_t off, const size_t n, const T *__restrict__ q, const T *__restrict__ p, const T *__restrict__ r,
               double *out)
  {
    double acc[4] = {0, 0, 0, 0};
    for (size_t j = blockIdx.x * (size_t)blockDim.x + threadIdx.x; j < n; j += (size_t)gridDim.x * blockDim.x)
      {
        const size_t i  = off + j;
        const double qi = q[i], pi = p[i], ri = r[i];
        acc[0] += qi * pi;
        acc[1] += ri * ri;
        acc[2] += qi * ri;
        acc[3] += qi * qi;
      }
    __shared__ double red[4][8];
    for (int s = 0; s < 4; ++s)
      {
        double v = acc[s];
        for (int o = 16; o > 0; o >>= 1)
          v += __shfl_down_sync(0xffffffffu, v, o);
        if ((threadIdx.x & 31) == 0)
          red[s][threadIdx.x >> 5] = v;
      }
    __syncthreads();
    if (threadIdx.x < 4)
      {
        double v = 0;
        for (int w = 0; w < (int)(blockDim.x >> 5); ++w)
          v += red[threadIdx.x][w];
        atomicAdd(&out[threadIdx.x], v);
      }
  }

  // ---- same-degree transfer between the DG space and FE_Q on the same mesh (laplace_operator_dg.h:1798-1819,
  //      1863-1894).  Both bases are Lagrange bases in the same Gauss-Lobatto points, so the 1-D change of basis of
  //      LocalBasisTransformer<dim,1,...> (:106-141) is the identity and the transfer is the cell-to-DoF map of FE_Q:
  //      prolongation gathers (constrained FE_Q entries read as zero, vector_access_reduced.h:174-179), restriction
  //      scatter-adds (constrained entries skipped, :431-433) -- written as a gather per FE_Q point over its up to
  //      8 cells, so it needs no atomics.
  template <typename T>
  __global__ void
  k_cg_to_dg_add(const LevelDesc L, const T *__restrict__ cg, T *__restrict__ dg)
  {
    const int    n = L.n, n3 = n * n * n;
    const size_t total = (size_t)L.n_cells * n3;
    for (size_t idx = blockIdx.x * (size_t)blockDim.x + threadIdx.x; idx < total; idx += (size_t)gridDim.x * blockDim.x)
      {
        const uint32_t c = idx / n3;
        const int      l = (int)(idx - (size_t)c * n3);
        bool           con;
        const uint32_t d = cell_dof(L, c, l % n, (l / n) % n, l / (n * n), con);
        if (!con)
          dg[idx] += cg[d];
      }
  }

  template <typename T>
  __global__ void
  k_dg_to_cg_add(const LevelDesc L, const T *__restrict__ dg, T *__restrict__ cg)
  {
    const int    p = L.p, n = L.n, n3 = n * n * n;
    const size_t total = L.n_dofs;
    for (size_t lex = blockIdx.x * (size_t)blockDim.x + threadIdx.x; lex < total; lex += (size_t)gridDim.x * blockDim.x)
      {
        const int    g0 = lex % L.nd[0];
        const size_t r  = lex / L.nd[0];
        const int    g1 = r % L.nd[1], g2 = r / L.nd[1];
        if (is_constrained(L, g0, g1, g2))
          continue;
        T sum = 0;
        for (int dz = 0; dz < 2; ++dz)
          {
            const int cz = g2 / p - dz, k = g2 - cz * p;
            if (cz < 0 || cz >= L.ncell[2] || k > p)
              continue;
            for (int dy = 0; dy < 2; ++dy)
              {
                const int cy = g1 / p - dy, j = g1 - cy * p;
                if (cy < 0 || cy >= L.ncell[1] || j > p)
                  continue;
                for (int dx = 0; dx < 2; ++dx)
                  {
                    const int cx = g0 / p - dx, i = g0 - cx * p;
                    if (cx < 0 || cx >= L.ncell[0] || i > p)
                      continue;
                    sum += dg[(size_t)cell_number(L, cx, cy, cz) * n3 + (k * n + j) * n + i];
                  }
              }
          }
        cg[dof_index(L, g0, g1, g2)] += sum;
      }
  }

  // Change of basis of a DG vector, cell by cell: out (+)= (M x M x M) in with the 1-D matrix M (TRANSPOSE: M^T);
  // LocalBasisTransformer<dim,1,...>::apply (laplace_operator_dg.h:106-141, 290-322) for the bases whose coefficients
  // are not the values in the Gauss-Lobatto points.  One cell per 128 threads, three sweeps in shared memory.
  struct DGMatrix1D
  {
    double m[81];
  };
  template <int n, typename T, bool TRANSPOSE, bool ADD>
  __global__ void
  __launch_bounds__(128) k_dg_basis_change(const size_t n_cells, const DGMatrix1D M, const T *in, T *out)
  {
    constexpr int n2 = n * n, n3 = n2 * n;
    __shared__ T  buf[2][n3];
    const size_t  c = blockIdx.x;
    if (c >= n_cells)
      return;
    for (int i = threadIdx.x; i < n3; i += 128)
      buf[0][i] = in[c * n3 + i];
    __syncthreads();
    // pencil-wise sweeps: a thread owns the n values of a line along the sweep direction
#pragma unroll
    for (int dir = 0; dir < 3; ++dir)
      {
        const T *a  = buf[dir & 1];
        T       *b  = buf[(dir + 1) & 1];
        const int st = dir == 0 ? 1 : (dir == 1 ? n : n2);
        for (int l = threadIdx.x; l < n2; l += 128)
          {
            // the two indices across the sweep direction
            const int u = l % n, v = l / n;
            const int base = dir == 0 ? (v * n + u) * n : (dir == 1 ? v * n2 + u : v * n + u);
            T         x[n];
#pragma unroll
            for (int k = 0; k < n; ++k)
              x[k] = a[base + k * st];
#pragma unroll
            for (int r = 0; r < n; ++r)
              {
                double s = 0;
#pragma unroll
                for (int k = 0; k < n; ++k)
                  s += (TRANSPOSE ? M.m[k * n + r] : M.m[r * n + k]) * (double)x[k];
                b[base + r * st] = (T)s;
              }
          }
        __syncthreads();
      }
    for (int i = threadIdx.x; i < n3; i += 128)
      {
        if (ADD)
          out[c * n3 + i] += buf[1][i];
        else
          out[c * n3 + i] = buf[1][i];
      }
  }

  struct DGOperator
  {
    mgb_dg_options opt;
    int            p = 0, n = 0, nc[3] = {0, 0, 0};
    size_t         n_cells = 0, n_dofs = 0; // LOCAL cells / vector length: owned layers + the ghost layers of the z-slab
    // z-slabs over the GPUs (one process per GPU): this rank owns `lz` of the nc[2] cell layers, starting at layer cz0;
    // a ghost layer of the rank below (zg) precedes and one of the rank above (ztop) follows them in every vector
    int         rank = 0, n_ranks = 1, lz = 0, cz0 = 0, zg = 0, ztop = 0;
    size_t      layer_dofs = 0, owned_off = 0, owned_n = 0, n_dofs_global = 0;
    mgb_comm_t *comm = nullptr;
    cudaStream_t   stream   = nullptr;
    bool           owns_stream = true;
    cudaEvent_t    ev[2]    = {nullptr, nullptr};
    int64_t        launches = 0;
    DGBasis        basis;
    DGGeom         geom;
    DGDiagTables   dtab;
    double        *scal_dev = nullptr, *scal_host = nullptr;
    void          *inv_diag[2] = {nullptr, nullptr}; // float, double
    void          *host_buf[2] = {nullptr, nullptr}; // device staging of the *_host entry points
    int            preconditioner = 0;                // 0 point Jacobi, 1 JacobiTransformed
    void          *jt_dinv[2] = {nullptr, nullptr};   // [64 cell types][n^3], float / double
    std::vector<double> jt_dinv_host;

    // `borrowed`: the communicator of the FE_Q solver the operator works with (DG multigrid solver); without it
    // and with comm_size > 1 the operator makes its own from o.comm_id
    explicit DGOperator(const mgb_dg_options &o, mgb_comm_t *borrowed = nullptr)
      : opt(o)
    {
      if (o.degree < 1 || o.degree > 8)
        throw std::runtime_error("polynomial degree must be between 1 and 8");
      if (o.n_refinements < 0 || o.n_refinements > 12)
        throw std::runtime_error("invalid number of refinements");
      p = o.degree;
      n = p + 1;
      double J[3][3], h[3];
      n_cells = 1;
      for (int d = 0; d < 3; ++d)
        {
          if (o.n_subdivisions[d] < 1 || !(o.upper[d] > o.lower[d]))
            throw std::runtime_error("invalid mesh description");
          nc[d] = o.n_subdivisions[d] << o.n_refinements;
          h[d]  = (o.upper[d] - o.lower[d]) / nc[d];
          n_cells *= (size_t)nc[d];
        }
      n_dofs_global = n_cells * n * n * n;
      n_ranks       = o.comm_size > 1 ? o.comm_size : 1;
      rank          = n_ranks > 1 ? o.comm_rank : 0;
      if (rank < 0 || rank >= n_ranks)
        throw std::runtime_error("invalid communicator rank");
      if (nc[2] % n_ranks != 0)
        throw std::runtime_error("the number of cell layers in z must be a multiple of the number of GPUs");
      lz         = nc[2] / n_ranks;
      cz0        = rank * lz;
      zg         = rank > 0 ? 1 : 0;
      ztop       = rank + 1 < n_ranks ? 1 : 0;
      layer_dofs = (size_t)nc[0] * nc[1] * n * n * n;
      n_cells    = (size_t)nc[0] * nc[1] * (lz + zg + ztop);
      n_dofs     = n_cells * n * n * n;
      owned_off  = zg * layer_dofs;
      owned_n    = lz * layer_dofs;
      if (n_cells >= 0xffffffffull)
        throw std::runtime_error("too many cells for one device");
      for (int k = 0; k < 3; ++k)
        for (int d = 0; d < 3; ++d)
          J[k][d] = o.trafo[k * 3 + d] * h[d]; // d x_k / d xi_d
      double det = J[0][0] * (J[1][1] * J[2][2] - J[1][2] * J[2][1]) - J[0][1] * (J[1][0] * J[2][2] - J[1][2] * J[2][0]) +
                   J[0][2] * (J[1][0] * J[2][1] - J[1][1] * J[2][0]);
      if (std::fabs(det) < 1e-300)
        throw std::runtime_error("singular vertex map");
      double Ji[3][3]; // Ji[d][k] = d xi_d / d x_k
      for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j)
          {
            const int i1 = (i + 1) % 3, i2 = (i + 2) % 3, j1 = (j + 1) % 3, j2 = (j + 2) % 3;
            Ji[j][i]     = (J[i1][j1] * J[i2][j2] - J[i1][j2] * J[i2][j1]) / det;
          }
      det = std::fabs(det);
      double Gm[3][3], gn[3];
      for (int d = 0; d < 3; ++d)
        for (int e = 0; e < 3; ++e)
          Gm[d][e] = Ji[d][0] * Ji[e][0] + Ji[d][1] * Ji[e][1] + Ji[d][2] * Ji[e][2];
      const int pair[6][2] = {{0, 0}, {1, 1}, {2, 2}, {0, 1}, {0, 2}, {1, 2}};
      for (int e = 0; e < 6; ++e)
        geom.gt[e] = det * Gm[pair[e][0]][pair[e][1]];
      for (int d = 0; d < 3; ++d)
        {
          gn[d]         = std::sqrt(Gm[d][d]);
          geom.sigma[d] = (p + 1.) * (p + 1.) * gn[d]; // penalty_factor = 1 (laplace_operator_dg.h:47,789-793)
          geom.area[d]  = det * gn[d];
          for (int e = 0; e < 3; ++e)
            geom.gs[d][e] = Gm[d][e] / gn[d];
        }
      if (o.basis < DG_BASIS_GAUSS_LOBATTO || o.basis > DG_BASIS_GAUSS)
        throw std::runtime_error("unknown DG basis");
      basis = make_dg_basis(p, o.basis);
      for (int i = 0; i < n; ++i)
        {
          double m = 0, k = 0, c = 0;
          for (int q = 0; q < n; ++q)
            {
              const double s = basis.S[q * n + i], g = basis.G[q * n + i];
              m += basis.wq[q] * s * s;
              k += basis.wq[q] * g * g;
              c += basis.wq[q] * g * s;
            }
          dtab.M[i]  = m;
          dtab.K[i]  = k;
          dtab.C[i]  = c;
          dtab.e0[i] = basis.e0[i];
          dtab.e1[i] = basis.e1[i];
          dtab.f0[i] = basis.f0[i];
          dtab.f1[i] = basis.f1[i];
        }
      int count = 0;
      if (cudaGetDeviceCount(&count) != cudaSuccess || count == 0)
        throw std::runtime_error("no CUDA device: the DG operator has no CPU fallback");
      MGB_DG_CHECK(cudaSetDevice(o.device));
      MGB_DG_CHECK(cudaStreamCreateWithFlags(&stream, cudaStreamNonBlocking));
      MGB_DG_CHECK(cudaEventCreate(&ev[0]));
      MGB_DG_CHECK(cudaEventCreate(&ev[1]));
      MGB_DG_CHECK(cudaMalloc(&scal_dev, 8 * sizeof(double)));
      MGB_DG_CHECK(cudaMallocHost(&scal_host, 8 * sizeof(double)));
      if (borrowed)
        comm = borrowed;
      else if (n_ranks > 1)
        comm = mgb_internal_comm_create(rank, n_ranks, o.comm_id, o.device);
    }

    ~DGOperator()
    {
      for (int i = 0; i < 2; ++i)
        {
          if (inv_diag[i])
            cudaFree(inv_diag[i]);
          if (host_buf[i])
            cudaFree(host_buf[i]);
          if (jt_dinv[i])
            cudaFree(jt_dinv[i]);
          if (ev[i])
            cudaEventDestroy(ev[i]);
        }
      for (int k = 0; k < 2; ++k)
        for (int i = 0; i < 3; ++i)
          if (cg_ws[k][i])
            cudaFree(cg_ws[k][i]);
      if (cg_partial)
        cudaFree(cg_partial);
      if (scal_dev)
        cudaFree(scal_dev);
      if (scal_host)
        cudaFreeHost(scal_host);
      if (stream && owns_stream)
        cudaStreamDestroy(stream);
      if (comm)
        mgb_internal_comm_destroy(comm);
      if (comm_stream)
        {
          cudaStreamDestroy(comm_stream);
          cudaEventDestroy(comm_ev[0]);
          cudaEventDestroy(comm_ev[1]);
        }
    }

    // update_ghost_values (the MPI face exchange of laplace_operator_dg.h:981-1058): the top owned cell layer goes to
    // the rank above, the bottom one to the rank below, one NCCL group on the operator's stream
    template <typename T>
    void
    update_ghosts(T *v, cudaStream_t on = nullptr)
    {
      if (n_ranks == 1)
        return;
      mgb_internal_comm_exchange(comm, on ? on : stream, v + owned_off + owned_n - layer_dofs, v, v + owned_off,
                                 v + owned_off + owned_n, layer_dofs, sizeof(T) == 8);
      ++exchanges;
    }
    int64_t exchanges = 0;

    unsigned grid(const size_t count) const { return (unsigned)std::min<size_t>((count + 255) / 256, 148 * 16); }

    // run on another stream (the FE_Q solver's, for the DG multigrid solver)
    void
    use_stream(cudaStream_t s)
    {
      if (owns_stream && stream)
        cudaStreamDestroy(stream);
      stream      = s;
      owns_stream = false;
    }
    template <typename F>
    void foreach (const size_t count, F f)
    {
      k_dg_foreach<<<grid(count), 256, 0, stream>>>(count, f);
      ++launches;
      MGB_DG_CHECK(cudaGetLastError());
    }
    // r[0..NS) = sums over i of what f adds to acc[]; one host synchronisation (the reference's MPI sum)
    template <int NS, typename F>
    void
    sums(double *r, F f)
    {
      static_assert(NS <= 4, "scal_dev holds four sums");
      MGB_DG_CHECK(cudaMemsetAsync(scal_dev, 0, 4 * sizeof(double), stream));
      k_dg_reduce<NS><<<grid(owned_n), 256, 0, stream>>>(owned_off, owned_n, f, scal_dev);
      ++launches;
      if (n_ranks > 1)
        mgb_internal_comm_allreduce(comm, stream, scal_dev, NS);
      MGB_DG_CHECK(cudaMemcpyAsync(scal_host, scal_dev, 4 * sizeof(double), cudaMemcpyDeviceToHost, stream));
      MGB_DG_CHECK(cudaStreamSynchronize(stream));
      for (int s = 0; s < NS; ++s)
        r[s] = scal_host[s];
    }

    template <typename T>
    T *
    get_inverse_diagonal()
    {
      const int k = sizeof(T) == 8;
      if (!inv_diag[k])
        {
          MGB_DG_CHECK(cudaMalloc(&inv_diag[k], n_dofs * sizeof(T)));
          k_dg_inverse_diagonal<T><<<grid(n_dofs), 256, 0, stream>>>(n, nc[0], nc[1], nc[2], lz + zg + ztop, cz0 - zg, geom, dtab,
                                                                     (T *)inv_diag[k]);
          ++launches;
          MGB_DG_CHECK(cudaGetLastError());
        }
      return (T *)inv_diag[k];
    }

    // JacobiTransformed::local_compute_diagonals (laplace_operator_dg.h:2098-2233): D = diag(V^T A_KK V) with the
    // cell-local matrix A_KK (cell term + the cell's own share of its six faces), which on this mesh depends on the
    // cell only through which of its faces lie on the boundary.  A_KK is probed with the operator kernel itself
    // on a mesh of at most 3 x 3 x 3 cells of the same size that contains every type: cells of equal parity are
    // no neighbours, so 8 colours x n^3 transformed unit vectors give all diagonals.
    void
    build_jacobi_transformed()
    {
      if (!jt_dinv_host.empty())
        return;
      const int      n3 = n * n * n;
      mgb_dg_options po = opt;
      int            ncp[3];
      for (int d = 0; d < 3; ++d)
        {
          ncp[d]               = std::min(nc[d], 3);
          po.n_subdivisions[d] = ncp[d];
          po.upper[d]          = opt.lower[d] + (opt.upper[d] - opt.lower[d]) / nc[d] * ncp[d];
        }
      po.n_refinements = 0;
      po.comm_rank     = 0;
      po.comm_size     = 1;
      DGOperator                probe(po);
      const std::vector<double> V = jacobi_transformed_eigenvectors(basis);
      const size_t              np = probe.n_dofs;
      double                   *du = nullptr, *dy = nullptr;
      MGB_DG_CHECK(cudaMalloc(&du, np * sizeof(double)));
      MGB_DG_CHECK(cudaMalloc(&dy, np * sizeof(double)));
      std::vector<double> u(np), y(np), vk(n3);
      jt_dinv_host.assign((size_t)64 * n3, 0.);
      for (int k = 0; k < n3; ++k)
        {
          const int k0 = k % n, k1 = (k / n) % n, k2 = k / (n * n);
          for (int l2 = 0; l2 < n; ++l2)
            for (int l1 = 0; l1 < n; ++l1)
              for (int l0 = 0; l0 < n; ++l0)
                vk[(l2 * n + l1) * n + l0] = V[k0 * n + l0] * V[k1 * n + l1] * V[k2 * n + l2];
          for (int colour = 0; colour < 8; ++colour)
            {
              std::fill(u.begin(), u.end(), 0.);
              bool any = false;
              for (size_t c = 0; c < probe.n_cells; ++c)
                {
                  const int cx = c % ncp[0], cy = (c / ncp[0]) % ncp[1], cz = c / (ncp[0] * ncp[1]);
                  if (((cx & 1) | ((cy & 1) << 1) | ((cz & 1) << 2)) != colour)
                    continue;
                  any = true;
                  std::copy(vk.begin(), vk.end(), u.begin() + c * n3);
                }
              if (!any)
                continue;
              MGB_DG_CHECK(cudaMemcpyAsync(du, u.data(), np * sizeof(double), cudaMemcpyHostToDevice, probe.stream));
              probe.apply<double>(DG_SET, dy, du);
              MGB_DG_CHECK(cudaMemcpyAsync(y.data(), dy, np * sizeof(double), cudaMemcpyDeviceToHost, probe.stream));
              MGB_DG_CHECK(cudaStreamSynchronize(probe.stream));
              for (size_t c = 0; c < probe.n_cells; ++c)
                {
                  const int cx = c % ncp[0], cy = (c / ncp[0]) % ncp[1], cz = c / (ncp[0] * ncp[1]);
                  if (((cx & 1) | ((cy & 1) << 1) | ((cz & 1) << 2)) != colour)
                    continue;
                  double dsum = 0;
                  for (int l = 0; l < n3; ++l)
                    dsum += vk[l] * y[c * n3 + l];
                  const int tx = (cx == 0 ? 1 : 0) + (cx == ncp[0] - 1 ? 2 : 0);
                  const int ty = (cy == 0 ? 1 : 0) + (cy == ncp[1] - 1 ? 2 : 0);
                  const int tz = (cz == 0 ? 1 : 0) + (cz == ncp[2] - 1 ? 2 : 0);
                  jt_dinv_host[(size_t)((tz * 4 + ty) * 4 + tx) * n3 + k] = 1. / dsum;
                }
            }
        }
      cudaFree(du);
      cudaFree(dy);
      launches += probe.launches;
      MGB_DG_CHECK(cudaSetDevice(opt.device));
    }

    template <typename T>
    const T *
    get_jt_dinv()
    {
      const int k = sizeof(T) == 8;
      if (!jt_dinv[k])
        {
          build_jacobi_transformed();
          std::vector<T> h(jt_dinv_host.begin(), jt_dinv_host.end());
          MGB_DG_CHECK(cudaMalloc(&jt_dinv[k], h.size() * sizeof(T)));
          MGB_DG_CHECK(cudaMemcpyAsync(jt_dinv[k], h.data(), h.size() * sizeof(T), cudaMemcpyHostToDevice, stream));
          MGB_DG_CHECK(cudaStreamSynchronize(stream));
        }
      return (const T *)jt_dinv[k];
    }

    template <typename T, bool TRANSPOSE, bool ADD>
    void
    basis_change(const std::vector<double> &M, const T *in, T *out)
    {
      DGMatrix1D m;
      for (int i = 0; i < n * n; ++i)
        m.m[i] = M[i];
      const size_t cells = n_dofs / ((size_t)n * n * n);
      switch (n)
        {
#define MGB_BC(N) \
  case N: k_dg_basis_change<N, T, TRANSPOSE, ADD><<<(unsigned)cells, 128, 0, stream>>>(cells, m, in, out); break;
          MGB_BC(2) MGB_BC(3) MGB_BC(4) MGB_BC(5) MGB_BC(6) MGB_BC(7) MGB_BC(8) MGB_BC(9)
#undef MGB_BC
          default: throw std::runtime_error("unsupported degree");
        }
      ++launches;
      MGB_DG_CHECK(cudaGetLastError());
    }

    void
    check_same_mesh(const LevelDesc &L) const
    {
      // on several GPUs both sides hold the same z-slab: the FE_Q box is the owned cell layers plus the ghost layer
      // below, so FE_Q cell c and DG cell c are the same cell (the DG vector has one more ghost layer on top)
      if (L.p != p || L.ncell[0] != nc[0] || L.ncell[1] != nc[1] || L.ncell[2] != lz + zg || L.zg != zg || L.ztop != ztop)
        throw std::runtime_error("the FE_Q level and the DG operator must share degree, mesh and partition");
    }
    // prolongate_add_cg_to_dg (laplace_operator_dg.h:1863-1894): dg += cg.  On several GPUs the caller has updated
    // the ghost planes of cg (the top plane of the last owned cell layer is a ghost)
    template <typename T>
    void
    prolongate_add_cg_to_dg(const LevelDesc &L, T *dg, const T *cg)
    {
      check_same_mesh(L);
      if (basis.kind == DG_BASIS_GAUSS_LOBATTO)
        k_cg_to_dg_add<T><<<grid(n_dofs), 256, 0, stream>>>(L, cg, dg);
      else
        { // values in the Gauss-Lobatto points of every cell, then the coefficients of the DG basis
          if (!host_buf[1])
            MGB_DG_CHECK(cudaMalloc(&host_buf[1], n_dofs * 8));
          T *tmp = (T *)host_buf[1];
          MGB_DG_CHECK(cudaMemsetAsync(tmp, 0, n_dofs * sizeof(T), stream));
          k_cg_to_dg_add<T><<<grid(n_dofs), 256, 0, stream>>>(L, cg, tmp);
          basis_change<T, false, true>(basis.Binv, tmp, dg);
        }
      ++launches;
      MGB_DG_CHECK(cudaGetLastError());
    }
    // vmult_residual_and_restrict_to_cg (laplace_operator_dg.h:851-860, action 1 :1798-1819): restricted = R (rhs - A lhs)
    template <typename T>
    void
    vmult_residual_and_restrict_to_cg(const LevelDesc &L, const T *rhs, const T *lhs, T *restricted)
    {
      check_same_mesh(L);
      for (int i = 0; i < 2; ++i)
        if (!host_buf[i])
          MGB_DG_CHECK(cudaMalloc(&host_buf[i], n_dofs * 8));
      T *res = (T *)host_buf[0];
      apply<T>(DG_RESIDUAL, res, lhs, rhs);
      update_ghosts(res); // the first owned FE_Q plane sums over the cell layer below (compress(add) of the reference)
      if (basis.kind != DG_BASIS_GAUSS_LOBATTO)
        basis_change<T, true, false>(basis.Binv, res, res); // transpose of the prolongation's change of basis
      MGB_DG_CHECK(cudaMemsetAsync(restricted, 0, (size_t)L.n_dofs * sizeof(T), stream));
      k_dg_to_cg_add<T><<<grid(L.n_dofs), 256, 0, stream>>>(L, res, restricted);
      ++launches;
      MGB_DG_CHECK(cudaGetLastError());
    }

    // 1-D tables of the operator's basis (kernel parameters), built once per degree, precision and basis
    template <int P, typename T>
    const DGTables<T, P + 1> &
    tables() const
    {
      static DGTables<T, P + 1> tb[3];
      static bool               made[3] = {false, false, false};
      if (!made[basis.kind])
        {
          tb[basis.kind]   = make_tables<T, P + 1>(make_dg_basis(P, basis.kind));
          made[basis.kind] = true;
        }
      return tb[basis.kind];
    }

    template <int P, typename T>
    void
    launch_jt(const DGArgs<T> &A, const T scale)
    {
      launch_dg_jacobi_transformed<P, T>(A, tables<P, T>(), scale, stream);
    }

    // dst = scale * P^-1 src with the selected preconditioner (JacobiTransformed::vmult, :2046-2083, or point Jacobi)
    template <typename T>
    void
    preconditioner_vmult(T *dst, const T *src, const double scale)
    {
      if (preconditioner == 0)
        {
          const T     *P = get_inverse_diagonal<T>();
          const T      f = (T)scale;
          const size_t N = n_dofs;
          k_dg_foreach<<<grid(N), 256, 0, stream>>>(N, [=] __device__(size_t i) { dst[i] = f * P[i] * src[i]; });
        }
      else
        {
          DGArgs<T> A;
          for (int d = 0; d < 3; ++d)
            A.nc[d] = nc[d];
          A.n_cells    = (uint32_t)((size_t)nc[0] * nc[1] * lz);
          A.cell_begin = (uint32_t)((size_t)nc[0] * nc[1] * zg);
          A.cz_shift   = cz0 - zg;
          A.src     = src;
          A.dst     = dst;
          A.jt_dinv = get_jt_dinv<T>();
          switch (p)
            {
              case 1: launch_jt<1, T>(A, (T)scale); break;
              case 2: launch_jt<2, T>(A, (T)scale); break;
              case 3: launch_jt<3, T>(A, (T)scale); break;
              case 4: launch_jt<4, T>(A, (T)scale); break;
              case 5: launch_jt<5, T>(A, (T)scale); break;
              case 6: launch_jt<6, T>(A, (T)scale); break;
              case 7: launch_jt<7, T>(A, (T)scale); break;
              default: launch_jt<8, T>(A, (T)scale); break;
            }
        }
      ++launches;
      MGB_DG_CHECK(cudaGetLastError());
    }

    template <int P, typename T>
    void
    launch(const DGArgs<T> &A)
    {
      static const bool generic_only = std::getenv("MGB200_DG_GENERIC") != nullptr; // A/B switch of the timing tools
      if (basis.kind == DG_BASIS_GAUSS_LOBATTO)
        launch_dg_apply<P, T, 0>(A, tables<P, T>(), stream);
      else if (basis.kind == DG_BASIS_GAUSS && !generic_only)
        launch_dg_apply<P, T, 2>(A, tables<P, T>(), stream);
      else
        launch_dg_apply<P, T, 1>(A, tables<P, T>(), stream);
    }

    template <typename T>
    void
    apply(const int mode, T *dst, const T *src, const T *rhs = nullptr, const double f1 = 0, const double f2 = 0,
          const T *diagonal_preconditioner = nullptr, double *partial = nullptr)
    {
      DGArgs<T> A;
      A.partial = partial;
      for (int d = 0; d < 3; ++d)
        {
          A.nc[d]    = nc[d];
          A.sigma[d] = (T)geom.sigma[d];
          A.area[d]  = (T)geom.area[d];
          for (int e = 0; e < 3; ++e)
            A.gs[d][e] = (T)geom.gs[d][e];
        }
      for (int e = 0; e < 6; ++e)
        A.gt[e] = (T)geom.gt[e];
      const size_t layer_cells = (size_t)nc[0] * nc[1];
      A.n_cells    = (uint32_t)(layer_cells * lz);
      A.cell_begin = (uint32_t)(layer_cells * zg);
      A.cz_shift   = cz0 - zg;
      A.src      = src;
      A.dst      = dst;
      A.rhs      = rhs;
      const bool chebyshev = mode == DG_CHEBYSHEV || mode == DG_CHEBYSHEV_FIRST;
      A.inv_diag = (chebyshev && preconditioner == 0) ? get_inverse_diagonal<T>() : nullptr;
      A.jt_dinv  = (chebyshev && preconditioner == 1) ? get_jt_dinv<T>() : nullptr;
      if (mode == DG_SET_SUMS)
        {
          if (n_ranks > 1)
            throw std::runtime_error("the interleaved CG runs on one GPU");
          A.inv_diag = diagonal_preconditioner;
        }
      A.f1       = (T)f1;
      A.f2       = (T)f2;
      A.mode     = mode;
      const auto run = [&](const DGArgs<T> &B) {
        switch (p)
          {
            case 1: launch<1, T>(B); break;
            case 2: launch<2, T>(B); break;
            case 3: launch<3, T>(B); break;
            case 4: launch<4, T>(B); break;
            case 5: launch<5, T>(B); break;
            case 6: launch<6, T>(B); break;
            case 7: launch<7, T>(B); break;
            default: launch<8, T>(B); break;
          }
        ++launches;
        MGB_DG_CHECK(cudaGetLastError());
      };
      if (n_ranks > 1 && overlap && lz >= 3)
        {
          // the exchange of the ghost cell layers runs on the communication stream while the cell layers that touch no
          // ghost layer are computed (laplace_operator_dg.h:981-1058 overlaps its MPI messages the same way); then the
          // first and the last owned layer
          if (!comm_stream)
            {
              MGB_DG_CHECK(cudaStreamCreateWithFlags(&comm_stream, cudaStreamNonBlocking));
              MGB_DG_CHECK(cudaEventCreateWithFlags(&comm_ev[0], cudaEventDisableTiming));
              MGB_DG_CHECK(cudaEventCreateWithFlags(&comm_ev[1], cudaEventDisableTiming));
            }
          MGB_DG_CHECK(cudaEventRecord(comm_ev[0], stream));
          MGB_DG_CHECK(cudaStreamWaitEvent(comm_stream, comm_ev[0], 0));
          update_ghosts(const_cast<T *>(src), comm_stream);
          MGB_DG_CHECK(cudaEventRecord(comm_ev[1], comm_stream));
          const int first = zg ? 1 : 0, last = ztop ? 1 : 0; // owned layers that read a ghost layer
          DGArgs<T> B   = A;
          B.cell_begin  = (uint32_t)(layer_cells * (zg + first));
          B.n_cells     = (uint32_t)(layer_cells * (lz - first - last));
          run(B);
          MGB_DG_CHECK(cudaStreamWaitEvent(stream, comm_ev[1], 0));
          if (first)
            {
              B.cell_begin = (uint32_t)(layer_cells * zg);
              B.n_cells    = (uint32_t)layer_cells;
              run(B);
            }
          if (last)
            {
              B.cell_begin = (uint32_t)(layer_cells * (zg + lz - 1));
              B.n_cells    = (uint32_t)layer_cells;
              run(B);
            }
          return;
        }
      update_ghosts(const_cast<T *>(src));
      run(A);
    }
    bool         overlap     = getenv("MGB200_DG_OVERLAP") == nullptr || getenv("MGB200_DG_OVERLAP")[0] != '0';
    cudaStream_t comm_stream = nullptr;
    cudaEvent_t  comm_ev[2]  = {nullptr, nullptr};

    // laplace_operator_dg.h:862-905
    template <typename T>
    void
    vmult_with_cg_update(const T alpha, const T beta, const T *r, T *q, T *pp, T *x, double sums[4])
    {
      const size_t N = n_dofs;
      if (alpha == T(0))
        k_dg_foreach<<<grid(N), 256, 0, stream>>>(N, [=] __device__(size_t i) { pp[i] = q[i]; });
      else
        k_dg_foreach<<<grid(N), 256, 0, stream>>>(N, [=] __device__(size_t i) {
          x[i] += alpha * pp[i];
          pp[i] = beta * pp[i] + q[i];
        });
      ++launches;
      apply<T>(DG_SET, q, pp);
      MGB_DG_CHECK(cudaMemsetAsync(scal_dev, 0, 4 * sizeof(double), stream));
      k_dg_cg_sums<T><<<grid(owned_n), 256, 0, stream>>>(owned_off, owned_n, q, pp, r, scal_dev);
      ++launches;
      if (n_ranks > 1)
        mgb_internal_comm_allreduce(comm, stream, scal_dev, 4);
      MGB_DG_CHECK(cudaMemcpyAsync(scal_host, scal_dev, 4 * sizeof(double), cudaMemcpyDeviceToHost, stream));
      MGB_DG_CHECK(cudaStreamSynchronize(stream));
      for (int s = 0; s < 4; ++s)
        sums[s] = scal_host[s];
    }

    // solver_dg/program.cc:180-262: SolverCG with IterationNumberControl(n_iterations, 1e-20) from x = 0 on this operator
    // with a diagonal preconditioner (there: the inverse of the lumped mass matrix).
    // interleaved = false, "cell-based loops" (:216-227): operator apply, then the vector updates and dot products as
    // separate passes.  interleaved = true, "interleaved loops" (LaplaceOperatorInterleave, :39-70, :237-248): deal.II's
    // SolverCG hands the vector updates to the operator's loop; here one vector kernel before the apply
    // (x += alpha p, r -= alpha v, p = D r + beta p) and the sums p.v, v.Dr, v.Dv, r.v, v.v in the apply's store stage,
    // from which alpha, beta = (r.Dr - 2 alpha v.Dr + alpha^2 v.Dv) / r.Dr and |r|^2 follow without another pass.
    // stats = {iterations done, last residual norm}
    void   *cg_ws[2][3] = {{nullptr, nullptr, nullptr}, {nullptr, nullptr, nullptr}};
    double *cg_partial  = nullptr;
    template <typename T>
    void
    solve_cg_diagonal(const unsigned int n_iterations, const bool interleaved, T *x, const T *b, const T *D, double stats[2])
    {
      const size_t N = n_dofs;
      const int    k = sizeof(T) == 8;
      for (int i = 0; i < 3; ++i)
        if (!cg_ws[k][i])
          MGB_DG_CHECK(cudaMalloc(&cg_ws[k][i], N * sizeof(T)));
      T *r = (T *)cg_ws[k][0], *pp = (T *)cg_ws[k][1], *v = (T *)cg_ws[k][2];
      MGB_DG_CHECK(cudaMemsetAsync(x, 0, N * sizeof(T), stream));
      double sum[4];
      sums<2>(sum, [=] __device__(size_t i, double *acc) {
        const T ri = b[i], zi = D[i] * ri;
        r[i]       = ri;
        pp[i]      = zi;
        acc[0] += (double)ri * zi;
        acc[1] += (double)ri * ri;
      });
      double       rz = sum[0], rr = sum[1];
      unsigned int it = 0;
      if (!interleaved)
        {
          while (it < n_iterations && std::sqrt(rr) > 1e-20)
            {
              ++it;
              apply<T>(DG_SET, v, pp);
              sums<1>(sum, [=] __device__(size_t i, double *acc) { acc[0] += (double)pp[i] * v[i]; });
              const T alpha = (T)(rz / sum[0]);
              sums<2>(sum, [=] __device__(size_t i, double *acc) {
                x[i] += alpha * pp[i];
                const T ri = r[i] - alpha * v[i];
                r[i]       = ri;
                acc[0] += (double)ri * (D[i] * ri);
                acc[1] += (double)ri * ri;
              });
              const T beta = (T)(sum[0] / rz);
              rz           = sum[0];
              rr           = sum[1];
              foreach (N, [=] __device__(size_t i) { pp[i] = D[i] * r[i] + beta * pp[i]; });
            }
        }
      else
        {
          if (n_ranks > 1)
            throw std::runtime_error("the interleaved CG runs on one GPU");
          const int    tpc = ((n * n + 31) / 32) * 32, cpb = tpc >= 128 ? 1 : 128 / tpc; // DGCfg<p>
          const size_t n_blocks = ((size_t)nc[0] * nc[1] * lz + cpb - 1) / cpb;
          if (!cg_partial)
            MGB_DG_CHECK(cudaMalloc(&cg_partial, n_blocks * 5 * sizeof(double)));
          double alpha = 0, beta = 0;
          while (it < n_iterations && std::sqrt(std::fabs(rr)) > 1e-20)
            {
              if (it > 0)
                {
                  const T a = (T)alpha, be = (T)beta;
                  foreach (N, [=] __device__(size_t i) {
                    const T pi = pp[i];
                    x[i] += a * pi;
                    const T ri = r[i] - a * v[i];
                    r[i]       = ri;
                    pp[i]      = D[i] * ri + be * pi;
                  });
                }
              ++it;
              apply<T>(DG_SET_SUMS, v, pp, r, 0., 0., D, cg_partial);
              MGB_DG_CHECK(cudaMemsetAsync(scal_dev, 0, 8 * sizeof(double), stream));
              k_dg_sum_partials<<<(unsigned)std::min<size_t>((n_blocks + 255) / 256, 296), 256, 0, stream>>>(n_blocks, cg_partial,
                                                                                                          scal_dev);
              ++launches;
              MGB_DG_CHECK(cudaMemcpyAsync(scal_host, scal_dev, 5 * sizeof(double), cudaMemcpyDeviceToHost, stream));
              MGB_DG_CHECK(cudaStreamSynchronize(stream));
              const double pv = scal_host[0], vDr = scal_host[1], vDv = scal_host[2], rv = scal_host[3], vv = scal_host[4];
              alpha               = rz / pv;
              const double rz_new = rz - 2. * alpha * vDr + alpha * alpha * vDv;
              rr                  = rr - 2. * alpha * rv + alpha * alpha * vv;
              beta                = rz_new / rz;
              rz                  = rz_new;
            }
          if (it > 0)
            {
              const T a = (T)alpha;
              foreach (N, [=] __device__(size_t i) {
                x[i] += a * pp[i];
                r[i] -= a * v[i];
              });
            }
        }
      MGB_DG_CHECK(cudaStreamSynchronize(stream));
      stats[0] = it;
      stats[1] = std::sqrt(std::fabs(rr));
    }

    // |det J| of the (affine) cells
    double
    cell_volume() const
    {
      double J[3][3];
      for (int k = 0; k < 3; ++k)
        for (int d = 0; d < 3; ++d)
          J[k][d] = opt.trafo[k * 3 + d] * (opt.upper[d] - opt.lower[d]) / nc[d];
      return std::fabs(J[0][0] * (J[1][1] * J[2][2] - J[1][2] * J[2][1]) - J[0][1] * (J[1][0] * J[2][2] - J[1][2] * J[2][0]) +
                       J[0][2] * (J[1][0] * J[2][1] - J[1][1] * J[2][0]));
    }

    // the diagonal of solver_dg/program.cc:134-148: 1 / (1, phi_i), the inverse of the lumped mass matrix
    template <typename T>
    void
    lumped_mass_inverse(T *out)
    {
      std::vector<double> col(n, 0.);
      for (int i = 0; i < n; ++i)
        for (int q = 0; q < n; ++q)
          col[i] += basis.wq[q] * basis.S[q * n + i];
      const int      n3 = n * n * n;
      std::vector<T> cellv(n3);
      const double   detJ = cell_volume();
      for (int i2 = 0; i2 < n; ++i2)
        for (int i1 = 0; i1 < n; ++i1)
          for (int i0 = 0; i0 < n; ++i0)
            cellv[(i2 * n + i1) * n + i0] = (T)(1. / (detJ * col[i0] * col[i1] * col[i2]));
      T *dev = nullptr;
      MGB_DG_CHECK(cudaMalloc(&dev, n3 * sizeof(T)));
      MGB_DG_CHECK(cudaMemcpyAsync(dev, cellv.data(), n3 * sizeof(T), cudaMemcpyHostToDevice, stream));
      const T *cv = dev;
      foreach (n_dofs, [=] __device__(size_t i) { out[i] = cv[i % n3]; });
      MGB_DG_CHECK(cudaStreamSynchronize(stream));
      cudaFree(dev);
    }

    // laplace_operator_dg.h:907-955 with the point-Jacobi preconditioner
    template <typename T>
    void
    chebyshev_update(const T *rhs, const int iteration_index, const double f1, const double f2, T *solution, T *solution_old)
    {
      if (iteration_index > 0)
        {
          apply<T>(iteration_index == 1 ? DG_CHEBYSHEV_FIRST : DG_CHEBYSHEV, solution_old, solution, rhs, f1, f2);
          return;
        }
      preconditioner_vmult<T>(solution, rhs, f2);
    }
  };

  // extreme eigenvalues of the Lanczos tridiagonal matrix of the eigenvalue-estimating CG (Sturm bisection)
  static void
  dg_tridiagonal_extreme_eigenvalues(const std::vector<double> &diag, const std::vector<double> &off, double &emin,
                                     double &emax)
  {
    const int k = (int)diag.size();
    emin = emax = 1.;
    if (k == 0)
      return;
    double lo = 1e300, hi = -1e300;
    for (int i = 0; i < k; ++i)
      {
        const double r = (i > 0 ? std::fabs(off[i - 1]) : 0.) + (i < k - 1 ? std::fabs(off[i]) : 0.);
        lo             = std::min(lo, diag[i] - r);
        hi             = std::max(hi, diag[i] + r);
      }
    const auto count_below = [&](const double x) {
      int    cnt = 0;
      double q   = 1;
      for (int i = 0; i < k; ++i)
        {
          q = diag[i] - x - (i > 0 ? off[i - 1] * off[i - 1] / q : 0.);
          if (q == 0)
            q = 1e-300;
          cnt += q < 0;
        }
      return cnt;
    };
    const auto kth = [&](const int idx) {
      double a = lo, c = hi;
      for (int it = 0; it < 200; ++it)
        {
          const double mid = 0.5 * (a + c);
          (count_below(mid) > idx ? c : a) = mid;
        }
      return 0.5 * (a + c);
    };
    emin = kth(0);
    emax = kth(k - 1);
  }
} // namespace mgb

#include "dg_solver.cuh"
#include "dg_plain_solver.cuh"

// ====================================================================== C ABI
using namespace mgb;

struct mgb_dg_t
{
  DGOperator *impl;
};

// error string shared with solver.cu's mgb_last_error()
void mgb_set_last_error(const char *msg);

#define DG_TRY try {
#define DG_CATCH                                   \
  }                                                \
  catch (const std::exception &e)                  \
  {                                                \
    mgb_set_last_error(e.what());                  \
    return 1;                                      \
  }                                                \
  catch (...)                                      \
  {                                                \
    mgb_set_last_error("unknown error");           \
    return 2;                                      \
  }                                                \
  return 0;

static DGOperator &
DGH(mgb_dg h)
{
  if (!h || !h->impl)
    throw std::runtime_error("null DG operator handle");
  return *h->impl;
}

struct mgb_dg_solver_t
{
  DGMultigridSolver *impl;
  mgb_dg_t           op; // view of impl->op for the operator-level entry points
};

static DGMultigridSolver &
DGS(mgb_dg_solver h)
{
  if (!h || !h->impl)
    throw std::runtime_error("null DG solver handle");
  return *h->impl;
}

struct mgb_dg_plain_solver_t
{
  DGPlainSolver        *impl;
  std::vector<mgb_dg_t> ops; // views of the level operators
};

static DGPlainSolver &
DGP(mgb_dg_plain_solver h)
{
  if (!h || !h->impl)
    throw std::runtime_error("null DG plain solver handle");
  return *h->impl;
}

static int
check_level(const DGPlainSolver &m, const int level)
{
  if (level < 0 || level >= m.n_levels)
    throw std::runtime_error("level out of range");
  return level;
}

static void
check_precision(const int precision)
{
  if (precision != MGB_F32 && precision != MGB_F64)
    throw std::runtime_error("precision must be MGB_F32 or MGB_F64");
}

extern "C"
{
  void
  mgb_dg_default_options(mgb_dg_options *o)
  {
    std::memset(o, 0, sizeof(*o));
    o->degree = 4;
    for (int d = 0; d < 3; ++d)
      {
        o->n_subdivisions[d] = 1;
        o->lower[d]          = -1. + 0.05 * (d + 1); // matvec_dg/program.cc:55-60
        o->upper[d]          = 0.95 - 0.06 * d;
        for (int e = 0; e < 3; ++e)
          o->trafo[d * 3 + e] = (d == e ? 1. : 0.) + 0.12 * (d + 1) * (e + 1); // :66-71
      }
  }

  int
  mgb_dg_create(const mgb_dg_options *opt, mgb_dg *out)
  {
    DG_TRY
    if (!opt || !out)
      throw std::runtime_error("null argument");
    *out         = new mgb_dg_t();
    (*out)->impl = nullptr;
    try
      {
        (*out)->impl = new DGOperator(*opt);
      }
    catch (...)
      {
        delete *out;
        *out = nullptr;
        throw;
      }
    DG_CATCH
  }

  int
  mgb_dg_destroy(mgb_dg h)
  {
    DG_TRY
    if (h)
      {
        delete h->impl;
        delete h;
      }
    DG_CATCH
  }

  int
  mgb_dg_sync(mgb_dg h)
  {
    DG_TRY
    MGB_DG_CHECK(cudaStreamSynchronize(DGH(h).stream));
    DG_CATCH
  }

  int64_t
  mgb_dg_n_dofs(mgb_dg h)
  {
    return h && h->impl ? (int64_t)h->impl->n_dofs : -1;
  }

  int
  mgb_dg_n_cells(mgb_dg h, int ncell[3])
  {
    DG_TRY
    for (int d = 0; d < 3; ++d)
      ncell[d] = DGH(h).nc[d];
    DG_CATCH
  }

  int
  mgb_dg_partition(mgb_dg h, int64_t info[6])
  {
    DG_TRY
    const DGOperator &o = DGH(h);
    info[0]             = (int64_t)o.n_dofs_global;
    info[1]             = (int64_t)o.n_dofs;
    info[2]             = (int64_t)o.owned_off;
    info[3]             = (int64_t)o.owned_n;
    info[4]             = o.cz0;
    info[5]             = o.lz;
    DG_CATCH
  }

  int64_t
  mgb_dg_exchange_count(mgb_dg h)
  {
    return h && h->impl ? h->impl->exchanges : -1;
  }

  int
  mgb_dg_penalty(mgb_dg h, double sigma[3])
  {
    DG_TRY
    for (int d = 0; d < 3; ++d)
      sigma[d] = DGH(h).geom.sigma[d];
    DG_CATCH
  }

  int
  mgb_dg_vmult(mgb_dg h, int precision, void *dst, const void *src)
  {
    DG_TRY
    check_precision(precision);
    if (precision == MGB_F64)
      DGH(h).apply<double>(DG_SET, (double *)dst, (const double *)src);
    else
      DGH(h).apply<float>(DG_SET, (float *)dst, (const float *)src);
    DG_CATCH
  }

  int
  mgb_dg_vmult_residual(mgb_dg h, int precision, const void *rhs, const void *lhs, void *residual)
  {
    DG_TRY
    check_precision(precision);
    if (precision == MGB_F64)
      DGH(h).apply<double>(DG_RESIDUAL, (double *)residual, (const double *)lhs, (const double *)rhs);
    else
      DGH(h).apply<float>(DG_RESIDUAL, (float *)residual, (const float *)lhs, (const float *)rhs);
    DG_CATCH
  }

  int
  mgb_dg_vmult_with_cg_update(mgb_dg h, int precision, double alpha, double beta, const void *r, void *q, void *p,
                              void *x, double sums[4])
  {
    DG_TRY
    check_precision(precision);
    if (precision == MGB_F64)
      DGH(h).vmult_with_cg_update<double>(alpha, beta, (const double *)r, (double *)q, (double *)p, (double *)x, sums);
    else
      DGH(h).vmult_with_cg_update<float>((float)alpha, (float)beta, (const float *)r, (float *)q, (float *)p, (float *)x,
                                         sums);
    DG_CATCH
  }

  int
  mgb_dg_inverse_diagonal(mgb_dg h, int precision, void *inv_diag)
  {
    DG_TRY
    check_precision(precision);
    DGOperator &o = DGH(h);
    if (precision == MGB_F64)
      MGB_DG_CHECK(cudaMemcpyAsync(inv_diag, o.get_inverse_diagonal<double>(), o.n_dofs * 8, cudaMemcpyDeviceToDevice, o.stream));
    else
      MGB_DG_CHECK(cudaMemcpyAsync(inv_diag, o.get_inverse_diagonal<float>(), o.n_dofs * 4, cudaMemcpyDeviceToDevice, o.stream));
    DG_CATCH
  }

  int
  mgb_dg_vmult_with_chebyshev_update(mgb_dg h, int precision, const void *rhs, int iteration_index, double factor1,
                                     double factor2, void *solution, void *solution_old)
  {
    DG_TRY
    check_precision(precision);
    if (precision == MGB_F64)
      DGH(h).chebyshev_update<double>((const double *)rhs, iteration_index, factor1, factor2, (double *)solution,
                                      (double *)solution_old);
    else
      DGH(h).chebyshev_update<float>((const float *)rhs, iteration_index, factor1, factor2, (float *)solution,
                                     (float *)solution_old);
    DG_CATCH
  }

  int
  mgb_dg_prolongate_add_cg_to_dg(mgb_dg h, mgb_solver s, int level, int precision, void *dg, const void *cg)
  {
    DG_TRY
    check_precision(precision);
    const LevelDesc L = *mgb_internal_level_desc(s, level);
    if (precision == MGB_F64)
      DGH(h).prolongate_add_cg_to_dg<double>(L, (double *)dg, (const double *)cg);
    else
      DGH(h).prolongate_add_cg_to_dg<float>(L, (float *)dg, (const float *)cg);
    DG_CATCH
  }

  int
  mgb_dg_vmult_residual_and_restrict_to_cg(mgb_dg h, mgb_solver s, int level, int precision, const void *rhs,
                                           const void *lhs, void *restricted_residual)
  {
    DG_TRY
    check_precision(precision);
    const LevelDesc L = *mgb_internal_level_desc(s, level);
    if (precision == MGB_F64)
      DGH(h).vmult_residual_and_restrict_to_cg<double>(L, (const double *)rhs, (const double *)lhs, (double *)restricted_residual);
    else
      DGH(h).vmult_residual_and_restrict_to_cg<float>(L, (const float *)rhs, (const float *)lhs, (float *)restricted_residual);
    DG_CATCH
  }

  int
  mgb_dg_set_preconditioner(mgb_dg h, int kind)
  {
    DG_TRY
    if (kind != 0 && kind != 1)
      throw std::runtime_error("preconditioner must be 0 (point Jacobi) or 1 (JacobiTransformed)");
    DGH(h).preconditioner = kind;
    if (kind == 1)
      DGH(h).build_jacobi_transformed();
    DG_CATCH
  }

  int
  mgb_dg_preconditioner_vmult(mgb_dg h, int precision, void *dst, const void *src)
  {
    DG_TRY
    check_precision(precision);
    if (precision == MGB_F64)
      DGH(h).preconditioner_vmult<double>((double *)dst, (const double *)src, 1.);
    else
      DGH(h).preconditioner_vmult<float>((float *)dst, (const float *)src, 1.);
    DG_CATCH
  }

  int
  mgb_dg_vmult_host(mgb_dg h, int precision, void *dst_host, const void *src_host)
  {
    DG_TRY
    check_precision(precision);
    DGOperator  &o     = DGH(h);
    const size_t bytes = o.n_dofs * (precision == MGB_F64 ? 8 : 4);
    for (int i = 0; i < 2; ++i)
      if (!o.host_buf[i])
        MGB_DG_CHECK(cudaMalloc(&o.host_buf[i], o.n_dofs * 8));
    MGB_DG_CHECK(cudaMemcpyAsync(o.host_buf[0], src_host, bytes, cudaMemcpyHostToDevice, o.stream));
    if (precision == MGB_F64)
      o.apply<double>(DG_SET, (double *)o.host_buf[1], (const double *)o.host_buf[0]);
    else
      o.apply<float>(DG_SET, (float *)o.host_buf[1], (const float *)o.host_buf[0]);
    MGB_DG_CHECK(cudaMemcpyAsync(dst_host, o.host_buf[1], bytes, cudaMemcpyDeviceToHost, o.stream));
    MGB_DG_CHECK(cudaStreamSynchronize(o.stream));
    DG_CATCH
  }

  int
  mgb_dg_timer_start(mgb_dg h)
  {
    DG_TRY
    MGB_DG_CHECK(cudaEventRecord(DGH(h).ev[0], DGH(h).stream));
    DG_CATCH
  }

  int
  mgb_dg_timer_stop(mgb_dg h, double *ms)
  {
    DG_TRY
    DGOperator &o = DGH(h);
    MGB_DG_CHECK(cudaEventRecord(o.ev[1], o.stream));
    MGB_DG_CHECK(cudaEventSynchronize(o.ev[1]));
    float t = 0;
    MGB_DG_CHECK(cudaEventElapsedTime(&t, o.ev[0], o.ev[1]));
    *ms = t;
    DG_CATCH
  }

  int
  mgb_dg_lumped_mass_inverse(mgb_dg h, int precision, void *out)
  {
    DG_TRY
    check_precision(precision);
    if (precision == MGB_F64)
      DGH(h).lumped_mass_inverse<double>((double *)out);
    else
      DGH(h).lumped_mass_inverse<float>((float *)out);
    DG_CATCH
  }

  int
  mgb_dg_solve_cg(mgb_dg h, int precision, unsigned int n_iterations, int interleaved, void *x, const void *b,
                  const void *diagonal_preconditioner, double stats[2])
  {
    DG_TRY
    check_precision(precision);
    if (!x || !b || !diagonal_preconditioner || !stats)
      throw std::runtime_error("null argument");
    if (precision == MGB_F64)
      DGH(h).solve_cg_diagonal<double>(n_iterations, interleaved != 0, (double *)x, (const double *)b,
                                       (const double *)diagonal_preconditioner, stats);
    else
      DGH(h).solve_cg_diagonal<float>(n_iterations, interleaved != 0, (float *)x, (const float *)b,
                                      (const float *)diagonal_preconditioner, stats);
    DG_CATCH
  }

  int64_t
  mgb_dg_launch_count(mgb_dg h)
  {
    return h && h->impl ? h->impl->launches : -1;
  }
  // ------------------------------------------------------------------ DG multigrid solver
  void
  mgb_dg_solver_default_options(mgb_dg_solver_options *o)
  {
    if (!o)
      return;
    std::memset(o, 0, sizeof(*o));
    o->degree = 4;
    for (int d = 0; d < 3; ++d)
      {
        o->n_subdivisions[d] = 1;
        o->lower[d]          = -1.; // poisson_dg/program.cc:357-376 (doubling mesh, remainder 0)
        o->upper[d]          = 1.;
      }
    o->n_refinements  = 3;
    o->degree_pre     = 3; // poisson_dg/program.cc:518-519
    o->degree_post    = 3;
    o->preconditioner = 1;
  }

  int
  mgb_dg_solver_create(const mgb_dg_solver_options *opt, mgb_dg_solver *out)
  {
    DG_TRY
    if (!opt || !out)
      throw std::runtime_error("null argument");
    *out      = nullptr;
    auto *h   = new mgb_dg_solver_t;
    h->impl   = nullptr;
    h->op.impl = nullptr;
    try
      {
        h->impl    = new DGMultigridSolver(*opt);
        h->op.impl = h->impl->op;
      }
    catch (...)
      {
        delete h;
        throw;
      }
    *out = h;
    DG_CATCH
  }

  int
  mgb_dg_solver_destroy(mgb_dg_solver h)
  {
    DG_TRY
    if (h)
      {
        delete h->impl;
        delete h;
      }
    DG_CATCH
  }

  int64_t
  mgb_dg_solver_n_dofs(mgb_dg_solver h)
  {
    return h && h->impl ? (int64_t)h->impl->N : -1;
  }

  int
  mgb_dg_solver_parts(mgb_dg_solver h, mgb_solver *fe, mgb_dg *dg)
  {
    DG_TRY
    DGMultigridSolver &m = DGS(h);
    if (fe)
      *fe = m.fe;
    if (dg)
      *dg = &h->op;
    DG_CATCH
  }

  int
  mgb_dg_solver_solve_cg(mgb_dg_solver h, double tolerance, double *its_estimate, double *rate, unsigned int *its)
  {
    DG_TRY
    if (!(tolerance > 0 && tolerance < 1))
      throw std::runtime_error("tolerance must be in (0, 1)");
    DGS(h).solve_cg(tolerance, its_estimate, rate, its);
    DG_CATCH
  }

  int
  mgb_dg_solver_residual_history(mgb_dg_solver h, double *history, int capacity, int *count)
  {
    DG_TRY
    DGMultigridSolver &m = DGS(h);
    const int          k = (int)m.residual_history.size();
    for (int i = 0; i < std::min(k, capacity); ++i)
      history[i] = m.residual_history[i];
    if (count)
      *count = k;
    DG_CATCH
  }

  int
  mgb_dg_solver_vmult(mgb_dg_solver h, double *dst, const double *src)
  {
    DG_TRY
    DGS(h).vmult(dst, src);
    DG_CATCH
  }

  int
  mgb_dg_solver_vmult_with_residual_update(mgb_dg_solver h, double *residual, double *update, double factor, double results[2])
  {
    DG_TRY
    DGS(h).vmult_with_residual_update(residual, update, factor, results);
    DG_CATCH
  }

  int
  mgb_dg_solver_l2_error(mgb_dg_solver h, double *error)
  {
    DG_TRY
    *error = DGS(h).compute_l2_error();
    DG_CATCH
  }

  int
  mgb_dg_solver_vector(mgb_dg_solver h, int which, void **dev_ptr)
  {
    DG_TRY
    DGMultigridSolver &m = DGS(h);
    switch (which)
      {
        case 0: *dev_ptr = m.solution; break;
        case 1: *dev_ptr = m.rhs; break;
        case 2: *dev_ptr = m.g; break;
        case 3: *dev_ptr = m.defect_dg; break;
        case 4: *dev_ptr = m.solution_update_dg; break;
        default: throw std::runtime_error("unknown vector id");
      }
    DG_CATCH
  }

  int
  mgb_dg_solver_get_solution(mgb_dg_solver h, double *host)
  {
    DG_TRY
    DGMultigridSolver &m = DGS(h);
    MGB_DG_CHECK(cudaMemcpyAsync(host, m.solution, m.N * sizeof(double), cudaMemcpyDeviceToHost, m.op->stream));
    MGB_DG_CHECK(cudaStreamSynchronize(m.op->stream));
    DG_CATCH
  }

  int
  mgb_dg_solver_smoother_info(mgb_dg_solver h, double info[6])
  {
    DG_TRY
    const DGChebyshev &c = DGS(h).smooth_dg;
    info[0]              = c.min_eig;
    info[1]              = c.max_eig;
    info[2]              = c.theta;
    info[3]              = c.delta;
    info[4]              = c.degree;
    info[5]              = c.cg_iterations;
    DG_CATCH
  }

  int
  mgb_dg_solver_do_matvec(mgb_dg_solver h)
  {
    DG_TRY
    DGMultigridSolver &m = DGS(h);
    m.op->apply<double>(DG_SET, m.g, m.solution);
    DG_CATCH
  }

  int
  mgb_dg_solver_do_matvec_smoother(mgb_dg_solver h)
  {
    DG_TRY
    DGMultigridSolver &m = DGS(h);
    m.op->apply<float>(DG_SET, m.solution_update_dg, m.defect_dg);
    DG_CATCH
  }

  int
  mgb_dg_solver_sync(mgb_dg_solver h)
  {
    DG_TRY
    MGB_DG_CHECK(cudaStreamSynchronize(DGS(h).op->stream));
    DG_CATCH
  }

  int
  mgb_dg_solver_solve_cg_host(mgb_dg_solver h, double tolerance, double *solution_host, double *its_estimate, double *rate,
                              unsigned int *its)
  {
    DG_TRY
    if (!(tolerance > 0 && tolerance < 1))
      throw std::runtime_error("tolerance must be in (0, 1)");
    DGMultigridSolver &m = DGS(h);
    m.solve_cg(tolerance, its_estimate, rate, its);
    if (solution_host)
      MGB_DG_CHECK(cudaMemcpyAsync(solution_host, m.solution, m.N * sizeof(double), cudaMemcpyDeviceToHost, m.op->stream));
    MGB_DG_CHECK(cudaStreamSynchronize(m.op->stream));
    DG_CATCH
  }

  int64_t
  mgb_dg_solver_launch_count(mgb_dg_solver h)
  {
    if (!h || !h->impl)
      return -1;
    return h->impl->op->launches + mgb_solver_launch_count(h->impl->fe);
  }

  // ------------------------------------------------------------------ pure DG h-multigrid solver
  int
  mgb_dg_plain_solver_create(const mgb_dg_solver_options *opt, mgb_dg_plain_solver *out)
  {
    DG_TRY
    if (!opt || !out)
      throw std::runtime_error("null argument");
    *out    = nullptr;
    auto *h = new mgb_dg_plain_solver_t;
    h->impl = nullptr;
    try
      {
        h->impl = new DGPlainSolver(*opt);
        for (auto &o : h->impl->ops)
          h->ops.push_back(mgb_dg_t{o.get()});
      }
    catch (...)
      {
        delete h;
        throw;
      }
    *out = h;
    DG_CATCH
  }

  int
  mgb_dg_plain_solver_destroy(mgb_dg_plain_solver h)
  {
    DG_TRY
    if (h)
      {
        delete h->impl;
        delete h;
      }
    DG_CATCH
  }

  int64_t
  mgb_dg_plain_solver_n_dofs(mgb_dg_plain_solver h)
  {
    return h && h->impl ? (int64_t)h->impl->N : -1;
  }

  int
  mgb_dg_plain_solver_n_levels(mgb_dg_plain_solver h)
  {
    return h && h->impl ? h->impl->n_levels : -1;
  }

  int
  mgb_dg_plain_solver_level_operator(mgb_dg_plain_solver h, int level, mgb_dg *dg)
  {
    DG_TRY
    *dg = &h->ops[check_level(DGP(h), level)];
    DG_CATCH
  }

  int
  mgb_dg_plain_solver_solve_cg(mgb_dg_plain_solver h, double tolerance, double *its_estimate, double *rate, unsigned int *its)
  {
    DG_TRY
    if (!(tolerance > 0 && tolerance < 1))
      throw std::runtime_error("tolerance must be in (0, 1)");
    DGP(h).solve_cg(tolerance, its_estimate, rate, its);
    DG_CATCH
  }

  int
  mgb_dg_plain_solver_residual_history(mgb_dg_plain_solver h, double *history, int capacity, int *count)
  {
    DG_TRY
    DGPlainSolver &m = DGP(h);
    const int      k = (int)m.residual_history.size();
    for (int i = 0; i < std::min(k, capacity); ++i)
      history[i] = m.residual_history[i];
    if (count)
      *count = k;
    DG_CATCH
  }

  int
  mgb_dg_plain_solver_vmult(mgb_dg_plain_solver h, double *dst, const double *src)
  {
    DG_TRY
    DGP(h).vmult(dst, src);
    DG_CATCH
  }

  int
  mgb_dg_plain_solver_vmult_with_residual_update(mgb_dg_plain_solver h, double *residual, double *update, double factor,
                                                 double results[2])
  {
    DG_TRY
    DGP(h).vmult_with_residual_update(residual, update, factor, results);
    DG_CATCH
  }

  int
  mgb_dg_plain_solver_l2_error(mgb_dg_plain_solver h, double *error)
  {
    DG_TRY
    *error = DGP(h).compute_l2_error();
    DG_CATCH
  }

  int
  mgb_dg_plain_solver_vector(mgb_dg_plain_solver h, int which, int level, void **dev_ptr)
  {
    DG_TRY
    DGPlainSolver &m = DGP(h);
    switch (which)
      {
        case 0: *dev_ptr = m.solution; break;
        case 1: *dev_ptr = m.rhs; break;
        case 2: *dev_ptr = m.g; break;
        case 3: *dev_ptr = m.defect[check_level(m, level)]; break;
        case 4: *dev_ptr = m.solution_update[check_level(m, level)]; break;
        case 5: *dev_ptr = m.t[check_level(m, level)]; break;
        default: throw std::runtime_error("unknown vector id");
      }
    DG_CATCH
  }

  int
  mgb_dg_plain_solver_get_vector(mgb_dg_plain_solver h, int which, int level, void *host)
  {
    DG_TRY
    DGPlainSolver &m   = DGP(h);
    void          *dev = nullptr;
    if (mgb_dg_plain_solver_vector(h, which, level, &dev) != 0)
      throw std::runtime_error(mgb_last_error());
    const size_t bytes = which < 3 ? m.N * sizeof(double) : m.ops[level]->n_dofs * sizeof(float);
    MGB_DG_CHECK(cudaMemcpyAsync(host, dev, bytes, cudaMemcpyDeviceToHost, m.stream));
    MGB_DG_CHECK(cudaStreamSynchronize(m.stream));
    DG_CATCH
  }

  int
  mgb_dg_plain_solver_get_solution(mgb_dg_plain_solver h, double *host)
  {
    DG_TRY
    DGPlainSolver &m = DGP(h);
    MGB_DG_CHECK(cudaMemcpyAsync(host, m.solution, m.N * sizeof(double), cudaMemcpyDeviceToHost, m.stream));
    MGB_DG_CHECK(cudaStreamSynchronize(m.stream));
    DG_CATCH
  }

  int
  mgb_dg_plain_solver_smoother_info(mgb_dg_plain_solver h, int level, double info[6])
  {
    DG_TRY
    DGPlainSolver     &m = DGP(h);
    const DGChebyshev &c = m.smooth[check_level(m, level)];
    info[0]              = c.min_eig;
    info[1]              = c.max_eig;
    info[2]              = c.theta;
    info[3]              = c.delta;
    info[4]              = c.degree;
    info[5]              = c.cg_iterations;
    DG_CATCH
  }

  int
  mgb_dg_plain_solver_prolongate(mgb_dg_plain_solver h, int to_level, float *dst, const float *src, int add)
  {
    DG_TRY
    DGPlainSolver &m = DGP(h);
    if (check_level(m, to_level) < 1)
      throw std::runtime_error("no level below level 0");
    if (add)
      m.transfer<false, true>(to_level, dst, src);
    else
      m.transfer<false, false>(to_level, dst, src);
    DG_CATCH
  }

  int
  mgb_dg_plain_solver_restrict(mgb_dg_plain_solver h, int from_level, float *dst, const float *src, int add)
  {
    DG_TRY
    DGPlainSolver &m = DGP(h);
    if (check_level(m, from_level) < 1)
      throw std::runtime_error("no level below level 0");
    if (add)
      m.transfer<true, true>(from_level, dst, src);
    else
      m.transfer<true, false>(from_level, dst, src);
    DG_CATCH
  }

  int
  mgb_dg_plain_solver_do_matvec(mgb_dg_plain_solver h)
  {
    DG_TRY
    DGPlainSolver &m = DGP(h);
    m.ops.back()->apply<double>(DG_SET, m.g, m.solution);
    DG_CATCH
  }

  int
  mgb_dg_plain_solver_do_matvec_smoother(mgb_dg_plain_solver h)
  {
    DG_TRY
    DGPlainSolver &m = DGP(h);
    m.ops.back()->apply<float>(DG_SET, m.solution_update.back(), m.defect.back());
    DG_CATCH
  }

  int
  mgb_dg_plain_solver_sync(mgb_dg_plain_solver h)
  {
    DG_TRY
    MGB_DG_CHECK(cudaStreamSynchronize(DGP(h).stream));
    DG_CATCH
  }

  int
  mgb_dg_plain_solver_solve_cg_host(mgb_dg_plain_solver h, double tolerance, double *solution_host, double *its_estimate,
                                    double *rate, unsigned int *its)
  {
    DG_TRY
    if (!(tolerance > 0 && tolerance < 1))
      throw std::runtime_error("tolerance must be in (0, 1)");
    DGPlainSolver &m = DGP(h);
    m.solve_cg(tolerance, its_estimate, rate, its);
    if (solution_host)
      MGB_DG_CHECK(cudaMemcpyAsync(solution_host, m.solution, m.N * sizeof(double), cudaMemcpyDeviceToHost, m.stream));
    MGB_DG_CHECK(cudaStreamSynchronize(m.stream));
    DG_CATCH
  }

  int64_t
  mgb_dg_plain_solver_launch_count(mgb_dg_plain_solver h)
  {
    return h && h->impl ? h->impl->launch_count() : -1;
  }
}
