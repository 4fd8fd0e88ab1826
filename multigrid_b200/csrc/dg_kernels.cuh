// Symmetric interior-penalty DG Laplacian on an affine box mesh: one kernel computes the cell integral and
// all six face integrals of a cell and the fused vector operation.
//
// Reference: multigrid::LaplaceOperatorCompactCombine::operation_on_cells<action>
// (common/laplace_operator_dg.h:1110-1861; specification common/laplace_operator_dg_face.h:58-160).
// For a cell K with neighbour N across a face with K's outward normal n, jump j = u_K - u_N and
// m = dn u_K + dn u_N (boundary faces: mirror principle u_N = -u_K, dn u_N = dn u_K, sigma_b = 2 sigma):
//     y_i += int_f  phi_i (sigma j - m/2)  -  dn phi_i  j/2
// which is the contribution of the face to the rows of K of  sigma [u][v] - {dn u}[v] - {dn v}[u].
// The metric is constant on the mesh, dn u = +-sum_e gs[d][e] d_e u with the reference derivatives d_e.
//
// Mapping: a cell is handled by (p+1)^2 threads (one warp for p <= 4); thread (a,b) owns the pencil of
// p+1 values along the sweep direction.  Cell term: collocation at the Gauss points (interpolate S,
// derivative D, metric, D^T, S^T: 12 sweeps).  Face terms per direction d, both faces at once, on the pencil
// along d in the collocation basis: own trace and normal reference derivative by 1-D sums, the neighbour's
// (its p+1 layers are read from global memory, i.e. L2) by a 2-D interpolation of two face arrays per side,
// flux at the face Gauss points, and the test side as a rank-one update of the residual pencil before the
// S^T sweeps (see k_dg_apply).
// DoFs are cell-wise contiguous, so all global accesses of a cell are one contiguous chunk.
#pragma once
#include <cuda_runtime.h>

#include <cstdint>
#include <cstdlib>

namespace mgb
{
  enum DGMode
  {
    DG_SET = 0,             // dst = A src                                   (action 0)
    DG_RESIDUAL = 1,        // dst = rhs - A src                             (action 4)
    DG_CHEBYSHEV = 2,       // dst = (1+f1) src - f1 dst + f2 P^-1 (rhs - A src)   (action 3, iteration_index > 1)
    DG_CHEBYSHEV_FIRST = 3, // dst = (1+f1) src + f2 P^-1 (rhs - A src)            (action 3, iteration_index == 1)
    DG_SET_SUMS = 4         // dst = v = A src and, per block, the five sums of the interleaved CG with the diagonal
                            // preconditioner D = inv_diag and the residual r = rhs (solver_dg/program.cc:39-70: the
                            // operation_after_loop of deal.II's SolverCG): src.v, v.D r, v.D v, r.v, v.v -> partial[block][5]
  };

  template <typename T, int n>
  struct DGTables
  {
    T S[n * n]; // S[q n + i] = phi_i(x_q)
    T G[n * n]; // phi_i'(x_q)
    T D[n * n]; // D[q n + r] = l_r'(x_q)
    T e0[n], e1[n]; // phi_i'(0), phi_i'(1)
    T f0[n], f1[n]; // phi_i(0), phi_i(1): unit vectors for the Gauss-Lobatto basis (kernel variant KIND = 0)
    T qf0[n], qf1[n], qe0[n], qe1[n]; // the same for the Lagrange polynomials in the Gauss points (collocation basis)
    T w[n];
    T V[n * n]; // V[k n + j] = component j of the k-th generalised eigenvector of (1-D SIP Laplace, 1-D mass)
  };

  template <typename T>
  struct DGArgs
  {
    int      nc[3];      // cells of the whole mesh (all ranks)
    uint32_t n_cells;    // cells this launch works on: the local cells cell_begin .. cell_begin + n_cells - 1
    uint32_t cell_begin; // z-slabs: the owned layers follow a ghost layer of the rank below (0 on one GPU)
    int      cz_shift;   // z-position of the first LOCAL cell layer in the whole mesh
    const T *src;
    T       *dst;
    const T *rhs;
    const T *inv_diag;
    const T *jt_dinv;  // JacobiTransformed: 1 / diag(V^T A_KK V) per cell type [64][n^3]; nullptr: point Jacobi
    T        f1, f2;
    int      mode;
    double  *partial;  // DG_SET_SUMS: [blocks][5]
    T        gt[6];    // detJ J^-1 J^-T: 00 11 22 01 02 12
    T        gs[3][3]; // gs[d][e] = grad xi_d . grad xi_e / |grad xi_d|
    T        sigma[3]; // (p+1)^2 |grad xi_d|
    T        area[3];  // detJ |grad xi_d|
  };

  template <int p>
  struct DGCfg
  {
    static constexpr int n   = p + 1;
    static constexpr int n2  = n * n;
    static constexpr int n3  = n * n * n;
    static constexpr int TPC = ((n2 + 31) / 32) * 32;        // threads per cell
    static constexpr int CPB = TPC >= 128 ? 1 : 128 / TPC;   // cells per block
    static constexpr int NT  = TPC * CPB;
    static constexpr int PER_CELL = 5 * n3 + 14 * n2;        // shared-memory elements per cell
    static constexpr int TABLES   = 2 * n2 + n;              // S, G, w indexed by the lane (block-wide copy)
  };

  template <int TPC>
  __device__ __forceinline__ void
  cell_sync(const int slot)
  {
    if (TPC == 32)
      __syncwarp();
    else
      asm volatile("bar.sync %0, %1;" ::"r"(slot + 1), "n"(TPC) : "memory");
  }

  // out[r] = sum_q M[q n + r] in[q]   (transposed product with a [q][r] table)
  template <int n, typename T>
  __device__ __forceinline__ void
  tprod(const T *__restrict__ M, const T (&in)[n], T (&out)[n])
  {
#pragma unroll
    for (int r = 0; r < n; ++r)
      {
        T s = 0;
#pragma unroll
        for (int q = 0; q < n; ++q)
          s += M[q * n + r] * in[q];
        out[r] = s;
      }
  }
  // out[q] = sum_r M[q n + r] in[r]
  template <int n, typename T>
  __device__ __forceinline__ void
  prod(const T *__restrict__ M, const T (&in)[n], T (&out)[n])
  {
#pragma unroll
    for (int q = 0; q < n; ++q)
      {
        T s = 0;
#pragma unroll
        for (int r = 0; r < n; ++r)
          s += M[q * n + r] * in[r];
        out[q] = s;
      }
  }

  // type of a cell for the cell-local matrix A_KK: per direction 0 interior, 1 low boundary, 2 high, 3 both
  __device__ __forceinline__ int
  dg_cell_type(const int cx, const int cy, const int cz, const int *nc)
  {
    const int tx = (cx == 0 ? 1 : 0) + (cx == nc[0] - 1 ? 2 : 0);
    const int ty = (cy == 0 ? 1 : 0) + (cy == nc[1] - 1 ? 2 : 0);
    const int tz = (cz == 0 ? 1 : 0) + (cz == nc[2] - 1 ? 2 : 0);
    return (tz * 4 + ty) * 4 + tx;
  }

  // multigrid::JacobiTransformed::do_local_operation (laplace_operator_dg.h:2085-2095): r <- V (D^-1 (V^T r)) with
  // the tensor-product transform V x V x V and D = diag(V^T A_KK V) of the cell's type.  `r` holds the input and
  // the result (n^3 values in shared memory), `a0`, `a1` are work arrays; thread (a,b) sweeps pencils.
  template <int n, int TPC, typename T>
  __device__ __forceinline__ void
  jacobi_transformed_local(T *r, T *a0, T *a1, const T *__restrict__ V, const T *__restrict__ dinv, const bool active,
                           const int a, const int b, const int slot)
  {
    constexpr int n2 = n * n;
    T             in[n], out[n];
    if (active)
      { // V^T along z: out[k] = sum_j V[k][j] in[j]
#pragma unroll
        for (int k = 0; k < n; ++k)
          in[k] = r[k * n2 + b * n + a];
        prod<n>(V, in, out);
#pragma unroll
        for (int k = 0; k < n; ++k)
          a0[k * n2 + b * n + a] = out[k];
      }
    cell_sync<TPC>(slot);
    if (active)
      { // along y
#pragma unroll
        for (int j = 0; j < n; ++j)
          in[j] = a0[(b * n + j) * n + a];
        prod<n>(V, in, out);
#pragma unroll
        for (int j = 0; j < n; ++j)
          a1[(b * n + j) * n + a] = out[j];
      }
    cell_sync<TPC>(slot);
    if (active)
      { // along x: to the eigenbasis, scale, and back along x
#pragma unroll
        for (int i = 0; i < n; ++i)
          in[i] = a1[(b * n + a) * n + i];
        prod<n>(V, in, out);
#pragma unroll
        for (int i = 0; i < n; ++i)
          out[i] *= dinv[(b * n + a) * n + i];
        tprod<n>(V, out, in);
#pragma unroll
        for (int i = 0; i < n; ++i)
          a0[(b * n + a) * n + i] = in[i];
      }
    cell_sync<TPC>(slot);
    if (active)
      { // V along y
#pragma unroll
        for (int j = 0; j < n; ++j)
          in[j] = a0[(b * n + j) * n + a];
        tprod<n>(V, in, out);
#pragma unroll
        for (int j = 0; j < n; ++j)
          a1[(b * n + j) * n + a] = out[j];
      }
    cell_sync<TPC>(slot);
    if (active)
      { // V along z
#pragma unroll
        for (int k = 0; k < n; ++k)
          in[k] = a1[k * n2 + b * n + a];
        tprod<n>(V, in, out);
#pragma unroll
        for (int k = 0; k < n; ++k)
          r[k * n2 + b * n + a] = out[k];
      }
    cell_sync<TPC>(slot);
  }

  // JacobiTransformed::vmult (laplace_operator_dg.h:2046-2083): dst = scale * P^-1 src, one cell per thread group
  template <int p, typename T>
  __global__ void
  __launch_bounds__(DGCfg<p>::NT)
    k_dg_jacobi_transformed(const DGArgs<T> A, const DGTables<T, p + 1> Tb, const T scale)
  {
    using C           = DGCfg<p>;
    constexpr int n   = C::n;
    constexpr int n2  = C::n2;
    constexpr int n3  = C::n3;
    constexpr int TPC = C::TPC;
    __shared__ T buf[C::CPB * 3 * n3];
    const int    slot = threadIdx.x / TPC, t = threadIdx.x % TPC;
    T           *r = buf + slot * 3 * n3, *a0 = r + n3, *a1 = a0 + n3;
    const uint32_t cell   = A.cell_begin + blockIdx.x * C::CPB + slot;
    const bool     valid  = cell < A.cell_begin + A.n_cells;
    const bool     active = valid && t < n2;
    const uint32_t c  = valid ? cell : 0;
    const int      cx = c % A.nc[0], cy = (c / A.nc[0]) % A.nc[1], cz = c / (A.nc[0] * A.nc[1]) + A.cz_shift;
    if (valid)
      for (int i = t; i < n3; i += TPC)
        r[i] = A.src[(size_t)c * n3 + i];
    cell_sync<TPC>(slot);
    jacobi_transformed_local<n, TPC, T>(r, a0, a1, Tb.V, A.jt_dinv + dg_cell_type(cx, cy, cz, A.nc) * n3, active, t % n,
                                        t / n, slot);
    if (valid)
      for (int i = t; i < n3; i += TPC)
        A.dst[(size_t)c * n3 + i] = scale * r[i];
  }

  // KIND selects the 1-D basis of the vectors: 0 FE_DGQ on the Gauss-Lobatto points, 1 any tensor-product basis through
  // its tables (FE_DGQHermite), 2 the basis collocated with the Gauss points (FE_DGQArbitraryNodes(QGauss), S = I: the
  // interpolation sweeps drop out, laplace_operator_dg.h:1179).
  // Everything
  // between the first interpolation and the last S^T sweeps happens in the collocation basis of the Gauss points, as in
  // the reference (laplace_operator_dg.h:1143-1790): the trace and the normal reference derivative of the OWN cell on a
  // face of direction d are 1-D sums along d of the values at the Gauss points (tables qf / qe), taken by the thread that
  // owns the pencil along d, and the test-function side is the transpose: the face fluxes are added to the residual at the
  // Gauss points of the pencil together with D_d^T of the cell flux, so that the final S^T sweeps of the cell term carry
  // them back to the basis of the vectors.  Only the NEIGHBOUR's trace and normal derivative (two face arrays per side,
  // given in the neighbour's nodal basis) go through a 2-D interpolation to the face Gauss points; the tangential
  // derivatives (non-orthogonal metric only) are collocation derivatives of face arrays.  Shared-memory accesses per
  // thread and direction: about 60 (orthogonal metric) / 105 (sheared); the first version of this kernel, which took the
  // traces in the basis of the vectors and interpolated jump / derivative / trace arrays of both sides in 2-D, needed about
  // 150 (measured on the 131 M DoF sheared mesh, double vmult: 5.84 -> 4.52 ms, profiles/r2_dg_timings.txt).
  template <int p, typename T, int KIND>
  __global__ void
  __launch_bounds__(DGCfg<p>::NT) k_dg_apply(const DGArgs<T> A, const DGTables<T, p + 1> Tb)
  {
    constexpr bool COLLOC = KIND == 2;
    using C           = DGCfg<p>;
    constexpr int n   = C::n;
    constexpr int n2  = C::n2;
    constexpr int n3  = C::n3;
    constexpr int TPC = C::TPC;
    extern __shared__ __align__(16) unsigned char dg_smem_raw[];
    const int slot = threadIdx.x / TPC, t = threadIdx.x % TPC;
    T *S_s = reinterpret_cast<T *>(dg_smem_raw), *D_s = S_s + n2, *w_s = D_s + n2;
    for (int i = threadIdx.x; i < n2; i += C::NT)
      {
        S_s[i] = Tb.S[i];
        D_s[i] = Tb.D[i];
      }
    if (threadIdx.x < n)
      w_s[threadIdx.x] = Tb.w[threadIdx.x];
    T *u_s = reinterpret_cast<T *>(dg_smem_raw) + C::TABLES + slot * C::PER_CELL;
    T *y_s = u_s + n3;
    T *w0  = y_s + n3;
    T *w1  = w0 + n3;
    T *w2  = w1 + n3;
    T *F   = w2 + n3;     // face arrays, 4 n^2 used
    T *H   = F + 6 * n2;  // face arrays, 4 n^2 used

    const uint32_t cell   = A.cell_begin + blockIdx.x * C::CPB + slot;
    const bool     valid  = cell < A.cell_begin + A.n_cells;
    const bool     active = valid && t < n2;
    const int      a = t % n, b = t / n;
    const uint32_t c  = valid ? cell : 0;
    const int      cx = c % A.nc[0], cy = (c / A.nc[0]) % A.nc[1], cz = c / (A.nc[0] * A.nc[1]) + A.cz_shift;
    const T *__restrict__ src = A.src;

    if (valid)
      for (int i = t; i < n3; i += TPC)
        u_s[i] = src[(size_t)c * n3 + i];
    __syncthreads(); // tables visible to all cells of the block

    // ---- values at the Gauss points: uq
    T        in[n], out[n];
    const T *uq = COLLOC ? u_s : w0;
    if (!COLLOC)
      {
        if (active)
          { // along z, thread (i,j)
#pragma unroll
            for (int k = 0; k < n; ++k)
              in[k] = u_s[k * n2 + b * n + a];
            prod<n>(Tb.S, in, out);
#pragma unroll
            for (int k = 0; k < n; ++k)
              w0[k * n2 + b * n + a] = out[k];
          }
        cell_sync<TPC>(slot);
        if (active)
          { // along y, thread (i,k)
#pragma unroll
            for (int j = 0; j < n; ++j)
              in[j] = w0[(b * n + j) * n + a];
            prod<n>(Tb.S, in, out);
#pragma unroll
            for (int j = 0; j < n; ++j)
              w1[(b * n + j) * n + a] = out[j];
          }
        cell_sync<TPC>(slot);
        if (active)
          { // along x, thread (j,k)
#pragma unroll
            for (int i = 0; i < n; ++i)
              in[i] = w1[(b * n + a) * n + i];
            prod<n>(Tb.S, in, out);
#pragma unroll
            for (int i = 0; i < n; ++i)
              w0[(b * n + a) * n + i] = out[i];
          }
        cell_sync<TPC>(slot);
      }
    // ---- reference gradient at the Gauss points: gx -> w1, gy -> w2, gz and the own z-pencil in registers of thread (i,j)
    T gz[n], uz[n];
    if (active)
      {
#pragma unroll
        for (int i = 0; i < n; ++i)
          in[i] = uq[(b * n + a) * n + i];
        prod<n>(Tb.D, in, out);
#pragma unroll
        for (int i = 0; i < n; ++i)
          w1[(b * n + a) * n + i] = out[i];
#pragma unroll
        for (int j = 0; j < n; ++j)
          in[j] = uq[(b * n + j) * n + a];
        prod<n>(Tb.D, in, out);
#pragma unroll
        for (int j = 0; j < n; ++j)
          w2[(b * n + j) * n + a] = out[j];
#pragma unroll
        for (int k = 0; k < n; ++k)
          uz[k] = uq[k * n2 + b * n + a];
        prod<n>(Tb.D, uz, gz);
      }
    cell_sync<TPC>(slot);
    // ---- metric at the points (a, b, k): fluxes fx -> w1, fy -> w2, fz in registers
    T fz[n];
    if (active)
      {
#pragma unroll
        for (int k = 0; k < n; ++k)
          {
            const int idx = k * n2 + b * n + a;
            const T   W   = w_s[a] * w_s[b] * Tb.w[k];
            const T   gx = w1[idx], gy = w2[idx], g2 = gz[k];
            w1[idx]       = W * (A.gt[0] * gx + A.gt[3] * gy + A.gt[4] * g2);
            w2[idx]       = W * (A.gt[3] * gx + A.gt[1] * gy + A.gt[5] * g2);
            fz[k]         = W * (A.gt[4] * gx + A.gt[5] * gy + A.gt[2] * g2);
          }
      }
    cell_sync<TPC>(slot);

    // ---- direction by direction (z, y, x): D_d^T of the cell flux plus the two faces, on the pencil along d
#pragma unroll
    for (int d = 2; d >= 0; --d)
      {
        const int  base = d == 2 ? b * n + a : (d == 1 ? b * n2 + a : (b * n + a) * n);
        const int  sd   = d == 2 ? n2 : (d == 1 ? n : 1);
        const int  da = d == 0 ? 1 : 0, db = d == 2 ? 1 : 2; // tangential directions of a and b
        const T    g_dd = A.gs[d][d], g_da = A.gs[d][da], g_db = A.gs[d][db];
        const bool sheared = g_da != T(0) || g_db != T(0);   // kernel-uniform
        const int  cd   = d == 0 ? cx : (d == 1 ? cy : cz);
        const long cstride = d == 0 ? 1 : (d == 1 ? A.nc[0] : (long)A.nc[0] * A.nc[1]);
        const bool inter[2] = {cd > 0, cd + 1 < A.nc[d]};
        T          own[n], res[n];
        T          tK[2] = {0, 0}, dK[2] = {0, 0}, tN[2] = {0, 0}, dN[2] = {0, 0};
        if (active)
          {
            // own pencil of values at the Gauss points, and D_d^T of the flux pencil
            if (d == 2)
              {
#pragma unroll
                for (int l = 0; l < n; ++l)
                  own[l] = uz[l];
                tprod<n>(Tb.D, fz, res);
              }
            else
              {
                const T *fl = d == 1 ? w2 : w1;
#pragma unroll
                for (int l = 0; l < n; ++l)
                  {
                    own[l] = uq[base + l * sd];
                    in[l]  = fl[base + l * sd];
                  }
                tprod<n>(Tb.D, in, res);
              }
#pragma unroll
            for (int l = 0; l < n; ++l)
              {
                tK[0] += Tb.qf0[l] * own[l];
                tK[1] += Tb.qf1[l] * own[l];
                dK[0] += Tb.qe0[l] * own[l];
                dK[1] += Tb.qe1[l] * own[l];
              }
            // neighbour: trace and normal reference derivative at the NODAL tangential index (a, b)
#pragma unroll
            for (int s = 0; s < 2; ++s)
              {
                T tn = 0, dn = 0;
                if (inter[s])
                  {
                    const T *eN = s ? Tb.e0 : Tb.e1, *fN = s ? Tb.f0 : Tb.f1;
                    const T *nb = src + (size_t)((long)c + (s ? cstride : -cstride)) * n3 + base;
#pragma unroll
                    for (int l = 0; l < n; ++l)
                      {
                        if (fN[l] == T(0) && eN[l] == T(0))
                          continue; // FE_DGQHermite: two layers of the neighbour (laplace_operator_dg.h:565,1015)
                        const T v = __ldg(nb + l * sd);
                        dn += eN[l] * v;
                        tn += fN[l] * v;
                      }
                  }
                if (COLLOC)
                  {
                    tN[s] = tn;
                    dN[s] = dn;
                  }
                else
                  {
                    F[(2 * s + 0) * n2 + t] = tn;
                    F[(2 * s + 1) * n2 + t] = dn;
                  }
              }
          }
        if (!COLLOC)
          {
            cell_sync<TPC>(slot);
            if (active)
              { // to the Gauss points along a: thread (qa, b)
#pragma unroll
                for (int m = 0; m < 4; ++m)
                  {
                    T v = 0;
#pragma unroll
                    for (int i = 0; i < n; ++i)
                      v += S_s[a * n + i] * F[m * n2 + b * n + i];
                    H[m * n2 + t] = v;
                  }
              }
            cell_sync<TPC>(slot);
            if (active)
              { // along b: thread (qa, qb)
                T v[4] = {0, 0, 0, 0};
#pragma unroll
                for (int i = 0; i < n; ++i)
                  {
                    const T Sv = S_s[b * n + i];
#pragma unroll
                    for (int m = 0; m < 4; ++m)
                      v[m] += Sv * H[m * n2 + i * n + a];
                  }
                tN[0] = v[0];
                dN[0] = v[1];
                tN[1] = v[2];
                dN[1] = v[3];
              }
          }
        // jump, sum of the normal reference derivatives along d, sum of the traces (for the tangential derivatives);
        // boundary faces: mirror principle (laplace_operator_dg_face.h:131-160)
        T jump[2], dsum[2], tsum[2], T1[2] = {0, 0}, T2[2] = {0, 0};
#pragma unroll
        for (int s = 0; s < 2; ++s)
          {
            jump[s] = inter[s] ? tK[s] - tN[s] : 2 * tK[s];
            dsum[s] = inter[s] ? dK[s] + dN[s] : 2 * dK[s];
            tsum[s] = inter[s] ? tK[s] + tN[s] : 2 * tK[s];
          }
        if (sheared)
          {
            if (active)
              {
                F[0 * n2 + t] = tsum[0];
                F[1 * n2 + t] = tsum[1];
              }
            cell_sync<TPC>(slot);
            if (active)
#pragma unroll
              for (int i = 0; i < n; ++i)
                {
                  const T Da = D_s[a * n + i], Db = D_s[b * n + i];
#pragma unroll
                  for (int s = 0; s < 2; ++s)
                    {
                      T1[s] += Da * F[s * n2 + b * n + i];
                      T2[s] += Db * F[s * n2 + i * n + a];
                    }
                }
          }
        T V[2], N[2];
        {
          const T wJ = A.area[d] * w_s[active ? a : 0] * w_s[active ? b : 0];
#pragma unroll
          for (int s = 0; s < 2; ++s)
            {
              const T sgn = s ? T(1) : T(-1);
              const T m   = sgn * (g_dd * dsum[s] + g_da * T1[s] + g_db * T2[s]);
              V[s]        = (A.sigma[d] * jump[s] - T(0.5) * m) * wJ; // tested with the value of the test function
              N[s]        = -T(0.5) * jump[s] * wJ * sgn;             // tested with its normal derivative
            }
        }
        if (sheared)
          {
            // tangential parts of the tested normal derivative: collocation derivatives of N, transposed
            if (active)
              {
                H[0 * n2 + t] = N[0];
                H[1 * n2 + t] = N[1];
              }
            cell_sync<TPC>(slot);
            if (active)
#pragma unroll
              for (int q = 0; q < n; ++q)
                {
                  const T Da = D_s[q * n + a], Db = D_s[q * n + b];
#pragma unroll
                  for (int s = 0; s < 2; ++s)
                    V[s] += g_da * Da * H[s * n2 + b * n + q] + g_db * Db * H[s * n2 + q * n + a];
                }
          }
        if (active)
          {
#pragma unroll
            for (int l = 0; l < n; ++l)
              res[l] += Tb.qf0[l] * V[0] + Tb.qf1[l] * V[1] + g_dd * (Tb.qe0[l] * N[0] + Tb.qe1[l] * N[1]);
            if (d == 2)
              {
#pragma unroll
                for (int l = 0; l < n; ++l)
                  y_s[base + l * sd] = res[l];
              }
            else if (d == 1)
              {
#pragma unroll
                for (int l = 0; l < n; ++l)
                  y_s[base + l * sd] += res[l];
              }
            else
              {
                // x is the last direction: the residual at the Gauss points is complete on this pencil; S^T along x
#pragma unroll
                for (int l = 0; l < n; ++l)
                  res[l] += y_s[base + l];
                if (COLLOC)
                  {
#pragma unroll
                    for (int l = 0; l < n; ++l)
                      y_s[base + l] = res[l];
                  }
                else
                  {
                    tprod<n>(Tb.S, res, out);
#pragma unroll
                    for (int l = 0; l < n; ++l)
                      w1[base + l] = out[l];
                  }
              }
          }
        cell_sync<TPC>(slot);
      }
    if (!COLLOC)
      {
        if (active)
          { // S^T along y, thread (i,k)
#pragma unroll
            for (int j = 0; j < n; ++j)
              in[j] = w1[(b * n + j) * n + a];
            tprod<n>(Tb.S, in, out);
#pragma unroll
            for (int j = 0; j < n; ++j)
              w2[(b * n + j) * n + a] = out[j];
          }
        cell_sync<TPC>(slot);
        if (active)
          { // S^T along z, thread (i,j)
#pragma unroll
            for (int k = 0; k < n; ++k)
              in[k] = w2[k * n2 + b * n + a];
            tprod<n>(Tb.S, in, out);
#pragma unroll
            for (int k = 0; k < n; ++k)
              y_s[k * n2 + b * n + a] = out[k];
          }
        cell_sync<TPC>(slot);
      }

    // ================= fused vector operation, coalesced over the cell
    if (A.mode >= DG_CHEBYSHEV && A.jt_dinv != nullptr)
      {
        const size_t off = (size_t)c * n3;
        if (valid)
          for (int i = t; i < n3; i += TPC)
            y_s[i] = A.rhs[off + i] - y_s[i];
        cell_sync<TPC>(slot);
        jacobi_transformed_local<n, TPC, T>(y_s, w0, w1, Tb.V, A.jt_dinv + dg_cell_type(cx, cy, cz, A.nc) * n3, active, a, b,
                                            slot);
        if (valid)
          for (int i = t; i < n3; i += TPC)
            {
              T r = A.f2 * y_s[i] + (T(1) + A.f1) * u_s[i];
              if (A.mode == DG_CHEBYSHEV)
                r -= A.f1 * A.dst[off + i];
              A.dst[off + i] = r;
            }
        return;
      }
    if (A.mode == DG_SET_SUMS)
      {
        double acc[5] = {0, 0, 0, 0, 0};
        if (valid)
          {
            const size_t off = (size_t)c * n3;
            for (int i = t; i < n3; i += TPC)
              {
                const double v = y_s[i], r = A.rhs[off + i], dv = (double)A.inv_diag[off + i] * v;
                acc[0] += (double)u_s[i] * v;
                acc[1] += dv * r;
                acc[2] += dv * v;
                acc[3] += r * v;
                acc[4] += v * v;
                A.dst[off + i] = y_s[i];
              }
          }
        __shared__ double red[5][C::NT / 32];
#pragma unroll
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
            for (int w = 0; w < C::NT / 32; ++w)
              v += red[threadIdx.x][w];
            A.partial[(size_t)blockIdx.x * 5 + threadIdx.x] = v;
          }
        return;
      }
    if (valid)
      {
        const size_t off = (size_t)c * n3;
        for (int i = t; i < n3; i += TPC)
          {
            const T v = y_s[i];
            T       r;
            if (A.mode == DG_SET)
              r = v;
            else if (A.mode == DG_RESIDUAL)
              r = A.rhs[off + i] - v;
            else
              {
                r = A.f2 * A.inv_diag[off + i] * (A.rhs[off + i] - v) + (T(1) + A.f1) * u_s[i];
                if (A.mode == DG_CHEBYSHEV)
                  r -= A.f1 * A.dst[off + i];
              }
            A.dst[off + i] = r;
          }
      }
  }

  template <int p, typename T, int KIND>
  void
  launch_dg_apply(const DGArgs<T> &A, const DGTables<T, p + 1> &Tb, cudaStream_t stream)
  {
    using C           = DGCfg<p>;
    const size_t smem = (size_t)(C::TABLES + C::PER_CELL * C::CPB) * sizeof(T);
    static bool  first = true;
    if (first)
      {
        cudaFuncSetAttribute(k_dg_apply<p, T, KIND>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        first = false;
      }
    const unsigned blocks = (A.n_cells + C::CPB - 1) / C::CPB;
    k_dg_apply<p, T, KIND><<<blocks, C::NT, smem, stream>>>(A, Tb);
  }

  template <int p, typename T>
  void
  launch_dg_jacobi_transformed(const DGArgs<T> &A, const DGTables<T, p + 1> &Tb, const T scale, cudaStream_t stream)
  {
    using C               = DGCfg<p>;
    const unsigned blocks = (A.n_cells + C::CPB - 1) / C::CPB;
    k_dg_jacobi_transformed<p, T><<<blocks, C::NT, 0, stream>>>(A, Tb, scale);
  }
} // namespace mgb
