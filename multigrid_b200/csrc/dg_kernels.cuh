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
// derivative D, metric, D^T, S^T: 12 sweeps).  Face terms per direction d, both faces at once: traces
// and normal reference derivatives of own and neighbour pencils (the neighbour's p+1 layers are read from
// global memory, i.e. L2), 2-D interpolation of the (p+1)^2 face arrays to the Gauss points, flux, and the
// transposed way back; the result is a rank-one update of the pencil along d plus a term on the face layer.
// DoFs are cell-wise contiguous, so all global accesses of a cell are one contiguous chunk.
#pragma once
#include <cuda_runtime.h>

#include <cstdint>

namespace mgb
{
  enum DGMode
  {
    DG_SET = 0,             // dst = A src                                   (action 0)
    DG_RESIDUAL = 1,        // dst = rhs - A src                             (action 4)
    DG_CHEBYSHEV = 2,       // dst = (1+f1) src - f1 dst + f2 P^-1 (rhs - A src)   (action 3, iteration_index > 1)
    DG_CHEBYSHEV_FIRST = 3  // dst = (1+f1) src + f2 P^-1 (rhs - A src)            (action 3, iteration_index == 1)
  };

  template <typename T, int n>
  struct DGTables
  {
    T S[n * n]; // S[q n + i] = phi_i(x_q)
    T G[n * n]; // phi_i'(x_q)
    T D[n * n]; // D[q n + r] = l_r'(x_q)
    T e0[n], e1[n]; // phi_i'(0), phi_i'(1)
    T f0[n], f1[n]; // phi_i(0), phi_i(1): unit vectors for the Gauss-Lobatto basis (kernel variant KIND = 0)
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

  // KIND = 0: nodal basis with a node on each end point (FE_DGQ on the Gauss-Lobatto points): the trace of a pencil
  // is its end value.  KIND = 1: any tensor-product basis (FE_DGQHermite, FE_DGQArbitraryNodes(QGauss)); traces and
  // their transposes go through the tables f0 / f1 (laplace_operator_dg.h:1179-1199, 1359-1389 use the structure of
  // each basis instead; the result is the same operator in that basis).
  // KIND = 2: generic traces and S = I (basis collocated with the Gauss points, FE_DGQArbitraryNodes(QGauss)): the six
  // interpolation sweeps of the cell term and the S-sums of the face stages drop out (laplace_operator_dg.h:1179).
  template <int p, typename T, int KIND>
  __global__ void
  __launch_bounds__(DGCfg<p>::NT) k_dg_apply(const DGArgs<T> A, const DGTables<T, p + 1> Tb)
  {
    constexpr bool GEN    = KIND != 0;
    constexpr bool COLLOC = KIND == 2;
    using C           = DGCfg<p>;
    constexpr int n   = C::n;
    constexpr int n2  = C::n2;
    constexpr int n3  = C::n3;
    constexpr int TPC = C::TPC;
    extern __shared__ __align__(16) unsigned char dg_smem_raw[];
    const int slot = threadIdx.x / TPC, t = threadIdx.x % TPC;
    // tables that the face sweeps index with the lane's (a,b): shared memory serves divergent indices,
    // the constant bank (kernel parameters) only uniform ones
    T *S_s = reinterpret_cast<T *>(dg_smem_raw), *G_s = S_s + n2, *w_s = G_s + n2;
    for (int i = threadIdx.x; i < n2; i += C::NT)
      {
        S_s[i] = Tb.S[i];
        G_s[i] = Tb.G[i];
      }
    if (threadIdx.x < n)
      w_s[threadIdx.x] = Tb.w[threadIdx.x];
    T        *u_s  = reinterpret_cast<T *>(dg_smem_raw) + C::TABLES + slot * C::PER_CELL;
    T        *y_s  = u_s + n3;
    T        *w0   = y_s + n3;
    T        *w1   = w0 + n3;
    T        *w2   = w1 + n3;
    T        *F    = w2 + n3; // 6 face arrays: [side][j, ds, ts]  (later R1, R2)
    T        *H    = F + 6 * n2; // 8 face arrays: [side][Sj, Sds, Sts, Gts]  (later A1, A2, A3)

    const uint32_t cell   = A.cell_begin + blockIdx.x * C::CPB + slot;
    const bool     valid  = cell < A.cell_begin + A.n_cells;
    const bool     active = valid && t < n2;
    const int      a = t % n, b = t / n;
    const uint32_t c  = valid ? cell : 0;
    const int      cx = c % A.nc[0], cy = (c / A.nc[0]) % A.nc[1], cz = c / (A.nc[0] * A.nc[1]) + A.cz_shift;
    const T *__restrict__ src = A.src;

    // ---- own values
    if (valid)
      for (int i = t; i < n3; i += TPC)
        u_s[i] = src[(size_t)c * n3 + i];
    __syncthreads(); // tables visible to all cells of the block

    // ================= cell integral: collocation at the Gauss points
    T yz[n]; // result pencil along z at (i,j) = (a,b)
    {
      T in[n], out[n];
      const T *uq = COLLOC ? u_s : w0; // values at the Gauss points
      if (!COLLOC)
      {
      if (active)
        { // interpolate along z, thread (i,j)
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
      // reference gradient at the Gauss points: gx -> w1, gy -> w2, gz in registers of thread (i,j)
      T gz[n];
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
            in[k] = uq[k * n2 + b * n + a];
          prod<n>(Tb.D, in, gz);
        }
      cell_sync<TPC>(slot);
      if (active)
        { // metric at the points (a, b, k) of thread (i,j); D^T along z goes to w0
          T fz[n];
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
          tprod<n>(Tb.D, fz, out);
#pragma unroll
          for (int k = 0; k < n; ++k)
            w0[k * n2 + b * n + a] = out[k];
        }
      cell_sync<TPC>(slot);
      if (active)
        { // D^T along y, thread (i,k)
#pragma unroll
          for (int j = 0; j < n; ++j)
            in[j] = w2[(b * n + j) * n + a];
          tprod<n>(Tb.D, in, out);
#pragma unroll
          for (int j = 0; j < n; ++j)
            w0[(b * n + j) * n + a] += out[j];
        }
      cell_sync<TPC>(slot);
      if (active)
        { // D^T along x, thread (j,k), then S^T along x -> w1
#pragma unroll
          for (int i = 0; i < n; ++i)
            in[i] = w1[(b * n + a) * n + i];
          tprod<n>(Tb.D, in, out);
#pragma unroll
          for (int i = 0; i < n; ++i)
            in[i] = w0[(b * n + a) * n + i] + out[i];
          if (COLLOC)
            {
#pragma unroll
              for (int i = 0; i < n; ++i)
                out[i] = in[i];
            }
          else
            tprod<n>(Tb.S, in, out);
#pragma unroll
          for (int i = 0; i < n; ++i)
            w1[(b * n + a) * n + i] = out[i];
        }
      cell_sync<TPC>(slot);
      if (COLLOC)
        {
          if (active)
            { // the result pencil along z, thread (i,j)
#pragma unroll
              for (int k = 0; k < n; ++k)
                yz[k] = w1[k * n2 + b * n + a];
            }
        }
      else
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
              tprod<n>(Tb.S, in, yz);
            }
        }
    }

    // ================= face integrals, direction by direction (d = 2: z, 1: y, 0: x)
#pragma unroll
    for (int d = 2; d >= 0; --d)
      {
        // thread (a,b) and its pencil along d: z-faces (a,b) = (i,j); y-faces (i,k); x-faces (j,k)
        const int base = d == 2 ? b * n + a : (d == 1 ? b * n2 + a : (b * n + a) * n);
        const int sd   = d == 2 ? n2 : (d == 1 ? n : 1);
        const int da = d == 0 ? 1 : 0, db = d == 2 ? 1 : 2; // tangential directions of a and b
        const T   g_dd = A.gs[d][d], g_da = A.gs[d][da], g_db = A.gs[d][db];
        const int cd   = d == 0 ? cx : (d == 1 ? cy : cz);
        const long cstride = d == 0 ? 1 : (d == 1 ? A.nc[0] : (long)A.nc[0] * A.nc[1]);
        if (active)
          {
            T own[n];
#pragma unroll
            for (int l = 0; l < n; ++l)
              own[l] = u_s[base + l * sd];
#pragma unroll
            for (int s = 0; s < 2; ++s)
              {
                const T *eK = s ? Tb.e1 : Tb.e0, *eN = s ? Tb.e0 : Tb.e1;
                const T *fK = s ? Tb.f1 : Tb.f0, *fN = s ? Tb.f0 : Tb.f1;
                T        tK = GEN ? T(0) : (s ? own[n - 1] : own[0]);
                T        dK = 0;
#pragma unroll
                for (int l = 0; l < n; ++l)
                  {
                    dK += eK[l] * own[l];
                    if (GEN)
                      tK += fK[l] * own[l];
                  }
                const bool interior = s ? (cd + 1 < A.nc[d]) : (cd > 0);
                T          jump, dsum, tsum;
                if (interior)
                  {
                    const T *nb = src + (size_t)((long)c + (s ? cstride : -cstride)) * n3 + base;
                    T        tN = 0, dN = 0;
#pragma unroll
                    for (int l = 0; l < n; ++l)
                      {
                        if (GEN && fN[l] == T(0) && eN[l] == T(0))
                          continue; // FE_DGQHermite: two layers of the neighbour (laplace_operator_dg.h:565,1015)
                        const T v = __ldg(nb + l * sd);
                        dN += eN[l] * v;
                        if (GEN)
                          tN += fN[l] * v;
                        else if (l == (s ? 0 : n - 1))
                          tN = v;
                      }
                    jump = tK - tN;
                    dsum = dK + dN;
                    tsum = tK + tN;
                  }
                else
                  { // mirror principle (laplace_operator_dg_face.h:131-160)
                    jump = 2 * tK;
                    dsum = 2 * dK;
                    tsum = 2 * tK;
                  }
                F[(3 * s + 0) * n2 + t] = jump;
                F[(3 * s + 1) * n2 + t] = dsum;
                F[(3 * s + 2) * n2 + t] = tsum;
              }
          }
        cell_sync<TPC>(slot);
        if (active)
          { // sweep along a: thread (qa, b)
#pragma unroll
            for (int s = 0; s < 2; ++s)
              {
                T sj = 0, sds = 0, sts = 0, gts = 0;
                if (COLLOC)
                  {
                    sj  = F[(3 * s + 0) * n2 + t];
                    sds = F[(3 * s + 1) * n2 + t];
                    sts = F[(3 * s + 2) * n2 + t];
#pragma unroll
                    for (int i = 0; i < n; ++i)
                      gts += G_s[a * n + i] * F[(3 * s + 2) * n2 + b * n + i];
                  }
                else
                  {
#pragma unroll
                    for (int i = 0; i < n; ++i)
                      {
                        const T Sv = S_s[a * n + i], Gv = G_s[a * n + i];
                        sj += Sv * F[(3 * s + 0) * n2 + b * n + i];
                        sds += Sv * F[(3 * s + 1) * n2 + b * n + i];
                        const T tv = F[(3 * s + 2) * n2 + b * n + i];
                        sts += Sv * tv;
                        gts += Gv * tv;
                      }
                  }
                H[(4 * s + 0) * n2 + t] = sj;
                H[(4 * s + 1) * n2 + t] = sds;
                H[(4 * s + 2) * n2 + t] = sts;
                H[(4 * s + 3) * n2 + t] = gts;
              }
          }
        cell_sync<TPC>(slot);
        if (active)
          { // sweep along b: thread (qa, qb); flux at the face Gauss point
            const T wJ = A.area[d] * w_s[a] * w_s[b];
#pragma unroll
            for (int s = 0; s < 2; ++s)
              {
                T J = 0, DS = 0, T1 = 0, T2 = 0;
                if (COLLOC)
                  {
                    J  = H[(4 * s + 0) * n2 + t];
                    DS = H[(4 * s + 1) * n2 + t];
                    T1 = H[(4 * s + 3) * n2 + t];
#pragma unroll
                    for (int i = 0; i < n; ++i)
                      T2 += G_s[b * n + i] * H[(4 * s + 2) * n2 + i * n + a];
                  }
                else
                  {
#pragma unroll
                    for (int i = 0; i < n; ++i)
                      {
                        const T Sv = S_s[b * n + i], Gv = G_s[b * n + i];
                        J += Sv * H[(4 * s + 0) * n2 + i * n + a];
                        DS += Sv * H[(4 * s + 1) * n2 + i * n + a];
                        T2 += Gv * H[(4 * s + 2) * n2 + i * n + a];
                        T1 += Sv * H[(4 * s + 3) * n2 + i * n + a];
                      }
                  }
                const T sgn = s ? T(1) : T(-1);
                const T m   = sgn * (g_dd * DS + g_da * T1 + g_db * T2);
                F[(3 * s + 0) * n2 + t] = (A.sigma[d] * J - T(0.5) * m) * wJ; // tested with phi
                F[(3 * s + 1) * n2 + t] = -T(0.5) * J * wJ * sgn;             // tested with dn phi (sign of n folded in)
              }
          }
        cell_sync<TPC>(slot);
        if (active)
          { // back along b: thread (qa, b')
#pragma unroll
            for (int s = 0; s < 2; ++s)
              {
                T a1 = 0, a2 = 0, a3 = 0;
                if (COLLOC)
                  {
                    a1 = F[(3 * s + 0) * n2 + t];
                    a2 = F[(3 * s + 1) * n2 + t];
#pragma unroll
                    for (int q = 0; q < n; ++q)
                      a3 += G_s[q * n + b] * F[(3 * s + 1) * n2 + q * n + a];
                  }
                else
                  {
#pragma unroll
                    for (int q = 0; q < n; ++q)
                      {
                        const T Sv = S_s[q * n + b], Gv = G_s[q * n + b];
                        const T r1 = F[(3 * s + 0) * n2 + q * n + a], r2 = F[(3 * s + 1) * n2 + q * n + a];
                        a1 += Sv * r1;
                        a2 += Sv * r2;
                        a3 += Gv * r2;
                      }
                  }
                H[(4 * s + 0) * n2 + t] = a1 + g_db * a3;
                H[(4 * s + 1) * n2 + t] = a2;
              }
          }
        cell_sync<TPC>(slot);
        if (active)
          { // back along a: thread (a', b'); rank-one update of the pencil along d plus the face layer
            T yd[n];
#pragma unroll
            for (int l = 0; l < n; ++l)
              yd[l] = d == 2 ? yz[l] : T(0);
#pragma unroll
            for (int s = 0; s < 2; ++s)
              {
                T V = 0, N = 0;
                if (COLLOC)
                  {
#pragma unroll
                    for (int q = 0; q < n; ++q)
                      V += G_s[q * n + a] * H[(4 * s + 1) * n2 + b * n + q];
                    V = H[(4 * s + 0) * n2 + t] + g_da * V;
                    N = H[(4 * s + 1) * n2 + t];
                  }
                else
                  {
#pragma unroll
                    for (int q = 0; q < n; ++q)
                      {
                        const T Sv = S_s[q * n + a], Gv = G_s[q * n + a];
                        const T h0 = H[(4 * s + 0) * n2 + b * n + q], h1 = H[(4 * s + 1) * n2 + b * n + q];
                        V += Sv * h0 + g_da * Gv * h1;
                        N += Sv * h1;
                      }
                  }
                N *= g_dd;
                const T *eK = s ? Tb.e1 : Tb.e0, *fK = s ? Tb.f1 : Tb.f0;
#pragma unroll
                for (int l = 0; l < n; ++l)
                  yd[l] += GEN ? N * eK[l] + V * fK[l] : N * eK[l];
                if (!GEN)
                  yd[s ? n - 1 : 0] += V;
              }
#pragma unroll
            for (int l = 0; l < n; ++l)
              {
                if (d == 2)
                  y_s[base + l * sd] = yd[l];
                else
                  y_s[base + l * sd] += yd[l];
              }
          }
        cell_sync<TPC>(slot);
      }

    // ================= fused vector operation, coalesced over the cell
    if (A.mode >= DG_CHEBYSHEV && A.jt_dinv != nullptr)
      {
        // Chebyshev update with the JacobiTransformed preconditioner (laplace_operator_dg.h:1839-1860): the residual
        // rhs - A x of the cell goes through V D^-1 V^T before the update (block-uniform branch)
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
