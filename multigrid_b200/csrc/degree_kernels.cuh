// Degree-templated sm_100a kernels.  Included by degree_tu.cu with MGB_P defined.
//
// k_apply_general   : FEEvaluation-style cell kernel (evaluate -> quadrature-point operation ->
//                     integrate) for arbitrary merged coefficients; restates
//                     common/laplace_operator.h:529-558 + :438-523 with the gather/scatter of
//                     common/vector_access_reduced.h on the patch entity table.
// k_apply_cartesian : the same operator on axis-aligned cells with constant coefficient,
//                     sum-factorised over a whole patch with 1-D mass/stiffness matrices and
//                     fused vector epilogues.
// k_transfer        : MGTransferMatrixFree prolongation / restriction per parent cell.
// k_diagonal        : diagonal of the cell matrices (laplace_operator.h:770-800).
// k_rhs_values      : int f phi   (value part of compute_residual, laplace_operator.h:828-841)
// k_l2_error        : multigrid_solver.h:298-343
#pragma once
#include <cuda_runtime.h>

#include "degree_kernels.h"
#include "functions.cuh"

namespace mgb
{
  template <int n>
  struct ConstTables
  {
    double S[n * n], D[n * n], G[n * n], M1[n * n], K1[n * n], w[n], xq[n], nodes[n];
    float2 MKf[n * n]; // (M1f, K1f) entry by entry, for the packed fp32 kernel
    float  Sf[n * n], Df[n * n], M1f[n * n], K1f[n * n], wf[n];
    double P1[(2 * n - 1) * n];
    float  P1f[(2 * n - 1) * n];
    double R1[n * (2 * n - 1)];
  };

  __constant__ ConstTables<MGB_P + 1> ctab;

  template <typename T>
  struct Tab;
  template <>
  struct Tab<double>
  {
    static __device__ __forceinline__ const double *S() { return ctab.S; }
    static __device__ __forceinline__ const double *D() { return ctab.D; }
    static __device__ __forceinline__ const double *M1() { return ctab.M1; }
    static __device__ __forceinline__ const double *K1() { return ctab.K1; }
    static __device__ __forceinline__ const double *w() { return ctab.w; }
  };
  template <>
  struct Tab<float>
  {
    static __device__ __forceinline__ const float *S() { return ctab.Sf; }
    static __device__ __forceinline__ const float *D() { return ctab.Df; }
    static __device__ __forceinline__ const float *M1() { return ctab.M1f; }
    static __device__ __forceinline__ const float *K1() { return ctab.K1f; }
    static __device__ __forceinline__ const float *w() { return ctab.wf; }
  };

  // out (+)= Mat in  (or Mat^T in); Mat in constant memory, all indices compile-time
  template <int n, typename T, bool TR, bool ADD>
  __device__ __forceinline__ void
  mat_vec(const T *__restrict__ Mc, const T (&in)[n], T (&out)[n])
  {
#pragma unroll
    for (int o = 0; o < n; ++o)
      {
        T s = ADD ? out[o] : T(0);
#pragma unroll
        for (int i = 0; i < n; ++i)
          s += (TR ? Mc[i * n + o] : Mc[o * n + i]) * in[i];
        out[o] = s;
      }
  }

  __host__ __device__ constexpr int
  cells_per_block(const int p)
  {
    return p == 1 ? 32 : p == 2 ? 14 : p == 3 ? 8 : p == 4 ? 8 : p == 5 ? 4 : p == 6 ? 3 : 2;
  }

  // ------------------------------------------------------------------ general cell kernel
  // FEEvaluation-style cell loop (laplace_operator.h:529-558): gather, evaluate the reference gradient at the Gauss
  // points by sum factorisation (S in x, y, z, collocation derivative D per direction), quadrature-point operation with the
  // merged coefficient (:438-523), integrate (the transposes), scatter-add.  (p+1)^2 threads per cell; thread (a, b) owns
  // the z-pencil (i, j) = (b, a) at both ends: the gathered values go into the first sweep from registers and the last
  // sweep leaves the result in registers for the scatter, so one DoF index is computed per thread (the p+1 points of a
  // pencil are one plane pitch apart) and the per-point coefficients are read in their storage order (x fastest across the
  // threads: coalesced).  Three shared-memory arrays per cell: values / residual U, and the x- and y-components G0, G1.
  template <int p, typename T, bool VAR>
  __global__ void
  __launch_bounds__(cells_per_block(p) * (p + 1) * (p + 1))
    k_apply_general(const ApplyArgs<T> A, T *__restrict__ acc)
  {
    constexpr int n    = p + 1;
    constexpr int npad = n | 1; // odd x-stride: conflict-free line access in every direction
    constexpr int NN   = n * n * npad;
    constexpr int tpc  = n * n;
    constexpr int CPB  = cells_per_block(p);
    extern __shared__ __align__(16) unsigned char smem_raw[];
    T *smem = reinterpret_cast<T *>(smem_raw);

    const int      cl     = threadIdx.x / tpc;
    const int      t      = threadIdx.x - cl * tpc;
    const int      a      = t / n;
    const int      b      = t - a * n;
    const uint32_t cell   = blockIdx.x * CPB + cl;
    const bool     active = cell < A.L.n_cells;
    T             *U      = smem + cl * NN; // [array][cell]: the cells of a block side by side (fewer bank conflicts
    T             *G0     = U + CPB * NN;   // where a warp straddles two cells)
    T             *G1     = G0 + CPB * NN;
#define IDX(i, j, k) (((k)*n + (j)) * npad + (i))

    const T *Sm = Tab<T>::S();
    const T *Dm = Tab<T>::D();
    T        r[n], s[n];

    // gather (vector_access_reduced.h:11-251): the z-pencil at (i, j) = (b, a); constrained entries read as zero
    const LevelDesc &L = A.L;
    int              cx = 0, cy = 0, cz = 0;
    if (active)
      cell_coords(L, cell, cx, cy, cz);
    const int      gi = cx * p + b, gj = cy * p + a;
    const uint32_t pitch = plane_pitch(L);
    const uint32_t base  = active ? dof_index(L, gi, gj, cz * p) : 0u;
    const bool     xycon = is_constrained_xy(L, gi, gj);
    const bool     zlo = cz == 0 && (L.dirichlet & 16u), zhi = cz == L.ncell[2] - 1 && (L.dirichlet & 32u);
#pragma unroll
    for (int m = 0; m < n; ++m)
      {
        const bool c = xycon || (m == 0 && zlo) || (m == n - 1 && zhi);
        r[m]         = (active && !(c && !A.read_plain)) ? A.src[base + (uint32_t)m * pitch] : T(0);
      }
    // the merged coefficients of this thread's quadrature points (b, a, k): requested now, used after the evaluation
    // sweeps (the stream that dominates the HBM traffic: its latency is covered by the sweeps instead of stalling the
    // quadrature-point loop)
    // float: into registers; double (registers are short: 60 more would halve the occupancy): asynchronous copies of the
    // block's contiguous coefficient range into shared memory
    constexpr bool PF_REG = VAR && sizeof(T) == 4, PF_SMEM = VAR && sizeof(T) == 8;
    T              cf[PF_REG ? 6 : 1][PF_REG ? n : 1];
    T             *cs = smem + ((3 * CPB * NN + 1) & ~1); // behind the three arrays (16-byte aligned): [cell][6][n^3]
    if (PF_REG)
      {
        const T *cq = A.coef_q + (size_t)(active ? cell : 0) * 6 * (n * n * n) + a * n + b;
#pragma unroll
        for (int e = 0; e < 6; ++e)
#pragma unroll
          for (int k = 0; k < n; ++k)
            cf[PF_REG ? e : 0][PF_REG ? k : 0] = active ? cq[e * n * n * n + k * n * n] : T(0);
      }
    if (PF_SMEM)
      {
        // 16-byte chunks; a cell's block of 6 n^3 doubles is a multiple of 16 bytes and 16-byte aligned
        const uint32_t first = blockIdx.x * CPB;
        const uint32_t ncell = A.L.n_cells - first < (uint32_t)CPB ? A.L.n_cells - first : (uint32_t)CPB;
        const uint32_t nchunk = ncell * (6 * n * n * n / 2);
        const T       *g      = A.coef_q + (size_t)first * 6 * (n * n * n);
        const uint32_t sb     = (uint32_t)__cvta_generic_to_shared(cs);
        for (uint32_t q = threadIdx.x; q < nchunk; q += blockDim.x)
          asm volatile("cp.async.cg.shared.global [%0], [%1], 16;\n" ::"r"(sb + q * 16u), "l"(g + 2 * q));
        asm volatile("cp.async.commit_group;\n" ::);
      }
    // evaluate: S in z from the registers, then y, then x together with the x-derivative
    mat_vec<n, T, false, false>(Sm, r, s);
#pragma unroll
    for (int k = 0; k < n; ++k)
      U[IDX(b, a, k)] = s[k];
    __syncthreads();
#pragma unroll
    for (int j = 0; j < n; ++j)
      r[j] = U[IDX(b, j, a)];
    mat_vec<n, T, false, false>(Sm, r, s);
#pragma unroll
    for (int j = 0; j < n; ++j)
      U[IDX(b, j, a)] = s[j];
    __syncthreads();
#pragma unroll
    for (int i = 0; i < n; ++i)
      r[i] = U[IDX(i, b, a)];
    mat_vec<n, T, false, false>(Sm, r, s);
    mat_vec<n, T, false, false>(Dm, s, r);
#pragma unroll
    for (int i = 0; i < n; ++i)
      {
        U[IDX(i, b, a)]  = s[i];
        G0[IDX(i, b, a)] = r[i];
      }
    __syncthreads();
    // y-derivative
#pragma unroll
    for (int j = 0; j < n; ++j)
      r[j] = U[IDX(b, j, a)];
    mat_vec<n, T, false, false>(Dm, r, s);
#pragma unroll
    for (int j = 0; j < n; ++j)
      G1[IDX(b, j, a)] = s[j];
    if (PF_SMEM)
      asm volatile("cp.async.wait_group 0;\n" ::: "memory"); // own copies landed; the barrier below publishes all of them
    __syncthreads();
    // z-derivative and the quadrature-point operation (laplace_operator.h:438-523) on the points (b, a, k); D_z^T of the
    // z-component stays with this thread and opens the residual in U
#pragma unroll
    for (int k = 0; k < n; ++k)
      r[k] = U[IDX(b, a, k)];
    mat_vec<n, T, false, false>(Dm, r, s);
    {
      const T *wq  = Tab<T>::w();
      const T  wab = wq[a] * wq[b];
#pragma unroll
      for (int k = 0; k < n; ++k)
        {
          T c0, c1, c2, c3, c4, c5;
          if (PF_REG)
            {
              c0 = cf[0][PF_REG ? k : 0];
              c1 = cf[PF_REG ? 1 : 0][PF_REG ? k : 0];
              c2 = cf[PF_REG ? 2 : 0][PF_REG ? k : 0];
              c3 = cf[PF_REG ? 3 : 0][PF_REG ? k : 0];
              c4 = cf[PF_REG ? 4 : 0][PF_REG ? k : 0];
              c5 = cf[PF_REG ? 5 : 0][PF_REG ? k : 0];
            }
          else if (PF_SMEM)
            {
              const T *c = cs + cl * 6 * (n * n * n) + k * n * n + t;
              c0         = active ? c[0 * n * n * n] : T(0);
              c1         = active ? c[1 * n * n * n] : T(0);
              c2         = active ? c[2 * n * n * n] : T(0);
              c3         = active ? c[3 * n * n * n] : T(0);
              c4         = active ? c[4 * n * n * n] : T(0);
              c5         = active ? c[5 * n * n * n] : T(0);
            }
          else
            {
              const T weight = wq[k] * wab;
              c0             = A.coef[0] * weight;
              c1             = A.coef[1] * weight;
              c2             = A.coef[2] * weight;
              c3             = A.coef[3] * weight;
              c4             = A.coef[4] * weight;
              c5             = A.coef[5] * weight;
            }
          const T t0 = G0[IDX(b, a, k)], t1 = G1[IDX(b, a, k)], t2 = s[k];
          G0[IDX(b, a, k)] = c0 * t0 + c3 * t1 + c4 * t2;
          G1[IDX(b, a, k)] = c3 * t0 + c1 * t1 + c5 * t2;
          r[k]             = c4 * t0 + c5 * t1 + c2 * t2;
        }
    }
    mat_vec<n, T, true, false>(Dm, r, s);
#pragma unroll
    for (int k = 0; k < n; ++k)
      U[IDX(b, a, k)] = s[k];
    __syncthreads();
    // integrate.  x-pencils: S_x^T (U + D_x^T G0) -> U, and S_x^T G1 in place (it commutes with the D_y^T to come)
#pragma unroll
    for (int i = 0; i < n; ++i)
      {
        r[i] = G0[IDX(i, b, a)];
        s[i] = U[IDX(i, b, a)];
      }
    mat_vec<n, T, true, true>(Dm, r, s);
    mat_vec<n, T, true, false>(Sm, s, r);
#pragma unroll
    for (int i = 0; i < n; ++i)
      {
        U[IDX(i, b, a)] = r[i];
        s[i]            = G1[IDX(i, b, a)];
      }
    mat_vec<n, T, true, false>(Sm, s, r);
#pragma unroll
    for (int i = 0; i < n; ++i)
      G1[IDX(i, b, a)] = r[i];
    __syncthreads();
    // y-pencils: S_y^T (U + D_y^T G1) -> U
#pragma unroll
    for (int j = 0; j < n; ++j)
      {
        r[j] = G1[IDX(b, j, a)];
        s[j] = U[IDX(b, j, a)];
      }
    mat_vec<n, T, true, true>(Dm, r, s);
    mat_vec<n, T, true, false>(Sm, s, r);
#pragma unroll
    for (int j = 0; j < n; ++j)
      U[IDX(b, j, a)] = r[j];
    __syncthreads();
    // z-pencils: S_z^T into registers, scatter-add (vector_access_reduced.h:255-505): constrained entries skipped
#pragma unroll
    for (int k = 0; k < n; ++k)
      r[k] = U[IDX(b, a, k)];
    mat_vec<n, T, true, false>(Sm, r, s);
    if (active)
      {
#pragma unroll
        for (int m = 0; m < n; ++m)
          {
            const bool c = xycon || (m == 0 && zlo) || (m == n - 1 && zhi);
            if (!c)
              atomicAdd(&acc[base + (uint32_t)m * pitch], s[m]);
          }
      }
#undef IDX
  }

  template <int p, typename T>
  void
  launch_apply_general(const ApplyArgs<T> &A, cudaStream_t stream)
  {
    constexpr int n    = p + 1;
    constexpr int NN   = n * n * (n | 1);
    constexpr int CPB  = cells_per_block(p);
    // three work arrays per cell; the double instance with per-point coefficients also stages them (6 n^3 per cell)
    const size_t  smem = (size_t)(((3 * CPB * NN + 1) & ~1) + (A.coef_q && sizeof(T) == 8 ? CPB * 6 * n * n * n : 0)) * sizeof(T);
    const unsigned grid = (A.L.n_cells + CPB - 1) / CPB;
    static bool    attr = false;
    if (!attr && A.coef_q)
      {
        cudaFuncSetAttribute(k_apply_general<p, T, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        attr = true;
      }
    if (A.coef_q)
      k_apply_general<p, T, true><<<grid, CPB * n * n, smem, stream>>>(A, A.dst);
    else
      k_apply_general<p, T, false><<<grid, CPB * n * n, smem, stream>>>(A, A.dst);
  }

  // ------------------------------------------------------------------ level transfer
  // one block per parent (coarse) cell; P1 = 1-D embedding (2p+1) x (p+1)
  template <int p, typename T>
  __global__ void
  k_transfer(const TransferArgs<T> A)
  {
    constexpr int n  = p + 1;
    constexpr int nf = 2 * p + 1;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    T            *P1   = reinterpret_cast<T *>(smem_raw);
    T            *buf0 = P1 + nf * n;
    T            *buf1 = buf0 + nf * nf * nf;
    const int     tid = threadIdx.x, nt = blockDim.x;
    for (int i = tid; i < nf * n; i += nt)
      P1[i] = (T)ctab.P1[i];

    const uint32_t ccell = blockIdx.x;
    int            cx, cy, cz;
    cell_coords(A.coarse, ccell, cx, cy, cz);

    if (A.mode < 2)
      {
        // ---- prolongation: u (n^3) -> x -> y -> z -> fine (nf^3)
        T *u = buf1; // n^3 <= nf*nf*n
        for (int l = tid; l < n * n * n; l += nt)
          {
            const int i = l % n, j = (l / n) % n, k = l / (n * n);
            bool      c;
            const uint32_t d = cell_dof(A.coarse, ccell, i, j, k, c);
            u[l]             = (c && A.constrained) ? T(0) : A.src[d];
          }
        __syncthreads();
        // x: a[k][j][o] in buf0 (n*n*nf)
        for (int l = tid; l < n * n * nf; l += nt)
          {
            const int o = l % nf, kj = l / nf;
            T         s = 0;
#pragma unroll
            for (int i = 0; i < n; ++i)
              s += P1[o * n + i] * u[kj * n + i];
            buf0[l] = s;
          }
        __syncthreads();
        // y: c[k][o][i] in buf1 (n*nf*nf)
        for (int l = tid; l < n * nf * nf; l += nt)
          {
            const int i = l % nf, o = (l / nf) % nf, k = l / (nf * nf);
            T         s = 0;
#pragma unroll
            for (int j = 0; j < n; ++j)
              s += P1[o * n + j] * buf0[(k * n + j) * nf + i];
            buf1[l] = s;
          }
        __syncthreads();
        // z + write: every fine DoF is written by exactly one parent (lower-owner rule)
        for (int l = tid; l < nf * nf * nf; l += nt)
          {
            const int I = l % nf, J = (l / nf) % nf, K = l / (nf * nf);
            const bool owned = (I < 2 * p || cx == A.coarse.ncell[0] - 1) &&
                               (J < 2 * p || cy == A.coarse.ncell[1] - 1) &&
                               (K < 2 * p || cz == A.coarse.ncell[2] - 1);
            if (!owned)
              continue;
            T s = 0;
#pragma unroll
            for (int k = 0; k < n; ++k)
              s += P1[K * n + k] * buf1[(k * nf + J) * nf + I];
            const int      fx = 2 * cx + (I > p), fy = 2 * cy + (J > p), fz = 2 * cz + (K > p);
            const uint32_t fcell = cell_number(A.fine, fx, fy, fz);
            bool           c;
            const uint32_t d =
              cell_dof(A.fine, fcell, I - (I > p) * p, J - (J > p) * p, K - (K > p) * p, c);
            if (A.mode == 1)
              A.dst[d] += s;
            else
              A.dst[d] = s;
          }
      }
    else
      {
        // ---- restriction: fine (nf^3, weighted) -> z^T -> y^T -> x^T -> coarse (n^3), add
        for (int l = tid; l < nf * nf * nf; l += nt)
          {
            const int I = l % nf, J = (l / nf) % nf, K = l / (nf * nf);
            const int fx = 2 * cx + (I > p), fy = 2 * cy + (J > p), fz = 2 * cz + (K > p);
            const uint32_t fcell = cell_number(A.fine, fx, fy, fz);
            bool           c;
            const uint32_t d =
              cell_dof(A.fine, fcell, I - (I > p) * p, J - (J > p) * p, K - (K > p) * p, c);
            T w = 1;
            if ((I == 0 && cx > 0) || (I == 2 * p && cx < A.coarse.ncell[0] - 1))
              w *= T(0.5);
            if ((J == 0 && cy > 0) || (J == 2 * p && cy < A.coarse.ncell[1] - 1))
              w *= T(0.5);
            if ((K == 0 && cz > 0) || (K == 2 * p && cz < A.coarse.ncell[2] - 1))
              w *= T(0.5);
            buf0[l] = w * A.src[d];
          }
        __syncthreads();
        // z^T: c[k][J][I] in buf1 (n*nf*nf)
        for (int l = tid; l < n * nf * nf; l += nt)
          {
            const int IJ = l % (nf * nf), k = l / (nf * nf);
            T         s  = 0;
#pragma unroll
            for (int K = 0; K < nf; ++K)
              s += P1[K * n + k] * buf0[K * nf * nf + IJ];
            buf1[l] = s;
          }
        __syncthreads();
        // y^T: a[k][j][I] in buf0 (n*n*nf)
        for (int l = tid; l < n * n * nf; l += nt)
          {
            const int I = l % nf, j = (l / nf) % n, k = l / (nf * n);
            T         s = 0;
#pragma unroll
            for (int J = 0; J < nf; ++J)
              s += P1[J * n + j] * buf1[(k * nf + J) * nf + I];
            buf0[l] = s;
          }
        __syncthreads();
        // x^T + scatter-add
        for (int l = tid; l < n * n * n; l += nt)
          {
            const int i = l % n, kj = l / n;
            T         s = 0;
#pragma unroll
            for (int I = 0; I < nf; ++I)
              s += P1[I * n + i] * buf0[kj * nf + I];
            bool           c;
            const uint32_t d = cell_dof(A.coarse, ccell, i, (l / n) % n, l / (n * n), c);
            if (!(c && A.constrained))
              atomicAdd(&A.dst[d], s);
          }
      }
  }

  template <int p, typename T>
  void
  launch_transfer(const TransferArgs<T> &A, cudaStream_t stream)
  {
    constexpr int nf = 2 * p + 1;
    constexpr int n  = p + 1;
    const int     nt = nf * nf < 64 ? 64 : (nf * nf > 256 ? 256 : ((nf * nf + 31) / 32) * 32);
    const size_t  smem = (size_t)(nf * n + nf * nf * nf + nf * nf * n) * sizeof(T);
    static bool   attr_set = false;
    if (!attr_set && smem > 48 * 1024)
      {
        cudaFuncSetAttribute(k_transfer<p, T>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        attr_set = true;
      }
    k_transfer<p, T><<<A.coarse.n_cells, nt, smem, stream>>>(A);
  }

  // ------------------------------------------------------------------ diagonal
  // diag_i = sum_q grad(phi_i)(q)^T C_q grad(phi_i)(q); one block per cell, thread per local DoF
  template <int p, typename T>
  __global__ void
  k_diagonal(const DiagArgs<T> A)
  {
    constexpr int  n    = p + 1;
    constexpr int  n3   = n * n * n;
    const uint32_t cell = blockIdx.x;
    for (int l = threadIdx.x; l < n3; l += blockDim.x)
      {
        const int i0 = l % n, i1 = (l / n) % n, i2 = l / (n * n);
        double    sum = 0;
        for (int q2 = 0; q2 < n; ++q2)
          for (int q1 = 0; q1 < n; ++q1)
            for (int q0 = 0; q0 < n; ++q0)
              {
                const double gx = ctab.G[q0 * n + i0] * ctab.S[q1 * n + i1] * ctab.S[q2 * n + i2];
                const double gy = ctab.S[q0 * n + i0] * ctab.G[q1 * n + i1] * ctab.S[q2 * n + i2];
                const double gz = ctab.S[q0 * n + i0] * ctab.S[q1 * n + i1] * ctab.G[q2 * n + i2];
                double       c[6];
                if (A.coef_q)
                  {
                    const T *cq = A.coef_q + (size_t)cell * 6 * n3 + (q2 * n + q1) * n + q0;
                    for (int e = 0; e < 6; ++e)
                      c[e] = cq[e * n3];
                  }
                else
                  {
                    const double weight = ctab.w[q0] * ctab.w[q1] * ctab.w[q2];
                    for (int e = 0; e < 6; ++e)
                      c[e] = (double)A.coef[e] * weight;
                  }
                sum += c[0] * gx * gx + c[1] * gy * gy + c[2] * gz * gz +
                       2. * (c[3] * gx * gy + c[4] * gx * gz + c[5] * gy * gz);
              }
        bool           con;
        const uint32_t d = cell_dof(A.L, cell, i0, i1, i2, con);
        if (!con)
          atomicAdd(&A.diag[d], (T)sum);
      }
  }

  template <int p, typename T>
  void
  launch_diagonal(const DiagArgs<T> &A, cudaStream_t stream)
  {
    constexpr int n3 = (p + 1) * (p + 1) * (p + 1);
    k_diagonal<p, T><<<A.L.n_cells, n3 < 256 ? ((n3 + 31) / 32) * 32 : 256, 0, stream>>>(A);
  }

  // ------------------------------------------------------------------ mapped cells (MappingQGeneric(p))
  // Images of the (p+1)^3 Gauss-Lobatto support points of cell (cx,cy,cz) under the geometry map, X[d * n^3 + l]
  // with l lexicographic (x fastest): the control points of the cell's polynomial geometry.  All threads of the
  // block take part; the caller synchronises before reading X.
  template <int p>
  __device__ __forceinline__ void
  cell_support_points(const Geometry &g, const LevelDesc &L, const int cx, const int cy, const int cz, double *X)
  {
    constexpr int n = p + 1, n3 = n * n * n;
    const double  ia = 1. / L.ncell[0], ib = 1. / L.ncell[1], ic = (double)L.p / (L.nd2_all - 1);
    const int     cz_global = cz + L.gk0 / L.p;
    for (int l = threadIdx.x; l < n3; l += blockDim.x)
      {
        const int i = l % n, j = (l / n) % n, k = l / (n * n);
        double    x[3];
        geometry_point(g, (cx + ctab.nodes[i]) * ia, (cy + ctab.nodes[j]) * ib, (cz_global + ctab.nodes[k]) * ic, x);
        X[l]          = x[0];
        X[n3 + l]     = x[1];
        X[2 * n3 + l] = x[2];
      }
  }

  // position x and Jacobian J[d][e] = d x_d / d xi_e of the cell's polynomial geometry at quadrature point (q0,q1,q2)
  template <int p>
  __device__ __forceinline__ void
  mapped_point(const double *X, const int q0, const int q1, const int q2, double x[3], double J[3][3])
  {
    constexpr int n = p + 1, n3 = n * n * n;
    for (int d = 0; d < 3; ++d)
      {
        x[d] = 0;
        for (int e = 0; e < 3; ++e)
          J[d][e] = 0;
      }
    for (int k = 0; k < n; ++k)
      for (int j = 0; j < n; ++j)
        {
          const double s2 = ctab.S[q2 * n + k], g2 = ctab.G[q2 * n + k];
          const double s1 = ctab.S[q1 * n + j], g1 = ctab.G[q1 * n + j];
          for (int i = 0; i < n; ++i)
            {
              const double s0 = ctab.S[q0 * n + i], g0 = ctab.G[q0 * n + i];
              const int    l  = (k * n + j) * n + i;
              const double v = s0 * s1 * s2, d0 = g0 * s1 * s2, d1 = s0 * g1 * s2, d2 = s0 * s1 * g2;
              for (int d = 0; d < 3; ++d)
                {
                  const double Xd = X[d * n3 + l];
                  x[d] += Xd * v;
                  J[d][0] += Xd * d0;
                  J[d][1] += Xd * d1;
                  J[d][2] += Xd * d2;
                }
            }
        }
  }

  __device__ __forceinline__ double
  det3(const double J[3][3])
  {
    return J[0][0] * (J[1][1] * J[2][2] - J[1][2] * J[2][1]) - J[0][1] * (J[1][0] * J[2][2] - J[1][2] * J[2][0]) +
           J[0][2] * (J[1][0] * J[2][1] - J[1][1] * J[2][0]);
  }

  // laplace_operator.h:388-430 on general cells: c(x_q) JxW J^-1 J^-T, six entries per quadrature point
  template <int p, typename T>
  __global__ void
  k_mapped_coefficient(const CoefArgs<T> A)
  {
    constexpr int     n = p + 1, n3 = n * n * n;
    __shared__ double X[3 * n3];
    const uint32_t    cell = blockIdx.x;
    int               cx, cy, cz;
    cell_coords(A.L, cell, cx, cy, cz);
    cell_support_points<p>(A.geom, A.L, cx, cy, cz, X);
    __syncthreads();
    for (int q = threadIdx.x; q < n3; q += blockDim.x)
      {
        const int q0 = q % n, q1 = (q / n) % n, q2 = q / (n * n);
        double    x[3], J[3][3];
        mapped_point<p>(X, q0, q1, q2, x, J);
        const double det = det3(J);
        // inverse Jacobian: I[e][d] = d xi_e / d x_d (cofactors)
        double I[3][3];
        I[0][0] = (J[1][1] * J[2][2] - J[1][2] * J[2][1]) / det;
        I[0][1] = (J[0][2] * J[2][1] - J[0][1] * J[2][2]) / det;
        I[0][2] = (J[0][1] * J[1][2] - J[0][2] * J[1][1]) / det;
        I[1][0] = (J[1][2] * J[2][0] - J[1][0] * J[2][2]) / det;
        I[1][1] = (J[0][0] * J[2][2] - J[0][2] * J[2][0]) / det;
        I[1][2] = (J[0][2] * J[1][0] - J[0][0] * J[1][2]) / det;
        I[2][0] = (J[1][0] * J[2][1] - J[1][1] * J[2][0]) / det;
        I[2][1] = (J[0][1] * J[2][0] - J[0][0] * J[2][1]) / det;
        I[2][2] = (J[0][0] * J[1][1] - J[0][1] * J[1][0]) / det;
        const double f = eval_fn(A.fn_id, x[0], x[1], x[2]) * ctab.w[q0] * ctab.w[q1] * ctab.w[q2] * fabs(det);
        auto         m = [&](const int e, const int g) { return f * (I[e][0] * I[g][0] + I[e][1] * I[g][1] + I[e][2] * I[g][2]); };
        T           *c = A.coef_q + (size_t)cell * 6 * n3 + q;
        c[0 * n3]      = (T)m(0, 0);
        c[1 * n3]      = (T)m(1, 1);
        c[2 * n3]      = (T)m(2, 2);
        c[3 * n3]      = (T)m(0, 1);
        c[4 * n3]      = (T)m(0, 2);
        c[5 * n3]      = (T)m(1, 2);
      }
  }

  template <int p, typename T>
  void
  launch_coefficient(const CoefArgs<T> &A, cudaStream_t stream)
  {
    constexpr int n3 = (p + 1) * (p + 1) * (p + 1);
    k_mapped_coefficient<p, T><<<A.L.n_cells, n3 < 128 ? ((n3 + 31) / 32) * 32 : 128, 0, stream>>>(A);
  }

  // MinimalSurfaceOperator::evaluate_coefficient(first_time, solution) (minimal_surface/program.cc:117-163): the merged
  // coefficient JxW J^-1 T J^-T from the gradient g of `u` at the quadrature points (all entries of u are read, Dirichlet
  // values included): mode 0 T = I; mode 1 T = (I - g g^T / (1 + |g|^2)) / sqrt(1 + |g|^2), the Newton linearisation of
  // the minimal surface operator; mode 2 T = I / sqrt(1 + |g|^2), the flux of the nonlinear residual (:188-192)
  template <int p, typename T>
  __global__ void
  k_solution_coefficient(const SolutionCoefArgs<T> A)
  {
    // one thread per quadrature point; reference derivatives of the three coordinate fields of the cell geometry and of
    // u by sum factorisation (x: S, G; y: three combinations; z in registers), field after field through shared memory
    constexpr int n = p + 1, n2 = n * n, n3 = n * n * n;
    extern __shared__ double sc_smem[];
    double        *X = sc_smem, *U = X + 3 * n3, *a_s = U + n3 /* x-stage: (S u, G u) */, *b_s = a_s + 2 * n3 /* y-stage */;
    const uint32_t cell = blockIdx.x;
    int            cx, cy, cz;
    cell_coords(A.L, cell, cx, cy, cz);
    const bool box = A.geom.kind == 0;
    if (!box)
      cell_support_points<p>(A.geom, A.L, cx, cy, cz, X);
    for (int l = threadIdx.x; l < n3; l += blockDim.x)
      {
        bool           con;
        const uint32_t d = cell_dof(A.L, cell, l % n, (l / n) % n, l / (n * n), con);
        U[l]             = (double)A.u[d];
      }
    __syncthreads();
    const int  q = threadIdx.x, q0 = q % n, q1 = (q / n) % n, q2 = q / n2;
    const bool act = q < n3;
    double     der[4][3]; // [field][reference direction]
    for (int f = box ? 3 : 0; f < 4; ++f)
      {
        const double *in = X + f * n3;
        if (act)
          { // x: thread (q0, j = q1, k = q2)
            double s = 0, g = 0;
            for (int i = 0; i < n; ++i)
              {
                const double v = in[(q2 * n + q1) * n + i];
                s += ctab.S[q0 * n + i] * v;
                g += ctab.G[q0 * n + i] * v;
              }
            a_s[q]      = s;
            a_s[n3 + q] = g;
          }
        __syncthreads();
        if (act)
          { // y: thread (q0, q1, k = q2): G S, S G, S S
            double gs = 0, sg = 0, ss = 0;
            for (int j = 0; j < n; ++j)
              {
                const double vs = a_s[(q2 * n + j) * n + q0], vg = a_s[n3 + (q2 * n + j) * n + q0];
                gs += ctab.S[q1 * n + j] * vg;
                sg += ctab.G[q1 * n + j] * vs;
                ss += ctab.S[q1 * n + j] * vs;
              }
            b_s[q]          = gs;
            b_s[n3 + q]     = sg;
            b_s[2 * n3 + q] = ss;
          }
        __syncthreads();
        if (act)
          { // z
            double d0 = 0, d1 = 0, d2 = 0;
            for (int k = 0; k < n; ++k)
              {
                const int l = (k * n + q1) * n + q0;
                d0 += ctab.S[q2 * n + k] * b_s[l];
                d1 += ctab.S[q2 * n + k] * b_s[n3 + l];
                d2 += ctab.G[q2 * n + k] * b_s[2 * n3 + l];
              }
            der[f][0] = d0;
            der[f][1] = d1;
            der[f][2] = d2;
          }
        // (the next field's x-stage writes a_s, which nobody reads any more; its y-stage writes b_s after the barrier
        // that follows the x-stage, by when every thread has left this z-stage)
      }
    if (act)
      {
        double J[3][3] = {{A.L.h[0], 0, 0}, {0, A.L.h[1], 0}, {0, 0, A.L.h[2]}};
        if (!box)
          for (int d = 0; d < 3; ++d)
            for (int e = 0; e < 3; ++e)
              J[d][e] = der[d][e];
        const double gr[3] = {der[3][0], der[3][1], der[3][2]};
        const double det = det3(J);
        double       I[3][3]; // I[e][d] = d xi_e / d x_d
        I[0][0] = (J[1][1] * J[2][2] - J[1][2] * J[2][1]) / det;
        I[0][1] = (J[0][2] * J[2][1] - J[0][1] * J[2][2]) / det;
        I[0][2] = (J[0][1] * J[1][2] - J[0][2] * J[1][1]) / det;
        I[1][0] = (J[1][2] * J[2][0] - J[1][0] * J[2][2]) / det;
        I[1][1] = (J[0][0] * J[2][2] - J[0][2] * J[2][0]) / det;
        I[1][2] = (J[0][2] * J[1][0] - J[0][0] * J[1][2]) / det;
        I[2][0] = (J[1][0] * J[2][1] - J[1][1] * J[2][0]) / det;
        I[2][1] = (J[0][1] * J[2][0] - J[0][0] * J[2][1]) / det;
        I[2][2] = (J[0][0] * J[1][1] - J[0][1] * J[1][0]) / det;
        double g[3];
        for (int d = 0; d < 3; ++d)
          g[d] = I[0][d] * gr[0] + I[1][d] * gr[1] + I[2][d] * gr[2];
        const double g2n = g[0] * g[0] + g[1] * g[1] + g[2] * g[2];
        const double isq = A.mode == 0 ? 1. : 1. / sqrt(1. + g2n);
        const double f   = ctab.w[q0] * ctab.w[q1] * ctab.w[q2] * fabs(det);
        // a_e = J^-1 g (row e of I times g): J^-1 (g g^T) J^-T = a a^T
        double a[3] = {0, 0, 0};
        if (A.mode == 1)
          for (int e = 0; e < 3; ++e)
            a[e] = I[e][0] * g[0] + I[e][1] * g[1] + I[e][2] * g[2];
        auto m = [&](const int e, const int h) {
          return f * isq * (I[e][0] * I[h][0] + I[e][1] * I[h][1] + I[e][2] * I[h][2] - a[e] * a[h] / (1. + g2n));
        };
        T *c      = A.coef_q + (size_t)cell * 6 * n3 + q;
        c[0 * n3] = (T)m(0, 0);
        c[1 * n3] = (T)m(1, 1);
        c[2 * n3] = (T)m(2, 2);
        c[3 * n3] = (T)m(0, 1);
        c[4 * n3] = (T)m(0, 2);
        c[5 * n3] = (T)m(1, 2);
      }
  }

  template <int p, typename T>
  void
  launch_solution_coefficient(const SolutionCoefArgs<T> &A, cudaStream_t stream)
  {
    constexpr int n3 = (p + 1) * (p + 1) * (p + 1);
    constexpr size_t smem = (size_t)9 * n3 * sizeof(double);
    static bool      attribute_set = false;
    if (!attribute_set && smem > 48 * 1024)
      {
        cudaFuncSetAttribute(k_solution_coefficient<p, T>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        attribute_set = true;
      }
    k_solution_coefficient<p, T><<<A.L.n_cells, ((n3 + 31) / 32) * 32, smem, stream>>>(A);
  }

  // The FE_Q restriction matrices applied to a level vector (minimal_surface/program.cc:417-456): the coarse vector
  // interpolates the fine function in the coarse support points.  One block per coarse cell: the (2p+1)^3 fine values of
  // its eight children, R1 along z, y, x, all (p+1)^3 coarse entries written (entries shared with a neighbour get the
  // same value from both: R1 selects the face value in the normal direction).
  template <int p>
  __global__ void
  k_interpolate_to_coarse(const TransferArgs<float> A)
  {
    constexpr int    n = p + 1, nf = 2 * p + 1;
    __shared__ float buf0[nf * nf * nf], buf1[n * nf * nf];
    const int        tid = threadIdx.x, nt = blockDim.x;
    const uint32_t   ccell = blockIdx.x;
    int              cx, cy, cz;
    cell_coords(A.coarse, ccell, cx, cy, cz);
    for (int l = tid; l < nf * nf * nf; l += nt)
      {
        const int      I = l % nf, J = (l / nf) % nf, K = l / (nf * nf);
        const int      fx = 2 * cx + (I > p), fy = 2 * cy + (J > p), fz = 2 * cz + (K > p);
        const uint32_t fcell = cell_number(A.fine, fx, fy, fz);
        bool           c;
        const uint32_t d = cell_dof(A.fine, fcell, I - (I > p) * p, J - (J > p) * p, K - (K > p) * p, c);
        buf0[l]          = A.src[d];
      }
    __syncthreads();
    for (int l = tid; l < n * nf * nf; l += nt) // z
      {
        const int IJ = l % (nf * nf), k = l / (nf * nf);
        double    s  = 0;
        for (int K = 0; K < nf; ++K)
          s += ctab.R1[k * nf + K] * (double)buf0[K * nf * nf + IJ];
        buf1[l] = (float)s;
      }
    __syncthreads();
    for (int l = tid; l < n * n * nf; l += nt) // y: buf0[(k n + j) nf + I]
      {
        const int I = l % nf, j = (l / nf) % n, k = l / (nf * n);
        double    s = 0;
        for (int Jf = 0; Jf < nf; ++Jf)
          s += ctab.R1[j * nf + Jf] * (double)buf1[(k * nf + Jf) * nf + I];
        buf0[l] = (float)s;
      }
    __syncthreads();
    for (int l = tid; l < n * n * n; l += nt) // x
      {
        const int i = l % n, kj = l / n;
        double    s = 0;
        for (int If = 0; If < nf; ++If)
          s += ctab.R1[i * nf + If] * (double)buf0[kj * nf + If];
        bool           c;
        const uint32_t d = cell_dof(A.coarse, ccell, i, kj % n, kj / n, c);
        A.dst[d]         = (float)s;
      }
  }

  template <int p>
  void
  launch_interpolate_to_coarse(const TransferArgs<float> &A, cudaStream_t stream)
  {
    k_interpolate_to_coarse<p><<<A.coarse.n_cells, 128, 0, stream>>>(A);
  }

  // ------------------------------------------------------------------ right-hand side values
  template <int p>
  __global__ void
  k_rhs_values(const RhsArgs A)
  {
    constexpr int     n  = p + 1;
    constexpr int     n3 = n * n * n;
    __shared__ double v[n3];
    __shared__ double X[3 * n3]; // mapped cells: support points of the cell geometry
    const uint32_t    cell = blockIdx.x;
    int               cx, cy, cz;
    cell_coords(A.L, cell, cx, cy, cz);
    const double detJ = A.L.h[0] * A.L.h[1] * A.L.h[2];
    const bool   mapped = A.geom.kind != 0;
    if (mapped)
      {
        cell_support_points<p>(A.geom, A.L, cx, cy, cz, X);
        __syncthreads();
      }
    for (int q = threadIdx.x; q < n3; q += blockDim.x)
      {
        const int    q0 = q % n, q1 = (q / n) % n, q2 = q / (n * n);
        if (mapped)
          {
            double x[3], J[3][3];
            mapped_point<p>(X, q0, q1, q2, x, J);
            const double f = A.values ? A.values[(size_t)cell * n3 + q] : eval_fn(A.fn_id, x[0], x[1], x[2]);
            v[q]           = f * ctab.w[q0] * ctab.w[q1] * ctab.w[q2] * fabs(det3(J));
            continue;
          }
        const double x = A.L.lo[0] + (cx + ctab.xq[q0]) * A.L.h[0];
        const double y = A.L.lo[1] + (cy + ctab.xq[q1]) * A.L.h[1];
        const double z = A.L.lo[2] + (cz + ctab.xq[q2]) * A.L.h[2];
        const double f = A.values ? A.values[(size_t)cell * n3 + q] : eval_fn(A.fn_id, x, y, z);
        v[q]           = f * ctab.w[q0] * ctab.w[q1] * ctab.w[q2] * detJ;
      }
    __syncthreads();
    for (int l = threadIdx.x; l < n3; l += blockDim.x)
      {
        const int i0 = l % n, i1 = (l / n) % n, i2 = l / (n * n);
        double    sum = 0;
        for (int q2 = 0; q2 < n; ++q2)
          for (int q1 = 0; q1 < n; ++q1)
            {
              const double s12 = ctab.S[q1 * n + i1] * ctab.S[q2 * n + i2];
              for (int q0 = 0; q0 < n; ++q0)
                sum += ctab.S[q0 * n + i0] * s12 * v[(q2 * n + q1) * n + q0];
            }
        bool           con;
        const uint32_t d = cell_dof(A.L, cell, i0, i1, i2, con);
        if (!con)
          atomicAdd(&A.dst[d], sum);
      }
  }

  template <int p>
  void
  launch_rhs_values(const RhsArgs &A, cudaStream_t stream)
  {
    constexpr int n3 = (p + 1) * (p + 1) * (p + 1);
    k_rhs_values<p><<<A.L.n_cells, n3 < 256 ? ((n3 + 31) / 32) * 32 : 256, 0, stream>>>(A);
  }

  // ------------------------------------------------------------------ quadrature points
  template <int p>
  __global__ void
  k_quadrature_points(const PointsArgs A)
  {
    constexpr int     n = p + 1, n3 = n * n * n;
    __shared__ double X[3 * n3];
    const uint32_t    cell = blockIdx.x;
    int               cx, cy, cz;
    cell_coords(A.L, cell, cx, cy, cz);
    const bool mapped = A.geom.kind != 0;
    if (mapped)
      {
        cell_support_points<p>(A.geom, A.L, cx, cy, cz, X);
        __syncthreads();
      }
    for (int q = threadIdx.x; q < n3; q += blockDim.x)
      {
        const int q0 = q % n, q1 = (q / n) % n, q2 = q / (n * n);
        double    x[3] = {A.L.lo[0] + (cx + ctab.xq[q0]) * A.L.h[0], A.L.lo[1] + (cy + ctab.xq[q1]) * A.L.h[1],
                          A.L.lo[2] + (cz + ctab.xq[q2]) * A.L.h[2]};
        if (mapped)
          {
            double J[3][3];
            mapped_point<p>(X, q0, q1, q2, x, J);
          }
        double *o = A.xyz + ((size_t)cell * n3 + q) * 3;
        o[0]      = x[0];
        o[1]      = x[1];
        o[2]      = x[2];
      }
  }

  template <int p>
  void
  launch_quadrature_points(const PointsArgs &A, cudaStream_t stream)
  {
    constexpr int n3 = (p + 1) * (p + 1) * (p + 1);
    k_quadrature_points<p><<<A.L.n_cells, n3 < 128 ? ((n3 + 31) / 32) * 32 : 128, 0, stream>>>(A);
  }

  // ------------------------------------------------------------------ L2 error
  template <int p>
  __global__ void
  k_l2_error(const ErrorArgs A)
  {
    constexpr int     n  = p + 1;
    constexpr int     n3 = n * n * n;
    __shared__ double u[n3];
    __shared__ double X[3 * n3]; // mapped cells: support points of the cell geometry
    __shared__ double red[2][32];
    const uint32_t    cell = blockIdx.x;
    int               cx, cy, cz;
    cell_coords(A.L, cell, cx, cy, cz);
    if (cz < A.L.zg)
      return; // ghost layer of a z-slab: integrated by the rank below
    const bool mapped = A.geom.kind != 0;
    if (mapped)
      cell_support_points<p>(A.geom, A.L, cx, cy, cz, X);
    for (int l = threadIdx.x; l < n3; l += blockDim.x)
      {
        bool           con;
        const uint32_t d = cell_dof(A.L, cell, l % n, (l / n) % n, l / (n * n), con);
        u[l]             = A.u[d];
      }
    __syncthreads();
    const double detJ = A.L.h[0] * A.L.h[1] * A.L.h[2];
    double       err = 0, vol = 0;
    for (int q = threadIdx.x; q < n3; q += blockDim.x)
      {
        const int q0 = q % n, q1 = (q / n) % n, q2 = q / (n * n);
        double    val = 0;
        for (int i2 = 0; i2 < n; ++i2)
          for (int i1 = 0; i1 < n; ++i1)
            {
              const double s12 = ctab.S[q1 * n + i1] * ctab.S[q2 * n + i2];
              for (int i0 = 0; i0 < n; ++i0)
                val += ctab.S[q0 * n + i0] * s12 * u[(i2 * n + i1) * n + i0];
            }
        double x    = A.L.lo[0] + (cx + ctab.xq[q0]) * A.L.h[0];
        double y    = A.L.lo[1] + (cy + ctab.xq[q1]) * A.L.h[1];
        double z    = A.L.lo[2] + (cz + ctab.xq[q2]) * A.L.h[2];
        double JxW  = ctab.w[q0] * ctab.w[q1] * ctab.w[q2] * detJ;
        if (mapped)
          {
            double xm[3], J[3][3];
            mapped_point<p>(X, q0, q1, q2, xm, J);
            x   = xm[0];
            y   = xm[1];
            z   = xm[2];
            JxW = ctab.w[q0] * ctab.w[q1] * ctab.w[q2] * fabs(det3(J));
          }
        const double diff = val - (A.values ? A.values[(size_t)cell * n3 + q] : eval_fn(A.fn_id, x, y, z));
        err += diff * diff * JxW;
        vol += JxW;
      }
    for (int o = 16; o > 0; o >>= 1)
      {
        err += __shfl_down_sync(0xffffffffu, err, o);
        vol += __shfl_down_sync(0xffffffffu, vol, o);
      }
    const int warp = threadIdx.x / 32, lane = threadIdx.x % 32;
    if (lane == 0)
      {
        red[0][warp] = err;
        red[1][warp] = vol;
      }
    __syncthreads();
    if (threadIdx.x == 0)
      {
        const int nw = (blockDim.x + 31) / 32;
        double    e = 0, v = 0;
        for (int i = 0; i < nw; ++i)
          {
            e += red[0][i];
            v += red[1][i];
          }
        atomicAdd(&A.accum[0], e);
        atomicAdd(&A.accum[1], v);
      }
  }

  template <int p>
  void
  launch_l2_error(const ErrorArgs &A, cudaStream_t stream)
  {
    constexpr int n3 = (p + 1) * (p + 1) * (p + 1);
    k_l2_error<p><<<A.L.n_cells, n3 < 256 ? ((n3 + 31) / 32) * 32 : 256, 0, stream>>>(A);
  }

  // ------------------------------------------------------------------ constant upload
  template <int p>
  void
  init_constants(const Basis &b)
  {
    constexpr int       n = p + 1;
    ConstTables<p + 1> *h = new ConstTables<p + 1>();
    for (int i = 0; i < n * n; ++i)
      {
        h->S[i]   = b.S[i];
        h->D[i]   = b.D[i];
        h->G[i]   = b.G[i];
        h->M1[i]  = b.M1[i];
        h->K1[i]  = b.K1[i];
        h->Sf[i]  = (float)b.S[i];
        h->Df[i]  = (float)b.D[i];
        h->M1f[i] = (float)b.M1[i];
        h->K1f[i] = (float)b.K1[i];
        h->MKf[i] = make_float2((float)b.M1[i], (float)b.K1[i]);
      }
    for (int i = 0; i < n; ++i)
      {
        h->w[i]     = b.wq[i];
        h->wf[i]    = (float)b.wq[i];
        h->xq[i]    = b.xq[i];
        h->nodes[i] = b.nodes[i];
      }
    for (int i = 0; i < (2 * p + 1) * n; ++i)
      {
        h->P1[i]  = b.P1[i];
        h->P1f[i] = (float)b.P1[i];
        h->R1[i]  = b.R1[i];
      }
    cudaMemcpyToSymbol(ctab, h, sizeof(*h));
    delete h;
  }
} // namespace mgb
