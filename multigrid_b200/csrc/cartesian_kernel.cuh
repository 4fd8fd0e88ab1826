// Cartesian constant-coefficient operator apply, one thread block per patch of M^3 cells.
//
// On axis-aligned cells with a constant coefficient the merged coefficient of
// common/laplace_operator.h:374-387 is one diagonal tensor for all cells, so the cell
// operator evaluate -> q-point op -> integrate (laplace_operator.h:544-551) collapses to
//     K_cell = c0 K1 x M1 x M1 + c1 M1 x K1 x M1 + c2 M1 x M1 x K1        (x fastest, left factor)
// with the 1-D reference matrices M1 = S^T W S, K1 = G^T W G and c_d = coef[d] (= detJ / h_d^2 * c).
// Because all cells of a patch share DoFs along common faces, the sum over the patch's cells is
// sum-factorised over whole patch lines: stage 1 applies {M1, c2 K1} cell by cell along every
// z-line of the (M p + 1)^3 patch grid, stage 2 {M1, c1 K1} along y, stage 3 {c0 K1, M1} along x:
//     a = Mz u, b = c2 Kz u;   c = My a, de = c1 Ky a + My b;   y = c0 Kx c + Mx de.
// All three stages run in shared memory, in place, with the line data in registers.
// Patch-interior DoFs are complete inside the block, so the vector operation of the caller
// (plain store, residual, Chebyshev update; cf. the pre/post range operations of
// laplace_operator.h:607-741) is applied in the same pass; patch-boundary DoFs are
// accumulated with red.global.add into `boundary_tmp` and finished by k_boundary_epilogue.
#pragma once
#include "degree_kernels.cuh"

namespace mgb
{
  template <int p, int M>
  struct CartCfg
  {
    static constexpr int n   = p + 1;
    static constexpr int NP  = M * p + 1;
    static constexpr int Q   = NP - 2;
    static constexpr int NPX = NP | 1;
    static constexpr int NT  = ((NP * NP + 31) / 32) * 32;
    static constexpr int NB  = NP * NP * NPX; // elements per buffer
  };

  // y[0..p] = Mat x  with Mat (n x n) in constant memory
  template <int n, typename T>
  __device__ __forceinline__ void
  mv(const T *__restrict__ Mc, const T (&x)[n], T (&y)[n])
  {
    mat_vec<n, T, false, false>(Mc, x, y);
  }

  template <int p, int M, typename T>
  __global__ void
  __launch_bounds__(CartCfg<p, M>::NT, sizeof(T) == 8 ? 2 : 3) k_apply_cartesian(const ApplyArgs<T> A)
  {
    using C           = CartCfg<p, M>;
    constexpr int n   = C::n;
    constexpr int NP  = C::NP;
    constexpr int Q   = C::Q;
    constexpr int NPX = C::NPX;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    T *buf0 = reinterpret_cast<T *>(smem_raw);
    T *buf1 = buf0 + C::NB;
#define PIDX(I, J, K) (((K)*NP + (J)) * NPX + (I))

    const int      tid   = threadIdx.x;
    const uint32_t patch = blockIdx.x;
    const uint32_t mask  = A.L.ent_mask[patch];
    const uint32_t *eb   = A.L.ent_base + (size_t)patch * 27;

    // ---- stage 0: coalesced gather of the 27 patch entities
#pragma unroll 1
    for (int e = 0; e < 27; ++e)
      {
        const int      ex = e % 3, ey = (e / 3) % 3, ez = e / 9;
        const int      sx = ex == 1 ? Q : 1, sy = ey == 1 ? Q : 1, sz = ez == 1 ? Q : 1;
        const int      cnt  = sx * sy * sz;
        const uint32_t base = eb[e];
        const bool     con  = (mask >> e) & 1u;
        for (int off = tid; off < cnt; off += C::NT)
          {
            const int i = off % sx, j = (off / sx) % sy, k = off / (sx * sy);
            const int I = ex == 0 ? 0 : (ex == 2 ? NP - 1 : 1 + i);
            const int J = ey == 0 ? 0 : (ey == 2 ? NP - 1 : 1 + j);
            const int K = ez == 0 ? 0 : (ez == 2 ? NP - 1 : 1 + k);
            buf0[PIDX(I, J, K)] = con ? T(0) : A.src[base + off];
          }
      }
    __syncthreads();

    const T *M1 = Tab<T>::M1();
    const T *K1 = Tab<T>::K1();
    const T  c0 = A.coef[0], c1 = A.coef[1], c2 = A.coef[2];
    const int la = tid % NP, lb = tid / NP; // line coordinates of this thread
    const bool has_line = tid < NP * NP;

    // ---- stage 1: z-lines, thread (I = la, J = lb)
    if (has_line)
      {
        T x[n], ya[n], yb[n];
        T carry_a = 0, carry_b = 0;
        x[0]      = buf0[PIDX(la, lb, 0)];
#pragma unroll
        for (int c = 0; c < M; ++c)
          {
#pragma unroll
            for (int k = 1; k < n; ++k)
              x[k] = buf0[PIDX(la, lb, c * p + k)];
            mv<n, T>(M1, x, ya);
            mv<n, T>(K1, x, yb);
            ya[0] += carry_a;
            yb[0] = c2 * yb[0] + carry_b;
#pragma unroll
            for (int k = 0; k < p; ++k)
              {
                buf0[PIDX(la, lb, c * p + k)] = ya[k];
                buf1[PIDX(la, lb, c * p + k)] = k == 0 ? yb[0] : c2 * yb[k];
              }
            carry_a = ya[p];
            carry_b = c2 * yb[p];
            x[0]    = x[p];
          }
        buf0[PIDX(la, lb, M * p)] = carry_a;
        buf1[PIDX(la, lb, M * p)] = carry_b;
      }
    __syncthreads();

    // ---- stage 2: y-lines, thread (I = la, K = lb)
    if (has_line)
      {
        T xa[n], xb[n], yc[n], yd[n], ye[n];
        T carry_c = 0, carry_d = 0;
        xa[0]     = buf0[PIDX(la, 0, lb)];
        xb[0]     = buf1[PIDX(la, 0, lb)];
#pragma unroll
        for (int c = 0; c < M; ++c)
          {
#pragma unroll
            for (int k = 1; k < n; ++k)
              {
                xa[k] = buf0[PIDX(la, c * p + k, lb)];
                xb[k] = buf1[PIDX(la, c * p + k, lb)];
              }
            mv<n, T>(M1, xa, yc);
            mv<n, T>(K1, xa, yd);
            mv<n, T>(M1, xb, ye);
#pragma unroll
            for (int k = 0; k < n; ++k)
              yd[k] = c1 * yd[k] + ye[k];
            yc[0] += carry_c;
            yd[0] += carry_d;
#pragma unroll
            for (int k = 0; k < p; ++k)
              {
                buf0[PIDX(la, c * p + k, lb)] = yc[k];
                buf1[PIDX(la, c * p + k, lb)] = yd[k];
              }
            carry_c = yc[p];
            carry_d = yd[p];
            xa[0]   = xa[p];
            xb[0]   = xb[p];
          }
        buf0[PIDX(la, M * p, lb)] = carry_c;
        buf1[PIDX(la, M * p, lb)] = carry_d;
      }
    __syncthreads();

    // ---- stage 3: x-lines, thread (J = la, K = lb)
    if (has_line)
      {
        T xc[n], xd[n], y0[n], y1[n];
        T carry = 0;
        xc[0]   = buf0[PIDX(0, la, lb)];
        xd[0]   = buf1[PIDX(0, la, lb)];
#pragma unroll
        for (int c = 0; c < M; ++c)
          {
#pragma unroll
            for (int k = 1; k < n; ++k)
              {
                xc[k] = buf0[PIDX(c * p + k, la, lb)];
                xd[k] = buf1[PIDX(c * p + k, la, lb)];
              }
            mv<n, T>(K1, xc, y0);
            mv<n, T>(M1, xd, y1);
#pragma unroll
            for (int k = 0; k < n; ++k)
              y0[k] = c0 * y0[k] + y1[k];
            y0[0] += carry;
#pragma unroll
            for (int k = 0; k < p; ++k)
              buf0[PIDX(c * p + k, la, lb)] = y0[k];
            carry = y0[p];
            xc[0] = xc[p];
            xd[0] = xd[p];
          }
        buf0[PIDX(M * p, la, lb)] = carry;
      }
    __syncthreads();

    // ---- stage 4: coalesced epilogue
#pragma unroll 1
    for (int e = 0; e < 27; ++e)
      {
        const bool con = (mask >> e) & 1u;
        if (con)
          continue;
        const int      ex = e % 3, ey = (e / 3) % 3, ez = e / 9;
        const int      sx = ex == 1 ? Q : 1, sy = ey == 1 ? Q : 1, sz = ez == 1 ? Q : 1;
        const int      cnt  = sx * sy * sz;
        const uint32_t base = eb[e];
        for (int off = tid; off < cnt; off += C::NT)
          {
            const int i = off % sx, j = (off / sx) % sy, k = off / (sx * sy);
            const int I = ex == 0 ? 0 : (ex == 2 ? NP - 1 : 1 + i);
            const int J = ey == 0 ? 0 : (ey == 2 ? NP - 1 : 1 + j);
            const int K = ez == 0 ? 0 : (ez == 2 ? NP - 1 : 1 + k);
            const T   y = buf0[PIDX(I, J, K)];
            const uint32_t d = base + off;
            if (e == 13)
              {
                if (A.mode == APPLY_SET)
                  A.dst[d] = y;
                else if (A.mode == APPLY_ADD)
                  A.dst[d] += y;
                else if (A.mode == APPLY_RESIDUAL)
                  A.dst[d] = A.rhs[d] - y;
                else
                  {
                    const T xk  = A.src[d];
                    const T upd = A.f2 * A.inv_diag[d] * (A.rhs[d] - y);
                    if (A.first_chebyshev)
                      A.dst[d] = (T(1) + A.f1) * xk + upd;
                    else
                      A.dst[d] = (T(1) + A.f1) * xk - A.f1 * A.dst[d] + upd;
                  }
              }
            else if (A.mode == APPLY_ADD)
              atomicAdd(&A.dst[d], y);
            else
              atomicAdd(&A.boundary_tmp[d - A.L.n_interior], y);
          }
      }
#undef PIDX
  }

  template <int p, int M, typename T>
  void
  launch_apply_cartesian_m(const ApplyArgs<T> &A, cudaStream_t stream)
  {
    using C            = CartCfg<p, M>;
    const size_t smem  = 2 * (size_t)C::NB * sizeof(T);
    static bool  first = true;
    if (first)
      {
        cudaFuncSetAttribute(k_apply_cartesian<p, M, T>,
                             cudaFuncAttributeMaxDynamicSharedMemorySize,
                             (int)smem);
        first = false;
      }
    k_apply_cartesian<p, M, T><<<A.L.n_patches, C::NT, smem, stream>>>(A);
  }

  template <int p, typename T>
  void
  launch_apply_cartesian(const ApplyArgs<T> &A, cudaStream_t stream)
  {
    if (A.L.M == 1)
      launch_apply_cartesian_m<p, 1, T>(A, stream);
    else if (A.L.M == 2)
      launch_apply_cartesian_m<p, 2, T>(A, stream);
    else if (p <= 4)
      launch_apply_cartesian_m<p, (p <= 4 ? 4 : 2), T>(A, stream);
  }
} // namespace mgb
