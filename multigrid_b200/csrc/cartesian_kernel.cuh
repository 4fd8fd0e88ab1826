// Cartesian constant-coefficient operator apply: one thread block streams one tile column.
//
// On axis-aligned cells with a constant coefficient the merged coefficient of
// common/laplace_operator.h:374-387 is one diagonal tensor for all cells, so the cell operator
// evaluate -> q-point op -> integrate (laplace_operator.h:544-551) collapses to
//     K_cell = c0 K1 x M1 x M1 + c1 M1 x K1 x M1 + c2 M1 x M1 x K1          (x fastest, left factor)
// with the 1-D reference matrices M1 = S^T W S, K1 = G^T W G and c_d = coef[d] (= detJ / h_d^2 * c),
// and the global operator to the same expression in the assembled 1-D matrices Mx, Kx, ... of the
// grid lines.  A block owns the TW x TW points of a tile (level.h) for a range of z-planes and
// computes the COMPLETE result there ("owner computes"): it reads the source on its tile plus a halo
// of one cell towards lower x/y and one point towards higher x/y, and evaluates per batch of p planes
//     x-stage:  mx = Mx u, kx = Kx u                 rows of the haloed tile        (shared memory)
//     y-stage:  P = My mx, Q = c0 My kx + c1 Ky mx   owned points                   (shared memory)
//     z-stage:  y = Mz Q + c2 Kz P                   cell layer by cell layer, the line data and
//                                                    the carry between layers in registers
// x- and y-stage are written row-wise ("p output rows of one cell plus the last row of the previous
// cell"), so neighbouring tasks never add into the same point.  Because every result is complete
// when it is stored, the vector operation of the caller (plain store, residual, Chebyshev update,
// cf. the pre/post range operations of laplace_operator.h:607-741) is fused into the same pass and
// each vector is read / written exactly once from HBM; there are no atomics and no index table.
#pragma once
#include "degree_kernels.cuh"

namespace mgb
{
  template <int p>
  struct TileCfg
  {
    static constexpr int n  = p + 1;
    static constexpr int CW = p == 1 ? 16 : p == 2 ? 8 : p == 3 ? 5 : p == 4 ? 4 : p == 5 ? 3 : 2;
    static constexpr int TW = CW * p;
    static constexpr int NI = TW + p + 1;                  // haloed points per direction: [-p, TW]
    static constexpr int NT = ((TW * TW + 31) / 32) * 32;  // threads per block
    static constexpr int NL = (NI * NI + NT - 1) / NT;     // source points per thread and plane
    // shared memory (elements): source planes | mx, kx | P, Q
    static constexpr int U_SIZE  = p * NI * NI;
    static constexpr int MX_SIZE = p * TW * NI;
    static constexpr int PQ_SIZE = p * TW * TW;
    static constexpr int SMEM    = U_SIZE + 2 * MX_SIZE + 2 * PQ_SIZE;
  };

  // p output rows of cell-row products: out[r] = sum_i A[r][i] x[p+i] (own cell, if it exists)
  //                                     out[0] += sum_i A[p][i] x[i]   (previous cell, if it exists)
  template <int p, typename T>
  __device__ __forceinline__ void
  row_product(const T *__restrict__ Ac, const T (&x)[2 * p + 1], const T wl, const T wr, T (&out)[p])
  {
    constexpr int n = p + 1;
#pragma unroll
    for (int r = 0; r < p; ++r)
      {
        T s = 0;
#pragma unroll
        for (int i = 0; i < n; ++i)
          s += Ac[r * n + i] * x[p + i];
        out[r] = wr * s;
      }
    T s = 0;
#pragma unroll
    for (int i = 0; i < n; ++i)
      s += Ac[p * n + i] * x[i];
    out[0] += wl * s;
  }

  template <int p, typename T>
  __global__ void
  __launch_bounds__(TileCfg<p>::NT) k_apply_tile(const ApplyArgs<T> A, const int layers_per_segment)
  {
    using C           = TileCfg<p>;
    constexpr int n   = C::n;
    constexpr int CW  = C::CW;
    constexpr int TW  = C::TW;
    constexpr int NI  = C::NI;
    constexpr int NT  = C::NT;
    constexpr int NL  = C::NL;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    T *u_s  = reinterpret_cast<T *>(smem_raw); // [plane][row J'+p][col I'+p]
    T *mx_s = u_s + C::U_SIZE;                 // [plane][I][row J'+p]
    T *kx_s = mx_s + C::MX_SIZE;
    T *P_s  = kx_s + C::MX_SIZE;               // [plane][J][I]
    T *Q_s  = P_s + C::PQ_SIZE;

    const LevelDesc &L   = A.L;
    const int        tid = threadIdx.x;
    const int        ntiles = L.nt[0] * L.nt[1];
    const int        tile = blockIdx.x % ntiles, seg = blockIdx.x / ntiles;
    const int        tx = tile % L.nt[0], ty = tile / L.nt[0];
    const int        gx0 = tx * TW, gy0 = ty * TW;
    const int        wx = tile_width(L, 0, tx), wy = tile_width(L, 1, ty);
    const int        ncz = L.ncell[2];
    const int        c0 = seg * layers_per_segment;
    const int        c1 = imin(c0 + layers_per_segment, ncz);
    const bool       plain = A.read_plain != 0;

    // ---- per-thread source points (fixed over the planes)
    uint32_t src_off[NL];
    int      src_stride[NL];
    bool     src_ok[NL];
#pragma unroll
    for (int l = 0; l < NL; ++l)
      {
        const int q  = tid + l * NT;
        const int gi = gx0 + q % NI - p, gj = gy0 + q / NI - p;
        bool      ok = q < NI * NI && gi >= 0 && gi < L.nd[0] && gj >= 0 && gj < L.nd[1];
        if (ok && !plain && is_constrained_xy(L, gi, gj))
          ok = false; // Dirichlet in x or y; the z-faces are handled per plane
        src_ok[l]     = ok;
        src_off[l]    = 0;
        src_stride[l] = 0;
        if (ok)
          {
            const int txs = gi / TW, tys = gj / TW;
            src_off[l]    = dof_index(L, gi, gj, 0);
            src_stride[l] = tile_width(L, 0, txs) * tile_width(L, 1, tys);
          }
      }
    const bool z_lo_dirichlet = (L.dirichlet & 16u) != 0, z_hi_dirichlet = (L.dirichlet & 32u) != 0;

    // ---- owned point of this thread
    const int  I = tid % TW, J = tid / TW;
    const bool owner = tid < TW * TW && I < wx && J < wy;
    const int  gi = gx0 + I, gj = gy0 + J;
    const bool xy_constrained = owner && is_constrained_xy(L, gi, gj);
    const uint32_t own_off    = tile_base(L, tx, ty) + (uint32_t)(J * wx + I);
    const int      own_stride = wx * wy;

    const T *M1 = Tab<T>::M1();
    const T *K1 = Tab<T>::K1();
    const T  cx = A.coef[0], cy = A.coef[1], cz = A.coef[2];

    T Pz[n], Qz[n]; // P, Q of this thread's column on the planes of the current cell layer
    T carry = 0;

    // two-dimensional stages for `np` planes starting at plane k_first; results land in P_s / Q_s
    auto planes_2d = [&](const int k_first, const int np) {
      // (a) source planes -> shared memory (constrained and outside points read as zero)
      for (int b = 0; b < np; ++b)
        {
          const int  gk     = k_first + b;
          const bool z_zero = !plain && ((gk == 0 && z_lo_dirichlet) || (gk == L.nd[2] - 1 && z_hi_dirichlet));
#pragma unroll
          for (int l = 0; l < NL; ++l)
            {
              const int q = tid + l * NT;
              if (q < NI * NI)
                u_s[b * NI * NI + q] =
                  (src_ok[l] && !z_zero) ? A.src[src_off[l] + (uint32_t)gk * (uint32_t)src_stride[l]] : T(0);
            }
        }
      __syncthreads();
      // (b) x-stage: task = (cell column cxl, row, plane)
      for (int t = tid; t < CW * NI * np; t += NT)
        {
          const int rp = t % (NI * np), cxl = t / (NI * np);
          const int b = rp / NI, row = rp - b * NI;
          const int gc = tx * CW + cxl; // global cell of the output rows
          const T   wl = gc - 1 >= 0 && gc - 1 < L.ncell[0] ? T(1) : T(0);
          const T   wr = gc < L.ncell[0] ? T(1) : T(0);
          T         x[2 * p + 1], om[p], ok[p];
          const T  *up = u_s + b * NI * NI + row * NI + cxl * p;
#pragma unroll
          for (int i = 0; i < 2 * p + 1; ++i)
            x[i] = up[i];
          row_product<p, T>(M1, x, wl, wr, om);
          row_product<p, T>(K1, x, wl, wr, ok);
#pragma unroll
          for (int r = 0; r < p; ++r)
            {
              mx_s[(b * TW + cxl * p + r) * NI + row] = om[r];
              kx_s[(b * TW + cxl * p + r) * NI + row] = ok[r];
            }
        }
      __syncthreads();
      // (c) y-stage: task = (I, cell row cyl, plane)
      for (int t = tid; t < TW * CW * np; t += NT)
        {
          const int Il = t % TW, rest = t / TW;
          const int cyl = rest % CW, b = rest / CW;
          const int gc  = ty * CW + cyl;
          const T   wl = gc - 1 >= 0 && gc - 1 < L.ncell[1] ? T(1) : T(0);
          const T   wr = gc < L.ncell[1] ? T(1) : T(0);
          T         xm[2 * p + 1], xk[2 * p + 1], oP[p], oQ[p], oR[p];
          const T  *mp = mx_s + (b * TW + Il) * NI + cyl * p;
          const T  *kp = kx_s + (b * TW + Il) * NI + cyl * p;
#pragma unroll
          for (int i = 0; i < 2 * p + 1; ++i)
            {
              xm[i] = mp[i];
              xk[i] = kp[i];
            }
          row_product<p, T>(M1, xm, wl, wr, oP);
          row_product<p, T>(M1, xk, wl, wr, oQ);
          row_product<p, T>(K1, xm, wl, wr, oR);
#pragma unroll
          for (int r = 0; r < p; ++r)
            {
              P_s[(b * TW + cyl * p + r) * TW + Il] = oP[r];
              Q_s[(b * TW + cyl * p + r) * TW + Il] = cx * oQ[r] + cy * oR[r];
            }
        }
      __syncthreads();
    };

    // ---- stores one completed plane with the fused vector operation
    auto store_plane = [&](const int gk, T y) {
      if (!owner)
        return;
      const uint32_t d = own_off + (uint32_t)gk * (uint32_t)own_stride;
      const bool constrained =
        xy_constrained || (gk == 0 && z_lo_dirichlet) || (gk == L.nd[2] - 1 && z_hi_dirichlet);
      if (A.mode == APPLY_ADD)
        {
          if (!constrained)
            A.dst[d] += y;
          return;
        }
      if (constrained)
        y = A.src[d]; // identity rows (laplace_operator.h:592-593, :632-633)
      if (A.mode == APPLY_SET)
        A.dst[d] = y;
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
    };

    // ---- march through the cell layers; layer c_begin < c0 only produces the carry
    const int c_begin = c0 > 0 ? c0 - 1 : 0;
    planes_2d(c_begin * p, 1);
    Pz[0] = owner ? P_s[J * TW + I] : T(0);
    Qz[0] = owner ? Q_s[J * TW + I] : T(0);
    for (int c = c_begin; c < c1; ++c)
      {
        planes_2d(c * p + 1, p);
#pragma unroll
        for (int b = 0; b < p; ++b)
          {
            Pz[b + 1] = owner ? P_s[(b * TW + J) * TW + I] : T(0);
            Qz[b + 1] = owner ? Q_s[(b * TW + J) * TW + I] : T(0);
          }
        T y[n];
#pragma unroll
        for (int r = 0; r < n; ++r)
          {
            T s = 0, k = 0;
#pragma unroll
            for (int i = 0; i < n; ++i)
              {
                s += M1[r * n + i] * Qz[i];
                k += K1[r * n + i] * Pz[i];
              }
            y[r] = s + cz * k;
          }
        y[0] += carry;
        carry = y[p];
        if (c >= c0)
          {
#pragma unroll
            for (int r = 0; r < p; ++r)
              store_plane(c * p + r, y[r]);
          }
        Pz[0] = Pz[p];
        Qz[0] = Qz[p];
      }
    if (c1 == ncz)
      store_plane(ncz * p, carry);
  }

  template <int p, typename T>
  void
  launch_apply_cartesian(const ApplyArgs<T> &A, cudaStream_t stream)
  {
    using C           = TileCfg<p>;
    const size_t smem = (size_t)C::SMEM * sizeof(T);
    static bool  first = true;
    if (first)
      {
        cudaFuncSetAttribute(k_apply_tile<p, T>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        first = false;
      }
    const int ntiles = A.L.nt[0] * A.L.nt[1];
    const int ncz    = A.L.ncell[2];
    // enough blocks for several waves on 148 SMs, but at least 4 cell layers per segment
    int nseg = (148 * 6 + ntiles - 1) / ntiles;
    if (nseg > ncz / 4)
      nseg = ncz / 4;
    if (nseg < 1)
      nseg = 1;
    const int lps = (ncz + nseg - 1) / nseg;
    nseg          = (ncz + lps - 1) / lps;
    k_apply_tile<p, T><<<ntiles * nseg, C::NT, smem, stream>>>(A, lps);
  }
} // namespace mgb
