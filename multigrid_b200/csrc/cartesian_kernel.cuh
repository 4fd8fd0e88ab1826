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
#include <algorithm>
#include <cstdlib>
#include <type_traits>
#include <utility>

#include "degree_kernels.cuh"

#ifndef MGB_MINB64
#define MGB_MINB64 2
#endif
#ifndef MGB_MINB32
#define MGB_MINB32 4
#endif
#ifndef MGB_MINB32C
#define MGB_MINB32C 4
#endif

namespace mgb
{
  // compile-time loop: f(std::integral_constant<int, 0>{}), ..., f(std::integral_constant<int, N-1>{})
  template <typename F, int... Is>
  __device__ __forceinline__ void
  static_for_impl(F &&f, std::integer_sequence<int, Is...>)
  {
    (f(std::integral_constant<int, Is>{}), ...);
  }
  template <int N, typename F>
  __device__ __forceinline__ void
  static_for(F &&f)
  {
    static_for_impl(f, std::make_integer_sequence<int, N>{});
  }

  template <int p>
  struct TileCfg
  {
    static constexpr int n  = p + 1;
    static constexpr int CW = p == 1 ? 16 : p == 2 ? 8 : p == 3 ? 5 : p == 4 ? 4 : p == 5 ? 3 : 2;
    static constexpr int TW = CW * p;
    static constexpr int NI = TW + p + 1;                  // haloed points per direction: [-p, TW]
    static constexpr int NT = ((TW * TW + 31) / 32) * 32;  // threads per block
    static constexpr int NL = (NI * NI + NT - 1) / NT;     // source points per thread and plane
    static constexpr int NXT = CW * NI * p;                // x-stage tasks per batch of p planes
    static constexpr int NXR = (NXT + NT - 1) / NT;        // x-stage rounds
    // shared memory (elements): source planes | mx, kx | P, Q
    static constexpr int U_SIZE  = p * NI * NI;
    static constexpr int MX_SIZE = p * TW * NI;
    static constexpr int PQP     = TW * TW + 16;           // plane pitch of P, Q (bank-conflict free)
    static constexpr int PQ_SIZE = p * PQP;
    static constexpr int SMEM    = 2 * U_SIZE + 2 * MX_SIZE + 2 * PQ_SIZE; // source planes double-buffered
    // packed float kernel: plane pitch of the (mx, kx) pairs such that the x-stage tasks of a warp, which run
    // through (row, plane, cell column), store to consecutive 8-byte slots modulo 16 across a change of plane
    static constexpr int MXP2  = TW * NI + (((NI - TW * NI) % 16) + 16) % 16;
    static constexpr int SMEM2 = 2 * U_SIZE + 2 * p * MXP2 + 2 * PQ_SIZE;  // in floats
  };

  // p output rows of cell-row products: out[r] = sum_i A[r][i] x[p+i] (own cell, if it exists)
  //                                     out[0] += sum_i A[p][i] x[i]   (previous cell, if it exists)
  // Points outside the grid are read as zero, so a missing cell only has to be masked at the shared
  // point x[p]: xl = x[p] if the previous cell exists else 0, xr likewise for the own cell.  (Rows
  // r > 0 of a missing own cell belong to points outside the grid and are never stored.)
  // M1 and K1 are symmetric and persymmetric (A[r][i] = A[i][r] = A[p-r][p-i]); addressing every
  // entry through one canonical representative leaves (p+1)(p+3)/4-ish distinct constants per matrix,
  // few enough to live in uniform registers (the 2 (p+1)^2 separate entries are not)
  template <int p>
  __host__ __device__ constexpr int
  canon(const int r, const int i)
  {
    int a = r, b = i;
    if (b < a)
      {
        const int t = a;
        a           = b;
        b           = t;
      }
    int c = p - i, d = p - r; // (p-r, p-i) transposed so that c <= d is tested below
    if (d < c)
      {
        const int t = c;
        c           = d;
        d           = t;
      }
    if (c < a || (c == a && d < b))
      {
        a = c;
        b = d;
      }
    return a * (p + 1) + b;
  }

  template <int p, typename T>
  __device__ __forceinline__ void
  row_product(const T *__restrict__ Ac, const T (&x)[2 * p + 1], const T xl, const T xr, T (&out)[p])
  {
#pragma unroll
    for (int r = 0; r < p; ++r)
      {
        T s = Ac[canon<p>(r, 0)] * xr;
#pragma unroll
        for (int i = 1; i <= p; ++i)
          s += Ac[canon<p>(r, i)] * x[p + i];
        out[r] = s;
      }
    T s = Ac[canon<p>(p, p)] * xl;
#pragma unroll
    for (int i = 0; i < p; ++i)
      s += Ac[canon<p>(p, i)] * x[i];
    out[0] += s;
  }

  // asynchronous copy of one value global -> shared (LDGSTS); `valid == false` writes zero
  template <typename T>
  __device__ __forceinline__ void
  cp_async_value(T *smem_dst, const T *gmem_src, const bool valid)
  {
    const uint32_t d  = (uint32_t)__cvta_generic_to_shared(smem_dst);
    const int      sz = valid ? (int)sizeof(T) : 0;
    asm volatile("cp.async.ca.shared.global [%0], [%1], %2, %3;\n" ::"r"(d), "l"(gmem_src), "n"(sizeof(T)), "r"(sz));
  }

  // internal mode of the kernel: the first Chebyshev iteration (x_{k-1} = 0 implied) is its own instance
  constexpr int APPLY_CHEBYSHEV_FIRST = 4;
  // dst = A src together with the sum of src[i] dst[i] (first sum of vmult_with_cg_update, laplace_operator.h:698)
  constexpr int APPLY_SET_DOT = 5;
  // Chebyshev with a zero initial guess (deal.II PreconditionChebyshev::vmult): the iterate x_1 = f0 P^-1 b is never
  // stored.  ZERO is iteration index 1 with the source x_1 formed on the fly from the staged planes of b and P^-1
  // (x_0 = 0 implied); DERIVED is iteration index 2, whose x_{k-1} = x_1 is recomputed from the operands b and P^-1
  // it loads anyway.  Same float operations as the stored variant, (f0 * P^-1) * b, so the results are bit-identical.
  constexpr int APPLY_CHEBYSHEV_ZERO    = 6;
  constexpr int APPLY_CHEBYSHEV_DERIVED = 7;
  // q = A p with the four sums of vmult_with_cg_update taken while the result is stored (laplace_operator.h:698-701):
  // q.p, r.r, q.r, q.q (r passed as `rhs`); Dirichlet rows of q stay zero as in the reference's merged loop
  constexpr int APPLY_SET_DOT4 = 8;

  // Operands of the fused vector operation (the post-range operation of laplace_operator.h:607-741) on
  // COUNT planes of one point column: loaded early (load_operands, before the y-stage) so that their
  // latency is covered by the stages, consumed when the planes are complete (store_planes).
  // `d0` = index of the first plane, consecutive planes are `stride` apart; bit r of `con` flags plane r
  // as a Dirichlet row (only looked at when BND).
  template <typename T, int COUNT>
  struct Operands
  {
    T xs[COUNT], rs[COUNT], ds[COUNT], os[COUNT];
  };

  template <int p, typename T, int MODE, int COUNT, bool BND>
  __device__ __forceinline__ void
  load_operands(const ApplyArgs<T> &A, Operands<T, COUNT> &o, const uint32_t d0, const uint32_t stride,
                const unsigned int con)
  {
    const T *__restrict__ src = A.src;
    const T *__restrict__ rhs = A.rhs;
    const T *__restrict__ inv = A.inv_diag;
    const T *__restrict__ dst = A.dst;
    constexpr bool CHEB = MODE == APPLY_CHEBYSHEV || MODE == APPLY_CHEBYSHEV_FIRST || MODE == APPLY_CHEBYSHEV_ZERO ||
                          MODE == APPLY_CHEBYSHEV_DERIVED;
#pragma unroll
    for (int r = 0; r < COUNT; ++r)
      {
        const uint32_t d = d0 + r * stride;
        if (MODE == APPLY_ADD)
          o.os[r] = dst[d];
        else
          {
            if (MODE != APPLY_CHEBYSHEV_ZERO &&
                (CHEB || MODE == APPLY_SET_DOT || MODE == APPLY_SET_DOT4 || (BND && ((con >> r) & 1u))))
              o.xs[r] = src[d];
            if (MODE != APPLY_SET && MODE != APPLY_SET_DOT)
              o.rs[r] = rhs[d];
            if (CHEB)
              o.ds[r] = inv[d];
            if (MODE == APPLY_CHEBYSHEV)
              o.os[r] = dst[d];
          }
      }
  }

  template <int p, typename T, int MODE, int COUNT, bool BND>
  __device__ __forceinline__ double
  store_planes(const ApplyArgs<T> &A, const Operands<T, COUNT> &o, const T (&y)[p + 1], const uint32_t d0,
               const uint32_t stride, const unsigned int con, double *sums4 = nullptr)
  {
    double dot = 0;
    T *__restrict__ dst = A.dst;
    const T f1 = A.f1, f2 = A.f2, f1p = T(1) + A.f1, f0 = A.f0;
#pragma unroll
    for (int r = 0; r < COUNT; ++r)
      {
        const uint32_t d = d0 + r * stride;
        const bool     c = BND && ((con >> r) & 1u);
        T              v = y[r];
        T              xs = o.xs[r], os = o.os[r];
        if (MODE == APPLY_CHEBYSHEV_ZERO)
          xs = (f0 * o.ds[r]) * o.rs[r]; // x_1 at this point, as staged for the stages
        if (MODE == APPLY_CHEBYSHEV_DERIVED)
          os = (f0 * o.ds[r]) * o.rs[r]; // x_{k-1} = x_1
        if (MODE == APPLY_ADD)
          {
            if (!c)
              dst[d] = o.os[r] + v;
            continue;
          }
        if (c)
          v = xs; // identity rows (laplace_operator.h:592-593, :632-633)
        if (MODE == APPLY_SET)
          dst[d] = v;
        else if (MODE == APPLY_SET_DOT)
          {
            dst[d] = v;
            dot += (double)xs * (double)v;
          }
        else if (MODE == APPLY_SET_DOT4)
          {
            const double q = c ? 0. : (double)y[r], rr = (double)o.rs[r];
            dst[d]         = (T)q;
            sums4[0] += q * (double)xs;
            sums4[1] += rr * rr;
            sums4[2] += q * rr;
            sums4[3] += q * q;
          }
        else if (MODE == APPLY_RESIDUAL)
          dst[d] = o.rs[r] - v;
        else
          {
            T t = f2 * o.ds[r] * (o.rs[r] - v) + f1p * xs;
            if (MODE == APPLY_CHEBYSHEV || MODE == APPLY_CHEBYSHEV_DERIVED)
              t -= f1 * os;
            dst[d] = t;
          }
      }
    return dot;
  }

  // asynchronous copy with an immediate shared-memory offset (bytes)
  template <typename T, int OFFSET>
  __device__ __forceinline__ void
  cp_async_imm(const uint32_t smem_base, const T *gmem_src, const int src_bytes)
  {
    asm volatile("cp.async.ca.shared.global [%0 + %1], [%2], %3, %4;\n" ::"r"(smem_base), "n"(OFFSET), "l"(gmem_src),
                 "n"(sizeof(T)), "r"(src_bytes));
  }

  template <int MODE, typename T>
  struct TileMinBlocks
  {
    static constexpr int value =
      sizeof(T) == 8 ? MGB_MINB64 : (MODE >= APPLY_CHEBYSHEV ? MGB_MINB32C : MGB_MINB32);
  };

  template <int p, typename T, int MODE>
  __global__ void
  __launch_bounds__(TileCfg<p>::NT, TileMinBlocks<MODE, T>::value)
    k_apply_tile(const ApplyArgs<T> A, const int layers_per_segment, const int layer_begin, const int layer_end)
  {
    using C           = TileCfg<p>;
    constexpr int n   = C::n;
    constexpr int CW  = C::CW;
    constexpr int TW  = C::TW;
    constexpr int NI  = C::NI;
    constexpr int NT  = C::NT;
    constexpr int NL  = C::NL;
    constexpr int NXR = C::NXR;
    constexpr int PQP = C::PQP;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    T *u_s  = reinterpret_cast<T *>(smem_raw); // [buffer][plane][row J'+p][col I'+p]
    T *mx_s = u_s + 2 * C::U_SIZE;             // [plane][I][row J'+p]
    T *kx_s = mx_s + C::MX_SIZE;
    T *P_s  = kx_s + C::MX_SIZE;               // [plane (pitch PQP)][J][I]
    T *Q_s  = P_s + C::PQ_SIZE;

    const LevelDesc &L      = A.L;
    const int        tid    = threadIdx.x;
    const int        ntiles = L.nt[0] * L.nt[1];
    const int        tile = blockIdx.x % ntiles, seg = blockIdx.x / ntiles;
    const int        tx = tile % L.nt[0], ty = tile / L.nt[0];
    const int        gx0 = tx * TW, gy0 = ty * TW;
    const int        wx = tile_width(L, 0, tx), wy = tile_width(L, 1, ty);
    const int        ncz = L.ncell[2], nd2 = L.nd[2];
    const bool       second = A.layer_begin2 >= 0 && seg == 1; // two single layers in one launch
    const int        c0 = second ? A.layer_begin2 : layer_begin + seg * layers_per_segment;
    const int        c1 = second ? c0 + 1 : imin(c0 + layers_per_segment, layer_end);
    const bool       plain = A.read_plain != 0;

    // ---- per-thread source points (fixed over the planes): running pointer, plane stride, copy size
    const T *src_ptr[NL];
    int      src_stride[NL], src_bytes[NL];
#pragma unroll
    for (int l = 0; l < NL; ++l)
      {
        const int q  = tid + l * NT;
        const int gi = gx0 + q % NI - p, gj = gy0 + q / NI - p;
        bool      ok = q < NI * NI && gi >= 0 && gi < L.nd[0] && gj >= 0 && gj < L.nd[1];
        if (ok && !plain && is_constrained_xy(L, gi, gj))
          ok = false; // Dirichlet in x or y; the z-faces are handled per plane
        src_bytes[l]  = ok ? (int)sizeof(T) : 0; // 0: the copy writes a zero
        src_stride[l] = ok ? (int)plane_pitch(L) : 0;
        src_ptr[l]    = A.src + (ok ? dof_index(L, gi, gj, 0) : 0u);
      }
    // planes that read as zero: outside the grid, or on a Dirichlet z-face
    const int k_lo = (!plain && (L.dirichlet & 16u)) ? 1 : 0;
    const int k_hi = (!plain && (L.dirichlet & 32u)) ? nd2 - 2 : nd2 - 1;

    // ---- x-stage tasks of this thread: (cell column, row, plane) -> p output rows
    int x_in[NXR], x_out[NXR];
    int x_flags = 0; // bit 2r: previous cell exists, bit 2r+1: own cell exists, round r
#pragma unroll
    for (int r = 0; r < NXR; ++r)
      {
        const int t  = tid + r * NT;
        const int rp = t % (NI * p), cxl = t / (NI * p);
        const int b = rp / NI, row = rp - b * NI;
        const int gc = tx * CW + cxl;
        x_in[r]      = t < C::NXT ? rp * NI + cxl * p : -1;
        x_out[r]     = (b * TW + cxl * p) * NI + row;
        if (gc - 1 >= 0 && gc - 1 < L.ncell[0])
          x_flags |= 1 << (2 * r);
        if (gc < L.ncell[0])
          x_flags |= 2 << (2 * r);
      }
    // ---- y-stage task of this thread: (I, plane, cell row) -> p output rows
    int y_in, y_out, y_flags = 0;
    {
      const int Il = tid % TW, rest = tid / TW;
      const int b = rest % p, cyl = rest / p;
      const int gc = ty * CW + cyl;
      y_in         = tid < TW * TW ? (b * TW + Il) * NI + cyl * p : -1;
      y_out        = b * PQP + cyl * p * TW + Il;
      if (gc - 1 >= 0 && gc - 1 < L.ncell[1])
        y_flags |= 1;
      if (gc < L.ncell[1])
        y_flags |= 2;
    }

    // ---- owned point of this thread
    const int  I = tid % TW, J = tid / TW;
    const bool owner = tid < TW * TW && I < wx && J < wy;
    const bool xy_constrained = owner && is_constrained_xy(L, gx0 + I, gy0 + J);
    // block-uniform: does this tile touch a Dirichlet face in x or y?
    const bool tile_bnd = ((L.dirichlet & 1u) && tx == 0) || ((L.dirichlet & 2u) && gx0 + wx == L.nd[0]) ||
                          ((L.dirichlet & 4u) && ty == 0) || ((L.dirichlet & 8u) && gy0 + wy == L.nd[1]);
    const uint32_t own_stride = plane_pitch(L);
    const uint32_t own_off    = tile_base(L, tx, ty) + (uint32_t)(J * wx + I);
    const int      pq_off     = J * TW + I;

    const T *M1 = Tab<T>::M1();
    const T *K1 = Tab<T>::K1();
    const T  cx = A.coef[0], cy = A.coef[1], cz = A.coef[2];

    T Pz[n], Qz[n]; // P, Q of this thread's column on the planes of the current cell layer
    T carry = 0;
    double dot_acc = 0; // APPLY_SET_DOT
    double sums4[MODE == APPLY_SET_DOT4 ? 4 : 1] = {0}; // APPLY_SET_DOT4

    const uint32_t u_base = (uint32_t)__cvta_generic_to_shared(u_s + tid);

    // source planes of batch c (planes c p + 1 .. c p + p) -> shared-memory buffer `buf`, asynchronously
    auto issue_loads = [&](const int c, const int buf) {
      const uint32_t sb = u_base + (uint32_t)(buf * C::U_SIZE * (int)sizeof(T));
      const int      k0 = c * p + 1;
      if (k0 >= k_lo && k0 + p - 1 <= k_hi)
        {
          static_for<NL>([&](auto lc) {
            constexpr int l = decltype(lc)::value;
            if (l + 1 < NL || tid + l * NT < NI * NI)
              {
                const T *g = src_ptr[l] + (size_t)k0 * src_stride[l];
                static_for<p>([&](auto bc) {
                  constexpr int b = decltype(bc)::value;
                  cp_async_imm<T, (b * NI * NI + l * NT) * (int)sizeof(T)>(sb, g + b * src_stride[l], src_bytes[l]);
                });
              }
          });
        }
      else
        {
#pragma unroll
          for (int b = 0; b < p; ++b)
            {
              const int  gk  = k0 + b;
              const bool okz = gk >= k_lo && gk <= k_hi;
#pragma unroll
              for (int l = 0; l < NL; ++l)
                if (tid + l * NT < NI * NI)
                  cp_async_value(u_s + (buf * p + b) * NI * NI + tid + l * NT,
                                 okz ? src_ptr[l] + (size_t)gk * src_stride[l] : A.src, okz && src_bytes[l] != 0);
            }
        }
      asm volatile("cp.async.commit_group;\n" ::);
    };

    // ---- march through the cell layers.  Batch c = planes c p + 1 .. c p + p.  The first batch only
    //      provides the lowest plane of the first layer; layer c_begin < c0 only produces the carry.
    const int c_begin = c0 > 0 ? c0 - 1 : 0;
    uint32_t  d_layer = own_off + (uint32_t)(c0 * p) * own_stride; // index of plane c p at this thread's point
    issue_loads(c_begin - 1, 0);
    for (int c = c_begin - 1; c < c1; ++c)
      {
        // (a) the planes of this batch have landed; prefetch the next batch into the other buffer
        const int cur = (c - c_begin + 1) & 1;
        asm volatile("cp.async.wait_group 0;\n" ::);
        __syncthreads();
        if (c + 1 < c1)
          issue_loads(c + 1, cur ^ 1);
        const T *u_c = u_s + cur * C::U_SIZE;

        // (b) x-stage: mx = Mx u, kx = Kx u
#pragma unroll
        for (int r = 0; r < NXR; ++r)
          if (x_in[r] >= 0)
            {
              T x[2 * p + 1], om[p], ok[p];
#pragma unroll
              for (int i = 0; i < 2 * p + 1; ++i)
                x[i] = u_c[x_in[r] + i];
              const T wl = (x_flags >> (2 * r)) & 1 ? x[p] : T(0);
              const T wr = (x_flags >> (2 * r)) & 2 ? x[p] : T(0);
              row_product<p, T>(M1, x, wl, wr, om);
              row_product<p, T>(K1, x, wl, wr, ok);
#pragma unroll
              for (int i = 0; i < p; ++i)
                {
                  mx_s[x_out[r] + i * NI] = om[i];
                  kx_s[x_out[r] + i * NI] = ok[i];
                }
            }
        __syncthreads();

        // operands of the fused update for this layer: issued now, consumed after the z-stage
        Operands<T, p> ops;
        const bool     bnd = tile_bnd || c == 0; // block-uniform: only then constraints are looked at
        unsigned int   con = 0;
        if (owner && c >= c0)
          {
            if (bnd)
              {
                con = xy_constrained ? 0xffu : 0u;
                if (c == 0 && (L.dirichlet & 16u))
                  con |= 1u;
                load_operands<p, T, MODE, p, true>(A, ops, d_layer, own_stride, con);
              }
            else
              load_operands<p, T, MODE, p, false>(A, ops, d_layer, own_stride, 0u);
          }

        // (c) y-stage: P = My mx, Q = c0 My kx + c1 Ky mx
        if (y_in >= 0)
          {
            T xm[2 * p + 1], oP[p], oQ[p], oR[p];
#pragma unroll
            for (int i = 0; i < 2 * p + 1; ++i)
              xm[i] = mx_s[y_in + i];
            T wl = y_flags & 1 ? xm[p] : T(0);
            T wr = y_flags & 2 ? xm[p] : T(0);
            row_product<p, T>(M1, xm, wl, wr, oP);
            row_product<p, T>(K1, xm, wl, wr, oR);
#pragma unroll
            for (int i = 0; i < p; ++i)
              P_s[y_out + i * TW] = oP[i];
#pragma unroll
            for (int i = 0; i < 2 * p + 1; ++i)
              xm[i] = kx_s[y_in + i];
            wl = y_flags & 1 ? xm[p] : T(0);
            wr = y_flags & 2 ? xm[p] : T(0);
            row_product<p, T>(M1, xm, wl, wr, oQ);
#pragma unroll
            for (int i = 0; i < p; ++i)
              Q_s[y_out + i * TW] = cx * oQ[i] + cy * oR[i];
          }
        __syncthreads();

        // (d) z-stage: y = Mz Q + c2 Kz P on the cell layer c, fused vector operation
        if (!owner)
          continue;
#pragma unroll
        for (int b = 0; b < p; ++b)
          {
            Pz[b + 1] = P_s[b * PQP + pq_off];
            Qz[b + 1] = Q_s[b * PQP + pq_off];
          }
        if (c >= c_begin)
          {
            T y[n];
#pragma unroll
            for (int r = 0; r < n; ++r)
              {
                T s = 0, k = 0;
#pragma unroll
                for (int i = 0; i < n; ++i)
                  {
                    s += M1[canon<p>(r, i)] * Qz[i];
                    k += K1[canon<p>(r, i)] * Pz[i];
                  }
                y[r] = s + cz * k;
              }
            y[0] += carry;
            carry = y[p];
            if (c >= c0)
              {
                if (bnd)
                  dot_acc += store_planes<p, T, MODE, p, true>(A, ops, y, d_layer, own_stride, con, sums4);
                else
                  dot_acc += store_planes<p, T, MODE, p, false>(A, ops, y, d_layer, own_stride, 0u, sums4);
                d_layer += p * own_stride;
              }
          }
        Pz[0] = Pz[p];
        Qz[0] = Qz[p];
      }
    if (c1 == ncz && owner && !L.ztop) // the top plane of the level (a ghost on all but the last z-slab)
      {
        T y[n];
        y[0] = carry;
        const unsigned int con = (xy_constrained || (L.dirichlet & 32u)) ? 1u : 0u;
        Operands<T, 1>     ops;
        load_operands<p, T, MODE, 1, true>(A, ops, d_layer, own_stride, con);
        dot_acc += store_planes<p, T, MODE, 1, true>(A, ops, y, d_layer, own_stride, con, sums4);
      }
    if (MODE == APPLY_SET_DOT4)
      {
        // block sums: warp shuffles, then one atomic per block and sum
        __syncthreads();
        double *red = reinterpret_cast<double *>(smem_raw); // [4][warps]
#pragma unroll
        for (int k = 0; k < 4; ++k)
          {
            double v = sums4[MODE == APPLY_SET_DOT4 ? k : 0];
            for (int o = 16; o > 0; o >>= 1)
              v += __shfl_down_sync(0xffffffffu, v, o);
            if ((tid & 31) == 0)
              red[k * (NT / 32) + (tid >> 5)] = v;
          }
        __syncthreads();
        if (tid < 4)
          {
            double t = 0;
            for (int w = 0; w < NT / 32; ++w)
              t += red[tid * (NT / 32) + w];
            atomicAdd(A.dot_out + tid, t);
          }
      }
    if (MODE == APPLY_SET_DOT)
      {
        // block sum of the per-thread partial sums: warp shuffles, one atomic per block
        __syncthreads(); // every warp is done with the stages: the staging memory is free
        double *red = reinterpret_cast<double *>(smem_raw);
        for (int o = 16; o > 0; o >>= 1)
          dot_acc += __shfl_down_sync(0xffffffffu, dot_acc, o);
        if ((tid & 31) == 0)
          red[tid >> 5] = dot_acc;
        __syncthreads();
        if (tid == 0)
          {
            double t = 0;
            for (int w = 0; w < NT / 32; ++w)
              t += red[w];
            atomicAdd(A.dot_out, t);
          }
      }
  }

  // cartesian_kernel_f32x2.cuh: the float instance on packed fp32 arithmetic
  template <int p, int MODE>
  __global__ void k_apply_tile_f32x2(const ApplyArgs<float> A, const int layers_per_segment, const int layer_begin,
                                     const int layer_end);

  // MGB200_F32X2=0 selects the scalar float kernel (A/B measurements)
  inline bool
  use_f32x2()
  {
    static const bool on = [] {
      const char *e = getenv("MGB200_F32X2");
      return !(e && e[0] == '0');
    }();
    return on;
  }

  // MGB200_SHORT_SEGMENTS=0 keeps segments of >= 4 cell layers on every level (A/B measurements)
  inline bool
  use_short_segments()
  {
    static const bool on = [] {
      const char *e = getenv("MGB200_SHORT_SEGMENTS");
      return !(e && e[0] == '0');
    }();
    return on;
  }

  template <int p, typename T, int MODE>
  void
  launch_apply_tile(const ApplyArgs<T> &A, cudaStream_t stream)
  {
    using C           = TileCfg<p>;
    // the zero-guess Chebyshev instance stages the planes of b and P^-1: one more double-buffered plane array
    const size_t smem  = (size_t)C::SMEM * sizeof(T),
                 smem2 = (size_t)(C::SMEM2 + (MODE == APPLY_CHEBYSHEV_ZERO ? 2 * C::U_SIZE : 0)) * sizeof(float);
    static bool  first = true;
    if (first)
      {
        if constexpr (MODE != APPLY_CHEBYSHEV_ZERO && MODE != APPLY_CHEBYSHEV_DERIVED)
          cudaFuncSetAttribute(k_apply_tile<p, T, MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if constexpr (std::is_same<T, float>::value && MODE != APPLY_SET_DOT4)
          cudaFuncSetAttribute(k_apply_tile_f32x2<p, MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem2);
        first = false;
      }
    const int ntiles = A.L.nt[0] * A.L.nt[1];
    // cell layers of this launch: by default all owned ones (the ghost layer below only feeds the carry)
    const int l0  = A.layer_begin >= 0 ? A.layer_begin : A.L.zg;
    const int l1  = A.layer_end >= 0 ? A.layer_end : A.L.ncell[2];
    const int ncz = l1 - l0;
    if (ncz <= 0)
      return;
    // enough blocks for several waves on 148 SMs, but at least 4 cell layers per segment
    int nseg = (148 * 8 + ntiles - 1) / ntiles;
    if (nseg > ncz / 4)
      nseg = ncz / 4;
    if (nseg < 1)
      nseg = 1;
    // small levels cannot fill the device with segments of >= 4 layers: a block is a serial chain of
    // (layers + 2) batches, so shorter segments (down to one layer, at the price of re-computing the carry
    // layer) cut the latency of the launch-bound coarse levels roughly in half
    const int slots = 148 * TileMinBlocks<MODE, T>::value;
    if (ntiles * nseg < slots && use_short_segments())
      nseg = std::max(nseg, std::min(ncz, slots / ntiles));
    int lps = (ncz + nseg - 1) / nseg;
    nseg    = (ncz + lps - 1) / lps;
    if (A.layer_begin2 >= 0) // the first and the last owned layer of a z-slab: one segment each
      {
        lps  = 1;
        nseg = 2;
      }
    if constexpr (std::is_same<T, float>::value && MODE != APPLY_SET_DOT4)
      if (use_f32x2() || MODE == APPLY_CHEBYSHEV_ZERO || MODE == APPLY_CHEBYSHEV_DERIVED)
        {
          k_apply_tile_f32x2<p, MODE><<<ntiles * nseg, C::NT, smem2, stream>>>(A, lps, l0, l1);
          return;
        }
    if constexpr (MODE != APPLY_CHEBYSHEV_ZERO && MODE != APPLY_CHEBYSHEV_DERIVED)
      k_apply_tile<p, T, MODE><<<ntiles * nseg, C::NT, smem, stream>>>(A, lps, l0, l1);
  }

  template <int p, typename T>
  void
  launch_apply_cartesian(const ApplyArgs<T> &A, cudaStream_t stream)
  {
    switch (A.mode)
      {
        case APPLY_ADD: launch_apply_tile<p, T, APPLY_ADD>(A, stream); break;
        case APPLY_SET:
          if (A.dot_out != nullptr && A.rhs != nullptr) // q = A p with the four sums of vmult_with_cg_update
            {
              launch_apply_tile<p, T, APPLY_SET_DOT4>(A, stream);
              break;
            }
          if constexpr (sizeof(T) == 8)
            if (A.dot_out != nullptr)
              {
                launch_apply_tile<p, T, APPLY_SET_DOT>(A, stream);
                break;
              }
          launch_apply_tile<p, T, APPLY_SET>(A, stream);
          break;
        case APPLY_RESIDUAL: launch_apply_tile<p, T, APPLY_RESIDUAL>(A, stream); break;
        default:
          if constexpr (std::is_same<T, float>::value)
            if (A.first_chebyshev >= 2)
              {
                // zero-guess variants (packed float kernel only; solver.cu never asks for them otherwise)
                if (A.first_chebyshev == 2)
                  launch_apply_tile<p, T, APPLY_CHEBYSHEV_ZERO>(A, stream);
                else
                  launch_apply_tile<p, T, APPLY_CHEBYSHEV_DERIVED>(A, stream);
                break;
              }
          if (A.first_chebyshev)
            launch_apply_tile<p, T, APPLY_CHEBYSHEV_FIRST>(A, stream);
          else
            launch_apply_tile<p, T, APPLY_CHEBYSHEV>(A, stream);
          break;
      }
  }
} // namespace mgb

#include "cartesian_kernel_f32x2.cuh"
