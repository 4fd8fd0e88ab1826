// float instance of k_apply_tile (cartesian_kernel.cuh) on packed fp32 arithmetic (FFMA2 / FMUL2, sm_100a):
// the 1-D products come in pairs that share one operand,
//     x-stage:  (mx, kx) = (Mx, Kx) u          coefficient pair x broadcast value
//     y-stage:  (P, S)   = My (mx, kx)         broadcast coefficient x value pair    (R = Ky mx stays scalar)
//     z-stage:  (Mz Q, Kz P)                   coefficient pair x value pair
// so two FMAs issue as one instruction and the pairs travel through shared memory as 8-byte words (half the
// LDS / STS).  The tile kernel is bound by instruction issue in float (profiles/r1_tile_kernel_ncu_full_summary.txt),
// not by the FMA pipe, which is what this addresses.  Same tiling, staging and fused epilogues as k_apply_tile.
#pragma once
#include "cartesian_kernel.cuh"

namespace mgb
{
  __device__ __forceinline__ float2
  bc2(const float v)
  {
    return make_float2(v, v);
  }

  // row_product (cartesian_kernel.cuh) for the pair of matrices (M1, K1): out[r] = (M row . x, K row . x)
  template <int p>
  __device__ __forceinline__ void
  row_product_mk(const float2 *__restrict__ MK, const float (&x)[2 * p + 1], const float xl, const float xr, float2 (&out)[p])
  {
#pragma unroll
    for (int r = 0; r < p; ++r)
      {
        float2 s = __fmul2_rn(MK[canon<p>(r, 0)], bc2(xr));
#pragma unroll
        for (int i = 1; i <= p; ++i)
          s = __ffma2_rn(MK[canon<p>(r, i)], bc2(x[p + i]), s);
        out[r] = s;
      }
    float2 s = __fmul2_rn(MK[canon<p>(p, p)], bc2(xl));
#pragma unroll
    for (int i = 0; i < p; ++i)
      s = __ffma2_rn(MK[canon<p>(p, i)], bc2(x[i]), s);
    out[0] = __fadd2_rn(out[0], s);
  }

  // row_product for one matrix applied to a pair of vectors: out[r] = A row . (x.x, x.y)
  template <int p>
  __device__ __forceinline__ void
  row_product_bc(const float *__restrict__ Ac, const float2 (&x)[2 * p + 1], const float2 xl, const float2 xr, float2 (&out)[p])
  {
#pragma unroll
    for (int r = 0; r < p; ++r)
      {
        float2 s = __fmul2_rn(bc2(Ac[canon<p>(r, 0)]), xr);
#pragma unroll
        for (int i = 1; i <= p; ++i)
          s = __ffma2_rn(bc2(Ac[canon<p>(r, i)]), x[p + i], s);
        out[r] = s;
      }
    float2 s = __fmul2_rn(bc2(Ac[canon<p>(p, p)]), xl);
#pragma unroll
    for (int i = 0; i < p; ++i)
      s = __ffma2_rn(bc2(Ac[canon<p>(p, i)]), x[i], s);
    out[0] = __fadd2_rn(out[0], s);
  }

  template <int p, int MODE>
  __global__ void
  __launch_bounds__(TileCfg<p>::NT, TileMinBlocks<MODE, float>::value)
    k_apply_tile_f32x2(const ApplyArgs<float> A, const int layers_per_segment, const int layer_begin, const int layer_end)
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
    using T       = float;
    T      *u_s   = reinterpret_cast<T *>(smem_raw);                    // [buffer][plane][row J'+p][col I'+p]
    float2 *mk_s  = reinterpret_cast<float2 *>(u_s + 2 * C::U_SIZE);    // (mx, kx) [plane][I][row J'+p]
    float2 *qp_s  = mk_s + p * C::MXP2;                                // (Q, P)  [plane (pitch PQP)][J][I]
    // zero-guess Chebyshev: u_s stages the planes of b, v_s (same shape, behind everything else) those of P^-1;
    // u_s is turned into x_1 = (f0 P^-1) b in place once the copies have landed
    constexpr bool ZERO  = MODE == APPLY_CHEBYSHEV_ZERO;
    constexpr int  V_OFF = C::SMEM2; // v_s - u_s in floats
    T             *v_s   = u_s + V_OFF;

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
        src_ptr[l]    = (ZERO ? A.rhs : A.src) + (ok ? dof_index(L, gi, gj, 0) : 0u);
      }
    const ptrdiff_t inv_delta = ZERO ? A.inv_diag - A.rhs : 0; // P^-1 sits at the same index as b
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
        x_out[r]     = b * C::MXP2 + cxl * p * NI + row;
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
      y_in         = tid < TW * TW ? b * C::MXP2 + Il * NI + cyl * p : -1;
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

    const T      *M1 = Tab<T>::M1();
    const T      *K1 = Tab<T>::K1();
    const float2 *MK = ctab.MKf; // (M1, K1) entry by entry
    const T  cx = A.coef[0], cy = A.coef[1], cz = A.coef[2];

    float2 QPz[n]; // (Q, P) of this thread's column on the planes of the current cell layer
    T carry = 0;

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
                  if (ZERO)
                    cp_async_imm<T, (V_OFF + b * NI * NI + l * NT) * (int)sizeof(T)>(sb, g + b * src_stride[l] + inv_delta,
                                                                                    src_bytes[l]);
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
                  {
                    const T *g = okz ? src_ptr[l] + (size_t)gk * src_stride[l] : src_ptr[l];
                    cp_async_value(u_s + (buf * p + b) * NI * NI + tid + l * NT, g, okz && src_bytes[l] != 0);
                    if (ZERO)
                      cp_async_value(v_s + (buf * p + b) * NI * NI + tid + l * NT, g + inv_delta, okz && src_bytes[l] != 0);
                  }
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
        asm volatile("cp.async.wait_group 0;\n" ::: "memory");
        if (ZERO)
          {
            // x_1 = (f0 P^-1) b on the elements this thread copied itself (they have landed for it)
            const T f0 = A.f0;
#pragma unroll
            for (int b = 0; b < p; ++b)
#pragma unroll
              for (int l = 0; l < NL; ++l)
                if (l + 1 < NL || tid + l * NT < NI * NI)
                  {
                    const int e = (cur * p + b) * NI * NI + tid + l * NT;
                    u_s[e]      = (f0 * v_s[e]) * u_s[e];
                  }
          }
        __syncthreads();
        if (c + 1 < c1)
          issue_loads(c + 1, cur ^ 1);
        const T *u_c = u_s + cur * C::U_SIZE;

        // (b) x-stage: mx = Mx u, kx = Kx u
#pragma unroll
        for (int r = 0; r < NXR; ++r)
          if (x_in[r] >= 0)
            {
              T      x[2 * p + 1];
              float2 o[p];
#pragma unroll
              for (int i = 0; i < 2 * p + 1; ++i)
                x[i] = u_c[x_in[r] + i];
              const T wl = (x_flags >> (2 * r)) & 1 ? x[p] : T(0);
              const T wr = (x_flags >> (2 * r)) & 2 ? x[p] : T(0);
              row_product_mk<p>(MK, x, wl, wr, o); // (M row . x, K row . x)
#pragma unroll
              for (int i = 0; i < p; ++i)
                mk_s[x_out[r] + i * NI] = o[i];
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

        // (c) y-stage: (P, S) = My (mx, kx) as pairs, R = Ky mx; stored as (Q, P) with Q = c0 S + c1 R
        if (y_in >= 0)
          {
            float2 xm[2 * p + 1], oPS[p];
            T      xs[2 * p + 1], oR[p];
#pragma unroll
            for (int i = 0; i < 2 * p + 1; ++i)
              {
                xm[i] = mk_s[y_in + i];
                xs[i] = xm[i].x;
              }
            const float2 zero2 = make_float2(0.f, 0.f);
            const float2 wl    = y_flags & 1 ? xm[p] : zero2;
            const float2 wr    = y_flags & 2 ? xm[p] : zero2;
            row_product_bc<p>(M1, xm, wl, wr, oPS);
            row_product<p, T>(K1, xs, wl.x, wr.x, oR);
#pragma unroll
            for (int i = 0; i < p; ++i)
              qp_s[y_out + i * TW] = make_float2(cx * oPS[i].y + cy * oR[i], oPS[i].x);
          }
        __syncthreads();

        // (d) z-stage: y = Mz Q + c2 Kz P on the cell layer c, fused vector operation
        if (!owner)
          continue;
#pragma unroll
        for (int b = 0; b < p; ++b)
          QPz[b + 1] = qp_s[b * PQP + pq_off];
        if (c >= c_begin)
          {
            T y[n];
#pragma unroll
            for (int r = 0; r < n; ++r)
              {
                float2 s = __fmul2_rn(MK[canon<p>(r, 0)], QPz[0]); // (Mz Q, Kz P)
#pragma unroll
                for (int i = 1; i < n; ++i)
                  s = __ffma2_rn(MK[canon<p>(r, i)], QPz[i], s);
                y[r] = s.x + cz * s.y;
              }
            y[0] += carry;
            carry = y[p];
            if (c >= c0)
              {
                if (bnd)
                  store_planes<p, T, MODE, p, true>(A, ops, y, d_layer, own_stride, con);
                else
                  store_planes<p, T, MODE, p, false>(A, ops, y, d_layer, own_stride, 0u);
                d_layer += p * own_stride;
              }
          }
        QPz[0] = QPz[p];
      }
    if (c1 == ncz && owner && !L.ztop) // the top plane of the level (a ghost on all but the last z-slab)
      {
        T y[n];
        y[0] = carry;
        const unsigned int con = (xy_constrained || (L.dirichlet & 32u)) ? 1u : 0u;
        Operands<T, 1>     ops;
        load_operands<p, T, MODE, 1, true>(A, ops, d_layer, own_stride, con);
        store_planes<p, T, MODE, 1, true>(A, ops, y, d_layer, own_stride, con);
      }
  }

} // namespace mgb
