// Float operator apply on Cartesian levels, second generation of the streaming tile kernel (cartesian_kernel.cuh /
// cartesian_kernel_f32x2.cuh): same mathematics, same summation order per 1-D product, same fused vector operations
// (laplace_operator.h:575-741), organised to move less data through shared memory -- the limiter of the first
// generation (profiles/r1_tile_f32x2_ncu_full_summary.txt: L1/shared data pipe 71 %, barrier stalls):
//
//   * a block owns a 2 x 2 group of layout tiles (32 x 32 points at p = 4) for a range of cell layers;
//   * x-stage: a task is (row, plane, half row) and produces 16 points = 4 cells from 21 staged values in registers
//     (the first generation read 9 values per 4 points);
//   * y- and z-stage are ONE stage: thread (I, cell row) computes its p rows of (Q, P) for the p planes of the layer
//     straight into registers and runs the z-products on them, so (Q, P) never travels through shared memory and one
//     of the three block barriers per batch is gone; a batch is 4096 instead of 1024 DoFs, i.e. six times fewer
//     barriers per DoF;
//   * shared-memory traffic per DoF: 39 instead of 63 bytes (float), source halo 1.34 instead of 1.72 values per DoF.
//
// Compiled for p = 2 and p = 4 (layout tile 16 points wide, a whole number of cells); other degrees and the
// double-precision operator keep the first generation.
#pragma once
#include "cartesian_kernel_f32x2.cuh"

namespace mgb
{
  template <int p>
  struct V5Cfg
  {
    using C                  = TileCfg<p>;
    static constexpr int n   = p + 1;
    static constexpr int TW  = C::TW;            // layout tile width (16)
    static constexpr int GW  = 2 * TW;           // points per block and direction (32)
    static constexpr int NI  = GW + p + 1;       // staged points per direction: [-p, GW]
    static constexpr int NCY = GW / p;           // cell rows per block = threads per point column
    static constexpr int NT  = GW * NCY;         // threads: (I, cell row)
    static constexpr int HALF_CELLS = GW / p / 2; // cells per x-stage task (half a row)
    static constexpr int HW  = GW / 2;           // points per x-stage task
    static constexpr int NXT = NI * p * 2;       // x-stage tasks per batch
    static constexpr int NXR = (NXT + NT - 1) / NT;
    static constexpr int NL  = (NI * NI + NT - 1) / NT; // staged points per thread and plane
    static constexpr int U_PLANE = NI * NI;
    static constexpr int U_SIZE  = p * U_PLANE;          // floats per staging buffer
    static constexpr int MKP     = GW * NI + 1;          // plane pitch of (mx, kx) in float2 (odd: conflict-free planes)
    static constexpr int SMEM_FLOATS = 2 * U_SIZE + 2 * p * MKP;
    static constexpr bool available  = (TW == 16) && (GW % p == 0) && (p == 2 || p == 4);
  };

  // (M, K) products of one line segment of NC cells: x holds the NC p + 1 points of the cells preceded by the p points of
  // the cell before; out[i], i = 0 .. NC p - 1, is the complete result at point i (first point of cell i / p shared with
  // the cell before).  Bit c of `own` / `prev`: cell c of the segment / the cell before it exists.
  template <int p, int NC>
  __device__ __forceinline__ void
  line_product_mk(const float2 *__restrict__ MK, const float (&x)[NC * p + p + 1], const unsigned own, const unsigned prev,
                  float2 (&out)[NC * p])
  {
#pragma unroll
    for (int c = 0; c < NC; ++c)
      {
        float w[2 * p + 1];
#pragma unroll
        for (int i = 0; i < 2 * p + 1; ++i)
          w[i] = x[c * p + i];
        const float xl = (prev >> c) & 1u ? w[p] : 0.f;
        const float xr = (own >> c) & 1u ? w[p] : 0.f;
        float2      o[p];
        row_product_mk<p>(MK, w, xl, xr, o);
#pragma unroll
        for (int i = 0; i < p; ++i)
          out[c * p + i] = o[i];
      }
  }

  template <int p, int MODE>
  __global__ void
  __launch_bounds__(V5Cfg<p>::NT, 2)
    k_apply_tile_v5(const ApplyArgs<float> A, const int layers_per_segment, const int layer_begin, const int layer_end)
  {
    using V           = V5Cfg<p>;
    using T           = float;
    constexpr int n   = V::n;
    constexpr int TW  = V::TW;
    constexpr int GW  = V::GW;
    constexpr int NI  = V::NI;
    constexpr int NT  = V::NT;
    constexpr int NL  = V::NL;
    constexpr int NXR = V::NXR;
    constexpr int HW  = V::HW;
    constexpr int HC  = V::HALF_CELLS;
    constexpr int MKP = V::MKP;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    T      *u_s  = reinterpret_cast<T *>(smem_raw);                 // [buffer][plane][row J' + p][col I' + p]
    float2 *mk_s = reinterpret_cast<float2 *>(u_s + 2 * V::U_SIZE); // (mx, kx) [plane (pitch MKP)][I][row J' + p]

    const LevelDesc &L   = A.L;
    const int        tid = threadIdx.x;
    const int        ng0 = (L.nt[0] + 1) / 2, ng1 = (L.nt[1] + 1) / 2, ngroups = ng0 * ng1;
    const int        group = blockIdx.x % ngroups, seg = blockIdx.x / ngroups;
    const int        bx = group % ng0, by = group / ng0;
    const int        gx0 = bx * GW, gy0 = by * GW;
    const int        ncz = L.ncell[2], nd2 = L.nd[2];
    const bool       second = A.layer_begin2 >= 0 && seg == 1; // two single layers in one launch
    const int        c0 = second ? A.layer_begin2 : layer_begin + seg * layers_per_segment;
    const int        c1 = second ? c0 + 1 : imin(c0 + layers_per_segment, layer_end);
    const bool       plain = A.read_plain != 0;
    const uint32_t   pitch = plane_pitch(L);

    // ---- staged source points of this thread (fixed over the planes): offset on plane 0, validity
    uint32_t src_off[NL];
    unsigned src_ok = 0;
#pragma unroll
    for (int l = 0; l < NL; ++l)
      {
        const int q  = tid + l * NT;
        const int gi = gx0 + q % NI - p, gj = gy0 + q / NI - p;
        bool      ok = q < NI * NI && gi >= 0 && gi < L.nd[0] && gj >= 0 && gj < L.nd[1];
        if (ok && !plain && is_constrained_xy(L, gi, gj))
          ok = false; // Dirichlet in x or y; the z-faces are handled per plane
        src_off[l] = ok ? dof_index(L, gi, gj, 0) : 0u;
        if (ok)
          src_ok |= 1u << l;
      }
    const int k_lo = (!plain && (L.dirichlet & 16u)) ? 1 : 0;
    const int k_hi = (!plain && (L.dirichlet & 32u)) ? nd2 - 2 : nd2 - 1;

    // ---- x-stage tasks: (row, plane, half)
    int      x_in[NXR], x_out[NXR];
    unsigned x_own[NXR], x_prev[NXR];
#pragma unroll
    for (int r = 0; r < NXR; ++r)
      {
        const int t    = tid + r * NT;
        const int row  = t % NI, rest = t / NI;
        const int b    = rest % p, half = rest / p;
        x_in[r]        = t < V::NXT ? (b * NI + row) * NI + half * HW : -1;
        x_out[r]       = b * MKP + half * HW * NI + row;
        x_own[r] = x_prev[r] = 0;
#pragma unroll
        for (int c = 0; c < HC; ++c)
          {
            const int gc = gx0 / p + half * HC + c;
            if (gc < L.ncell[0])
              x_own[r] |= 1u << c;
            if (gc - 1 >= 0 && gc - 1 < L.ncell[0])
              x_prev[r] |= 1u << c;
          }
      }

    // ---- y/z-stage: thread (I, cell row cy) owns the points (I, cy p + r), r = 0 .. p-1
    const int  I = tid % GW, cy = tid / GW;
    const int  gi = gx0 + I;
    const int  gcy = gy0 / p + cy;
    const bool y_own = gcy < L.ncell[1], y_prev = gcy - 1 >= 0 && gcy - 1 < L.ncell[1];
    const int  y_in = I * NI + cy * p;
    uint32_t   own_off[p];
    unsigned   own_ok = 0, xy_con = 0;
#pragma unroll
    for (int r = 0; r < p; ++r)
      {
        const int  gj = gy0 + cy * p + r;
        const bool ok = gi < L.nd[0] && gj < L.nd[1];
        own_off[r]    = ok ? dof_index(L, gi, gj, 0) : 0u;
        if (ok)
          own_ok |= 1u << r;
        if (ok && is_constrained_xy(L, gi, gj))
          xy_con |= 1u << r;
      }
    // block-uniform: does this group touch a Dirichlet face in x or y?
    const bool group_bnd = ((L.dirichlet & 1u) && bx == 0) || ((L.dirichlet & 2u) && gx0 + GW >= L.nd[0]) ||
                           ((L.dirichlet & 4u) && by == 0) || ((L.dirichlet & 8u) && gy0 + GW >= L.nd[1]);

    const T      *M1 = Tab<T>::M1();
    const T      *K1 = Tab<T>::K1();
    const float2 *MK = ctab.MKf;
    const T       cx = A.coef[0], cyc = A.coef[1], cz = A.coef[2];

    float2 QP0[p]; // (Q, P) on the lowest plane of the current layer, per owned row
    T      carry[p];
#pragma unroll
    for (int r = 0; r < p; ++r)
      {
        QP0[r]   = make_float2(0.f, 0.f);
        carry[r] = 0;
      }

    // source planes of batch c (planes c p + 1 .. c p + p) -> buffer `buf`, asynchronously
    auto issue_loads = [&](const int c, const int buf) {
      const int k0 = c * p + 1;
#pragma unroll
      for (int b = 0; b < p; ++b)
        {
          const int  gk  = k0 + b;
          const bool okz = gk >= k_lo && gk <= k_hi;
          const T   *g   = A.src + (okz ? (size_t)gk * pitch : 0);
#pragma unroll
          for (int l = 0; l < NL; ++l)
            if (l + 1 < NL || tid + l * NT < NI * NI)
              cp_async_value(u_s + (buf * p + b) * V::U_PLANE + tid + l * NT, g + src_off[l], okz && ((src_ok >> l) & 1u));
        }
      asm volatile("cp.async.commit_group;\n" ::);
    };

    const int c_begin = c0 > 0 ? c0 - 1 : 0;
    uint32_t  d_plane = (uint32_t)(c0 * p) * pitch; // first plane of the layer being stored
    issue_loads(c_begin - 1, 0);
    for (int c = c_begin - 1; c < c1; ++c)
      {
        const int cur = (c - c_begin + 1) & 1;
        asm volatile("cp.async.wait_group 0;\n" ::: "memory");
        __syncthreads(); // planes of this batch visible; everybody is done with mk_s of the previous batch
        if (c + 1 < c1)
          issue_loads(c + 1, cur ^ 1);
        const T *u_c = u_s + cur * V::U_SIZE;

        // (a) x-stage: (mx, kx) = (Mx, Kx) u on half rows
#pragma unroll
        for (int r = 0; r < NXR; ++r)
          if (x_in[r] >= 0)
            {
              T x[HW + p + 1];
#pragma unroll
              for (int i = 0; i < HW + p + 1; ++i)
                x[i] = u_c[x_in[r] + i];
              float2 o[HW];
              line_product_mk<p, HC>(MK, x, x_own[r], x_prev[r], o);
#pragma unroll
              for (int i = 0; i < HW; ++i)
                mk_s[x_out[r] + i * NI] = o[i];
            }
        __syncthreads();

        // (b) y-stage into registers: (Q, P) of the p owned rows on the p planes of this batch
        float2 QP[p][p]; // [row][plane]
#pragma unroll
        for (int b = 0; b < p; ++b)
          {
            float2 xm[2 * p + 1], oPS[p];
            T      xs[2 * p + 1], oR[p];
#pragma unroll
            for (int i = 0; i < 2 * p + 1; ++i)
              {
                xm[i] = mk_s[b * MKP + y_in + i];
                xs[i] = xm[i].x;
              }
            const float2 zero2 = make_float2(0.f, 0.f);
            const float2 wl    = y_prev ? xm[p] : zero2;
            const float2 wr    = y_own ? xm[p] : zero2;
            row_product_bc<p>(M1, xm, wl, wr, oPS);
            row_product<p, T>(K1, xs, wl.x, wr.x, oR);
#pragma unroll
            for (int r = 0; r < p; ++r)
              QP[r][b] = make_float2(cx * oPS[r].y + cyc * oR[r], oPS[r].x);
          }

        // (c) z-stage on the layer c and the fused vector operation, row by row
        if (c >= c_begin)
          {
            const bool bnd = group_bnd || c == 0; // block-uniform: only then constraints are looked at
#pragma unroll
            for (int r = 0; r < p; ++r)
              {
                T y[n];
#pragma unroll
                for (int o = 0; o < n; ++o)
                  {
                    float2 s = __fmul2_rn(MK[canon<p>(o, 0)], QP0[r]); // (Mz Q, Kz P)
#pragma unroll
                    for (int i = 1; i < n; ++i)
                      s = __ffma2_rn(MK[canon<p>(o, i)], QP[r][i - 1], s);
                    y[o] = s.x + cz * s.y;
                  }
                y[0] += carry[r];
                carry[r] = y[p];
                if (c >= c0 && ((own_ok >> r) & 1u))
                  {
                    const uint32_t d0 = own_off[r] + d_plane;
                    Operands<T, p> ops;
                    if (bnd)
                      {
                        unsigned con = (xy_con >> r) & 1u ? 0xffu : 0u;
                        if (c == 0 && (L.dirichlet & 16u))
                          con |= 1u;
                        load_operands<p, T, MODE, p, true>(A, ops, d0, pitch, con);
                        store_planes<p, T, MODE, p, true>(A, ops, y, d0, pitch, con);
                      }
                    else
                      {
                        load_operands<p, T, MODE, p, false>(A, ops, d0, pitch, 0u);
                        store_planes<p, T, MODE, p, false>(A, ops, y, d0, pitch, 0u);
                      }
                  }
              }
            if (c >= c0)
              d_plane += p * pitch;
          }
#pragma unroll
        for (int r = 0; r < p; ++r)
          QP0[r] = QP[r][p - 1];
      }
    if (c1 == ncz && !L.ztop) // the top plane of the level (a ghost on all but the last z-slab)
      {
#pragma unroll
        for (int r = 0; r < p; ++r)
          if ((own_ok >> r) & 1u)
            {
              T y[n];
              y[0] = carry[r];
              const unsigned  con = (((xy_con >> r) & 1u) || (L.dirichlet & 32u)) ? 1u : 0u;
              const uint32_t  d0  = own_off[r] + d_plane;
              Operands<T, 1>  ops;
              load_operands<p, T, MODE, 1, true>(A, ops, d0, pitch, con);
              store_planes<p, T, MODE, 1, true>(A, ops, y, d0, pitch, con);
            }
      }
  }

  // MGB200_V5=0 keeps the first-generation float kernel (A/B measurements)
  inline bool
  use_v5()
  {
    static const bool on = [] {
      const char *e = getenv("MGB200_V5");
      return !(e && e[0] == '0');
    }();
    return on;
  }

  // returns false if this instance is not served by the second-generation kernel
  template <int p, int MODE>
  bool
  launch_apply_tile_v5(const ApplyArgs<float> &A, cudaStream_t stream)
  {
    using V = V5Cfg<p>;
    if constexpr (!V::available || MODE == APPLY_CHEBYSHEV_ZERO || MODE == APPLY_SET_DOT)
      return false;
    else
      {
        if (!use_v5() || !use_f32x2())
          return false;
        const size_t smem  = (size_t)V::SMEM_FLOATS * sizeof(float);
        static bool  first = true;
        if (first)
          {
            cudaFuncSetAttribute(k_apply_tile_v5<p, MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
            first = false;
          }
        const int ngroups = ((A.L.nt[0] + 1) / 2) * ((A.L.nt[1] + 1) / 2);
        const int l0      = A.layer_begin >= 0 ? A.layer_begin : A.L.zg;
        const int l1      = A.layer_end >= 0 ? A.layer_end : A.L.ncell[2];
        const int ncz     = l1 - l0;
        if (ncz <= 0)
          return true;
        // blocks for a few waves at 2 blocks per SM; a segment re-computes one layer for the carry
        int nseg = (148 * 2 * 3 + ngroups - 1) / ngroups;
        if (nseg > ncz / 4)
          nseg = ncz / 4;
        if (nseg < 1)
          nseg = 1;
        if (ngroups * nseg < 148 * 2 && use_short_segments())
          nseg = std::max(nseg, std::min(ncz, 148 * 2 / ngroups));
        int lps = (ncz + nseg - 1) / nseg;
        nseg    = (ncz + lps - 1) / lps;
        if (A.layer_begin2 >= 0)
          {
            lps  = 1;
            nseg = 2;
          }
        k_apply_tile_v5<p, MODE><<<ngroups * nseg, V::NT, smem, stream>>>(A, lps, l0, l1);
        return true;
      }
  }
} // namespace mgb
