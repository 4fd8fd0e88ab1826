// Level transfer in the style of MGTransferMatrixFree (deal.II; call sites common/multigrid_solver.h:
// 213-221, 415, 668, 674), streaming over the tile columns of the FINE level (level.h).
//
// Global refinement halves every cell, so in each direction the prolongation is the 1-D embedding
// P1 ((2p+1) x (p+1): coarse GLL basis at the two children's GLL nodes) applied parent cell by parent
// cell:  fine(f) = sum_i P1[r][i] coarse(p c + i),  c = min(f / 2p, nc - 1),  r = f - 2p c.
// Fine points on a parent face coincide with a coarse node (rows 0 and 2p of P1 are unit vectors), so
// assigning each fine point to ONE parent defines the global matrix; restriction is its transpose,
// i.e. every fine entry is counted once (deal.II's weighting by the number of adjacent parents).
//
//   k_prolongate_tile : block = fine tile column; per parent z-layer the p new coarse planes are
//                       interpolated to the tile in x and y through shared memory, the z-interpolation
//                       runs on registers; every fine value is written once (set or add).
//   k_restrict_tile   : block = fine tile column; each thread folds its own fine column into the p+1
//                       coarse planes of the parent z-layer (registers), the x/y restriction of those
//                       planes runs through shared memory and the few coarse values per block are
//                       added with red.global.add (the coarse vector is 8x smaller than the fine one).
// Dirichlet DoFs of the coarse level are read as zero / not written when `constrained` is set
// (MGTransferMatrixFree::initialize_constraints, multigrid_solver.h:220).
#pragma once
#include "cartesian_kernel.cuh"

namespace mgb
{
  template <int p>
  struct XferCfg
  {
    static constexpr int n  = p + 1;
    static constexpr int TW = TileCfg<p>::TW;
    static constexpr int NT = TileCfg<p>::NT;
    static constexpr int NC = p * ((TW - 1) / (2 * p) + 2) + 1; // coarse nodes per direction under a fine tile
    static constexpr int NLC = (NC * NC + NT - 1) / NT;         // coarse points per thread and plane
  };

  // parent cell and row of P1 of fine point f
  template <int p>
  __device__ __forceinline__ void
  parent_of(const int f, const int nc_coarse, int &c, int &r)
  {
    c = f / (2 * p);
    if (c > nc_coarse - 1)
      c = nc_coarse - 1;
    r = f - 2 * p * c;
  }

  template <typename T>
  __device__ __forceinline__ const T *p1_table();
  template <>
  __device__ __forceinline__ const double *p1_table<double>()
  {
    return ctab.P1;
  }
  template <>
  __device__ __forceinline__ const float *p1_table<float>()
  {
    return ctab.P1f;
  }

  // ------------------------------------------------------------------ prolongation
  template <int p, typename T, bool ADD>
  __global__ void
  __launch_bounds__(XferCfg<p>::NT) k_prolongate_tile(const TransferArgs<T> A, const int layers_per_segment)
  {
    using C           = XferCfg<p>;
    constexpr int n   = C::n;
    constexpr int TW  = C::TW;
    constexpr int NT  = C::NT;
    constexpr int NC  = C::NC;
    constexpr int NLC = C::NLC;
    __shared__ T P1s[(2 * p + 1) * n];
    __shared__ T c_s[p * NC * NC];  // [plane][cj][ci]
    __shared__ T t_s[p * NC * TW];  // [plane][cj][fi]

    const LevelDesc &Lf = A.fine, &Lc = A.coarse;
    const int        tid    = threadIdx.x;
    const int        ntiles = Lf.nt[0] * Lf.nt[1];
    const int        tile = blockIdx.x % ntiles, seg = blockIdx.x / ntiles;
    const int        tx = tile % Lf.nt[0], ty = tile / Lf.nt[0];
    const int        gx0 = tx * TW, gy0 = ty * TW;
    const int        wx = tile_width(Lf, 0, tx), wy = tile_width(Lf, 1, ty);
    const int        cz0 = A.cz_begin + seg * layers_per_segment;
    const int        cz1 = imin(cz0 + layers_per_segment, A.cz_end);
    const int        fsh = A.f_shift;
    const T *__restrict__ src = A.src;
    T *__restrict__ dst       = A.dst;

    for (int i = tid; i < (2 * p + 1) * n; i += NT)
      P1s[i] = p1_table<T>()[i];

    // coarse nodes under this tile
    int cax, cay, r_unused;
    parent_of<p>(gx0, Lc.ncell[0], cax, r_unused);
    parent_of<p>(gy0, Lc.ncell[1], cay, r_unused);
    const int bx = p * cax, by = p * cay;

    // per-thread coarse source points (fixed over the planes)
    uint32_t c_off[NLC];
    int      c_stride[NLC];
    bool     c_ok[NLC];
#pragma unroll
    for (int l = 0; l < NLC; ++l)
      {
        const int q  = tid + l * NT;
        const int gi = bx + q % NC, gj = by + q / NC;
        bool      ok = q < NC * NC && gi < Lc.nd[0] && gj < Lc.nd[1];
        if (ok && A.constrained && is_constrained_xy(Lc, gi, gj))
          ok = false;
        c_ok[l]     = ok;
        c_off[l]    = ok ? dof_index(Lc, gi, gj, 0) : 0u;
        c_stride[l] = ok ? (int)plane_pitch(Lc) : 0;
      }
    const int k_lo = (A.constrained && (Lc.dirichlet & 16u)) ? 1 : 0;
    const int k_hi = (A.constrained && (Lc.dirichlet & 32u)) ? Lc.nd[2] - 2 : Lc.nd[2] - 1;

    // owned fine point
    const int  I = tid % TW, J = tid / TW;
    const bool owner = tid < TW * TW && I < wx && J < wy;
    int        cyo, ryo;
    parent_of<p>(gy0 + J, Lc.ncell[1], cyo, ryo);
    const int      cj0        = p * (cyo - cay);
    const uint32_t own_off    = tile_base(Lf, tx, ty) + (uint32_t)(J * wx + I);
    const uint32_t own_stride = plane_pitch(Lf);

    T W[n]; // x/y-interpolated coarse planes of the current parent layer at this thread's point

    // coarse planes k_first .. k_first + np - 1 under the tile -> registers (requested one parent layer ahead)
    T    creg[p][NLC];
    auto load_coarse = [&](const int k_first, const int np) {
#pragma unroll
      for (int b = 0; b < p; ++b)
        {
          const int  gk  = k_first + b;
          const bool okz = b < np && gk >= k_lo && gk <= k_hi;
#pragma unroll
          for (int l = 0; l < NLC; ++l)
            creg[b][l] = (c_ok[l] && okz) ? __ldg(src + c_off[l] + (uint32_t)gk * (uint32_t)c_stride[l]) : T(0);
        }
    };
    // full tiles of a whole number of parent cells (see k_restrict_tile): a line is interpolated parent cell by
    // parent cell from registers
    constexpr bool ALIGNED = TW % (2 * p) == 0;
    constexpr int  NPC     = ALIGNED ? TW / (2 * p) : 1;
    constexpr int  NCF     = p * NPC + 1;
    const bool     fast    = ALIGNED && wx == TW && wy == TW;

    // interpolates the np coarse planes held in creg in x and y; thread (I,J) gets W[w0 + b]
    auto planes_2d = [&](const int np, const int w0) {
#pragma unroll
      for (int b = 0; b < p; ++b)
        if (b < np)
#pragma unroll
          for (int l = 0; l < NLC; ++l)
            if (tid + l * NT < NC * NC)
              c_s[b * NC * NC + tid + l * NT] = creg[b][l];
      __syncthreads();
      if (fast)
        for (int t = tid; t < np * NCF; t += NT)
          {
            const int cj = t % NCF, b = t / NCF;
            T         cv[NCF];
#pragma unroll
            for (int ci = 0; ci < NCF; ++ci)
              cv[ci] = c_s[(b * NC + cj) * NC + ci];
#pragma unroll
            for (int c = 0; c < NPC; ++c)
#pragma unroll
              for (int r = 0; r < 2 * p; ++r)
                {
                  T v = 0;
#pragma unroll
                  for (int i = 0; i < n; ++i)
                    v += p1_table<T>()[r * n + i] * cv[p * c + i];
                  t_s[(b * NC + cj) * TW + 2 * p * c + r] = v;
                }
          }
      else
      for (int t = tid; t < TW * NC * np; t += NT)
        {
          const int fi = t % TW, rest = t / TW;
          const int cj = rest % NC, b = rest / NC;
          int       c, r;
          parent_of<p>(gx0 + fi, Lc.ncell[0], c, r);
          const T *cp = c_s + (b * NC + cj) * NC + p * (c - cax);
          T        s  = 0;
          if (fi < wx)
#pragma unroll
            for (int i = 0; i < n; ++i)
              s += P1s[r * n + i] * cp[i];
          t_s[(b * NC + cj) * TW + fi] = s;
        }
      __syncthreads();
      if (owner)
        for (int b = 0; b < np; ++b)
          {
            T s = 0;
#pragma unroll
            for (int j = 0; j < n; ++j)
              s += P1s[ryo * n + j] * t_s[(b * NC + cj0 + j) * TW + I];
#pragma unroll
            for (int k = 0; k < n; ++k)
              if (k == w0 + b)
                W[k] = s;
          }
    };

    load_coarse(p * cz0, 1);
    planes_2d(1, 0);
    load_coarse(p * cz0 + 1, p);
    for (int cz = cz0; cz < cz1; ++cz)
      {
        planes_2d(p, 1);
        if (cz + 1 < cz1)
          load_coarse(p * (cz + 1) + 1, p);
        if (owner)
          {
            const int nr = (cz == A.cz_end - 1 && A.close_top) ? 2 * p + 1 : 2 * p; // the top plane belongs to the last layer
            T         old[2 * p + 1];
            if (ADD)
              {
#pragma unroll
                for (int r = 0; r < 2 * p + 1; ++r)
                  if (r < nr)
                    old[r] = dst[own_off + (uint32_t)(2 * p * cz + r + fsh) * own_stride];
              }
#pragma unroll
            for (int r = 0; r < 2 * p + 1; ++r)
              if (r < nr)
                {
                  T s = 0;
#pragma unroll
                  for (int k = 0; k < n; ++k)
                    s += p1_table<T>()[r * n + k] * W[k];
                  const uint32_t d = own_off + (uint32_t)(2 * p * cz + r + fsh) * own_stride;
                  dst[d]           = ADD ? old[r] + s : s;
                }
          }
        W[0] = W[p];
      }
  }

  // ------------------------------------------------------------------ restriction (adds into the coarse vector)
  template <int p, typename T>
  __global__ void
  __launch_bounds__(XferCfg<p>::NT) k_restrict_tile(const TransferArgs<T> A, const int layers_per_segment)
  {
    using C          = XferCfg<p>;
    constexpr int n  = C::n;
    constexpr int TW = C::TW;
    constexpr int NT = C::NT;
    constexpr int NC = C::NC;
    __shared__ T Wx[NC * TW], Wy[NC * TW]; // Wx[ci][fi] = weight of fine column fi in coarse node ci
    __shared__ __align__(16) T z_s[p * TW * TW]; // [plane][fj][fi]
    __shared__ T x_s[p * TW * NC];         // [plane][fj][ci]

    const LevelDesc &Lf = A.fine, &Lc = A.coarse;
    const int        tid    = threadIdx.x;
    const int        ntiles = Lf.nt[0] * Lf.nt[1];
    const int        tile = blockIdx.x % ntiles, seg = blockIdx.x / ntiles;
    const int        tx = tile % Lf.nt[0], ty = tile / Lf.nt[0];
    const int        gx0 = tx * TW, gy0 = ty * TW;
    const int        wx = tile_width(Lf, 0, tx), wy = tile_width(Lf, 1, ty);
    const int        cz0 = A.cz_begin + seg * layers_per_segment;
    const int        cz1 = imin(cz0 + layers_per_segment, A.cz_end);
    const int        fsh = A.f_shift;
    const T *__restrict__ src = A.src;
    T *__restrict__ dst       = A.dst;

    int cax, cay, r_unused;
    parent_of<p>(gx0, Lc.ncell[0], cax, r_unused);
    parent_of<p>(gy0, Lc.ncell[1], cay, r_unused);
    const int bx = p * cax, by = p * cay;

    // weight tables
    for (int i = tid; i < NC * TW; i += NT)
      Wx[i] = Wy[i] = T(0);
    __syncthreads();
    for (int t = tid; t < 2 * TW * n; t += NT)
      {
        const int  dir = t / (TW * n), rem = t - dir * TW * n;
        const int  f = rem / n, i = rem - f * n;
        const int  g0 = dir == 0 ? gx0 : gy0, w = dir == 0 ? wx : wy, ca = dir == 0 ? cax : cay;
        if (f < w)
          {
            int c, r;
            parent_of<p>(g0 + f, Lc.ncell[dir], c, r);
            (dir == 0 ? Wx : Wy)[(p * (c - ca) + i) * TW + f] = p1_table<T>()[r * n + i];
          }
      }
    // number of coarse nodes actually touched
    int ncx, ncy;
    {
      int c, r;
      parent_of<p>(gx0 + wx - 1, Lc.ncell[0], c, r);
      ncx = p * (c - cax) + p + 1;
      parent_of<p>(gy0 + wy - 1, Lc.ncell[1], c, r);
      ncy = p * (c - cay) + p + 1;
    }

    const int  I = tid % TW, J = tid / TW;
    const bool owner = tid < TW * TW && I < wx && J < wy;
    const uint32_t own_off    = tile_base(Lf, tx, ty) + (uint32_t)(J * wx + I);
    const uint32_t own_stride = plane_pitch(Lf);

    // Full tiles whose width is a whole number of parent cells (all but the ragged last tile row / column, and only
    // degrees with TW % 2p == 0): every line of the tile restricts parent cell by parent cell with the 2p x (p+1)
    // block of P1 as compile-time constants -- 2p (p+1) FMAs per parent instead of a dense TW-term sum per coarse
    // node, the line in registers instead of two shared-memory reads per term.
    constexpr bool ALIGNED = TW % (2 * p) == 0;
    constexpr int  NPC     = ALIGNED ? TW / (2 * p) : 1; // parent cells across a tile
    constexpr int  NCF     = p * NPC + 1;                // coarse nodes across a tile
    const bool     fast    = ALIGNED && wx == TW && wy == TW;
    auto           line_restrict = [&](const T(&f)[TW], T(&out)[NCF]) {
#pragma unroll
      for (int i = 0; i < NCF; ++i)
        out[i] = 0;
#pragma unroll
      for (int c = 0; c < NPC; ++c)
#pragma unroll
        for (int r = 0; r < 2 * p; ++r)
#pragma unroll
          for (int i = 0; i < n; ++i)
            out[p * c + i] += p1_table<T>()[r * n + i] * f[2 * p * c + r];
    };

    // x/y restriction of `np` planes held in z_s and addition into the coarse planes k_first, ...
    auto planes_2d = [&](const int k_first, const int np) {
      __syncthreads();
      if (fast)
        {
          for (int t = tid; t < np * TW; t += NT)
            {
              const int fj = t % TW, b = t / TW;
              T         f[TW], out[NCF];
#pragma unroll
              for (int fi = 0; fi < TW; ++fi)
                f[fi] = z_s[(b * TW + fj) * TW + fi];
              line_restrict(f, out);
#pragma unroll
              for (int ci = 0; ci < NCF; ++ci)
                x_s[(b * TW + fj) * NC + ci] = out[ci];
            }
          __syncthreads();
          for (int t = tid; t < np * NCF; t += NT)
            {
              const int ci = t % NCF, b = t / NCF;
              T         f[TW], out[NCF];
#pragma unroll
              for (int fj = 0; fj < TW; ++fj)
                f[fj] = x_s[(b * TW + fj) * NC + ci];
              line_restrict(f, out);
              const int gi = bx + ci, gk = k_first + b;
#pragma unroll
              for (int cj = 0; cj < NCF; ++cj)
                if (!(A.constrained && is_constrained(Lc, gi, by + cj, gk)))
                  atomicAdd(dst + dof_index(Lc, gi, by + cj, gk), out[cj]);
            }
          return;
        }
      for (int t = tid; t < np * TW * ncx; t += NT)
        {
          const int ci = t % ncx, rest = t / ncx;
          const int fj = rest % TW, b = rest / TW;
          const T  *zp = z_s + (b * TW + fj) * TW;
          const T  *wp = Wx + ci * TW;
          T         s  = 0;
#pragma unroll
          for (int fi = 0; fi < TW; ++fi)
            s += wp[fi] * zp[fi];
          x_s[(b * TW + fj) * NC + ci] = s;
        }
      __syncthreads();
      for (int t = tid; t < np * ncx * ncy; t += NT)
        {
          const int ci = t % ncx, rest = t / ncx;
          const int cj = rest % ncy, b = rest / ncy;
          const T  *wp = Wy + cj * TW;
          T         s  = 0;
#pragma unroll
          for (int fj = 0; fj < TW; ++fj)
            s += wp[fj] * x_s[(b * TW + fj) * NC + ci];
          const int gi = bx + ci, gj = by + cj, gk = k_first + b;
          if (!(A.constrained && is_constrained(Lc, gi, gj, gk)))
            atomicAdd(dst + dof_index(Lc, gi, gj, gk), s);
        }
    };

    T carry = 0;
    for (int cz = cz0; cz < cz1; ++cz)
      {
        const int nr = (cz == A.cz_end - 1 && A.close_top) ? 2 * p + 1 : 2 * p; // fine planes assigned to this parent layer
        T         v[2 * p + 1], Z[n];
#pragma unroll
        for (int r = 0; r < 2 * p + 1; ++r)
          v[r] = (owner && r < nr) ? __ldg(src + own_off + (uint32_t)(2 * p * cz + r + fsh) * own_stride) : T(0);
#pragma unroll
        for (int k = 0; k < n; ++k)
          {
            T s = k == 0 ? carry : T(0);
#pragma unroll
            for (int r = 0; r < 2 * p + 1; ++r)
              s += p1_table<T>()[r * n + k] * v[r];
            Z[k] = s;
          }
        carry = Z[p];
        // (z_s is free: every thread has passed the barrier after the previous x-stage)
        if (tid < TW * TW)
#pragma unroll
          for (int k = 0; k < p; ++k)
            z_s[k * TW * TW + tid] = Z[k];
        planes_2d(p * cz, p);
      }
    // the plane on top of the segment
    if (tid < TW * TW)
      z_s[tid] = carry;
    planes_2d(p * cz1, 1);
  }

  template <int p, typename T>
  void
  launch_transfer_tile(const TransferArgs<T> &A, cudaStream_t stream)
  {
    using C          = XferCfg<p>;
    const int ntiles = A.fine.nt[0] * A.fine.nt[1];
    const int ncz    = A.cz_end - A.cz_begin;
    if (ncz <= 0)
      return;
    int       nseg   = (148 * 8 + ntiles - 1) / ntiles;
    if (nseg > ncz / 2)
      nseg = ncz / 2;
    if (nseg < 1)
      nseg = 1;
    // small levels: one parent layer per block rather than an idle device (as launch_apply_tile)
    if (ntiles * nseg < 148 * 4 && use_short_segments())
      nseg = std::max(nseg, std::min(ncz, 148 * 4 / ntiles));
    const int lps = (ncz + nseg - 1) / nseg;
    nseg          = (ncz + lps - 1) / lps;
    if (A.mode == 0)
      k_prolongate_tile<p, T, false><<<ntiles * nseg, C::NT, 0, stream>>>(A, lps);
    else if (A.mode == 1)
      k_prolongate_tile<p, T, true><<<ntiles * nseg, C::NT, 0, stream>>>(A, lps);
    else
      k_restrict_tile<p, T><<<ntiles * nseg, C::NT, 0, stream>>>(A, lps);
  }
} // namespace mgb
