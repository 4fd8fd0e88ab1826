// The parts of the adaptive loop of poisson_l/program.cc:497-602 around the solver, on the forest of
// adaptive.cu; included by adaptive.cu after adaptive_kernels.cuh.
//   compute_residual          rhs = (f, phi) - K u for a vector u that carries the Dirichlet data (:545-549; the operator
//                             applied through the table that keeps Dirichlet DoFs, scattered through the one that drops them)
//   KellyErrorEstimator       eta_K^2 = diameter_K / 24 * sum over the faces of K of int_F [du/dn]^2 (:515-531, deal.II's
//                             default strategy cell_diameter_over_24, no Neumann data: boundary faces contribute nothing);
//                             on a face between cells of different size the integral runs over the faces of the small
//                             cells and is added to both sides
//   refine_and_coarsen_fixed_number + execute_coarsening_and_refinement (:532-541): AdaptiveMesh::refine_and_coarsen
//   integrate_difference      L2 norm with QGauss(p+2) and H1 seminorm with QGauss(p+1) of u_h - u (:556-575) for the
//                             functions of the program: zero and the L-shape singularity r^(2/3) sin(2 phi / 3)
#pragma once

namespace mgb
{
  constexpr int AD_EST_MAXN = 11; // 1-D points: up to p + 2 = 10 quadrature points

  struct AdEstimateTables
  {
    double e[2][AD_EST_MAXN];                   // phi_i'(0), phi_i'(1) of the nodal basis
    double S[AD_EST_MAXN * AD_EST_MAXN];        // S[q n + i]  = phi_i(x_q), n_q rows
    double G[AD_EST_MAXN * AD_EST_MAXN];        // G[q n + i]  = phi_i'(x_q)
    double Sh[2][AD_EST_MAXN * AD_EST_MAXN];    // Sh[o][q n + i] = phi_i((o + x_q) / 2): a fine face seen from the coarse cell
    double w[AD_EST_MAXN], x[AD_EST_MAXN];      // quadrature weights and points on [0, 1]
    int    nq;
  };

  struct AdEstimateArgs
  {
    int            n_cells;
    const int32_t *cell_index; // the table that keeps Dirichlet DoFs
    const int32_t *row_start, *row_dof;
    const double  *row_w;
    const uint8_t *cell_level;
    const int32_t *cell_pos; // [cell][3]
    const int8_t  *nbr_type; // [cell][6]
    const int32_t *nbr;      // [cell][6][4]
    const double  *u;
    double         lower[3], root_size[3];
    double        *face_int; // [cell][6]
    double        *out;      // [cell]
    int            function_id, norm;
  };

  __device__ __forceinline__ double
  ad_est_read(const AdEstimateArgs &A, const int32_t entry)
  {
    if (entry >= 0)
      return A.u[entry];
    if (entry == -1)
      return 0.;
    const int row = -2 - entry;
    double    s   = 0;
    for (int t = A.row_start[row]; t < A.row_start[row + 1]; ++t)
      s += A.row_w[t] * A.u[A.row_dof[t]];
    return s;
  }

  // int_F [du/dn]^2 over every face of a cell whose neighbour is of the same size or larger; one block per cell
  template <int p>
  __global__ void
  __launch_bounds__(128) k_ad_face_integrals(const AdEstimateArgs A, const AdEstimateTables Tb)
  {
    constexpr int     n = p + 1, n2 = n * n, n3 = n2 * n;
    __shared__ double U[n3], V[n3], Do[n2], Dn[n2], red[4];
    const int         cell = blockIdx.x, lev = A.cell_level[cell];
    for (int l = threadIdx.x; l < n3; l += blockDim.x)
      U[l] = ad_est_read(A, A.cell_index[(size_t)cell * n3 + l]);
    double h[3];
    for (int d = 0; d < 3; ++d)
      h[d] = A.root_size[d] / (double)(1 << lev);
    for (int f = 0; f < 6; ++f)
      {
        const int type = A.nbr_type[cell * 6 + f];
        double    sum  = 0;
        __syncthreads();
        if (type == 0 || type == 1)
          {
            const int d = f / 2, side = f % 2, t1 = d == 0 ? 1 : 0, t2 = d == 2 ? 1 : 2;
            const int sd = d == 0 ? 1 : (d == 1 ? n : n2), s1 = t1 == 0 ? 1 : n, s2 = t2 == 1 ? n : n2;
            const int nb = A.nbr[(cell * 6 + f) * 4];
            for (int l = threadIdx.x; l < n3; l += blockDim.x)
              V[l] = ad_est_read(A, A.cell_index[(size_t)nb * n3 + l]);
            __syncthreads();
            // d/dx_d at the face nodes, own side and neighbour's side (its opposite face)
            const double hn = type == 1 ? 2. * h[d] : h[d];
            for (int l = threadIdx.x; l < n2; l += blockDim.x)
              {
                const int a = l % n, b = l / n;
                double    so = 0, sn = 0;
                for (int i = 0; i < n; ++i)
                  {
                    so += Tb.e[side][i] * U[a * s1 + b * s2 + i * sd];
                    sn += Tb.e[1 - side][i] * V[a * s1 + b * s2 + i * sd];
                  }
                Do[l] = so / h[d];
                Dn[l] = sn / hn;
              }
            __syncthreads();
            // position of this face inside the neighbour's face (coarser neighbour): parity of the cell's position
            const int     o1 = A.cell_pos[cell * 3 + t1] & 1, o2 = A.cell_pos[cell * 3 + t2] & 1;
            const double *N1 = type == 1 ? Tb.Sh[o1] : Tb.S, *N2 = type == 1 ? Tb.Sh[o2] : Tb.S;
            for (int l = threadIdx.x; l < n2; l += blockDim.x)
              {
                const int q1 = l % n, q2 = l / n;
                double    jo = 0, jn = 0;
                for (int b = 0; b < n; ++b)
                  {
                    double ro = 0, rn = 0;
                    for (int a = 0; a < n; ++a)
                      {
                        ro += Tb.S[q1 * n + a] * Do[b * n + a];
                        rn += N1[q1 * n + a] * Dn[b * n + a];
                      }
                    jo += Tb.S[q2 * n + b] * ro;
                    jn += N2[q2 * n + b] * rn;
                  }
                const double jump = jo - jn;
                sum += Tb.w[q1] * Tb.w[q2] * h[t1] * h[t2] * jump * jump;
              }
          }
        for (int o = 16; o > 0; o >>= 1)
          sum += __shfl_down_sync(0xffffffffu, sum, o);
        if ((threadIdx.x & 31) == 0)
          red[threadIdx.x >> 5] = sum;
        __syncthreads();
        if (threadIdx.x == 0)
          A.face_int[cell * 6 + f] = red[0] + red[1] + red[2] + red[3];
      }
  }

  // eta_K = sqrt(diameter / 24 * sum of the face integrals); faces with smaller neighbours take the neighbours' integrals
  __global__ void
  k_ad_kelly_sum(const AdEstimateArgs A)
  {
    for (int cell = blockIdx.x * blockDim.x + threadIdx.x; cell < A.n_cells; cell += gridDim.x * blockDim.x)
      {
        const int lev = A.cell_level[cell];
        double    d2 = 0, sum = 0;
        for (int d = 0; d < 3; ++d)
          {
            const double h = A.root_size[d] / (double)(1 << lev);
            d2 += h * h;
          }
        for (int f = 0; f < 6; ++f)
          {
            const int type = A.nbr_type[cell * 6 + f];
            if (type == 0 || type == 1)
              sum += A.face_int[cell * 6 + f];
            else if (type == 2)
              for (int s = 0; s < 4; ++s)
                sum += A.face_int[A.nbr[(cell * 6 + f) * 4 + s] * 6 + (f ^ 1)];
          }
        A.out[cell] = sqrt(sqrt(d2) / 24. * sum);
      }
  }

  // the functions of poisson_l/program.cc:85-135: 0 = zero, 1 = Functions::LSingularityFunction in (x, y)
  __host__ __device__ inline void
  ad_exact(const int id, const double x, const double y, double &value, double grad[3])
  {
    value   = 0;
    grad[0] = grad[1] = grad[2] = 0;
    if (id != 1 || (x >= 0 && y >= 0))
      return;
    const double phi = atan2(y, -x) + 3.14159265358979323846, r2 = x * x + y * y, r43 = pow(r2, 2. / 3.);
    value   = pow(r2, 1. / 3.) * sin(2. / 3. * phi);
    grad[0] = 2. / 3. * (sin(2. / 3. * phi) * x + cos(2. / 3. * phi) * y) / r43;
    grad[1] = 2. / 3. * (sin(2. / 3. * phi) * y - cos(2. / 3. * phi) * x) / r43;
  }

  // out[cell] = int_K (u_h - u)^2 (norm 0) or int_K |grad u_h - grad u|^2 (norm 1) with Tb.nq^3 Gauss points
  template <int p>
  __global__ void
  __launch_bounds__(128) k_ad_integrate_difference(const AdEstimateArgs A, const AdEstimateTables Tb)
  {
    constexpr int     n = p + 1, n3 = n * n * n;
    __shared__ double U[n3], red[4];
    const int         cell = blockIdx.x, lev = A.cell_level[cell], nq = Tb.nq;
    for (int l = threadIdx.x; l < n3; l += blockDim.x)
      U[l] = ad_est_read(A, A.cell_index[(size_t)cell * n3 + l]);
    __syncthreads();
    double h[3], lo[3];
    for (int d = 0; d < 3; ++d)
      {
        h[d]  = A.root_size[d] / (double)(1 << lev);
        lo[d] = A.lower[d] + h[d] * A.cell_pos[cell * 3 + d];
      }
    double sum = 0;
    for (int q = threadIdx.x; q < nq * nq * nq; q += blockDim.x)
      {
        const int q0 = q % nq, q1 = (q / nq) % nq, q2 = q / (nq * nq);
        double    v = 0, g[3] = {0, 0, 0};
        for (int k = 0; k < n; ++k)
          for (int j = 0; j < n; ++j)
            {
              const double s12 = Tb.S[q1 * n + j] * Tb.S[q2 * n + k];
              const double g1 = Tb.G[q1 * n + j] * Tb.S[q2 * n + k], g2 = Tb.S[q1 * n + j] * Tb.G[q2 * n + k];
              for (int i = 0; i < n; ++i)
                {
                  const double u = U[(k * n + j) * n + i], s0 = Tb.S[q0 * n + i];
                  v += u * s0 * s12;
                  g[0] += u * Tb.G[q0 * n + i] * s12;
                  g[1] += u * s0 * g1;
                  g[2] += u * s0 * g2;
                }
            }
        double ev, eg[3];
        ad_exact(A.function_id, lo[0] + h[0] * Tb.x[q0], lo[1] + h[1] * Tb.x[q1], ev, eg);
        const double JxW = Tb.w[q0] * Tb.w[q1] * Tb.w[q2] * h[0] * h[1] * h[2];
        if (A.norm == 0)
          sum += JxW * (v - ev) * (v - ev);
        else
          for (int d = 0; d < 3; ++d)
            sum += JxW * (g[d] / h[d] - eg[d]) * (g[d] / h[d] - eg[d]);
      }
    for (int o = 16; o > 0; o >>= 1)
      sum += __shfl_down_sync(0xffffffffu, sum, o);
    if ((threadIdx.x & 31) == 0)
      red[threadIdx.x >> 5] = sum;
    __syncthreads();
    if (threadIdx.x == 0)
      A.out[cell] = red[0] + red[1] + red[2] + red[3];
  }

  // dst += f (1, phi_i) through the operator's table: one block per cell adds the cell's share f |K| m_i m_j m_k with the
  // 1-D integrals m of the nodal basis
  template <int p>
  __global__ void
  k_ad_rhs_constant(const AdaptiveArgs<double> A, const double f, const AdEstimateTables Tb, const double root_volume)
  {
    constexpr int n = p + 1, n3 = n * n * n;
    const int     cell = blockIdx.x, lev = A.cell_level[cell];
    const double  vol  = root_volume / (double)(1 << (3 * lev));
    for (int l = threadIdx.x; l < n3; l += blockDim.x)
      {
        double m[3] = {0, 0, 0};
        const int idx[3] = {l % n, (l / n) % n, l / (n * n)};
        for (int d = 0; d < 3; ++d)
          for (int q = 0; q < Tb.nq; ++q)
            m[d] += Tb.w[q] * Tb.S[q * n + idx[d]];
        ad_add(A, A.cell_index[(size_t)cell * n3 + l], f * vol * m[0] * m[1] * m[2]);
      }
  }

  __global__ void
  k_ad_axpy(const size_t n, double *u, const double *v)
  {
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
      u[i] += v[i];
  }

  __global__ void
  k_ad_negate(const size_t n, double *v)
  {
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
      v[i] = -v[i];
  }

  struct AdEstimateData
  {
    int32_t *cell_index_plain = nullptr, *cell_pos = nullptr, *nbr = nullptr;
    int8_t  *nbr_type = nullptr;
    double  *face_int = nullptr, *cell_out = nullptr, *u_buf = nullptr, *v_buf = nullptr;
    ~AdEstimateData()
    {
      for (void *ptr : {(void *)cell_index_plain, (void *)cell_pos, (void *)nbr, (void *)nbr_type, (void *)face_int,
                        (void *)cell_out, (void *)u_buf, (void *)v_buf})
        if (ptr)
          cudaFree(ptr);
    }
  };

  inline AdEstimateData &
  ad_estimate_data(AdaptiveOperator &o)
  {
    o.to_device();
    if (!o.est)
      {
        AdaptiveNumbering &N = o.numbering();
        auto              *d = new AdEstimateData();
        o.est                = d;
        o.est_deleter        = [](void *ptr) { delete static_cast<AdEstimateData *>(ptr); };
        std::vector<int8_t>  type;
        std::vector<int32_t> ids, pos(o.mesh.cells.size() * 3);
        o.mesh.face_neighbours(type, ids);
        for (size_t c = 0; c < o.mesh.cells.size(); ++c)
          for (int k = 0; k < 3; ++k)
            pos[3 * c + k] = o.mesh.cells[c].c[k];
        d->cell_index_plain = AdaptiveOperator::upload(N.cell_index_plain, o.stream);
        d->cell_pos         = AdaptiveOperator::upload(pos, o.stream);
        d->nbr              = AdaptiveOperator::upload(ids, o.stream);
        d->nbr_type         = AdaptiveOperator::upload(type, o.stream);
        MGB_AD_CHECK(cudaMalloc(&d->face_int, o.mesh.cells.size() * 6 * sizeof(double)));
        MGB_AD_CHECK(cudaMalloc(&d->cell_out, o.mesh.cells.size() * sizeof(double)));
        MGB_AD_CHECK(cudaMalloc(&d->u_buf, N.n_dofs() * sizeof(double)));
        MGB_AD_CHECK(cudaMalloc(&d->v_buf, N.n_dofs() * sizeof(double)));
        MGB_AD_CHECK(cudaStreamSynchronize(o.stream));
      }
    return *static_cast<AdEstimateData *>(o.est);
  }

  // tables of the nodal basis of degree p at nq Gauss points
  inline AdEstimateTables
  ad_estimate_tables(const int p, const int nq)
  {
    if (p + 1 > AD_EST_MAXN || nq > AD_EST_MAXN || nq < 1)
      throw std::runtime_error("too many quadrature points");
    const Basis                      b = make_basis(p), bq = make_basis(nq - 1); // Gauss points of the nq-point rule
    std::vector<detail::ld>          nodes(b.nodes.begin(), b.nodes.end());
    AdEstimateTables                 T;
    std::memset(&T, 0, sizeof(T));
    const int n = p + 1;
    T.nq        = nq;
    for (int i = 0; i < n; ++i)
      {
        T.e[0][i] = (double)detail::dlag(nodes, i, 0.L);
        T.e[1][i] = (double)detail::dlag(nodes, i, 1.L);
      }
    for (int q = 0; q < nq; ++q)
      {
        T.x[q] = bq.xq[q];
        T.w[q] = bq.wq[q];
        for (int i = 0; i < n; ++i)
          {
            T.S[q * n + i] = (double)detail::lag(nodes, i, (detail::ld)bq.xq[q]);
            T.G[q * n + i] = (double)detail::dlag(nodes, i, (detail::ld)bq.xq[q]);
            for (int o = 0; o < 2; ++o)
              T.Sh[o][q * n + i] = (double)detail::lag(nodes, i, ((detail::ld)o + (detail::ld)bq.xq[q]) / 2);
          }
      }
    return T;
  }

  inline AdEstimateArgs
  ad_estimate_args(AdaptiveOperator &o, const double *u)
  {
    AdEstimateData &d = ad_estimate_data(o);
    AdEstimateArgs  A;
    A.n_cells    = (int)o.mesh.cells.size();
    A.cell_index = d.cell_index_plain;
    A.row_start  = o.d_row_start;
    A.row_dof    = o.d_row_dof;
    A.row_w      = (const double *)o.d_row_w[1];
    A.cell_level = o.d_cell_level;
    A.cell_pos   = d.cell_pos;
    A.nbr_type   = d.nbr_type;
    A.nbr        = d.nbr;
    A.u          = u;
    for (int k = 0; k < 3; ++k)
      {
        A.lower[k]     = o.mesh.lower[k];
        A.root_size[k] = o.mesh.root_size[k];
      }
    A.face_int    = d.face_int;
    A.out         = d.cell_out;
    A.function_id = 0;
    A.norm        = 0;
    return A;
  }

#define MGB_AD_DEGREE_SWITCH(P, CALL)            \
  switch (P)                                     \
    {                                            \
      case 1: { constexpr int p_ = 1; CALL; } break; \
      case 2: { constexpr int p_ = 2; CALL; } break; \
      case 3: { constexpr int p_ = 3; CALL; } break; \
      case 4: { constexpr int p_ = 4; CALL; } break; \
      case 5: { constexpr int p_ = 5; CALL; } break; \
      case 6: { constexpr int p_ = 6; CALL; } break; \
      case 7: { constexpr int p_ = 7; CALL; } break; \
      default: { constexpr int p_ = 8; CALL; } break; \
    }

  // KellyErrorEstimator<3>::estimate (poisson_l/program.cc:524-531) on a device vector u that carries the boundary values
  inline void
  ad_kelly(AdaptiveOperator &o, const double *u, float *indicators_host)
  {
    AdEstimateArgs         A  = ad_estimate_args(o, u);
    const AdEstimateTables Tb = ad_estimate_tables(o.mesh.p, o.mesh.p + 1);
    MGB_AD_DEGREE_SWITCH(o.mesh.p, (k_ad_face_integrals<p_><<<A.n_cells, 128, 0, o.stream>>>(A, Tb)));
    k_ad_kelly_sum<<<std::min((A.n_cells + 255) / 256, 148 * 8), 256, 0, o.stream>>>(A);
    o.launches += 2;
    MGB_AD_CHECK(cudaGetLastError());
    std::vector<double> eta(A.n_cells);
    MGB_AD_CHECK(cudaMemcpyAsync(eta.data(), A.out, eta.size() * sizeof(double), cudaMemcpyDeviceToHost, o.stream));
    MGB_AD_CHECK(cudaStreamSynchronize(o.stream));
    for (int c = 0; c < A.n_cells; ++c)
      indicators_host[c] = (float)eta[c];
  }

  // VectorTools::integrate_difference: per-cell integrals (squared) to the host; returns the norm
  inline double
  ad_integrate_difference(AdaptiveOperator &o, const double *u, const int function_id, const int norm, const int nq,
                          double *cell_squares_host)
  {
    if (norm != 0 && norm != 1)
      throw std::runtime_error("norm: 0 = L2, 1 = H1 seminorm");
    if (function_id != 0 && function_id != 1)
      throw std::runtime_error("function: 0 = zero, 1 = L-shape singularity");
    AdEstimateArgs A = ad_estimate_args(o, u);
    A.function_id    = function_id;
    A.norm           = norm;
    const AdEstimateTables Tb = ad_estimate_tables(o.mesh.p, nq);
    MGB_AD_DEGREE_SWITCH(o.mesh.p, (k_ad_integrate_difference<p_><<<A.n_cells, 128, 0, o.stream>>>(A, Tb)));
    ++o.launches;
    MGB_AD_CHECK(cudaGetLastError());
    std::vector<double> sq(A.n_cells);
    MGB_AD_CHECK(cudaMemcpyAsync(sq.data(), A.out, sq.size() * sizeof(double), cudaMemcpyDeviceToHost, o.stream));
    MGB_AD_CHECK(cudaStreamSynchronize(o.stream));
    double total = 0;
    for (double v : sq)
      total += v;
    if (cell_squares_host)
      std::copy(sq.begin(), sq.end(), cell_squares_host);
    return std::sqrt(total);
  }

  // LaplaceOperator::compute_residual(dst, src, ConstantFunction(f)) (laplace_operator.h:806-845, poisson_l/program.cc:
  // 546-549): dst = (f, phi) - K src on the rows without Dirichlet condition, src carrying the boundary values
  inline void
  ad_compute_residual(AdaptiveOperator &o, double *dst, const double *src, const double f)
  {
    AdEstimateData &d = ad_estimate_data(o);
    const size_t    N = o.numbering().n_dofs();
    MGB_AD_CHECK(cudaMemsetAsync(dst, 0, N * sizeof(double), o.stream));
    AdaptiveArgs<double> A = o.args<double>(src, dst);
    A.cell_index           = d.cell_index_plain;
    A.cell_index_out       = o.d_cell_index;
    dispatch_adaptive<double>(o.mesh.p, A, false, o.stream);
    k_ad_negate<<<(unsigned)std::min<size_t>((N + 255) / 256, 148 * 8), 256, 0, o.stream>>>(N, dst);
    o.launches += 2;
    if (f != 0.)
      {
        A.cell_index           = o.d_cell_index;
        A.cell_index_out       = nullptr;
        const AdEstimateTables Tb = ad_estimate_tables(o.mesh.p, o.mesh.p + 1);
        const double           rv = o.mesh.root_size[0] * o.mesh.root_size[1] * o.mesh.root_size[2];
        MGB_AD_DEGREE_SWITCH(o.mesh.p, (k_ad_rhs_constant<p_><<<A.n_cells, 128, 0, o.stream>>>(A, f, Tb, rv)));
        ++o.launches;
      }
    MGB_AD_CHECK(cudaGetLastError());
  }
} // namespace mgb

extern "C"
{
  int
  mgb_adaptive_refine_and_coarsen(mgb_adaptive h, const unsigned char *refine_flags, const unsigned char *coarsen_flags)
  {
    AD_TRY
    AdaptiveOperator &o = ADH(h);
    if (!refine_flags)
      throw std::runtime_error("null argument");
    o.invalidate();
    o.mesh.refine_and_coarsen(refine_flags, coarsen_flags);
    AD_CATCH
  }

  // GridRefinement::refine_and_coarsen_fixed_number (serial semantics): the int(top_fraction N) cells with the largest
  // indicators are refined, the int(bottom_fraction N) with the smallest coarsened; ties at the two thresholds are flagged
  int
  mgb_adaptive_refine_and_coarsen_fixed_number(mgb_adaptive h, const float *indicators, double top_fraction, double bottom_fraction)
  {
    AD_TRY
    AdaptiveOperator &o = ADH(h);
    if (!indicators)
      throw std::runtime_error("null argument");
    if (!(top_fraction >= 0 && top_fraction <= 1 && bottom_fraction >= 0 && bottom_fraction <= 1 && top_fraction + bottom_fraction <= 1))
      throw std::runtime_error("fractions must lie in [0, 1] and sum to at most 1");
    const size_t         N = o.mesh.cells.size();
    const int            n_ref = (int)(top_fraction * (double)N), n_coa = (int)(bottom_fraction * (double)N);
    std::vector<uint8_t> ref(N, 0), coa(N, 0);
    std::vector<float>   tmp(indicators, indicators + N);
    if (n_ref > 0)
      {
        std::nth_element(tmp.begin(), tmp.begin() + n_ref - 1, tmp.end(), std::greater<float>());
        const float threshold = tmp[n_ref - 1];
        for (size_t c = 0; c < N; ++c)
          ref[c] = indicators[c] >= threshold;
      }
    if (n_coa > 0)
      {
        std::nth_element(tmp.begin(), tmp.begin() + n_coa - 1, tmp.end());
        const float threshold = tmp[n_coa - 1];
        for (size_t c = 0; c < N; ++c)
          coa[c] = !ref[c] && indicators[c] <= threshold;
      }
    o.invalidate();
    o.mesh.refine_and_coarsen(ref.data(), coa.data());
    AD_CATCH
  }

  int
  mgb_adaptive_face_neighbours(mgb_adaptive h, signed char *type, int32_t *ids)
  {
    AD_TRY
    std::vector<int8_t>  t;
    std::vector<int32_t> i;
    ADH(h).mesh.face_neighbours(t, i);
    if (type)
      std::copy(t.begin(), t.end(), type);
    if (ids)
      std::copy(i.begin(), i.end(), ids);
    AD_CATCH
  }

  int
  mgb_adaptive_plain_table(mgb_adaptive h, int32_t *cell_index_plain)
  {
    AD_TRY
    AdaptiveNumbering &N = ADH(h).numbering();
    std::copy(N.cell_index_plain.begin(), N.cell_index_plain.end(), cell_index_plain);
    AD_CATCH
  }

  int
  mgb_adaptive_interpolate_boundary_values(mgb_adaptive h, int function_id, double *u_dev)
  {
    AD_TRY
    AdaptiveOperator &o = ADH(h);
    if (function_id != 0 && function_id != 1)
      throw std::runtime_error("function: 0 = zero, 1 = L-shape singularity");
    o.to_device();
    AdaptiveNumbering  &N = o.numbering();
    std::vector<double> u(N.n_dofs(), 0.);
    for (size_t g = 0; g < N.n_dofs(); ++g)
      if (N.dirichlet[g])
        {
          double grad[3];
          ad_exact(function_id, N.points[3 * g], N.points[3 * g + 1], u[g], grad);
        }
    MGB_AD_CHECK(cudaMemcpyAsync(u_dev, u.data(), u.size() * sizeof(double), cudaMemcpyHostToDevice, o.stream));
    MGB_AD_CHECK(cudaStreamSynchronize(o.stream));
    AD_CATCH
  }

  int
  mgb_adaptive_compute_residual(mgb_adaptive h, double *dst_dev, const double *src_dev, double f)
  {
    AD_TRY
    ad_compute_residual(ADH(h), dst_dev, src_dev, f);
    AD_CATCH
  }

  int
  mgb_adaptive_estimate_kelly(mgb_adaptive h, const double *u_dev, float *indicators_host)
  {
    AD_TRY
    if (!u_dev || !indicators_host)
      throw std::runtime_error("null argument");
    ad_kelly(ADH(h), u_dev, indicators_host);
    AD_CATCH
  }

  int
  mgb_adaptive_integrate_difference(mgb_adaptive h, const double *u_dev, int function_id, int norm, int n_q_points,
                                    double *result, double *cell_squares_host)
  {
    AD_TRY
    if (!u_dev || !result)
      throw std::runtime_error("null argument");
    *result = ad_integrate_difference(ADH(h), u_dev, function_id, norm, n_q_points, cell_squares_host);
    AD_CATCH
  }

  // one cycle of LaplaceProblem::run after the mesh is set (poisson_l/program.cc:545-575): boundary values, right-hand
  // side, CG with the local-smoothing V-cycle for the update, solution = boundary values + update; the solution stays on
  // the device in a buffer owned by the operator.  stats = {CG iterations, reduction rate, L2 error (QGauss(p+2)),
  // H1-seminorm error (QGauss(p+1))}
  int
  mgb_adaptive_solve_cycle(mgb_adaptive h, int function_id, double f, int n_smoother, double stats[4], double **solution_dev)
  {
    AD_TRY
    AdaptiveOperator &o = ADH(h);
    if (mgb_adaptive_mg_setup(h, n_smoother) != 0)
      throw std::runtime_error(mgb_last_error());
    AdEstimateData &d = ad_estimate_data(o);
    const size_t    N = o.numbering().n_dofs();
    if (mgb_adaptive_interpolate_boundary_values(h, function_id, d.u_buf) != 0)
      throw std::runtime_error(mgb_last_error());
    for (int k = 0; k < 2; ++k)
      if (!o.d_buf[k])
        MGB_AD_CHECK(cudaMalloc(&o.d_buf[k], N * 8));
    double *rhs = (double *)o.d_buf[0], *upd = (double *)o.d_buf[1];
    ad_compute_residual(o, rhs, d.u_buf, f);
    unsigned int its  = 0;
    double       rate = 0;
    if (mgb_adaptive_solve_cg(h, upd, rhs, 1e-9, &its, &rate, nullptr, nullptr) != 0)
      throw std::runtime_error(mgb_last_error());
    // constraints.distribute(update) sets the Dirichlet rows of the update to zero: solve_cg leaves them at rhs = 0
    {
      double       *u = d.u_buf;
      const double *v = upd;
      k_ad_axpy<<<(unsigned)std::min<size_t>((N + 255) / 256, 148 * 8), 256, 0, o.stream>>>(N, u, v);
      ++o.launches;
      MGB_AD_CHECK(cudaGetLastError());
    }
    stats[0] = its;
    stats[1] = rate;
    stats[2] = ad_integrate_difference(o, d.u_buf, function_id, 0, o.mesh.p + 2, nullptr);
    stats[3] = ad_integrate_difference(o, d.u_buf, function_id, 1, o.mesh.p + 1, nullptr);
    if (solution_dev)
      *solution_dev = d.u_buf;
    AD_CATCH
  }
}
