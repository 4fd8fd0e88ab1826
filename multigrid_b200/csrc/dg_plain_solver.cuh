// Pure DG h-multigrid: multigrid::MultigridSolverDGPlain<3, p, float, double, type> (common/multigrid_solver_dg_plain.h):
// the SIP-DG operator on EVERY level of the refinement hierarchy, Chebyshev over JacobiTransformed as smoother
// (:192-213: range 20, degree degree_pre on the inner levels and max(1, degree_pre - 1) on the finest, 15 CG steps for the
// eigenvalue estimate; coarsest level: tolerance 1e-5, automatic degree, m() CG steps), the cell-wise embedding of the DG
// space into the children as level transfer (MGTransferMatrixFree on FE_DGQ / FE_DGQHermite / FE_DGQArbitraryNodes,
// restriction = transpose), one float V-cycle (:456-497) as preconditioner of CG in double (:304-316).
// Included by dg.cu after dg_solver.cuh.  One GPU (the levels are whole meshes, not z-slabs).
#pragma once

namespace mgb
{
  // Level transfer of a cell-wise DG vector, one block of 128 threads per cell of the level the result lives on.
  // RESTRICT = false: fine cell <- (E_ox x E_oy x E_oz) parent cell (prolongate; ADD: prolongate_and_add)
  // RESTRICT = true : coarse cell <- sum over its 8 children of (E^T x E^T x E^T) child (restrict_and_add into a zeroed vector)
  template <int n, typename T, bool RESTRICT, bool ADD>
  __global__ void
  __launch_bounds__(128)
    k_dg_level_transfer(const int ncx, const int ncy, const int ncz /* cells of the COARSE level */, const DGMatrix1D E0,
                        const DGMatrix1D E1, const T *__restrict__ src, T *__restrict__ dst)
  {
    constexpr int n2 = n * n, n3 = n2 * n;
    __shared__ T  buf[2][n3];
    __shared__ T  acc[RESTRICT ? n3 : 1];
    const size_t  cell = blockIdx.x;
    // pencil-wise sweep of buf[0] -> buf[1] -> buf[0] -> buf[1] with the matrices of the octant
    auto sweeps = [&](const int oct) {
#pragma unroll
      for (int dir = 0; dir < 3; ++dir)
        {
          const T          *a  = buf[dir & 1];
          T                *b  = buf[(dir + 1) & 1];
          const int         st = dir == 0 ? 1 : (dir == 1 ? n : n2);
          const DGMatrix1D &E  = ((oct >> dir) & 1) ? E1 : E0;
          for (int l = threadIdx.x; l < n2; l += 128)
            {
              const int u = l % n, v = l / n;
              const int base = dir == 0 ? (v * n + u) * n : (dir == 1 ? v * n2 + u : v * n + u);
              T         x[n];
#pragma unroll
              for (int k = 0; k < n; ++k)
                x[k] = a[base + k * st];
#pragma unroll
              for (int r = 0; r < n; ++r)
                {
                  double s = 0;
#pragma unroll
                  for (int k = 0; k < n; ++k)
                    s += (RESTRICT ? E.m[k * n + r] : E.m[r * n + k]) * (double)x[k];
                  b[base + r * st] = (T)s;
                }
            }
          __syncthreads();
        }
    };
    if (!RESTRICT)
      {
        const int    nfx = 2 * ncx, nfy = 2 * ncy;
        const int    fx = (int)(cell % nfx), fy = (int)((cell / nfx) % nfy), fz = (int)(cell / ((size_t)nfx * nfy));
        const size_t parent = ((size_t)(fz / 2) * ncy + fy / 2) * ncx + fx / 2;
        for (int i = threadIdx.x; i < n3; i += 128)
          buf[0][i] = src[parent * n3 + i];
        __syncthreads();
        sweeps((fx & 1) | ((fy & 1) << 1) | ((fz & 1) << 2));
        for (int i = threadIdx.x; i < n3; i += 128)
          {
            if (ADD)
              dst[cell * n3 + i] += buf[1][i];
            else
              dst[cell * n3 + i] = buf[1][i];
          }
      }
    else
      {
        const int cx = (int)(cell % ncx), cy = (int)((cell / ncx) % ncy), cz = (int)(cell / ((size_t)ncx * ncy));
        for (int i = threadIdx.x; i < n3; i += 128)
          acc[RESTRICT ? i : 0] = 0;
        for (int oct = 0; oct < 8; ++oct)
          {
            const size_t child = ((size_t)(2 * cz + (oct >> 2)) * (2 * ncy) + 2 * cy + ((oct >> 1) & 1)) * (2 * ncx) + 2 * cx + (oct & 1);
            __syncthreads();
            for (int i = threadIdx.x; i < n3; i += 128)
              buf[0][i] = src[child * n3 + i];
            __syncthreads();
            sweeps(oct);
            for (int i = threadIdx.x; i < n3; i += 128)
              acc[RESTRICT ? i : 0] += buf[1][i];
          }
        for (int i = threadIdx.x; i < n3; i += 128)
          {
            if (ADD)
              dst[cell * n3 + i] += acc[RESTRICT ? i : 0];
            else
              dst[cell * n3 + i] = acc[RESTRICT ? i : 0];
          }
      }
  }

  struct DGPlainSolver
  {
    mgb_dg_solver_options                    opt;
    int                                      n_levels = 0;
    cudaStream_t                             stream   = nullptr;
    std::vector<std::unique_ptr<DGOperator>> ops; // matrix[level]; the finest one also serves as matrix_dg_dp
    std::vector<DGChebyshev>                 smooth;
    std::vector<float *>                     defect, t, solution_update;
    DGMatrix1D                               E[2];
    size_t                                   N = 0;
    double             *solution = nullptr, *rhs = nullptr, *g = nullptr, *d = nullptr, *h = nullptr, *basis_dev = nullptr;
    std::vector<double> residual_history;

    explicit DGPlainSolver(const mgb_dg_solver_options &o)
      : opt(o)
    {
      if (o.degree_pre != o.degree_post)
        throw std::runtime_error("pre- and post-smoother degree must agree (multigrid_solver_dg_plain.h:80-82)");
      if (o.degree_pre < 1)
        throw std::runtime_error("smoother degree must be at least 1");
      if (o.comm_size > 1)
        throw std::runtime_error("the pure DG multigrid solver runs on one GPU");
      if (o.n_refinements < 0 || o.n_refinements > 10)
        throw std::runtime_error("invalid number of refinements");
      try
        {
          setup();
        }
      catch (...)
        {
          cleanup();
          throw;
        }
    }
    ~DGPlainSolver() { cleanup(); }

    void
    setup()
    {
      n_levels = opt.n_refinements + 1;
      smooth.resize(n_levels);
      for (int l = 0; l < n_levels; ++l)
        {
          mgb_dg_options dopt;
          std::memset(&dopt, 0, sizeof(dopt));
          dopt.degree        = opt.degree;
          dopt.n_refinements = l;
          for (int k = 0; k < 3; ++k)
            {
              dopt.n_subdivisions[k] = opt.n_subdivisions[k];
              dopt.lower[k]          = opt.lower[k];
              dopt.upper[k]          = opt.upper[k];
              dopt.trafo[4 * k]      = 1.;
            }
          dopt.device    = opt.device;
          dopt.basis     = opt.basis;
          dopt.comm_size = 1;
          ops.emplace_back(new DGOperator(dopt));
          if (l == 0)
            stream = ops[0]->stream;
          else
            ops[l]->use_stream(stream);
          ops[l]->preconditioner = opt.preconditioner;
          const size_t nl = ops[l]->n_dofs;
          float       *v[3];
          for (int k = 0; k < 3; ++k)
            {
              MGB_DG_CHECK(cudaMalloc(&v[k], nl * sizeof(float)));
              MGB_DG_CHECK(cudaMemsetAsync(v[k], 0, nl * sizeof(float), stream));
            }
          defect.push_back(v[0]);
          t.push_back(v[1]);
          solution_update.push_back(v[2]);
        }
      DGOperator &fine = *ops.back();
      N                = fine.n_dofs;
      for (double **v : {&solution, &rhs, &g, &d, &h})
        {
          MGB_DG_CHECK(cudaMalloc(v, N * sizeof(double)));
          MGB_DG_CHECK(cudaMemsetAsync(*v, 0, N * sizeof(double), stream));
        }
      const int n = fine.n;
      for (int o = 0; o < 2; ++o)
        for (int i = 0; i < n * n; ++i)
          E[o].m[i] = fine.basis.E[o][i];
      std::vector<double> hb(fine.basis.S);
      hb.insert(hb.end(), fine.basis.xq.begin(), fine.basis.xq.end());
      hb.insert(hb.end(), fine.basis.wq.begin(), fine.basis.wq.end());
      MGB_DG_CHECK(cudaMalloc(&basis_dev, hb.size() * sizeof(double)));
      MGB_DG_CHECK(cudaMemcpyAsync(basis_dev, hb.data(), hb.size() * sizeof(double), cudaMemcpyHostToDevice, stream));
      MGB_DG_CHECK(cudaStreamSynchronize(stream));
      cell_integrals(0, FN_CUBE_RHS, nullptr, rhs); // (f, phi_i), multigrid_solver_dg_plain.h:162-185
      // smoothers (:192-213)
      for (int l = 0; l < n_levels; ++l)
        {
          DGChebyshev &s = smooth[l];
          if (l > 0)
            {
              s.smoothing_range     = 20.;
              s.degree              = l < n_levels - 1 ? (unsigned int)opt.degree_pre : (unsigned int)std::max(1, opt.degree_pre - 1);
              s.eig_cg_n_iterations = 15;
            }
          else
            {
              s.smoothing_range     = 1e-5;
              s.degree              = DGChebyshev::invalid_degree;
              s.eig_cg_n_iterations = (unsigned int)ops[0]->n_dofs;
            }
          s.initialize(*ops[l]);
        }
    }

    void
    cleanup()
    {
      for (double *v : {solution, rhs, g, d, h, basis_dev})
        if (v)
          cudaFree(v);
      solution = rhs = g = d = h = basis_dev = nullptr;
      for (auto *list : {&defect, &t, &solution_update})
        {
          for (float *v : *list)
            if (v)
              cudaFree(v);
          list->clear();
        }
      smooth.clear(); // before the operators they point to
      ops.clear();
    }

    void
    cell_integrals(const int mode, const int fn, const double *u, double *out)
    {
      DGOperator &op = *ops.back();
      DGCellGeom  G;
      G.n        = op.n;
      double det = 1;
      for (int k = 0; k < 3; ++k)
        {
          G.nc[k]    = op.nc[k];
          G.lower[k] = opt.lower[k];
          G.h[k]     = (opt.upper[k] - opt.lower[k]) / op.nc[k];
          det *= G.h[k];
        }
      for (int k = 0; k < 9; ++k)
        G.T[k] = (k % 4 == 0) ? 1. : 0.;
      G.jxw                 = det;
      G.cell_begin          = 0;
      G.n_cells             = (size_t)op.nc[0] * op.nc[1] * op.nc[2];
      G.cz_shift            = 0;
      const int      n      = op.n;
      const size_t   smem   = (size_t)(n * n * n + n * n) * sizeof(double);
      const unsigned blocks = (unsigned)std::min<size_t>(G.n_cells, 148 * 32);
      k_dg_cell_integrals<<<blocks, 128, smem, stream>>>(G, mode, fn, basis_dev, basis_dev + n * n, basis_dev + n * n + n, u, out);
      ++op.launches;
      MGB_DG_CHECK(cudaGetLastError());
    }

    double
    compute_l2_error()
    {
      DGOperator &op = *ops.back();
      MGB_DG_CHECK(cudaMemsetAsync(op.scal_dev, 0, 4 * sizeof(double), stream));
      cell_integrals(1, FN_CUBE_SOLUTION, solution, op.scal_dev);
      MGB_DG_CHECK(cudaMemcpyAsync(op.scal_host, op.scal_dev, 4 * sizeof(double), cudaMemcpyDeviceToHost, stream));
      MGB_DG_CHECK(cudaStreamSynchronize(stream));
      return std::sqrt(op.scal_host[0] / op.scal_host[1]);
    }

    template <bool RESTRICT, bool ADD>
    void
    transfer(const int fine_level, float *dst, const float *src)
    {
      const DGOperator &c     = *ops[fine_level - 1];
      const size_t      cells = RESTRICT ? (size_t)c.nc[0] * c.nc[1] * c.nc[2] : 8 * (size_t)c.nc[0] * c.nc[1] * c.nc[2];
      switch (c.n)
        {
#define MGB_DGT(NN)                                                                                                      \
  case NN:                                                                                                               \
    k_dg_level_transfer<NN, float, RESTRICT, ADD><<<(unsigned)cells, 128, 0, stream>>>(c.nc[0], c.nc[1], c.nc[2], E[0], E[1], src, dst); \
    break;
          MGB_DGT(2) MGB_DGT(3) MGB_DGT(4) MGB_DGT(5) MGB_DGT(6) MGB_DGT(7) MGB_DGT(8) MGB_DGT(9)
#undef MGB_DGT
          default: throw std::runtime_error("unsupported degree");
        }
      ++ops[fine_level]->launches;
      MGB_DG_CHECK(cudaGetLastError());
    }

    // multigrid_solver_dg_plain.h:456-497 with my_n_cycles = 1
    void
    v_cycle(const int level)
    {
      if (level == 0)
        {
          smooth[0].vmult(solution_update[0], defect[0]); // MGCoarseFromSmoother
          return;
        }
      smooth[level].vmult(solution_update[level], defect[level]);
      ops[level]->apply<float>(DG_RESIDUAL, t[level], solution_update[level], defect[level]);
      transfer<true, false>(level, defect[level - 1], t[level]); // defect[level-1] = 0; restrict_and_add
      v_cycle(level - 1);
      transfer<false, true>(level, solution_update[level], solution_update[level - 1]); // prolongate_and_add
      smooth[level].step(solution_update[level], defect[level]);
    }

    // preconditioner interface (:321-337): dst = M^-1 src in double, the cycle in float
    void
    vmult(double *dst, const double *src)
    {
      DGOperator &op  = *ops.back();
      float      *def = defect.back();
      op.foreach (N, [=] __device__(size_t i) { def[i] = (float)src[i]; });
      v_cycle(n_levels - 1);
      const float *upd = solution_update.back();
      op.foreach (N, [=] __device__(size_t i) { dst[i] = (double)upd[i]; });
    }

    // :340-427
    void
    vmult_with_residual_update(double *residual, double *update, const double factor, double results[2])
    {
      dg_vmult_with_residual_update(
        *ops.back(), N, residual, update, factor, defect.back(),
        [&]() -> const float * {
          v_cycle(n_levels - 1);
          return solution_update.back();
        },
        results);
    }

    void
    solve_cg(const double tolerance, double *iterations_estimate, double *reduction_rate, unsigned int *n_iterations)
    {
      double             res0 = 0, res = 0;
      const unsigned int it   = dg_solve_cg(
        *ops.back(), N, solution, rhs, g, d, h, defect.back(),
        [&]() -> const float * {
          v_cycle(n_levels - 1);
          return solution_update.back();
        },
        tolerance, residual_history, res0, res);
      const double rate = it > 0 ? std::pow(res / res0, 1. / it) : 0.;
      if (reduction_rate)
        *reduction_rate = rate;
      if (iterations_estimate)
        *iterations_estimate = it > 0 ? std::log(tolerance) / std::log(rate) : 0.;
      if (n_iterations)
        *n_iterations = it;
    }

    int64_t
    launch_count() const
    {
      int64_t s = 0;
      for (const auto &o : ops)
        s += o->launches;
      return s;
    }
  };
} // namespace mgb
