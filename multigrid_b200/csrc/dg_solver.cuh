// DG multigrid solver: included by dg.cu after DGOperator.
//   DGMultigridSolver  <->  multigrid::MultigridSolverDG<3,p,float,double>   common/multigrid_solver_dg.h:44-735
//   DGChebyshev        <->  PreconditionChebyshev<LaplaceOperatorCompactCombine<float>, Vector<float>, JacobiTransformed>
//                           (multigrid_solver_dg.h:294-304, 709-712)
// The FE_Q hierarchy below the DG space is a mgb_solver (solver.cu) driven through the C ABI; the DG operator runs
// on that solver's stream, so one cycle is one stream-ordered sequence of kernels with no host synchronisation.
#pragma once

namespace mgb
{
  struct DGCellGeom
  {
    int    n, nc[3];
    size_t cell_begin, n_cells; // owned cells of this rank in its local numbering
    int    cz_shift;            // global cell layer of local layer 0
    double lower[3], h[3], T[9], jxw; // x = T (lower + h (c + xi)), jxw = |det(T diag(h))|
  };

  // One block per cell.  mode 0: rhs_i = sum_q phi_i(x_q) f(x_q) JxW_q   (multigrid_solver_dg.h:243-265)
  //                      mode 1: out[0] += sum_q (u_h(x_q) - u(x_q))^2 JxW_q, out[1] += sum_q JxW_q   (:328-366)
  // S[q n + i] = phi_i(x_q), xq / wq Gauss points and weights on [0,1]; all in global memory (set-up path).
  __global__ void
  k_dg_cell_integrals(const DGCellGeom g, const int mode, const int fn, const double *__restrict__ S,
                      const double *__restrict__ xq, const double *__restrict__ wq, const double *__restrict__ u,
                      double *__restrict__ out)
  {
    extern __shared__ double sm[];
    const int                n = g.n, n3 = n * n * n;
    double                  *vq = sm;     // [n^3] values at quadrature points
    double                  *Ss = sm + n3; // [n^2]
    for (int i = threadIdx.x; i < n * n; i += blockDim.x)
      Ss[i] = S[i];
    double       err = 0, vol = 0;
    for (size_t c = g.cell_begin + blockIdx.x; c < g.cell_begin + g.n_cells; c += gridDim.x)
      {
        const int cx = (int)(c % g.nc[0]), cy = (int)((c / g.nc[0]) % g.nc[1]),
                  cz = (int)(c / ((size_t)g.nc[0] * g.nc[1])) + g.cz_shift;
        __syncthreads();
        for (int q = threadIdx.x; q < n3; q += blockDim.x)
          {
            const int    q0 = q % n, q1 = (q / n) % n, q2 = q / (n * n);
            const double r0 = g.lower[0] + g.h[0] * (cx + xq[q0]), r1 = g.lower[1] + g.h[1] * (cy + xq[q1]),
                         r2 = g.lower[2] + g.h[2] * (cz + xq[q2]);
            const double x = g.T[0] * r0 + g.T[1] * r1 + g.T[2] * r2, y = g.T[3] * r0 + g.T[4] * r1 + g.T[5] * r2,
                         z = g.T[6] * r0 + g.T[7] * r1 + g.T[8] * r2;
            const double jxw = wq[q0] * wq[q1] * wq[q2] * g.jxw;
            if (mode == 0)
              vq[q] = eval_fn(fn, x, y, z) * jxw;
            else
              {
                double v = 0;
                for (int i2 = 0; i2 < n; ++i2)
                  for (int i1 = 0; i1 < n; ++i1)
                    {
                      double row = 0;
                      for (int i0 = 0; i0 < n; ++i0)
                        row += Ss[q0 * n + i0] * u[c * n3 + (i2 * n + i1) * n + i0];
                      v += Ss[q2 * n + i2] * Ss[q1 * n + i1] * row;
                    }
                const double e = v - eval_fn(fn, x, y, z);
                err += e * e * jxw;
                vol += jxw;
              }
          }
        if (mode == 0)
          {
            __syncthreads();
            for (int i = threadIdx.x; i < n3; i += blockDim.x)
              {
                const int i0 = i % n, i1 = (i / n) % n, i2 = i / (n * n);
                double    v = 0;
                for (int q2 = 0; q2 < n; ++q2)
                  for (int q1 = 0; q1 < n; ++q1)
                    {
                      double row = 0;
                      for (int q0 = 0; q0 < n; ++q0)
                        row += Ss[q0 * n + i0] * vq[(q2 * n + q1) * n + q0];
                      v += Ss[q2 * n + i2] * Ss[q1 * n + i1] * row;
                    }
                out[c * n3 + i] = v;
              }
          }
      }
    if (mode == 1)
      {
        __shared__ double red[2][8];
        for (int o = 16; o > 0; o >>= 1)
          {
            err += __shfl_down_sync(0xffffffffu, err, o);
            vol += __shfl_down_sync(0xffffffffu, vol, o);
          }
        if ((threadIdx.x & 31) == 0)
          {
            red[0][threadIdx.x >> 5] = err;
            red[1][threadIdx.x >> 5] = vol;
          }
        __syncthreads();
        if (threadIdx.x < 2)
          {
            double v = 0;
            for (int w = 0; w < (int)(blockDim.x >> 5); ++w)
              v += red[threadIdx.x][w];
            atomicAdd(&out[threadIdx.x], v);
          }
      }
  }

  // deal.II PreconditionChebyshev over the float DG operator; the preconditioner (JacobiTransformed or point
  // Jacobi) is the one selected on the operator.  vmult / step are the fused vmult_with_chebyshev_update calls.
  struct DGChebyshev
  {
    static constexpr unsigned int invalid_degree = 0xffffffffu;
    DGOperator  *A = nullptr;
    double       smoothing_range = 20.;
    unsigned int degree = 3, eig_cg_n_iterations = 15, cg_iterations = 0;
    double       min_eig = 1, max_eig = 1, theta = 1, delta = 1;
    float       *solution_old = nullptr;

    ~DGChebyshev()
    {
      if (solution_old)
        cudaFree(solution_old);
    }

    double
    dot(const float *a, const float *b)
    {
      double r[2];
      A->sums<2>(r, [=] __device__(size_t i, double *acc) { acc[0] += (double)a[i] * (double)b[i]; });
      return r[0];
    }

    // estimate_eigenvalues of deal.II: eig_cg_n_iterations of CG on P^-1 A from "index mod 11, mean removed"
    void
    initialize(DGOperator &op)
    {
      A              = &op;
      const size_t N = op.n_dofs;
      float       *g, *h, *d;
      MGB_DG_CHECK(cudaMalloc(&solution_old, N * sizeof(float)));
      MGB_DG_CHECK(cudaMalloc(&g, N * sizeof(float)));
      MGB_DG_CHECK(cudaMalloc(&h, N * sizeof(float)));
      MGB_DG_CHECK(cudaMalloc(&d, N * sizeof(float)));
      double       sum[2];
      const size_t shift = (size_t)(op.cz0 - op.zg) * op.layer_dofs; // global index of local entry 0
      op.sums<2>(sum, [=] __device__(size_t i, double *acc) { acc[0] += (double)((i + shift) % 11); });
      const float mean = (float)(sum[0] / (double)op.n_dofs_global);
      op.foreach (N, [=] __device__(size_t i) { g[i] = -((float)((i + shift) % 11) - mean); });
      std::vector<double> diag, off;
      double              res = std::sqrt(dot(g, g));
      cg_iterations           = 0;
      if (res > 1e-10 && eig_cg_n_iterations > 0)
        {
          op.preconditioner_vmult<float>(h, g, 1.);
          op.foreach (N, [=] __device__(size_t i) { d[i] = -h[i]; });
          double gh = dot(g, h), alpha_old = 0, beta_old = 0;
          for (unsigned int it = 1;; ++it)
            {
              op.apply<float>(DG_SET, h, d);
              const double alpha = gh / dot(d, h);
              const float  al    = (float)alpha;
              op.foreach (N, [=] __device__(size_t i) { g[i] += al * h[i]; });
              res = std::sqrt(dot(g, g));
              if (it == 1)
                diag.push_back(1. / alpha);
              else
                {
                  diag.push_back(1. / alpha + beta_old / alpha_old);
                  off.push_back(std::sqrt(beta_old) / alpha_old);
                }
              cg_iterations = it;
              if (it >= eig_cg_n_iterations || res <= 1e-10)
                break;
              op.preconditioner_vmult<float>(h, g, 1.);
              const double gh_new = dot(g, h);
              const double beta   = gh_new / gh;
              gh                  = gh_new;
              const float be      = (float)beta;
              op.foreach (N, [=] __device__(size_t i) { d[i] = be * d[i] - h[i]; });
              alpha_old = alpha;
              beta_old  = beta;
            }
          dg_tridiagonal_extreme_eigenvalues(diag, off, min_eig, max_eig);
        }
      else
        min_eig = max_eig = 1.;
      const double max_est = 1.2 * max_eig;
      const double alpha   = smoothing_range > 1. ? max_est / smoothing_range : std::min(0.9 * max_est, min_eig);
      if (degree == invalid_degree)
        {
          // coarse solver: smoothing_range is the tolerance (deal.II PreconditionChebyshev, multigrid_solver_dg_plain.h:204-208)
          const double actual_range = max_est / alpha;
          const double sigma        = (1. - std::sqrt(1. / actual_range)) / (1. + std::sqrt(1. / actual_range));
          const double eps          = smoothing_range;
          degree = 1 + static_cast<unsigned int>(std::log(1. / eps + std::sqrt(1. / eps / eps - 1.)) / std::log(1. / sigma));
        }
      theta                = (max_est + alpha) * 0.5;
      delta                = (max_est - alpha) * 0.5;
      MGB_DG_CHECK(cudaStreamSynchronize(op.stream));
      cudaFree(g);
      cudaFree(h);
      cudaFree(d);
    }

    void
    loop(float *&x, const float *b, const bool first_is_index_one)
    {
      if (degree < 2)
        return;
      double       rhok  = delta / theta;
      const double sigma = theta / delta;
      for (unsigned int k = 0; k < degree - 1; ++k)
        {
          const double rhokp = 1. / (2. * sigma - rhok);
          const double f1 = rhokp * rhok, f2 = 2. * rhokp / delta;
          rhok = rhokp;
          A->chebyshev_update<float>(b, first_is_index_one && k == 0 ? 1 : 2, f1, f2, x, solution_old);
          std::swap(x, solution_old);
        }
    }
    // zero initial guess
    void
    vmult(float *&x, const float *b)
    {
      A->chebyshev_update<float>(b, 0, 0., 1. / theta, x, solution_old);
      loop(x, b, true);
    }
    // nonzero initial guess
    void
    step(float *&x, const float *b)
    {
      A->chebyshev_update<float>(b, 1, 0., 1. / theta, x, solution_old);
      std::swap(x, solution_old);
      loop(x, b, false);
    }
  };

  // deal.II SolverCG with ReductionControl(100, 1e-16, tolerance) from a zero start vector, in double on the DG operator
  // `op`, preconditioned by `cycle()` which maps the float vector `defect` to the float vector it returns
  // (multigrid_solver_dg.h:410-424, multigrid_solver_dg_plain.h:304-316).  Returns the iteration count and the final
  // residual norm; `history` receives the residual norms.
  inline unsigned int
  dg_solve_cg(DGOperator &op, const size_t N, double *x, const double *b, double *gp, double *dp, double *hp, float *def,
              const std::function<const float *()> &cycle, const double tolerance, std::vector<double> &history, double &res0_out, double &res_out)
  {
    double r[2];
    history.clear();
    MGB_DG_CHECK(cudaMemsetAsync(x, 0, N * sizeof(double), op.stream));
    op.foreach (N, [=] __device__(size_t i) { gp[i] = -b[i]; }); // g = A x - b with x = 0
    op.sums<2>(r, [=] __device__(size_t i, double *acc) { acc[0] += gp[i] * gp[i]; });
    const double res0 = std::sqrt(r[0]);
    double       res  = res0;
    history.push_back(res0);
    unsigned int it = 0;
    double       gh = 0;
    const auto   converged = [&](const double v) { return v <= 1e-16 || v <= tolerance * res0; };
    if (!converged(res0))
      for (it = 1;; ++it)
        {
          // h = M^-1 g, gh = g.h fused into the conversion back to double
          op.foreach (N, [=] __device__(size_t i) { def[i] = (float)gp[i]; });
          const float *upd = cycle();
          op.sums<2>(r, [=] __device__(size_t i, double *acc) {
            const double hi = (double)upd[i];
            hp[i]           = hi;
            acc[0] += gp[i] * hi;
          });
          const double beta = it > 1 ? r[0] / gh : 0.;
          gh                = r[0];
          if (it > 1)
            op.foreach (N, [=] __device__(size_t i) { dp[i] = beta * dp[i] - hp[i]; });
          else
            op.foreach (N, [=] __device__(size_t i) { dp[i] = -hp[i]; });
          op.apply<double>(DG_SET, hp, dp);
          op.sums<2>(r, [=] __device__(size_t i, double *acc) { acc[0] += dp[i] * hp[i]; });
          const double alpha = gh / r[0];
          op.sums<2>(r, [=] __device__(size_t i, double *acc) {
            x[i] += alpha * dp[i];
            const double gi = gp[i] + alpha * hp[i];
            gp[i]           = gi;
            acc[0] += gi * gi;
          });
          res = std::sqrt(r[0]);
          history.push_back(res);
          if (converged(res))
            break;
          if (it >= 100)
            throw std::runtime_error("SolverCG: no convergence within 100 iterations");
        }
    res0_out = res0;
    res_out  = res;
    return it;
  }

  // vmult_with_residual_update of the two DG solvers (multigrid_solver_dg.h:445-527, multigrid_solver_dg_plain.h:340-427),
  // the preconditioner call of a CG solver with merged vector operations: defect = residual + factor * update in float,
  // one cycle, then residual += factor * update, update = M^-1 defect and the two sums
  // {update_new . residual_new, update_new . (factor * update_old)} in the same sweep
  inline void
  dg_vmult_with_residual_update(DGOperator &op, const size_t N, double *residual, double *update, const double factor,
                                float *def, const std::function<const float *()> &cycle, double results[2])
  {
    if (factor != 0.)
      op.foreach (N, [=] __device__(size_t i) { def[i] = (float)(residual[i] + factor * update[i]); });
    else
      op.foreach (N, [=] __device__(size_t i) { def[i] = (float)residual[i]; });
    const float *mg = cycle();
    double       r[2];
    if (factor != 0.)
      op.sums<2>(r, [=] __device__(size_t i, double *acc) {
        const double m = (double)mg[i], u = update[i] * factor, res = residual[i] + u;
        acc[0] += m * res;
        acc[1] += m * u;
        residual[i] = res;
        update[i]   = m;
      });
    else
      {
        op.sums<2>(r, [=] __device__(size_t i, double *acc) {
          const double m = (double)mg[i];
          acc[0] += m * residual[i];
          update[i] = m;
        });
        r[1] = r[0];
      }
    results[0] = r[0];
    results[1] = r[1];
  }

  struct DGMultigridSolver
  {
    mgb_dg_solver_options opt;
    mgb_solver            fe = nullptr; // FE_Q hierarchy: matrix[l], smooth[l], transfer, coarse (multigrid_solver_dg.h:155-293)
    DGOperator           *op = nullptr; // matrix_dg (float applies) and matrix_dg_dp (double applies)
    DGChebyshev           smooth_dg;
    LevelDesc             Lfine;
    int                   maxlevel = 0;
    size_t                N = 0;
    double               *solution = nullptr, *rhs = nullptr, *g = nullptr, *d = nullptr, *h = nullptr;
    float                *defect_dg = nullptr, *solution_update_dg = nullptr;
    double               *basis_dev = nullptr; // S, xq, wq
    std::vector<double>   residual_history;

    explicit DGMultigridSolver(const mgb_dg_solver_options &o)
      : opt(o)
    {
      if (o.degree_pre != o.degree_post)
        throw std::runtime_error("pre- and post-smoother degree must agree (multigrid_solver_dg.h:83-85)");
      if (o.degree_pre < 1)
        throw std::runtime_error("smoother degree must be at least 1");
      if (o.preconditioner != 0 && o.preconditioner != 1)
        throw std::runtime_error("preconditioner: 0 = point Jacobi, 1 = JacobiTransformed");
      if (o.basis < DG_BASIS_GAUSS_LOBATTO || o.basis > DG_BASIS_GAUSS)
        throw std::runtime_error("unknown DG basis");
      mgb_options fo;
      mgb_default_options(&fo);
      fo.degree        = o.degree;
      fo.n_refinements = o.n_refinements;
      for (int k = 0; k < 3; ++k)
        {
          fo.n_subdivisions[k] = o.n_subdivisions[k];
          fo.lower[k]          = o.lower[k];
          fo.upper[k]          = o.upper[k];
        }
      fo.degree_pre  = o.degree_pre;
      fo.degree_post = o.degree_post;
      fo.n_cycles    = 1;
      fo.problem     = MGB_PROBLEM_CUBE_SINE;
      fo.device      = o.device;
      fo.comm_rank   = o.comm_rank;
      fo.comm_size   = o.comm_size > 1 ? o.comm_size : 1;
      std::memcpy(fo.comm_id, o.comm_id, sizeof(fo.comm_id));
      if (mgb_solver_create(&fo, &fe) != 0)
        throw std::runtime_error(mgb_last_error());
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
    ~DGMultigridSolver() { cleanup(); }

    void
    fe_check(const int code) const
    {
      if (code != 0)
        throw std::runtime_error(mgb_last_error());
    }

    void
    setup()
    {
      maxlevel = mgb_solver_n_levels(fe) - 1;
      // smoothers of the FE_Q levels (multigrid_solver_dg.h:272-293): one degree less on the finest FE_Q level,
      // coarse solver with range 2e-3
      fe_check(mgb_smoother_reinit(fe, 0, 2e-3, -1, -1));
      if (maxlevel > 0)
        fe_check(mgb_smoother_reinit(fe, maxlevel, 20., std::max(1, opt.degree_pre - 1), 15));
      Lfine = *mgb_internal_level_desc(fe, maxlevel);

      mgb_dg_options dopt;
      std::memset(&dopt, 0, sizeof(dopt));
      dopt.degree        = opt.degree;
      dopt.n_refinements = opt.n_refinements;
      for (int k = 0; k < 3; ++k)
        {
          dopt.n_subdivisions[k] = opt.n_subdivisions[k];
          dopt.lower[k]          = opt.lower[k];
          dopt.upper[k]          = opt.upper[k];
          dopt.trafo[4 * k]      = 1.;
        }
      dopt.device    = opt.device;
      dopt.basis     = opt.basis;
      dopt.comm_rank = opt.comm_rank;
      dopt.comm_size = opt.comm_size > 1 ? opt.comm_size : 1;
      // one process per GPU: the DG operator shares the z-slabs and the NCCL communicator of the FE_Q hierarchy
      op = new DGOperator(dopt, dopt.comm_size > 1 ? mgb_internal_comm_borrow(fe) : nullptr);
      op->use_stream(mgb_internal_stream(fe));
      op->preconditioner = opt.preconditioner;
      op->check_same_mesh(Lfine);
      N = op->n_dofs;
      for (double **v : {&solution, &rhs, &g, &d, &h})
        {
          MGB_DG_CHECK(cudaMalloc(v, N * sizeof(double)));
          MGB_DG_CHECK(cudaMemsetAsync(*v, 0, N * sizeof(double), op->stream));
        }
      for (float **v : {&defect_dg, &solution_update_dg})
        {
          MGB_DG_CHECK(cudaMalloc(v, N * sizeof(float)));
          MGB_DG_CHECK(cudaMemsetAsync(*v, 0, N * sizeof(float), op->stream));
        }
      const int           n = op->n;
      std::vector<double> hb(op->basis.S);
      hb.insert(hb.end(), op->basis.xq.begin(), op->basis.xq.end());
      hb.insert(hb.end(), op->basis.wq.begin(), op->basis.wq.end());
      MGB_DG_CHECK(cudaMalloc(&basis_dev, hb.size() * sizeof(double)));
      MGB_DG_CHECK(cudaMemcpyAsync(basis_dev, hb.data(), hb.size() * sizeof(double), cudaMemcpyHostToDevice, op->stream));
      MGB_DG_CHECK(cudaStreamSynchronize(op->stream));
      (void)n;
      cell_integrals(0, FN_CUBE_RHS, nullptr, rhs); // right-hand side: (f, phi_i) only (multigrid_solver_dg.h:243-265)
      smooth_dg.smoothing_range     = 20.;
      smooth_dg.degree              = (unsigned int)opt.degree_pre;
      smooth_dg.eig_cg_n_iterations = 15;
      smooth_dg.initialize(*op);
    }

    void
    cleanup()
    {
      for (double *v : {solution, rhs, g, d, h, basis_dev})
        if (v)
          cudaFree(v);
      for (float *v : {defect_dg, solution_update_dg})
        if (v)
          cudaFree(v);
      solution = rhs = g = d = h = basis_dev = nullptr;
      defect_dg = solution_update_dg = nullptr;
      delete op;
      op = nullptr;
      if (fe)
        mgb_solver_destroy(fe);
      fe = nullptr;
    }

    void
    cell_integrals(const int mode, const int fn, const double *u, double *out)
    {
      DGCellGeom G;
      G.n = op->n;
      double det = 1;
      for (int k = 0; k < 3; ++k)
        {
          G.nc[k]    = op->nc[k];
          G.lower[k] = opt.lower[k];
          G.h[k]     = (opt.upper[k] - opt.lower[k]) / op->nc[k];
          det *= G.h[k];
        }
      for (int k = 0; k < 9; ++k)
        G.T[k] = (k % 4 == 0) ? 1. : 0.;
      G.jxw             = det;
      G.cell_begin      = (size_t)op->nc[0] * op->nc[1] * op->zg;
      G.n_cells         = (size_t)op->nc[0] * op->nc[1] * op->lz;
      G.cz_shift        = op->cz0 - op->zg;
      const int    n    = op->n;
      const size_t smem = (size_t)(n * n * n + n * n) * sizeof(double);
      const unsigned blocks = (unsigned)std::min<size_t>(G.n_cells, 148 * 32);
      k_dg_cell_integrals<<<blocks, 128, smem, op->stream>>>(G, mode, fn, basis_dev, basis_dev + n * n, basis_dev + n * n + n,
                                                              u, out);
      ++op->launches;
      MGB_DG_CHECK(cudaGetLastError());
    }

    // multigrid_solver_dg.h:328-366
    double
    compute_l2_error()
    {
      MGB_DG_CHECK(cudaMemsetAsync(op->scal_dev, 0, 4 * sizeof(double), op->stream));
      cell_integrals(1, FN_CUBE_SOLUTION, solution, op->scal_dev);
      if (op->n_ranks > 1)
        mgb_internal_comm_allreduce(op->comm, op->stream, op->scal_dev, 2);
      MGB_DG_CHECK(cudaMemcpyAsync(op->scal_host, op->scal_dev, 4 * sizeof(double), cudaMemcpyDeviceToHost, op->stream));
      MGB_DG_CHECK(cudaStreamSynchronize(op->stream));
      return std::sqrt(op->scal_host[0] / op->scal_host[1]);
    }

    // multigrid_solver_dg.h:605-633 with my_n_cycles = 1: solution_update_dg = M^-1 defect_dg
    void
    dg_v_cycle()
    {
      void *fe_defect = nullptr, *fe_update = nullptr;
      smooth_dg.vmult(solution_update_dg, defect_dg);
      fe_check(mgb_solver_vector(fe, 3, maxlevel, &fe_defect));
      op->vmult_residual_and_restrict_to_cg<float>(Lfine, defect_dg, solution_update_dg, (float *)fe_defect);
      fe_check(mgb_solver_vcycle_inplace(fe));
      fe_check(mgb_solver_vector(fe, 5, maxlevel, &fe_update)); // the smoother swaps vectors: ask after the cycle
      if (op->n_ranks > 1)
        fe_check(mgb_vec_update_ghost_values(fe, maxlevel, MGB_F32, fe_update));
      op->prolongate_add_cg_to_dg<float>(Lfine, solution_update_dg, (const float *)fe_update);
      smooth_dg.step(solution_update_dg, defect_dg);
    }

    // preconditioner interface (multigrid_solver_dg.h:429-441): dst = M^-1 src in double, the cycle in float
    void
    vmult(double *dst, const double *src)
    {
      float *def = defect_dg;
      op->foreach (N, [=] __device__(size_t i) { def[i] = (float)src[i]; });
      dg_v_cycle();
      const float *upd = solution_update_dg;
      op->foreach (N, [=] __device__(size_t i) { dst[i] = (double)upd[i]; });
    }

    // multigrid_solver_dg.h:445-527
    void
    vmult_with_residual_update(double *residual, double *update, const double factor, double results[2])
    {
      dg_vmult_with_residual_update(
        *op, N, residual, update, factor, defect_dg,
        [&]() -> const float * {
          dg_v_cycle();
          return solution_update_dg;
        },
        results);
    }

    // multigrid_solver_dg.h:410-424: deal.II SolverCG with ReductionControl(100, 1e-16, tolerance), zero start
    void
    solve_cg(const double tolerance, double *iterations_estimate, double *reduction_rate, unsigned int *n_iterations)
    {
      double             res0 = 0, res = 0;
      const unsigned int it   = dg_solve_cg(
        *op, N, solution, rhs, g, d, h, defect_dg,
        [&]() -> const float * {
          dg_v_cycle();
          return solution_update_dg;
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
  };
} // namespace mgb
