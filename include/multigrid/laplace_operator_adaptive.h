// multigrid::LaplaceOperator on adaptively refined meshes with hanging nodes -- the operator that
// poisson_l/program.cc:232-260 builds from a MatrixFree object with AffineConstraints (hanging nodes + zero boundary
// values), over the C ABI (include/mgb200.h, mgb_adaptive_*).  The mesh (Triangulation + refine flags of
// poisson_l/program.cc:465-541) is owned by the operator: a forest of root cells, refined cell by cell and kept 2:1
// balanced.  The vector holds the unconstrained and the Dirichlet DoFs; hanging-node values are not stored.
#ifndef MULTIGRID_LAPLACE_OPERATOR_ADAPTIVE_H
#define MULTIGRID_LAPLACE_OPERATOR_ADAPTIVE_H

#include "laplace_operator_dg.h" // DGVector: plain device array

namespace multigrid
{
  template <int dim, int fe_degree, typename number>
  class AdaptiveLaplaceOperator
  {
    static_assert(dim == 3, "the B200 path implements dim = 3");

  public:
    typedef DGVector<number> VectorType;
    AdaptiveLaplaceOperator() = default;
    AdaptiveLaplaceOperator(const AdaptiveLaplaceOperator &) = delete;
    ~AdaptiveLaplaceOperator()
    {
      if (h_)
        mgb_adaptive_destroy(h_);
    }
    // GridGenerator::subdivided_hyper_rectangle / hyper_L (root_mask[z][y][x] == 0 removes a root cell) + refine_global
    void
    reinit(const int n_roots[3], const unsigned char *root_mask, const double lower[3], const double upper[3],
           const int n_global_refinements, const int device = 0)
    {
      if (h_)
        mgb_adaptive_destroy(h_);
      h_ = nullptr;
      check(mgb_adaptive_create(fe_degree, n_roots, root_mask, lower, upper, n_global_refinements, device, &h_));
    }
    std::size_t n_active_cells() const { return (std::size_t)mgb_adaptive_n_cells(h_); }
    // level and position (in cells of that level) of the active cells, in the order the refine flags refer to
    void
    get_cells(std::vector<int> &level, std::vector<int> &position) const
    {
      level.resize(n_active_cells());
      position.resize(3 * n_active_cells());
      check(mgb_adaptive_cells(h_, level.data(), position.data()));
    }
    // cell->set_refine_flag() + execute_coarsening_and_refinement()   poisson_l/program.cc:508-541
    void execute_refinement(const std::vector<unsigned char> &flags) { check(mgb_adaptive_refine(h_, flags.data())); }
    std::size_t
    m() const
    {
      const int64_t n = mgb_adaptive_n_dofs(h_);
      if (n < 0)
        throw std::runtime_error(std::string("mgb200: ") + mgb_last_error());
      return (std::size_t)n;
    }
    std::size_t n() const { return m(); }
    // 1 for the DoFs on the Dirichlet boundary (identity rows of the operator)
    std::vector<unsigned char>
    dirichlet_mask() const
    {
      std::vector<unsigned char> mask(m());
      check(mgb_adaptive_dirichlet(h_, mask.data()));
      return mask;
    }
    void        initialize_dof_vector(VectorType &v) const { v.reinit(m()); }
    // laplace_operator.h:575-601 with the constraints applied in the kernel
    void vmult(VectorType &dst, const VectorType &src) const { check(mgb_adaptive_vmult(h_, prec, dst.begin(), src.begin())); }
    // compute_diagonal + get_matrix_diagonal_inverse   laplace_operator.h:747-800
    void compute_diagonal(VectorType &inverse_diagonal) const { check(mgb_adaptive_inverse_diagonal(h_, prec, inverse_diagonal.begin())); }
    void synchronize() const { check(mgb_adaptive_sync(h_)); }

    // ---- the loop around the solver, LaplaceProblem::run of poisson_l/program.cc:497-602; the functions of the program
    //      are selected by id: 0 = zero, 1 = Functions::LSingularityFunction (the exact solution on the extruded L)
    // interpolate_boundary_values (:321-333): u = g on the Dirichlet DoFs, zero elsewhere
    void
    interpolate_boundary_values(DGVector<double> &u, const int function_id) const
    {
      check(mgb_adaptive_interpolate_boundary_values(h_, function_id, u.begin()));
    }
    // compute_residual(dst, src, ConstantFunction(f)) (laplace_operator.h:806-845): src carries the boundary values
    void
    compute_residual(DGVector<double> &dst, const DGVector<double> &src, const double f) const
    {
      check(mgb_adaptive_compute_residual(h_, dst.begin(), src.begin(), f));
    }
    // KellyErrorEstimator<dim>::estimate(dof_handler, QGauss<dim-1>(p+1), {}, solution, estimated_error_per_cell) :524-531
    std::vector<float>
    estimate_kelly(const DGVector<double> &solution) const
    {
      std::vector<float> eta(n_active_cells());
      check(mgb_adaptive_estimate_kelly(h_, solution.begin(), eta.data()));
      return eta;
    }
    // GridRefinement::refine_and_coarsen_fixed_number + execute_coarsening_and_refinement :532-541
    void
    refine_and_coarsen_fixed_number(const std::vector<float> &indicators, const double top_fraction, const double bottom_fraction)
    {
      if (indicators.size() != n_active_cells())
        throw std::runtime_error("one indicator per active cell");
      check(mgb_adaptive_refine_and_coarsen_fixed_number(h_, indicators.data(), top_fraction, bottom_fraction));
    }
    void
    execute_coarsening_and_refinement(const std::vector<unsigned char> &refine_flags, const std::vector<unsigned char> &coarsen_flags)
    {
      check(mgb_adaptive_refine_and_coarsen(h_, refine_flags.data(), coarsen_flags.empty() ? nullptr : coarsen_flags.data()));
    }
    // VectorTools::integrate_difference (:556-575): norm 0 = L2_norm, 1 = H1_seminorm, n_q_points per direction
    double
    integrate_difference(const DGVector<double> &solution, const int function_id, const int norm, const int n_q_points) const
    {
      double r = 0;
      check(mgb_adaptive_integrate_difference(h_, solution.begin(), function_id, norm, n_q_points, &r, nullptr));
      return r;
    }
    // one cycle on the current mesh (:545-575): {CG iterations, rate, L2 error (QGauss(p+2)), H1-seminorm error (QGauss(p+1))};
    // the solution stays on the device (solution_device(), valid until the mesh changes)
    struct CycleResult
    {
      unsigned int iterations;
      double       rate, l2_error, h1_error;
      std::size_t  n_cells, n_dofs;
    };
    CycleResult
    solve_cycle(const int function_id, const double f, const unsigned int n_smoother)
    {
      double stats[4];
      check(mgb_adaptive_solve_cycle(h_, function_id, f, (int)n_smoother, stats, &solution_));
      CycleResult r;
      r.iterations = (unsigned int)stats[0];
      r.rate       = stats[1];
      r.l2_error   = stats[2];
      r.h1_error   = stats[3];
      r.n_cells    = n_active_cells();
      r.n_dofs     = m();
      return r;
    }
    const double *solution_device() const { return solution_; }
    // LaplaceProblem::run (:497-602) from the current mesh: the convergence table of the program
    std::vector<CycleResult>
    run(const unsigned int n_cycles, const std::size_t max_size, const unsigned int n_smoother, const bool use_hyper_l)
    {
      std::vector<CycleResult> table;
      for (unsigned int cycle = 0; cycle < n_cycles; ++cycle)
        {
          if (cycle > 0)
            {
              std::vector<float> eta(n_active_cells());
              check(mgb_adaptive_estimate_kelly(h_, solution_, eta.data()));
              refine_and_coarsen_fixed_number(eta, use_hyper_l ? 0.15 : 0.3, 0.03);
            }
          table.push_back(solve_cycle(use_hyper_l ? 0 : 1, use_hyper_l ? 1. : 0., n_smoother));
          if (table.back().n_dofs > max_size)
            break;
        }
      return table;
    }

    // ---- the solver of poisson_l/program.cc:337-420 on this mesh (double vectors only): level operators with constrained
    // refinement edges, MGInterfaceOperator edge matrices, MGTransferMatrixFree, Chebyshev smoothers of degree n_smoother
    void setup_multigrid(const unsigned int n_smoother) { check(mgb_adaptive_mg_setup(h_, (int)n_smoother)); }
    unsigned int n_levels() const { return (unsigned int)mgb_adaptive_mg_n_levels(h_); }
    // PreconditionMG::vmult: one local-smoothing V-cycle in float
    void
    precondition(DGVector<double> &dst, const DGVector<double> &src) const
    {
      check(mgb_adaptive_mg_vmult(h_, dst.begin(), src.begin()));
    }
    // the level operator and the edge matrices of level `level` of the local-smoothing hierarchy on float level vectors
    // (mg_matrices[level].vmult, vmult_interface_down, vmult_interface_up; poisson_l/program.cc:283-291,374-383)
    std::size_t
    level_size(const unsigned int level) const
    {
      double info[12];
      check(mgb_adaptive_mg_level_info(h_, (int)level, info));
      return (std::size_t)info[0];
    }
    void level_vmult(const unsigned int level, DGVector<float> &dst, const DGVector<float> &src) const
    {
      check(mgb_adaptive_mg_level_vmult(h_, (int)level, 0, dst.begin(), src.begin()));
    }
    void vmult_interface_down(const unsigned int level, DGVector<float> &dst, const DGVector<float> &src) const
    {
      check(mgb_adaptive_mg_level_vmult(h_, (int)level, 1, dst.begin(), src.begin()));
    }
    void vmult_interface_up(const unsigned int level, DGVector<float> &dst, const DGVector<float> &src) const
    {
      check(mgb_adaptive_mg_level_vmult(h_, (int)level, 2, dst.begin(), src.begin()));
    }
    // SolverCG with ReductionControl(m(), 1e-13, rel_tol) from a zero start vector; (iterations, reduction per iteration)
    std::pair<unsigned int, double>
    solve_cg(DGVector<double> &x, const DGVector<double> &rhs, const double rel_tol = 1e-9) const
    {
      unsigned int its  = 0;
      double       rate = 0;
      check(mgb_adaptive_solve_cg(h_, x.begin(), rhs.begin(), rel_tol, &its, &rate, nullptr, nullptr));
      return std::make_pair(its, rate);
    }

  private:
    static constexpr int prec = precision_of<number>::value;
    mgb_adaptive         h_   = nullptr;
    double              *solution_ = nullptr; // owned by the library handle
  };
} // namespace multigrid
#endif
