// multigrid::MultigridSolverDGPlain<dim, fe_degree, Number, Number2, type> -- class, template parameters and method
// names of common/multigrid_solver_dg_plain.h:54-590 of the reference; bodies are calls into the C ABI
// (include/mgb200.h, mgb_dg_plain_solver_*).  The constructor's DoFHandler / Function arguments (:58-64) become the
// MeshDescription of multigrid_solver.h; the problem data are those of poisson_dg_plain/program.cc:70-135.
#ifndef MULTIGRID_MULTIGRID_SOLVER_DG_PLAIN_H
#define MULTIGRID_MULTIGRID_SOLVER_DG_PLAIN_H

#include <array>
#include <utility>

#include "laplace_operator_dg.h"
#include "multigrid_solver.h"

namespace multigrid
{
  template <int dim, int fe_degree, typename Number = float, typename Number2 = double, int type = 0>
  class MultigridSolverDGPlain
  {
    static_assert(dim == 3, "the B200 path implements dim = 3");
    static_assert(std::is_same<Number, float>::value && std::is_same<Number2, double>::value,
                  "mixed precision as the reference: float cycle, double outer solver (poisson_dg_plain/program.cc:73-74)");
    static_assert(type >= 0 && type <= 2, "type: 0 FE_DGQHermite, 1 FE_DGQ, 2 FE_DGQArbitraryNodes(QGauss)");

  public:
    // :58-64; n_cycles is accepted and, as in the reference (v_cycle(maxlevel, 1), :329), not used
    MultigridSolverDGPlain(const MeshDescription &mesh, const unsigned int degree_pre, const unsigned int degree_post,
                           const unsigned int /*n_cycles*/ = 1)
    {
      if (mesh.comm_size > 1)
        throw std::runtime_error("MultigridSolverDGPlain runs on one GPU");
      mgb_dg_solver_options o;
      mgb_dg_solver_default_options(&o);
      o.degree = fe_degree;
      for (int d = 0; d < 3; ++d)
        {
          o.n_subdivisions[d] = mesh.n_subdivisions[d];
          o.lower[d]          = mesh.lower[d];
          o.upper[d]          = mesh.upper[d];
        }
      o.n_refinements = mesh.n_refinements;
      o.degree_pre    = (int)degree_pre;
      o.degree_post   = (int)degree_post;
      o.device        = mesh.device;
      o.basis         = MGB_DG_BASIS_OF_TYPE(type);
      check(mgb_dg_plain_solver_create(&o, &s_));
    }
    MultigridSolverDGPlain(const MultigridSolverDGPlain &) = delete;
    MultigridSolverDGPlain &operator=(const MultigridSolverDGPlain &) = delete;
    ~MultigridSolverDGPlain()
    {
      if (s_)
        mgb_dg_plain_solver_destroy(s_);
    }

    mgb_dg_plain_solver handle() const { return s_; }
    std::size_t         m() const { return (std::size_t)mgb_dg_plain_solver_n_dofs(s_); }
    unsigned int        n_levels() const { return (unsigned int)mgb_dg_plain_solver_n_levels(s_); }

    // :304-316
    std::pair<double, double>
    solve_cg(const double tolerance)
    {
      double est = 0, rate = 0;
      check(mgb_dg_plain_solver_solve_cg(s_, tolerance, &est, &rate, &last_step_));
      return std::make_pair(est, rate);
    }
    unsigned int last_step() const { return last_step_; } // solver_control.last_step()
    // :321-337
    void
    vmult(DGVector<Number2> &dst, const DGVector<Number2> &src) const
    {
      check(mgb_dg_plain_solver_vmult(s_, dst.begin(), src.begin()));
    }
    // :340-427: the preconditioner call of a CG solver with merged vector operations
    std::array<Number2, 2>
    vmult_with_residual_update(DGVector<Number2> &residual, DGVector<Number2> &update, const Number2 factor) const
    {
      std::array<Number2, 2> results;
      check(mgb_dg_plain_solver_vmult_with_residual_update(s_, residual.begin(), update.begin(), factor, results.data()));
      return results;
    }
    // :221-259
    double
    compute_l2_error() const
    {
      double e = 0;
      check(mgb_dg_plain_solver_l2_error(s_, &e));
      return e;
    }
    // host copy of `solution`, cell-wise contiguous DoFs, cells lexicographic
    std::vector<double>
    get_solution() const
    {
      std::vector<double> sol(m());
      check(mgb_dg_plain_solver_get_solution(s_, sol.data()));
      return sol;
    }
    void do_matvec() { check(mgb_dg_plain_solver_do_matvec(s_)); }                   // :432-436
    void do_matvec_smoother() { check(mgb_dg_plain_solver_do_matvec_smoother(s_)); } // :441-445
    void synchronize() const { check(mgb_dg_plain_solver_sync(s_)); }

  private:
    mgb_dg_plain_solver s_         = nullptr;
    unsigned int        last_step_ = 0;
  };
} // namespace multigrid
#endif
