// multigrid::MultigridSolverDG<dim, fe_degree, Number, Number2> -- class, template parameters and method names of
// common/multigrid_solver_dg.h:44-735 of the reference; bodies are calls into the C ABI (include/mgb200.h,
// mgb_dg_solver_*).  The constructor's DoFHandler / Function arguments (multigrid_solver_dg.h:59-65) become the
// MeshDescription of multigrid_solver.h; the problem data are those of poisson_dg/program.cc:70-135.
#ifndef MULTIGRID_MULTIGRID_SOLVER_DG_H
#define MULTIGRID_MULTIGRID_SOLVER_DG_H

#include <array>
#include <utility>

#include "laplace_operator_dg.h"
#include "multigrid_solver.h"

namespace multigrid
{
  template <int dim, int fe_degree, typename Number = float, typename Number2 = double>
  class MultigridSolverDG
  {
    static_assert(dim == 3, "the B200 path implements dim = 3");
    static_assert(std::is_same<Number, float>::value && std::is_same<Number2, double>::value,
                  "mixed precision as the reference: float cycle, double outer solver (poisson_dg/program.cc:60-61)");

  public:
    // multigrid_solver_dg.h:59-65; n_cycles is accepted and, as in the reference (dg_v_cycle(1), :436), not used
    // basis: MGB_DG_BASIS_* of the DG space; the default is the reference's: its solver runs on FE_DGQHermite
    // (LaplaceOperatorCompactCombine<dim, fe_degree, Number> with the default type 0, multigrid_solver_dg.h:700-701;
    // poisson_dg/program.cc:174), so right-hand side and solution are coefficients of that basis
    MultigridSolverDG(const MeshDescription &mesh, const unsigned int degree_pre, const unsigned int degree_post,
                      const unsigned int /*n_cycles*/ = 1, const int basis = MGB_DG_BASIS_HERMITE)
    {
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
      o.basis         = basis;
      o.comm_rank     = mesh.comm_rank; // one process per GPU: z-slabs shared by the DG operator and the FE_Q hierarchy
      o.comm_size     = mesh.comm_size;
      for (int i = 0; i < 128; ++i)
        o.comm_id[i] = mesh.comm_id[i];
      check(mgb_dg_solver_create(&o, &s_));
    }
    MultigridSolverDG(const MultigridSolverDG &) = delete;
    MultigridSolverDG &operator=(const MultigridSolverDG &) = delete;
    ~MultigridSolverDG()
    {
      if (s_)
        mgb_dg_solver_destroy(s_);
    }

    mgb_dg_solver handle() const { return s_; }
    std::size_t   m() const { return (std::size_t)mgb_dg_solver_n_dofs(s_); }

    // :410-424
    std::pair<double, double>
    solve_cg(const double tolerance)
    {
      double est = 0, rate = 0;
      check(mgb_dg_solver_solve_cg(s_, tolerance, &est, &rate, &last_step_));
      return std::make_pair(est, rate);
    }
    unsigned int last_step() const { return last_step_; } // solver_control.last_step()
    // :429-441
    void vmult(DGVector<Number2> &dst, const DGVector<Number2> &src) const { check(mgb_dg_solver_vmult(s_, dst.begin(), src.begin())); }
    // :445-527: the preconditioner call of a CG solver with merged vector operations
    std::array<Number2, 2>
    vmult_with_residual_update(DGVector<Number2> &residual, DGVector<Number2> &update, const Number2 factor) const
    {
      std::array<Number2, 2> results;
      check(mgb_dg_solver_vmult_with_residual_update(s_, residual.begin(), update.begin(), factor, results.data()));
      return results;
    }
    // :328-366
    double
    compute_l2_error() const
    {
      double e = 0;
      check(mgb_dg_solver_l2_error(s_, &e));
      return e;
    }
    // :399-403 (host copy, cell-wise contiguous DoFs, cells lexicographic)
    std::vector<double>
    get_solution() const
    {
      std::vector<double> sol(m());
      check(mgb_dg_solver_get_solution(s_, sol.data()));
      return sol;
    }
    void do_matvec() { check(mgb_dg_solver_do_matvec(s_)); }                   // :539-543
    void do_matvec_smoother() { check(mgb_dg_solver_do_matvec_smoother(s_)); } // :548-552
    void synchronize() const { check(mgb_dg_solver_sync(s_)); }

  private:
    mgb_dg_solver s_         = nullptr;
    unsigned int  last_step_ = 0;
  };
} // namespace multigrid
#endif
