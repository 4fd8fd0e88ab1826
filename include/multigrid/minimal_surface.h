// multigrid::MinimalSurfaceProblem<dim, fe_degree> -- the LaplaceProblem class of minimal_surface/program.cc:200-660 of the
// reference (its MinimalSurfaceOperator, :102-198, is inside the library): Newton's method with line search for
// -div(grad u / sqrt(1 + |grad u|^2)) = 0 with Dirichlet data, the linear systems solved by CG with a multigrid V-cycle on
// the linearised operator whose six-entry coefficient is re-evaluated from the iterate on every level in every step.
// Bodies are calls into the C ABI (include/mgb200.h, mgb_solver_minimal_surface_*).  The reference runs it in 2-D on
// hyper_ball; here it runs in 3-D on the meshes of MeshDescription (box, sheared box, shell sector).
#ifndef MULTIGRID_MINIMAL_SURFACE_H
#define MULTIGRID_MINIMAL_SURFACE_H

#include <vector>

#include "multigrid_solver.h"

namespace multigrid
{
  template <int dim, int fe_degree>
  class MinimalSurfaceProblem
  {
    static_assert(dim == 3, "the B200 path implements dim = 3");

  public:
    struct StepInfo
    {
      double       initial_residual_norm, final_residual_norm, step_length;
      unsigned int linear_iterations, line_search_steps;
    };

    // the smoother of program.cc:458-486: Chebyshev of degree 2 on the levels, the coarse solver with tolerance 1e-3
    explicit MinimalSurfaceProblem(const MeshDescription &mesh)
      : solver_(mesh, 2, 2, 1)
    {}

    MultigridSolver<dim, fe_degree, float, double> &solver() { return solver_; }

    // interpolate_boundary_values (:396-408): Dirichlet data g(x, y, z); the iterate is g on the boundary, zero inside
    template <typename BoundaryValues>
    void
    interpolate_boundary_values(const BoundaryValues &g)
    {
      const int                 level = (int)solver_.n_levels() - 1;
      const std::vector<double> xyz   = solver_.support_points(level);
      std::vector<double>       values(xyz.size() / 3);
      for (std::size_t i = 0; i < values.size(); ++i)
        values[i] = g(xyz[3 * i], xyz[3 * i + 1], xyz[3 * i + 2]);
      check(mgb_solver_minimal_surface_set_boundary_values(solver_.handle(), values.data()));
    }
    // the analytic solution of MeshDescription::problem as Dirichlet data
    void interpolate_boundary_values() { check(mgb_solver_minimal_surface_set_boundary_values(solver_.handle(), nullptr)); }

    // solve(first_time) (:412-572)
    StepInfo
    solve(const bool first_time)
    {
      double info[5];
      check(mgb_solver_minimal_surface_step(solver_.handle(), first_time ? 1 : 0, info));
      StepInfo s;
      s.initial_residual_norm = info[0];
      s.final_residual_norm   = info[1];
      s.linear_iterations     = (unsigned int)info[2];
      s.line_search_steps     = (unsigned int)info[3];
      s.step_length           = info[4];
      return s;
    }

    // the inner loop of run() (:652-659): Newton steps until the residual norm is below 1e-12 (100 at most)
    std::vector<StepInfo>
    run(const unsigned int n_inner_iterations = 100, const double tolerance = 1e-12)
    {
      std::vector<StepInfo> steps;
      for (unsigned int it = 0; it < n_inner_iterations; ++it)
        {
          steps.push_back(solve(it == 0));
          if (steps.back().final_residual_norm < tolerance)
            break;
        }
      return steps;
    }

    std::vector<double> get_solution() const { return solver_.get_solution(); }

  private:
    MultigridSolver<dim, fe_degree, float, double> solver_;
  };
} // namespace multigrid
#endif
