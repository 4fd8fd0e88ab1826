// multigrid::MultigridSolver<dim, fe_degree, Number, Number2> -- same name, template parameters and method names
// as common/multigrid_solver.h:96-782 of the reference; bodies are calls into the C ABI (include/mgb200.h).
// DoFHandler / Function objects of the reference's constructor (multigrid_solver.h:100-106) are replaced by a
// MeshDescription (box generator + refinements, poisson_cube/program.cc:498-570) and a problem id whose analytic
// solution, right-hand side and coefficient are evaluated on the device (poisson_cube/program.cc:79-144).
#ifndef MULTIGRID_MULTIGRID_SOLVER_H
#define MULTIGRID_MULTIGRID_SOLVER_H

#include <cstdio>
#include <array>
#include <utility>
#include <vector>

#include "laplace_operator.h"

namespace multigrid
{
  struct MeshDescription
  {
    int    n_subdivisions[3] = {1, 1, 1}; // GridGenerator::subdivided_hyper_cube / _rectangle
    int    n_refinements     = 0;         // Triangulation::refine_global
    double lower[3] = {-0.9, -0.9, -0.9}, upper[3] = {1., 1., 1.};
    int    problem = MGB_PROBLEM_CUBE_SINE;
    // geometry of the structured mesh: MGB_GEOMETRY_BOX, _SHELL_SECTOR (params r0, r1, opening half-angle: one coarse
    // cell of GridGenerator::hyper_shell(.., 6), poisson_shell/program.cc:426) or _AFFINE (params: 3 x 3 matrix)
    int    geometry = MGB_GEOMETRY_BOX;
    double geometry_params[12] = {0.5, 1.0, 0.78539816339744830962, 0, 0, 0, 0, 0, 0, 0, 0, 0};
    // one process per GPU (MPI_COMM_WORLD of the reference): rank, size and the id of mgb_comm_unique_id()
    int           comm_rank = 0, comm_size = 1;
    unsigned char comm_id[128] = {};
    int           device = 0;
  };

  template <int dim, int fe_degree, typename Number = float, typename Number2 = double>
  class MultigridSolver
  {
    static_assert(dim == 3, "the B200 path implements dim = 3");
    static_assert(std::is_same<Number, float>::value && std::is_same<Number2, double>::value,
                  "mixed precision as the reference: float V-cycle, double outer solver (poisson_cube/program.cc:76-77)");

  public:
    typedef DeviceVector<Number2> VectorType;

    // multigrid_solver.h:100-106
    MultigridSolver(const MeshDescription &mesh, const unsigned int degree_pre, const unsigned int degree_post,
                    const unsigned int n_cycles = 1)
    {
      mgb_options o;
      mgb_default_options(&o);
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
      o.n_cycles      = (int)n_cycles;
      o.problem       = mesh.problem;
      o.geometry      = mesh.geometry;
      for (int i = 0; i < 12; ++i)
        o.geometry_params[i] = mesh.geometry_params[i];
      o.device        = mesh.device;
      o.comm_rank     = mesh.comm_rank;
      o.comm_size     = mesh.comm_size;
      for (int i = 0; i < 128; ++i)
        o.comm_id[i] = mesh.comm_id[i];
      check(mgb_solver_create(&o, &s_));
    }
    MultigridSolver(const MultigridSolver &) = delete;
    MultigridSolver &operator=(const MultigridSolver &) = delete;
    ~MultigridSolver()
    {
      if (s_)
        mgb_solver_destroy(s_);
    }

    mgb_solver   handle() const { return s_; }
    unsigned int n_levels() const { return (unsigned int)mgb_solver_n_levels(s_); }

    // ---- the Function<dim> arguments of the reference's constructor (boundary_values, right_hand_side, coefficient,
    // multigrid_solver.h:100-106) as callables double(double x, double y, double z): evaluated here at the library's
    // quadrature points (coefficient, right-hand side, exact solution for compute_l2_error) and support points (boundary
    // values) of every level and handed over as samples.  Any of them may be null (keep what the problem id set).
    template <typename Coefficient, typename RightHandSide, typename BoundaryValues>
    void
    set_problem(const Coefficient *coefficient, const RightHandSide *right_hand_side, const BoundaryValues *boundary_values,
                const bool boundary_values_are_exact_solution = true)
    {
      const int           nl = (int)n_levels();
      std::vector<double> pts, val, bc;
      for (int l = 0; l < nl && coefficient; ++l)
        {
          sample(l, true, *coefficient, pts, val);
          check(mgb_solver_set_coefficient_values(s_, l, val.data()));
        }
      if (coefficient)
        check(mgb_solver_reinit_smoothers(s_));
      for (int l = 0; l < nl; ++l)
        {
          if (right_hand_side)
            {
              sample(l, true, *right_hand_side, pts, val);
              if (boundary_values)
                sample(l, false, *boundary_values, pts, bc);
              check(mgb_solver_set_rhs_values(s_, l, val.data(), boundary_values ? bc.data() : nullptr));
            }
          if (boundary_values && boundary_values_are_exact_solution)
            {
              sample(l, true, *boundary_values, pts, val);
              check(mgb_solver_set_exact_solution_values(s_, l, val.data()));
            }
        }
    }
    // FEEvaluation::quadrature_point of all cells of a level, [cell][q][3] (cells lexicographic, x fastest)
    std::vector<double>
    quadrature_points(const int level) const
    {
      std::vector<double> xyz(3 * (std::size_t)mgb_solver_n_quadrature_points(s_, level));
      check(mgb_solver_quadrature_points(s_, level, xyz.data()));
      return xyz;
    }
    // support points of the DoFs of a level in the library's numbering, [dof][3]
    std::vector<double>
    support_points(const int level) const
    {
      std::vector<double> xyz(3 * (std::size_t)mgb_solver_n_dofs(s_, level));
      check(mgb_solver_support_points(s_, level, xyz.data()));
      return xyz;
    }

    // level operators: matrix[level] (float) and matrix_dp[level] (double), multigrid_solver.h:740-745
    LaplaceOperator<dim, fe_degree, Number>
    get_matrix(const int level = -1) const
    {
      LaplaceOperator<dim, fe_degree, Number> op;
      op.initialize(s_, level);
      return op;
    }
    LaplaceOperator<dim, fe_degree, Number2>
    get_matrix_dp(const int level = -1) const
    {
      LaplaceOperator<dim, fe_degree, Number2> op;
      op.initialize(s_, level);
      return op;
    }

    // multigrid_solver.h:387-476: full multigrid cycle, returns the residual reduction rate on the finest level
    double
    solve(const bool do_analyze)
    {
      double rate = 0;
      check(mgb_solver_solve(s_, do_analyze ? 1 : 0, &rate, nullptr));
      return rate;
    }
    // multigrid_solver.h:483-493: CG preconditioned by one V-cycle; (iterations, reduction per iteration)
    std::pair<unsigned int, double>
    solve_cg()
    {
      unsigned int its  = 0;
      double       rate = 0;
      check(mgb_solver_solve_cg(s_, &its, &rate));
      return std::make_pair(its, rate);
    }
    // multigrid_solver.h:498-510: one V-cycle as preconditioner
    void vmult(VectorType &dst, const VectorType &src) const { check(mgb_solver_vcycle(s_, dst.begin(), src.begin())); }
    // multigrid_solver.h:516-619: preconditioner entry of the merged CG
    std::array<Number2, 2>
    vmult_with_residual_update(VectorType &residual, VectorType &update, const Number2 factor) const
    {
      double sums[2] = {0., 0.};
      check(mgb_solver_vmult_with_residual_update(s_, residual.begin(), update.begin(), (double)factor, sums));
      return {{(Number2)sums[0], (Number2)sums[1]}};
    }
    // multigrid_solver.h:624-637
    void do_matvec() { check(mgb_solver_do_matvec(s_)); }
    void do_matvec_smoother() { check(mgb_solver_do_matvec_smoother(s_)); }
    // multigrid_solver.h:298-343 (volume-normalised L2 error against the analytic solution)
    double
    compute_l2_error(const int level = -1) const
    {
      double e = 0;
      check(mgb_solver_l2_error(s_, level, &e));
      return e;
    }
    // multigrid_solver.h:348-371
    void
    print_wall_times()
    {
      const unsigned int  nl = n_levels();
      std::vector<double> t(6 * nl);
      check(mgb_solver_timings(s_, t.data(), 1));
      std::printf("level  smoother    mg_mv       mg_vec      restrict    prolongate  inhomBC\n");
      for (unsigned int l = 1; l < nl; ++l)
        std::printf("L%-2u    %-12.4g%-12.4g%-12.4g%-12.4g%-12.4g%-12.4g\n", l, t[6 * l + 5], t[6 * l + 0], t[6 * l + 4],
                    t[6 * l + 1], t[6 * l + 2], t[6 * l + 3]);
    }
    // multigrid_solver.h:376-382 (host copy in the library's numbering)
    std::vector<double>
    get_solution() const
    {
      std::vector<double> sol((std::size_t)mgb_solver_n_dofs(s_, -1));
      check(mgb_solver_get_solution(s_, sol.data()));
      return sol;
    }
    void synchronize() const { check(mgb_solver_sync(s_)); }

  private:
    template <typename F>
    void
    sample(const int level, const bool at_quadrature_points, const F &f, std::vector<double> &pts, std::vector<double> &val) const
    {
      pts = at_quadrature_points ? quadrature_points(level) : support_points(level);
      val.resize(pts.size() / 3);
      for (std::size_t i = 0; i < val.size(); ++i)
        val[i] = f(pts[3 * i], pts[3 * i + 1], pts[3 * i + 2]);
    }

    mgb_solver s_ = nullptr;
  };
} // namespace multigrid
#endif
