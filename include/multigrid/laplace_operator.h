// multigrid::LaplaceOperator<dim, fe_degree, number> -- same name, template parameters, method names and
// argument order as common/laplace_operator.h:56-164 of the reference; every body is one call into the C ABI
// (include/mgb200.h).  The operator is a view of `matrix[level]` (float) or `matrix_dp[level]` (double) of a
// multigrid::MultigridSolver, which owns the device data (multigrid_solver.h:740-745).
//
// What cannot cross to the device: the two std::function<void(unsigned,unsigned)> range callbacks of
// vmult(dst, src, before, after) (laplace_operator.h:96-101).  Every call site of the reference passes one of a
// few vector updates (laplace_operator.h:617-631, 655-710; PreconditionChebyshev in deal.II), so they become an
// enum of fused epilogues with scalar parameters.
#ifndef MULTIGRID_LAPLACE_OPERATOR_H
#define MULTIGRID_LAPLACE_OPERATOR_H

#include <array>

#include "device_vector.h"

namespace multigrid
{
  template <int dim, int fe_degree, typename number>
  class LaplaceOperator
  {
    static_assert(dim == 3, "the B200 path implements dim = 3");
    static_assert(fe_degree >= 1 && fe_degree <= 8, "degrees 1..8 are compiled");

  public:
    typedef number               value_type;
    typedef DeviceVector<number> VectorType;

    LaplaceOperator() = default;
    // replaces initialize(matrix_free, constraints / mg_constrained_dofs, level)   laplace_operator.h:65-74
    void
    initialize(mgb_solver solver, const int level)
    {
      s_     = solver;
      level_ = level < 0 ? mgb_solver_n_levels(solver) + level : level;
    }
    void clear() { s_ = nullptr; }

    std::size_t
    m() const
    {
      int64_t info[12];
      check(mgb_solver_partition(s_, level_, info));
      return (std::size_t)info[1];
    }
    std::size_t n() const { return m(); }
    void initialize_dof_vector(VectorType &v) const { v.reinit(s_, level_); }

    // laplace_operator.h:79-81 (575-601)
    void vmult(VectorType &dst, const VectorType &src) const { check(mgb_op_vmult(s_, level_, prec, dst.begin(), src.begin())); }
    // MatrixFreeOperators::Base
    void vmult_add(VectorType &dst, const VectorType &src) const { check(mgb_op_vmult_add(s_, level_, prec, dst.begin(), src.begin())); }
    void Tvmult(VectorType &dst, const VectorType &src) const { check(mgb_op_Tvmult(s_, level_, prec, dst.begin(), src.begin())); }
    void Tvmult_add(VectorType &dst, const VectorType &src) const { vmult_add(dst, src); }
    // laplace_operator.h:83-86 (607-634)
    void
    vmult_residual(const VectorType &rhs, const VectorType &lhs, VectorType &residual) const
    {
      check(mgb_op_vmult_residual(s_, level_, prec, rhs.begin(), lhs.begin(), residual.begin()));
    }
    // laplace_operator.h:88-94 (640-719)
    std::array<number, 4>
    vmult_with_cg_update(const number alpha, const number beta, const VectorType &r, VectorType &q, VectorType &p,
                         VectorType &x) const
    {
      double sums[4];
      check(mgb_op_vmult_with_cg_update(s_, level_, prec, alpha, beta, r.begin(), q.begin(), p.begin(), x.begin(), sums));
      return {{(number)sums[0], (number)sums[1], (number)sums[2], (number)sums[3]}};
    }
    // laplace_operator.h:96-101 (725-741) with the Chebyshev epilogue that PreconditionChebyshev passes:
    // one smoother step x <- x + ... fused with the operator apply (float level operators)
    void
    vmult_chebyshev_step(VectorType &x, const VectorType &b) const
    {
      static_assert(std::is_same<number, float>::value, "the smoother runs on the float level operators");
      check(mgb_smoother_step(s_, level_, x.begin(), b.begin()));
    }
    // laplace_operator.h:111-112 (747-800): compute_diagonal() + get_matrix_diagonal_inverse()
    void
    compute_diagonal(VectorType &inverse_diagonal) const
    {
      if (inverse_diagonal.begin() == nullptr)
        inverse_diagonal.reinit(s_, level_);
      check(mgb_op_inverse_diagonal(s_, level_, prec, inverse_diagonal.begin()));
    }

    // get_matrix_diagonal_inverse(): the vector of compute_diagonal(), computed on first use
    const VectorType &
    get_matrix_diagonal_inverse() const
    {
      if (inverse_diagonal_.begin() == nullptr)
        compute_diagonal(inverse_diagonal_);
      return inverse_diagonal_;
    }
    // MatrixFreeOperators::Base::el(row, col): only diagonal entries exist
    number
    el(const std::size_t row, const std::size_t col) const
    {
      if (row != col)
        throw std::runtime_error("matrix-free operator: only diagonal entries are available");
      double v = 0;
      check(mgb_op_el(s_, level_, prec, (int64_t)row, &v));
      return (number)v;
    }
    // vmult_interface_down / vmult_interface_up (MatrixFreeOperators::Base): the coupling over refinement edges; the
    // levels of a globally refined mesh have none, and deal.II returns dst = 0 there
    void vmult_interface_down(VectorType &dst, const VectorType &) const { dst = number(0); }
    void vmult_interface_up(VectorType &dst, const VectorType &) const { dst = number(0); }
    // laplace_operator.h:103-106 (806-845) for the double operator: dst = (f, phi) - K src on the rows without Dirichlet
    // condition, src carrying the boundary values; rhs_function(x, y, z) is sampled at the library's quadrature points
    template <typename RightHandSide>
    void
    compute_residual(VectorType &dst, const VectorType &src, const RightHandSide &rhs_function) const
    {
      static_assert(std::is_same<number, double>::value, "compute_residual runs on the double operator (matrix_dp)");
      const int64_t nq = mgb_solver_n_quadrature_points(s_, level_);
      if (nq < 0)
        throw std::runtime_error(std::string("mgb200: ") + mgb_last_error());
      std::vector<double> xyz(3 * (std::size_t)nq), f((std::size_t)nq);
      check(mgb_solver_quadrature_points(s_, level_, xyz.data()));
      for (std::size_t q = 0; q < f.size(); ++q)
        f[q] = rhs_function(xyz[3 * q], xyz[3 * q + 1], xyz[3 * q + 2]);
      check(mgb_op_compute_residual(s_, level_, dst.begin(), src.begin(), f.data()));
    }
    // laplace_operator.h:108-109 (359-432): the coefficient sampled at the quadrature points; sets it on the float and
    // the double operator of the level (call MultigridSolver::reinit_smoothers() after the last level is set)
    template <typename Coefficient>
    void
    evaluate_coefficient(const Coefficient &coefficient)
    {
      const int64_t nq = mgb_solver_n_quadrature_points(s_, level_);
      if (nq < 0)
        throw std::runtime_error(std::string("mgb200: ") + mgb_last_error());
      std::vector<double> xyz(3 * (std::size_t)nq), c((std::size_t)nq);
      check(mgb_solver_quadrature_points(s_, level_, xyz.data()));
      for (std::size_t q = 0; q < c.size(); ++q)
        c[q] = coefficient(xyz[3 * q], xyz[3 * q + 1], xyz[3 * q + 2]);
      check(mgb_solver_set_coefficient_values(s_, level_, c.data()));
      inverse_diagonal_.clear();
    }

  private:
    static constexpr int prec = precision_of<number>::value;
    mutable VectorType   inverse_diagonal_;
    mgb_solver           s_     = nullptr;
    int                  level_ = 0;
  };
} // namespace multigrid
#endif
