// multigrid::LaplaceOperatorCompactCombine<dim, fe_degree, Number, type> -- names and argument order of
// common/laplace_operator_dg.h:360-363,771-955 over the C ABI (include/mgb200.h, mgb_dg_*).  type = 0 (FE_DGQHermite),
// 1 (FE_DGQ on Gauss-Lobatto points) and 2 (FE_DGQArbitraryNodes(QGauss)) as in the reference (:346-348); vectors hold
// the coefficients of that basis.  The Chebyshev preconditioner is the point-Jacobi diagonal or JacobiTransformed.
#ifndef MULTIGRID_LAPLACE_OPERATOR_DG_H
#define MULTIGRID_LAPLACE_OPERATOR_DG_H

#include <array>

#include "device_vector.h"

namespace multigrid
{
  // plain device array for the DG operator (cell-wise contiguous DoFs, one GPU)
  template <typename Number>
  class DGVector
  {
  public:
    DGVector() = default;
    DGVector(const DGVector &) = delete;
    DGVector &operator=(const DGVector &) = delete;
    ~DGVector() { clear(); }
    void
    reinit(const std::size_t n)
    {
      clear();
      n_ = n;
      check(mgb_device_alloc(nullptr, n * sizeof(Number), reinterpret_cast<void **>(&data_)));
    }
    void
    clear()
    {
      if (data_)
        mgb_device_free(nullptr, data_);
      data_ = nullptr;
    }
    Number       *begin() { return data_; }
    const Number *begin() const { return data_; }
    std::size_t   size() const { return n_; }
    void swap(DGVector &o) noexcept
    {
      std::swap(data_, o.data_);
      std::swap(n_, o.n_);
    }
    void import_from_host(const std::vector<Number> &h) { check(mgb_copy_to_device(nullptr, data_, h.data(), n_ * sizeof(Number))); }
    std::vector<Number>
    export_to_host() const
    {
      std::vector<Number> h(n_);
      check(mgb_copy_to_host(nullptr, h.data(), data_, n_ * sizeof(Number)));
      return h;
    }

  private:
    Number     *data_ = nullptr;
    std::size_t n_    = 0;
  };

  template <int dim, int fe_degree, typename Number, int type = 0>
  class LaplaceOperatorCompactCombine
  {
    static_assert(dim == 3 && type >= 0 && type <= 2, "3-D, type 0 (Hermite), 1 (Gauss-Lobatto) or 2 (Gauss)");

  public:
    typedef DGVector<Number> VectorType;
    LaplaceOperatorCompactCombine() = default;
    LaplaceOperatorCompactCombine(const LaplaceOperatorCompactCombine &) = delete;
    ~LaplaceOperatorCompactCombine()
    {
      if (h_)
        mgb_dg_destroy(h_);
    }
    // reinit(matrix_free, dof_index)  laplace_operator_dg.h:360-769; the mesh comes as options (matvec_dg/program.cc:55-78)
    void
    reinit(mgb_dg_options o)
    {
      o.degree = fe_degree;
      o.basis  = MGB_DG_BASIS_OF_TYPE(type);
      check(mgb_dg_create(&o, &h_));
    }
    std::size_t m() const { return (std::size_t)mgb_dg_n_dofs(h_); }
    void        initialize_dof_vector(VectorType &v) const { v.reinit(m()); }
    Number
    get_penalty(const unsigned int /*cell*/, const unsigned int face) const
    {
      double s[3];
      check(mgb_dg_penalty(h_, s));
      return (Number)s[face / 2];
    }
    void vmult(VectorType &dst, const VectorType &src) const { check(mgb_dg_vmult(h_, prec, dst.begin(), src.begin())); }
    void
    vmult_residual(const VectorType &rhs, const VectorType &lhs, VectorType &residual) const
    {
      check(mgb_dg_vmult_residual(h_, prec, rhs.begin(), lhs.begin(), residual.begin()));
    }
    std::array<Number, 4>
    vmult_with_cg_update(const Number alpha, const Number beta, const VectorType &r, VectorType &q, VectorType &p,
                         VectorType &x) const
    {
      double s[4];
      check(mgb_dg_vmult_with_cg_update(h_, prec, alpha, beta, r.begin(), q.begin(), p.begin(), x.begin(), s));
      return {{(Number)s[0], (Number)s[1], (Number)s[2], (Number)s[3]}};
    }
    // laplace_operator_dg.h:907-955 (the preconditioner argument is the operator's own point-Jacobi diagonal; like
    // the reference the vectors are swapped for iteration_index > 0)
    void
    vmult_with_chebyshev_update(const VectorType &rhs, const unsigned int iteration_index, const Number factor1,
                                const Number factor2, VectorType &solution, VectorType &solution_old) const
    {
      check(mgb_dg_vmult_with_chebyshev_update(h_, prec, rhs.begin(), (int)iteration_index, factor1, factor2,
                                               solution.begin(), solution_old.begin()));
      if (iteration_index > 0)
        solution.swap(solution_old);
    }
    // preconditioner of the Chebyshev update: point Jacobi, or the JacobiTransformed object that the reference
    // passes as `prec` (laplace_operator_dg.h:2028-2256, matvec_dg_cheby/program.cc:108)
    enum Preconditioner
    {
      point_jacobi       = 0,
      jacobi_transformed = 1
    };
    void set_preconditioner(const Preconditioner kind) { check(mgb_dg_set_preconditioner(h_, (int)kind)); }
    // JacobiTransformed::vmult  :2046-2083 (or the point-Jacobi scaling)
    void
    preconditioner_vmult(VectorType &dst, const VectorType &src) const
    {
      check(mgb_dg_preconditioner_vmult(h_, prec, dst.begin(), src.begin()));
    }
    // :1863-1894 and :851-860: the FE_Q space is level `level` of a multigrid::MultigridSolver on the same mesh
    void
    prolongate_add_cg_to_dg(VectorType &dst, const DeviceVector<Number> &src_cg) const
    {
      check(mgb_dg_prolongate_add_cg_to_dg(h_, src_cg.solver(), src_cg.level(), prec, dst.begin(), src_cg.begin()));
    }
    void
    vmult_residual_and_restrict_to_cg(const VectorType &rhs, const VectorType &lhs, DeviceVector<Number> &residual_cg) const
    {
      check(mgb_dg_vmult_residual_and_restrict_to_cg(h_, residual_cg.solver(), residual_cg.level(), prec, rhs.begin(),
                                                     lhs.begin(), residual_cg.begin()));
    }
    void synchronize() const { check(mgb_dg_sync(h_)); }

  private:
    static constexpr int prec = precision_of<Number>::value;
    mgb_dg               h_   = nullptr;
  };
} // namespace multigrid
#endif
