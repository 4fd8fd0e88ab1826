// multigrid::DeviceVector<Number> -- the subset of dealii::LinearAlgebra::distributed::Vector<Number> that the
// reference's hot path touches (common/multigrid_solver.h:54-67,387-619; poisson_cube/program.cc:262-380),
// over device memory owned through the C ABI (include/mgb200.h).  Driver code ports by changing a typedef.
//
// Semantics that differ from deal.II and why:
//  * the vector belongs to one level of one solver (it needs the solver's partition for sums and ghosts);
//  * update_ghost_values()/compress() exist for source compatibility; every library call refreshes the ghosts
//    it reads itself (the reference's vmult does the same inside cell_loop, laplace_operator.h:590).
#ifndef MULTIGRID_DEVICE_VECTOR_H
#define MULTIGRID_DEVICE_VECTOR_H

#include <cmath>
#include <cstddef>
#include <stdexcept>
#include <string>
#include <type_traits>
#include <utility>
#include <vector>

#include "../mgb200.h"

namespace multigrid
{
  // AssertThrow of the reference (poisson_cube/program.cc:717-737): C ABI error codes become exceptions
  inline void
  check(const int code)
  {
    if (code != 0)
      throw std::runtime_error(std::string("mgb200: ") + mgb_last_error());
  }

  template <typename Number>
  struct precision_of
  {
    static_assert(std::is_same<Number, float>::value || std::is_same<Number, double>::value, "float or double");
    static constexpr int value = std::is_same<Number, double>::value ? MGB_F64 : MGB_F32;
  };

  template <typename Number>
  class DeviceVector
  {
  public:
    typedef Number value_type;

    DeviceVector() = default;
    DeviceVector(const DeviceVector &) = delete;
    DeviceVector &operator=(const DeviceVector &) = delete;
    DeviceVector(DeviceVector &&o) noexcept { swap(o); }
    ~DeviceVector() { clear(); }

    // initialize_dof_vector of the operator: zero vector of the level's local size
    void
    reinit(mgb_solver solver, const int level)
    {
      clear();
      s_     = solver;
      level_ = level;
      int64_t info[12];
      check(mgb_solver_partition(s_, level_, info));
      global_size_ = (std::size_t)info[1];
      local_size_  = (std::size_t)info[2];
      owned_size_  = (std::size_t)info[4];
      void *p      = nullptr;
      check(mgb_device_alloc(s_, local_size_ * sizeof(Number), &p));
      data_ = static_cast<Number *>(p);
    }
    void
    reinit(const DeviceVector &other)
    {
      reinit(other.s_, other.level_);
    }
    void
    clear()
    {
      if (data_)
        mgb_device_free(s_, data_);
      data_ = nullptr;
    }

    Number       *begin() { return data_; }
    const Number *begin() const { return data_; }
    std::size_t   size() const { return global_size_; }                // all ranks
    std::size_t   locally_owned_size() const { return owned_size_; }   // this rank, without ghost planes
    std::size_t   local_storage_size() const { return local_size_; }   // owned + ghost planes
    mgb_solver    solver() const { return s_; }
    int           level() const { return level_; }

    DeviceVector &
    operator=(const Number value)
    {
      check(mgb_vec_set(s_, level_, precision_of<Number>::value, data_, (double)value));
      return *this;
    }
    // Vector::operator=(const Vector<OtherNumber> &): precision-converting copy (same solver and level)
    template <typename Other>
    typename std::enable_if<!std::is_same<Other, Number>::value, DeviceVector &>::type
    operator=(const DeviceVector<Other> &x)
    {
      if (data_ == nullptr)
        reinit(x.solver(), x.level());
      check(mgb_vec_convert(s_, level_, precision_of<Number>::value, data_, x.begin()));
      return *this;
    }
    Number
    operator*(const DeviceVector &other) const
    {
      double r = 0;
      check(mgb_vec_dot(s_, level_, precision_of<Number>::value, data_, other.data_, &r));
      return (Number)r;
    }
    Number l2_norm() const { return (Number)std::sqrt((double)((*this) * (*this))); }
    // this = s * this + a * x
    void
    sadd(const Number s, const Number a, const DeviceVector &x)
    {
      check(mgb_vec_sadd(s_, level_, precision_of<Number>::value, data_, (double)s, (double)a, x.data_));
    }
    void add(const Number a, const DeviceVector &x) { sadd(Number(1), a, x); }
    void
    copy_locally_owned_data_from(const DeviceVector &x)
    {
      sadd(Number(0), Number(1), x);
    }
    void
    swap(DeviceVector &o) noexcept
    {
      std::swap(s_, o.s_);
      std::swap(level_, o.level_);
      std::swap(data_, o.data_);
      std::swap(global_size_, o.global_size_);
      std::swap(local_size_, o.local_size_);
      std::swap(owned_size_, o.owned_size_);
    }
    void update_ghost_values() const
    {
      check(mgb_vec_update_ghost_values(s_, level_, precision_of<Number>::value, data_));
    }
    void zero_out_ghost_values() const
    {
      check(mgb_vec_zero_out_ghost_values(s_, level_, precision_of<Number>::value, data_));
    }
    // compress(VectorOperation::add): ghost entries are added to their owners on the neighbouring z-slabs and
    // zeroed (no-op on one GPU and on replicated levels)
    void compress() { check(mgb_vec_compress_add(s_, level_, precision_of<Number>::value, data_)); }

    // host <-> device in the library's numbering (mgb_solver_lex_map gives the permutation)
    void
    import_from_host(const std::vector<Number> &host)
    {
      if (host.size() != local_size_)
        throw std::runtime_error("mgb200: host vector has the wrong size");
      check(mgb_copy_to_device(s_, data_, host.data(), local_size_ * sizeof(Number)));
    }
    std::vector<Number>
    export_to_host() const
    {
      std::vector<Number> host(local_size_);
      check(mgb_solver_sync(s_));
      check(mgb_copy_to_host(s_, host.data(), data_, local_size_ * sizeof(Number)));
      return host;
    }

  private:
    mgb_solver  s_ = nullptr;
    int         level_ = 0;
    Number     *data_ = nullptr;
    std::size_t global_size_ = 0, local_size_ = 0, owned_size_ = 0;
  };
} // namespace multigrid
#endif
