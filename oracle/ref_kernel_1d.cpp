// ORACLE side (test infrastructure): C entry points around the reference's own 1-D even-odd product
// `apply_1d_matvec_kernel` (/root/reference/common/matrix_vector_kernel.h:30-216), the only piece of the
// reference that compiles without deal.II.  The header is included from where it lies under /root/reference
// (never copied); the recipe is oracle/Makefile target `_ref/libref1d.so`, the output stays in oracle/_ref/.
// tests/test_oracle.py::test_reference_even_odd_kernel checks the oracle's 1-D tables (oracle/basis1d.h) against
// it: the reference's kernel applied to the even-odd compressed tables must reproduce the plain dense products
// the oracle (and the CUDA kernels' host tables) are built from.
#define ALWAYS_INLINE
#include <matrix_vector_kernel.h>

namespace
{
  // one-lane stand-in for deal.II's VectorizedArray<double>: the kernel's (untaken) non-temporal branch needs the
  // members streaming_store and operator[] to exist
  struct Lane
  {
    double v;
    Lane(double x = 0.)
      : v(x)
    {}
    double &
    operator[](int)
    {
      return v;
    }
    void
    streaming_store(double *p) const
    {
      *p = v;
    }
    Lane &
    operator+=(const Lane &o)
    {
      v += o.v;
      return *this;
    }
  };
  inline Lane operator+(const Lane &a, const Lane &b) { return Lane(a.v + b.v); }
  inline Lane operator-(const Lane &a, const Lane &b) { return Lane(a.v - b.v); }
  inline Lane operator*(const Lane &a, const Lane &b) { return Lane(a.v * b.v); }

  template <int nn, int type, bool rows>
  void
  run(const double *eo, const double *in, double *out)
  {
    Lane c[nn * ((nn + 1) / 2)], x[nn], y[nn];
    for (int i = 0; i < nn * ((nn + 1) / 2); ++i)
      c[i] = eo[i];
    for (int i = 0; i < nn; ++i)
      x[i] = in[i];
    apply_1d_matvec_kernel<nn, 1, type, rows, false, Lane>(c, x, y);
    for (int i = 0; i < nn; ++i)
      out[i] = y[i].v;
  }

  template <int nn>
  int
  dispatch(int type, int rows, const double *eo, const double *in, double *out)
  {
    if (type == 0 && rows)
      run<nn, 0, true>(eo, in, out);
    else if (type == 0)
      run<nn, 0, false>(eo, in, out);
    else if (type == 1 && rows)
      run<nn, 1, true>(eo, in, out);
    else if (type == 1)
      run<nn, 1, false>(eo, in, out);
    else
      return 1;
    return 0;
  }
} // namespace

extern "C" int
ref_apply_1d(int nn, int type, int contract_over_rows, const double *coefficients_eo, const double *in, double *out)
{
  switch (nn)
    {
      case 2: return dispatch<2>(type, contract_over_rows, coefficients_eo, in, out);
      case 3: return dispatch<3>(type, contract_over_rows, coefficients_eo, in, out);
      case 4: return dispatch<4>(type, contract_over_rows, coefficients_eo, in, out);
      case 5: return dispatch<5>(type, contract_over_rows, coefficients_eo, in, out);
      case 6: return dispatch<6>(type, contract_over_rows, coefficients_eo, in, out);
      case 7: return dispatch<7>(type, contract_over_rows, coefficients_eo, in, out);
      case 8: return dispatch<8>(type, contract_over_rows, coefficients_eo, in, out);
      case 9: return dispatch<9>(type, contract_over_rows, coefficients_eo, in, out);
      default: return 2;
    }
}
