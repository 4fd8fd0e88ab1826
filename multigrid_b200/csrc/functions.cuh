// Analytic problem data evaluated on the device at set-up time.
// Mirrors Solution / RightHandSide / Coefficient of poisson_cube/program.cc:79-144 and
// poisson_shell/program.cc:70-225.
#pragma once

namespace mgb
{
  enum FnId
  {
    FN_ZERO = 0,
    FN_ONE,
    FN_CUBE_SOLUTION,
    FN_CUBE_RHS,
    FN_SHELL_SOLUTION,
    FN_SHELL_RHS,
    FN_SHELL_COEFFICIENT,
    FN_ONE_PER_POINT // the value 1 stored per quadrature point (to be scaled by caller-supplied samples)
  };

  __host__ __device__ inline double
  eval_fn(const int id, const double x, const double y, const double z)
  {
    const double pi = 3.14159265358979323846;
    switch (id)
      {
        case FN_ONE:
        case FN_ONE_PER_POINT:
          return 1.;
        case FN_CUBE_SOLUTION: // poisson_cube/program.cc:98-104, wave_number = 3
          return sin(pi * x * 3.) * sin(pi * y * 3.) * sin(pi * z * 3.);
        case FN_CUBE_RHS: // :140-144
          return 3 * pi * 3. * pi * 3. * (sin(pi * x * 3.) * sin(pi * y * 3.) * sin(pi * z * 3.));
        case FN_SHELL_SOLUTION: // poisson_shell/program.cc:97-101
          return sin(2 * pi * (x + y));
        case FN_SHELL_COEFFICIENT: // :159-171
          {
            const double c0 = cos(2 * pi * x), c1 = cos(2 * pi * y + 0.1), c2 = cos(2 * pi * z + 0.2);
            return 1.0 + 1.0e6 * (c0 * c0) * (c1 * c1) * (c2 * c2);
          }
        case FN_SHELL_RHS: // :219-225: -(lap u * c + grad c . grad u)
          {
            const double c0 = cos(2 * pi * x), c1 = cos(2 * pi * y + 0.1), c2 = cos(2 * pi * z + 0.2);
            const double s0 = sin(2 * pi * x), s1 = sin(2 * pi * y + 0.1);
            const double coef = 1.0 + 1.0e6 * (c0 * c0) * (c1 * c1) * (c2 * c2);
            const double lap  = -2 * 2 * pi * 2 * pi * sin(2 * pi * (x + y));
            const double gsol = 2 * pi * cos(2 * pi * (x + y));
            const double gc0  = 1.0e6 * (-4 * pi * c0 * s0) * (c1 * c1) * (c2 * c2);
            const double gc1  = 1.0e6 * (c0 * c0) * (-4 * pi * c1 * s1) * (c2 * c2);
            return -(lap * coef + (gc0 + gc1) * gsol);
          }
        default:
          return 0.;
      }
  }
} // namespace mgb
