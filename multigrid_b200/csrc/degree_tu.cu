// One translation unit per polynomial degree: compile with -DMGB_P=<p>, p = 1..8
// (the reference precompiles p = 1..9 the same way: poisson_cube/program.cc:614-643).
#ifndef MGB_P
#error "define MGB_P"
#endif
#include "degree_kernels.cuh"
#include "cartesian_kernel.cuh"
#include "transfer_kernels.cuh"

namespace mgb
{
#define MGB_CAT2(a, b) a##b
#define MGB_CAT(a, b) MGB_CAT2(a, b)
  const DegreeKernels &MGB_CAT(degree_kernels_, MGB_P)()
  {
    static const DegreeKernels k = {&init_constants<MGB_P>,
                                    &launch_apply_general<MGB_P, float>,
                                    &launch_apply_general<MGB_P, double>,
                                    &launch_apply_cartesian<MGB_P, float>,
                                    &launch_apply_cartesian<MGB_P, double>,
                                    &launch_transfer<MGB_P, float>,
                                    &launch_transfer<MGB_P, double>,
                                    &launch_transfer_tile<MGB_P, float>,
                                    &launch_transfer_tile<MGB_P, double>,
                                    &launch_diagonal<MGB_P, float>,
                                    &launch_diagonal<MGB_P, double>,
                                    &launch_coefficient<MGB_P, float>,
                                    &launch_coefficient<MGB_P, double>,
                                    &launch_rhs_values<MGB_P>,
                                    &launch_quadrature_points<MGB_P>,
                                    &launch_l2_error<MGB_P>,
                                    &launch_solution_coefficient<MGB_P, float>,
                                    &launch_solution_coefficient<MGB_P, double>,
                                    &launch_interpolate_to_coarse<MGB_P>};
    return k;
  }
} // namespace mgb
