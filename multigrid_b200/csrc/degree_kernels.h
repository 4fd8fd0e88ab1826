// Interface between the degree-independent host code (solver.cu) and the per-degree
// translation units (degree_tu.cu compiled once per p = 1..8).
#pragma once
#include <cuda_runtime.h>

#include "basis.h"
#include "geometry.h"
#include "level.h"

namespace mgb
{
  // operator apply:  what happens with the result y = K x of the cell loop
  enum ApplyMode
  {
    APPLY_ADD = 0,      // dst += y                       (vmult_add / apply_add, laplace_operator.h:562-569)
    APPLY_SET = 1,      // dst  = y, dst_B = src_B        (vmult, :575-601)
    APPLY_RESIDUAL = 2, // dst  = rhs - y, dst_B -= src_B (vmult_residual, :607-634)
    APPLY_CHEBYSHEV = 3 // dst  = (1+f1) src - f1 dst + f2 P^-1 (rhs - y)   (PreconditionChebyshev via
                        //        vmult(dst,src,before,after), :725-741; dst holds x_{k-1} on entry)
  };

  template <typename T>
  struct ApplyArgs
  {
    LevelDesc L;
    const T  *src;
    T        *dst;
    const T  *rhs;      // APPLY_RESIDUAL / APPLY_CHEBYSHEV
    const T  *inv_diag; // APPLY_CHEBYSHEV
    const T  *coef_q;   // per-quadrature-point merged coefficient [cell][6][n^3] or nullptr
    T         coef[6];  // affine constant-coefficient merged tensor (without quadrature weight)
    T         f1, f2;
    T         f0;            // zero-guess Chebyshev (first_chebyshev >= 2): x_1 = f0 P^-1 rhs
    int       mode;
    int       read_plain;    // 1: do not zero constrained entries on read (compute_residual)
    int       first_chebyshev; // 1: x_{k-1} = 0 implied (iteration_index == 1); tile kernel, float: 2 = index 1 of a zero
                               // initial guess (src = f0 P^-1 rhs formed on the fly, `src` unused), 3 = index 2 of it
                               // (x_{k-1} = f0 P^-1 rhs recomputed instead of read from dst)
    int       layer_begin, layer_end; // tile kernel: local cell layers of this launch (-1: all owned layers)
    int       layer_begin2;           // tile kernel: if >= 0 the launch works on the two single layers layer_begin and layer_begin2
    double   *dot_out;   // APPLY_SET on the tile kernel: if set, sum of src[i] * dst[i] over the stored points is added here
  };

  template <typename T>
  struct TransferArgs
  {
    LevelDesc fine, coarse;
    const T  *src;
    T        *dst;
    int       mode;        // 0 prolongate (overwrite), 1 prolongate_and_add, 2 restrict_and_add
    int       constrained; // MGTransferMatrixFree::initialize_constraints called
    // z-slabs (tile kernels): the parent layers [cz_begin, cz_end) in the local numbering of the coarse
    // level are processed; the children of local coarse plane 0 start at local fine plane `f_shift`
    // (may be negative), and `close_top` says that the last layer also owns the top fine plane
    int cz_begin, cz_end, f_shift, close_top;
  };

  template <typename T>
  struct DiagArgs
  {
    LevelDesc L;
    T        *diag; // zero on entry; accumulates the unconstrained diagonal
    const T  *coef_q;
    T         coef[6];
  };

  struct RhsArgs
  {
    LevelDesc L;
    Geometry  geom;
    double   *dst; // += int f phi
    int       fn_id;
    const double *values = nullptr; // caller-supplied f at the quadrature points [cell][n^3] instead of fn_id
  };

  struct ErrorArgs
  {
    LevelDesc     L;
    Geometry      geom;
    const double *u;
    int           fn_id;
    double       *accum; // [0] += error^2, [1] += volume
    const double *values = nullptr; // caller-supplied exact solution at the quadrature points [cell][n^3] instead of fn_id
  };

  // quadrature points of all cells, [cell][n^3][3] (FEEvaluation::quadrature_point, laplace_operator.h:400-407)
  struct PointsArgs
  {
    LevelDesc L;
    Geometry  geom;
    double   *xyz;
  };

  // merged coefficient c(x_q) JxW J^-1 J^-T per quadrature point on a mapped mesh (laplace_operator.h:388-430)
  template <typename T>
  struct CoefArgs
  {
    LevelDesc L;
    Geometry  geom;
    T        *coef_q; // [cell][6][n^3], entries 00, 11, 22, 01, 02, 12
    int       fn_id;
  };

  // merged coefficient from the gradient of a level vector (MinimalSurfaceOperator::evaluate_coefficient,
  // minimal_surface/program.cc:117-163); mode 0 identity tensor, 1 Newton linearisation, 2 residual flux
  template <typename T>
  struct SolutionCoefArgs
  {
    LevelDesc L;
    Geometry  geom;
    T        *coef_q;
    const T  *u;
    int       mode;
  };

  struct DegreeKernels
  {
    void (*init)(const Basis &);
    void (*apply_general_f)(const ApplyArgs<float> &, cudaStream_t);
    void (*apply_general_d)(const ApplyArgs<double> &, cudaStream_t);
    void (*apply_cartesian_f)(const ApplyArgs<float> &, cudaStream_t);
    void (*apply_cartesian_d)(const ApplyArgs<double> &, cudaStream_t);
    void (*transfer_f)(const TransferArgs<float> &, cudaStream_t);
    void (*transfer_d)(const TransferArgs<double> &, cudaStream_t);
    void (*transfer_tile_f)(const TransferArgs<float> &, cudaStream_t);
    void (*transfer_tile_d)(const TransferArgs<double> &, cudaStream_t);
    void (*diagonal_f)(const DiagArgs<float> &, cudaStream_t);
    void (*diagonal_d)(const DiagArgs<double> &, cudaStream_t);
    void (*coefficient_f)(const CoefArgs<float> &, cudaStream_t);
    void (*coefficient_d)(const CoefArgs<double> &, cudaStream_t);
    void (*rhs_values)(const RhsArgs &, cudaStream_t);
    void (*quadrature_points)(const PointsArgs &, cudaStream_t);
    void (*l2_error)(const ErrorArgs &, cudaStream_t);
    void (*solution_coefficient_f)(const SolutionCoefArgs<float> &, cudaStream_t);
    void (*solution_coefficient_d)(const SolutionCoefArgs<double> &, cudaStream_t);
    void (*interpolate_to_coarse)(const TransferArgs<float> &, cudaStream_t);
  };

  const DegreeKernels &degree_kernels(int p);

#define MGB_DECLARE_DEGREE(P) const DegreeKernels &degree_kernels_##P();
  MGB_DECLARE_DEGREE(1)
  MGB_DECLARE_DEGREE(2)
  MGB_DECLARE_DEGREE(3)
  MGB_DECLARE_DEGREE(4)
  MGB_DECLARE_DEGREE(5)
  MGB_DECLARE_DEGREE(6)
  MGB_DECLARE_DEGREE(7)
  MGB_DECLARE_DEGREE(8)
} // namespace mgb
