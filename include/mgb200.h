/* mgb200 — C ABI of the B200-native matrix-free multigrid path.
 *
 * The reference (kronbichler/multigrid) has no FFI layer: its boundary is the set of C++
 * duck-typed concepts that deal.II's solvers and its own MultigridSolver call on the
 * operator.  Every entry point below names the reference interface it replaces
 * (file:line under /root/reference).  All functions return 0 on success and a non-zero
 * code on failure; mgb_last_error() gives the message (the header adapters re-throw it,
 * mirroring AssertThrow, poisson_cube/program.cc:717-737).  No C++ types, no torch types.
 *
 * Conventions
 *   - handles are opaque pointers; a solver owns its level operators, smoothers,
 *     transfers and level vectors (multigrid_solver.h:709-751).
 *   - "device pointer" = plain CUDA device address (e.g. torch.Tensor.data_ptr()).
 *     "host pointer"   = ordinary host memory; the call copies inside.
 *   - DoF vectors use the library's renumbered layout (z-plane major; inside a plane tile by tile,
 *     16 x 16 points per tile); mgb_solver_lex_map() gives the permutation to the
 *     lexicographic (x fastest) numbering of the (N_d p + 1)^3 grid.
 *   - several GPUs (one process per GPU, comm_size > 1): the mesh is cut into z-slabs of whole cell
 *     layers, one per rank (the role of the p4est partition, poisson_cube/program.cc:193-195).  A
 *     vector of a distributed level holds the local box = ghost planes below, owned planes, one ghost
 *     plane on top (mgb_solver_partition()); every call refreshes the ghosts it reads
 *     (update_ghost_values / compress of LinearAlgebra::distributed::Vector) through NCCL, and
 *     sums and norms run over the owned planes of all ranks.  Levels too coarse to be cut are held
 *     (and computed) completely by every rank.
 *   - precision: MGB_F32 selects the V-cycle (float) level operator `matrix[level]`,
 *     MGB_F64 the outer operator `matrix_dp[level]` (multigrid_solver.h:740-745).
 *   - all work is ordered on one CUDA stream per solver; mgb_solver_sync() waits for it.
 */
#ifndef MGB200_H
#define MGB200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct mgb_solver_t *mgb_solver;

enum { MGB_F32 = 0, MGB_F64 = 1 };

/* analytic problem definitions evaluated on the device at set-up time */
enum {
  MGB_PROBLEM_CUBE_SINE = 0, /* poisson_cube/program.cc:79-144, coefficient 1            */
  MGB_PROBLEM_SHELL_COEF = 1 /* poisson_shell/program.cc:70-225 functions (variable coeff) */
};

/* geometry of the structured mesh: the map from the unit box of mesh coordinates to physical space; inside a cell
 * the geometry is the degree-p polynomial through the mapped support points (MappingQGeneric, multigrid_solver.h:138-139) */
enum {
  MGB_GEOMETRY_BOX = 0,          /* [lower, upper], axis-aligned cells (subdivided_hyper_rectangle)                        */
  MGB_GEOMETRY_SHELL_SECTOR = 1, /* one of the six coarse cells of hyper_shell(0, r0, r1, 6), poisson_shell/program.cc:426,
                                    and its refinements; geometry_params = {r0, r1, opening half-angle (pi/4: whole face)} */
  MGB_GEOMETRY_AFFINE = 2        /* x = lower + A ((upper - lower) o xi), geometry_params[0..8] = A row-major (sheared cells) */
};

/* kernel selection for the operator apply */
enum {
  MGB_KERNEL_AUTO = 0,     /* Cartesian streaming tile kernel where admissible              */
  MGB_KERNEL_GENERAL = 1   /* FEEvaluation-style evaluate / q-point op / integrate per cell */
};

typedef struct mgb_options {
  int degree;              /* FE_Q degree p, 1..8                 (LaplaceRunTime, poisson_cube/program.cc:614-643) */
  int n_subdivisions[3];   /* coarse cells per direction           (subdivided_hyper_cube/_rectangle, :528,:542)     */
  int n_refinements;       /* global refinements = n_levels - 1    (:570)                                            */
  double lower[3], upper[3];
  int degree_pre;          /* Chebyshev degree pre-smoother        (multigrid_solver.h:104)                          */
  int degree_post;         /* must equal degree_pre                (multigrid_solver.h:126-128)                      */
  int n_cycles;            /* V-cycles per FMG level               (multigrid_solver.h:106)                          */
  int problem;             /* MGB_PROBLEM_*                                                                          */
  int kernel;              /* MGB_KERNEL_*                                                                           */
  int device;              /* CUDA device ordinal                                                                    */
  int tile_cells;          /* cells per tile and direction (0 = automatic: about 16 points per tile direction)       */
  int comm_rank;           /* this process' rank, 0 .. comm_size-1  (MPI_Comm_rank of the reference's MPI_COMM_WORLD) */
  int comm_size;           /* number of ranks = GPUs (0 or 1: one GPU, no communicator)                              */
  unsigned char comm_id[128]; /* mgb_comm_unique_id() of rank 0, distributed to all ranks by the caller              */
  int geometry;            /* MGB_GEOMETRY_*  (MGB_GEOMETRY_BOX unless set)                                          */
  double geometry_params[12];
} mgb_options;

const char *mgb_last_error(void);
int mgb_version(void);
void mgb_default_options(mgb_options *opt);

/* Communicator bootstrap (replaces MPI_Init / MPI_COMM_WORLD of poisson_cube/program.cc:704-712): rank 0 calls
 * this, the caller broadcasts the 128 bytes by its own means (torch.distributed, MPI, a file) and every rank
 * passes them in mgb_options.comm_id; mgb_solver_create() is then collective over the ranks.             */
int mgb_comm_unique_id(unsigned char id[128]);

/* MultigridSolver<dim,p,float,double>::MultigridSolver   multigrid_solver.h:100-292 */
int mgb_solver_create(const mgb_options *opt, mgb_solver *out);
int mgb_solver_destroy(mgb_solver s);
int mgb_solver_sync(mgb_solver s);

int mgb_solver_n_levels(mgb_solver s);
int64_t mgb_solver_n_dofs(mgb_solver s, int level);        /* length of a vector of this rank (one GPU: LaplaceOperator::m()) */
/* The z-slab of this rank on one level (Utilities::MPI::Partitioner of the reference's vectors), 12 values:
 * 0 distributed (0/1), 1 global DoFs = LaplaceOperator::m(), 2 local vector length, 3 first owned index,
 * 4 number of owned entries, 5 global z-plane of local plane 0, 6 local planes, 7 planes of the level,
 * 8 ghost cell layers below, 9 ghost plane on top (0/1), 10 first owned global cell layer, 11 one past the last */
int mgb_solver_partition(mgb_solver s, int level, int64_t info[12]);
/* ghost exchanges and all-reduces issued so far (0 on one GPU) */
int64_t mgb_solver_exchange_count(mgb_solver s);
int mgb_solver_dof_grid(mgb_solver s, int level, int nd[3]);
int mgb_solver_tile_cells(mgb_solver s, int level);
/* lex_to_dof[i] = library index of the DoF with lexicographic index i (host pointer, n_dofs entries) */
int mgb_solver_lex_map(mgb_solver s, int level, uint32_t *lex_to_dof);

/* double MultigridSolver::solve(bool do_analyze)          multigrid_solver.h:387-476
 * trace (host, may be NULL): 4 doubles per level = error start, residual start, residual end, error end */
int mgb_solver_solve(mgb_solver s, int do_analyze, double *reduction_rate, double *trace);
/* pair<unsigned,double> MultigridSolver::solve_cg()       multigrid_solver.h:483-493 */
int mgb_solver_solve_cg(mgb_solver s, unsigned int *iterations, double *reduction_rate);
/* same, with the right-hand side taken from / the solution returned to HOST buffers in
 * library numbering (rhs may be NULL = keep the assembled one; solution may be NULL)      */
int mgb_solver_solve_cg_host(mgb_solver s, const double *rhs_host, double *solution_host,
                             unsigned int *iterations, double *reduction_rate);
/* residual history of the last solve_cg (host, up to n entries; returns count via *n)  */
int mgb_solver_cg_history(mgb_solver s, double *history, int *n);
/* double MultigridSolver::compute_l2_error(level)         multigrid_solver.h:298-343 */
int mgb_solver_l2_error(mgb_solver s, int level, double *error);
/* MultigridSolver::vmult(dst, src)  (one float V-cycle as preconditioner)  multigrid_solver.h:498-510 */
int mgb_solver_vcycle(mgb_solver s, double *dst_dev, const double *src_dev);
/* std::array<Number2,2> MultigridSolver::vmult_with_residual_update(residual, update, factor)   multigrid_solver.h:516-619
 * The preconditioner entry of a merged CG: defect = float(residual + factor update), one V-cycle, then in one pass
 * residual += factor update, update = V-cycle result, sums[0] = <update_new, residual_new>, sums[1] = <update_new,
 * factor update_old> over all ranks (factor == 0: residual untouched, sums[1] = sums[0]); Dirichlet rows are identity rows. */
int mgb_solver_vmult_with_residual_update(mgb_solver s, double *residual_dev, double *update_dev, double factor,
                                          double sums[2]);
/* Problem data supplied by the caller instead of the built-in MGB_PROBLEM_* functions -- what the reference passes as
 * Function<dim> objects (MultigridSolver(dof_handler, boundary_values, right_hand_side, coefficient, ...),
 * multigrid_solver.h:100-106; LaplaceOperator::evaluate_coefficient(const Function &) and compute_residual(dst, src,
 * rhs_function), laplace_operator.h:103-109).  The caller evaluates its functions at the points the library reports:
 *   quadrature points   [cell][q][3], cells of this rank's box in lexicographic order (x fastest), q = (q2 n + q1) n + q0
 *                       (FEEvaluation::quadrature_point); n_quadrature_points = cells (p+1)^3
 *   support points      [dof][3] in the library's numbering (VectorTools::interpolate_boundary_values positions)
 * set_coefficient_values: c at the quadrature points of one level (call it for every level, then reinit_smoothers, which
 * recomputes the diagonals and the Chebyshev eigenvalue estimates); set_rhs_values: f at the quadrature points and the
 * Dirichlet values at the support points (one value per DoF, non-boundary entries ignored; NULL = homogeneous) of one
 * level -> rhs = int f phi - K u_bc (solve_cg needs the finest level, the FMG solve() all levels);
 * set_exact_solution_values: u at the quadrature points, used by mgb_solver_l2_error instead of the built-in solution. */
int64_t mgb_solver_n_quadrature_points(mgb_solver s, int level);
int mgb_solver_quadrature_points(mgb_solver s, int level, double *xyz_host);
int mgb_solver_support_points(mgb_solver s, int level, double *xyz_host);
int mgb_solver_set_coefficient_values(mgb_solver s, int level, const double *c_host);
int mgb_solver_reinit_smoothers(mgb_solver s);
int mgb_solver_set_rhs_values(mgb_solver s, int level, const double *f_host, const double *u_boundary_host);
int mgb_solver_set_exact_solution_values(mgb_solver s, int level, const double *u_host);

/* ---- Minimal surface problem: minimal_surface/program.cc, here in 3-D on the solver's mesh, one GPU.
 * Newton's method with line search for -div(grad u / sqrt(1 + |grad u|^2)) = 0 with Dirichlet data.  One step
 * (LaplaceProblem::solve(first_time), :412-572): the solution is carried to every level by the FE_Q restriction
 * matrices (:417-456); MinimalSurfaceOperator::evaluate_coefficient (:117-163) fills the merged coefficient
 * JxW J^-1 T J^-T on every level from the gradient of that vector -- T = I the first time, else the Newton tensor
 * (I - g g^T / (1 + |g|^2)) / sqrt(1 + |g|^2), all six entries per quadrature point; diagonals and Chebyshev
 * smoothers are rebuilt (:458-486; create the solver with degree_pre = degree_post = 2 for the reference's setting);
 * the update solves A du = r by CG with the V-cycle, ReductionControl(m(), 1e-13, 1e-4), r the nonlinear residual of
 * MinimalSurfaceOperator::compute_residual (:169-198); the step length is halved until the residual norm falls
 * (:543-557).  The iterate is the solver's finest-level solution vector (mgb_solver_get_solution / mgb_solver_vector). */
/* interpolate_boundary_values :396-408: Dirichlet values from u_host (one value per DoF of the finest level, library
 * numbering; only boundary entries are used) or, if NULL, from the analytic solution of options.problem; the iterate
 * is set to these values on the boundary and to zero inside                                                        */
int mgb_solver_minimal_surface_set_boundary_values(mgb_solver s, const double *u_host);
/* info = {initial residual norm, final residual norm, CG iterations, line search halvings, step length}            */
int mgb_solver_minimal_surface_step(mgb_solver s, int first_time, double info[5]);
/* compute_residual(dst, u, first_time) :169-198 on double device vectors of the finest level; *norm = |dst|_2      */
int mgb_solver_minimal_surface_residual(mgb_solver s, double *dst, const double *u, int first_time, double *norm);
/* coefficient_solutions[level] of the last step (float device vector of the level), :417-456                       */
int mgb_solver_minimal_surface_level_solution(mgb_solver s, int level, float **dev_ptr);
/* MultigridSolver::do_matvec / do_matvec_smoother         multigrid_solver.h:624-637 */
int mgb_solver_do_matvec(mgb_solver s);
int mgb_solver_do_matvec_smoother(mgb_solver s);
/* MultigridSolver::print_wall_times                       multigrid_solver.h:348-371 : 6 doubles per level
 * (seconds; columns mg_mv, restrict, prolongate, inhomBC, mg_vec, smoother as timings[level][0..5], :776);
 * measured with CUDA events on the solver stream while mgb_solver_profile(s, 1) is active       */
int mgb_solver_timings(mgb_solver s, double *timings, int reset);
/* switch the CUDA-event ranges on/off (the LIKWID regions of poisson_cube/program.cc:281-376)    */
int mgb_solver_profile(mgb_solver s, int on);
/* accumulated milliseconds / number of ranges per class on one level (16 entries each):
 * 0..5 as above, 6 cell-loop kernel double, 7 cell-loop kernel float, 8 cell-loop kernel float with the
 * fused Chebyshev update, 9 patch-boundary epilogue kernel, 10 outer CG vector work, 11 other    */
int mgb_solver_class_times(mgb_solver s, int level, double *ms, int64_t *count, int reset);
/* CUDA-event stopwatch on the solver's stream (Timer of poisson_cube/program.cc:283-322)         */
int mgb_timer_start(mgb_solver s);
int mgb_timer_stop(mgb_solver s, double *ms);
/* device address of a solver-owned vector: which = 0 solution, 1 rhs, 2 residual (double);
 * 3 defect, 4 t, 5 solution_update, 6 inverse diagonal (float)     multigrid_solver.h:709-735 */
int mgb_solver_vector(mgb_solver s, int which, int level, void **dev_ptr);
/* const Vector &MultigridSolver::get_solution()           multigrid_solver.h:376-382 (host copy) */
int mgb_solver_get_solution(mgb_solver s, double *solution_host);

/* LaplaceOperator<3,p,number> on level `level`            laplace_operator.h:56-164; vectors are device pointers
 * of the operator's number type */
int mgb_op_vmult(mgb_solver s, int level, int precision, void *dst, const void *src);           /* :575-601 */
int mgb_op_vmult_add(mgb_solver s, int level, int precision, void *dst, const void *src);       /* Base::vmult_add */
int mgb_op_Tvmult(mgb_solver s, int level, int precision, void *dst, const void *src);          /* Base::Tvmult (symmetric) */
/* vmult with HOST vectors in library numbering: copy in, apply, copy out (end-to-end call)      */
int mgb_op_vmult_host(mgb_solver s, int level, int precision, void *dst_host, const void *src_host);
int mgb_op_vmult_residual(mgb_solver s, int level, int precision, const void *rhs, const void *lhs,
                          void *residual);                                                       /* :607-634 */
/* compute_residual(dst, src, rhs_function) :103-106 (806-845), double operator: dst = (f, phi) - K src on the rows without
 * Dirichlet condition; src is read with its Dirichlet entries (the boundary values); f sampled at the quadrature
 * points of mgb_solver_quadrature_points (host array) or NULL for the right-hand side function of options.problem   */
int mgb_op_compute_residual(mgb_solver s, int level, double *dst, const double *src, const double *f_values_host);
/* MatrixFreeOperators::Base::el(row, row): the diagonal entry (1 on Dirichlet rows)                               */
int mgb_op_el(mgb_solver s, int level, int precision, int64_t row, double *value);
/* std::array<number,4> vmult_with_cg_update(alpha,beta,r,q,p,x)                                    :640-719 */
int mgb_op_vmult_with_cg_update(mgb_solver s, int level, int precision, double alpha, double beta,
                                const void *r, void *q, void *p, void *x, double sums[4]);
/* compute_diagonal() + get_matrix_diagonal_inverse(): copies 1/diag to a device buffer            :747-800 */
int mgb_op_inverse_diagonal(mgb_solver s, int level, int precision, void *inv_diag);
int mgb_op_set_kernel(mgb_solver s, int kernel);

/* PreconditionChebyshev<LaplaceOperator<float>>  (deal.II; call sites multigrid_solver.h:271-288,657-678) */
/* info[6] = min_eig, max_eig (Lanczos estimates), theta, delta, degree, eigenvalue-CG iterations */
int mgb_smoother_info(mgb_solver s, int level, double info[6]);
/* PreconditionChebyshev::initialize with other AdditionalData than the FE_Q solver's (multigrid_solver_dg.h:271-291):
 * degree < 0 = automatic (numbers::invalid_unsigned_int), eig_cg_n_iterations < 0 = m() */
int mgb_smoother_reinit(mgb_solver s, int level, double smoothing_range, int degree, int eig_cg_n_iterations);
/* v_cycle(maxlevel, 1) on the solver's own vectors: solution_update[maxlevel] = V-cycle(defect[maxlevel])
 * (multigrid_solver_dg.h:575-603; get the vectors with mgb_solver_vector, after the call)                      */
int mgb_solver_vcycle_inplace(mgb_solver s);
int mgb_smoother_vmult(mgb_solver s, int level, float *x, const float *b);  /* zero initial guess   */
int mgb_smoother_step(mgb_solver s, int level, float *x, const float *b);   /* nonzero guess        */

/* MGTransferMatrixFree<3,number>  (deal.II; call sites multigrid_solver.h:213-221,415,668,674); level = fine level */
int mgb_transfer_prolongate(mgb_solver s, int level, float *fine, const float *coarse);
int mgb_transfer_prolongate_and_add(mgb_solver s, int level, float *fine, const float *coarse);
int mgb_transfer_restrict_and_add(mgb_solver s, int level, float *coarse, const float *fine);
int mgb_transfer_prolongate_no_boundary(mgb_solver s, int level, double *fine, const double *coarse);

/* ---------------------------------------------------------------------------------------------------------
 * Discontinuous Galerkin: symmetric interior-penalty Laplacian, FE_DGQ(p) on an affine box mesh
 * (multigrid::LaplaceOperatorCompactCombine<dim,p,Number,type=1>, common/laplace_operator_dg.h:360-2026, verified by
 * the reference against MFReference::LaplaceOperatorFaceBased, common/laplace_operator_dg_face.h:13-166; mesh and
 * protocol of matvec_dg/program.cc:40-210 and matvec_dg_cheby/program.cc:40-200).
 * DoFs are cell-wise contiguous (dof_indices_contiguous, laplace_operator_dg.h:1141-1142): cells lexicographic with
 * x fastest, (k n + j) n + i inside a cell.  One kernel computes cell and face integrals of a cell and the fused
 * vector operation ("actions" 0, 3, 4 of vmult_with_merged_ops, laplace_operator_dg.h:957-962).               */
typedef struct mgb_dg_t *mgb_dg;

#define MGB_DG_BASIS_GAUSS_LOBATTO 0 /* FE_DGQ(p): Lagrange in the Gauss-Lobatto points         (reference type 1) */
#define MGB_DG_BASIS_HERMITE 1       /* FE_DGQHermite(p): Hermite-like interpolation            (reference type 0) */
#define MGB_DG_BASIS_GAUSS 2         /* FE_DGQArbitraryNodes(QGauss(p+1)): collocated with the quadrature (type 2) */
/* MGB_DG_BASIS_* of the reference's template parameter `type` (0, 1, 2) */
#define MGB_DG_BASIS_OF_TYPE(type) ((type) == 0 ? MGB_DG_BASIS_HERMITE : ((type) == 1 ? MGB_DG_BASIS_GAUSS_LOBATTO : MGB_DG_BASIS_GAUSS))

typedef struct mgb_dg_options {
  int degree;             /* 1..8                                                                   */
  int n_subdivisions[3];  /* coarse cells per direction      (matvec_dg/program.cc:61-64)           */
  int n_refinements;      /* global refinements              (:78)                                  */
  double lower[3], upper[3]; /* box before the vertex map    (:55-60)                               */
  double trafo[9];        /* row-major constant matrix applied to the vertices (:66-76)             */
  int device;
  /* one process per GPU (the reference's MPI ranks): z-slabs of whole cell layers, the number of layers must be
   * a multiple of comm_size; comm_id from mgb_comm_unique_id() of rank 0.  comm_size <= 1: one GPU             */
  int comm_rank, comm_size;
  unsigned char comm_id[128];
  /* 1-D basis of the DG space (template parameter `type` of LaplaceOperatorCompactCombine,
   * laplace_operator_dg.h:346-350; matvec_dg/program.cc:44-49): one of MGB_DG_BASIS_*.  Zero, the default, is
   * FE_DGQ; poisson_dg uses FE_DGQHermite (poisson_dg/program.cc:174).  Vectors hold the coefficients of this basis */
  int basis;
} mgb_dg_options;

void mgb_dg_default_options(mgb_dg_options *opt);        /* the mesh of matvec_dg/program.cc, degree 4 */
/* LaplaceOperatorCompactCombine::reinit                          laplace_operator_dg.h:360-769 */
int mgb_dg_create(const mgb_dg_options *opt, mgb_dg *out);
int mgb_dg_destroy(mgb_dg h);
int mgb_dg_sync(mgb_dg h);
int64_t mgb_dg_n_dofs(mgb_dg h);                                   /* m()                        :795-799 */
int mgb_dg_n_cells(mgb_dg h, int ncell[3]);
/* z-slab of this rank: info = {DoFs of the whole mesh, local vector length (owned + ghost cell layers), offset of
 * the owned range, owned DoFs, first owned cell layer, owned cell layers}.  Vectors passed to mgb_dg_* have the
 * local length; their ghost layers are refreshed by the library (the MPI face exchange of :981-1058)           */
int mgb_dg_partition(mgb_dg h, int64_t info[6]);
int64_t mgb_dg_exchange_count(mgb_dg h);
int mgb_dg_penalty(mgb_dg h, double sigma[3]);                     /* get_penalty, per face direction :789-793 */
/* vectors are device pointers of the chosen precision (MGB_F32 / MGB_F64) */
int mgb_dg_vmult(mgb_dg h, int precision, void *dst, const void *src);                       /* :827-841 */
int mgb_dg_vmult_residual(mgb_dg h, int precision, const void *rhs, const void *lhs, void *residual); /* :843-849 */
/* std::array<Number,4> vmult_with_cg_update(alpha,beta,r,q,p,x)                                  :862-905 */
int mgb_dg_vmult_with_cg_update(mgb_dg h, int precision, double alpha, double beta, const void *r, void *q,
                                void *p, void *x, double sums[4]);
/* exact inverse diagonal of the SIP matrix (point Jacobi) */
int mgb_dg_inverse_diagonal(mgb_dg h, int precision, void *inv_diag);
/* preconditioner of the Chebyshev update: 0 = point Jacobi (default), 1 = multigrid::JacobiTransformed
 * (laplace_operator_dg.h:2028-2256 with LocalBasisTransformer type 0, :176-222: cell-wise V D^-1 V^T in the tensor basis of
 * the generalised eigenvectors of the 1-D SIP Laplacian w.r.t. the 1-D mass matrix, D = diag(V^T A_KK V))     */
int mgb_dg_set_preconditioner(mgb_dg h, int kind);
/* JacobiTransformed::vmult (:2046-2083) or the point-Jacobi scaling: dst = P^-1 src */
int mgb_dg_preconditioner_vmult(mgb_dg h, int precision, void *dst, const void *src);
/* vmult_with_chebyshev_update(preconditioner, rhs, iteration_index, factor1, factor2, solution, solution_old)
 * :907-955 with the selected preconditioner P: iteration_index 0: solution = factor2 P^-1 rhs; otherwise
 * solution_old = (1+factor1) solution - factor1 solution_old + factor2 P^-1 (rhs - A solution)  (the caller swaps) */
int mgb_dg_vmult_with_chebyshev_update(mgb_dg h, int precision, const void *rhs, int iteration_index, double factor1,
                                       double factor2, void *solution, void *solution_old);
/* Same-degree transfer between the DG space and the FE_Q space of level `level` of solver `s` on the same mesh
 * (both vectors of the given precision, the FE_Q vector in the solver's numbering; on several GPUs operator and
 * level must hold the same z-slab and the caller refreshes the ghost planes of src_cg before prolongating, as
 * mgb_dg_solver does):
 * prolongate_add_cg_to_dg(dst_dg, src_cg)                                   laplace_operator_dg.h:1863-1894: dg += P cg
 * vmult_residual_and_restrict_to_cg(rhs, lhs, restricted_residual)           :851-860 (action 1, :1798-1819):
 *                                                                           restricted = P^T (rhs - A lhs)          */
int mgb_dg_prolongate_add_cg_to_dg(mgb_dg h, mgb_solver s, int level, int precision, void *dg, const void *cg);
int mgb_dg_vmult_residual_and_restrict_to_cg(mgb_dg h, mgb_solver s, int level, int precision, const void *rhs,
                                             const void *lhs, void *restricted_residual);
/* end-to-end: host vectors in, apply, host vector out */
int mgb_dg_vmult_host(mgb_dg h, int precision, void *dst_host, const void *src_host);
/* The CG solver of solver_dg/program.cc:180-262 on this operator: SolverCG with IterationNumberControl(n_iterations,
 * 1e-20) from x = 0 and a diagonal preconditioner (device vector; mgb_dg_lumped_mass_inverse gives the program's, the
 * inverse of (1, phi_i), :134-148).  interleaved = 0: "cell-based loops" (operator apply, then vector updates and dot
 * products as separate passes, :216-227); interleaved != 0: "interleaved loops" (LaplaceOperatorInterleave, :39-70 --
 * the vector updates ride in one kernel before the apply and the five sums in the apply's store stage; one GPU).
 * stats = {iterations done, last residual norm}                                                                    */
int mgb_dg_lumped_mass_inverse(mgb_dg h, int precision, void *out);
int mgb_dg_solve_cg(mgb_dg h, int precision, unsigned int n_iterations, int interleaved, void *x, const void *b,
                    const void *diagonal_preconditioner, double stats[2]);
int mgb_dg_timer_start(mgb_dg h);
int mgb_dg_timer_stop(mgb_dg h, double *ms);
int64_t mgb_dg_launch_count(mgb_dg h);

/* ---- DG multigrid solver: multigrid::MultigridSolverDG<3,p,float,double>     common/multigrid_solver_dg.h:44-735
 * Outer CG in double on the SIP-DG operator, preconditioned by one float cycle: Chebyshev(JacobiTransformed) on the
 * DG space, same-degree transfer to FE_Q on the same mesh, the FE_Q V-cycle of mgb_solver below it (finest FE_Q
 * smoother one degree lower, coarse range 2e-3, :272-293), transfer back, Chebyshev step (:605-633).
 * Mesh: box [lower, upper], n_subdivisions coarse cells, n_refinements (poisson_dg/program.cc:356-409); problem
 * data of poisson_dg/program.cc:70-135 (the same sine solution as poisson_cube).  One GPU.                    */
typedef struct mgb_dg_solver_t *mgb_dg_solver;

typedef struct mgb_dg_solver_options {
  int degree;               /* 1..8                                                                      */
  int n_subdivisions[3];
  int n_refinements;
  double lower[3], upper[3];
  int degree_pre, degree_post; /* Chebyshev degree (equal, multigrid_solver_dg.h:83-85); default 3       */
  int preconditioner;       /* of the DG smoother: 1 = JacobiTransformed (reference, :294-304), 0 = point Jacobi */
  int device;
  /* one process per GPU (the reference's MPI_COMM_WORLD): as in mgb_options; the DG operator and the finest FE_Q
   * level share the z-slabs (cell layers in z a multiple of comm_size, at least two per GPU) and one communicator;
   * vectors are then the local DG vectors of mgb_dg_partition() on the operator of mgb_dg_solver_parts()        */
  int comm_rank, comm_size;
  unsigned char comm_id[128];
  int basis;                /* MGB_DG_BASIS_* of the DG space (default FE_DGQ; poisson_dg: MGB_DG_BASIS_HERMITE)  */
} mgb_dg_solver_options;

void mgb_dg_solver_default_options(mgb_dg_solver_options *opt);
int mgb_dg_solver_create(const mgb_dg_solver_options *opt, mgb_dg_solver *out);    /* constructor :59-322 */
int mgb_dg_solver_destroy(mgb_dg_solver h);
int64_t mgb_dg_solver_n_dofs(mgb_dg_solver h);
/* the two parts, for inspection and for the operator-level entry points; owned by the solver */
int mgb_dg_solver_parts(mgb_dg_solver h, mgb_solver *fe_q_hierarchy, mgb_dg *dg_operator);
/* std::pair<double,double> solve_cg(tolerance) :410-424: (log(tol)/log(rate), rate); also the iteration count.
 * SolverCG with ReductionControl(100, 1e-16, tolerance), zero initial guess                                  */
int mgb_dg_solver_solve_cg(mgb_dg_solver h, double tolerance, double *iterations_estimate, double *reduction_rate,
                           unsigned int *n_iterations);
/* residual norms of the last solve_cg: initial value and one per iteration; returns how many in *count       */
int mgb_dg_solver_residual_history(mgb_dg_solver h, double *history, int capacity, int *count);
/* vmult(dst, src) :429-441: one cycle as preconditioner, double device vectors of mgb_dg_solver_n_dofs()     */
int mgb_dg_solver_vmult(mgb_dg_solver h, double *dst, const double *src);
/* vmult_with_residual_update(residual, update, factor) :445-527, the preconditioner call of a CG solver with merged
 * vector operations: one cycle on residual + factor * update; afterwards residual += factor * update, update = the
 * cycle's result, results = {update_new . residual_new, update_new . (factor * update_old)} (factor 0: both the
 * first).  Double device vectors                                                                             */
int mgb_dg_solver_vmult_with_residual_update(mgb_dg_solver h, double *residual, double *update, double factor,
                                             double results[2]);
int mgb_dg_solver_l2_error(mgb_dg_solver h, double *error);                        /* compute_l2_error :328-366 */
/* device pointers of the solver's vectors: 0 solution, 1 rhs, 2 residual (double); 3 defect_dg,
 * 4 solution_update_dg (float; the smoother swaps vectors, ask again after a cycle)              :663-693 */
int mgb_dg_solver_vector(mgb_dg_solver h, int which, void **dev_ptr);
int mgb_dg_solver_get_solution(mgb_dg_solver h, double *solution_host);            /* get_solution :399-403 */
/* info = {min_eig, max_eig, theta, delta, degree, cg_iterations} of smooth_dg                     :294-304 */
int mgb_dg_solver_smoother_info(mgb_dg_solver h, double info[6]);
int mgb_dg_solver_do_matvec(mgb_dg_solver h);                                      /* :539-543 (double)  */
int mgb_dg_solver_do_matvec_smoother(mgb_dg_solver h);                             /* :548-552 (float)   */
int mgb_dg_solver_sync(mgb_dg_solver h);
/* end-to-end call: solve_cg, then the solution copied to a host buffer (may be NULL)                       */
int mgb_dg_solver_solve_cg_host(mgb_dg_solver h, double tolerance, double *solution_host, double *iterations_estimate,
                                double *reduction_rate, unsigned int *n_iterations);
int64_t mgb_dg_solver_launch_count(mgb_dg_solver h);      /* kernels of the DG part + the FE_Q hierarchy */

/* ---- Pure DG h-multigrid: multigrid::MultigridSolverDGPlain<3,p,float,double,type>
 *                                                                  common/multigrid_solver_dg_plain.h:54-590
 * The SIP-DG operator on every level of the refinement hierarchy (level l = the coarse mesh refined l times), no
 * FE_Q space: Chebyshev over JacobiTransformed on each level (:192-213: range 20, degree degree_pre on inner levels,
 * max(1, degree_pre - 1) on the finest, 15 CG steps for the eigenvalues; level 0: Chebyshev run as solver with
 * tolerance 1e-5, its degree chosen by deal.II's rule, m() CG steps), the embedding of the DG space of a cell into
 * its eight children as transfer (MGTransferMatrixFree; restriction its transpose), V-cycle :456-497 in float under
 * CG in double (:304-316).  Driver: poisson_dg_plain/program.cc:232-330.  Options and problem data as mgb_dg_solver
 * (options.basis = the reference's template argument `type`; preconditioner is 1 in the reference).  One GPU.   */
typedef struct mgb_dg_plain_solver_t *mgb_dg_plain_solver;

int mgb_dg_plain_solver_create(const mgb_dg_solver_options *opt, mgb_dg_plain_solver *out);   /* constructor :58-215 */
int mgb_dg_plain_solver_destroy(mgb_dg_plain_solver h);
int64_t mgb_dg_plain_solver_n_dofs(mgb_dg_plain_solver h);
int mgb_dg_plain_solver_n_levels(mgb_dg_plain_solver h);
/* matrix[level] (:550), owned by the solver, for the operator-level entry points                              */
int mgb_dg_plain_solver_level_operator(mgb_dg_plain_solver h, int level, mgb_dg *dg_operator);
/* solve_cg(tolerance) :304-316, as mgb_dg_solver_solve_cg                                                      */
int mgb_dg_plain_solver_solve_cg(mgb_dg_plain_solver h, double tolerance, double *iterations_estimate,
                                 double *reduction_rate, unsigned int *n_iterations);
int mgb_dg_plain_solver_residual_history(mgb_dg_plain_solver h, double *history, int capacity, int *count);
int mgb_dg_plain_solver_vmult(mgb_dg_plain_solver h, double *dst, const double *src);          /* vmult :321-337 */
/* vmult_with_residual_update :340-427, as mgb_dg_solver_vmult_with_residual_update                             */
int mgb_dg_plain_solver_vmult_with_residual_update(mgb_dg_plain_solver h, double *residual, double *update, double factor,
                                                   double results[2]);
int mgb_dg_plain_solver_l2_error(mgb_dg_plain_solver h, double *error);             /* compute_l2_error :221-259 */
/* device pointers: 0 solution, 1 rhs, 2 residual (double, finest level; `level` ignored); 3 defect[level],
 * 4 solution_update[level], 5 t[level] (float; the smoother swaps vectors, ask again after a cycle)   :535-545 */
int mgb_dg_plain_solver_vector(mgb_dg_plain_solver h, int which, int level, void **dev_ptr);
/* the same vector copied to a host buffer of its length and type                                              */
int mgb_dg_plain_solver_get_vector(mgb_dg_plain_solver h, int which, int level, void *host);
int mgb_dg_plain_solver_get_solution(mgb_dg_plain_solver h, double *solution_host);
/* info = {min_eig, max_eig, theta, delta, degree, cg_iterations} of smooth[level]                     :192-213 */
int mgb_dg_plain_solver_smoother_info(mgb_dg_plain_solver h, int level, double info[6]);
/* the level transfer on float device vectors (transfer.prolongate / restrict_and_add, :476-486): `level` is the
 * finer of the two levels; add != 0 adds to dst instead of overwriting it                                     */
int mgb_dg_plain_solver_prolongate(mgb_dg_plain_solver h, int to_level, float *dst, const float *src, int add);
int mgb_dg_plain_solver_restrict(mgb_dg_plain_solver h, int from_level, float *dst, const float *src, int add);
int mgb_dg_plain_solver_do_matvec(mgb_dg_plain_solver h);                           /* :432-436 (double)  */
int mgb_dg_plain_solver_do_matvec_smoother(mgb_dg_plain_solver h);                  /* :441-445 (float)   */
int mgb_dg_plain_solver_sync(mgb_dg_plain_solver h);
int mgb_dg_plain_solver_solve_cg_host(mgb_dg_plain_solver h, double tolerance, double *solution_host,
                                      double *iterations_estimate, double *reduction_rate, unsigned int *n_iterations);
int64_t mgb_dg_plain_solver_launch_count(mgb_dg_plain_solver h);

/* ---- Adaptive meshes with hanging nodes (poisson_l/program.cc:232-260,465-541)
 * FE_Q(p) Laplacian on a forest of n_roots[0] x n_roots[1] x n_roots[2] root cells filling [lower, upper]
 * (root_mask[z][y][x] != 0 keeps a root cell; NULL keeps all: GridGenerator::hyper_L is a 2x2x2 grid with one
 * root removed), refined globally and then cell by cell; the mesh is kept 2:1 balanced over faces, edges and
 * corners (p4est).  Hanging-node constraints (AffineConstraints of DoFTools::make_hanging_node_constraints, with
 * chains resolved) and the homogeneous Dirichlet condition on the whole boundary are applied inside the kernel
 * while it gathers and scatters; constrained rows are identity rows (laplace_operator.h:592-593).
 * The vector holds the unconstrained and the Dirichlet DoFs; hanging-node values are not stored.
 * Mesh and numbering are host data (no GPU needed for create / refine / n_dofs / info / tables); vmult and
 * inverse_diagonal are kernels and fail without a device.                                                   */
typedef struct mgb_adaptive_t *mgb_adaptive;
int mgb_adaptive_create(int degree, const int n_roots[3], const unsigned char *root_mask, const double lower[3],
                        const double upper[3], int n_global_refinements, int device, mgb_adaptive *out);
int mgb_adaptive_destroy(mgb_adaptive h);
int64_t mgb_adaptive_n_cells(mgb_adaptive h);
/* active cells in the library's order: level[n_cells], position[n_cells][3] in units of cells of that level */
int mgb_adaptive_cells(mgb_adaptive h, int *level, int *position);
/* Triangulation::execute_coarsening_and_refinement with refine flags (poisson_l/program.cc:508-541): flags[n_cells] */
int mgb_adaptive_refine(mgb_adaptive h, const unsigned char *flags);
int64_t mgb_adaptive_n_dofs(mgb_adaptive h);
/* info = {n_dofs, n_dirichlet, hanging local nodes, constraint rows, constraint entries, finest level} */
int mgb_adaptive_info(mgb_adaptive h, int64_t info[6]);
int mgb_adaptive_dof_points(mgb_adaptive h, double *xyz /* [n_dofs][3] */);
int mgb_adaptive_dirichlet(mgb_adaptive h, unsigned char *mask /* [n_dofs] */);
/* host copies of the tables the kernel reads: cell_index[n_cells][(p+1)^3] (>= 0 DoF, -1 zero, <= -2 row -2-e),
 * row_start[rows+1], row_dof / row_w[entries]; any pointer may be NULL                                          */
int mgb_adaptive_tables(mgb_adaptive h, int32_t *cell_index, int32_t *row_start, int32_t *row_dof, double *row_w);
/* LaplaceOperator::vmult on the active mesh                                          laplace_operator.h:575-601 */
int mgb_adaptive_vmult(mgb_adaptive h, int precision, void *dst, const void *src);
/* compute_diagonal + inverse, exact (cell operator applied to every constraint column)   :747-800 */
int mgb_adaptive_inverse_diagonal(mgb_adaptive h, int precision, void *inv_diag);
int mgb_adaptive_vmult_host(mgb_adaptive h, int precision, void *dst_host, const void *src_host);
int mgb_adaptive_sync(mgb_adaptive h);
/* Local-smoothing multigrid on the forest and the CG solver of poisson_l (program.cc:283-420): level operators with the
 * refinement-edge DoFs constrained, edge matrices (MGInterfaceOperator: vmult_interface_down / _up), matrix-free
 * transfer between partially refined levels, Chebyshev smoothers (range 15, degree n_smoother; level 0: tolerance 1e-3),
 * PreconditionMG::vmult, and SolverCG in double on the operator with hanging-node constraints.
 * level_info: 0 DoFs, 1 cells, 2 refinement-edge DoFs, 3 cells at the edge, 4 DoFs copied from/to the active vector,
 * 5 constrained DoFs, 6 Chebyshev degree, 7 / 8 smallest / largest eigenvalue estimate, 9 theta, 10 delta, 11 CG steps. */
int mgb_adaptive_mg_setup(mgb_adaptive h, int n_smoother);
int mgb_adaptive_mg_n_levels(mgb_adaptive h);
int mgb_adaptive_mg_level_info(mgb_adaptive h, int level, double info[12]);
/* the level operator and the edge matrices of one level on float device vectors of that level (length: info[0] of
 * mgb_adaptive_mg_level_info): which 0 = matrix[level].vmult (refinement-edge and boundary rows identity), 1 =
 * vmult_interface_down, 2 = vmult_interface_up (MGInterfaceOperator, poisson_l/program.cc:374-383)                */
int mgb_adaptive_mg_level_vmult(mgb_adaptive h, int level, int which, float *dst_dev, const float *src_dev);
int mgb_adaptive_mg_vmult(mgb_adaptive h, double *dst_dev, const double *src_dev);
/* x = 0; CG to |r| <= rel_tol |r0| (or 1e-13 absolute); history (may be NULL): residual norms, *n_history in: capacity */
int mgb_adaptive_solve_cg(mgb_adaptive h, double *x_dev, const double *rhs_dev, double rel_tol, unsigned int *iterations,
                          double *reduction_rate, double *history, int *n_history);
int64_t mgb_adaptive_launch_count(mgb_adaptive h);

/* ---- The adaptive loop around the solver (LaplaceProblem::run, poisson_l/program.cc:497-602)
 * Vectors here carry the Dirichlet data in their boundary entries (the reference's `solution`).
 * refine_and_coarsen: Triangulation::execute_coarsening_and_refinement (:533-541) -- flagged cells are refined (2:1
 *   balance kept), then every family of eight active siblings all flagged for coarsening is replaced by its parent unless
 *   the parent would touch a cell more than one level finer; coarsen_flags may be NULL.
 * refine_and_coarsen_fixed_number: GridRefinement::refine_and_coarsen_fixed_number (:532-533; serial semantics: the
 *   int(top_fraction N) largest indicators refined, the int(bottom_fraction N) smallest coarsened, ties at the
 *   thresholds flagged), followed by refine_and_coarsen.  indicators: host array, one float per active cell.
 * face_neighbours: type[cell][6] (-1 boundary, 0 same level, 1 coarser, 2 four finer) and ids[cell][6][4], faces -x +x
 *   -y +y -z +z.   plain_table: the cell -> DoF table that keeps Dirichlet DoFs, [cell][(p+1)^3] (rows as mgb_adaptive_tables)
 * interpolate_boundary_values (:321-333): u = g on the Dirichlet DoFs, zero elsewhere; g: 0 = zero, 1 =
 *   Functions::LSingularityFunction in (x, y) (program.cc:85-135)
 * compute_residual (laplace_operator.h:806-845, program.cc:546-549): dst = (f, phi) - K src on rows without Dirichlet
 *   condition (zero on the others), f constant
 * estimate_kelly: KellyErrorEstimator<3>::estimate (:524-531) with QGauss<2>(p+1), no Neumann data, strategy
 *   cell_diameter_over_24: eta_K = sqrt(diameter_K / 24 * sum_F int_F [du/dn]^2); indicators to a host array
 * integrate_difference: VectorTools::integrate_difference (:556-575) against g; norm 0 = L2, 1 = H1 seminorm, with
 *   n_q_points Gauss points per direction; cell_squares_host (may be NULL): the cells' squared contributions
 * solve_cycle: :545-575 for the current mesh -- boundary values, right-hand side, CG with the local-smoothing V-cycle
 *   (mgb_adaptive_mg_setup(n_smoother) inside), solution = boundary values + update; stats = {CG iterations, rate, L2
 *   error with QGauss(p+2), H1-seminorm error with QGauss(p+1)}; *solution_dev: device vector owned by the handle, valid
 *   until the mesh changes                                                                                            */
int mgb_adaptive_refine_and_coarsen(mgb_adaptive h, const unsigned char *refine_flags, const unsigned char *coarsen_flags);
int mgb_adaptive_refine_and_coarsen_fixed_number(mgb_adaptive h, const float *indicators, double top_fraction,
                                                 double bottom_fraction);
int mgb_adaptive_face_neighbours(mgb_adaptive h, signed char *type, int32_t *ids);
int mgb_adaptive_plain_table(mgb_adaptive h, int32_t *cell_index_plain);
int mgb_adaptive_interpolate_boundary_values(mgb_adaptive h, int function_id, double *u_dev);
int mgb_adaptive_compute_residual(mgb_adaptive h, double *dst_dev, const double *src_dev, double f);
int mgb_adaptive_estimate_kelly(mgb_adaptive h, const double *u_dev, float *indicators_host);
int mgb_adaptive_integrate_difference(mgb_adaptive h, const double *u_dev, int function_id, int norm, int n_q_points,
                                      double *result, double *cell_squares_host);
int mgb_adaptive_solve_cycle(mgb_adaptive h, int function_id, double f, int n_smoother, double stats[4],
                             double **solution_dev);

/* The subset of LinearAlgebra::distributed::Vector<Number> the drivers use (poisson_cube/program.cc:262-380,
 * multigrid_solver.h:54-67): device vectors of one level in the precision given; sums run over the owned
 * entries of all ranks.  dot: a . b (l2_norm = sqrt(dot(v,v)));  sadd: y = sy y + ax x;  set: y = value;
 * update_ghost_values: refresh the ghost planes of a z-slab (no-op on one GPU).                       */
int mgb_vec_dot(mgb_solver s, int level, int precision, const void *a, const void *b, double *result);
int mgb_vec_sadd(mgb_solver s, int level, int precision, void *y, double sy, double ax, const void *x);
/* dst = src converted to dst_precision (src has the other precision): Vector::operator=(const Vector<OtherNumber> &)  */
int mgb_vec_convert(mgb_solver s, int level, int dst_precision, void *dst, const void *src);
int mgb_vec_set(mgb_solver s, int level, int precision, void *y, double value);
int mgb_vec_update_ghost_values(mgb_solver s, int level, int precision, void *v);
/* compress(VectorOperation::add) and zero_out_ghost_values of LinearAlgebra::distributed::Vector (the calls around
 * every cell loop that writes ghosts, laplace_operator.h:590,843): ghost planes are added to their owners on the
 * neighbouring z-slabs and then zeroed / ghost planes are zeroed.  No-ops on one GPU and on replicated levels. */
int mgb_vec_compress_add(mgb_solver s, int level, int precision, void *v);
int mgb_vec_zero_out_ghost_values(mgb_solver s, int level, int precision, void *v);

/* ---- Allocation-free dot product.  No counterpart in the reference: added for
 * callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * *result_dev = sum_i x[i] y[i] over n entries of two device vectors of the given precision, accumulated in double;
 * result_dev is a device pointer, the work is queued on `cuda_stream` (a cudaStream_t, NULL = the default stream) and the
 * call does not synchronise.  It needs no solver handle and allocates nothing: the partial sums live in a static
 * __device__ array of 512.25 KiB inside the library (held by every GPU that loads it until exit), which bounds the
 * length: n <= MGB_NOALLOC_DOT_MAX_N, larger calls fail with a message.  All calls on a GPU share that array: calls on
 * one stream are ordered, calls that overlap on the GPU (other streams, other threads) race.  The first call on a device
 * looks up the array's device address (cudaGetSymbolAddress, with that device current); n = 0 does only that -- call it
 * once before starting a capture.  Fixed summation order: bitwise reproducible.  (csrc/noalloc.cu)                     */
#define MGB_NOALLOC_DOT_MAX_N 268435456LL
int mgb_noalloc_dot(int precision, int64_t n, const void *x_dev, const void *y_dev, double *result_dev, void *cuda_stream);

/* plain host<->device helpers so that a C/C++ caller needs no CUDA runtime of its own; `s` may be NULL (vectors
 * of the DG operator): current device, synchronous */
int mgb_device_alloc(mgb_solver s, size_t bytes, void **dev_ptr);
int mgb_device_free(mgb_solver s, void *dev_ptr);
int mgb_copy_to_device(mgb_solver s, void *dst_dev, const void *src_host, size_t bytes);
int mgb_copy_to_host(mgb_solver s, void *dst_host, const void *src_dev, size_t bytes);

/* number of kernels of this library launched so far on the solver's stream */
int64_t mgb_solver_launch_count(mgb_solver s);

/* Host-only helpers (no GPU needed; used by the CPU test-suite for the host logic).
 * 1-D ingredients of FE_Q(p)/QGauss(p+1) (deal.II ShapeInfo as consumed by FEEvaluation,
 * laplace_operator.h:535): S,G,D,M1,K1 are (p+1)^2, P1 is (2p+1)(p+1), nodes/xq/wq are p+1 */
int mgb_host_basis(int p, double *S, double *G, double *D, double *M1, double *K1, double *P1,
                   double *nodes, double *xq, double *wq);
/* tile-column numbering of one box level, all faces Dirichlet (the role of extract_compressed_indices,
 * laplace_operator.h:226-353, which becomes a closed formula on box meshes); lex_to_dof may be NULL */
int mgb_host_numbering(int p, const int ncell[3], int tile_cells, int64_t *n_dofs, int *tile_cells_used,
                       int64_t *n_tiles, int64_t *n_constrained, uint32_t *lex_to_dof);
/* z-slab partition and ghost-exchange plan of `rank` of `n_ranks` on multigrid level `level` of a box with
 * n_subdivisions coarse cells: info[12] as mgb_solver_partition(); plan[8] = {offset, count} (in values) of
 * send_up, recv_dn, send_dn, recv_up: update_ghost_values sends send_up to rank+1 (its recv_dn) and send_dn to
 * rank-1 (its recv_up); compress(add) sends recv_up to rank+1, which adds it to its send_dn range.
 * min_dofs: levels with fewer DoFs are held completely by every rank (the solver's default is 3 000 000,
 * environment MGB200_MIN_SLAB_DOFS).
 * lex_to_dof (may be NULL): local vector index of each point of the LOCAL box in lexicographic order     */
int mgb_host_partition(int p, const int n_subdivisions[3], int level, int tile_cells, int n_ranks, int rank,
                       int64_t min_dofs, int64_t info[12], int64_t plan[8], uint32_t *lex_to_dof);

#ifdef __cplusplus
}
#endif
#endif
