// What a driver of the reference looks like after switching to the B200 library: the calls of
// poisson_cube/program.cc:262-380 (solve, solve_cg, matvec timing, L2 error) on multigrid::MultigridSolver, a CG
// iteration by hand on multigrid::LaplaceOperator / DeviceVector, and the DG operator of matvec_dg/program.cc.
// Built by tests/test_cpp_adapter.py with g++ (no nvcc, no CUDA headers): only include/ and libmgb200.so.
#include <cstdio>
#include <cstdlib>

#include <cmath>
#include "multigrid/laplace_operator_adaptive.h"
#include "multigrid/laplace_operator_dg.h"
#include "multigrid/minimal_surface.h"
#include "multigrid/multigrid_solver_dg.h"
#include "multigrid/multigrid_solver_dg_plain.h"
#include "multigrid/multigrid_solver.h"

// A third problem, defined only here (the Function<dim> objects a driver of the reference would write):
// -div(c grad u) = f with u = exp(x) sin(2y) cos(z), c = 1 + x^2 + sin(3y) / 2, on the curved shell sector.
static double
third_solution(const double x, const double y, const double z)
{
  return std::exp(x) * std::sin(2 * y) * std::cos(z);
}
static double
third_coefficient(const double x, const double y, const double)
{
  return 1 + x * x + 0.5 * std::sin(3 * y);
}
static double
third_rhs(const double x, const double y, const double z)
{
  const double u = third_solution(x, y, z), ux = u, uy = 2 * std::exp(x) * std::cos(2 * y) * std::cos(z);
  return -(third_coefficient(x, y, z) * (ux - 5 * u) + 2 * x * ux + 1.5 * std::cos(3 * y) * uy);
}

int
main(int argc, char **argv)
{
  try
    {
      constexpr int               degree = 4;
      multigrid::MeshDescription mesh;
      mesh.n_refinements = argc > 1 ? std::atoi(argv[1]) : 3;
      multigrid::MultigridSolver<3, degree, float, double> solver(mesh, 3, 3, 2);
      const double                                         rate = solver.solve(false);
      const double                                         e_fmg = solver.compute_l2_error();
      const std::pair<unsigned int, double>                cg = solver.solve_cg();
      const double                                         e_cg = solver.compute_l2_error();
      std::printf("fmg_reduction %.6e fmg_l2 %.6e cg_its %u cg_reduction %.6e cg_l2 %.6e\n", rate, e_fmg, cg.first, cg.second, e_cg);

      // operator concept: q = A p, <p, q> and a residual through the level operator
      auto                         A = solver.get_matrix_dp();
      multigrid::DeviceVector<double> p, q, r;
      A.initialize_dof_vector(p);
      A.initialize_dof_vector(q);
      A.initialize_dof_vector(r);
      p = 1.0;
      A.vmult(q, p);
      const double pq = p * q;
      A.vmult_residual(q, p, r); // r = q - A p = 0
      std::printf("operator m %zu pAp %.12e residual_norm %.3e\n", A.m(), pq, (double)r.l2_norm());

      // the rest of the operator concept: diagonal entry, precision-converting assignment, interface matrices (zero on a
      // globally refined mesh), compute_residual with a right-hand side function of the caller
      {
        const double                 d0 = A.el(A.m() / 2, A.m() / 2);
        multigrid::DeviceVector<float> pf;
        pf = p; // double -> float
        multigrid::DeviceVector<double> back;
        back = pf;
        back.add(-1.0, p);
        A.vmult_interface_down(q, p);
        const double                 idn = q.l2_norm();
        multigrid::DeviceVector<double> zero, res;
        A.initialize_dof_vector(zero);
        A.initialize_dof_vector(res);
        zero = 0.0;
        A.compute_residual(res, zero, [](const double x, const double y, const double z) { return 1. + x * y - z; });
        const double                 load = res * p; // sum of (f, phi_i) over the interior rows (p = 1)
        std::printf("el_mid %.12e convert_roundtrip %.3e interface_down_norm %.3e load_sum %.12e diag_inv_norm %.12e\n", d0,
                    (double)back.l2_norm(), idn, load, (double)A.get_matrix_diagonal_inverse().l2_norm());
      }

      // the third problem on one coarse cell of hyper_shell(0, 0.5, 1, 6) with 2 refinements (curved cells, six-entry
      // coefficient), handed over as samples at the library's quadrature / support points
      {
        multigrid::MeshDescription shell;
        shell.n_refinements = 2;
        shell.geometry      = MGB_GEOMETRY_SHELL_SECTOR;
        multigrid::MultigridSolver<3, degree, float, double> third(shell, 3, 3, 2);
        double (*c)(double, double, double) = third_coefficient, (*f)(double, double, double) = third_rhs,
                                   (*g)(double, double, double) = third_solution;
        third.set_problem(&c, &f, &g);
        const std::pair<unsigned int, double> cg3 = third.solve_cg();
        std::printf("third_cg_its %u third_cg_rate %.6e third_l2 %.6e\n", cg3.first, cg3.second, third.compute_l2_error());
      }

      // DG operator on the mesh of matvec_dg/program.cc, 2 refinements
      mgb_dg_options o;
      mgb_dg_default_options(&o);
      o.n_refinements = 2;
      multigrid::LaplaceOperatorCompactCombine<3, degree, double, 1> dg;
      dg.reinit(o);
      multigrid::DGVector<double> x, y;
      dg.initialize_dof_vector(x);
      dg.initialize_dof_vector(y);
      std::vector<double> hx(dg.m(), 1.0);
      x.import_from_host(hx);
      dg.vmult(y, x);
      dg.synchronize();
      const std::vector<double> hy = y.export_to_host();
      double                    xAx = 0;
      for (std::size_t i = 0; i < hy.size(); ++i)
        xAx += hx[i] * hy[i];
      std::printf("dg m %zu penalty_x %.6e xAx %.12e\n", dg.m(), (double)dg.get_penalty(0, 0), xAx);
      dg.set_preconditioner(dg.jacobi_transformed);
      dg.preconditioner_vmult(y, x); // y = P^-1 x with JacobiTransformed
      dg.synchronize();
      double xPx = 0;
      {
        const std::vector<double> hp = y.export_to_host();
        for (std::size_t i = 0; i < hp.size(); ++i)
          xPx += hx[i] * hp[i];
      }
      std::printf("xPx %.12e\n", xPx);
      {
        // the same operator with the reference's default template argument (type 0, FE_DGQHermite) and on the Gauss
        // basis (type 2): z^T A z of one coefficient vector (not a constant function) in each basis
        multigrid::LaplaceOperatorCompactCombine<3, degree, double> dg_hermite;
        multigrid::LaplaceOperatorCompactCombine<3, degree, double, 2> dg_gauss;
        dg_hermite.reinit(o);
        dg_gauss.reinit(o);
        std::vector<double> hz(dg.m());
        for (std::size_t i = 0; i < hz.size(); ++i)
          hz[i] = 0.1 + (i % 7) / 7.;
        multigrid::DGVector<double> z;
        dg.initialize_dof_vector(z);
        z.import_from_host(hz);
        double e[2] = {0, 0};
        dg_hermite.vmult(y, z);
        dg_hermite.synchronize();
        {
          const std::vector<double> h = y.export_to_host();
          for (std::size_t i = 0; i < h.size(); ++i)
            e[0] += hz[i] * h[i];
        }
        dg_gauss.vmult(y, z);
        dg_gauss.synchronize();
        {
          const std::vector<double> h = y.export_to_host();
          for (std::size_t i = 0; i < h.size(); ++i)
            e[1] += hz[i] * h[i];
        }
        std::printf("xAx_hermite %.12e xAx_gauss %.12e\n", e[0], e[1]);
      }
      if (argc > 99) // compiled, not run: the transfer needs an FE_Q level on the same (unsheared) mesh
        {
          dg.prolongate_add_cg_to_dg(y, p);
          dg.vmult_residual_and_restrict_to_cg(x, y, q);
        }

      // operator on an adaptively refined hyper_L (poisson_l/program.cc:465-541): one global refinement, then the cells
      // next to the re-entrant edge x = y = 0 once more; x^T A x of the all-ones vector counts the boundary layer
      {
        const int           roots[3] = {2, 2, 2};
        const unsigned char mask[8]  = {1, 1, 1, 1, 1, 1, 1, 0};
        const double        lo[3] = {-1, -1, -1}, hi[3] = {1, 1, 1};
        multigrid::AdaptiveLaplaceOperator<3, 2, double> ad;
        ad.reinit(roots, mask, lo, hi, 1);
        std::vector<int> lev, pos;
        ad.get_cells(lev, pos);
        std::vector<unsigned char> flags(lev.size(), 0);
        for (std::size_t c = 0; c < lev.size(); ++c)
          {
            const double hc = 1. / (1 << lev[c]), cx = -1 + (pos[3 * c] + 0.5) * hc, cy = -1 + (pos[3 * c + 1] + 0.5) * hc;
            flags[c]        = cx * cx + cy * cy < 0.55 * 0.55;
          }
        ad.execute_refinement(flags);
        multigrid::DGVector<double> ax, ay;
        ad.initialize_dof_vector(ax);
        ad.initialize_dof_vector(ay);
        std::vector<double> ones(ad.m(), 1.0);
        ax.import_from_host(ones);
        ad.vmult(ay, ax);
        ad.synchronize();
        const std::vector<double> r = ay.export_to_host();
        double                    e = 0;
        for (double v : r)
          e += v;
        std::printf("adaptive_cells %zu adaptive_m %zu adaptive_xAx %.12e\n", ad.n_active_cells(), ad.m(), e);
        // the solve of poisson_l/program.cc:337-420 on this mesh: A x = A 1 by CG with the local-smoothing V-cycle
        // (right-hand side zero on the Dirichlet rows, as an assembled one is: the solution is 1 inside, 0 on the boundary)
        ad.setup_multigrid(3);
        const std::vector<unsigned char> bnd = ad.dirichlet_mask();
        std::vector<double>              rhs = r;
        for (std::size_t i = 0; i < rhs.size(); ++i)
          if (bnd[i])
            rhs[i] = 0;
        multigrid::DGVector<double> sol, b;
        ad.initialize_dof_vector(sol);
        ad.initialize_dof_vector(b);
        b.import_from_host(rhs);
        const std::pair<unsigned int, double> cg = ad.solve_cg(sol, b);
        ad.synchronize();
        const std::vector<double> xs = sol.export_to_host();
        double                    err = 0;
        for (std::size_t i = 0; i < xs.size(); ++i)
          err = std::max(err, std::fabs(xs[i] - (bnd[i] ? 0.0 : 1.0)));
        std::printf("adaptive_levels %u adaptive_cg_its %u adaptive_cg_rate %.6e adaptive_solution_error %.3e\n", ad.n_levels(),
                    cg.first, cg.second, err);
      }

      // the adaptive loop of poisson_l/program.cc:497-602 in its default case: the L-shaped domain extruded in z, exact
      // solution r^(2/3) sin(2 phi / 3) as Dirichlet data, Kelly estimator, 30 % refined / 3 % coarsened per cycle
      {
        const int           roots[3] = {2, 2, 1};
        const unsigned char mask[4]  = {1, 1, 1, 0};
        const double        lo[3] = {-1, -1, 0}, hi[3] = {1, 1, 1};
        multigrid::AdaptiveLaplaceOperator<3, 2, double> lp;
        lp.reinit(roots, mask, lo, hi, 1);
        const auto table = lp.run(4, 1000000, 3, false);
        for (std::size_t c = 0; c < table.size(); ++c)
          std::printf("loop%zu_cells %zu loop%zu_dofs %zu loop%zu_its %u loop%zu_l2 %.6e loop%zu_h1 %.6e\n", c, table[c].n_cells, c,
                      table[c].n_dofs, c, table[c].iterations, c, table[c].l2_error, c, table[c].h1_error);
      }

      // DG multigrid solver as poisson_dg/program.cc:227-263: [-1,1]^3, 2 refinements, tolerance 1e-3
      multigrid::MeshDescription dmesh;
      dmesh.n_refinements = 2;
      for (int d = 0; d < 3; ++d)
        dmesh.lower[d] = -1.;
      multigrid::MultigridSolverDG<3, degree, float, double> dg_solver(dmesh, 3, 3, 1);
      const std::pair<double, double>                         dg_cg = dg_solver.solve_cg(1e-3);
      std::printf("dg_cg_its %u dg_cg_rate %.6e dg_l2 %.6e\n", dg_solver.last_step(), dg_cg.second,
                  dg_solver.compute_l2_error());

      // the minimal surface problem of minimal_surface/program.cc on [-1,1]^3 with 2 refinements: Scherk's surface
      // log(cos x / cos y) solves the equation, so the Newton iteration started from its boundary values must find it
      {
        multigrid::MeshDescription ms_mesh;
        ms_mesh.n_refinements = 2;
        for (int d = 0; d < 3; ++d)
          ms_mesh.lower[d] = -1.;
        multigrid::MinimalSurfaceProblem<3, 3> minimal(ms_mesh);
        const auto scherk = [](const double x, const double y, const double) { return std::log(std::cos(x) / std::cos(y)); };
        minimal.interpolate_boundary_values(scherk);
        const auto                steps = minimal.run(20, 1e-11);
        const std::vector<double> u     = minimal.get_solution();
        const std::vector<double> xyz   = minimal.solver().support_points((int)minimal.solver().n_levels() - 1);
        double                    err   = 0;
        unsigned int              its   = 0;
        for (std::size_t i = 0; i < u.size(); ++i)
          err = std::max(err, std::fabs(u[i] - scherk(xyz[3 * i], xyz[3 * i + 1], xyz[3 * i + 2])));
        for (const auto &st : steps)
          its += st.linear_iterations;
        std::printf("minimal_steps %zu minimal_linear_its %u minimal_residual %.3e minimal_error %.6e\n", steps.size(), its,
                    steps.back().final_residual_norm, err);
      }

      // pure DG h-multigrid as poisson_dg_plain/program.cc:232-263 on the same mesh, FE_DGQArbitraryNodes (type 2); one
      // preconditioned CG step by hand through vmult_with_residual_update (residual r = b - A 0 = b is not available
      // to a driver, so the step runs on r = 1: update = M^-1 r, results[0] = r . M^-1 r > 0)
      multigrid::MultigridSolverDGPlain<3, degree, float, double, 2> plain(dmesh, 3, 3, 1);
      const std::pair<double, double>                                 plain_cg = plain.solve_cg(1e-3);
      std::printf("plain_levels %u plain_cg_its %u plain_cg_rate %.6e plain_l2 %.6e\n", plain.n_levels(), plain.last_step(),
                  plain_cg.second, plain.compute_l2_error());
      {
        multigrid::DGVector<double> res, upd;
        dg.initialize_dof_vector(res);
        dg.initialize_dof_vector(upd);
        std::vector<double> ones(plain.m(), 1.0), zeros(plain.m(), 0.0);
        res.import_from_host(ones);
        upd.import_from_host(zeros);
        const std::array<double, 2> ip = plain.vmult_with_residual_update(res, upd, 0.);
        plain.synchronize();
        const std::vector<double> m = upd.export_to_host();
        double                    s = 0;
        for (double v : m)
          s += v;
        std::printf("plain_rMr %.10e plain_rMr_check %.10e\n", ip[0], s);
      }
    }
  catch (const std::exception &e)
    {
      std::fprintf(stderr, "Exception: %s\n", e.what());
      return 1;
    }
  return 0;
}
