// Geometry of the structured mesh: the map F from the unit box of mesh coordinates (a, b, c) in [0,1]^3 to physical
// space.  The cells are the images of the uniform n_x x n_y x n_z subdivision of the unit box; inside a cell the
// geometry is the degree-p polynomial through the images of the cell's Gauss-Lobatto support points, which is what
// deal.II's MappingQGeneric(p) builds from a manifold description (multigrid_solver.h:138-139: MappingQGeneric of
// degree min(fe_degree, 10)) -- Jacobians and quadrature points come from that polynomial, not from F itself.
//
//   MGB_GEOMETRY_BOX           x = lower + (upper - lower) o (a, b, c): axis-aligned cells, the fast Cartesian path
//   MGB_GEOMETRY_SHELL_SECTOR  one cell of GridGenerator::hyper_shell(center 0, r0, r1, 6) (poisson_shell/program.cc:
//                              418-431: a cube face blown up radially between the spheres r0 and r1) and its uniform
//                              refinements: direction = normalize(tan(alpha) (2a-1), tan(alpha) (2b-1), 1), radius
//                              r0 + (r1 - r0) c, alpha = opening half-angle (pi/4 = a whole cube face)
//   MGB_GEOMETRY_AFFINE        x = lower + A ((upper - lower) o (a, b, c)) with a full 3 x 3 matrix A (sheared cells
//                              with a constant but non-diagonal metric, as matvec_dg/program.cc:55-77)
#pragma once

namespace mgb
{
  struct Geometry
  {
    int    kind;
    double lo[3], hi[3];
    double par[12]; // SHELL_SECTOR: r0, r1, alpha;  AFFINE: A row-major in par[0..8]
  };

  __host__ __device__ inline void
  geometry_point(const Geometry &g, const double a, const double b, const double c, double x[3])
  {
    if (g.kind == 1)
      {
        const double ta = tan(g.par[2]);
        const double s = ta * (2. * a - 1.), t = ta * (2. * b - 1.);
        const double inv = 1. / sqrt(s * s + t * t + 1.);
        const double r   = g.par[0] + (g.par[1] - g.par[0]) * c;
        x[0]             = r * s * inv;
        x[1]             = r * t * inv;
        x[2]             = r * inv;
        return;
      }
    const double u[3] = {(g.hi[0] - g.lo[0]) * a, (g.hi[1] - g.lo[1]) * b, (g.hi[2] - g.lo[2]) * c};
    if (g.kind == 2)
      {
        for (int d = 0; d < 3; ++d)
          x[d] = g.lo[d] + g.par[3 * d] * u[0] + g.par[3 * d + 1] * u[1] + g.par[3 * d + 2] * u[2];
        return;
      }
    for (int d = 0; d < 3; ++d)
      x[d] = g.lo[d] + u[d];
  }
} // namespace mgb
