// Level description shared by host and device: the renumbered cell-to-DoF map.
//
// The reference renumbers DoFs for matrix-free data locality
// (DoFRenumbering::compute_matrix_free_data_locality, poisson_cube/program.cc:229-246) and stores,
// per cell, 3^dim "compressed" entity start indices (common/laplace_operator.h:226-353, consumed by
// common/vector_access_reduced.h:11-251,255-505).  On the uniformly refined box meshes of this path
// the map is a closed formula, so no index table is stored or streamed at all:
//
//   the (nd0 x nd1 x nd2) grid of DoFs (nd = ncell * p + 1) is cut into columns of TW x TW points
//   (TW = CW * p, CW cells per tile and direction); tile (tx,ty) owns the points
//   [TW tx, TW tx + TW) x [TW ty, TW ty + TW) x [0, nd2) (the last tile in a direction is narrower,
//   possibly one point wide).  The numbering is z-plane major over the whole level, tile by tile
//   inside a plane:
//
//       index = gk * (nd0 nd1) + offset(tx,ty) + ly * wx + lx ,   offset = ty TW nd0 + tx TW wy
//
// so every z-plane of a tile is one contiguous chunk of wx wy values, and a slab of z-planes is one
// contiguous range of the vector: the ghost planes of a z-slab decomposition over several GPUs sit at
// the two ends of the local vector and the owned DoFs in between (see `zg` below).  A thread
// block streams a tile column plane by plane with coalesced loads and owns the result at all of
// its points (see cartesian_kernel.cuh): no scatter, no atomics.  Dirichlet DoFs are the points on the
// flagged faces of the box: read as 0, never written by the cell loop
// (vector_access_reduced.h:174-179,431-433).
#pragma once
#include <cstdint>

#if defined(__CUDACC__)
#define MGB_HD __host__ __device__ __forceinline__
#else
#define MGB_HD inline
#endif

namespace mgb
{
  struct LevelDesc
  {
    int      p, n;      // degree, p+1
    int      CW;        // cells per tile and direction (x and y)
    int      TW;        // points per tile and direction = CW p
    int      nt[2];     // tiles in x and y
    int      ncell[3];  // cells per direction
    int      nd[3];     // DoFs per direction
    uint32_t n_cells, n_dofs;
    uint32_t dirichlet; // bit 2d: low face of direction d is Dirichlet, bit 2d+1: high face
    double   lo[3], h[3];
    // z-slab of a level distributed over several GPUs (all zero on one GPU and on replicated levels):
    // ncell[2], nd[2], n_dofs, lo[2] and the z bits of `dirichlet` describe the LOCAL box, which is the
    // owned cell layers plus `zg` ghost layers below; the planes [0, zg p) are copies of the last planes
    // of the rank below, and with `ztop` the plane nd[2]-1 is a copy of the first owned plane of the rank
    // above (the role of the ghost range of LinearAlgebra::distributed::Vector in the reference)
    int      zg;      // ghost cell layers below (0 or 1)
    int      ztop;    // 1: the top plane is a ghost
    int      gk0;     // global index of the local plane 0
    int      nd2_all; // planes of the whole level
  };

  // first owned plane, one past the last owned plane
  MGB_HD int
  owned_plane_begin(const LevelDesc &L)
  {
    return L.zg * L.p;
  }
  MGB_HD int
  owned_plane_end(const LevelDesc &L)
  {
    return L.nd[2] - L.ztop;
  }

  MGB_HD int
  imin(const int a, const int b)
  {
    return a < b ? a : b;
  }

  // width of tile t in direction d (0 or 1)
  MGB_HD int
  tile_width(const LevelDesc &L, const int d, const int t)
  {
    return imin(L.TW, L.nd[d] - t * L.TW);
  }

  // distance between two z-planes of the level
  MGB_HD uint32_t
  plane_pitch(const LevelDesc &L)
  {
    return (uint32_t)L.nd[0] * (uint32_t)L.nd[1];
  }

  // index of the first point of tile (tx,ty) on plane 0
  MGB_HD uint32_t
  tile_base(const LevelDesc &L, const int tx, const int ty)
  {
    const int wy = tile_width(L, 1, ty);
    return (uint32_t)ty * (uint32_t)(L.TW * L.nd[0]) + (uint32_t)tx * (uint32_t)(L.TW * wy);
  }

  // library index of the DoF at grid point (gi,gj,gk)
  MGB_HD uint32_t
  dof_index(const LevelDesc &L, const int gi, const int gj, const int gk)
  {
    const int tx = gi / L.TW, ty = gj / L.TW;
    const int lx = gi - tx * L.TW, ly = gj - ty * L.TW;
    const int wx = tile_width(L, 0, tx);
    return (uint32_t)gk * plane_pitch(L) + tile_base(L, tx, ty) + (uint32_t)(ly * wx + lx);
  }

  MGB_HD bool
  is_constrained(const LevelDesc &L, const int gi, const int gj, const int gk)
  {
    const uint32_t m = L.dirichlet;
    return (gi == 0 && (m & 1u)) || (gi == L.nd[0] - 1 && (m & 2u)) || (gj == 0 && (m & 4u)) ||
           (gj == L.nd[1] - 1 && (m & 8u)) || (gk == 0 && (m & 16u)) || (gk == L.nd[2] - 1 && (m & 32u));
  }

  // constrained through a Dirichlet face in x or y (independent of the plane)
  MGB_HD bool
  is_constrained_xy(const LevelDesc &L, const int gi, const int gj)
  {
    const uint32_t m = L.dirichlet;
    return (gi == 0 && (m & 1u)) || (gi == L.nd[0] - 1 && (m & 2u)) || (gj == 0 && (m & 4u)) ||
           (gj == L.nd[1] - 1 && (m & 8u));
  }

  // cells are numbered lexicographically, x fastest
  MGB_HD void
  cell_coords(const LevelDesc &L, const uint32_t cell, int &cx, int &cy, int &cz)
  {
    cx               = cell % L.ncell[0];
    const uint32_t r = cell / L.ncell[0];
    cy               = r % L.ncell[1];
    cz               = r / L.ncell[1];
  }

  MGB_HD uint32_t
  cell_number(const LevelDesc &L, const int cx, const int cy, const int cz)
  {
    return ((uint32_t)cz * L.ncell[1] + cy) * L.ncell[0] + cx;
  }

  // DoF (i,j,k) of a cell, and whether it is Dirichlet-constrained
  MGB_HD uint32_t
  cell_dof(const LevelDesc &L, const uint32_t cell, const int i, const int j, const int k, bool &constrained)
  {
    int cx, cy, cz;
    cell_coords(L, cell, cx, cy, cz);
    const int gi = cx * L.p + i, gj = cy * L.p + j, gk = cz * L.p + k;
    constrained  = is_constrained(L, gi, gj, gk);
    return dof_index(L, gi, gj, gk);
  }
} // namespace mgb
