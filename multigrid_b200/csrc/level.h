// Level description shared by host and device: the renumbered cell-to-DoF map.
//
// The reference stores, per cell, 3^dim "compressed" entity start indices (8 vertices,
// 12 lines, 6 quads, 1 hex; DoFs inside an entity contiguous) -- common/laplace_operator.h:
// 226-353, consumed by common/vector_access_reduced.h:11-251,255-505.  Here the same
// 27-entry scheme is applied to a *patch* of M x M x M cells (M = 4, 2 or 1): a patch is
// treated as one macro element with (M p - 1)^3 interior, 6 x (M p - 1)^2 face, 12 x (M p - 1)
// edge and 8 vertex DoFs.  Numbering: all patch interiors first (contiguous per patch,
// lexicographic), then patch faces, edges, vertices.  A thread block gathers a whole patch
// with coalesced loads; DoFs interior to a patch are touched by exactly one block, so
// fused vector updates on them need no second pass over HBM.  M = 1 is exactly the
// reference layout.  Constrained (Dirichlet) entities are flagged per patch in a 27-bit
// mask: read as 0, never written (vector_access_reduced.h:174-179,431-433).
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
    int      p, n;       // degree, p+1
    int      M;          // cells per patch and direction
    int      np;         // points per patch and direction = M p + 1
    int      q;          // interior points per patch and direction = M p - 1
    int      npatch[3];  // patches per direction
    int      ncell[3];   // cells per direction
    int      nd[3];      // DoFs per direction
    uint32_t n_patches, n_cells, n_dofs;
    uint32_t n_interior; // DoFs interior to patches = n_patches q^3 (they come first)
    double   lo[3], h[3];
    const uint32_t *ent_base; // [n_patches][27] first DoF of every patch entity
    const uint32_t *ent_mask; // [n_patches]     bit e set: entity e is Dirichlet-constrained
  };

  // entity number (ex + 3 ey + 9 ez, e_d in {0 low, 1 interior, 2 high}) and offset inside
  // the entity of patch point (I,J,K)
  MGB_HD void
  patch_point_entity(const int np, const int q, const int I, const int J, const int K, int &e, int &off)
  {
    const int ex = I == 0 ? 0 : (I == np - 1 ? 2 : 1);
    const int ey = J == 0 ? 0 : (J == np - 1 ? 2 : 1);
    const int ez = K == 0 ? 0 : (K == np - 1 ? 2 : 1);
    e            = ex + 3 * ey + 9 * ez;
    const int sx = ex == 1 ? q : 1, sy = ey == 1 ? q : 1;
    off          = (ex == 1 ? I - 1 : 0) + sx * ((ey == 1 ? J - 1 : 0) + sy * (ez == 1 ? K - 1 : 0));
  }

  MGB_HD uint32_t
  patch_dof(const LevelDesc &L, const uint32_t patch, const int I, const int J, const int K, bool &constrained)
  {
    int e, off;
    patch_point_entity(L.np, L.q, I, J, K, e, off);
    constrained = (L.ent_mask[patch] >> e) & 1u;
    return L.ent_base[patch * 27 + e] + (uint32_t)off;
  }

  // cells are numbered patch-major: cell = patch * M^3 + (lz M + ly) M + lx
  MGB_HD uint32_t
  cell_dof(const LevelDesc &L, const uint32_t cell, const int i, const int j, const int k, bool &constrained)
  {
    const int      M3    = L.M * L.M * L.M;
    const uint32_t patch = cell / M3;
    const int      lc    = cell - patch * M3;
    const int      lx = lc % L.M, ly = (lc / L.M) % L.M, lz = lc / (L.M * L.M);
    return patch_dof(L, patch, lx * L.p + i, ly * L.p + j, lz * L.p + k, constrained);
  }

  // global cell coordinates of a patch-major cell number
  MGB_HD void
  cell_coords(const LevelDesc &L, const uint32_t cell, int &cx, int &cy, int &cz)
  {
    const int      M3    = L.M * L.M * L.M;
    const uint32_t patch = cell / M3;
    const int      lc    = cell - patch * M3;
    const int      px = patch % L.npatch[0], py = (patch / L.npatch[0]) % L.npatch[1],
              pz = patch / (L.npatch[0] * L.npatch[1]);
    cx = px * L.M + lc % L.M;
    cy = py * L.M + (lc / L.M) % L.M;
    cz = pz * L.M + lc / (L.M * L.M);
  }

  // patch-major cell number of global cell coordinates
  MGB_HD uint32_t
  cell_number(const LevelDesc &L, const int cx, const int cy, const int cz)
  {
    const uint32_t patch =
      ((uint32_t)(cz / L.M) * L.npatch[1] + (cy / L.M)) * L.npatch[0] + (cx / L.M);
    return patch * (L.M * L.M * L.M) + ((cz % L.M) * L.M + (cy % L.M)) * L.M + (cx % L.M);
  }
} // namespace mgb
