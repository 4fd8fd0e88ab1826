// Host-side construction of the tile-column numbering for refined box meshes
// (the role of DoFHandler::distribute_mg_dofs + DoFRenumbering + extract_compressed_indices in the
// reference: poisson_cube/program.cc:208-251, common/laplace_operator.h:226-353).
// Pure host code (no CUDA): exercised by the CPU tests through mgb_host_numbering().
#pragma once
#include <cstdint>
#include <stdexcept>
#include <vector>

#include "level.h"

namespace mgb
{
  struct HostLevel
  {
    LevelDesc             d;
    std::vector<uint32_t> constrained; // sorted list of Dirichlet DoFs
  };

  // cells per tile and direction: tiles of about 16 x 16 points (one thread per point column)
  inline int
  choose_tile_cells(const int p, const int requested)
  {
    static const int table[9] = {0, 16, 8, 5, 4, 3, 2, 2, 2};
    int              cw       = table[p];
    if (requested > 0 && requested < cw)
      cw = requested;
    return cw;
  }

  inline void
  build_box_level(const int      p,
                  const int      ncell[3],
                  const int      tile_cells,
                  const double   lo[3],
                  const double   hi[3],
                  const uint32_t dirichlet,
                  HostLevel     &out)
  {
    LevelDesc &d   = out.d;
    d.p            = p;
    d.n            = p + 1;
    d.CW           = tile_cells;
    d.TW           = tile_cells * p;
    d.dirichlet    = dirichlet;
    uint64_t ndofs = 1, ncells = 1;
    for (int c = 0; c < 3; ++c)
      {
        if (ncell[c] < 1)
          throw std::runtime_error("a level needs at least one cell per direction");
        d.ncell[c] = ncell[c];
        d.nd[c]    = ncell[c] * p + 1;
        d.lo[c]    = lo[c];
        d.h[c]     = (hi[c] - lo[c]) / ncell[c];
        ndofs *= d.nd[c];
        ncells *= ncell[c];
      }
    if (ndofs >= 0xffffffffull)
      throw std::runtime_error("level exceeds 2^32 DoFs per device");
    for (int c = 0; c < 2; ++c)
      d.nt[c] = (d.nd[c] + d.TW - 1) / d.TW;
    d.n_cells = (uint32_t)ncells;
    d.n_dofs  = (uint32_t)ndofs;

    // Dirichlet DoFs: all points of the flagged faces
    out.constrained.clear();
    for (int f = 0; f < 6; ++f)
      {
        if (!((dirichlet >> f) & 1u))
          continue;
        const int dn = f / 2, a = (dn + 1) % 3, b = (dn + 2) % 3;
        int       g[3];
        g[dn] = (f & 1) ? d.nd[dn] - 1 : 0;
        for (int v = 0; v < d.nd[b]; ++v)
          for (int u = 0; u < d.nd[a]; ++u)
            {
              g[a] = u;
              g[b] = v;
              out.constrained.push_back(dof_index(d, g[0], g[1], g[2]));
            }
      }
    std::sort(out.constrained.begin(), out.constrained.end());
    out.constrained.erase(std::unique(out.constrained.begin(), out.constrained.end()),
                          out.constrained.end());
  }

  // lex_to_dof[lexicographic index] = library DoF index
  inline void
  build_lex_map(const HostLevel &lvl, uint32_t *lex_to_dof)
  {
    const LevelDesc &d = lvl.d;
    size_t           l = 0;
    for (int gk = 0; gk < d.nd[2]; ++gk)
      for (int gj = 0; gj < d.nd[1]; ++gj)
        for (int gi = 0; gi < d.nd[0]; ++gi)
          lex_to_dof[l++] = dof_index(d, gi, gj, gk);
  }
} // namespace mgb
