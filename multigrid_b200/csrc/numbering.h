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

  // z-slab of one rank on one level: global cell layers [c0, c1) of `ncz_all`
  struct Slab
  {
    bool distributed = false;
    int  rank = 0, n_ranks = 1;
    int  ncz_all = 0, c0 = 0, c1 = 0;
    int  zg = 0, ztop = 0; // ghost layer below / ghost plane on top
  };

  // Level l has n0 2^l cell layers in z.  It is distributed when it has at least `min_dofs` DoFs (below
  // that a ghost exchange per operator apply costs more than computing the whole level on every rank),
  // every rank gets the same number (>= min_layers) of whole layers AND the slabs nest with the level below (the children of a rank's layers are its
  // layers on the finer level): that needs the level below to be distributed too, or an even number of
  // layers per rank.  All other levels are held completely by every rank (the coarse levels: "the
  // coarse-grid solve stays on a single GPU", here computed redundantly so that the correction needs
  // no broadcast).
  inline Slab
  make_slab(const int  p,
            const int  nsub[3],
            const int  level,
            const int  n_ranks,
            const int  rank,
            const int  min_layers = 2,
            const long min_dofs   = 3000000)
  {
    const int n0   = nsub[2];
    bool      dist = false;
    for (int l = 0; l <= level; ++l)
      {
        const long ncz = (long)n0 << l;
        long       n   = 1;
        for (int c = 0; c < 3; ++c)
          n *= ((long)nsub[c] << l) * p + 1;
        dist = n_ranks > 1 && ncz % n_ranks == 0 && ncz / n_ranks >= min_layers &&
               (dist || (n >= min_dofs && (l == 0 || (ncz / n_ranks) % 2 == 0)));
      }
    Slab s;
    s.rank    = rank;
    s.n_ranks = n_ranks;
    s.ncz_all = n0 << level;
    s.c0      = 0;
    s.c1      = s.ncz_all;
    if (dist)
      {
        const int lz  = s.ncz_all / n_ranks;
        s.distributed = true;
        s.c0          = rank * lz;
        s.c1          = s.c0 + lz;
        s.zg          = rank > 0 ? 1 : 0;
        s.ztop        = rank + 1 < n_ranks ? 1 : 0;
      }
    return s;
  }

  // contiguous index ranges (in values) of the ghost exchange of a distributed level:
  //   update_ghost_values: send_up -> rank+1's recv_dn ; send_dn -> rank-1's recv_up
  //   compress(add) of the top ghost plane: recv_up -> added into rank+1's send_dn range
  struct ExchangePlan
  {
    uint64_t send_up[2] = {0, 0}, recv_dn[2] = {0, 0}, send_dn[2] = {0, 0}, recv_up[2] = {0, 0}; // {offset, count}
    uint64_t owned[2]   = {0, 0};
  };

  inline ExchangePlan
  make_exchange_plan(const LevelDesc &d)
  {
    ExchangePlan   e;
    const uint64_t pitch = (uint64_t)d.nd[0] * d.nd[1];
    const int      b = owned_plane_begin(d), t = owned_plane_end(d);
    e.owned[0]       = b * pitch;
    e.owned[1]       = (uint64_t)(t - b) * pitch;
    if (d.ztop)
      {
        e.send_up[0] = (uint64_t)(t - d.p) * pitch; // the p planes below the top ghost plane
        e.send_up[1] = d.p * pitch;
        e.recv_up[0] = (uint64_t)t * pitch;
        e.recv_up[1] = pitch;
      }
    if (d.zg)
      {
        e.recv_dn[0] = 0;
        e.recv_dn[1] = d.p * pitch;
        e.send_dn[0] = (uint64_t)b * pitch;
        e.send_dn[1] = pitch;
      }
    return e;
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
    d.zg = d.ztop = d.gk0 = 0;
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
    d.nd2_all = d.nd[2];

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

  // the local box of one rank: the owned layers [c0, c1) of the slab plus the ghost layer below
  inline void
  build_slab_level(const int      p,
                   const int      ncell_all[3],
                   const int      tile_cells,
                   const double   lo[3],
                   const double   hi[3],
                   const uint32_t dirichlet,
                   const Slab    &slab,
                   HostLevel     &out)
  {
    if (!slab.distributed)
      {
        build_box_level(p, ncell_all, tile_cells, lo, hi, dirichlet, out);
        return;
      }
    const double hz     = (hi[2] - lo[2]) / ncell_all[2];
    const int    first  = slab.c0 - slab.zg;
    int          nc[3]  = {ncell_all[0], ncell_all[1], slab.c1 - first};
    double       llo[3] = {lo[0], lo[1], lo[2] + first * hz};
    double       lhi[3] = {hi[0], hi[1], lo[2] + slab.c1 * hz};
    uint32_t     dir    = dirichlet & 0xfu;
    if (slab.c0 == 0)
      dir |= dirichlet & 16u;
    if (slab.c1 == slab.ncz_all)
      dir |= dirichlet & 32u;
    build_box_level(p, nc, tile_cells, llo, lhi, dir, out);
    out.d.h[2]    = hz;
    out.d.zg      = slab.zg;
    out.d.ztop    = slab.ztop;
    out.d.gk0     = first * p;
    out.d.nd2_all = ncell_all[2] * p + 1;
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
