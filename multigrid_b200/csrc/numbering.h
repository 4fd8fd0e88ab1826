// Host-side construction of the patch entity numbering for refined box meshes
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
    std::vector<uint32_t> ent_base; // [n_patches][27]
    std::vector<uint32_t> ent_mask; // [n_patches]
    std::vector<uint32_t> constrained; // sorted list of Dirichlet DoFs
    std::vector<uint8_t>  bflag;       // per patch-boundary DoF (index - n_interior): 1 = Dirichlet
  };

  // largest admissible patch size: M in {4,2,1}, M p + 1 <= 17, M divides every ncell[d]
  inline int
  choose_patch_cells(const int p, const int ncell[3], const int max_m)
  {
    int M = 4;
    if (max_m > 0 && max_m < M)
      M = max_m >= 2 ? 2 : 1;
    while (M > 1 && (M * p + 1 > 17 || ncell[0] % M || ncell[1] % M || ncell[2] % M))
      M /= 2;
    return M;
  }

  inline void
  build_box_level(const int    p,
                  const int    ncell[3],
                  const int    M,
                  const double lo[3],
                  const double hi[3],
                  HostLevel   &out)
  {
    LevelDesc &d = out.d;
    d.p          = p;
    d.n          = p + 1;
    d.M          = M;
    d.np         = M * p + 1;
    d.q          = d.np - 2;
    uint64_t ndofs = 1;
    for (int c = 0; c < 3; ++c)
      {
        if (ncell[c] % M)
          throw std::runtime_error("patch size must divide the number of cells");
        d.ncell[c]  = ncell[c];
        d.npatch[c] = ncell[c] / M;
        d.nd[c]     = ncell[c] * p + 1;
        d.lo[c]     = lo[c];
        d.h[c]      = (hi[c] - lo[c]) / ncell[c];
        ndofs *= d.nd[c];
      }
    if (ndofs >= 0xffffffffull)
      throw std::runtime_error("level exceeds 2^32 DoFs per device");
    const int      npx = d.npatch[0], npy = d.npatch[1], npz = d.npatch[2];
    const uint64_t q = d.q;
    d.n_patches  = (uint32_t)npx * npy * npz;
    d.n_cells    = d.n_patches * M * M * M;
    d.n_dofs     = (uint32_t)ndofs;
    d.n_interior = (uint32_t)(d.n_patches * q * q * q);
    d.ent_base   = nullptr;
    d.ent_mask   = nullptr;

    // entity blocks: interiors | faces x,y,z | edges x,y,z | vertices
    const uint64_t nface[3] = {(uint64_t)(npx + 1) * npy * npz,
                               (uint64_t)npx * (npy + 1) * npz,
                               (uint64_t)npx * npy * (npz + 1)};
    const uint64_t nedge[3] = {(uint64_t)npx * (npy + 1) * (npz + 1),
                               (uint64_t)(npx + 1) * npy * (npz + 1),
                               (uint64_t)(npx + 1) * (npy + 1) * npz};
    uint64_t off_face[3], off_edge[3], off_vert;
    uint64_t off = d.n_interior;
    for (int c = 0; c < 3; ++c)
      {
        off_face[c] = off;
        off += nface[c] * q * q;
      }
    for (int c = 0; c < 3; ++c)
      {
        off_edge[c] = off;
        off += nedge[c] * q;
      }
    off_vert = off;
    off += (uint64_t)(npx + 1) * (npy + 1) * (npz + 1);
    if (off != ndofs)
      throw std::runtime_error("internal error: DoF count mismatch in numbering");

    out.ent_base.assign((size_t)d.n_patches * 27, 0);
    out.ent_mask.assign(d.n_patches, 0);
    out.bflag.assign(d.n_dofs - d.n_interior, 0);
    const int np[3] = {npx, npy, npz};
    for (int pz = 0; pz < npz; ++pz)
      for (int py = 0; py < npy; ++py)
        for (int px = 0; px < npx; ++px)
          {
            const uint32_t patch = ((uint32_t)pz * npy + py) * npx + px;
            const int      pp[3] = {px, py, pz};
            uint32_t       mask  = 0;
            for (int e = 0; e < 27; ++e)
              {
                const int ee[3] = {e % 3, (e / 3) % 3, e / 9};
                int       g[3], n_free = 0;
                bool      boundary = false;
                for (int c = 0; c < 3; ++c)
                  {
                    g[c] = pp[c] + (ee[c] == 2);
                    n_free += ee[c] == 1;
                    if ((ee[c] == 0 && pp[c] == 0) || (ee[c] == 2 && pp[c] == np[c] - 1))
                      boundary = true;
                  }
                uint64_t base = 0, count = 1;
                if (n_free == 3)
                  {
                    base  = (uint64_t)patch * q * q * q;
                    count = q * q * q;
                  }
                else if (n_free == 2)
                  {
                    const int dn = ee[0] != 1 ? 0 : (ee[1] != 1 ? 1 : 2); // normal direction
                    uint64_t  id;
                    if (dn == 0)
                      id = ((uint64_t)pz * npy + py) * (npx + 1) + g[0];
                    else if (dn == 1)
                      id = ((uint64_t)pz * (npy + 1) + g[1]) * npx + px;
                    else
                      id = ((uint64_t)g[2] * npy + py) * npx + px;
                    base  = off_face[dn] + id * q * q;
                    count = q * q;
                  }
                else if (n_free == 1)
                  {
                    const int da = ee[0] == 1 ? 0 : (ee[1] == 1 ? 1 : 2); // edge direction
                    uint64_t  id;
                    if (da == 0)
                      id = ((uint64_t)g[2] * (npy + 1) + g[1]) * npx + px;
                    else if (da == 1)
                      id = ((uint64_t)g[2] * npy + py) * (npx + 1) + g[0];
                    else
                      id = ((uint64_t)pz * (npy + 1) + g[1]) * (npx + 1) + g[0];
                    base  = off_edge[da] + id * q;
                    count = q;
                  }
                else
                  {
                    base  = off_vert + ((uint64_t)g[2] * (npy + 1) + g[1]) * (npx + 1) + g[0];
                    count = 1;
                  }
                out.ent_base[(size_t)patch * 27 + e] = (uint32_t)base;
                if (boundary)
                  {
                    mask |= 1u << e;
                    for (uint64_t i = 0; i < count; ++i)
                      out.bflag[base + i - d.n_interior] = 1;
                  }
              }
            out.ent_mask[patch] = mask;
          }
    out.constrained.clear();
    for (size_t i = 0; i < out.bflag.size(); ++i)
      if (out.bflag[i])
        out.constrained.push_back((uint32_t)(i + d.n_interior));
  }

  // lex_to_dof[lexicographic index] = library DoF index
  inline void
  build_lex_map(const HostLevel &lvl, uint32_t *lex_to_dof)
  {
    LevelDesc d = lvl.d;
    d.ent_base  = lvl.ent_base.data();
    d.ent_mask  = lvl.ent_mask.data();
    const int mp = d.M * d.p;
    for (uint32_t patch = 0; patch < d.n_patches; ++patch)
      {
        const int px = patch % d.npatch[0], py = (patch / d.npatch[0]) % d.npatch[1],
                  pz = patch / (d.npatch[0] * d.npatch[1]);
        for (int K = 0; K < d.np; ++K)
          for (int J = 0; J < d.np; ++J)
            for (int I = 0; I < d.np; ++I)
              {
                bool           c;
                const uint32_t dof = patch_dof(d, patch, I, J, K, c);
                const size_t   lex =
                  ((size_t)(pz * mp + K) * d.nd[1] + (py * mp + J)) * d.nd[0] + (px * mp + I);
                lex_to_dof[lex] = dof;
              }
      }
  }
} // namespace mgb
