// FE_Q(p) Laplacian on adaptively refined box meshes with hanging nodes:
//   AdaptiveMesh      <->  parallel::distributed::Triangulation after refine_global + execute_coarsening_and_refinement
//                          (poisson_l/program.cc:465-541), 2:1 balanced over faces, edges and corners as p4est does
//   AdaptiveOperator  <->  LaplaceOperator on the active cells with hanging-node constraints
//                          (poisson_l/program.cc:232-260): the reference reads and writes through deal.II's generic
//                          read_dof_values / distribute_local_to_global with AffineConstraints, i.e. a constrained
//                          value is sum_k w_k u[master_k] on read and its contribution is distributed to the masters
//                          on write; Dirichlet rows are identity rows (laplace_operator.h:592-593)
// Here the constraints are resolved once on the host (chains included) into a per-cell table: entry >= 0 is a DoF,
// -1 reads as zero (Dirichlet), <= -2 is a row of a small CSR list of (DoF, weight); the kernel applies them while it
// gathers and scatters, so the hanging-node values are never stored in a vector.
// There is no CPU fallback: vmult / diagonal run as kernels; the mesh and the numbering are host data structures.
#include <cuda_runtime.h>

#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <map>
#include <stdexcept>
#include <string>
#include <unordered_map>
#include <vector>

#include "../../include/mgb200.h"
#include "basis.h"

void mgb_set_last_error(const char *msg);

namespace mgb
{
#define MGB_AD_CHECK(call)                                                                      \
  do                                                                                            \
    {                                                                                           \
      cudaError_t err__ = (call);                                                               \
      if (err__ != cudaSuccess)                                                                 \
        throw std::runtime_error(std::string("CUDA error: ") + cudaGetErrorString(err__) +     \
                                 " at " + __FILE__ + ":" + std::to_string(__LINE__));           \
    }                                                                                           \
  while (0)

  struct ACell
  {
    int level;
    int c[3]; // position in units of cells of this level
  };

  static inline uint64_t
  pack_key(const int level, const int64_t x, const int64_t y, const int64_t z)
  {
    // 5 bits level, 19 bits per coordinate (node indices up to 524287)
    return ((uint64_t)level << 57) | ((uint64_t)x << 38) | ((uint64_t)y << 19) | (uint64_t)z;
  }

  struct AdaptiveMesh
  {
    int                 p = 1, n = 2;
    int                 nroot[3] = {1, 1, 1};
    std::vector<uint8_t> root_mask;
    double              lower[3] = {0, 0, 0}, root_size[3] = {1, 1, 1};
    std::vector<ACell>  cells;
    std::unordered_map<uint64_t, int> active; // (level, position) -> index in cells
    int                 max_level = 0;

    bool
    inside(const int level, const int c[3]) const
    {
      for (int d = 0; d < 3; ++d)
        if (c[d] < 0 || c[d] >= (nroot[d] << level))
          return false;
      const int r0 = c[0] >> level, r1 = c[1] >> level, r2 = c[2] >> level;
      return root_mask[(size_t)(r2 * nroot[1] + r1) * nroot[0] + r0] != 0;
    }
    int
    find(const int level, const int c[3]) const
    {
      auto it = active.find(pack_key(level, c[0], c[1], c[2]));
      return it == active.end() ? -1 : it->second;
    }

    void
    rebuild_index()
    {
      // deterministic order: by the lower corner on the finest lattice (z, y, x), coarser cell first
      max_level = 0;
      for (const ACell &k : cells)
        max_level = std::max(max_level, k.level);
      const int L = max_level;
      for (int d = 0; d < 3; ++d) // node indices are packed into 19 bits per direction (pack_key)
        if (((int64_t)nroot[d] << L) * p + 1 >= (int64_t)1 << 19)
          throw std::runtime_error("mesh too fine for the node index range: reduce the root grid, the degree or the number of levels");
      std::sort(cells.begin(), cells.end(), [L](const ACell &a, const ACell &b) {
        for (int d = 2; d >= 0; --d)
          {
            const int64_t xa = (int64_t)a.c[d] << (L - a.level), xb = (int64_t)b.c[d] << (L - b.level);
            if (xa != xb)
              return xa < xb;
          }
        return a.level < b.level;
      });
      active.clear();
      for (size_t i = 0; i < cells.size(); ++i)
        active[pack_key(cells[i].level, cells[i].c[0], cells[i].c[1], cells[i].c[2])] = (int)i;
    }

    void
    init(const int degree, const int nr[3], const uint8_t *mask, const double lo[3], const double hi[3])
    {
      p = degree;
      n = p + 1;
      size_t nm = 1;
      for (int d = 0; d < 3; ++d)
        {
          if (nr[d] < 1 || nr[d] > 64 || !(hi[d] > lo[d]))
            throw std::runtime_error("invalid root grid");
          nroot[d]     = nr[d];
          lower[d]     = lo[d];
          root_size[d] = (hi[d] - lo[d]) / nr[d];
          nm *= (size_t)nr[d];
        }
      root_mask.assign(nm, 1);
      if (mask)
        std::copy(mask, mask + nm, root_mask.begin());
      cells.clear();
      for (int z = 0; z < nroot[2]; ++z)
        for (int y = 0; y < nroot[1]; ++y)
          for (int x = 0; x < nroot[0]; ++x)
            if (root_mask[(size_t)(z * nroot[1] + y) * nroot[0] + x])
              cells.push_back(ACell{0, {x, y, z}});
      if (cells.empty())
        throw std::runtime_error("empty domain");
      rebuild_index();
    }

    // refine the flagged active cells; cells are added to the set until no two cells that share a face, an edge or
    // a corner differ by more than one level
    void
    refine(const uint8_t *flags)
    {
      std::vector<uint8_t> mark(flags, flags + cells.size());
      std::vector<int>     work;
      for (size_t i = 0; i < cells.size(); ++i)
        if (mark[i])
          work.push_back((int)i);
      while (!work.empty())
        {
          const ACell k = cells[work.back()];
          work.pop_back();
          if (k.level == 0)
            continue;
          if (k.level >= 12)
            throw std::runtime_error("more than 12 refinement levels");
          for (int dz = -1; dz <= 1; ++dz)
            for (int dy = -1; dy <= 1; ++dy)
              for (int dx = -1; dx <= 1; ++dx)
                {
                  const int nb[3] = {k.c[0] + dx, k.c[1] + dy, k.c[2] + dz};
                  if ((dx | dy | dz) == 0 || !inside(k.level, nb))
                    continue;
                  const int par[3] = {nb[0] >> 1, nb[1] >> 1, nb[2] >> 1};
                  const int j      = find(k.level - 1, par); // a coarser cell there must be refined as well
                  if (j >= 0 && !mark[j])
                    {
                      mark[j] = 1;
                      work.push_back(j);
                    }
                }
        }
      std::vector<ACell> next;
      next.reserve(cells.size() * 2);
      for (size_t i = 0; i < cells.size(); ++i)
        if (!mark[i])
          next.push_back(cells[i]);
        else
          for (int ch = 0; ch < 8; ++ch)
            next.push_back(ACell{cells[i].level + 1,
                                 {2 * cells[i].c[0] + (ch & 1), 2 * cells[i].c[1] + ((ch >> 1) & 1), 2 * cells[i].c[2] + (ch >> 2)}});
      cells.swap(next);
      rebuild_index();
    }

    // Triangulation::execute_coarsening_and_refinement (poisson_l/program.cc:533-541): refine the cells flagged for
    // refinement (keeping the 2:1 balance, which may refine further cells), then replace every family of eight active
    // siblings that are all flagged for coarsening by their parent, provided the parent would not touch a cell more than
    // one level finer than itself (judged on the mesh after the refinement, so the outcome does not depend on the order
    // in which families are visited).  A coarsening flag on a cell that is refined, or whose family is incomplete, lapses.
    void
    refine_and_coarsen(const uint8_t *refine_flags, const uint8_t *coarsen_flags)
    {
      std::unordered_map<uint64_t, int> flagged; // keys of the cells flagged for coarsening
      for (size_t i = 0; i < cells.size(); ++i)
        if (coarsen_flags && coarsen_flags[i] && !refine_flags[i] && cells[i].level > 0)
          flagged[pack_key(cells[i].level, cells[i].c[0], cells[i].c[1], cells[i].c[2])] = 1;
      refine(refine_flags);
      std::unordered_map<uint64_t, int> family; // parent key -> flagged children that are still active
      for (const ACell &k : cells)
        if (k.level > 0 && flagged.count(pack_key(k.level, k.c[0], k.c[1], k.c[2])))
          ++family[pack_key(k.level - 1, k.c[0] >> 1, k.c[1] >> 1, k.c[2] >> 1)];
      std::vector<ACell> parents;
      for (const ACell &k : cells)
        {
          if (k.level == 0 || (k.c[0] & 1) || (k.c[1] & 1) || (k.c[2] & 1)) // visit a family through its first child
            continue;
          const int par[3] = {k.c[0] >> 1, k.c[1] >> 1, k.c[2] >> 1};
          auto      it     = family.find(pack_key(k.level - 1, par[0], par[1], par[2]));
          if (it == family.end() || it->second != 8)
            continue;
          // the shell of child-level positions around the family: each one inside the domain must be covered by an
          // active cell of the children's level or coarser
          bool ok = true;
          for (int dz = -1; dz <= 2 && ok; ++dz)
            for (int dy = -1; dy <= 2 && ok; ++dy)
              for (int dx = -1; dx <= 2 && ok; ++dx)
                {
                  if (dx >= 0 && dx <= 1 && dy >= 0 && dy <= 1 && dz >= 0 && dz <= 1)
                    continue;
                  const int q[3] = {k.c[0] + dx, k.c[1] + dy, k.c[2] + dz};
                  if (!inside(k.level, q))
                    continue;
                  bool covered = false;
                  int  r[3]    = {q[0], q[1], q[2]};
                  for (int l = k.level; l >= 0 && !covered; --l)
                    {
                      covered = find(l, r) >= 0;
                      for (int d = 0; d < 3; ++d)
                        r[d] >>= 1;
                    }
                  ok = covered;
                }
          if (ok)
            parents.push_back(ACell{k.level - 1, {par[0], par[1], par[2]}});
        }
      if (parents.empty())
        return;
      std::unordered_map<uint64_t, int> gone;
      for (const ACell &P : parents)
        gone[pack_key(P.level, P.c[0], P.c[1], P.c[2])] = 1;
      std::vector<ACell> next;
      next.reserve(cells.size());
      for (const ACell &k : cells)
        if (k.level == 0 || !gone.count(pack_key(k.level - 1, k.c[0] >> 1, k.c[1] >> 1, k.c[2] >> 1)))
          next.push_back(k);
      next.insert(next.end(), parents.begin(), parents.end());
      cells.swap(next);
      rebuild_index();
    }

    // face neighbours of every active cell: faces 0..5 = -x, +x, -y, +y, -z, +z.  type -1: boundary; 0: one neighbour of the
    // same level; 1: one neighbour twice as large; 2: four neighbours half as large (in the order of their position in
    // the two tangential directions, first tangential direction fastest).  ids[cell][face][4].
    void
    face_neighbours(std::vector<int8_t> &type, std::vector<int32_t> &ids) const
    {
      type.assign(cells.size() * 6, -1);
      ids.assign(cells.size() * 24, -1);
      for (size_t c = 0; c < cells.size(); ++c)
        {
          const ACell &K = cells[c];
          for (int f = 0; f < 6; ++f)
            {
              const int d = f / 2, side = f % 2;
              int       q[3] = {K.c[0], K.c[1], K.c[2]};
              q[d] += side ? 1 : -1;
              if (!inside(K.level, q))
                continue;
              int j = find(K.level, q);
              if (j >= 0)
                {
                  type[c * 6 + f]    = 0;
                  ids[c * 24 + f * 4] = j;
                  continue;
                }
              const int par[3] = {q[0] >> 1, q[1] >> 1, q[2] >> 1};
              j                = K.level > 0 ? find(K.level - 1, par) : -1;
              if (j >= 0)
                {
                  type[c * 6 + f]    = 1;
                  ids[c * 24 + f * 4] = j;
                  continue;
                }
              const int t1 = d == 0 ? 1 : 0, t2 = d == 2 ? 1 : 2;
              type[c * 6 + f] = 2;
              for (int s = 0; s < 4; ++s)
                {
                  int r[3];
                  r[d]  = 2 * q[d] + (side ? 0 : 1);
                  r[t1] = 2 * q[t1] + (s & 1);
                  r[t2] = 2 * q[t2] + (s >> 1);
                  const int ch = find(K.level + 1, r);
                  if (ch < 0)
                    throw std::runtime_error("the mesh is not 2:1 balanced across a face");
                  ids[c * 24 + f * 4 + s] = ch;
                }
            }
        }
    }
  };

  // ------------------------------------------------------------------ numbering and constraints
  struct Term
  {
    int    dof;
    double w;
  };

  struct AdaptiveNumbering
  {
    const AdaptiveMesh *m = nullptr;
    std::vector<double> gll;                       // Gauss-Lobatto points on [0,1]
    std::vector<long double> gll_ld;
    std::unordered_map<uint64_t, int> dof_of_key;  // canonical (level, node index) -> DoF
    std::vector<uint8_t> dirichlet;                // per DoF
    std::vector<double>  points;                   // per DoF: x, y, z
    std::vector<int32_t> cell_index;               // [cell][n^3], Dirichlet DoFs read as zero (the operator's table)
    std::vector<int32_t> cell_index_plain;         // [cell][n^3], Dirichlet DoFs kept: values of a vector that carries the
                                                   // boundary data (compute_residual, error estimator, error norms)
    std::vector<int32_t> row_start;                // constraint rows
    std::vector<int32_t> row_dof;
    std::vector<double>  row_w;
    std::map<std::vector<std::pair<int, int64_t>>, int> row_cache; // identical rows are stored once
    int64_t n_hanging = 0;

    // positions (level-l cells) whose closure contains node index nd in one direction
    void
    cells_at(const int nd, int out[2], int &count) const
    {
      const int p = m->p;
      if (nd % p == 0)
        {
          out[0] = nd / p - 1;
          out[1] = nd / p;
          count  = 2;
        }
      else
        {
          out[0] = nd / p;
          count  = 1;
        }
    }

    // coordinate of node index nd in units of level-l cells
    long double
    coordinate(const int nd) const
    {
      const int p = m->p;
      if (nd % p == 0)
        return (long double)(nd / p);
      return (long double)(nd / p) + gll_ld[nd % p];
    }

    bool
    on_boundary(const int level, const int nd[3]) const
    {
      int c[3][2], cnt[3];
      for (int d = 0; d < 3; ++d)
        cells_at(nd[d], c[d], cnt[d]);
      for (int a = 0; a < cnt[0]; ++a)
        for (int b = 0; b < cnt[1]; ++b)
          for (int e = 0; e < cnt[2]; ++e)
            {
              const int pos[3] = {c[0][a], c[1][b], c[2][e]};
              if (!m->inside(level, pos))
                return true;
            }
      return false;
    }

    // an active cell of level-1 whose closure contains the node (level, nd), or -1
    int
    coarser_cell_at(const int level, const int nd[3], int pos_out[3]) const
    {
      if (level == 0)
        return -1;
      const int p = m->p;
      int       c[3][2], cnt[3];
      for (int d = 0; d < 3; ++d)
        {
          if (nd[d] % p != 0)
            {
              c[d][0] = (nd[d] / p) >> 1;
              cnt[d]  = 1;
            }
          else
            {
              const int v = nd[d] / p;
              if (v % 2 == 0)
                {
                  c[d][0] = v / 2 - 1;
                  c[d][1] = v / 2;
                  cnt[d]  = 2;
                }
              else
                {
                  c[d][0] = (v - 1) / 2;
                  cnt[d]  = 1;
                }
            }
        }
      for (int a = 0; a < cnt[0]; ++a)
        for (int b = 0; b < cnt[1]; ++b)
          for (int e = 0; e < cnt[2]; ++e)
            {
              const int pos[3] = {c[0][a], c[1][b], c[2][e]};
              const int j      = m->find(level - 1, pos);
              if (j >= 0)
                {
                  for (int d = 0; d < 3; ++d)
                    pos_out[d] = pos[d];
                  return j;
                }
            }
      return -1;
    }

    int
    regular_dof(int level, int nd[3])
    {
      const int p = m->p;
      // vertices exist on every coarser level they are a vertex of: one key for all of them
      while (level > 0 && nd[0] % p == 0 && nd[1] % p == 0 && nd[2] % p == 0 && (nd[0] / p) % 2 == 0 &&
             (nd[1] / p) % 2 == 0 && (nd[2] / p) % 2 == 0)
        {
          for (int d = 0; d < 3; ++d)
            nd[d] /= 2;
          --level;
        }
      const uint64_t key = pack_key(level, nd[0], nd[1], nd[2]);
      auto           it  = dof_of_key.find(key);
      if (it != dof_of_key.end())
        return it->second;
      const int id    = (int)dirichlet.size();
      dof_of_key[key] = id;
      dirichlet.push_back(on_boundary(level, nd) ? 1 : 0);
      for (int d = 0; d < 3; ++d)
        points.push_back(m->lower[d] + m->root_size[d] * (double)(coordinate(nd[d]) / (long double)(1 << level)));
      return id;
    }

    // value at node (level, nd) as a combination of DoFs (Dirichlet DoFs included; build() drops them for the operator)
    void
    resolve(const int level, const int nd[3], const double weight, std::vector<Term> &out)
    {
      const int p = m->p, n = m->n;
      int       cp[3];
      if (coarser_cell_at(level, nd, cp) < 0)
        {
          int       q[3] = {nd[0], nd[1], nd[2]};
          const int id   = regular_dof(level, q);
          out.push_back(Term{id, weight});
          return;
        }
      // weights of the coarse cell's nodes: 1-D Lagrange polynomials at the node's position inside the coarse cell
      std::vector<long double> w[3];
      for (int d = 0; d < 3; ++d)
        {
          const long double xi = coordinate(nd[d]) / 2 - (long double)cp[d];
          w[d].resize(n);
          for (int j = 0; j < n; ++j)
            {
              long double v = detail::lag(gll_ld, j, xi);
              if (std::fabs((double)v) < 1e-14)
                v = 0;
              w[d][j] = v;
            }
        }
      for (int k = 0; k < n; ++k)
        if (w[2][k] != 0)
          for (int j = 0; j < n; ++j)
            if (w[1][j] != 0)
              for (int i = 0; i < n; ++i)
                if (w[0][i] != 0)
                  {
                    const int master[3] = {cp[0] * p + i, cp[1] * p + j, cp[2] * p + k};
                    resolve(level - 1, master, weight * (double)(w[0][i] * w[1][j] * w[2][k]), out);
                  }
    }

    void
    build(const AdaptiveMesh &mesh)
    {
      m = &mesh;
      const int p = mesh.p, n = mesh.n, n3 = n * n * n;
      const Basis b = make_basis(p);
      gll           = b.nodes;
      gll_ld.assign(b.nodes.begin(), b.nodes.end());
      cell_index.assign(mesh.cells.size() * (size_t)n3, -1);
      cell_index_plain.assign(mesh.cells.size() * (size_t)n3, -1);
      row_start.assign(1, 0);
      std::vector<Term> terms, merged, kept;
      // table entry of a combination of DoFs: -1 (nothing), a DoF (weight one) or a row of the CSR list
      const auto encode = [&](const std::vector<Term> &t) -> int32_t {
        if (t.empty())
          return -1;
        if (t.size() == 1 && std::fabs(t[0].w - 1.) < 1e-13)
          return t[0].dof;
        std::vector<std::pair<int, int64_t>> sig;
        for (const Term &e : t)
          sig.emplace_back(e.dof, (int64_t)std::llround(e.w * 1e13));
        auto it = row_cache.find(sig);
        if (it != row_cache.end())
          return -2 - it->second;
        const int row = (int)row_start.size() - 1;
        for (const Term &e : t)
          {
            row_dof.push_back(e.dof);
            row_w.push_back(e.w);
          }
        row_start.push_back((int32_t)row_dof.size());
        row_cache[sig] = row;
        return -2 - row;
      };
      for (size_t c = 0; c < mesh.cells.size(); ++c)
        {
          const ACell &K = mesh.cells[c];
          for (int k = 0; k < n; ++k)
            for (int j = 0; j < n; ++j)
              for (int i = 0; i < n; ++i)
                {
                  const int nd[3] = {K.c[0] * p + i, K.c[1] * p + j, K.c[2] * p + k};
                  terms.clear();
                  int          cp[3];
                  const bool   hanging = coarser_cell_at(K.level, nd, cp) >= 0;
                  const size_t slot    = c * (size_t)n3 + (size_t)(k * n + j) * n + i;
                  if (!hanging)
                    {
                      int       q[3] = {nd[0], nd[1], nd[2]};
                      const int id   = regular_dof(K.level, q);
                      cell_index[slot]       = dirichlet[id] ? -1 : id;
                      cell_index_plain[slot] = id;
                      continue;
                    }
                  ++n_hanging;
                  resolve(K.level, nd, 1., terms);
                  // merge equal DoFs
                  std::sort(terms.begin(), terms.end(), [](const Term &a, const Term &b) { return a.dof < b.dof; });
                  merged.clear();
                  for (const Term &t : terms)
                    if (!merged.empty() && merged.back().dof == t.dof)
                      merged.back().w += t.w;
                    else
                      merged.push_back(t);
                  terms.clear();
                  kept.clear();
                  for (const Term &t : merged)
                    if (std::fabs(t.w) > 1e-14)
                      {
                        terms.push_back(t);
                        if (!dirichlet[t.dof])
                          kept.push_back(t);
                      }
                  cell_index[slot]       = encode(kept);
                  cell_index_plain[slot] = kept.size() == terms.size() ? cell_index[slot] : encode(terms);
                }
        }
      row_cache.clear();
      check_size();
    }
    size_t n_dofs() const { return dirichlet.size(); }
    void
    check_size() const
    {
      if (dirichlet.size() >= 0x7fffffffull || row_dof.size() >= 0x7fffffffull)
        throw std::runtime_error("too many DoFs for 32-bit indices");
    }
  };
} // namespace mgb

#include "adaptive_kernels.cuh"
#include "adaptive_solver.cuh"
#include "adaptive_estimate.cuh"
