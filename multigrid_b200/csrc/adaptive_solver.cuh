// Local-smoothing multigrid preconditioner and CG solver on the adaptive forest (config 5 of BASELINE.json;
// poisson_l/program.cc:283-420); included at the end of adaptive.cu.
//
//   level l                 all cells of size root / 2^l that are active or refined
//                           (parallel::distributed::Triangulation with construct_multigrid_hierarchy)
//   level operator          mg_matrices[level] (poisson_l/program.cc:283-305): no hanging nodes inside a level; the
//                           DoFs on the domain boundary and on the refinement edge (faces towards coarser active
//                           cells) are constrained to zero, identity rows
//   edge matrices           MGInterfaceOperator (:374-383; semantics spelled out at gpu/poisson_l.cu:390-436):
//                           vmult_interface_down = the refinement-edge rows of the level's unconstrained cell loop
//                           applied to a vector with the constrained entries zeroed; vmult_interface_up = the cell loop
//                           applied to the refinement-edge part of a vector, constrained rows dropped.  Only the cells
//                           that touch the refinement edge are visited.
//   transfer                MGTransferMatrixFree with initialize_constraints (:312-313): per child cell the parent's
//                           polynomial at the child's support points, boundary DoFs zero; restriction = transpose,
//                           fine values weighted by 1 / (number of level cells sharing the DoF)
//   copy_to_mg / from_mg    DoFs of the active cells of a level except those on its refinement edge
//   smoother                PreconditionChebyshev over point Jacobi, range 15, 15 CG steps for the eigenvalue estimate
//                           (:346-363); level 0: tolerance 1e-3, automatic degree, m() CG steps
//   cycle                   deal.II Multigrid::level_v_step with edge_out = edge_in = the interface matrices
//   outer solver            SolverCG in double on the active-mesh operator with hanging-node constraints
//                           (adaptive_kernels.cuh), ReductionControl(m(), 1e-13, rel_tol) (:385-392); V-cycle in float
// The level numbering is by support point, z-major (then y, x), so that deal.II's start vector of the eigenvalue
// estimate ("index mod 11") is defined independently of how the tables were built.
#pragma once

namespace mgb
{
  template <typename F>
  __global__ void
  k_ad_foreach(const size_t n, F f)
  {
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
      f(i);
  }

  template <typename F>
  __global__ void
  k_ad_sum(const size_t n, F f, double *out)
  {
    double acc = 0;
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
      acc += f(i);
    for (int o = 16; o > 0; o >>= 1)
      acc += __shfl_down_sync(0xffffffffu, acc, o);
    __shared__ double red[8];
    if ((threadIdx.x & 31) == 0)
      red[threadIdx.x >> 5] = acc;
    __syncthreads();
    if (threadIdx.x == 0)
      {
        double t = 0;
        for (int w = 0; w < (int)(blockDim.x >> 5); ++w)
          t += red[w];
        atomicAdd(out, t);
      }
  }

  // diag[dof] += diagonal entry of the (level-uniform) cell matrix, for every unconstrained local node
  template <typename T>
  __global__ void
  k_ad_level_diagonal(const int n_cells, const int n3, const int32_t *__restrict__ idx, const T *__restrict__ kdiag, T *diag)
  {
    const size_t total = (size_t)n_cells * n3;
    for (size_t t = blockIdx.x * (size_t)blockDim.x + threadIdx.x; t < total; t += (size_t)gridDim.x * blockDim.x)
      {
        const int32_t e = idx[t];
        if (e >= 0)
          atomicAdd(diag + e, kdiag[t % n3]);
      }
  }

  // Level transfer between a level and the one below, one block per FINE cell (child `octant` of `parent`).
  // RESTRICT = false: dst[fine] = sum_coarse P src[coarse] (overwrites; cells sharing a DoF write the same value).
  // RESTRICT = true:  dst[coarse] += P^T (weight o src[fine]).  Entries -1 of the tables (boundary DoFs) read as zero and
  // are never written.  P1[o (p+1) + i] = coarse 1-D basis function i at fine point o of the parent, o = 0 .. 2p.
  template <int p, typename T, bool RESTRICT>
  __global__ void
  __launch_bounds__((p + 1) * (p + 1))
    k_ad_transfer(const int32_t *__restrict__ fine_idx, const int32_t *__restrict__ coarse_idx,
                  const int32_t *__restrict__ parent, const uint8_t *__restrict__ octant, const T *__restrict__ P1,
                  const T *__restrict__ weight, const T *__restrict__ src, T *dst)
  {
    constexpr int n = p + 1, n3 = n * n * n;
    __shared__ T  A[n3], B[n3], Ps[3][n * n];
    const int     cell = blockIdx.x;
    const int     oc   = octant[cell];
    const int     i = threadIdx.x % n, j = threadIdx.x / n;
    const int32_t *fi = fine_idx + (size_t)cell * n3, *ci = coarse_idx + (size_t)parent[cell] * n3;
    // the (p+1) x (p+1) block of the embedding for this child, per direction: Ps[d][o n + c]
    for (int t = threadIdx.x; t < 3 * n * n; t += blockDim.x)
      {
        const int d = t / (n * n), r = t % (n * n);
        Ps[d][r]    = P1[(((oc >> d) & 1) * p + r / n) * n + r % n];
      }
    for (int k = 0; k < n; ++k)
      {
        const int     l = (k * n + j) * n + i;
        const int32_t e = RESTRICT ? fi[l] : ci[l];
        A[l]            = e >= 0 ? (RESTRICT ? weight[e] * src[e] : src[e]) : T(0);
      }
    __syncthreads();
    // x: thread (j, k) = (i, j) of the block
    for (int o = 0; o < n; ++o)
      {
        T s = 0;
#pragma unroll
        for (int c = 0; c < n; ++c)
          s += (RESTRICT ? Ps[0][c * n + o] : Ps[0][o * n + c]) * A[(j * n + i) * n + c];
        B[(j * n + i) * n + o] = s;
      }
    __syncthreads();
    // y: thread (i, k)
    for (int o = 0; o < n; ++o)
      {
        T s = 0;
#pragma unroll
        for (int c = 0; c < n; ++c)
          s += (RESTRICT ? Ps[1][c * n + o] : Ps[1][o * n + c]) * B[(j * n + c) * n + i];
        A[(j * n + o) * n + i] = s;
      }
    __syncthreads();
    // z: thread (i, j), result to global
    for (int o = 0; o < n; ++o)
      {
        T s = 0;
#pragma unroll
        for (int c = 0; c < n; ++c)
          s += (RESTRICT ? Ps[2][c * n + o] : Ps[2][o * n + c]) * A[(c * n + j) * n + i];
        const int     l = (o * n + j) * n + i;
        const int32_t e = RESTRICT ? ci[l] : fi[l];
        if (e >= 0)
          {
            if (RESTRICT)
              atomicAdd(dst + e, s);
            else
              dst[e] = s;
          }
      }
  }

  struct AdLevel
  {
    int                level = 0;
    std::vector<ACell> cells;
    int                n_dofs = 0, n_edge = 0, n_edge_cells = 0, n_copy = 0, n_con = 0;
    // device
    int32_t *idx_con = nullptr, *idx_full = nullptr; // [cell][n^3]: DoF, or -1 if constrained (boundary | edge) / if boundary
    int32_t *edge_in = nullptr, *edge_out = nullptr; // edge cells only: unconstrained DoF or -1 / refinement-edge DoF or -1
    int32_t *con_list = nullptr, *parent = nullptr, *copy_level = nullptr, *copy_active = nullptr;
    uint8_t *octant   = nullptr;
    float   *weight = nullptr, *inv_diag = nullptr;
    float   *defect = nullptr, *t = nullptr, *solution = nullptr, *tmp = nullptr, *tmp2 = nullptr;
    // smoother
    double       smoothing_range = 15., theta = 1, delta = 1, min_eig = 1, max_eig = 1;
    unsigned int degree = 1, eig_cg_n_iterations = 15, cg_iterations = 0;
  };

  struct AdaptiveSolver
  {
    AdaptiveOperator    *op = nullptr;
    cudaStream_t         stream = nullptr;
    int                  p = 1, n = 2, n3 = 8;
    std::vector<AdLevel> levels;
    float               *d_P1 = nullptr;
    double              *d_scal = nullptr, *h_scal = nullptr;
    double              *cg_g = nullptr, *cg_h = nullptr, *cg_d = nullptr;
    std::vector<double>  history;
    std::vector<void *>  owned;

    template <typename S>
    S *
    up(const std::vector<S> &h)
    {
      S *d = AdaptiveOperator::upload(h, stream);
      owned.push_back(d);
      return d;
    }
    template <typename S>
    S *
    zeros(const size_t count)
    {
      S *d = nullptr;
      MGB_AD_CHECK(cudaMalloc(&d, std::max<size_t>(count, 1) * sizeof(S)));
      MGB_AD_CHECK(cudaMemsetAsync(d, 0, std::max<size_t>(count, 1) * sizeof(S), stream));
      owned.push_back(d);
      return d;
    }

    ~AdaptiveSolver()
    {
      for (void *q : owned)
        cudaFree(q);
      if (h_scal)
        cudaFreeHost(h_scal);
    }

    template <typename F>
    void
    foreach (const size_t count, F f)
    {
      if (count == 0)
        return;
      const unsigned g = (unsigned)std::min<size_t>((count + 255) / 256, 148 * 8);
      k_ad_foreach<<<g, 256, 0, stream>>>(count, f);
      ++op->launches;
    }
    template <typename F>
    double
    sum(const size_t count, F f)
    {
      MGB_AD_CHECK(cudaMemsetAsync(d_scal, 0, sizeof(double), stream));
      const unsigned g = (unsigned)std::min<size_t>((count + 255) / 256, 148 * 4);
      k_ad_sum<<<std::max(g, 1u), 256, 0, stream>>>(count, f, d_scal);
      ++op->launches;
      MGB_AD_CHECK(cudaMemcpyAsync(h_scal, d_scal, sizeof(double), cudaMemcpyDeviceToHost, stream));
      MGB_AD_CHECK(cudaStreamSynchronize(stream));
      return h_scal[0];
    }

    // ------------------------------------------------------------------ set-up
    void
    build(AdaptiveOperator &o, const unsigned int n_smoother)
    {
      op = &o;
      o.to_device();
      stream = o.stream;
      AdaptiveMesh      &mesh = o.mesh;
      AdaptiveNumbering &num  = o.numbering();
      p                       = mesh.p;
      n                       = p + 1;
      n3                      = n * n * n;
      const int L             = mesh.max_level;
      MGB_AD_CHECK(cudaMalloc(&d_scal, 4 * sizeof(double)));
      owned.push_back(d_scal);
      MGB_AD_CHECK(cudaMallocHost(&h_scal, 4 * sizeof(double)));
      {
        const Basis        b = make_basis(p);
        std::vector<float> pf(b.P1.begin(), b.P1.end());
        d_P1 = up(pf);
      }
      // ---- cells of every level: active cells and their ancestors, z-major order
      std::vector<std::unordered_map<uint64_t, int>> cell_at(L + 1);
      levels.resize(L + 1);
      {
        std::vector<std::vector<ACell>> set(L + 1);
        for (const ACell &k : mesh.cells)
          {
            ACell a = k;
            for (;;)
              {
                const uint64_t key = pack_key(a.level, a.c[0], a.c[1], a.c[2]);
                if (cell_at[a.level].count(key))
                  break;
                cell_at[a.level][key] = -1;
                set[a.level].push_back(a);
                if (a.level == 0)
                  break;
                a = ACell{a.level - 1, {a.c[0] >> 1, a.c[1] >> 1, a.c[2] >> 1}};
              }
          }
        for (int l = 0; l <= L; ++l)
          {
            std::sort(set[l].begin(), set[l].end(), [](const ACell &a, const ACell &b) {
              for (int d = 2; d >= 0; --d)
                if (a.c[d] != b.c[d])
                  return a.c[d] < b.c[d];
              return false;
            });
            for (size_t c = 0; c < set[l].size(); ++c)
              cell_at[l][pack_key(l, set[l][c].c[0], set[l][c].c[1], set[l][c].c[2])] = (int)c;
            levels[l].level = l;
            levels[l].cells = set[l];
          }
      }
      // ---- numbering, constraints and tables level by level
      std::vector<std::vector<int32_t>> full_tab(L + 1);
      for (int l = 0; l <= L; ++l)
        {
          AdLevel     &lv = levels[l];
          const size_t nc = lv.cells.size();
          // support points by integer node index; the level numbering is their z-major order
          std::vector<uint64_t> keys(nc * n3);
          for (size_t c = 0; c < nc; ++c)
            for (int k = 0; k < n; ++k)
              for (int j = 0; j < n; ++j)
                for (int i = 0; i < n; ++i)
                  keys[c * n3 + (k * n + j) * n + i] =
                    ((uint64_t)(lv.cells[c].c[2] * p + k) << 38) | ((uint64_t)(lv.cells[c].c[1] * p + j) << 19) |
                    (uint64_t)(lv.cells[c].c[0] * p + i);
          std::vector<uint64_t> uniq(keys);
          std::sort(uniq.begin(), uniq.end());
          uniq.erase(std::unique(uniq.begin(), uniq.end()), uniq.end());
          lv.n_dofs = (int)uniq.size();
          std::vector<uint8_t> boundary(lv.n_dofs, 0), edge(lv.n_dofs, 0);
          std::vector<int>     valence(lv.n_dofs, 0);
          for (int g = 0; g < lv.n_dofs; ++g)
            {
              const int nd[3] = {(int)(uniq[g] & 0x7ffff), (int)((uniq[g] >> 19) & 0x7ffff), (int)(uniq[g] >> 38)};
              int       c[3][2], cnt[3];
              for (int d = 0; d < 3; ++d)
                num.cells_at(nd[d], c[d], cnt[d]);
              bool bnd = false, edg = false;
              for (int a = 0; a < cnt[0]; ++a)
                for (int b = 0; b < cnt[1]; ++b)
                  for (int e = 0; e < cnt[2]; ++e)
                    {
                      const int pos[3] = {c[0][a], c[1][b], c[2][e]};
                      if (!mesh.inside(l, pos))
                        bnd = true;
                      else if (!cell_at[l].count(pack_key(l, pos[0], pos[1], pos[2])))
                        edg = true;
                      else
                        ++valence[g];
                    }
              boundary[g] = bnd;
              edge[g]     = edg && !bnd;
            }
          std::vector<int32_t> con(nc * n3), full(nc * n3), clist, ein, eout;
          for (size_t t = 0; t < keys.size(); ++t)
            {
              const int g = (int)(std::lower_bound(uniq.begin(), uniq.end(), keys[t]) - uniq.begin());
              full[t]     = boundary[g] ? -1 : g;
              con[t]      = (boundary[g] || edge[g]) ? -1 : g;
            }
          for (int g = 0; g < lv.n_dofs; ++g)
            if (boundary[g] || edge[g])
              clist.push_back(g);
          for (size_t c = 0; c < nc; ++c)
            {
              bool touches = false;
              for (int t = 0; t < n3; ++t)
                if (full[c * n3 + t] >= 0 && edge[full[c * n3 + t]])
                  touches = true;
              if (!touches)
                continue;
              for (int t = 0; t < n3; ++t)
                {
                  const int32_t g = full[c * n3 + t];
                  ein.push_back(con[c * n3 + t]);
                  eout.push_back(g >= 0 && edge[g] ? g : -1);
                }
            }
          lv.n_edge       = (int)std::count(edge.begin(), edge.end(), (uint8_t)1);
          lv.n_edge_cells = (int)(ein.size() / n3);
          lv.n_con        = (int)clist.size();
          lv.idx_con      = up(con);
          lv.idx_full     = up(full);
          lv.con_list     = up(clist);
          lv.edge_in      = up(ein);
          lv.edge_out     = up(eout);
          std::vector<float> w(lv.n_dofs);
          for (int g = 0; g < lv.n_dofs; ++g)
            w[g] = 1.f / (float)std::max(valence[g], 1);
          lv.weight = up(w);
          if (l > 0)
            {
              std::vector<int32_t> par(nc);
              std::vector<uint8_t> oct(nc);
              for (size_t c = 0; c < nc; ++c)
                {
                  const ACell &k = lv.cells[c];
                  par[c]         = cell_at[l - 1].at(pack_key(l - 1, k.c[0] >> 1, k.c[1] >> 1, k.c[2] >> 1));
                  oct[c]         = (uint8_t)((k.c[0] & 1) | ((k.c[1] & 1) << 1) | ((k.c[2] & 1) << 2));
                }
              lv.parent = up(par);
              lv.octant = up(oct);
            }
          // copy indices: DoFs of the active cells of this level, off the refinement edge, unconstrained on the active mesh
          std::vector<std::pair<int32_t, int32_t>> pairs;
          for (size_t ca = 0; ca < mesh.cells.size(); ++ca)
            if (mesh.cells[ca].level == l)
              {
                const int c = cell_at[l].at(pack_key(l, mesh.cells[ca].c[0], mesh.cells[ca].c[1], mesh.cells[ca].c[2]));
                for (int t = 0; t < n3; ++t)
                  {
                    const int32_t a = num.cell_index[ca * (size_t)n3 + t], g = con[(size_t)c * n3 + t];
                    if (a >= 0 && g >= 0)
                      pairs.emplace_back(g, a);
                  }
              }
          std::sort(pairs.begin(), pairs.end());
          pairs.erase(std::unique(pairs.begin(), pairs.end()), pairs.end());
          std::vector<int32_t> cl, ca;
          for (auto &q : pairs)
            {
              cl.push_back(q.first);
              ca.push_back(q.second);
            }
          lv.n_copy      = (int)cl.size();
          lv.copy_level  = up(cl);
          lv.copy_active = up(ca);
          lv.defect      = zeros<float>(lv.n_dofs);
          lv.t           = zeros<float>(lv.n_dofs);
          lv.solution    = zeros<float>(lv.n_dofs);
          lv.tmp         = zeros<float>(lv.n_dofs);
          lv.tmp2        = zeros<float>(lv.n_dofs);
          lv.inv_diag    = zeros<float>(lv.n_dofs);
          full_tab[l].swap(full);
          // ---- inverse diagonal: the cell matrix is the same for all cells of a level
          {
            const Basis        b = make_basis(p);
            double             h[3], vol = 1;
            std::vector<float> kd(n3);
            for (int d = 0; d < 3; ++d)
              {
                h[d] = mesh.root_size[d] / (double)(1 << l);
                vol *= h[d];
              }
            for (int k = 0; k < n; ++k)
              for (int j = 0; j < n; ++j)
                for (int i = 0; i < n; ++i)
                  kd[(k * n + j) * n + i] = (float)(vol * (b.K1[i * n + i] * b.M1[j * n + j] * b.M1[k * n + k] / (h[0] * h[0]) +
                                                           b.M1[i * n + i] * b.K1[j * n + j] * b.M1[k * n + k] / (h[1] * h[1]) +
                                                           b.M1[i * n + i] * b.M1[j * n + j] * b.K1[k * n + k] / (h[2] * h[2])));
            float         *d_kd = up(kd);
            const unsigned g    = (unsigned)std::min<size_t>((nc * n3 + 255) / 256, 148 * 8);
            k_ad_level_diagonal<float><<<g, 256, 0, stream>>>((int)nc, n3, lv.idx_con, d_kd, lv.inv_diag);
            float         *dg = lv.inv_diag;
            const int32_t *cl2 = lv.con_list;
            foreach (lv.n_con, [=] __device__(size_t i) { dg[cl2[i]] = 1.f; });
            foreach (lv.n_dofs, [=] __device__(size_t i) { dg[i] = 1.f / dg[i]; });
          }
        }
      // ---- smoothers (poisson_l/program.cc:346-363)
      for (int l = 0; l <= L; ++l)
        {
          AdLevel &lv = levels[l];
          if (l > 0)
            {
              lv.smoothing_range     = 15.;
              lv.degree              = n_smoother;
              lv.eig_cg_n_iterations = 15;
            }
          else
            {
              lv.smoothing_range     = 1e-3;
              lv.degree              = 0xffffffffu;
              lv.eig_cg_n_iterations = (unsigned int)lv.n_dofs;
            }
          estimate_eigenvalues(lv);
        }
      const size_t N = num.n_dofs();
      cg_g           = zeros<double>(N);
      cg_h           = zeros<double>(N);
      cg_d           = zeros<double>(N);
      MGB_AD_CHECK(cudaStreamSynchronize(stream));
    }

    AdaptiveArgs<float>
    level_args(const AdLevel &lv, const float *src, float *dst)
    {
      AdaptiveArgs<float> A = op->args<float>(src, dst);
      A.n_cells             = (int)lv.cells.size();
      A.cell_index          = lv.idx_con;
      A.cell_index_out      = nullptr;
      A.uniform_level       = lv.level;
      A.row_start           = nullptr;
      A.row_dof             = nullptr;
      A.row_w               = nullptr;
      return A;
    }

    // dst = A_level src (constrained rows: identity)
    void
    apply_level(AdLevel &lv, float *dst, const float *src)
    {
      MGB_AD_CHECK(cudaMemsetAsync(dst, 0, lv.n_dofs * sizeof(float), stream));
      dispatch_adaptive<float>(p, level_args(lv, src, dst), false, stream);
      ++op->launches;
      const int32_t *cl = lv.con_list;
      foreach (lv.n_con, [=] __device__(size_t i) { dst[cl[i]] = src[cl[i]]; });
    }

    // dst += (cell loop over the cells at the refinement edge), read through `in`, scattered through `out`
    void
    apply_edge(AdLevel &lv, float *dst, const float *src, const int32_t *in, const int32_t *out)
    {
      if (lv.n_edge_cells == 0)
        return;
      AdaptiveArgs<float> A = level_args(lv, src, dst);
      A.n_cells             = lv.n_edge_cells;
      A.cell_index          = in;
      A.cell_index_out      = out;
      dispatch_adaptive<float>(p, A, false, stream);
      ++op->launches;
    }

    template <bool RESTRICT>
    void
    transfer(AdLevel &fine, AdLevel &coarse, float *dst, const float *src)
    {
      const int nc = (int)fine.cells.size();
#define MGB_AD_XFER(P)                                                                                              \
  case P:                                                                                                           \
    k_ad_transfer<P, float, RESTRICT><<<nc, (P + 1) * (P + 1), 0, stream>>>(fine.idx_full, coarse.idx_full, fine.parent, \
                                                                            fine.octant, d_P1, fine.weight, src, dst); \
    break;
      switch (p)
        {
          MGB_AD_XFER(1)
          MGB_AD_XFER(2)
          MGB_AD_XFER(3)
          MGB_AD_XFER(4)
          MGB_AD_XFER(5)
          MGB_AD_XFER(6)
          MGB_AD_XFER(7)
          MGB_AD_XFER(8)
        }
#undef MGB_AD_XFER
      ++op->launches;
      MGB_AD_CHECK(cudaGetLastError());
    }

    // deal.II PreconditionChebyshev::estimate_eigenvalues on the level operator (as Chebyshev<T>::initialize, solver.cu)
    void
    estimate_eigenvalues(AdLevel &lv)
    {
      const size_t   N = lv.n_dofs;
      float         *g = lv.defect, *h = lv.t, *d = lv.solution;
      const float   *P = lv.inv_diag;
      const int32_t *cl = lv.con_list;
      const double   total = sum(N, [=] __device__(size_t i) { return (double)(float)(i % 11); });
      const float    mean  = (float)(total / (double)N);
      foreach (N, [=] __device__(size_t i) { g[i] = -((float)(i % 11) - mean); });
      foreach (lv.n_con, [=] __device__(size_t i) { g[cl[i]] = 0.f; });
      std::vector<double> diag, off;
      double              res = std::sqrt(sum(N, [=] __device__(size_t i) { return (double)g[i] * g[i]; }));
      lv.cg_iterations        = 0;
      lv.min_eig = lv.max_eig = 1.;
      if (res > 1e-10 && lv.eig_cg_n_iterations > 0)
        {
          foreach (N, [=] __device__(size_t i) {
            h[i] = P[i] * g[i];
            d[i] = -h[i];
          });
          double gh = sum(N, [=] __device__(size_t i) { return (double)g[i] * h[i]; });
          double alpha_old = 0, beta_old = 0;
          for (unsigned int it = 1;; ++it)
            {
              apply_level(lv, h, d);
              const double alpha = gh / sum(N, [=] __device__(size_t i) { return (double)d[i] * h[i]; });
              const float  al    = (float)alpha;
              foreach (N, [=] __device__(size_t i) { g[i] += al * h[i]; });
              res = std::sqrt(sum(N, [=] __device__(size_t i) { return (double)g[i] * g[i]; }));
              if (it == 1)
                diag.push_back(1. / alpha);
              else
                {
                  diag.push_back(1. / alpha + beta_old / alpha_old);
                  off.push_back(std::sqrt(beta_old) / alpha_old);
                }
              lv.cg_iterations = it;
              if (it >= lv.eig_cg_n_iterations || res <= 1e-10)
                break;
              foreach (N, [=] __device__(size_t i) { h[i] = P[i] * g[i]; });
              const double gh_new = sum(N, [=] __device__(size_t i) { return (double)g[i] * h[i]; });
              const double beta   = gh_new / gh;
              gh                  = gh_new;
              const float be      = (float)beta;
              foreach (N, [=] __device__(size_t i) { d[i] = be * d[i] - h[i]; });
              alpha_old = alpha;
              beta_old  = beta;
            }
          tridiagonal_extremes(diag, off, lv.min_eig, lv.max_eig);
        }
      const double max_est = 1.2 * lv.max_eig;
      const double alpha   = lv.smoothing_range > 1. ? max_est / lv.smoothing_range : std::min(0.9 * max_est, lv.min_eig);
      if (lv.degree == 0xffffffffu)
        {
          const double actual_range = max_est / alpha;
          const double sigma        = (1. - std::sqrt(1. / actual_range)) / (1. + std::sqrt(1. / actual_range));
          const double eps          = lv.smoothing_range;
          lv.degree = 1 + static_cast<unsigned int>(std::log(1. / eps + std::sqrt(1. / eps / eps - 1.)) / std::log(1. / sigma));
        }
      lv.theta = (max_est + alpha) * 0.5;
      lv.delta = (max_est - alpha) * 0.5;
      for (float *v : {lv.defect, lv.t, lv.solution})
        MGB_AD_CHECK(cudaMemsetAsync(v, 0, N * sizeof(float), stream));
    }

    // extreme eigenvalues of the Lanczos tridiagonal matrix by Sturm-sequence bisection
    static void
    tridiagonal_extremes(const std::vector<double> &diag, const std::vector<double> &off, double &emin, double &emax)
    {
      const int k = (int)diag.size();
      if (k == 0)
        {
          emin = emax = 1.;
          return;
        }
      double lo = 1e300, hi = -1e300;
      for (int i = 0; i < k; ++i)
        {
          const double r = (i > 0 ? std::fabs(off[i - 1]) : 0.) + (i < k - 1 ? std::fabs(off[i]) : 0.);
          lo             = std::min(lo, diag[i] - r);
          hi             = std::max(hi, diag[i] + r);
        }
      auto below = [&](const double x) {
        int    count = 0;
        double q     = 1;
        for (int i = 0; i < k; ++i)
          {
            q = diag[i] - x - (i > 0 ? off[i - 1] * off[i - 1] / q : 0.);
            if (q == 0)
              q = 1e-300;
            if (q < 0)
              ++count;
          }
        return count;
      };
      auto kth = [&](const int idx) {
        double a = lo, c = hi;
        for (int it = 0; it < 200; ++it)
          {
            const double mid = 0.5 * (a + c);
            if (below(mid) > idx)
              c = mid;
            else
              a = mid;
          }
        return 0.5 * (a + c);
      };
      emin = kth(0);
      emax = kth(k - 1);
    }

    // PreconditionChebyshev: x = p(A) b from a zero start (zero_guess) or one more sweep on the iterate in x.
    // x_{k+1} = x_k + f1 (x_k - x_{k-1}) + f2 D^-1 (b - A x_k); the new iterate lands in lv.tmp, then the roles swap.
    void
    chebyshev(AdLevel &lv, float *&x, const float *b, const bool zero_guess)
    {
      const size_t N   = lv.n_dofs;
      const float *P   = lv.inv_diag;
      float       *xo  = lv.tmp, *r = lv.tmp2, *xc = x;
      const float  f0  = (float)(1. / lv.theta);
      if (zero_guess)
        foreach (N, [=] __device__(size_t i) { xc[i] = (f0 * P[i]) * b[i]; });
      else
        {
          apply_level(lv, r, xc);
          foreach (N, [=] __device__(size_t i) { xo[i] = xc[i] + (f0 * P[i]) * (b[i] - r[i]); });
          std::swap(xc, xo);
        }
      double       rhok  = lv.delta / lv.theta;
      const double sigma = lv.theta / lv.delta;
      for (unsigned int k = 0; k + 1 < lv.degree; ++k)
        {
          const double rhokp = 1. / (2. * sigma - rhok);
          const float  f1 = (float)(rhokp * rhok), f2 = (float)(2. * rhokp / lv.delta);
          rhok             = rhokp;
          const bool first = zero_guess && k == 0; // x_{k-1} = 0
          apply_level(lv, r, xc);
          float *xn = xo, *xk = xc;
          foreach (N, [=] __device__(size_t i) {
            const float upd = f2 * P[i] * (b[i] - r[i]);
            xn[i]           = first ? (1.f + f1) * xk[i] + upd : (1.f + f1) * xk[i] - f1 * xn[i] + upd;
          });
          std::swap(xc, xo);
        }
      // hand the buffers back: x is the iterate, lv.tmp the spare
      lv.tmp = xo;
      x      = xc;
    }

    // deal.II Multigrid::level_v_step with edge_out = edge_in = the interface matrices
    void
    level_v_step(const int l)
    {
      AdLevel &lv = levels[l];
      chebyshev(lv, lv.solution, lv.defect, true);
      if (l == 0)
        return;
      AdLevel     &lc = levels[l - 1];
      const size_t N  = lv.n_dofs;
      // t = defect - (A solution + interface_down(solution))
      apply_level(lv, lv.t, lv.solution);
      apply_edge(lv, lv.t, lv.solution, lv.edge_in, lv.edge_out);
      {
        float       *t = lv.t;
        const float *df = lv.defect;
        foreach (N, [=] __device__(size_t i) { t[i] = df[i] - t[i]; });
      }
      transfer<true>(lv, lc, lc.defect, lv.t);
      level_v_step(l - 1);
      // solution += P solution_coarse
      transfer<false>(lv, lc, lv.t, lc.solution);
      {
        float       *s = lv.solution, *df = lv.defect, *e = lv.tmp2;
        const float *t = lv.t;
        foreach (N, [=] __device__(size_t i) { s[i] += t[i]; });
        // defect -= interface_up(solution): the coupling of the refinement-edge values into the interior rows
        if (lv.n_edge_cells)
          {
            MGB_AD_CHECK(cudaMemsetAsync(e, 0, N * sizeof(float), stream));
            apply_edge(lv, e, lv.solution, lv.edge_out, lv.edge_in);
            foreach (N, [=] __device__(size_t i) { df[i] -= e[i]; });
          }
      }
      chebyshev(lv, lv.solution, lv.defect, false);
    }

    // PreconditionMG::vmult: copy_to_mg, one V-cycle, copy_from_mg
    void
    vmult(double *dst, const double *src)
    {
      const size_t N = op->numbering().n_dofs();
      for (AdLevel &lv : levels)
        {
          MGB_AD_CHECK(cudaMemsetAsync(lv.defect, 0, lv.n_dofs * sizeof(float), stream));
          float         *df = lv.defect;
          const int32_t *cl = lv.copy_level, *ca = lv.copy_active;
          foreach (lv.n_copy, [=] __device__(size_t i) { df[cl[i]] = (float)src[ca[i]]; });
        }
      level_v_step((int)levels.size() - 1);
      MGB_AD_CHECK(cudaMemsetAsync(dst, 0, N * sizeof(double), stream));
      for (AdLevel &lv : levels)
        {
          const float   *s  = lv.solution;
          const int32_t *cl = lv.copy_level, *ca = lv.copy_active;
          foreach (lv.n_copy, [=] __device__(size_t i) { dst[ca[i]] = (double)s[cl[i]]; });
        }
    }

    // deal.II SolverCG, ReductionControl(m(), 1e-13, rel_tol), zero start vector
    std::pair<unsigned int, double>
    solve_cg(double *x, const double *b, const double rel_tol)
    {
      const size_t N = op->numbering().n_dofs();
      double      *g = cg_g, *h = cg_h, *d = cg_d;
      history.clear();
      MGB_AD_CHECK(cudaMemsetAsync(x, 0, N * sizeof(double), stream));
      foreach (N, [=] __device__(size_t i) { g[i] = -b[i]; });
      double       res  = std::sqrt(sum(N, [=] __device__(size_t i) { return g[i] * g[i]; }));
      const double res0 = res;
      history.push_back(res);
      unsigned int it = 0;
      if (res > 1e-13)
        {
          vmult(h, g);
          foreach (N, [=] __device__(size_t i) { d[i] = -h[i]; });
          double gh = sum(N, [=] __device__(size_t i) { return g[i] * h[i]; });
          for (;;)
            {
              ++it;
              op->vmult<double>(h, d);
              const double alpha = gh / sum(N, [=] __device__(size_t i) { return d[i] * h[i]; });
              foreach (N, [=] __device__(size_t i) {
                x[i] += alpha * d[i];
                g[i] += alpha * h[i];
              });
              res = std::sqrt(sum(N, [=] __device__(size_t i) { return g[i] * g[i]; }));
              history.push_back(res);
              if (res <= 1e-13 || res <= rel_tol * res0 || it >= N)
                break;
              vmult(h, g);
              const double gh_new = sum(N, [=] __device__(size_t i) { return g[i] * h[i]; });
              const double beta   = gh_new / gh;
              gh                  = gh_new;
              foreach (N, [=] __device__(size_t i) { d[i] = beta * d[i] - h[i]; });
            }
        }
      MGB_AD_CHECK(cudaStreamSynchronize(stream));
      return std::make_pair(it, it > 0 ? std::pow(res / res0, 1. / it) : 0.);
    }
  };

  static void
  delete_adaptive_solver(void *q)
  {
    delete static_cast<AdaptiveSolver *>(q);
  }

  static AdaptiveSolver &
  ADS(mgb_adaptive h)
  {
    AdaptiveOperator &o = ADH(h);
    if (!o.mg)
      throw std::runtime_error("call mgb_adaptive_mg_setup first");
    return *static_cast<AdaptiveSolver *>(o.mg);
  }
} // namespace mgb

extern "C"
{
  int
  mgb_adaptive_mg_setup(mgb_adaptive h, int n_smoother)
  {
    AD_TRY
    AdaptiveOperator &o = ADH(h);
    if (n_smoother < 1 || n_smoother > 100)
      throw std::runtime_error("smoother degree must be between 1 and 100");
    o.drop_multigrid();
    AdaptiveSolver *s = new AdaptiveSolver();
    o.mg              = s;
    o.mg_deleter      = &delete_adaptive_solver;
    s->build(o, (unsigned int)n_smoother);
    AD_CATCH
  }

  int
  mgb_adaptive_mg_n_levels(mgb_adaptive h)
  {
    try
      {
        return (int)ADS(h).levels.size();
      }
    catch (const std::exception &e)
      {
        mgb_set_last_error(e.what());
        return -1;
      }
  }

  int
  mgb_adaptive_mg_level_info(mgb_adaptive h, int level, double info[12])
  {
    AD_TRY
    AdaptiveSolver &s = ADS(h);
    if (level < 0 || level >= (int)s.levels.size())
      throw std::runtime_error("level out of range");
    const AdLevel &lv = s.levels[level];
    info[0]           = lv.n_dofs;
    info[1]           = (double)lv.cells.size();
    info[2]           = lv.n_edge;
    info[3]           = lv.n_edge_cells;
    info[4]           = lv.n_copy;
    info[5]           = lv.n_con;
    info[6]           = lv.degree;
    info[7]           = lv.min_eig;
    info[8]           = lv.max_eig;
    info[9]           = lv.theta;
    info[10]          = lv.delta;
    info[11]          = lv.cg_iterations;
    AD_CATCH
  }

  // level operator and edge matrices of the local smoothing on float device vectors of the level (mg_level_info: n_dofs):
  // which = 0: matrix[level].vmult; 1: vmult_interface_down; 2: vmult_interface_up (MatrixFreeOperators::Base, used
  // through MGInterfaceOperator, poisson_l/program.cc:374-383); dst is overwritten
  int
  mgb_adaptive_mg_level_vmult(mgb_adaptive h, int level, int which, float *dst_dev, const float *src_dev)
  {
    AD_TRY
    AdaptiveSolver &s = ADS(h);
    if (level < 0 || level >= (int)s.levels.size())
      throw std::runtime_error("level out of range");
    if (which < 0 || which > 2 || !dst_dev || !src_dev)
      throw std::runtime_error("which: 0 vmult, 1 vmult_interface_down, 2 vmult_interface_up");
    AdLevel &lv = s.levels[level];
    if (which == 0)
      s.apply_level(lv, dst_dev, src_dev);
    else
      {
        MGB_AD_CHECK(cudaMemsetAsync(dst_dev, 0, (size_t)lv.n_dofs * sizeof(float), s.stream));
        if (which == 1)
          s.apply_edge(lv, dst_dev, src_dev, lv.edge_in, lv.edge_out);
        else
          s.apply_edge(lv, dst_dev, src_dev, lv.edge_out, lv.edge_in);
      }
    MGB_AD_CHECK(cudaStreamSynchronize(s.stream));
    AD_CATCH
  }

  int
  mgb_adaptive_mg_vmult(mgb_adaptive h, double *dst_dev, const double *src_dev)
  {
    AD_TRY
    ADS(h).vmult(dst_dev, src_dev);
    AD_CATCH
  }

  int
  mgb_adaptive_solve_cg(mgb_adaptive h, double *x_dev, const double *rhs_dev, double rel_tol, unsigned int *iterations,
                        double *reduction_rate, double *history, int *n_history)
  {
    AD_TRY
    AdaptiveSolver &s = ADS(h);
    const auto      r = s.solve_cg(x_dev, rhs_dev, rel_tol);
    if (iterations)
      *iterations = r.first;
    if (reduction_rate)
      *reduction_rate = r.second;
    if (history && n_history)
      {
        const int m = std::min<int>(*n_history, (int)s.history.size());
        std::copy(s.history.begin(), s.history.begin() + m, history);
        *n_history = m;
      }
    AD_CATCH
  }
}
