// Kernels, device object and C ABI of the adaptive-mesh operator; included at the end of adaptive.cu.
#pragma once

namespace mgb
{
  constexpr int AD_MAX_LEVELS = 13;

  template <typename T>
  struct AdaptiveArgs
  {
    int            n_cells;
    const int32_t *cell_index; // [cell][n^3]
    const int32_t *cell_index_out = nullptr; // scatter through another table (edge matrices of the local smoothing); null: the same
    const uint8_t *cell_level;
    int            uniform_level = -1;       // >= 0: all cells are of this level (level operators), cell_level unused
    const int32_t *row_start, *row_dof;
    const T       *row_w;
    const T       *M1, *K1;    // reference-cell 1-D mass and stiffness matrices, (p+1)^2
    const T       *src;
    T             *dst;
    T              coef[AD_MAX_LEVELS][3]; // per level: |J| / h_d^2
    // diagonal kernel
    const int32_t *cell_dof_start, *cell_dofs;
  };

  // value of local node `entry` of a cell: DoF, zero, or constraint row
  template <typename T>
  __device__ __forceinline__ T
  ad_read(const AdaptiveArgs<T> &A, const int32_t entry)
  {
    if (entry >= 0)
      return A.src[entry];
    if (entry == -1)
      return T(0);
    const int row = -2 - entry;
    T         s   = 0;
    for (int t = A.row_start[row]; t < A.row_start[row + 1]; ++t)
      s += A.row_w[t] * A.src[A.row_dof[t]];
    return s;
  }

  template <typename T>
  __device__ __forceinline__ void
  ad_add(const AdaptiveArgs<T> &A, const int32_t entry, const T v)
  {
    if (entry >= 0)
      atomicAdd(A.dst + entry, v);
    else if (entry <= -2)
      {
        const int row = -2 - entry;
        for (int t = A.row_start[row]; t < A.row_start[row + 1]; ++t)
          atomicAdd(A.dst + A.row_dof[t], A.row_w[t] * v);
      }
  }

  // K_cell u = (c0 K x M x M + c1 M x K x M + c2 M x M x K) u for the cell whose values sit in U (shared memory);
  // thread (i, j) of the cell returns its column out[k].  T1, T2, Q are scratch arrays of n^3.
  template <int p, typename T>
  __device__ __forceinline__ void
  ad_cell_apply(T *U, T *T1, T *T2, T *Q, const T *Ms, const T *Ks, const T c0, const T c1, const T c2, const int i,
                const int j, const bool active, T (&out)[p + 1])
  {
    constexpr int n = p + 1;
    if (active)
      for (int k = 0; k < n; ++k)
        {
          T        m = 0, s = 0;
          const T *u = U + (k * n + j) * n;
#pragma unroll
          for (int a = 0; a < n; ++a)
            {
              m += Ms[i * n + a] * u[a];
              s += Ks[i * n + a] * u[a];
            }
          T1[(k * n + j) * n + i] = m;
          T2[(k * n + j) * n + i] = s;
        }
    __syncthreads();
    if (active)
      for (int k = 0; k < n; ++k)
        {
          T mm = 0, km = 0, mk = 0;
#pragma unroll
          for (int a = 0; a < n; ++a)
            {
              const T t1 = T1[(k * n + a) * n + i], t2 = T2[(k * n + a) * n + i];
              mm += Ms[j * n + a] * t1;
              km += Ks[j * n + a] * t1;
              mk += Ms[j * n + a] * t2;
            }
          U[(k * n + j) * n + i] = mm;                // My Mx u
          Q[(k * n + j) * n + i] = c0 * mk + c1 * km; // c0 My Kx u + c1 Ky Mx u
        }
    __syncthreads();
    if (active)
      {
        T a[n], q[n];
#pragma unroll
        for (int k = 0; k < n; ++k)
          {
            a[k] = U[(k * n + j) * n + i];
            q[k] = Q[(k * n + j) * n + i];
          }
#pragma unroll
        for (int k = 0; k < n; ++k)
          {
            T s = 0, t = 0;
#pragma unroll
            for (int b = 0; b < n; ++b)
              {
                s += Ms[k * n + b] * q[b];
                t += Ks[k * n + b] * a[b];
              }
            out[k] = s + c2 * t;
          }
      }
  }

  template <int p>
  struct AdCfg
  {
    static constexpr int n   = p + 1;
    static constexpr int TPC = n * n;
    static constexpr int CPB = TPC >= 64 ? 1 : 64 / TPC;
  };

  // dst += C^T K C src over all active cells (dst zeroed by the caller)
  template <int p, typename T>
  __global__ void
  __launch_bounds__(AdCfg<p>::TPC *AdCfg<p>::CPB) k_adaptive_apply(const AdaptiveArgs<T> A)
  {
    constexpr int n = p + 1, n3 = n * n * n, TPC = AdCfg<p>::TPC, CPB = AdCfg<p>::CPB;
    extern __shared__ __align__(16) unsigned char ad_smem[];
    T        *Ms = reinterpret_cast<T *>(ad_smem), *Ks = Ms + n * n;
    const int lc = threadIdx.x / TPC, lt = threadIdx.x % TPC;
    T        *U = Ks + n * n + lc * 4 * n3, *T1 = U + n3, *T2 = T1 + n3, *Q = T2 + n3;
    for (int t = threadIdx.x; t < n * n; t += blockDim.x)
      {
        Ms[t] = A.M1[t];
        Ks[t] = A.K1[t];
      }
    const int  cell = blockIdx.x * CPB + lc;
    const bool active = cell < A.n_cells;
    const int  i = lt % n, j = lt / n;
    const int32_t *idx = A.cell_index + (size_t)(active ? cell : 0) * n3;
    int32_t        e[n];
    if (active)
      for (int k = 0; k < n; ++k)
        {
          e[k]                   = idx[(k * n + j) * n + i];
          U[(k * n + j) * n + i] = ad_read(A, e[k]);
        }
    __syncthreads();
    const int lev = A.uniform_level >= 0 ? A.uniform_level : (active ? A.cell_level[cell] : 0);
    T         out[n];
    ad_cell_apply<p, T>(U, T1, T2, Q, Ms, Ks, A.coef[lev][0], A.coef[lev][1], A.coef[lev][2], i, j, active, out);
    if (active)
      {
        if (A.cell_index_out != nullptr)
          {
            const int32_t *odx = A.cell_index_out + (size_t)cell * n3;
            for (int k = 0; k < n; ++k)
              ad_add(A, odx[(k * n + j) * n + i], out[k]);
          }
        else
          for (int k = 0; k < n; ++k)
            ad_add(A, e[k], out[k]);
      }
  }

  // diag[g] += v^T K_cell v with v = column g of the cell's constraint matrix, for every DoF g the cell refers to
  // (the reference applies the cell operator to unit vectors as well: laplace_operator.h:770-800)
  template <int p, typename T>
  __global__ void
  __launch_bounds__((AdCfg<p>::TPC + 31) / 32 * 32) k_adaptive_diagonal(const AdaptiveArgs<T> A)
  {
    constexpr int n = p + 1, n3 = n * n * n, TPC = AdCfg<p>::TPC;
    extern __shared__ __align__(16) unsigned char ad_smem[];
    T *Ms = reinterpret_cast<T *>(ad_smem), *Ks = Ms + n * n;
    T *U = Ks + n * n, *T1 = U + n3, *T2 = T1 + n3, *Q = T2 + n3;
    __shared__ double red[32];
    for (int t = threadIdx.x; t < n * n; t += blockDim.x)
      {
        Ms[t] = A.M1[t];
        Ks[t] = A.K1[t];
      }
    const int      cell = blockIdx.x;
    const bool     act  = threadIdx.x < TPC; // the block is padded to whole warps
    const int      i = threadIdx.x % n, j = act ? threadIdx.x / n : 0;
    const int32_t *idx = A.cell_index + (size_t)cell * n3;
    const int      lev = A.cell_level[cell];
    int32_t        e[n];
    for (int k = 0; k < n; ++k)
      e[k] = act ? idx[(k * n + j) * n + i] : -1;
    for (int t = A.cell_dof_start[cell]; t < A.cell_dof_start[cell + 1]; ++t)
      {
        const int g = A.cell_dofs[t];
        T         v[n];
        __syncthreads();
        for (int k = 0; k < n; ++k)
          {
            T w = 0;
            if (e[k] == g)
              w = 1;
            else if (e[k] <= -2)
              {
                const int row = -2 - e[k];
                for (int r = A.row_start[row]; r < A.row_start[row + 1]; ++r)
                  if (A.row_dof[r] == g)
                    w = A.row_w[r];
              }
            v[k] = w;
            if (act)
              U[(k * n + j) * n + i] = w;
          }
        __syncthreads();
        T out[n];
        ad_cell_apply<p, T>(U, T1, T2, Q, Ms, Ks, A.coef[lev][0], A.coef[lev][1], A.coef[lev][2], i, j, act, out);
        double s = 0;
        if (act)
          for (int k = 0; k < n; ++k)
            s += (double)v[k] * (double)out[k];
        for (int o = 16; o > 0; o >>= 1)
          s += __shfl_down_sync(0xffffffffu, s, o);
        __syncthreads();
        if ((threadIdx.x & 31) == 0)
          red[threadIdx.x >> 5] = s;
        __syncthreads();
        if (threadIdx.x == 0)
          {
            double tot = 0;
            for (int w = 0; w < (TPC + 31) / 32; ++w)
              tot += red[w];
            atomicAdd(A.dst + g, (T)tot);
          }
      }
  }

  template <typename T>
  __global__ void
  k_adaptive_fix_dirichlet(const int count, const int32_t *__restrict__ list, const T *__restrict__ src, T *__restrict__ dst,
                           const int mode)
  {
    // mode 0: identity rows dst = src (laplace_operator.h:592-593); mode 1: diagonal entry 1
    for (int t = blockIdx.x * blockDim.x + threadIdx.x; t < count; t += gridDim.x * blockDim.x)
      dst[list[t]] = mode == 0 ? src[list[t]] : T(1);
  }

  template <typename T>
  __global__ void
  k_adaptive_invert(const size_t n, T *__restrict__ d)
  {
    for (size_t t = blockIdx.x * (size_t)blockDim.x + threadIdx.x; t < n; t += (size_t)gridDim.x * blockDim.x)
      d[t] = T(1) / d[t];
  }

  template <int p, typename T>
  void
  launch_adaptive(const AdaptiveArgs<T> &A, const bool diagonal, cudaStream_t stream)
  {
    constexpr int n = p + 1, n3 = n * n * n;
    if (diagonal)
      {
        const size_t smem = (size_t)(2 * n * n + 4 * n3) * sizeof(T);
        cudaFuncSetAttribute(k_adaptive_diagonal<p, T>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        k_adaptive_diagonal<p, T><<<A.n_cells, (AdCfg<p>::TPC + 31) / 32 * 32, smem, stream>>>(A);
      }
    else
      {
        constexpr int CPB  = AdCfg<p>::CPB;
        const size_t  smem = (size_t)(2 * n * n + CPB * 4 * n3) * sizeof(T);
        cudaFuncSetAttribute(k_adaptive_apply<p, T>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        k_adaptive_apply<p, T><<<(A.n_cells + CPB - 1) / CPB, AdCfg<p>::TPC * CPB, smem, stream>>>(A);
      }
  }

  template <typename T>
  void
  dispatch_adaptive(const int p, const AdaptiveArgs<T> &A, const bool diagonal, cudaStream_t stream)
  {
    switch (p)
      {
        case 1: launch_adaptive<1, T>(A, diagonal, stream); break;
        case 2: launch_adaptive<2, T>(A, diagonal, stream); break;
        case 3: launch_adaptive<3, T>(A, diagonal, stream); break;
        case 4: launch_adaptive<4, T>(A, diagonal, stream); break;
        case 5: launch_adaptive<5, T>(A, diagonal, stream); break;
        case 6: launch_adaptive<6, T>(A, diagonal, stream); break;
        case 7: launch_adaptive<7, T>(A, diagonal, stream); break;
        default: launch_adaptive<8, T>(A, diagonal, stream); break;
      }
  }
} // namespace mgb

// ====================================================================== host object and C ABI
namespace mgb
{
  struct AdaptiveOperator
  {
    AdaptiveMesh       mesh;
    AdaptiveNumbering *num = nullptr;
    int                device = 0;
    // device side (built on first use)
    bool         on_device = false;
    cudaStream_t stream    = nullptr;
    int32_t     *d_cell_index = nullptr, *d_row_start = nullptr, *d_row_dof = nullptr, *d_dir_list = nullptr;
    int32_t     *d_cell_dof_start = nullptr, *d_cell_dofs = nullptr;
    uint8_t     *d_cell_level = nullptr;
    void        *d_row_w[2] = {nullptr, nullptr}, *d_M1[2] = {nullptr, nullptr}, *d_K1[2] = {nullptr, nullptr};
    void        *d_buf[2]   = {nullptr, nullptr};
    int          n_dirichlet = 0;
    int64_t      launches    = 0;
    // local-smoothing multigrid built on this mesh (adaptive_solver.cuh); dropped whenever the mesh changes
    void *mg                  = nullptr;
    void (*mg_deleter)(void *) = nullptr;

    void
    drop_multigrid()
    {
      if (mg && mg_deleter)
        mg_deleter(mg);
      mg = nullptr;
    }
    // device tables of the error estimator and the error norms (adaptive_estimate.cuh); dropped with the device data
    void *est                   = nullptr;
    void (*est_deleter)(void *) = nullptr;
    void
    drop_estimate()
    {
      if (est && est_deleter)
        est_deleter(est);
      est = nullptr;
    }

    ~AdaptiveOperator()
    {
      drop_multigrid();
      release_device();
      delete num;
    }

    void
    invalidate()
    {
      drop_multigrid();
      release_device();
      delete num;
      num = nullptr;
    }

    AdaptiveNumbering &
    numbering()
    {
      if (!num)
        {
          num = new AdaptiveNumbering();
          num->build(mesh);
        }
      return *num;
    }

    void
    release_device()
    {
      drop_estimate();
      if (!on_device)
        return;
      for (void *ptr : {(void *)d_cell_index, (void *)d_row_start, (void *)d_row_dof, (void *)d_dir_list, (void *)d_cell_dof_start,
                        (void *)d_cell_dofs, (void *)d_cell_level, d_row_w[0], d_row_w[1], d_M1[0], d_M1[1], d_K1[0], d_K1[1],
                        d_buf[0], d_buf[1]})
        if (ptr)
          cudaFree(ptr);
      d_cell_index = d_row_start = d_row_dof = d_dir_list = d_cell_dof_start = d_cell_dofs = nullptr;
      d_cell_level                                                                          = nullptr;
      for (int k = 0; k < 2; ++k)
        d_row_w[k] = d_M1[k] = d_K1[k] = d_buf[k] = nullptr;
      if (stream)
        cudaStreamDestroy(stream);
      stream    = nullptr;
      on_device = false;
    }

    template <typename S>
    static S *
    upload(const std::vector<S> &h, cudaStream_t st)
    {
      S *d = nullptr;
      MGB_AD_CHECK(cudaMalloc(&d, std::max<size_t>(h.size(), 1) * sizeof(S)));
      if (!h.empty())
        MGB_AD_CHECK(cudaMemcpyAsync(d, h.data(), h.size() * sizeof(S), cudaMemcpyHostToDevice, st));
      return d;
    }

    void
    to_device()
    {
      if (on_device)
        return;
      int count = 0;
      if (cudaGetDeviceCount(&count) != cudaSuccess || count == 0)
        throw std::runtime_error("no CUDA device: the adaptive operator has no CPU fallback");
      AdaptiveNumbering &N = numbering();
      MGB_AD_CHECK(cudaSetDevice(device));
      MGB_AD_CHECK(cudaStreamCreateWithFlags(&stream, cudaStreamNonBlocking));
      on_device    = true;
      d_cell_index = upload(N.cell_index, stream);
      d_row_start  = upload(N.row_start, stream);
      d_row_dof    = upload(N.row_dof, stream);
      std::vector<uint8_t> lev(mesh.cells.size());
      for (size_t c = 0; c < mesh.cells.size(); ++c)
        lev[c] = (uint8_t)mesh.cells[c].level;
      d_cell_level = upload(lev, stream);
      std::vector<int32_t> dl;
      for (size_t g = 0; g < N.n_dofs(); ++g)
        if (N.dirichlet[g])
          dl.push_back((int32_t)g);
      n_dirichlet = (int)dl.size();
      d_dir_list  = upload(dl, stream);
      // DoFs referred to by each cell (for the diagonal)
      const int            n3 = mesh.n * mesh.n * mesh.n;
      std::vector<int32_t> start(1, 0), list, tmp;
      for (size_t c = 0; c < mesh.cells.size(); ++c)
        {
          tmp.clear();
          for (int l = 0; l < n3; ++l)
            {
              const int32_t e = N.cell_index[c * (size_t)n3 + l];
              if (e >= 0)
                tmp.push_back(e);
              else if (e <= -2)
                for (int r = N.row_start[-2 - e]; r < N.row_start[-2 - e + 1]; ++r)
                  tmp.push_back(N.row_dof[r]);
            }
          std::sort(tmp.begin(), tmp.end());
          tmp.erase(std::unique(tmp.begin(), tmp.end()), tmp.end());
          list.insert(list.end(), tmp.begin(), tmp.end());
          start.push_back((int32_t)list.size());
        }
      d_cell_dof_start = upload(start, stream);
      d_cell_dofs      = upload(list, stream);
      const Basis         b = make_basis(mesh.p);
      std::vector<float>  wf(N.row_w.begin(), N.row_w.end()), mf(b.M1.begin(), b.M1.end()), kf(b.K1.begin(), b.K1.end());
      d_row_w[0] = upload(wf, stream);
      d_row_w[1] = upload(N.row_w, stream);
      d_M1[0]    = upload(mf, stream);
      d_M1[1]    = upload(b.M1, stream);
      d_K1[0]    = upload(kf, stream);
      d_K1[1]    = upload(b.K1, stream);
      MGB_AD_CHECK(cudaStreamSynchronize(stream));
    }

    template <typename T>
    AdaptiveArgs<T>
    args(const T *src, T *dst)
    {
      const int       k = sizeof(T) == 8;
      AdaptiveArgs<T> A;
      A.n_cells        = (int)mesh.cells.size();
      A.cell_index     = d_cell_index;
      A.cell_level     = d_cell_level;
      A.row_start      = d_row_start;
      A.row_dof        = d_row_dof;
      A.row_w          = (const T *)d_row_w[k];
      A.M1             = (const T *)d_M1[k];
      A.K1             = (const T *)d_K1[k];
      A.src            = src;
      A.dst            = dst;
      A.cell_dof_start = d_cell_dof_start;
      A.cell_dofs      = d_cell_dofs;
      for (int l = 0; l < AD_MAX_LEVELS; ++l)
        {
          double h[3], vol = 1;
          for (int d = 0; d < 3; ++d)
            {
              h[d] = mesh.root_size[d] / (double)(1 << l);
              vol *= h[d];
            }
          for (int d = 0; d < 3; ++d)
            A.coef[l][d] = (T)(vol / (h[d] * h[d]));
        }
      return A;
    }

    // dst = A src: hanging-node constraints resolved in the gather / scatter, Dirichlet rows identity
    template <typename T>
    void
    vmult(T *dst, const T *src)
    {
      to_device();
      const size_t N = numbering().n_dofs();
      MGB_AD_CHECK(cudaMemsetAsync(dst, 0, N * sizeof(T), stream));
      dispatch_adaptive<T>(mesh.p, args<T>(src, dst), false, stream);
      ++launches;
      if (n_dirichlet)
        {
          k_adaptive_fix_dirichlet<T><<<(n_dirichlet + 255) / 256, 256, 0, stream>>>(n_dirichlet, d_dir_list, src, dst, 0);
          ++launches;
        }
      MGB_AD_CHECK(cudaGetLastError());
    }

    template <typename T>
    void
    inverse_diagonal(T *inv)
    {
      to_device();
      const size_t N = numbering().n_dofs();
      MGB_AD_CHECK(cudaMemsetAsync(inv, 0, N * sizeof(T), stream));
      dispatch_adaptive<T>(mesh.p, args<T>(nullptr, inv), true, stream);
      ++launches;
      if (n_dirichlet)
        k_adaptive_fix_dirichlet<T><<<(n_dirichlet + 255) / 256, 256, 0, stream>>>(n_dirichlet, d_dir_list, inv, inv, 1);
      k_adaptive_invert<T><<<(unsigned)std::min<size_t>((N + 255) / 256, 148 * 8), 256, 0, stream>>>(N, inv);
      launches += 2;
      MGB_AD_CHECK(cudaGetLastError());
    }
  };
} // namespace mgb

using namespace mgb;

struct mgb_adaptive_t
{
  AdaptiveOperator *impl;
};

#define AD_TRY try {
#define AD_CATCH                                   \
  }                                                \
  catch (const std::exception &e)                  \
  {                                                \
    mgb_set_last_error(e.what());                  \
    return 1;                                      \
  }                                                \
  catch (...)                                      \
  {                                                \
    mgb_set_last_error("unknown error");           \
    return 2;                                      \
  }                                                \
  return 0;

static AdaptiveOperator &
ADH(mgb_adaptive h)
{
  if (!h || !h->impl)
    throw std::runtime_error("null adaptive operator handle");
  return *h->impl;
}

extern "C"
{
  int
  mgb_adaptive_create(int degree, const int n_roots[3], const unsigned char *root_mask, const double lower[3],
                      const double upper[3], int n_global_refinements, int device, mgb_adaptive *out)
  {
    AD_TRY
    if (!out || !n_roots || !lower || !upper)
      throw std::runtime_error("null argument");
    if (degree < 1 || degree > 8)
      throw std::runtime_error("polynomial degree must be between 1 and 8");
    if (n_global_refinements < 0 || n_global_refinements > 8)
      throw std::runtime_error("invalid number of refinements");
    *out    = nullptr;
    auto *o = new AdaptiveOperator();
    try
      {
        o->device = device;
        o->mesh.init(degree, n_roots, root_mask, lower, upper);
        for (int r = 0; r < n_global_refinements; ++r)
          {
            std::vector<uint8_t> all(o->mesh.cells.size(), 1);
            o->mesh.refine(all.data());
          }
      }
    catch (...)
      {
        delete o;
        throw;
      }
    *out         = new mgb_adaptive_t;
    (*out)->impl = o;
    AD_CATCH
  }

  int
  mgb_adaptive_destroy(mgb_adaptive h)
  {
    AD_TRY
    if (h)
      {
        delete h->impl;
        delete h;
      }
    AD_CATCH
  }

  int64_t
  mgb_adaptive_n_cells(mgb_adaptive h)
  {
    return h && h->impl ? (int64_t)h->impl->mesh.cells.size() : -1;
  }

  int
  mgb_adaptive_cells(mgb_adaptive h, int *level, int *position)
  {
    AD_TRY
    const AdaptiveMesh &m = ADH(h).mesh;
    for (size_t c = 0; c < m.cells.size(); ++c)
      {
        level[c] = m.cells[c].level;
        for (int d = 0; d < 3; ++d)
          position[3 * c + d] = m.cells[c].c[d];
      }
    AD_CATCH
  }

  int
  mgb_adaptive_refine(mgb_adaptive h, const unsigned char *flags)
  {
    AD_TRY
    AdaptiveOperator &o = ADH(h);
    if (!flags)
      throw std::runtime_error("null argument");
    o.invalidate();
    o.mesh.refine(flags);
    AD_CATCH
  }

  int64_t
  mgb_adaptive_n_dofs(mgb_adaptive h)
  {
    try
      {
        return (int64_t)ADH(h).numbering().n_dofs();
      }
    catch (const std::exception &e)
      {
        mgb_set_last_error(e.what());
        return -1;
      }
  }

  int
  mgb_adaptive_info(mgb_adaptive h, int64_t info[6])
  {
    AD_TRY
    AdaptiveOperator  &o = ADH(h);
    AdaptiveNumbering &N = o.numbering();
    int64_t            nd = 0;
    for (uint8_t f : N.dirichlet)
      nd += f;
    info[0] = (int64_t)N.n_dofs();
    info[1] = nd;
    info[2] = N.n_hanging;
    info[3] = (int64_t)N.row_start.size() - 1;
    info[4] = (int64_t)N.row_dof.size();
    info[5] = o.mesh.max_level;
    AD_CATCH
  }

  int
  mgb_adaptive_dof_points(mgb_adaptive h, double *xyz)
  {
    AD_TRY
    AdaptiveNumbering &N = ADH(h).numbering();
    std::copy(N.points.begin(), N.points.end(), xyz);
    AD_CATCH
  }

  int
  mgb_adaptive_dirichlet(mgb_adaptive h, unsigned char *mask)
  {
    AD_TRY
    AdaptiveNumbering &N = ADH(h).numbering();
    std::copy(N.dirichlet.begin(), N.dirichlet.end(), mask);
    AD_CATCH
  }

  int
  mgb_adaptive_tables(mgb_adaptive h, int32_t *cell_index, int32_t *row_start, int32_t *row_dof, double *row_w)
  {
    AD_TRY
    AdaptiveNumbering &N = ADH(h).numbering();
    if (cell_index)
      std::copy(N.cell_index.begin(), N.cell_index.end(), cell_index);
    if (row_start)
      std::copy(N.row_start.begin(), N.row_start.end(), row_start);
    if (row_dof)
      std::copy(N.row_dof.begin(), N.row_dof.end(), row_dof);
    if (row_w)
      std::copy(N.row_w.begin(), N.row_w.end(), row_w);
    AD_CATCH
  }

  int
  mgb_adaptive_vmult(mgb_adaptive h, int precision, void *dst, const void *src)
  {
    AD_TRY
    if (precision == MGB_F64)
      ADH(h).vmult<double>((double *)dst, (const double *)src);
    else if (precision == MGB_F32)
      ADH(h).vmult<float>((float *)dst, (const float *)src);
    else
      throw std::runtime_error("precision must be MGB_F32 or MGB_F64");
    AD_CATCH
  }

  int
  mgb_adaptive_inverse_diagonal(mgb_adaptive h, int precision, void *inv_diag)
  {
    AD_TRY
    if (precision == MGB_F64)
      ADH(h).inverse_diagonal<double>((double *)inv_diag);
    else if (precision == MGB_F32)
      ADH(h).inverse_diagonal<float>((float *)inv_diag);
    else
      throw std::runtime_error("precision must be MGB_F32 or MGB_F64");
    AD_CATCH
  }

  int
  mgb_adaptive_vmult_host(mgb_adaptive h, int precision, void *dst_host, const void *src_host)
  {
    AD_TRY
    AdaptiveOperator &o = ADH(h);
    if (precision != MGB_F64 && precision != MGB_F32)
      throw std::runtime_error("precision must be MGB_F32 or MGB_F64");
    o.to_device();
    const size_t bytes = o.numbering().n_dofs() * (precision == MGB_F64 ? 8 : 4);
    for (int k = 0; k < 2; ++k)
      if (!o.d_buf[k])
        MGB_AD_CHECK(cudaMalloc(&o.d_buf[k], o.numbering().n_dofs() * 8));
    MGB_AD_CHECK(cudaMemcpyAsync(o.d_buf[0], src_host, bytes, cudaMemcpyHostToDevice, o.stream));
    if (precision == MGB_F64)
      o.vmult<double>((double *)o.d_buf[1], (const double *)o.d_buf[0]);
    else
      o.vmult<float>((float *)o.d_buf[1], (const float *)o.d_buf[0]);
    MGB_AD_CHECK(cudaMemcpyAsync(dst_host, o.d_buf[1], bytes, cudaMemcpyDeviceToHost, o.stream));
    MGB_AD_CHECK(cudaStreamSynchronize(o.stream));
    AD_CATCH
  }

  int
  mgb_adaptive_sync(mgb_adaptive h)
  {
    AD_TRY
    AdaptiveOperator &o = ADH(h);
    if (o.on_device)
      MGB_AD_CHECK(cudaStreamSynchronize(o.stream));
    AD_CATCH
  }

  int64_t
  mgb_adaptive_launch_count(mgb_adaptive h)
  {
    return h && h->impl ? h->impl->launches : -1;
  }
}
