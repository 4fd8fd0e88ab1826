// Host orchestration of the matrix-free multigrid path on one B200 + degree-independent kernels.
//
//   LaplaceOperator<T>   <-> multigrid::LaplaceOperator<3,p,T>        common/laplace_operator.h:56-164
//   Chebyshev<T>         <-> dealii::PreconditionChebyshev<...>       call sites common/multigrid_solver.h:271-288
//   Transfer<T>          <-> dealii::MGTransferMatrixFree<3,T>        call sites common/multigrid_solver.h:213-221
//   MultigridSolver      <-> multigrid::MultigridSolver<3,p,float,double>  common/multigrid_solver.h:96-782
//
// There is no CPU fallback: every numerical step below is a kernel launch on the solver's stream.
#include <cuda_runtime.h>
#include <dlfcn.h>
#include <nccl.h> // types and enumerators only: the library is bound at run time (see Nccl below)

#include <algorithm>
#include <array>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <memory>
#include <stdexcept>
#include <string>
#include <vector>

#include "../../include/mgb200.h"
#include "basis.h"
#include "degree_kernels.h"
#include "functions.cuh"
#include "level.h"
#include "numbering.h"

namespace mgb
{
#define MGB_CUDA_CHECK(call)                                                                   \
  do                                                                                           \
    {                                                                                          \
      cudaError_t err__ = (call);                                                              \
      if (err__ != cudaSuccess)                                                                \
        throw std::runtime_error(std::string("CUDA error: ") + cudaGetErrorString(err__) +    \
                                 " at " + __FILE__ + ":" + std::to_string(__LINE__));          \
    }                                                                                          \
  while (0)

  const DegreeKernels &
  degree_kernels(const int p)
  {
    switch (p)
      {
        case 1: return degree_kernels_1();
        case 2: return degree_kernels_2();
        case 3: return degree_kernels_3();
        case 4: return degree_kernels_4();
        case 5: return degree_kernels_5();
        case 6: return degree_kernels_6();
        case 7: return degree_kernels_7();
        case 8: return degree_kernels_8();
      }
    throw std::runtime_error("polynomial degree must be between 1 and 8");
  }

  // ------------------------------------------------------------------ NCCL (several GPUs only)
  // The ghost exchange between z-slabs and the scalar all-reduces go through NCCL over NVLink.  The
  // library is bound with dlopen on first use, so one GPU needs no NCCL at all and a host process
  // that already carries a libnccl.so.2 (torch does) shares it.
  struct Nccl
  {
    void *handle = nullptr;
    ncclResult_t (*GetUniqueId)(ncclUniqueId *)                                                       = nullptr;
    ncclResult_t (*CommInitRank)(ncclComm_t *, int, ncclUniqueId, int)                                = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t)                                                           = nullptr;
    ncclResult_t (*AllReduce)(const void *, void *, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t,
                              cudaStream_t)                                                           = nullptr;
    ncclResult_t (*Send)(const void *, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t)         = nullptr;
    ncclResult_t (*Recv)(void *, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t)               = nullptr;
    ncclResult_t (*AllGather)(const void *, void *, size_t, ncclDataType_t, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*GroupStart)()                                                                      = nullptr;
    ncclResult_t (*GroupEnd)()                                                                        = nullptr;
    const char *(*GetErrorString)(ncclResult_t)                                                       = nullptr;

    static Nccl &
    get()
    {
      static Nccl n;
      if (n.handle)
        return n;
      const char *names[] = {getenv("MGB200_NCCL_LIB"), "libnccl.so.2", "libnccl.so"};
      for (const char *name : names)
        if (name && (n.handle = dlopen(name, RTLD_NOW | RTLD_GLOBAL)))
          break;
      if (!n.handle)
        throw std::runtime_error("more than one GPU requested but libnccl.so.2 cannot be loaded");
      auto sym = [&](const char *name) {
        void *f = dlsym(n.handle, name);
        if (!f)
          throw std::runtime_error(std::string("NCCL symbol missing: ") + name);
        return f;
      };
      n.GetUniqueId    = reinterpret_cast<decltype(n.GetUniqueId)>(sym("ncclGetUniqueId"));
      n.CommInitRank   = reinterpret_cast<decltype(n.CommInitRank)>(sym("ncclCommInitRank"));
      n.CommDestroy    = reinterpret_cast<decltype(n.CommDestroy)>(sym("ncclCommDestroy"));
      n.AllReduce      = reinterpret_cast<decltype(n.AllReduce)>(sym("ncclAllReduce"));
      n.AllGather      = reinterpret_cast<decltype(n.AllGather)>(sym("ncclAllGather"));
      n.Send           = reinterpret_cast<decltype(n.Send)>(sym("ncclSend"));
      n.Recv           = reinterpret_cast<decltype(n.Recv)>(sym("ncclRecv"));
      n.GroupStart     = reinterpret_cast<decltype(n.GroupStart)>(sym("ncclGroupStart"));
      n.GroupEnd       = reinterpret_cast<decltype(n.GroupEnd)>(sym("ncclGroupEnd"));
      n.GetErrorString = reinterpret_cast<decltype(n.GetErrorString)>(sym("ncclGetErrorString"));
      return n;
    }
    void
    check(const ncclResult_t r, const char *what)
    {
      if (r != ncclSuccess)
        throw std::runtime_error(std::string("NCCL error in ") + what + ": " + GetErrorString(r));
    }
  };

  // ------------------------------------------------------------------ peer-to-peer exchange over NVLink
  // Every rank owns an "arena" (one cudaMalloc, IPC-mapped by all ranks of the node): counters, all-reduce
  // slots and double-buffered receive buffers for the ghost planes.  A ghost exchange is two kernels per
  // rank, no NCCL call and no host round trip:
  //   k_p2p (push): waits for the neighbour's acknowledgement of the message before last, stores the
  //                 boundary planes STRAIGHT INTO THE NEIGHBOUR'S receive buffer (remote stores over
  //                 NVLink), fences, and the last block bumps the neighbour's data counter;
  //   k_p2p (pull): waits on the own data counters, copies the planes into the ghost range of the vector
  //                 and acknowledges.
  // The all-reduce of the CG scalars is one single-warp kernel: every rank stores its partial sums into its
  // slot on every peer, bumps a counter there, waits for all peers and adds the slots in rank order (so all
  // ranks get bit-identical sums).  All waits are bounded (about 2 s) and raise an error flag instead of
  // hanging the GPU.
  namespace p2p
  {
    constexpr size_t FLAG_STRIDE = 128;  // one counter per 128-byte line
    constexpr int    F_UP_DATA = 0, F_DN_DATA = 1, F_UP_ACK = 2, F_DN_ACK = 3, F_RED = 8; // F_RED + peer
    constexpr size_t RED_OFFSET = 8192;  // [parity][rank][4] doubles
    constexpr size_t BUF_OFFSET = 16384; // up[0], up[1], dn[0], dn[1]
    constexpr int    MAX_RANKS  = 16;
    constexpr long long TIMEOUT_CYCLES = 4000000000ll;

    struct Msg
    {
      const uint32_t    *src = nullptr;
      uint32_t          *dst = nullptr;
      size_t             words = 0;
      volatile uint32_t *wait_flag = nullptr;   // local counter that must reach wait_value first
      uint32_t           wait_value = 0;
      volatile uint32_t *signal_flag = nullptr; // counter (usually on the peer) set afterwards
      uint32_t           signal_value = 0;
    };

    struct Peers
    {
      unsigned char *arena[MAX_RANKS];
    };

    __device__ int g_failed = 0; // sticky: after the first time-out no wait spins again

    __device__ __forceinline__ void
    spin_until(volatile uint32_t *flag, const uint32_t value, int *error)
    {
      const long long t0 = clock64();
      while (*flag < value)
        if (*reinterpret_cast<volatile int *>(&g_failed) || clock64() - t0 > TIMEOUT_CYCLES)
          {
            g_failed = 1;
            *error   = 1;
            break;
          }
    }

    // 16-byte stores to the (256-byte aligned) buffers; the vector side is only element-aligned and is read
    // or written with scalar accesses
    __device__ __forceinline__ void
    copy_words(const Msg &m)
    {
      const size_t start = blockIdx.x * (size_t)blockDim.x + threadIdx.x, step = (size_t)gridDim.x * blockDim.x;
      const size_t quads = m.words / 4;
      if ((reinterpret_cast<uintptr_t>(m.dst) & 15) == 0)
        {
          uint4 *d4 = reinterpret_cast<uint4 *>(m.dst);
          for (size_t i = start; i < quads; i += step)
            d4[i] = make_uint4(m.src[4 * i], m.src[4 * i + 1], m.src[4 * i + 2], m.src[4 * i + 3]);
        }
      else if ((reinterpret_cast<uintptr_t>(m.src) & 15) == 0)
        {
          const uint4 *s4 = reinterpret_cast<const uint4 *>(m.src);
          for (size_t i = start; i < quads; i += step)
            {
              const uint4 v    = s4[i];
              m.dst[4 * i]     = v.x;
              m.dst[4 * i + 1] = v.y;
              m.dst[4 * i + 2] = v.z;
              m.dst[4 * i + 3] = v.w;
            }
        }
      else
        for (size_t i = start; i < 4 * quads; i += step)
          m.dst[i] = m.src[i];
      for (size_t i = 4 * quads + start; i < m.words; i += step)
        m.dst[i] = m.src[i];
    }

    __global__ void
    k_p2p(const Msg a, const Msg b, unsigned int *done, int *error)
    {
      if (threadIdx.x == 0)
        {
          if (a.wait_flag)
            spin_until(a.wait_flag, a.wait_value, error);
          if (b.wait_flag)
            spin_until(b.wait_flag, b.wait_value, error);
          __threadfence_system();
        }
      __syncthreads();
      copy_words(a);
      copy_words(b);
      __syncthreads();
      if (threadIdx.x == 0)
        {
          __threadfence_system(); // cumulative: orders the stores of the whole block (bar.sync above) before the ticket
          const unsigned int ticket = atomicAdd(done, 1u);
          if (ticket == gridDim.x - 1)
            {
              *done = 0;
              __threadfence_system();
              if (a.signal_flag)
                *a.signal_flag = a.signal_value;
              if (b.signal_flag)
                *b.signal_flag = b.signal_value;
            }
        }
    }

    // one warp; `k` = 1-based number of this all-reduce
    __global__ void
    k_allreduce(const Peers peers, const int rank, const int n_ranks, const int count, const uint32_t k, double *inout,
                int *error)
    {
      const int j      = threadIdx.x;
      const int parity = k & 1u;
      if (j < n_ranks)
        {
          volatile double *slot =
            reinterpret_cast<volatile double *>(peers.arena[j] + RED_OFFSET) + (parity * MAX_RANKS + rank) * 4;
          for (int c = 0; c < count; ++c)
            slot[c] = inout[c];
          __threadfence_system();
          *reinterpret_cast<volatile uint32_t *>(peers.arena[j] + (F_RED + rank) * FLAG_STRIDE) = k;
          spin_until(reinterpret_cast<volatile uint32_t *>(peers.arena[rank] + (F_RED + j) * FLAG_STRIDE), k, error);
          __threadfence_system();
        }
      __syncthreads();
      if (j < count)
        {
          const volatile double *slots =
            reinterpret_cast<const volatile double *>(peers.arena[rank] + RED_OFFSET) + parity * MAX_RANKS * 4;
          double sum = 0;
          for (int r = 0; r < n_ranks; ++r)
            sum += slots[r * 4 + j];
          inout[j] = sum;
        }
    }
  } // namespace p2p

  // ------------------------------------------------------------------ context
  struct Ctx
  {
    int          device   = 0;
    int          rank = 0, n_ranks = 1; // z-slab decomposition over the GPUs of one node
    ncclComm_t   comm     = nullptr;
    int64_t      exchanges = 0;          // ghost exchanges + all-reduces issued so far
    // peer-to-peer path (MGB200_P2P=1; the default is NCCL send/recv, which measured faster)
    bool           p2p_on = false;
    p2p::Peers     peers;
    size_t         p2p_buf_bytes = 0;
    uint32_t       n_up_sent = 0, n_dn_sent = 0, n_up_recv = 0, n_dn_recv = 0, n_reduce = 0;
    unsigned int  *p2p_done  = nullptr; // ticket counter of k_p2p
    int           *p2p_error = nullptr; // mapped pinned host memory: set by a wait that timed out
    cudaStream_t comm_stream = nullptr;  // ghost exchange overlapped with the interior cell layers
    cudaEvent_t  ev_fork = nullptr, ev_join = nullptr;
    bool         overlap = true;
    cudaStream_t stream   = nullptr;
    int64_t      launches = 0;
    double      *scal_dev = nullptr; // 16 device scalars for reductions
    double      *scal_host = nullptr; // pinned mirror
    bool         profile  = false;

    // ---- optional CUDA-event ranges on the solver stream (README-style per-level table,
    //      multigrid_solver.h:348-371, and per-kernel-class launch durations for bench.py)
    static constexpr int n_classes_per_level = 16;
    struct Range
    {
      cudaEvent_t a, b;
      int         cls;
    };
    std::vector<Range>       open_ranges;
    int                      range_depth = 0; // ranges begun but not yet ended (they nest)
    std::vector<cudaEvent_t> event_pool;
    std::vector<double>      class_ms;
    std::vector<int64_t>     class_count;
    cudaEvent_t              timer_ev[2] = {nullptr, nullptr};

    cudaEvent_t
    get_event()
    {
      cudaEvent_t e;
      if (!event_pool.empty())
        {
          e = event_pool.back();
          event_pool.pop_back();
        }
      else
        MGB_CUDA_CHECK(cudaEventCreate(&e));
      return e;
    }
    // returns a handle (index) or -1 when profiling is off
    int
    range_begin(const int level, const int kind)
    {
      if (!profile)
        return -1;
      // flush only between top-level ranges: an enclosing range has no end event yet and its
      // handle (an index into open_ranges) must stay valid until its range_end
      if (open_ranges.size() > 200000 && range_depth == 0)
        collect_ranges();
      Range r;
      r.a   = get_event();
      r.b   = get_event();
      r.cls = level * n_classes_per_level + kind;
      MGB_CUDA_CHECK(cudaEventRecord(r.a, stream));
      open_ranges.push_back(r);
      ++range_depth;
      return (int)open_ranges.size() - 1;
    }
    void
    range_end(const int handle)
    {
      if (handle >= 0 && (size_t)handle < open_ranges.size())
        {
          cudaEventRecord(open_ranges[handle].b, stream); // called from destructors: never throws
          --range_depth;
        }
    }
    void
    collect_ranges()
    {
      if (open_ranges.empty() || range_depth != 0)
        return;
      MGB_CUDA_CHECK(cudaStreamSynchronize(stream));
      for (const Range &r : open_ranges)
        {
          float ms = 0;
          MGB_CUDA_CHECK(cudaEventElapsedTime(&ms, r.a, r.b));
          if ((size_t)r.cls >= class_ms.size())
            {
              class_ms.resize(r.cls + 1, 0.);
              class_count.resize(r.cls + 1, 0);
            }
          class_ms[r.cls] += ms;
          class_count[r.cls] += 1;
          event_pool.push_back(r.a);
          event_pool.push_back(r.b);
        }
      open_ranges.clear();
    }

    void
    init(const int dev)
    {
      device = dev;
      MGB_CUDA_CHECK(cudaSetDevice(dev));
      MGB_CUDA_CHECK(cudaStreamCreateWithFlags(&stream, cudaStreamNonBlocking));
      MGB_CUDA_CHECK(cudaMalloc(&scal_dev, 16 * sizeof(double)));
      MGB_CUDA_CHECK(cudaMallocHost(&scal_host, 16 * sizeof(double)));
    }
    void
    init_comm(const int rank_, const int n_ranks_, const unsigned char *id)
    {
      rank    = rank_;
      n_ranks = n_ranks_;
      if (n_ranks <= 1)
        return;
      if (rank < 0 || rank >= n_ranks)
        throw std::runtime_error("invalid rank");
      Nccl        &n = Nccl::get();
      ncclUniqueId uid;
      static_assert(sizeof(uid) == 128, "ncclUniqueId is 128 bytes");
      std::memcpy(&uid, id, sizeof(uid));
      n.check(n.CommInitRank(&comm, n_ranks, uid, rank), "ncclCommInitRank");
      int lo = 0, hi = 0;
      MGB_CUDA_CHECK(cudaDeviceGetStreamPriorityRange(&lo, &hi));
      MGB_CUDA_CHECK(cudaStreamCreateWithPriority(&comm_stream, cudaStreamNonBlocking, hi));
      MGB_CUDA_CHECK(cudaEventCreateWithFlags(&ev_fork, cudaEventDisableTiming));
      MGB_CUDA_CHECK(cudaEventCreateWithFlags(&ev_join, cudaEventDisableTiming));
      const char *ov = getenv("MGB200_OVERLAP");
      overlap        = !(ov && ov[0] == '0');
    }
    // the comm stream continues after everything issued so far on the solver stream ...
    void
    fork_comm()
    {
      MGB_CUDA_CHECK(cudaEventRecord(ev_fork, stream));
      MGB_CUDA_CHECK(cudaStreamWaitEvent(comm_stream, ev_fork, 0));
    }
    // ... and the solver stream after everything issued so far on the comm stream
    void
    join_comm()
    {
      MGB_CUDA_CHECK(cudaEventRecord(ev_join, comm_stream));
      MGB_CUDA_CHECK(cudaStreamWaitEvent(stream, ev_join, 0));
    }
    unsigned char *flag_at(const int r, const int f) const { return peers.arena[r] + f * p2p::FLAG_STRIDE; }
    unsigned char *buffer_at(const int r, const int channel /*0 up, 1 down*/, const uint32_t k) const
    {
      return peers.arena[r] + p2p::BUF_OFFSET + (size_t)(2 * channel + (k & 1u)) * p2p_buf_bytes;
    }

    // allocate and zero the arena, exchange the IPC handles through NCCL, map the peers' arenas
    void
    init_p2p(const size_t max_message_bytes)
    {
      // opt-in: measured on 2 x B200 (profiles/r1_exchange_latency.txt) one exchange costs 21-29 us this way
      // and 11-24 us through NCCL's grouped send/recv, so NCCL stays the default
      const char *env = getenv("MGB200_P2P");
      if (n_ranks <= 1 || !(env && env[0] == '1') || n_ranks > p2p::MAX_RANKS)
        return;
      Nccl &n         = Nccl::get();
      p2p_buf_bytes   = (max_message_bytes + 255) / 256 * 256;
      const size_t sz = p2p::BUF_OFFSET + 4 * p2p_buf_bytes;
      unsigned char *mine = nullptr;
      MGB_CUDA_CHECK(cudaMalloc(&mine, sz));
      MGB_CUDA_CHECK(cudaMemset(mine, 0, sz));
      MGB_CUDA_CHECK(cudaMalloc(&p2p_done, sizeof(unsigned int)));
      MGB_CUDA_CHECK(cudaMemset(p2p_done, 0, sizeof(unsigned int)));
      MGB_CUDA_CHECK(cudaHostAlloc(&p2p_error, sizeof(int), cudaHostAllocMapped));
      *p2p_error = 0;
      MGB_CUDA_CHECK(cudaDeviceSynchronize());
      cudaIpcMemHandle_t handle;
      MGB_CUDA_CHECK(cudaIpcGetMemHandle(&handle, mine));
      static_assert(sizeof(handle) == 64, "cudaIpcMemHandle_t is 64 bytes");
      unsigned char *all_dev = nullptr;
      MGB_CUDA_CHECK(cudaMalloc(&all_dev, 64 * n_ranks));
      MGB_CUDA_CHECK(cudaMemcpy(all_dev + 64 * rank, &handle, 64, cudaMemcpyHostToDevice));
      n.check(n.AllGather(all_dev + 64 * rank, all_dev, 64, ncclInt8, comm, stream), "ncclAllGather");
      MGB_CUDA_CHECK(cudaStreamSynchronize(stream));
      std::vector<cudaIpcMemHandle_t> all(n_ranks);
      MGB_CUDA_CHECK(cudaMemcpy(all.data(), all_dev, 64 * n_ranks, cudaMemcpyDeviceToHost));
      MGB_CUDA_CHECK(cudaFree(all_dev));
      bool ok = true;
      for (int r = 0; r < n_ranks; ++r)
        {
          peers.arena[r] = nullptr;
          if (r == rank)
            peers.arena[r] = mine;
          else if (cudaIpcOpenMemHandle(reinterpret_cast<void **>(&peers.arena[r]), all[r], cudaIpcMemLazyEnablePeerAccess) !=
                   cudaSuccess)
            {
              cudaGetLastError();
              ok = false;
            }
        }
      // all ranks must agree (one failed mapping anywhere -> everybody uses NCCL)
      double flag = ok ? 0. : 1.;
      MGB_CUDA_CHECK(cudaMemcpy(scal_dev, &flag, sizeof(double), cudaMemcpyHostToDevice));
      n.check(n.AllReduce(scal_dev, scal_dev, 1, ncclDouble, ncclSum, comm, stream), "ncclAllReduce");
      MGB_CUDA_CHECK(cudaStreamSynchronize(stream));
      MGB_CUDA_CHECK(cudaMemcpy(&flag, scal_dev, sizeof(double), cudaMemcpyDeviceToHost));
      p2p_on = flag == 0.;
    }

    void
    check_p2p_error() const
    {
      if (p2p_error && *p2p_error)
        throw std::runtime_error("a peer-to-peer wait timed out: the ranks are out of step (or a peer died)");
    }

    // sum of `count` doubles over the ranks, in place, ordered on the solver stream
    void
    all_reduce_sum(double *dev, const int count)
    {
      if (p2p_on && count <= 4)
        {
          ++n_reduce;
          p2p::k_allreduce<<<1, 32, 0, stream>>>(peers, rank, n_ranks, count, n_reduce, dev, p2p_error);
          check_launch();
          ++exchanges;
          return;
        }
      Nccl &n = Nccl::get();
      n.check(n.AllReduce(dev, dev, count, ncclDouble, ncclSum, comm, stream), "ncclAllReduce");
      ++exchanges;
    }
    void
    destroy()
    {
      if (p2p_buf_bytes)
        {
          cudaDeviceSynchronize();
          for (int r = 0; r < n_ranks; ++r)
            if (peers.arena[r])
              {
                if (r == rank)
                  cudaFree(peers.arena[r]);
                else
                  cudaIpcCloseMemHandle(peers.arena[r]);
              }
          if (p2p_done)
            cudaFree(p2p_done);
          if (p2p_error)
            cudaFreeHost(p2p_error);
          p2p_buf_bytes = 0;
          p2p_on        = false;
        }
      if (comm)
        Nccl::get().CommDestroy(comm);
      comm = nullptr;
      if (comm_stream)
        cudaStreamDestroy(comm_stream);
      if (ev_fork)
        cudaEventDestroy(ev_fork);
      if (ev_join)
        cudaEventDestroy(ev_join);
      comm_stream = nullptr;
      ev_fork = ev_join = nullptr;
      if (scal_dev)
        cudaFree(scal_dev);
      if (scal_host)
        cudaFreeHost(scal_host);
      for (const Range &r : open_ranges)
        {
          cudaEventDestroy(r.a);
          cudaEventDestroy(r.b);
        }
      open_ranges.clear();
      for (cudaEvent_t e : event_pool)
        cudaEventDestroy(e);
      event_pool.clear();
      for (int i = 0; i < 2; ++i)
        if (timer_ev[i])
          cudaEventDestroy(timer_ev[i]);
      timer_ev[0] = timer_ev[1] = nullptr;
      if (stream)
        cudaStreamDestroy(stream);
      scal_dev = scal_host = nullptr;
      stream               = nullptr;
    }
    void
    sync()
    {
      MGB_CUDA_CHECK(cudaStreamSynchronize(stream));
    }
    void
    check_launch()
    {
      ++launches;
      MGB_CUDA_CHECK(cudaGetLastError());
    }
  };

  // kinds of timed ranges (per level).  0..5 are the columns of print_wall_times
  // (multigrid_solver.h:655-679,776); the rest are kernel classes
  enum RangeKind
  {
    RK_MG_MV = 0,
    RK_RESTRICT,
    RK_PROLONGATE,
    RK_INHOM_BC,
    RK_MG_VEC,
    RK_SMOOTHER,
    RK_APPLY_F64,       // cell-loop kernel, double
    RK_APPLY_F32,       // cell-loop kernel, float, set/add/residual
    RK_APPLY_F32_CHEB,  // cell-loop kernel, float, fused Chebyshev update
    RK_APPLY_EPILOGUE,  // patch-boundary epilogue kernel
    RK_CG_VEC,          // outer CG vector updates and dot products
    RK_OTHER
  };

  struct ScopedRange
  {
    Ctx &ctx;
    int  h;
    ScopedRange(Ctx &c, const int level, const int kind)
      : ctx(c)
      , h(c.range_begin(level, kind))
    {}
    ~ScopedRange() { ctx.range_end(h); }
  };

  template <typename T>
  struct DVec
  {
    T     *p = nullptr;
    size_t n = 0;
    DVec()   = default;
    DVec(const DVec &) = delete;
    DVec &operator=(const DVec &) = delete;
    DVec(DVec &&o) noexcept
      : p(o.p)
      , n(o.n)
    {
      o.p = nullptr;
      o.n = 0;
    }
    ~DVec() { release(); }
    void
    alloc(const size_t count, cudaStream_t stream)
    {
      release();
      n = count;
      if (n)
        {
          MGB_CUDA_CHECK(cudaMalloc(&p, n * sizeof(T)));
          MGB_CUDA_CHECK(cudaMemsetAsync(p, 0, n * sizeof(T), stream));
        }
    }
    void
    release()
    {
      if (p)
        cudaFree(p);
      p = nullptr;
      n = 0;
    }
    void
    swap(DVec &o)
    {
      std::swap(p, o.p);
      std::swap(n, o.n);
    }
  };

  // ------------------------------------------------------------------ generic vector kernels
  template <typename F>
  __global__ void
  k_foreach(const size_t n, F f)
  {
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
      f(i);
  }

  static inline unsigned
  grid_for(const size_t n, const int block = 256)
  {
    const size_t g = (n + block - 1) / block;
    return (unsigned)std::min<size_t>(std::max<size_t>(g, 1), 148 * 16);
  }

  template <typename F>
  void
  foreach (Ctx &ctx, const size_t n, F f)
  {
    if (n == 0)
      return;
    k_foreach<<<grid_for(n), 256, 0, ctx.stream>>>(n, f);
    ctx.check_launch();
  }

  // up to 4 simultaneous sums of products; warp-shuffle + shared reduction, one red.add.f64 per block
  template <int NS, typename F>
  __global__ void
  k_reduce(const size_t off, const size_t n, F f, double *out)
  {
    double acc[NS];
#pragma unroll
    for (int s = 0; s < NS; ++s)
      acc[s] = 0;
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
      f(i + off, acc);
    __shared__ double red[NS][8];
#pragma unroll
    for (int s = 0; s < NS; ++s)
      {
        double v = acc[s];
        for (int o = 16; o > 0; o >>= 1)
          v += __shfl_down_sync(0xffffffffu, v, o);
        if ((threadIdx.x & 31) == 0)
          red[s][threadIdx.x >> 5] = v;
      }
    __syncthreads();
    if (threadIdx.x < NS)
      {
        double v = 0;
        for (int w = 0; w < (int)(blockDim.x >> 5); ++w)
          v += red[threadIdx.x][w];
        atomicAdd(&out[threadIdx.x], v);
      }
  }

  // The entries of a vector that enter sums: on a distributed level the owned planes of the z-slab
  // (one contiguous range), summed over the ranks; otherwise the whole vector.  Element-wise updates
  // (foreach) run over all local entries instead: ghost entries are never trusted, every consumer
  // (operator apply, prolongation) refreshes them first.
  struct Span
  {
    size_t off = 0, n = 0;
    bool   all_ranks = false;
    Span() = default;
    Span(const size_t count)
      : off(0)
      , n(count)
      , all_ranks(false)
    {}
    Span(const size_t o, const size_t count, const bool ar)
      : off(o)
      , n(count)
      , all_ranks(ar)
    {}
  };

  // returns the sums on the host (synchronises the stream)
  template <int NS, typename F>
  void
  reduce(Ctx &ctx, const Span sp, F f, double *result)
  {
    MGB_CUDA_CHECK(cudaMemsetAsync(ctx.scal_dev, 0, NS * sizeof(double), ctx.stream));
    k_reduce<NS><<<grid_for(sp.n), 256, 0, ctx.stream>>>(sp.off, sp.n, f, ctx.scal_dev);
    ctx.check_launch();
    if (sp.all_ranks)
      ctx.all_reduce_sum(ctx.scal_dev, NS);
    MGB_CUDA_CHECK(cudaMemcpyAsync(
      ctx.scal_host, ctx.scal_dev, NS * sizeof(double), cudaMemcpyDeviceToHost, ctx.stream));
    ctx.sync();
    ctx.check_p2p_error();
    for (int s = 0; s < NS; ++s)
      result[s] = ctx.scal_host[s];
  }

  template <typename A, typename B>
  double
  dot(Ctx &ctx, const A *a, const B *b, const Span sp)
  {
    double r;
    reduce<1>(
      ctx, sp, [=] __device__(size_t i, double *acc) { acc[0] += (double)a[i] * (double)b[i]; }, &r);
    return r;
  }

  template <typename A, typename B>
  void
  convert(Ctx &ctx, A *dst, const B *src, const size_t n)
  {
    foreach (ctx, n, [=] __device__(size_t i) { dst[i] = (A)src[i]; });
  }

  // ------------------------------------------------------------------ level (device side)
  struct Level
  {
    HostLevel      host;
    LevelDesc      d;
    Slab           slab;
    ExchangePlan   plan;
    Ctx           *ctx   = nullptr;
    int            index = 0; // multigrid level number
    uint64_t       n_global = 0; // DoFs of the whole level (m() of the operator)
    DVec<uint32_t> constrained;
    DVec<double>   plane_buf; // one z-plane: receive buffer of compress_add()

    void
    upload(Ctx &c)
    {
      ctx = &c;
      constrained.alloc(host.constrained.size(), c.stream);
      if (constrained.n)
        MGB_CUDA_CHECK(cudaMemcpyAsync(constrained.p, host.constrained.data(), constrained.n * 4, cudaMemcpyHostToDevice, c.stream));
      MGB_CUDA_CHECK(cudaStreamSynchronize(c.stream));
      d        = host.d;
      plan     = make_exchange_plan(d);
      n_global = (uint64_t)d.nd[0] * d.nd[1] * d.nd2_all;
      if (slab.distributed)
        plane_buf.alloc(plane_pitch(d), c.stream);
    }

    // the entries that enter dot products and norms
    Span owned() const { return Span(plan.owned[0], plan.owned[1], slab.distributed); }

    // LinearAlgebra::distributed::Vector::update_ghost_values (laplace_operator.h:590,812): the p planes
    // below the slab come from the rank below, the plane on top from the rank above; one NCCL group of
    // at most two sends and two receives on the solver stream, contiguous ranges thanks to the
    // plane-major numbering (level.h)
    template <typename T>
    void
    update_ghosts(T *v, cudaStream_t on = nullptr) const
    {
      if (!slab.distributed)
        return;
      if (ctx->p2p_on)
        {
          p2p_exchange(v + plan.send_up[0], d.ztop ? plan.send_up[1] : 0, v + plan.send_dn[0], d.zg ? plan.send_dn[1] : 0,
                       v + plan.recv_dn[0], d.zg ? plan.recv_dn[1] : 0, v + plan.recv_up[0], d.ztop ? plan.recv_up[1] : 0);
          return;
        }
      const cudaStream_t stream = on ? on : ctx->stream;
      Nccl &n = Nccl::get();
      n.check(n.GroupStart(), "ncclGroupStart");
      if (d.ztop)
        {
          n.check(n.Send(v + plan.send_up[0], plan.send_up[1] * sizeof(T), ncclInt8, ctx->rank + 1, ctx->comm, stream), "ncclSend");
          n.check(n.Recv(v + plan.recv_up[0], plan.recv_up[1] * sizeof(T), ncclInt8, ctx->rank + 1, ctx->comm, stream), "ncclRecv");
        }
      if (d.zg)
        {
          n.check(n.Send(v + plan.send_dn[0], plan.send_dn[1] * sizeof(T), ncclInt8, ctx->rank - 1, ctx->comm, stream), "ncclSend");
          n.check(n.Recv(v + plan.recv_dn[0], plan.recv_dn[1] * sizeof(T), ncclInt8, ctx->rank - 1, ctx->comm, stream), "ncclRecv");
        }
      n.check(n.GroupEnd(), "ncclGroupEnd");
      ++ctx->exchanges;
    }

    // One exchange over peer memory (see namespace p2p): `n_up` values at `send_up` are stored into the
    // receive buffer of the rank above, `n_dn` values at `send_dn` into that of the rank below; then the
    // message from below (`n_below` values) is copied to `from_below` and the one from above to `from_above`.
    // A count of zero means "no message on this channel".  The message counters advance identically on both
    // ends of a channel because all ranks issue the same sequence of exchanges.
    template <typename T>
    void
    p2p_exchange(const T *send_up, const uint64_t n_up, const T *send_dn, const uint64_t n_dn, T *from_below,
                 const uint64_t n_below, T *from_above, const uint64_t n_above) const
    {
      const bool do_send_up = n_up > 0, do_send_dn = n_dn > 0, do_recv_below = n_below > 0, do_recv_above = n_above > 0;
      Ctx      &c = *ctx;
      const int r = c.rank;
      auto flag   = [&](const int rank, const int f) { return reinterpret_cast<volatile uint32_t *>(c.flag_at(rank, f)); };
      auto launch = [&](const p2p::Msg &a, const p2p::Msg &b) {
        const size_t words = a.words + b.words;
        if (words == 0)
          return;
        const unsigned blocks = (unsigned)std::min<size_t>(std::max<size_t>(words / 8192, 1), 148);
        p2p::k_p2p<<<blocks, 256, 0, c.stream>>>(a, b, c.p2p_done, c.p2p_error);
        c.check_launch();
      };
      p2p::Msg up, dn, fb, fa;
      if (do_send_up)
        {
          up.src   = reinterpret_cast<const uint32_t *>(send_up);
          up.words = n_up * sizeof(T) / 4;
          up.dst   = reinterpret_cast<uint32_t *>(c.buffer_at(r + 1, 0, c.n_up_sent));
          if (c.n_up_sent >= 2)
            {
              up.wait_flag  = flag(r, p2p::F_UP_ACK);
              up.wait_value = c.n_up_sent - 1;
            }
          up.signal_flag  = flag(r + 1, p2p::F_UP_DATA);
          up.signal_value = ++c.n_up_sent;
        }
      if (do_send_dn)
        {
          dn.src   = reinterpret_cast<const uint32_t *>(send_dn);
          dn.words = n_dn * sizeof(T) / 4;
          dn.dst   = reinterpret_cast<uint32_t *>(c.buffer_at(r - 1, 1, c.n_dn_sent));
          if (c.n_dn_sent >= 2)
            {
              dn.wait_flag  = flag(r, p2p::F_DN_ACK);
              dn.wait_value = c.n_dn_sent - 1;
            }
          dn.signal_flag  = flag(r - 1, p2p::F_DN_DATA);
          dn.signal_value = ++c.n_dn_sent;
        }
      if (up.words * 4 > c.p2p_buf_bytes || dn.words * 4 > c.p2p_buf_bytes)
        throw std::runtime_error("ghost message larger than the peer-to-peer buffer");
      launch(up, dn);
      if (do_recv_below)
        {
          fb.src          = reinterpret_cast<const uint32_t *>(c.buffer_at(r, 0, c.n_up_recv));
          fb.dst          = reinterpret_cast<uint32_t *>(from_below);
          fb.words        = n_below * sizeof(T) / 4;
          fb.wait_flag    = flag(r, p2p::F_UP_DATA);
          fb.wait_value   = c.n_up_recv + 1;
          fb.signal_flag  = flag(r - 1, p2p::F_UP_ACK);
          fb.signal_value = ++c.n_up_recv;
        }
      if (do_recv_above)
        {
          fa.src          = reinterpret_cast<const uint32_t *>(c.buffer_at(r, 1, c.n_dn_recv));
          fa.dst          = reinterpret_cast<uint32_t *>(from_above);
          fa.words        = n_above * sizeof(T) / 4;
          fa.wait_flag    = flag(r, p2p::F_DN_DATA);
          fa.wait_value   = c.n_dn_recv + 1;
          fa.signal_flag  = flag(r + 1, p2p::F_DN_ACK);
          fa.signal_value = ++c.n_dn_recv;
        }
      launch(fb, fa);
      ++c.exchanges;
    }

    // Vector::compress(VectorOperation::add) for the one ghost plane that is ever written (the top plane,
    // by the restriction): it is added to the first owned plane of the rank above
    template <typename T>
    void
    compress_add(T *v)
    {
      if (!slab.distributed)
        return;
      T *buf = reinterpret_cast<T *>(plane_buf.p);
      if (ctx->p2p_on)
        {
          // upward channel only: the plane on top travels to the rank above
          p2p_exchange<T>(v + plan.recv_up[0], d.ztop ? plan.recv_up[1] : 0, nullptr, 0, buf, d.zg ? plan.send_dn[1] : 0,
                          nullptr, 0);
          if (d.zg)
            {
              T *own = v + plan.send_dn[0];
              foreach (*ctx, plan.send_dn[1], [=] __device__(size_t i) { own[i] += buf[i]; });
            }
          return;
        }
      Nccl &n   = Nccl::get();
      n.check(n.GroupStart(), "ncclGroupStart");
      if (d.ztop)
        n.check(n.Send(v + plan.recv_up[0], plan.recv_up[1] * sizeof(T), ncclInt8, ctx->rank + 1, ctx->comm, ctx->stream), "ncclSend");
      if (d.zg)
        n.check(n.Recv(buf, plan.send_dn[1] * sizeof(T), ncclInt8, ctx->rank - 1, ctx->comm, ctx->stream), "ncclRecv");
      n.check(n.GroupEnd(), "ncclGroupEnd");
      ++ctx->exchanges;
      if (d.zg)
        {
          T *own = v + plan.send_dn[0];
          foreach (*ctx, plan.send_dn[1], [=] __device__(size_t i) { own[i] += buf[i]; });
        }
    }
  };

  // LinearAlgebra::distributed::Vector::compress(VectorOperation::add) in full (the C-ABI entry behind
  // multigrid::DeviceVector::compress): BOTH ghost ranges travel to their owners and are added there -- the
  // plane on top into the first owned plane of the rank above, the p planes below into the top p owned planes
  // of the rank below -- and the ghost entries are zeroed afterwards, as deal.II leaves them.
  template <typename T>
  void
  compress_add_all(Level &l, T *v)
  {
    if (!l.slab.distributed)
      return;
    Ctx               &c    = *l.ctx;
    const ExchangePlan &pl  = l.plan;
    const bool         up = l.d.ztop != 0, dn = l.d.zg != 0;
    DVec<T>            buf_up, buf_dn; // what arrives from below (for send_dn range) / from above (for send_up range)
    if (dn)
      buf_up.alloc(pl.send_dn[1], c.stream);
    if (up)
      buf_dn.alloc(pl.send_up[1], c.stream);
    Nccl &n = Nccl::get();
    n.check(n.GroupStart(), "ncclGroupStart");
    if (up)
      {
        n.check(n.Send(v + pl.recv_up[0], pl.recv_up[1] * sizeof(T), ncclInt8, c.rank + 1, c.comm, c.stream), "ncclSend");
        n.check(n.Recv(buf_dn.p, pl.send_up[1] * sizeof(T), ncclInt8, c.rank + 1, c.comm, c.stream), "ncclRecv");
      }
    if (dn)
      {
        n.check(n.Send(v + pl.recv_dn[0], pl.recv_dn[1] * sizeof(T), ncclInt8, c.rank - 1, c.comm, c.stream), "ncclSend");
        n.check(n.Recv(buf_up.p, pl.send_dn[1] * sizeof(T), ncclInt8, c.rank - 1, c.comm, c.stream), "ncclRecv");
      }
    n.check(n.GroupEnd(), "ncclGroupEnd");
    ++c.exchanges;
    if (dn)
      {
        T *own = v + pl.send_dn[0], *b = buf_up.p, *ghost = v + pl.recv_dn[0];
        foreach (c, pl.send_dn[1], [=] __device__(size_t i) { own[i] += b[i]; });
        foreach (c, pl.recv_dn[1], [=] __device__(size_t i) { ghost[i] = T(0); });
      }
    if (up)
      {
        T *own = v + pl.send_up[0], *b = buf_dn.p, *ghost = v + pl.recv_up[0];
        foreach (c, pl.send_up[1], [=] __device__(size_t i) { own[i] += b[i]; });
        foreach (c, pl.recv_up[1], [=] __device__(size_t i) { ghost[i] = T(0); });
      }
    c.sync(); // the temporaries are released on return
  }

  template <typename T>
  void
  zero_ghosts(Level &l, T *v)
  {
    if (!l.slab.distributed)
      return;
    Ctx &c = *l.ctx;
    if (l.d.zg)
      {
        T *ghost = v + l.plan.recv_dn[0];
        foreach (c, l.plan.recv_dn[1], [=] __device__(size_t i) { ghost[i] = T(0); });
      }
    if (l.d.ztop)
      {
        T *ghost = v + l.plan.recv_up[0];
        foreach (c, l.plan.recv_up[1], [=] __device__(size_t i) { ghost[i] = T(0); });
      }
  }

  struct Nodes1D
  {
    double nodes[9], xq[9], w[9];
  };

  // visits every grid point once: f(dof, constrained, gi, gj, gk)
  template <typename F>
  __global__ void
  k_grid_points(const LevelDesc L, F f)
  {
    const size_t n = L.n_dofs;
    for (size_t l = blockIdx.x * (size_t)blockDim.x + threadIdx.x; l < n; l += (size_t)gridDim.x * blockDim.x)
      {
        const int    gi = l % L.nd[0];
        const size_t r  = l / L.nd[0];
        const int    gj = r % L.nd[1], gk = r / L.nd[1];
        f(dof_index(L, gi, gj, gk), is_constrained(L, gi, gj, gk), gi, gj, gk);
      }
  }

  template <typename F>
  void
  patch_points(Ctx &ctx, const LevelDesc &L, F f)
  {
    k_grid_points<<<grid_for(L.n_dofs), 256, 0, ctx.stream>>>(L, f);
    ctx.check_launch();
  }

  __device__ __forceinline__ double
  support_point(const LevelDesc &L, const Nodes1D &nd, const int dir, const int gi)
  {
    int cell = gi / L.p;
    if (cell > L.ncell[dir] - 1)
      cell = L.ncell[dir] - 1;
    return L.lo[dir] + (cell + nd.nodes[gi - cell * L.p]) * L.h[dir];
  }

  // mesh coordinate in [0,1] of grid point gi in direction dir (z: of the whole level, not of the z-slab)
  __device__ __forceinline__ double
  unit_coordinate(const LevelDesc &L, const Nodes1D &nd, const int dir, const int gi)
  {
    int cell = gi / L.p;
    if (cell > L.ncell[dir] - 1)
      cell = L.ncell[dir] - 1;
    const double node = nd.nodes[gi - cell * L.p];
    if (dir == 2)
      return (cell + L.gk0 / L.p + node) * L.p / (L.nd2_all - 1);
    return (cell + node) / L.ncell[dir];
  }

  // ------------------------------------------------------------------ operator
  template <typename T>
  struct LaplaceOperator
  {
    Ctx                 *ctx   = nullptr;
    Level               *level = nullptr;
    const DegreeKernels *dk    = nullptr;
    T                    coef[6];
    DVec<T>              coef_q;
    bool                 cartesian = true; // axis-aligned cells and constant coefficient
    int                  kernel    = MGB_KERNEL_AUTO;
    DVec<T>              inverse_diagonal;
    DVec<T>              tmp;          // full-length scratch (general kernel epilogues)

    size_t m() const { return level->d.n_dofs; }          // local vector length (owned + ghost planes)
    size_t m_global() const { return level->n_global; } // MatrixFreeOperators::Base::m()

    void
    initialize(Ctx &c, Level &l, const Basis &basis, const int coefficient_fn, const Geometry &geom)
    {
      ctx   = &c;
      level = &l;
      dk    = &degree_kernels(l.d.p);
      evaluate_coefficient(basis, coefficient_fn, geom);
    }

    // laplace_operator.h:359-432 (merged coefficient = c JxW J^-1 J^-T)
    void
    evaluate_coefficient(const Basis &basis, const int fn, const Geometry &geom)
    {
      const LevelDesc &L    = level->d;
      const double     detJ = L.h[0] * L.h[1] * L.h[2];
      if (geom.kind == MGB_GEOMETRY_AFFINE && fn == FN_ONE)
        {
          // affine cells, constant coefficient: one tensor for all cells (:374-387), here with all six entries.
          // J = A diag(h): coef = |det J| J^-1 J^-T without the quadrature weight
          cartesian = false;
          double J[3][3], I[3][3];
          for (int d = 0; d < 3; ++d)
            for (int e = 0; e < 3; ++e)
              J[d][e] = geom.par[3 * d + e] * L.h[e];
          const double det = J[0][0] * (J[1][1] * J[2][2] - J[1][2] * J[2][1]) - J[0][1] * (J[1][0] * J[2][2] - J[1][2] * J[2][0]) +
                             J[0][2] * (J[1][0] * J[2][1] - J[1][1] * J[2][0]);
          I[0][0] = (J[1][1] * J[2][2] - J[1][2] * J[2][1]) / det;
          I[0][1] = (J[0][2] * J[2][1] - J[0][1] * J[2][2]) / det;
          I[0][2] = (J[0][1] * J[1][2] - J[0][2] * J[1][1]) / det;
          I[1][0] = (J[1][2] * J[2][0] - J[1][0] * J[2][2]) / det;
          I[1][1] = (J[0][0] * J[2][2] - J[0][2] * J[2][0]) / det;
          I[1][2] = (J[0][2] * J[1][0] - J[0][0] * J[1][2]) / det;
          I[2][0] = (J[1][0] * J[2][1] - J[1][1] * J[2][0]) / det;
          I[2][1] = (J[0][1] * J[2][0] - J[0][0] * J[2][1]) / det;
          I[2][2] = (J[0][0] * J[1][1] - J[0][1] * J[1][0]) / det;
          auto m = [&](const int e, const int g) {
            return std::fabs(det) * (I[e][0] * I[g][0] + I[e][1] * I[g][1] + I[e][2] * I[g][2]);
          };
          coef[0] = (T)m(0, 0);
          coef[1] = (T)m(1, 1);
          coef[2] = (T)m(2, 2);
          coef[3] = (T)m(0, 1);
          coef[4] = (T)m(0, 2);
          coef[5] = (T)m(1, 2);
          return;
        }
      if (geom.kind != MGB_GEOMETRY_BOX)
        {
          // general cells: one tensor per quadrature point from the Jacobian of the polynomial cell geometry (:388-430)
          cartesian   = false;
          const int n = L.n, n3 = n * n * n;
          coef_q.alloc((size_t)L.n_cells * 6 * n3, ctx->stream);
          CoefArgs<T> A;
          A.L      = L;
          A.geom   = geom;
          A.coef_q = coef_q.p;
          A.fn_id  = fn;
          if (sizeof(T) == 8)
            dk->coefficient_d(reinterpret_cast<const CoefArgs<double> &>(A), ctx->stream);
          else
            dk->coefficient_f(reinterpret_cast<const CoefArgs<float> &>(A), ctx->stream);
          ctx->check_launch();
          return;
        }
      if (fn == FN_ONE)
        {
          cartesian = true;
          for (int d = 0; d < 3; ++d)
            coef[d] = (T)(detJ / (L.h[d] * L.h[d]));
          coef[3] = coef[4] = coef[5] = 0;
          return;
        }
      cartesian   = false;
      const int n = L.n, n3 = n * n * n;
      coef_q.alloc((size_t)L.n_cells * 6 * n3, ctx->stream);
      Nodes1D nd;
      for (int i = 0; i < n; ++i)
        {
          nd.xq[i] = basis.xq[i];
          nd.w[i]  = basis.wq[i];
        }
      T              *cq = coef_q.p;
      const LevelDesc Lc = L;
      foreach (*ctx, (size_t)L.n_cells * n3, [=] __device__(size_t idx) {
        const uint32_t cell = idx / n3;
        const int      q    = idx - (size_t)cell * n3;
        const int      q0 = q % n, q1 = (q / n) % n, q2 = q / (n * n);
        int            cx, cy, cz;
        cell_coords(Lc, cell, cx, cy, cz);
        const double x = Lc.lo[0] + (cx + nd.xq[q0]) * Lc.h[0];
        const double y = Lc.lo[1] + (cy + nd.xq[q1]) * Lc.h[1];
        const double z = Lc.lo[2] + (cz + nd.xq[q2]) * Lc.h[2];
        const double f = eval_fn(fn, x, y, z) * (nd.w[q0] * nd.w[q1] * nd.w[q2]) * detJ;
        T           *c = cq + (size_t)cell * 6 * n3 + q;
        c[0 * n3]      = (T)(f / (Lc.h[0] * Lc.h[0]));
        c[1 * n3]      = (T)(f / (Lc.h[1] * Lc.h[1]));
        c[2 * n3]      = (T)(f / (Lc.h[2] * Lc.h[2]));
        c[3 * n3] = c[4 * n3] = c[5 * n3] = T(0);
      });
    }

    // evaluate_coefficient(const Function &) with the function sampled by the caller (laplace_operator.h:108-109,
    // :388-430): c[cell][q] at the quadrature points; the geometry part JxW J^-1 J^-T comes from the mesh
    void
    set_coefficient_values(const Basis &basis, const Geometry &geom, const double *c_dev)
    {
      const LevelDesc &L  = level->d;
      const int        n3 = L.n * L.n * L.n;
      evaluate_coefficient(basis, FN_ONE_PER_POINT, geom); // per-point tensors of the geometry alone
      cartesian   = false;
      T *cq       = coef_q.p;
      foreach (*ctx, (size_t)L.n_cells * n3, [=] __device__(size_t idx) {
        const size_t cell = idx / n3, q = idx - cell * n3;
        const T      c    = (T)c_dev[idx];
        T           *e    = cq + cell * 6 * n3 + q;
        for (int k = 0; k < 6; ++k)
          e[k * n3] *= c;
      });
    }

    // MinimalSurfaceOperator::evaluate_coefficient(first_time, solution) (minimal_surface/program.cc:117-163): the merged
    // coefficient from the gradient of the level vector u (mode 0 identity, 1 Newton linearisation, 2 residual flux)
    void
    evaluate_solution_coefficient(const Geometry &geom, const int mode, const T *u)
    {
      const LevelDesc &L  = level->d;
      const size_t     nc = (size_t)L.n_cells * 6 * L.n * L.n * L.n;
      cartesian           = false;
      if (coef_q.n != nc)
        coef_q.alloc(nc, ctx->stream);
      SolutionCoefArgs<T> A;
      A.L      = L;
      A.geom   = geom;
      A.coef_q = coef_q.p;
      A.u      = u;
      A.mode   = mode;
      if (sizeof(T) == 8)
        dk->solution_coefficient_d(reinterpret_cast<const SolutionCoefArgs<double> &>(A), ctx->stream);
      else
        dk->solution_coefficient_f(reinterpret_cast<const SolutionCoefArgs<float> &>(A), ctx->stream);
      ctx->check_launch();
    }

    bool use_cartesian() const
    {
      // the streaming kernel is compiled for the default tile width of the degree
      return cartesian && kernel == MGB_KERNEL_AUTO && level->d.CW == choose_tile_cells(level->d.p, 0);
    }

    void
    fill_args(ApplyArgs<T> &A, const T *src, T *dst) const
    {
      A.L               = level->d;
      A.src             = src;
      A.dst             = dst;
      A.rhs             = nullptr;
      A.inv_diag        = nullptr;
      A.coef_q          = coef_q.p;
      for (int i = 0; i < 6; ++i)
        A.coef[i] = coef[i];
      A.f1 = A.f2       = 0;
      A.f0              = 0;
      A.mode            = APPLY_ADD;
      A.read_plain      = 0;
      A.first_chebyshev = 0;
      A.layer_begin = A.layer_end = -1;
      A.layer_begin2 = -1;
      A.dot_out     = nullptr;
    }

    int
    range_kind(const int mode) const
    {
      return sizeof(T) == 8 ? RK_APPLY_F64 : (mode == APPLY_CHEBYSHEV ? RK_APPLY_F32_CHEB : RK_APPLY_F32);
    }

    void
    launch_general(const ApplyArgs<T> &A) const
    {
      ScopedRange range(*ctx, level->index, range_kind(A.mode));
      if (sizeof(T) == 8)
        dk->apply_general_d(reinterpret_cast<const ApplyArgs<double> &>(A), ctx->stream);
      else
        dk->apply_general_f(reinterpret_cast<const ApplyArgs<float> &>(A), ctx->stream);
      ctx->check_launch();
    }

    void
    launch_cartesian(const ApplyArgs<T> &A) const
    {
      if (sizeof(T) == 8)
        dk->apply_cartesian_d(reinterpret_cast<const ApplyArgs<double> &>(A), ctx->stream);
      else
        dk->apply_cartesian_f(reinterpret_cast<const ApplyArgs<float> &>(A), ctx->stream);
      ctx->check_launch();
    }

    // the one entry point behind vmult / vmult_add / vmult_residual / Chebyshev step
    void
    apply(const int mode, T *dst, const T *src, const T *rhs = nullptr, const T *inv_diag = nullptr,
          const T f1 = 0, const T f2 = 0, const int first = 0, const bool read_plain = false,
          double *dot_out = nullptr, const T f0 = 0)
    {
      ApplyArgs<T> A;
      fill_args(A, src, dst);
      A.dot_out         = dot_out;
      A.mode            = mode;
      A.rhs             = rhs;
      A.inv_diag        = inv_diag;
      A.f1              = f1;
      A.f2              = f2;
      A.f0              = f0;
      A.first_chebyshev = first;
      A.read_plain      = read_plain;
      // zero-guess Chebyshev (first == 2): the source x_1 = f0 P^-1 rhs is formed inside the tile kernel, so it is
      // the ghost planes of rhs that have to be current (those of P^-1 are, compute_diagonal)
      if (first == 2)
        src = rhs;
      // update_ghost_values of the source (laplace_operator.h:590): like the reference's vmult, this
      // mutates the ghost range of the const source vector
      if (use_cartesian())
        {
          // streaming tile kernel: the vector operation is fused, every DoF is written once
          ScopedRange      range(*ctx, level->index, range_kind(A.mode));
          const LevelDesc &L  = level->d;
          const int        i0 = L.zg ? 2 : 0, i1 = L.ztop ? L.ncell[2] - 1 : L.ncell[2];
          if (level->slab.distributed && ctx->overlap && !ctx->p2p_on && i1 > i0)
            {
              // ghost exchange on the comm stream while the layers that read no ghost plane are
              // computed; then the first owned layer (reads the p planes from below through the carry
              // of the ghost layer) and the last one (reads the plane on top)
              ctx->fork_comm();
              level->update_ghosts(const_cast<T *>(src), ctx->comm_stream);
              A.layer_begin = i0;
              A.layer_end   = i1;
              launch_cartesian(A);
              ctx->join_comm();
              if (L.zg && L.ztop)
                {
                  // both boundary layers in one launch
                  A.layer_begin  = 1;
                  A.layer_end    = 2;
                  A.layer_begin2 = L.ncell[2] - 1;
                  launch_cartesian(A);
                }
              else if (L.zg)
                {
                  A.layer_begin = 1;
                  A.layer_end   = 2;
                  launch_cartesian(A);
                }
              else if (L.ztop)
                {
                  A.layer_begin = L.ncell[2] - 1;
                  A.layer_end   = L.ncell[2];
                  launch_cartesian(A);
                }
              return;
            }
          level->update_ghosts(const_cast<T *>(src));
          launch_cartesian(A);
          return;
        }
      level->update_ghosts(const_cast<T *>(src));
      const LevelDesc &L = level->d;
      const size_t     N = L.n_dofs;

      // general kernel: accumulate K src with red.global.add, then the vector operation
      if (mode == APPLY_ADD)
        {
          launch_general(A);
          return;
        }
      T *acc = dst;
      if (mode != APPLY_SET)
        {
          if (tmp.n != N)
            tmp.alloc(N, ctx->stream);
          acc = tmp.p;
        }
      MGB_CUDA_CHECK(cudaMemsetAsync(acc, 0, N * sizeof(T), ctx->stream));
      A.dst = acc;
      launch_general(A);
      const uint32_t *clist = level->constrained.p;
      const size_t    nc    = level->constrained.n;
      if (mode == APPLY_SET)
        {
          foreach (*ctx, nc, [=] __device__(size_t i) { dst[clist[i]] = src[clist[i]]; });
          return;
        }
      foreach (*ctx, nc, [=] __device__(size_t i) { acc[clist[i]] = src[clist[i]]; });
      if (mode == APPLY_RESIDUAL)
        foreach (*ctx, N, [=] __device__(size_t d) { dst[d] = rhs[d] - acc[d]; });
      else
        foreach (*ctx, N, [=] __device__(size_t d) {
          const T upd = f2 * inv_diag[d] * (rhs[d] - acc[d]);
          dst[d] = first != 0 ? (T(1) + f1) * src[d] + upd : (T(1) + f1) * src[d] - f1 * dst[d] + upd;
        });
    }

    void vmult(T *dst, const T *src) { apply(APPLY_SET, dst, src); }
    // dst = A src and the dot product src . dst (first sum of vmult_with_cg_update, laplace_operator.h:698)
    void
    vmult_dot(T *dst, const T *src, double *result)
    {
      if (sizeof(T) == 8 && use_cartesian() && fuse_dot)
        {
          // the tile kernel adds src[i] dst[i] of the points it stores (exactly the owned ones) to a device scalar
          MGB_CUDA_CHECK(cudaMemsetAsync(ctx->scal_dev, 0, sizeof(double), ctx->stream));
          apply(APPLY_SET, dst, src, nullptr, nullptr, 0, 0, false, false, ctx->scal_dev);
          if (level->owned().all_ranks)
            ctx->all_reduce_sum(ctx->scal_dev, 1);
          MGB_CUDA_CHECK(cudaMemcpyAsync(ctx->scal_host, ctx->scal_dev, sizeof(double), cudaMemcpyDeviceToHost, ctx->stream));
          ctx->sync();
          ctx->check_p2p_error();
          *result = ctx->scal_host[0];
          return;
        }
      apply(APPLY_SET, dst, src);
      ScopedRange range(*ctx, level->index, RK_CG_VEC);
      *result = dot(*ctx, src, dst, level->owned());
    }
    bool fuse_dot = getenv("MGB200_FUSE_DOT") == nullptr || getenv("MGB200_FUSE_DOT")[0] != '0'; // A/B switch
    void vmult_add(T *dst, const T *src) { apply(APPLY_ADD, dst, src); }
    void vmult_residual(const T *rhs, const T *lhs, T *residual) { apply(APPLY_RESIDUAL, residual, lhs, rhs); }

    // laplace_operator.h:640-719 (q = A p after the vector update, 4 dot products).  Two passes over the vectors on
    // Cartesian levels: the pre-range update x += alpha p, p = beta p + q as one vector kernel (it must be complete
    // on the halo of every tile before any block reads p, which is why it cannot ride in the apply's load stage
    // without a second copy of q), then the tile kernel stores q = A p and takes the four sums from the values it
    // holds (mode SET_DOT4): 40 + 24 bytes per DoF in double instead of the 96 of update / memset / apply-add / sums.
    void
    vmult_with_cg_update(const T alpha, const T beta, const T *r, T *q, T *pp, T *x, double sums[4])
    {
      const size_t N = m();
      if (alpha == T(0))
        foreach (*ctx, N, [=] __device__(size_t i) { pp[i] = q[i]; });
      else
        foreach (*ctx, N, [=] __device__(size_t i) {
          x[i] += alpha * pp[i];
          pp[i] = beta * pp[i] + q[i];
        });
      if (use_cartesian())
        {
          MGB_CUDA_CHECK(cudaMemsetAsync(ctx->scal_dev, 0, 4 * sizeof(double), ctx->stream));
          apply(APPLY_SET, q, pp, r, nullptr, 0, 0, 0, false, ctx->scal_dev);
          if (level->owned().all_ranks)
            ctx->all_reduce_sum(ctx->scal_dev, 4);
          MGB_CUDA_CHECK(cudaMemcpyAsync(ctx->scal_host, ctx->scal_dev, 4 * sizeof(double), cudaMemcpyDeviceToHost, ctx->stream));
          ctx->sync();
          ctx->check_p2p_error();
          for (int k = 0; k < 4; ++k)
            sums[k] = ctx->scal_host[k];
          return;
        }
      // the reference leaves constrained rows of q at zero here (no identity fix-up)
      MGB_CUDA_CHECK(cudaMemsetAsync(q, 0, N * sizeof(T), ctx->stream));
      apply(APPLY_ADD, q, pp);
      reduce<4>(
        *ctx,
        level->owned(),
        [=] __device__(size_t i, double *acc) {
          const double qi = q[i], pi = pp[i], ri = r[i];
          acc[0] += qi * pi;
          acc[1] += ri * ri;
          acc[2] += qi * ri;
          acc[3] += qi * qi;
        },
        sums);
    }

    // laplace_operator.h:747-800
    void
    compute_diagonal()
    {
      const size_t N = m();
      inverse_diagonal.alloc(N, ctx->stream);
      DiagArgs<T> A;
      A.L      = level->d;
      A.diag   = inverse_diagonal.p;
      A.coef_q = coef_q.p;
      for (int i = 0; i < 6; ++i)
        A.coef[i] = coef[i];
      if (sizeof(T) == 8)
        dk->diagonal_d(reinterpret_cast<const DiagArgs<double> &>(A), ctx->stream);
      else
        dk->diagonal_f(reinterpret_cast<const DiagArgs<float> &>(A), ctx->stream);
      ctx->check_launch();
      T              *dg    = inverse_diagonal.p;
      const uint32_t *clist = level->constrained.p;
      foreach (*ctx, level->constrained.n, [=] __device__(size_t i) { dg[clist[i]] = T(1); });
      foreach (*ctx, N, [=] __device__(size_t i) { dg[i] = T(1) / dg[i]; });
      // the cell loop above ran over the local box: the owned planes are complete (their cells are the
      // owned layers and the ghost layer), the ghost planes are not
      level->update_ghosts(dg);
    }
  };

  // ------------------------------------------------------------------ Chebyshev smoother
  static void
  tridiagonal_extreme_eigenvalues(const std::vector<double> &diag,
                                  const std::vector<double> &off,
                                  double                    &emin,
                                  double                    &emax)
  {
    const int k = (int)diag.size();
    if (k == 0)
      {
        emin = emax = 1.;
        return;
      }
    double glo = 1e300, ghi = -1e300;
    for (int i = 0; i < k; ++i)
      {
        const double r = (i > 0 ? std::fabs(off[i - 1]) : 0.) + (i < k - 1 ? std::fabs(off[i]) : 0.);
        glo            = std::min(glo, diag[i] - r);
        ghi            = std::max(ghi, diag[i] + r);
      }
    auto n_below = [&](const double x) { // Sturm count
      int    cnt = 0;
      double d   = 1;
      for (int i = 0; i < k; ++i)
        {
          d = diag[i] - x - (i > 0 ? off[i - 1] * off[i - 1] / d : 0.);
          if (d == 0)
            d = 1e-300;
          cnt += d < 0;
        }
      return cnt;
    };
    auto bisect = [&](const int idx) {
      double a = glo, c = ghi;
      for (int it = 0; it < 200; ++it)
        {
          const double mid = 0.5 * (a + c);
          (n_below(mid) > idx ? c : a) = mid;
        }
      return 0.5 * (a + c);
    };
    emin = bisect(0);
    emax = bisect(k - 1);
  }

  template <typename T>
  struct Chebyshev
  {
    static constexpr unsigned int invalid_degree = 0xffffffffu;
    Ctx                *ctx = nullptr;
    LaplaceOperator<T> *A   = nullptr;
    double              smoothing_range     = 20.;
    unsigned int        degree              = 1;
    unsigned int        eig_cg_n_iterations = 15;
    double              min_eig = 1, max_eig = 1, theta = 1, delta = 1;
    unsigned int        cg_iterations = 0;
    DVec<T>             solution_old;

    // deal.II PreconditionChebyshev::estimate_eigenvalues: CG/Lanczos on D^-1 A started from
    // "lexicographic index mod 11, mean removed, constrained entries zero"
    void
    initialize(Ctx &c, LaplaceOperator<T> &op, const Basis &)
    {
      ctx            = &c;
      A              = &op;
      const size_t N = op.m();
      solution_old.alloc(N, c.stream);
      DVec<T> g, h, d;
      g.alloc(N, c.stream);
      h.alloc(N, c.stream);
      d.alloc(N, c.stream);
      const LevelDesc L   = op.level->d;
      const Span      own = op.level->owned();
      T              *gp = g.p, *hp = h.p, *dp = d.p;
      const T        *P = op.inverse_diagonal.p;
      patch_points(c, L, [=] __device__(uint32_t dof, bool con, int gi, int gj, int gk) {
        const size_t lex = ((size_t)(gk + L.gk0) * L.nd[1] + gj) * L.nd[0] + gi;
        gp[dof]          = con ? T(0) : (T)(lex % 11);
      });
      // mean over all entries before constraining (cheap closed form avoided: do it on the device)
      double sum;
      {
        // unconstrained pattern sum: recompute with constrained entries included
        DVec<T> full;
        full.alloc(N, c.stream);
        T *fp = full.p;
        patch_points(c, L, [=] __device__(uint32_t dof, bool, int gi, int gj, int gk) {
          const size_t lex = ((size_t)(gk + L.gk0) * L.nd[1] + gj) * L.nd[0] + gi;
          fp[dof]          = (T)(lex % 11);
        });
        reduce<1>(
          c, own, [=] __device__(size_t i, double *acc) { acc[0] += (double)fp[i]; }, &sum);
      }
      const T mean = (T)(sum / op.m_global());
      // g = -(pattern - mean) with constrained entries zero
      foreach (c, N, [=] __device__(size_t i) { gp[i] = -(gp[i] - mean); });
      {
        const uint32_t *clist = op.level->constrained.p;
        foreach (c, op.level->constrained.n, [=] __device__(size_t i) { gp[clist[i]] = T(0); });
      }
      std::vector<double> diag, off;
      double              res = std::sqrt(dot(c, gp, gp, own));
      cg_iterations           = 0;
      if (res > 1e-10 && eig_cg_n_iterations > 0)
        {
          foreach (c, N, [=] __device__(size_t i) {
            hp[i] = P[i] * gp[i];
            dp[i] = -hp[i];
          });
          double gh = dot(c, gp, hp, own);
          double alpha_old = 0, beta_old = 0;
          for (unsigned int it = 1;; ++it)
            {
              op.vmult(hp, dp);
              const double alpha = gh / dot(c, dp, hp, own);
              const T      al    = (T)alpha;
              foreach (c, N, [=] __device__(size_t i) { gp[i] += al * hp[i]; });
              res = std::sqrt(dot(c, gp, gp, own));
              if (it == 1)
                diag.push_back(1. / alpha);
              else
                {
                  diag.push_back(1. / alpha + beta_old / alpha_old);
                  off.push_back(std::sqrt(beta_old) / alpha_old);
                }
              cg_iterations = it;
              if (it >= eig_cg_n_iterations || res <= 1e-10)
                break;
              foreach (c, N, [=] __device__(size_t i) { hp[i] = P[i] * gp[i]; });
              const double gh_new = dot(c, gp, hp, own);
              const double beta   = gh_new / gh;
              gh                  = gh_new;
              const T be          = (T)beta;
              foreach (c, N, [=] __device__(size_t i) { dp[i] = be * dp[i] - hp[i]; });
              alpha_old = alpha;
              beta_old  = beta;
            }
          tridiagonal_extreme_eigenvalues(diag, off, min_eig, max_eig);
        }
      else
        min_eig = max_eig = 1.;

      const double max_est = 1.2 * max_eig;
      const double alpha =
        smoothing_range > 1. ? max_est / smoothing_range : std::min(0.9 * max_est, min_eig);
      if (degree == invalid_degree)
        {
          const double actual_range = max_est / alpha;
          const double sigma =
            (1. - std::sqrt(1. / actual_range)) / (1. + std::sqrt(1. / actual_range));
          const double eps = smoothing_range;
          degree           = 1 + static_cast<unsigned int>(
                         std::log(1. / eps + std::sqrt(1. / eps / eps - 1.)) / std::log(1. / sigma));
        }
      theta = (max_est + alpha) * 0.5;
      delta = (max_est - alpha) * 0.5;
      c.sync();
    }

    // x_{k+1} = (1+f1) x_k - f1 x_{k-1} + f2 P^-1 (b - A x_k), fused with the operator apply;
    // the result lands in solution_old and the two vectors are swapped (as deal.II does)
    void
    fused_step(DVec<T> &x, const T *b, const double f1, const double f2, const int first, const T f0 = 0)
    {
      A->apply(APPLY_CHEBYSHEV, solution_old.p, x.p, b, A->inverse_diagonal.p, (T)f1, (T)f2, first, false, nullptr, f0);
      x.swap(solution_old);
    }

    // `zero_guess`: the iterate x_1 = f0 P^-1 b was not stored (vmult on the packed float tile kernel): step k = 0
    // forms it on the fly, step k = 1 recomputes it as x_{k-1}
    void
    loop(DVec<T> &x, const T *b, const bool first_is_index_one, const bool zero_guess = false)
    {
      if (degree < 2)
        return;
      double       rhok  = delta / theta;
      const double sigma = theta / delta;
      const T      f0    = (T)(1. / theta);
      for (unsigned int k = 0; k < degree - 1; ++k)
        {
          const double rhokp = 1. / (2. * sigma - rhok);
          const double f1 = rhokp * rhok, f2 = 2. * rhokp / delta;
          rhok = rhokp;
          if (zero_guess && k < 2)
            fused_step(x, b, f1, f2, k == 0 ? 2 : 3, f0);
          else
            fused_step(x, b, f1, f2, first_is_index_one && k == 0 ? 1 : 0);
        }
    }

    // MGB200_FUSE_ZERO_GUESS=0: store x_1 with a separate pass as before (A/B measurements)
    // (only the packed float tile kernel has the zero-guess instances: MGB200_F32X2=0 switches them off as well)
    bool fuse_zero_guess = (getenv("MGB200_FUSE_ZERO_GUESS") == nullptr || getenv("MGB200_FUSE_ZERO_GUESS")[0] != '0') &&
                           (getenv("MGB200_F32X2") == nullptr || getenv("MGB200_F32X2")[0] != '0');

    // zero initial guess (deal.II vmult): x = 1/theta P^-1 b, then degree-1 fused steps
    void
    vmult(DVec<T> &x, const T *b)
    {
      T       *xp = x.p;
      const T *P  = A->inverse_diagonal.p;
      const T  f2 = (T)(1. / theta);
      if (std::is_same<T, float>::value && degree >= 2 && fuse_zero_guess && A->use_cartesian())
        {
          // x_1 never touches HBM: 28 instead of 48 bytes per DoF for the three-step pre-smoother
          loop(x, b, true, true);
          return;
        }
      foreach (*ctx, x.n, [=] __device__(size_t i) { xp[i] = (f2 * P[i]) * b[i]; });
      loop(x, b, true);
    }

    // nonzero initial guess (deal.II step): degree fused steps
    void
    step(DVec<T> &x, const T *b)
    {
      fused_step(x, b, 0., 1. / theta, true);
      loop(x, b, false);
    }
  };

  // ------------------------------------------------------------------ transfer
  template <typename T>
  struct Transfer
  {
    Ctx                 *ctx = nullptr;
    Level               *fine = nullptr, *coarse = nullptr;
    const DegreeKernels *dk = nullptr;
    bool                 constrained = false;
    int                  kernel      = MGB_KERNEL_AUTO;

    DVec<T>              partial; // restriction onto a replicated coarse level: this rank's share

    void
    run(const int mode, T *dst, const T *src)
    {
      TransferArgs<T> A;
      A.fine        = fine->d;
      A.coarse      = coarse->d;
      A.src         = src;
      A.dst         = dst;
      A.mode        = mode;
      A.constrained = constrained;
      // z-slabs: this rank handles the parents of its owned fine layers.  Local coarse plane 0 is the
      // global coarse plane gk0_c, whose children start at the global fine plane 2 gk0_c.
      const bool fd = fine->slab.distributed, cd = coarse->slab.distributed;
      A.cz_begin    = cd ? coarse->d.zg : (fd ? fine->slab.c0 / 2 : 0);
      A.cz_end      = cd || !fd ? coarse->d.ncell[2] : fine->slab.c1 / 2;
      A.f_shift     = 2 * coarse->d.gk0 - fine->d.gk0;
      A.close_top   = fine->d.ztop ? 0 : 1;
      // streaming tile kernels are compiled for the default tile width of the degree
      const bool tiles = fine->d.CW == choose_tile_cells(fine->d.p, 0) && kernel == MGB_KERNEL_AUTO;
      if (!tiles && fd)
        throw std::runtime_error("the cell-wise transfer kernel does not support z-slabs; use the default tile width");
      if (sizeof(T) == 8)
        (tiles ? dk->transfer_tile_d : dk->transfer_d)(reinterpret_cast<const TransferArgs<double> &>(A), ctx->stream);
      else
        (tiles ? dk->transfer_tile_f : dk->transfer_f)(reinterpret_cast<const TransferArgs<float> &>(A), ctx->stream);
      ctx->check_launch();
    }
    // the coarse plane on top of the slab is read: update_ghost_values of the coarse vector first
    void
    prolongate(T *f, const T *c)
    {
      coarse->update_ghosts(const_cast<T *>(c));
      run(0, f, c);
    }
    void
    prolongate_and_add(T *f, const T *c)
    {
      coarse->update_ghosts(const_cast<T *>(c));
      run(1, f, c);
    }
    // every rank restricts its owned fine planes; what lands on the coarse plane on top of the slab
    // belongs to the rank above (compress(add)), and a replicated coarse level gets the sum over all ranks
    void
    restrict_and_add(T *c, const T *f)
    {
      if (!fine->slab.distributed)
        {
          run(2, c, f);
          return;
        }
      if (coarse->slab.distributed)
        {
          if (coarse->d.ztop)
            MGB_CUDA_CHECK(cudaMemsetAsync(c + coarse->plan.recv_up[0], 0, coarse->plan.recv_up[1] * sizeof(T), ctx->stream));
          run(2, c, f);
          coarse->compress_add(c);
          return;
        }
      const size_t nc = coarse->d.n_dofs;
      if (partial.n != nc)
        partial.alloc(nc, ctx->stream);
      else
        MGB_CUDA_CHECK(cudaMemsetAsync(partial.p, 0, nc * sizeof(T), ctx->stream));
      run(2, partial.p, f);
      Nccl &n = Nccl::get();
      n.check(n.AllReduce(partial.p, partial.p, nc, sizeof(T) == 8 ? ncclDouble : ncclFloat, ncclSum, ctx->comm, ctx->stream),
              "ncclAllReduce");
      ++ctx->exchanges;
      const T *part = partial.p;
      foreach (*ctx, nc, [=] __device__(size_t i) { c[i] += part[i]; });
    }
  };

  // ------------------------------------------------------------------ solver
  struct MultigridSolver
  {
    mgb_options  opt;
    Ctx          ctx;
    Basis        basis;
    int          n_levels = 0;
    int          fn_solution, fn_rhs, fn_coef;
    Geometry     geom;
    std::vector<std::unique_ptr<Level>>                   levels;
    std::vector<std::unique_ptr<LaplaceOperator<float>>>  matrix;
    std::vector<std::unique_ptr<LaplaceOperator<double>>> matrix_dp;
    std::vector<std::unique_ptr<Chebyshev<float>>>        smooth;
    std::vector<Transfer<float>>                          transfer;
    std::vector<Transfer<double>>                         mg_transfer_no_boundary;
    std::vector<std::unique_ptr<DVec<double>>>            solution, rhs, residual, bc_values;
    std::vector<std::unique_ptr<DVec<float>>>             defect, t, solution_update;
    DVec<double>                                          cg_g, cg_h, cg_d;
    DVec<double>                                          con_scratch; // vmult_with_residual_update: parked identity rows
    std::vector<double>                                   cg_history;

    explicit MultigridSolver(const mgb_options &o)
      : opt(o)
    {
      setup(o);
    }

    // body of the constructor (extended device lambdas are not allowed inside constructors)
    void
    setup(const mgb_options &o)
    {
      if (o.degree < 1 || o.degree > 8)
        throw std::runtime_error("polynomial degree must be between 1 and 8");
      if (o.degree_pre != o.degree_post)
        throw std::runtime_error("Change of pre- and post-smoother degree currently not possible");
      if (o.n_refinements < 0 || o.n_subdivisions[0] < 1 || o.n_subdivisions[1] < 1 ||
          o.n_subdivisions[2] < 1)
        throw std::runtime_error("invalid mesh description");
      ctx.init(o.device);
      ctx.init_comm(o.comm_rank, o.comm_size < 1 ? 1 : o.comm_size, o.comm_id);
      basis = make_basis(o.degree);
      degree_kernels(o.degree).init(basis);
      if (o.problem == MGB_PROBLEM_CUBE_SINE)
        {
          fn_solution = FN_CUBE_SOLUTION;
          fn_rhs      = FN_CUBE_RHS;
          fn_coef     = FN_ONE;
        }
      else if (o.problem == MGB_PROBLEM_SHELL_COEF)
        {
          fn_solution = FN_SHELL_SOLUTION;
          fn_rhs      = FN_SHELL_RHS;
          fn_coef     = FN_SHELL_COEFFICIENT;
        }
      else
        throw std::runtime_error("unknown problem id");
      if (o.geometry < MGB_GEOMETRY_BOX || o.geometry > MGB_GEOMETRY_AFFINE)
        throw std::runtime_error("unknown geometry id");
      geom.kind = o.geometry;
      for (int d = 0; d < 3; ++d)
        {
          geom.lo[d] = o.lower[d];
          geom.hi[d] = o.upper[d];
        }
      for (int i = 0; i < 12; ++i)
        geom.par[i] = o.geometry_params[i];
      if (geom.kind == MGB_GEOMETRY_SHELL_SECTOR &&
          !(geom.par[0] > 0 && geom.par[1] > geom.par[0] && geom.par[2] > 0 && geom.par[2] < 1.5))
        throw std::runtime_error("shell sector: need 0 < r0 < r1 and an opening half-angle in (0, 1.5)");

      n_levels = o.n_refinements + 1;
      transfer.resize(n_levels);
      mg_transfer_no_boundary.resize(n_levels);
      const DegreeKernels *dk = &degree_kernels(o.degree);
      for (int l = 0; l < n_levels; ++l)
        {
          int nc[3];
          for (int d = 0; d < 3; ++d)
            nc[d] = o.n_subdivisions[d] << l;
          levels.emplace_back(new Level());
          Level &lv = *levels.back();
          lv.index  = l;
          // small levels cost a ghost exchange per operator apply and save next to nothing: by default a
          // level is cut only from 3 M DoFs on (measured: profiles/r1_mgpu_levels.txt)
          const char *ml = getenv("MGB200_MIN_SLAB_LAYERS"), *md = getenv("MGB200_MIN_SLAB_DOFS");
          lv.slab = make_slab(o.degree, o.n_subdivisions, l, ctx.n_ranks, ctx.rank, ml ? std::max(1, atoi(ml)) : 2,
                              md ? atol(md) : 3000000L);
          build_slab_level(o.degree, nc, choose_tile_cells(o.degree, o.tile_cells), o.lower, o.upper, 0x3fu, lv.slab, lv.host);
          lv.upload(ctx);
          const size_t N = lv.d.n_dofs;
          matrix.emplace_back(new LaplaceOperator<float>());
          matrix_dp.emplace_back(new LaplaceOperator<double>());
          matrix[l]->kernel    = o.kernel;
          matrix_dp[l]->kernel = o.kernel;
          matrix[l]->initialize(ctx, lv, basis, fn_coef, geom);
          matrix_dp[l]->initialize(ctx, lv, basis, fn_coef, geom);
          auto newd = [&]() { auto v = new DVec<double>(); v->alloc(N, ctx.stream); return std::unique_ptr<DVec<double>>(v); };
          auto newf = [&]() { auto v = new DVec<float>(); v->alloc(N, ctx.stream); return std::unique_ptr<DVec<float>>(v); };
          solution.push_back(newd());
          rhs.push_back(newd());
          residual.push_back(newd());
          bc_values.push_back(newd());
          defect.push_back(newf());
          t.push_back(newf());
          solution_update.push_back(newf());
          if (l > 0)
            {
              transfer[l].ctx = mg_transfer_no_boundary[l].ctx = &ctx;
              transfer[l].fine = mg_transfer_no_boundary[l].fine = levels[l].get();
              transfer[l].coarse = mg_transfer_no_boundary[l].coarse = levels[l - 1].get();
              transfer[l].dk = mg_transfer_no_boundary[l].dk = dk;
              transfer[l].constrained                = true;
              mg_transfer_no_boundary[l].constrained = false;
              transfer[l].kernel = mg_transfer_no_boundary[l].kernel = o.kernel;
            }
        }
      {
        // peer-to-peer ghost exchange: the receive buffers hold the largest message, p planes of doubles
        size_t max_msg = 0;
        for (int l = 0; l < n_levels; ++l)
          if (levels[l]->slab.distributed)
            max_msg = std::max<size_t>(max_msg, (size_t)o.degree * plane_pitch(levels[l]->d) * sizeof(double));
        if (max_msg)
          ctx.init_p2p(max_msg);
      }
      // inhomogeneous boundary values and right-hand side (multigrid_solver.h:224-262)
      Nodes1D nd;
      for (int i = 0; i <= o.degree; ++i)
        nd.nodes[i] = basis.nodes[i];
      for (int l = 0; l < n_levels; ++l)
        {
          const LevelDesc L   = levels[l]->d;
          double         *bc  = bc_values[l]->p;
          const int       fns = fn_solution;
          const Geometry  gm  = geom;
          patch_points(ctx, L, [=] __device__(uint32_t dof, bool con, int gi, int gj, int gk) {
            if (!con)
              return;
            if (gm.kind == MGB_GEOMETRY_BOX)
              bc[dof] = eval_fn(fns,
                                support_point(L, nd, 0, gi),
                                support_point(L, nd, 1, gj),
                                support_point(L, nd, 2, gk));
            else
              {
                // mapped support point: the geometry map at the node's mesh coordinates (what the polynomial cell
                // geometry interpolates)
                double x[3];
                geometry_point(gm, unit_coordinate(L, nd, 0, gi), unit_coordinate(L, nd, 1, gj),
                               unit_coordinate(L, nd, 2, gk), x);
                bc[dof] = eval_fn(fns, x[0], x[1], x[2]);
              }
          });
          // rhs = int f phi - K u_bc  (laplace_operator.h:806-845): general kernel, plain read
          const size_t N   = L.n_dofs;
          double      *neg = residual[l]->p;
          foreach (ctx, N, [=] __device__(size_t i) { neg[i] = -bc[i]; });
          MGB_CUDA_CHECK(cudaMemsetAsync(rhs[l]->p, 0, N * sizeof(double), ctx.stream));
          matrix_dp[l]->apply(APPLY_ADD, rhs[l]->p, neg, nullptr, nullptr, 0, 0, false, /*read_plain=*/true);
          RhsArgs R;
          R.L     = L;
          R.geom  = geom;
          R.dst   = rhs[l]->p;
          R.fn_id = fn_rhs;
          dk->rhs_values(R, ctx.stream);
          ctx.check_launch();
          levels[l]->update_ghosts(rhs[l]->p); // the cell loop completed the owned planes only
          MGB_CUDA_CHECK(cudaMemsetAsync(residual[l]->p, 0, N * sizeof(double), ctx.stream));
        }
      // smoothers (multigrid_solver.h:268-289)
      for (int l = 0; l < n_levels; ++l)
        {
          smooth.emplace_back(new Chebyshev<float>());
          Chebyshev<float> &s = *smooth[l];
          if (l > 0)
            {
              s.smoothing_range     = 20.;
              s.degree              = o.degree_pre;
              s.eig_cg_n_iterations = 15;
            }
          else
            {
              s.smoothing_range     = 1e-3;
              s.degree              = Chebyshev<float>::invalid_degree;
              s.eig_cg_n_iterations = (unsigned int)matrix[0]->m_global();
            }
          matrix[l]->compute_diagonal();
          s.initialize(ctx, *matrix[l], basis);
        }
      choose_graph_level();
      const size_t N = levels.back()->d.n_dofs;
      cg_g.alloc(N, ctx.stream);
      cg_h.alloc(N, ctx.stream);
      cg_d.alloc(N, ctx.stream);
      ctx.sync();
    }

    ~MultigridSolver()
    {
      for (CycleGraph &g : cycle_graphs)
        if (g.exec)
          cudaGraphExecDestroy(g.exec);
      ctx.destroy();
    }

    void
    set_boundary_values(const int level, const bool zero)
    {
      double         *sol   = solution[level]->p;
      const double   *bc    = bc_values[level]->p;
      const uint32_t *clist = levels[level]->constrained.p;
      foreach (ctx, levels[level]->constrained.n, [=] __device__(size_t i) {
        const uint32_t d = clist[i];
        sol[d]           = zero ? 0. : bc[d];
      });
    }

    // ---- problem data supplied by the caller instead of the built-in functions (what evaluate_coefficient(const
    //      Function &) and compute_residual(dst, src, rhs_function) of laplace_operator.h:103-109 take): samples at the
    //      quadrature points, cells in lexicographic order (x fastest) of this rank's box, points x fastest inside a cell
    DVec<double> caller_values;          // staging for one level
    std::vector<std::unique_ptr<DVec<double>>> exact_q; // caller's exact solution at the quadrature points, per level
    bool         smoothers_stale = false;

    size_t
    n_quadrature_points(const int level) const
    {
      const LevelDesc &L = levels[level]->d;
      return (size_t)L.n_cells * L.n * L.n * L.n;
    }

    void
    quadrature_points(const int level, double *xyz_host)
    {
      const size_t nq = n_quadrature_points(level);
      DVec<double> buf;
      buf.alloc(3 * nq, ctx.stream);
      PointsArgs A;
      A.L    = levels[level]->d;
      A.geom = geom;
      A.xyz  = buf.p;
      degree_kernels(opt.degree).quadrature_points(A, ctx.stream);
      ctx.check_launch();
      MGB_CUDA_CHECK(cudaMemcpyAsync(xyz_host, buf.p, 3 * nq * sizeof(double), cudaMemcpyDeviceToHost, ctx.stream));
      ctx.sync();
    }

    void
    support_points(const int level, double *xyz_host)
    {
      const LevelDesc L = levels[level]->d;
      DVec<double>    buf;
      buf.alloc(3 * (size_t)L.n_dofs, ctx.stream);
      Nodes1D nd;
      for (int i = 0; i <= opt.degree; ++i)
        nd.nodes[i] = basis.nodes[i];
      double        *out = buf.p;
      const Geometry gm  = geom;
      patch_points(ctx, L, [=] __device__(uint32_t dof, bool, int gi, int gj, int gk) {
        double x[3] = {support_point(L, nd, 0, gi), support_point(L, nd, 1, gj), support_point(L, nd, 2, gk)};
        if (gm.kind != MGB_GEOMETRY_BOX)
          geometry_point(gm, unit_coordinate(L, nd, 0, gi), unit_coordinate(L, nd, 1, gj), unit_coordinate(L, nd, 2, gk), x);
        out[3 * (size_t)dof]     = x[0];
        out[3 * (size_t)dof + 1] = x[1];
        out[3 * (size_t)dof + 2] = x[2];
      });
      MGB_CUDA_CHECK(cudaMemcpyAsync(xyz_host, buf.p, 3 * (size_t)L.n_dofs * sizeof(double), cudaMemcpyDeviceToHost, ctx.stream));
      ctx.sync();
    }

    void
    stage_values(const double *host, const size_t count)
    {
      if (caller_values.n < count)
        caller_values.alloc(count, ctx.stream);
      MGB_CUDA_CHECK(cudaMemcpyAsync(caller_values.p, host, count * sizeof(double), cudaMemcpyHostToDevice, ctx.stream));
    }

    void
    set_coefficient_values(const int level, const double *c_host)
    {
      stage_values(c_host, n_quadrature_points(level));
      matrix[level]->set_coefficient_values(basis, geom, caller_values.p);
      matrix_dp[level]->set_coefficient_values(basis, geom, caller_values.p);
      ctx.sync();
      smoothers_stale = true;
      cycle_graphs_invalidate();
    }

    // diagonals and Chebyshev eigenvalue estimates after the coefficient changed (multigrid_solver.h:268-289)
    void
    reinit_smoothers()
    {
      for (int l = 0; l < n_levels; ++l)
        {
          matrix[l]->compute_diagonal();
          Chebyshev<float> &s = *smooth[l];
          if (l == 0)
            {
              s.degree              = Chebyshev<float>::invalid_degree;
              s.eig_cg_n_iterations = (unsigned int)matrix[0]->m_global();
            }
          s.initialize(ctx, *matrix[l], basis);
        }
      cycle_graphs_invalidate();
      choose_graph_level();
      smoothers_stale = false;
      ctx.sync();
    }

    // rhs = int f phi - K u_bc with f sampled at the quadrature points and the boundary values at the support points
    // (u_bc_host: one value per DoF of the level, only the Dirichlet entries are used; null = homogeneous)
    void
    set_rhs_values(const int level, const double *f_host, const double *u_bc_host)
    {
      const LevelDesc L = levels[level]->d;
      const size_t    N = L.n_dofs;
      double         *bc = bc_values[level]->p;
      MGB_CUDA_CHECK(cudaMemsetAsync(bc, 0, N * sizeof(double), ctx.stream));
      if (u_bc_host)
        {
          stage_values(u_bc_host, N);
          const double   *v     = caller_values.p;
          const uint32_t *clist = levels[level]->constrained.p;
          foreach (ctx, levels[level]->constrained.n, [=] __device__(size_t i) { bc[clist[i]] = v[clist[i]]; });
        }
      double *neg = residual[level]->p;
      foreach (ctx, N, [=] __device__(size_t i) { neg[i] = -bc[i]; });
      MGB_CUDA_CHECK(cudaMemsetAsync(rhs[level]->p, 0, N * sizeof(double), ctx.stream));
      matrix_dp[level]->apply(APPLY_ADD, rhs[level]->p, neg, nullptr, nullptr, 0, 0, 0, /*read_plain=*/true);
      ctx.sync(); // the staging buffer is reused below
      stage_values(f_host, n_quadrature_points(level));
      RhsArgs R;
      R.L      = L;
      R.geom   = geom;
      R.dst    = rhs[level]->p;
      R.fn_id  = FN_ZERO;
      R.values = caller_values.p;
      degree_kernels(opt.degree).rhs_values(R, ctx.stream);
      ctx.check_launch();
      levels[level]->update_ghosts(rhs[level]->p);
      MGB_CUDA_CHECK(cudaMemsetAsync(residual[level]->p, 0, N * sizeof(double), ctx.stream));
      ctx.sync();
    }

    // LaplaceOperator::compute_residual(dst, src, rhs_function) (laplace_operator.h:103-106, 806-845) on the double operator
    // of `level`: dst = (f, phi) - K src on the rows without Dirichlet condition, src read with its Dirichlet entries; f by
    // samples at the quadrature points (host) or, if null, the right-hand side function of the solver's problem
    void
    compute_residual(const int level, double *dst, const double *src, const double *f_host)
    {
      const LevelDesc L = levels[level]->d;
      const size_t    N = L.n_dofs;
      if (caller_scratch.n < N)
        caller_scratch.alloc(N, ctx.stream);
      double *neg = caller_scratch.p;
      foreach (ctx, N, [=] __device__(size_t i) { neg[i] = -src[i]; });
      MGB_CUDA_CHECK(cudaMemsetAsync(dst, 0, N * sizeof(double), ctx.stream));
      matrix_dp[level]->apply(APPLY_ADD, dst, neg, nullptr, nullptr, 0, 0, 0, /*read_plain=*/true);
      RhsArgs R;
      R.L     = L;
      R.geom  = geom;
      R.dst   = dst;
      R.fn_id = fn_rhs;
      if (f_host)
        {
          stage_values(f_host, n_quadrature_points(level));
          R.fn_id  = FN_ZERO;
          R.values = caller_values.p;
        }
      degree_kernels(opt.degree).rhs_values(R, ctx.stream);
      ctx.check_launch();
      levels[level]->update_ghosts(dst);
      ctx.sync();
    }
    DVec<double> caller_scratch;

    // the exact solution for compute_l2_error, sampled at the quadrature points of `level`
    void
    set_exact_solution_values(const int level, const double *u_host)
    {
      if ((int)exact_q.size() < n_levels)
        exact_q.resize(n_levels);
      if (!exact_q[level])
        exact_q[level].reset(new DVec<double>());
      const size_t nq = n_quadrature_points(level);
      exact_q[level]->alloc(nq, ctx.stream);
      MGB_CUDA_CHECK(cudaMemcpyAsync(exact_q[level]->p, u_host, nq * sizeof(double), cudaMemcpyHostToDevice, ctx.stream));
      ctx.sync();
    }

    // multigrid_solver.h:298-343
    double
    compute_l2_error(const int level)
    {
      if (smoothers_stale)
        throw std::runtime_error("the coefficient changed: call mgb_solver_reinit_smoothers before solving");
      set_boundary_values(level, false);
      levels[level]->update_ghosts(solution[level]->p);
      MGB_CUDA_CHECK(cudaMemsetAsync(ctx.scal_dev, 0, 2 * sizeof(double), ctx.stream));
      ErrorArgs E;
      E.L     = levels[level]->d;
      E.geom  = geom;
      E.u     = solution[level]->p;
      E.fn_id = fn_solution;
      E.accum = ctx.scal_dev;
      if ((int)exact_q.size() > level && exact_q[level])
        E.values = exact_q[level]->p;
      degree_kernels(opt.degree).l2_error(E, ctx.stream);
      ctx.check_launch();
      if (levels[level]->slab.distributed)
        ctx.all_reduce_sum(ctx.scal_dev, 2);
      MGB_CUDA_CHECK(cudaMemcpyAsync(ctx.scal_host, ctx.scal_dev, 2 * sizeof(double), cudaMemcpyDeviceToHost, ctx.stream));
      ctx.sync();
      return std::sqrt(ctx.scal_host[0] / ctx.scal_host[1]);
    }

    // ---- CUDA graphs for the launch-bound coarse part of the V-cycle.  Below `graph_level` every kernel runs
    //      for a few microseconds, so the cycle costs what the host needs to launch it; one captured graph
    //      replaces ~25 launches per level.  The Chebyshev steps swap vectors, so the device addresses a cycle
    //      works on alternate: graphs are cached by the address state they were captured for, and launching
    //      one puts the host-side pointers into the state the captured code left them in.
    struct CycleGraph
    {
      std::vector<void *> pre, post;
      cudaGraphExec_t     exec = nullptr;
    };
    std::vector<CycleGraph> cycle_graphs;
    int                     graph_level = -1; // v_cycle(graph_level, 1) is one graph launch; -1: off
    int64_t                 graph_launches = 0;
    unsigned int            v_cycle_calls = 0;

    std::vector<void *>
    pointer_state(const int level) const
    {
      std::vector<void *> st;
      for (int l = 0; l <= level; ++l)
        {
          st.push_back(solution_update[l]->p);
          st.push_back(smooth[l]->solution_old.p);
        }
      return st;
    }

    void
    cycle_graphs_invalidate()
    {
      ctx.sync();
      for (CycleGraph &g : cycle_graphs)
        if (g.exec)
          cudaGraphExecDestroy(g.exec);
      cycle_graphs.clear();
    }

    void
    choose_graph_level()
    {
      const char *env   = getenv("MGB200_GRAPH_DOFS");
      const long  limit = env ? atol(env) : 3000000L;
      graph_level       = -1;
      for (int l = 0; l < n_levels - 1; ++l)
        if (!levels[l]->slab.distributed && matrix[l]->use_cartesian() && (long)levels[l]->d.n_dofs <= limit &&
            transfer_uses_tiles(l))
          graph_level = l;
        else
          break;
      if (graph_level < 1)
        graph_level = -1;
    }
    bool
    transfer_uses_tiles(const int l) const
    {
      return l == 0 || (levels[l]->d.CW == choose_tile_cells(levels[l]->d.p, 0) && transfer[l].kernel == MGB_KERNEL_AUTO);
    }

    void
    v_cycle_graph(const int level)
    {
      const std::vector<void *> pre = pointer_state(level);
      for (CycleGraph &g : cycle_graphs)
        if (g.pre == pre)
          {
            MGB_CUDA_CHECK(cudaGraphLaunch(g.exec, ctx.stream));
            ++graph_launches;
            ctx.launches += graph_kernels; // the kernels inside the graph
            for (int l = 0; l <= level; ++l)
              {
                solution_update[l]->p       = static_cast<float *>(g.post[2 * l]);
                smooth[l]->solution_old.p = static_cast<float *>(g.post[2 * l + 1]);
              }
            return;
          }
      if (cycle_graphs.size() >= 8)
        { // unexpected number of address states: fall back to plain launches
          v_cycle_plain(level, 1);
          return;
        }
      CycleGraph   g;
      cudaGraph_t  graph   = nullptr;
      const int64_t before = ctx.launches;
      g.pre                = pre;
      MGB_CUDA_CHECK(cudaStreamBeginCapture(ctx.stream, cudaStreamCaptureModeRelaxed));
      try
        {
          v_cycle_plain(level, 1);
        }
      catch (...)
        {
          cudaStreamEndCapture(ctx.stream, &graph);
          if (graph)
            cudaGraphDestroy(graph);
          throw;
        }
      MGB_CUDA_CHECK(cudaStreamEndCapture(ctx.stream, &graph));
      MGB_CUDA_CHECK(cudaGraphInstantiate(&g.exec, graph, 0));
      MGB_CUDA_CHECK(cudaGraphDestroy(graph));
      g.post          = pointer_state(level);
      graph_kernels   = ctx.launches - before; // kernels inside one graph launch (for the launch count)
      MGB_CUDA_CHECK(cudaGraphLaunch(g.exec, ctx.stream));
      ++graph_launches;
      cycle_graphs.push_back(g);
    }
    int64_t graph_kernels = 0;

    void
    v_cycle(const int level, const unsigned int my_n_cycles)
    {
      // the first cycles run plainly (kernel attributes are set on first launch); profiling needs the
      // per-range events, which a graph does not record
      if (level == graph_level && my_n_cycles == 1 && !ctx.profile && v_cycle_calls >= 2)
        {
          v_cycle_graph(level);
          return;
        }
      if (level == graph_level)
        ++v_cycle_calls;
      v_cycle_plain(level, my_n_cycles);
    }

    // multigrid_solver.h:641-681 (all in float)
    void
    v_cycle_plain(const int level, const unsigned int my_n_cycles)
    {
      if (level == 0)
        {
          ScopedRange range(ctx, 0, RK_SMOOTHER);
          smooth[0]->vmult(*solution_update[0], defect[0]->p);
          return;
        }
      for (unsigned int c = 0; c < my_n_cycles; ++c)
        {
          {
            ScopedRange range(ctx, level, RK_SMOOTHER);
            if (c == 0)
              smooth[level]->vmult(*solution_update[level], defect[level]->p);
            else
              smooth[level]->step(*solution_update[level], defect[level]->p);
          }
          {
            ScopedRange range(ctx, level, RK_MG_MV);
            matrix[level]->vmult_residual(defect[level]->p, solution_update[level]->p, t[level]->p);
          }
          {
            ScopedRange range(ctx, level, RK_RESTRICT);
            MGB_CUDA_CHECK(cudaMemsetAsync(defect[level - 1]->p, 0, defect[level - 1]->n * sizeof(float), ctx.stream));
            transfer[level].restrict_and_add(defect[level - 1]->p, t[level]->p);
          }
          if (level - 1 == graph_level)
            v_cycle(level - 1, 1);
          else
            v_cycle_plain(level - 1, 1);
          {
            ScopedRange range(ctx, level, RK_PROLONGATE);
            transfer[level].prolongate_and_add(solution_update[level]->p, solution_update[level - 1]->p);
          }
          {
            ScopedRange range(ctx, level, RK_SMOOTHER);
            smooth[level]->step(*solution_update[level], defect[level]->p);
          }
        }
    }

    // multigrid_solver.h:387-476
    double
    solve(const bool do_analyze, double *trace)
    {
      double reduction_rate = 1.;
      convert(ctx, defect[0]->p, rhs[0]->p, rhs[0]->n);
      smooth[0]->vmult(*t[0], defect[0]->p);
      smooth[0]->step(*t[0], defect[0]->p);
      convert(ctx, solution[0]->p, t[0]->p, t[0]->n);
      for (int level = 1; level < n_levels; ++level)
        {
          const size_t N = levels[level]->d.n_dofs;
          set_boundary_values(level - 1, false);
          mg_transfer_no_boundary[level].prolongate(solution[level]->p, solution[level - 1]->p);
          double init_residual = 1.;
          if (do_analyze && trace)
            trace[level * 4 + 0] = compute_l2_error(level);
          set_boundary_values(level, true);
          matrix_dp[level]->vmult_residual(rhs[level]->p, solution[level]->p, residual[level]->p);
          convert(ctx, defect[level]->p, residual[level]->p, N);
          if (do_analyze)
            {
              init_residual = std::sqrt(dot(ctx, residual[level]->p, residual[level]->p, levels[level]->owned()));
              if (trace)
                trace[level * 4 + 1] = init_residual;
            }
          v_cycle(level, opt.n_cycles);
          {
            double      *sol = solution[level]->p;
            const float *upd = solution_update[level]->p;
            foreach (ctx, N, [=] __device__(size_t i) { sol[i] += upd[i]; }); // internal::add_vector
          }
          if (do_analyze)
            {
              set_boundary_values(level, true);
              matrix_dp[level]->vmult(residual[level]->p, solution[level]->p);
              double       *res = residual[level]->p;
              const double *b   = rhs[level]->p;
              foreach (ctx, N, [=] __device__(size_t i) { res[i] = b[i] - res[i]; });
              const double res_norm = std::sqrt(dot(ctx, res, res, levels[level]->owned()));
              reduction_rate        = std::pow(res_norm / init_residual, 1. / opt.n_cycles);
              if (trace)
                {
                  trace[level * 4 + 2] = res_norm;
                  trace[level * 4 + 3] = compute_l2_error(level);
                }
            }
        }
      ctx.sync();
      return reduction_rate;
    }

    // multigrid_solver.h:498-510
    void
    vmult(double *dst, const double *src)
    {
      const int    L = n_levels - 1;
      const size_t N = levels[L]->d.n_dofs;
      {
        ScopedRange range(ctx, L, RK_MG_VEC);
        convert(ctx, defect[L]->p, src, N);
      }
      v_cycle(L, 1);
      {
        ScopedRange range(ctx, L, RK_MG_VEC);
        convert(ctx, dst, solution_update[L]->p, N);
      }
    }

    // multigrid_solver.h:516-619: the preconditioner entry of a merged CG.  defect = float(residual + factor update);
    // V-cycle; then in ONE pass over the vectors  residual += factor update,  update = double(V-cycle result) and the
    // sums  <z, residual_new>, <z, factor update_old>  (z = the new update) over the owned entries of all ranks.
    // Dirichlet rows are identity rows of the preconditioner (:570-577, :598-603): z = residual_new there.
    // factor == 0 skips the residual update; both sums are then <z, residual>.
    std::array<double, 2>
    vmult_with_residual_update(double *residual, double *update, const double factor)
    {
      const int    L   = n_levels - 1;
      Level       &lvl = *levels[L];
      const size_t N   = lvl.d.n_dofs;
      const Span   own = lvl.owned();
      float       *def = defect[L]->p;
      {
        ScopedRange range(ctx, L, RK_MG_VEC);
        if (factor != 0.)
          foreach (ctx, N, [=] __device__(size_t i) { def[i] = (float)(residual[i] + factor * update[i]); });
        else
          convert(ctx, def, residual, N);
      }
      v_cycle(L, 1);
      float         *z     = solution_update[L]->p;
      const uint32_t *clist = lvl.constrained.p;
      const size_t   nc    = lvl.constrained.n;
      const size_t   o0 = own.off, o1 = own.off + own.n;
      double         sums[2];
      ScopedRange    range(ctx, L, RK_MG_VEC);
      // identity rows: park (residual_new, factor update_old) of the constrained DoFs and zero the V-cycle result
      // there, so that the main pass contributes nothing for them
      if (con_scratch.n < 2 * nc)
        con_scratch.alloc(2 * nc, ctx.stream);
      double *park = con_scratch.p;
      foreach (ctx, nc, [=] __device__(size_t k) {
        const uint32_t i = clist[k];
        const double   fu = factor != 0. ? factor * update[i] : 0.;
        park[2 * k]       = residual[i] + fu;
        park[2 * k + 1]   = fu;
        z[i]              = 0.f;
      });
      reduce<2>(
        ctx,
        own,
        [=] __device__(size_t i, double *acc) {
          const double mg = (double)z[i], fu = factor != 0. ? factor * update[i] : 0.;
          const double res = residual[i] + fu;
          residual[i]      = res;
          update[i]        = mg;
          acc[0] += mg * res;
          acc[1] += mg * fu;
        },
        sums);
      if (nc)
        {
          double fix[2];
          reduce<2>(
            ctx,
            Span(0, nc, own.all_ranks),
            [=] __device__(size_t k, double *acc) {
              const uint32_t i = clist[k];
              const double   res = park[2 * k];
              update[i]          = res;
              residual[i]        = res;
              if (i >= o0 && i < o1)
                {
                  acc[0] += res * res;
                  acc[1] += res * park[2 * k + 1];
                }
            },
            fix);
          sums[0] += fix[0];
          sums[1] += fix[1];
        }
      else if (own.all_ranks)
        {
          // every rank must take part in the all-reduce of the fix-up even without constrained DoFs of its own
          double fix[2];
          reduce<2>(ctx, Span(0, 0, true), [=] __device__(size_t, double *) {}, fix);
          sums[0] += fix[0];
          sums[1] += fix[1];
        }
      if (factor == 0.)
        sums[1] = sums[0];
      // (ghost planes of a z-slab are left alone: every consumer refreshes them, section 7 of DESIGN.md)
      return {{sums[0], sums[1]}};
    }

    // ---- minimal_surface/program.cc: Newton's method with line search for -div(grad u / sqrt(1 + |grad u|^2)) = 0
    //      with Dirichlet data; the linear systems by CG with the V-cycle on the linearised operator
    //      (LaplaceProblem::solve, :412-572).  Here in 3-D on the solver's mesh; one GPU.
    std::vector<std::unique_ptr<DVec<float>>> coefficient_solutions;
    std::unique_ptr<LaplaceOperator<double>>  ms_residual_op;
    DVec<double>                              ms_search, ms_tentative, ms_rhs, ms_current;

    void
    ms_require_one_gpu() const
    {
      if (ctx.n_ranks > 1)
        throw std::runtime_error("the minimal surface solver runs on one GPU");
    }

    // interpolate_boundary_values (:396-408): the Dirichlet data in the boundary entries of `solution`, zero inside.
    // u_host: one value per DoF of the finest level in library numbering (only the Dirichlet entries are used), or
    // null for the solver's analytic solution
    void
    ms_set_boundary_values(const double *u_host)
    {
      ms_require_one_gpu();
      const int    L = n_levels - 1;
      const size_t N = levels[L]->d.n_dofs;
      if (u_host)
        {
          stage_values(u_host, N);
          double         *bc    = bc_values[L]->p;
          const double   *v     = caller_values.p;
          const uint32_t *clist = levels[L]->constrained.p;
          foreach (ctx, levels[L]->constrained.n, [=] __device__(size_t i) { bc[clist[i]] = v[clist[i]]; });
        }
      MGB_CUDA_CHECK(cudaMemsetAsync(solution[L]->p, 0, N * sizeof(double), ctx.stream));
      set_boundary_values(L, false);
      if (!ms_residual_op)
        {
          ms_residual_op.reset(new LaplaceOperator<double>());
          ms_residual_op->initialize(ctx, *levels[L], basis, FN_ONE, geom);
          ms_residual_op->kernel = MGB_KERNEL_GENERAL;
          coefficient_solutions.clear();
          for (int l = 0; l < n_levels; ++l)
            {
              coefficient_solutions.emplace_back(new DVec<float>());
              coefficient_solutions[l]->alloc(levels[l]->d.n_dofs, ctx.stream);
            }
          for (DVec<double> *v : {&ms_search, &ms_tentative, &ms_rhs, &ms_current})
            v->alloc(N, ctx.stream);
        }
      ctx.sync();
    }

    // MinimalSurfaceOperator::compute_residual (:169-198): dst_i = -(grad phi_i, grad u / sqrt(1 + |grad u|^2)) on the
    // rows without Dirichlet condition (first_time: the Laplace residual); returns the l2 norm
    double
    ms_residual(double *dst, const double *u, const bool first_time)
    {
      const int    L = n_levels - 1;
      const size_t N = levels[L]->d.n_dofs;
      if (!ms_residual_op)
        throw std::runtime_error("call mgb_solver_minimal_surface_set_boundary_values first");
      ms_residual_op->evaluate_solution_coefficient(geom, first_time ? 0 : 2, u);
      double *neg = residual[L]->p;
      foreach (ctx, N, [=] __device__(size_t i) { neg[i] = -u[i]; });
      MGB_CUDA_CHECK(cudaMemsetAsync(dst, 0, N * sizeof(double), ctx.stream));
      ms_residual_op->apply(APPLY_ADD, dst, neg, nullptr, nullptr, 0, 0, 0, /*read_plain=*/true);
      return std::sqrt(dot(ctx, dst, dst, levels[L]->owned()));
    }

    // LaplaceProblem::solve(first_time): info = {initial residual, final residual, linear iterations, line search
    // steps, step length}
    void
    ms_newton_step(const bool first_time, double info[5])
    {
      ms_require_one_gpu();
      if (!ms_residual_op)
        throw std::runtime_error("call mgb_solver_minimal_surface_set_boundary_values first");
      const int    L   = n_levels - 1;
      const size_t N   = levels[L]->d.n_dofs;
      double      *sol = solution[L]->p;
      // the solution on every level: conversion to float, then the FE_Q restriction matrices level by level (:417-456)
      convert(ctx, coefficient_solutions[L]->p, sol, N);
      for (int l = L; l > 0; --l)
        {
          TransferArgs<float> A;
          A.fine        = levels[l]->d;
          A.coarse      = levels[l - 1]->d;
          A.src         = coefficient_solutions[l]->p;
          A.dst         = coefficient_solutions[l - 1]->p;
          A.mode        = 0;
          A.constrained = 0;
          A.cz_begin = A.cz_end = A.f_shift = A.close_top = 0;
          degree_kernels(opt.degree).interpolate_to_coarse(A, ctx.stream);
          ctx.check_launch();
        }
      const int mode = first_time ? 0 : 1;
      matrix_dp[L]->evaluate_solution_coefficient(geom, mode, sol);
      for (int l = 0; l < n_levels; ++l)
        matrix[l]->evaluate_solution_coefficient(geom, mode, coefficient_solutions[l]->p);
      reinit_smoothers(); // diagonals, eigenvalue estimates (:458-486)
      const double initial_residual_norm = ms_residual(ms_rhs.p, sol, first_time);
      // SolverCG with ReductionControl(m(), 1e-13, 1e-4) from a zero start (:505-535); the solver works in solution / rhs
      MGB_CUDA_CHECK(cudaMemcpyAsync(ms_current.p, sol, N * sizeof(double), cudaMemcpyDeviceToDevice, ctx.stream));
      MGB_CUDA_CHECK(cudaMemcpyAsync(rhs[L]->p, ms_rhs.p, N * sizeof(double), cudaMemcpyDeviceToDevice, ctx.stream));
      const auto cg = solve_cg((unsigned int)matrix_dp[L]->m_global(), 1e-13, 1e-4);
      MGB_CUDA_CHECK(cudaMemcpyAsync(ms_search.p, sol, N * sizeof(double), cudaMemcpyDeviceToDevice, ctx.stream));
      MGB_CUDA_CHECK(cudaMemcpyAsync(sol, ms_current.p, N * sizeof(double), cudaMemcpyDeviceToDevice, ctx.stream));
      {
        double         *sd    = ms_search.p;
        const uint32_t *clist = levels[L]->constrained.p;
        foreach (ctx, levels[L]->constrained.n, [=] __device__(size_t i) { sd[clist[i]] = 0.; }); // constraints.distribute
      }
      // line search (:543-557)
      double       final_residual_norm = initial_residual_norm, alpha = 1.;
      unsigned int n_steps             = 0;
      while (n_steps < 100)
        {
          double       *tent = ms_tentative.p;
          const double *sd   = ms_search.p;
          const double  a    = alpha;
          foreach (ctx, N, [=] __device__(size_t i) { tent[i] = sol[i] + a * sd[i]; });
          final_residual_norm = ms_residual(ms_rhs.p, tent, false);
          if (final_residual_norm < initial_residual_norm)
            break;
          alpha /= 2.;
          ++n_steps;
        }
      MGB_CUDA_CHECK(cudaMemcpyAsync(sol, ms_tentative.p, N * sizeof(double), cudaMemcpyDeviceToDevice, ctx.stream));
      ctx.sync();
      info[0] = initial_residual_norm;
      info[1] = final_residual_norm;
      info[2] = cg.first;
      info[3] = n_steps;
      info[4] = alpha;
    }

    // multigrid_solver.h:483-493 + deal.II SolverCG with ReductionControl(1000, 1e-16, 1e-9)
    std::pair<unsigned int, double>
    solve_cg(const unsigned int max_it = 1000, const double abs_tol = 1e-16, const double rel_tol = 1e-9)
    {
      if (smoothers_stale)
        throw std::runtime_error("the coefficient changed: call mgb_solver_reinit_smoothers before solving");
      // deal.II SolverCG (x0 = 0): g = -b; h = M^-1 g; d = -h; loop { h = A d; alpha = g.h_old / d.h;
      // x += alpha d; g += alpha h; check |g|; h = M^-1 g; beta = g.h / g.h_old; d = beta d - h }.
      // Vector work is fused as in the reference's merged loops (laplace_operator.h:640-719,
      // multigrid_solver.h:516-619): the residual update also produces |g|^2 and the float defect of the
      // V-cycle, and the preconditioned residual h = double(solution_update) is never materialised.
      const int     L = n_levels - 1;
      const size_t  N = levels[L]->d.n_dofs;
      const Span    own = levels[L]->owned();
      double       *x = solution[L]->p, *g = cg_g.p, *h = cg_h.p, *d = cg_d.p;
      const double *b = rhs[L]->p;
      float        *def = defect[L]->p;
      cg_history.clear();
      MGB_CUDA_CHECK(cudaMemsetAsync(x, 0, N * sizeof(double), ctx.stream));
      double r2;
      {
        ScopedRange range(ctx, L, RK_CG_VEC);
        reduce<1>(
          ctx,
          own,
          [=] __device__(size_t i, double *acc) {
            const double gi = -b[i];
            g[i]            = gi;
            def[i]          = (float)gi;
            acc[0] += gi * gi;
          },
          &r2);
      }
      double       res  = std::sqrt(r2);
      const double res0 = res;
      cg_history.push_back(res);
      unsigned int it = 0;
      if (res > abs_tol)
        {
          double gh = 0;
          // x += alpha d is postponed into the next direction update (which reads d anyway) or the final pass:
          // 8 bytes per DoF and iteration less than updating x together with the residual
          double alpha_pending = 0;
          for (;;)
            {
              v_cycle(L, 1); // solution_update[L] = M^-1 defect[L]
              const float *upd = solution_update[L]->p;
              {
                ScopedRange range(ctx, L, RK_CG_VEC);
                double      gh_new;
                reduce<1>(
                  ctx, own, [=] __device__(size_t i, double *acc) { acc[0] += g[i] * (double)upd[i]; }, &gh_new);
                if (it == 0)
                  foreach (ctx, N, [=] __device__(size_t i) { d[i] = -(double)upd[i]; });
                else
                  {
                    const double beta = gh_new / gh, ap = alpha_pending;
                    foreach (ctx, N, [=] __device__(size_t i) {
                      const double di = d[i];
                      x[i] += ap * di;
                      d[i] = beta * di - (double)upd[i];
                    });
                  }
                gh = gh_new;
              }
              ++it;
              double dh;
              matrix_dp[L]->vmult_dot(h, d, &dh);
              {
                ScopedRange  range(ctx, L, RK_CG_VEC);
                const double alpha = gh / dh;
                alpha_pending      = alpha;
                reduce<1>(
                  ctx,
                  own,
                  [=] __device__(size_t i, double *acc) {
                    const double gi = g[i] + alpha * h[i];
                    g[i]            = gi;
                    def[i]          = (float)gi;
                    acc[0] += gi * gi;
                  },
                  &r2);
              }
              res = std::sqrt(r2);
              cg_history.push_back(res);
              if (res <= abs_tol || res <= rel_tol * res0 || it >= max_it)
                break;
            }
          {
            ScopedRange  range(ctx, L, RK_CG_VEC);
            const double ap = alpha_pending;
            foreach (ctx, N, [=] __device__(size_t i) { x[i] += ap * d[i]; });
          }
        }
      ctx.sync();
      return std::make_pair(it, it > 0 ? std::pow(res / res0, 1. / it) : 0.);
    }
  };
  // ---- the subset of LinearAlgebra::distributed::Vector the drivers use (poisson_cube/program.cc, multigrid_solver.h)
  template <typename T>
  void
  vec_ops(MultigridSolver &m, const int level, const int op, void *yv, const void *xv, const double a, const double b,
          double *result)
  {
    Level       &lv = *m.levels[level];
    const size_t N  = lv.d.n_dofs;
    T           *y  = static_cast<T *>(yv);
    const T     *x  = static_cast<const T *>(xv);
    const T      ta = (T)a, tb = (T)b;
    switch (op)
      {
        case 0: *result = dot(m.ctx, y, x, lv.owned()); break;                                        // y . x
        case 1: foreach (m.ctx, N, [=] __device__(size_t i) { y[i] = ta * y[i] + tb * x[i]; }); break; // sadd
        case 2: foreach (m.ctx, N, [=] __device__(size_t i) { y[i] = ta; }); break;                    // = a
        default: throw std::runtime_error("unknown vector operation");
      }
  }

} // namespace mgb

// ====================================================================== C ABI
using namespace mgb;

static thread_local std::string g_last_error;

// used by the other translation units of the library (dg.cu)
void
mgb_set_last_error(const char *msg)
{
  g_last_error = msg;
}

// level description, stream and V-cycle vectors of a solver for dg.cu (same-degree DG <-> FE_Q transfer and the DG
// solver built on top of the FE_Q hierarchy); not part of the public C ABI
struct mgb_solver_t;
const mgb::LevelDesc *mgb_internal_level_desc(mgb_solver_t *s, int level);
cudaStream_t          mgb_internal_stream(mgb_solver_t *s);

struct mgb_solver_t
{
  MultigridSolver *impl;
};

// one-process-per-GPU communicator for the other translation units (the DG operator's cell-layer exchange): the
// same dlopen'ed NCCL as the solver; not part of the public C ABI
struct mgb_comm_t
{
  ncclComm_t comm = nullptr;
  int        rank = 0, size = 1;
  bool       owned = true;
};

mgb_comm_t *
mgb_internal_comm_create(const int rank, const int size, const unsigned char id[128], const int device)
{
  if (size < 1 || rank < 0 || rank >= size)
    throw std::runtime_error("invalid communicator rank / size");
  mgb_comm_t *c = new mgb_comm_t;
  c->rank       = rank;
  c->size       = size;
  if (size > 1)
    {
      Nccl &n = Nccl::get();
      static_assert(sizeof(ncclUniqueId) == 128, "ncclUniqueId is 128 bytes");
      ncclUniqueId uid;
      std::memcpy(&uid, id, sizeof(uid));
      MGB_CUDA_CHECK(cudaSetDevice(device));
      n.check(n.CommInitRank(&c->comm, size, uid, rank), "ncclCommInitRank");
    }
  return c;
}

void
mgb_internal_comm_destroy(mgb_comm_t *c)
{
  if (c && c->comm && c->owned)
    Nccl::get().CommDestroy(c->comm);
  delete c;
}

// update_ghost_values of a z-slab vector: `count` values go up to rank + 1 (its recv_dn) and down to rank - 1
// (its recv_up) in one NCCL group
void
mgb_internal_comm_exchange(mgb_comm_t *c, cudaStream_t stream, const void *send_up, void *recv_dn, const void *send_dn,
                           void *recv_up, const size_t count, const int is_double)
{
  if (c->size == 1)
    return;
  Nccl                &n  = Nccl::get();
  const ncclDataType_t dt = is_double ? ncclDouble : ncclFloat;
  n.check(n.GroupStart(), "ncclGroupStart");
  if (c->rank + 1 < c->size)
    {
      n.check(n.Send(send_up, count, dt, c->rank + 1, c->comm, stream), "ncclSend");
      n.check(n.Recv(recv_up, count, dt, c->rank + 1, c->comm, stream), "ncclRecv");
    }
  if (c->rank > 0)
    {
      n.check(n.Send(send_dn, count, dt, c->rank - 1, c->comm, stream), "ncclSend");
      n.check(n.Recv(recv_dn, count, dt, c->rank - 1, c->comm, stream), "ncclRecv");
    }
  n.check(n.GroupEnd(), "ncclGroupEnd");
}

void
mgb_internal_comm_allreduce(mgb_comm_t *c, cudaStream_t stream, double *dev, const int count)
{
  if (c->size == 1)
    return;
  Nccl &n = Nccl::get();
  n.check(n.AllReduce(dev, dev, count, ncclDouble, ncclSum, c->comm, stream), "ncclAllReduce");
}

#define MGB_TRY try {
#define MGB_CATCH                                  \
  }                                                \
  catch (const std::exception &e)                  \
  {                                                \
    g_last_error = e.what();                       \
    return 1;                                      \
  }                                                \
  catch (...)                                      \
  {                                                \
    g_last_error = "unknown error";                \
    return 2;                                      \
  }                                                \
  return 0;

template <typename LevelLike>
static void
fill_partition_info(const LevelLike &lv, int64_t info[12])
{
  const LevelDesc &d = lv.d;
  const ExchangePlan e = make_exchange_plan(d);
  info[0]  = lv.slab.distributed;
  info[1]  = (int64_t)d.nd[0] * d.nd[1] * d.nd2_all;
  info[2]  = d.n_dofs;
  info[3]  = e.owned[0];
  info[4]  = e.owned[1];
  info[5]  = d.gk0;
  info[6]  = d.nd[2];
  info[7]  = d.nd2_all;
  info[8]  = d.zg;
  info[9]  = d.ztop;
  info[10] = lv.slab.c0;
  info[11] = lv.slab.c1;
}

static MultigridSolver &
S(mgb_solver s)
{
  if (!s || !s->impl)
    throw std::runtime_error("null solver handle");
  return *s->impl;
}

static int
check_level(MultigridSolver &s, int level)
{
  if (level < 0)
    level += s.n_levels;
  if (level < 0 || level >= s.n_levels)
    throw std::runtime_error("level out of range");
  return level;
}

cudaStream_t
mgb_internal_stream(mgb_solver_t *s)
{
  return S(s).ctx.stream;
}

// the local box of a level; on several GPUs the z-slab of this rank (the DG operator checks that its own slab is
// the same one)
const mgb::LevelDesc *
mgb_internal_level_desc(mgb_solver_t *s, int level)
{
  MultigridSolver &m = S(s);
  Level           &l = *m.levels[check_level(m, level)];
  if (m.ctx.n_ranks > 1 && !l.slab.distributed)
    throw std::runtime_error("the DG <-> FE_Q transfer needs a distributed finest FE_Q level (too few cell layers per GPU)");
  m.ctx.sync();
  return &l.d;
}

// the solver's communicator for the DG operator of the DG multigrid solver (one NCCL communicator per solver);
// the returned object does not own it
mgb_comm_t *
mgb_internal_comm_borrow(mgb_solver_t *s)
{
  MultigridSolver &m = S(s);
  mgb_comm_t      *c = new mgb_comm_t;
  c->comm            = m.ctx.comm;
  c->rank            = m.ctx.rank;
  c->size            = m.ctx.n_ranks;
  c->owned           = false;
  return c;
}

extern "C"
{
  const char *mgb_last_error(void) { return g_last_error.c_str(); }
  int mgb_version(void) { return 100; }

  void
  mgb_default_options(mgb_options *o)
  {
    std::memset(o, 0, sizeof(*o));
    o->degree    = 4;
    o->comm_size = 1;
    for (int d = 0; d < 3; ++d)
      {
        o->n_subdivisions[d] = 1;
        o->lower[d]          = -0.9;
        o->upper[d]          = 1.0;
      }
    o->n_refinements = 3;
    o->degree_pre = o->degree_post = 3;
    o->n_cycles                    = 2;
    o->problem                     = MGB_PROBLEM_CUBE_SINE;
    o->geometry                    = MGB_GEOMETRY_BOX;
    for (int i = 0; i < 12; ++i)
      o->geometry_params[i] = 0.;
    o->geometry_params[0] = 0.5;                    // hyper_shell(center, 0.5, 1.0, 6): poisson_shell/program.cc:426
    o->geometry_params[1] = 1.0;
    o->geometry_params[2] = 0.78539816339744830962; // a whole cube face
    o->kernel                      = MGB_KERNEL_AUTO;
    o->device                      = 0;
    o->tile_cells                  = 0;
  }

  int
  mgb_solver_create(const mgb_options *opt, mgb_solver *out)
  {
    MGB_TRY
    if (!opt || !out)
      throw std::runtime_error("null argument");
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0)
      throw std::runtime_error("no CUDA device available: mgb200 has no CPU fallback");
    *out         = new mgb_solver_t{nullptr};
    (*out)->impl = new MultigridSolver(*opt);
    MGB_CATCH
  }

  int
  mgb_solver_destroy(mgb_solver s)
  {
    MGB_TRY
    if (s)
      {
        delete s->impl;
        delete s;
      }
    MGB_CATCH
  }

  int
  mgb_solver_sync(mgb_solver s)
  {
    MGB_TRY
    S(s).ctx.sync();
    MGB_CATCH
  }

  int
  mgb_solver_n_levels(mgb_solver s)
  {
    return s && s->impl ? s->impl->n_levels : -1;
  }

  int64_t
  mgb_solver_n_dofs(mgb_solver s, int level)
  {
    try
      {
        MultigridSolver &m = S(s);
        return m.levels[check_level(m, level)]->d.n_dofs;
      }
    catch (const std::exception &e)
      {
        g_last_error = e.what();
        return -1;
      }
  }

  int
  mgb_solver_partition(mgb_solver s, int level, int64_t info[12])
  {
    MGB_TRY
    MultigridSolver &m = S(s);
    fill_partition_info(*m.levels[check_level(m, level)], info);
    MGB_CATCH
  }

  int64_t
  mgb_solver_exchange_count(mgb_solver s)
  {
    return s && s->impl ? s->impl->ctx.exchanges : -1;
  }

  int
  mgb_comm_unique_id(unsigned char id[128])
  {
    MGB_TRY
    Nccl        &n = Nccl::get();
    ncclUniqueId uid;
    n.check(n.GetUniqueId(&uid), "ncclGetUniqueId");
    std::memcpy(id, &uid, 128);
    MGB_CATCH
  }

  int
  mgb_solver_dof_grid(mgb_solver s, int level, int nd[3])
  {
    MGB_TRY
    MultigridSolver &m = S(s);
    level              = check_level(m, level);
    for (int d = 0; d < 3; ++d)
      nd[d] = m.levels[level]->d.nd[d];
    MGB_CATCH
  }

  int
  mgb_solver_tile_cells(mgb_solver s, int level)
  {
    try
      {
        MultigridSolver &m = S(s);
        return m.levels[check_level(m, level)]->d.CW;
      }
    catch (const std::exception &e)
      {
        g_last_error = e.what();
        return -1;
      }
  }

  int
  mgb_solver_lex_map(mgb_solver s, int level, uint32_t *lex_to_dof)
  {
    MGB_TRY
    MultigridSolver &m = S(s);
    build_lex_map(m.levels[check_level(m, level)]->host, lex_to_dof);
    MGB_CATCH
  }

  int
  mgb_solver_solve(mgb_solver s, int do_analyze, double *reduction_rate, double *trace)
  {
    MGB_TRY
    const double r = S(s).solve(do_analyze != 0, trace);
    if (reduction_rate)
      *reduction_rate = r;
    MGB_CATCH
  }

  int
  mgb_solver_solve_cg(mgb_solver s, unsigned int *iterations, double *reduction_rate)
  {
    MGB_TRY
    auto r = S(s).solve_cg();
    if (iterations)
      *iterations = r.first;
    if (reduction_rate)
      *reduction_rate = r.second;
    MGB_CATCH
  }

  int
  mgb_solver_solve_cg_host(mgb_solver s, const double *rhs_host, double *solution_host,
                           unsigned int *iterations, double *reduction_rate)
  {
    MGB_TRY
    MultigridSolver &m = S(s);
    const int        L = m.n_levels - 1;
    const size_t     N = m.levels[L]->d.n_dofs;
    if (rhs_host)
      MGB_CUDA_CHECK(cudaMemcpyAsync(m.rhs[L]->p, rhs_host, N * sizeof(double), cudaMemcpyHostToDevice, m.ctx.stream));
    auto r = m.solve_cg();
    if (solution_host)
      {
        m.set_boundary_values(L, false);
        MGB_CUDA_CHECK(cudaMemcpyAsync(solution_host, m.solution[L]->p, N * sizeof(double), cudaMemcpyDeviceToHost, m.ctx.stream));
        m.ctx.sync();
      }
    if (iterations)
      *iterations = r.first;
    if (reduction_rate)
      *reduction_rate = r.second;
    MGB_CATCH
  }

  int
  mgb_solver_cg_history(mgb_solver s, double *history, int *n)
  {
    MGB_TRY
    MultigridSolver &m   = S(s);
    const int        cnt = std::min<int>(*n, (int)m.cg_history.size());
    for (int i = 0; i < cnt; ++i)
      history[i] = m.cg_history[i];
    *n = cnt;
    MGB_CATCH
  }

  int
  mgb_solver_l2_error(mgb_solver s, int level, double *error)
  {
    MGB_TRY
    MultigridSolver &m = S(s);
    *error             = m.compute_l2_error(check_level(m, level));
    MGB_CATCH
  }

  int
  mgb_solver_vcycle(mgb_solver s, double *dst_dev, const double *src_dev)
  {
    MGB_TRY
    S(s).vmult(dst_dev, src_dev);
    MGB_CATCH
  }

  int
  mgb_solver_vmult_with_residual_update(mgb_solver s, double *residual_dev, double *update_dev, double factor, double sums[2])
  {
    MGB_TRY
    const std::array<double, 2> r = S(s).vmult_with_residual_update(residual_dev, update_dev, factor);
    sums[0]                       = r[0];
    sums[1]                       = r[1];
    MGB_CATCH
  }

  int64_t
  mgb_solver_n_quadrature_points(mgb_solver s, int level)
  {
    try
      {
        MultigridSolver &m = S(s);
        return (int64_t)m.n_quadrature_points(check_level(m, level));
      }
    catch (const std::exception &e)
      {
        g_last_error = e.what();
        return -1;
      }
  }

  int
  mgb_solver_quadrature_points(mgb_solver s, int level, double *xyz_host)
  {
    MGB_TRY
    MultigridSolver &m = S(s);
    m.quadrature_points(check_level(m, level), xyz_host);
    MGB_CATCH
  }

  int
  mgb_solver_support_points(mgb_solver s, int level, double *xyz_host)
  {
    MGB_TRY
    MultigridSolver &m = S(s);
    m.support_points(check_level(m, level), xyz_host);
    MGB_CATCH
  }

  int
  mgb_solver_set_coefficient_values(mgb_solver s, int level, const double *c_host)
  {
    MGB_TRY
    MultigridSolver &m = S(s);
    m.set_coefficient_values(check_level(m, level), c_host);
    MGB_CATCH
  }

  int
  mgb_solver_reinit_smoothers(mgb_solver s)
  {
    MGB_TRY
    S(s).reinit_smoothers();
    MGB_CATCH
  }

  int
  mgb_solver_set_rhs_values(mgb_solver s, int level, const double *f_host, const double *u_boundary_host)
  {
    MGB_TRY
    MultigridSolver &m = S(s);
    m.set_rhs_values(check_level(m, level), f_host, u_boundary_host);
    MGB_CATCH
  }

  int
  mgb_solver_set_exact_solution_values(mgb_solver s, int level, const double *u_host)
  {
    MGB_TRY
    MultigridSolver &m = S(s);
    m.set_exact_solution_values(check_level(m, level), u_host);
    MGB_CATCH
  }

  int
  mgb_solver_minimal_surface_set_boundary_values(mgb_solver s, const double *u_host)
  {
    MGB_TRY
    S(s).ms_set_boundary_values(u_host);
    MGB_CATCH
  }

  int
  mgb_solver_minimal_surface_step(mgb_solver s, int first_time, double info[5])
  {
    MGB_TRY
    S(s).ms_newton_step(first_time != 0, info);
    MGB_CATCH
  }

  int
  mgb_solver_minimal_surface_residual(mgb_solver s, double *dst, const double *u, int first_time, double *norm)
  {
    MGB_TRY
    MultigridSolver &m = S(s);
    m.ms_require_one_gpu();
    const double r = m.ms_residual(dst, u, first_time != 0);
    m.ctx.sync();
    if (norm)
      *norm = r;
    MGB_CATCH
  }

  int
  mgb_solver_minimal_surface_level_solution(mgb_solver s, int level, float **dev_ptr)
  {
    MGB_TRY
    MultigridSolver &m = S(s);
    check_level(m, level);
    if ((int)m.coefficient_solutions.size() != m.n_levels)
      throw std::runtime_error("call mgb_solver_minimal_surface_set_boundary_values first");
    *dev_ptr = m.coefficient_solutions[level]->p;
    MGB_CATCH
  }

  int
  mgb_solver_do_matvec(mgb_solver s)
  {
    MGB_TRY
    MultigridSolver &m = S(s);
    const int        L = m.n_levels - 1;
    m.matrix_dp[L]->vmult(m.residual[L]->p, m.solution[L]->p);
    MGB_CATCH
  }

  int
  mgb_solver_do_matvec_smoother(mgb_solver s)
  {
    MGB_TRY
    MultigridSolver &m = S(s);
    const int        L = m.n_levels - 1;
    m.matrix[L]->vmult(m.solution_update[L]->p, m.defect[L]->p);
    MGB_CATCH
  }

  int
  mgb_solver_timings(mgb_solver s, double *timings, int reset)
  {
    MGB_TRY
    MultigridSolver &m = S(s);
    m.ctx.collect_ranges();
    const int nk = Ctx::n_classes_per_level;
    for (int l = 0; l < m.n_levels; ++l)
      for (int j = 0; j < 6; ++j)
        {
          const size_t c     = (size_t)l * nk + j;
          timings[l * 6 + j] = c < m.ctx.class_ms.size() ? 1e-3 * m.ctx.class_ms[c] : 0.;
        }
    if (reset)
      {
        std::fill(m.ctx.class_ms.begin(), m.ctx.class_ms.end(), 0.);
        std::fill(m.ctx.class_count.begin(), m.ctx.class_count.end(), 0);
      }
    MGB_CATCH
  }

  int
  mgb_solver_profile(mgb_solver s, int on)
  {
    MGB_TRY
    MultigridSolver &m = S(s);
    m.ctx.collect_ranges();
    m.ctx.profile = on != 0;
    MGB_CATCH
  }

  int
  mgb_solver_class_times(mgb_solver s, int level, double *ms, int64_t *count, int reset)
  {
    MGB_TRY
    MultigridSolver &m = S(s);
    level              = check_level(m, level);
    m.ctx.collect_ranges();
    const int nk = Ctx::n_classes_per_level;
    for (int j = 0; j < nk; ++j)
      {
        const size_t c = (size_t)level * nk + j;
        ms[j]          = c < m.ctx.class_ms.size() ? m.ctx.class_ms[c] : 0.;
        count[j]       = c < m.ctx.class_count.size() ? m.ctx.class_count[c] : 0;
        if (reset && c < m.ctx.class_ms.size())
          {
            m.ctx.class_ms[c]    = 0;
            m.ctx.class_count[c] = 0;
          }
      }
    MGB_CATCH
  }

  int
  mgb_timer_start(mgb_solver s)
  {
    MGB_TRY
    Ctx &c = S(s).ctx;
    for (int i = 0; i < 2; ++i)
      if (!c.timer_ev[i])
        MGB_CUDA_CHECK(cudaEventCreate(&c.timer_ev[i]));
    MGB_CUDA_CHECK(cudaEventRecord(c.timer_ev[0], c.stream));
    MGB_CATCH
  }

  int
  mgb_timer_stop(mgb_solver s, double *ms)
  {
    MGB_TRY
    Ctx &c = S(s).ctx;
    if (!c.timer_ev[0])
      throw std::runtime_error("mgb_timer_stop without mgb_timer_start");
    MGB_CUDA_CHECK(cudaEventRecord(c.timer_ev[1], c.stream));
    MGB_CUDA_CHECK(cudaEventSynchronize(c.timer_ev[1]));
    float f = 0;
    MGB_CUDA_CHECK(cudaEventElapsedTime(&f, c.timer_ev[0], c.timer_ev[1]));
    *ms = f;
    MGB_CATCH
  }

  int
  mgb_solver_vector(mgb_solver s, int which, int level, void **dev_ptr)
  {
    MGB_TRY
    MultigridSolver &m = S(s);
    level              = check_level(m, level);
    switch (which)
      {
        case 0: *dev_ptr = m.solution[level]->p; break;
        case 1: *dev_ptr = m.rhs[level]->p; break;
        case 2: *dev_ptr = m.residual[level]->p; break;
        case 3: *dev_ptr = m.defect[level]->p; break;
        case 4: *dev_ptr = m.t[level]->p; break;
        case 5: *dev_ptr = m.solution_update[level]->p; break;
        case 6: *dev_ptr = m.matrix[level]->inverse_diagonal.p; break;
        default: throw std::runtime_error("unknown vector id");
      }
    MGB_CATCH
  }

  int
  mgb_solver_get_solution(mgb_solver s, double *solution_host)
  {
    MGB_TRY
    MultigridSolver &m = S(s);
    const int        L = m.n_levels - 1;
    m.set_boundary_values(L, false);
    MGB_CUDA_CHECK(cudaMemcpyAsync(solution_host, m.solution[L]->p, m.solution[L]->n * sizeof(double), cudaMemcpyDeviceToHost, m.ctx.stream));
    m.ctx.sync();
    MGB_CATCH
  }

  int
  mgb_op_vmult(mgb_solver s, int level, int precision, void *dst, const void *src)
  {
    MGB_TRY
    MultigridSolver &m = S(s);
    level              = check_level(m, level);
    if (precision == MGB_F64)
      m.matrix_dp[level]->vmult((double *)dst, (const double *)src);
    else
      m.matrix[level]->vmult((float *)dst, (const float *)src);
    MGB_CATCH
  }

  int
  mgb_op_vmult_add(mgb_solver s, int level, int precision, void *dst, const void *src)
  {
    MGB_TRY
    MultigridSolver &m = S(s);
    level              = check_level(m, level);
    // Base::vmult_add = apply_add, then identity rows added for the constrained DoFs
    if (precision == MGB_F64)
      {
        m.matrix_dp[level]->vmult_add((double *)dst, (const double *)src);
        double         *d = (double *)dst;
        const double   *x = (const double *)src;
        const uint32_t *c = m.levels[level]->constrained.p;
        foreach (m.ctx, m.levels[level]->constrained.n, [=] __device__(size_t i) { d[c[i]] += x[c[i]]; });
      }
    else
      {
        m.matrix[level]->vmult_add((float *)dst, (const float *)src);
        float          *d = (float *)dst;
        const float    *x = (const float *)src;
        const uint32_t *c = m.levels[level]->constrained.p;
        foreach (m.ctx, m.levels[level]->constrained.n, [=] __device__(size_t i) { d[c[i]] += x[c[i]]; });
      }
    MGB_CATCH
  }

  int
  mgb_op_Tvmult(mgb_solver s, int level, int precision, void *dst, const void *src)
  {
    return mgb_op_vmult(s, level, precision, dst, src); // symmetric operator
  }

  int
  mgb_op_vmult_host(mgb_solver s, int level, int precision, void *dst_host, const void *src_host)
  {
    MGB_TRY
    MultigridSolver &m  = S(s);
    level               = check_level(m, level);
    const size_t N      = m.levels[level]->d.n_dofs;
    const size_t bytes  = N * (precision == MGB_F64 ? sizeof(double) : sizeof(float));
    void        *src    = precision == MGB_F64 ? (void *)m.solution[level]->p : (void *)m.defect[level]->p;
    void        *dst    = precision == MGB_F64 ? (void *)m.residual[level]->p : (void *)m.t[level]->p;
    MGB_CUDA_CHECK(cudaMemcpyAsync(src, src_host, bytes, cudaMemcpyHostToDevice, m.ctx.stream));
    if (precision == MGB_F64)
      m.matrix_dp[level]->vmult((double *)dst, (const double *)src);
    else
      m.matrix[level]->vmult((float *)dst, (const float *)src);
    MGB_CUDA_CHECK(cudaMemcpyAsync(dst_host, dst, bytes, cudaMemcpyDeviceToHost, m.ctx.stream));
    m.ctx.sync();
    MGB_CATCH
  }

  int
  mgb_op_vmult_residual(mgb_solver s, int level, int precision, const void *rhs, const void *lhs, void *residual)
  {
    MGB_TRY
    MultigridSolver &m = S(s);
    level              = check_level(m, level);
    if (precision == MGB_F64)
      m.matrix_dp[level]->vmult_residual((const double *)rhs, (const double *)lhs, (double *)residual);
    else
      m.matrix[level]->vmult_residual((const float *)rhs, (const float *)lhs, (float *)residual);
    MGB_CATCH
  }

  int
  mgb_op_vmult_with_cg_update(mgb_solver s, int level, int precision, double alpha, double beta,
                              const void *r, void *q, void *p, void *x, double sums[4])
  {
    MGB_TRY
    MultigridSolver &m = S(s);
    level              = check_level(m, level);
    if (precision == MGB_F64)
      m.matrix_dp[level]->vmult_with_cg_update(alpha, beta, (const double *)r, (double *)q, (double *)p, (double *)x, sums);
    else
      m.matrix[level]->vmult_with_cg_update((float)alpha, (float)beta, (const float *)r, (float *)q, (float *)p, (float *)x, sums);
    MGB_CATCH
  }

  int
  mgb_op_inverse_diagonal(mgb_solver s, int level, int precision, void *inv_diag)
  {
    MGB_TRY
    MultigridSolver &m = S(s);
    level              = check_level(m, level);
    if (precision == MGB_F64)
      {
        m.matrix_dp[level]->compute_diagonal();
        MGB_CUDA_CHECK(cudaMemcpyAsync(inv_diag, m.matrix_dp[level]->inverse_diagonal.p, m.matrix_dp[level]->m() * sizeof(double), cudaMemcpyDeviceToDevice, m.ctx.stream));
      }
    else
      MGB_CUDA_CHECK(cudaMemcpyAsync(inv_diag, m.matrix[level]->inverse_diagonal.p, m.matrix[level]->m() * sizeof(float), cudaMemcpyDeviceToDevice, m.ctx.stream));
    MGB_CATCH
  }

  int
  mgb_op_set_kernel(mgb_solver s, int kernel)
  {
    MGB_TRY
    MultigridSolver &m = S(s);
    for (int l = 0; l < m.n_levels; ++l)
      {
        m.matrix[l]->kernel = m.matrix_dp[l]->kernel = kernel;
        m.transfer[l].kernel = m.mg_transfer_no_boundary[l].kernel = kernel;
      }
    m.ctx.sync();
    for (auto &g : m.cycle_graphs) // captured with the previous kernel selection
      if (g.exec)
        cudaGraphExecDestroy(g.exec);
    m.cycle_graphs.clear();
    m.choose_graph_level();
    MGB_CATCH
  }

  int
  mgb_smoother_reinit(mgb_solver s, int level, double smoothing_range, int degree, int eig_cg_n_iterations)
  {
    MGB_TRY
    MultigridSolver &m = S(s);
    level              = check_level(m, level);
    m.ctx.sync();
    for (auto &g : m.cycle_graphs) // captured with the previous smoother
      if (g.exec)
        cudaGraphExecDestroy(g.exec);
    m.cycle_graphs.clear();
    Chebyshev<float> &c   = *m.smooth[level];
    c.smoothing_range     = smoothing_range;
    c.degree              = degree < 0 ? Chebyshev<float>::invalid_degree : (unsigned int)degree;
    c.eig_cg_n_iterations = eig_cg_n_iterations < 0 ? (unsigned int)m.matrix[level]->m_global() : (unsigned int)eig_cg_n_iterations;
    c.initialize(m.ctx, *m.matrix[level], m.basis);
    MGB_CATCH
  }

  int
  mgb_solver_vcycle_inplace(mgb_solver s)
  {
    MGB_TRY
    MultigridSolver &m = S(s);
    m.v_cycle(m.n_levels - 1, 1);
    MGB_CATCH
  }

  int
  mgb_smoother_info(mgb_solver s, int level, double info[6])
  {
    MGB_TRY
    MultigridSolver &m        = S(s);
    const Chebyshev<float> &c = *m.smooth[check_level(m, level)];
    info[0]                   = c.min_eig;
    info[1]                   = c.max_eig;
    info[2]                   = c.theta;
    info[3]                   = c.delta;
    info[4]                   = c.degree;
    info[5]                   = c.cg_iterations;
    MGB_CATCH
  }

  static void
  smoother_apply(MultigridSolver &m, int level, float *x, const float *b, bool step)
  {
    // run on a solver-owned vector (the smoother swaps buffers), then copy back
    DVec<float> &work = *m.t[level];
    const size_t N    = work.n;
    MGB_CUDA_CHECK(cudaMemcpyAsync(work.p, x, N * sizeof(float), cudaMemcpyDeviceToDevice, m.ctx.stream));
    if (step)
      m.smooth[level]->step(work, b);
    else
      m.smooth[level]->vmult(work, b);
    MGB_CUDA_CHECK(cudaMemcpyAsync(x, work.p, N * sizeof(float), cudaMemcpyDeviceToDevice, m.ctx.stream));
  }

  int
  mgb_smoother_vmult(mgb_solver s, int level, float *x, const float *b)
  {
    MGB_TRY
    MultigridSolver &m = S(s);
    smoother_apply(m, check_level(m, level), x, b, false);
    MGB_CATCH
  }

  int
  mgb_smoother_step(mgb_solver s, int level, float *x, const float *b)
  {
    MGB_TRY
    MultigridSolver &m = S(s);
    smoother_apply(m, check_level(m, level), x, b, true);
    MGB_CATCH
  }

  int
  mgb_transfer_prolongate(mgb_solver s, int level, float *fine, const float *coarse)
  {
    MGB_TRY
    MultigridSolver &m = S(s);
    level              = check_level(m, level);
    if (level == 0)
      throw std::runtime_error("no transfer to level 0");
    m.transfer[level].prolongate(fine, coarse);
    MGB_CATCH
  }

  int
  mgb_transfer_prolongate_and_add(mgb_solver s, int level, float *fine, const float *coarse)
  {
    MGB_TRY
    MultigridSolver &m = S(s);
    level              = check_level(m, level);
    if (level == 0)
      throw std::runtime_error("no transfer to level 0");
    m.transfer[level].prolongate_and_add(fine, coarse);
    MGB_CATCH
  }

  int
  mgb_transfer_restrict_and_add(mgb_solver s, int level, float *coarse, const float *fine)
  {
    MGB_TRY
    MultigridSolver &m = S(s);
    level              = check_level(m, level);
    if (level == 0)
      throw std::runtime_error("no transfer from level 0");
    m.transfer[level].restrict_and_add(coarse, fine);
    MGB_CATCH
  }

  int
  mgb_transfer_prolongate_no_boundary(mgb_solver s, int level, double *fine, const double *coarse)
  {
    MGB_TRY
    MultigridSolver &m = S(s);
    level              = check_level(m, level);
    if (level == 0)
      throw std::runtime_error("no transfer to level 0");
    m.mg_transfer_no_boundary[level].prolongate(fine, coarse);
    MGB_CATCH
  }

  int
  mgb_vec_dot(mgb_solver s, int level, int precision, const void *a, const void *b, double *result)
  {
    MGB_TRY
    MultigridSolver &m = S(s);
    level              = check_level(m, level);
    if (precision == MGB_F64)
      vec_ops<double>(m, level, 0, const_cast<void *>(a), b, 0, 0, result);
    else
      vec_ops<float>(m, level, 0, const_cast<void *>(a), b, 0, 0, result);
    MGB_CATCH
  }

  int
  mgb_vec_convert(mgb_solver s, int level, int dst_precision, void *dst, const void *src)
  {
    MGB_TRY
    MultigridSolver &m = S(s);
    level              = check_level(m, level);
    const size_t N     = m.levels[level]->d.n_dofs;
    if (dst_precision == MGB_F64)
      convert(m.ctx, static_cast<double *>(dst), static_cast<const float *>(src), N);
    else if (dst_precision == MGB_F32)
      convert(m.ctx, static_cast<float *>(dst), static_cast<const double *>(src), N);
    else
      throw std::runtime_error("precision must be MGB_F32 or MGB_F64");
    MGB_CATCH
  }

  int
  mgb_op_compute_residual(mgb_solver s, int level, double *dst, const double *src, const double *f_values_host)
  {
    MGB_TRY
    MultigridSolver &m = S(s);
    if (!dst || !src)
      throw std::runtime_error("null argument");
    m.compute_residual(check_level(m, level), dst, src, f_values_host);
    MGB_CATCH
  }

  int
  mgb_op_el(mgb_solver s, int level, int precision, int64_t row, double *value)
  {
    MGB_TRY
    MultigridSolver &m = S(s);
    level              = check_level(m, level);
    if (row < 0 || (size_t)row >= m.levels[level]->d.n_dofs || !value)
      throw std::runtime_error("row out of range");
    // MatrixFreeOperators::Base::el(row, row) = 1 / inverse_diagonal(row)
    if (precision == MGB_F64)
      {
        if (!m.matrix_dp[level]->inverse_diagonal.p)
          m.matrix_dp[level]->compute_diagonal();
        double v;
        MGB_CUDA_CHECK(cudaMemcpyAsync(&v, m.matrix_dp[level]->inverse_diagonal.p + row, sizeof(double), cudaMemcpyDeviceToHost, m.ctx.stream));
        m.ctx.sync();
        *value = 1. / v;
      }
    else if (precision == MGB_F32)
      {
        if (!m.matrix[level]->inverse_diagonal.p)
          m.matrix[level]->compute_diagonal();
        float v;
        MGB_CUDA_CHECK(cudaMemcpyAsync(&v, m.matrix[level]->inverse_diagonal.p + row, sizeof(float), cudaMemcpyDeviceToHost, m.ctx.stream));
        m.ctx.sync();
        *value = 1. / (double)v;
      }
    else
      throw std::runtime_error("precision must be MGB_F32 or MGB_F64");
    MGB_CATCH
  }

  int
  mgb_vec_sadd(mgb_solver s, int level, int precision, void *y, double sy, double ax, const void *x)
  {
    MGB_TRY
    MultigridSolver &m = S(s);
    level              = check_level(m, level);
    double dummy;
    if (precision == MGB_F64)
      vec_ops<double>(m, level, 1, y, x, sy, ax, &dummy);
    else
      vec_ops<float>(m, level, 1, y, x, sy, ax, &dummy);
    MGB_CATCH
  }

  int
  mgb_vec_set(mgb_solver s, int level, int precision, void *y, double value)
  {
    MGB_TRY
    MultigridSolver &m = S(s);
    level              = check_level(m, level);
    double dummy;
    if (precision == MGB_F64)
      vec_ops<double>(m, level, 2, y, nullptr, value, 0, &dummy);
    else
      vec_ops<float>(m, level, 2, y, nullptr, value, 0, &dummy);
    MGB_CATCH
  }

  int
  mgb_vec_update_ghost_values(mgb_solver s, int level, int precision, void *v)
  {
    MGB_TRY
    MultigridSolver &m = S(s);
    level              = check_level(m, level);
    if (precision == MGB_F64)
      m.levels[level]->update_ghosts(static_cast<double *>(v));
    else
      m.levels[level]->update_ghosts(static_cast<float *>(v));
    MGB_CATCH
  }

  int
  mgb_vec_compress_add(mgb_solver s, int level, int precision, void *v)
  {
    MGB_TRY
    MultigridSolver &m = S(s);
    level              = check_level(m, level);
    if (precision == MGB_F64)
      compress_add_all(*m.levels[level], static_cast<double *>(v));
    else
      compress_add_all(*m.levels[level], static_cast<float *>(v));
    MGB_CATCH
  }

  int
  mgb_vec_zero_out_ghost_values(mgb_solver s, int level, int precision, void *v)
  {
    MGB_TRY
    MultigridSolver &m = S(s);
    level              = check_level(m, level);
    if (precision == MGB_F64)
      zero_ghosts(*m.levels[level], static_cast<double *>(v));
    else
      zero_ghosts(*m.levels[level], static_cast<float *>(v));
    MGB_CATCH
  }

  int
  mgb_device_alloc(mgb_solver s, size_t bytes, void **dev_ptr)
  {
    MGB_TRY
    if (!s) // no solver (DG operator vectors): current device, synchronous
      {
        MGB_CUDA_CHECK(cudaMalloc(dev_ptr, bytes));
        MGB_CUDA_CHECK(cudaMemset(*dev_ptr, 0, bytes));
        return 0;
      }
    MGB_CUDA_CHECK(cudaSetDevice(S(s).ctx.device));
    MGB_CUDA_CHECK(cudaMalloc(dev_ptr, bytes));
    MGB_CUDA_CHECK(cudaMemsetAsync(*dev_ptr, 0, bytes, S(s).ctx.stream));
    MGB_CATCH
  }

  int
  mgb_device_free(mgb_solver s, void *dev_ptr)
  {
    MGB_TRY
    if (s)
      S(s).ctx.sync();
    else
      MGB_CUDA_CHECK(cudaDeviceSynchronize());
    MGB_CUDA_CHECK(cudaFree(dev_ptr));
    MGB_CATCH
  }

  int
  mgb_copy_to_device(mgb_solver s, void *dst_dev, const void *src_host, size_t bytes)
  {
    MGB_TRY
    if (!s)
      {
        MGB_CUDA_CHECK(cudaDeviceSynchronize());
        MGB_CUDA_CHECK(cudaMemcpy(dst_dev, src_host, bytes, cudaMemcpyHostToDevice));
        return 0;
      }
    MGB_CUDA_CHECK(cudaMemcpyAsync(dst_dev, src_host, bytes, cudaMemcpyHostToDevice, S(s).ctx.stream));
    S(s).ctx.sync();
    MGB_CATCH
  }

  int
  mgb_copy_to_host(mgb_solver s, void *dst_host, const void *src_dev, size_t bytes)
  {
    MGB_TRY
    if (!s)
      {
        MGB_CUDA_CHECK(cudaDeviceSynchronize());
        MGB_CUDA_CHECK(cudaMemcpy(dst_host, src_dev, bytes, cudaMemcpyDeviceToHost));
        return 0;
      }
    MGB_CUDA_CHECK(cudaMemcpyAsync(dst_host, src_dev, bytes, cudaMemcpyDeviceToHost, S(s).ctx.stream));
    S(s).ctx.sync();
    MGB_CATCH
  }

  int64_t
  mgb_solver_launch_count(mgb_solver s)
  {
    return s && s->impl ? s->impl->ctx.launches : -1;
  }

  // ---- host-only helpers (no GPU needed): used by the CPU test-suite for the host logic
  int
  mgb_host_basis(int p, double *S, double *G, double *D, double *M1, double *K1, double *P1,
                 double *nodes, double *xq, double *wq)
  {
    MGB_TRY
    if (p < 1 || p > 8)
      throw std::runtime_error("polynomial degree must be between 1 and 8");
    const Basis b = make_basis(p);
    std::copy(b.S.begin(), b.S.end(), S);
    std::copy(b.G.begin(), b.G.end(), G);
    std::copy(b.D.begin(), b.D.end(), D);
    std::copy(b.M1.begin(), b.M1.end(), M1);
    std::copy(b.K1.begin(), b.K1.end(), K1);
    std::copy(b.P1.begin(), b.P1.end(), P1);
    std::copy(b.nodes.begin(), b.nodes.end(), nodes);
    std::copy(b.xq.begin(), b.xq.end(), xq);
    std::copy(b.wq.begin(), b.wq.end(), wq);
    MGB_CATCH
  }

  // numbering of one box level with all faces Dirichlet: sizes, and lex_to_dof (n_dofs entries) when non-NULL
  int
  mgb_host_numbering(int p, const int ncell[3], int tile_cells, int64_t *n_dofs, int *tile_cells_used,
                     int64_t *n_tiles, int64_t *n_constrained, uint32_t *lex_to_dof)
  {
    MGB_TRY
    if (p < 1 || p > 8)
      throw std::runtime_error("polynomial degree must be between 1 and 8");
    HostLevel    lvl;
    const double lo[3] = {0, 0, 0}, hi[3] = {1, 1, 1};
    build_box_level(p, ncell, choose_tile_cells(p, tile_cells), lo, hi, 0x3fu, lvl);
    *n_dofs          = lvl.d.n_dofs;
    *tile_cells_used = lvl.d.CW;
    *n_tiles         = (int64_t)lvl.d.nt[0] * lvl.d.nt[1];
    *n_constrained   = (int64_t)lvl.constrained.size();
    if (lex_to_dof)
      build_lex_map(lvl, lex_to_dof);
    MGB_CATCH
  }

  int
  mgb_host_partition(int p, const int n_subdivisions[3], int level, int tile_cells, int n_ranks, int rank,
                     int64_t min_dofs, int64_t info[12], int64_t plan[8], uint32_t *lex_to_dof)
  {
    MGB_TRY
    if (p < 1 || p > 8)
      throw std::runtime_error("polynomial degree must be between 1 and 8");
    if (n_ranks < 1 || rank < 0 || rank >= n_ranks || level < 0 || level > 20)
      throw std::runtime_error("invalid rank or level");
    struct
    {
      HostLevel host;
      LevelDesc d;
      Slab      slab;
    } lv;
    int nc[3];
    for (int c = 0; c < 3; ++c)
      nc[c] = n_subdivisions[c] << level;
    const double lo[3] = {0, 0, 0}, hi[3] = {1, 1, 1};
    lv.slab            = make_slab(p, n_subdivisions, level, n_ranks, rank, 2, min_dofs);
    build_slab_level(p, nc, choose_tile_cells(p, tile_cells), lo, hi, 0x3fu, lv.slab, lv.host);
    lv.d = lv.host.d;
    fill_partition_info(lv, info);
    const ExchangePlan e = make_exchange_plan(lv.d);
    const uint64_t    *r[4] = {e.send_up, e.recv_dn, e.send_dn, e.recv_up};
    for (int i = 0; i < 4; ++i)
      {
        plan[2 * i]     = r[i][0];
        plan[2 * i + 1] = r[i][1];
      }
    if (lex_to_dof)
      build_lex_map(lv.host, lex_to_dof);
    MGB_CATCH
  }
}
