// mgb_noalloc_dot: the dot product of two device vectors (the reduction behind mgb_vec_dot / the sums of the CG loop)
// for callers that may not allocate: code under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc.
// The library's own reduction (k_reduce in solver.cu) accumulates into a scratch buffer that every solver handle
// allocates with cudaMalloc at set-up; this variant keeps its scratch -- one partial sum per chunk of the vectors and a
// ticket counter -- in a file-scope __device__ array, so the call needs no handle and allocates nothing: no cudaMalloc,
// no cudaFree, no allocating library routine.  No counterpart in the
// reference.
//
// Limits and rules
//   * n <= MGB_NOALLOC_DOT_MAX_N = 65536 chunks x 4096 entries = 268 435 456 entries; larger calls are rejected.
//   * the array (512.25 KiB) is part of the library image: every GPU that loads libmgb200.so holds it until exit.
//   * every call on a device shares the array and it keeps the last call's contents.  The only part that must start at
//     zero, the ticket counter, is cleared on the caller's stream at the start of every call; the partial sums of a call
//     are all written before they are read, so stale ones are never used.
//   * two calls that overlap on the same GPU (different streams, or two host threads) race on the array.  Calls on one
//     stream are ordered and safe.
//   * the address of the array is looked up once per device with cudaGetSymbolAddress, with that device current (in host
//     code the array's name is a host-side stand-in, not a device address).  The first call on a device does this; a
//     call with n = 0 does nothing else and is the way to get it done before a capture starts.
//   * summation order is fixed (chunk by chunk, then the chunks in order), so the result is bitwise reproducible and does
//     not depend on the GPU's occupancy, unlike the atomicAdd accumulation of mgb_vec_dot.
#include <cuda_runtime.h>

#include <cstdint>
#include <mutex>
#include <string>

#include "../../include/mgb200.h"

void mgb_set_last_error(const char *msg); // solver.cu

constexpr int     CHUNK        = 4096; // entries per block: 256 threads x 16
constexpr int     MAX_PARTIALS = 65536;
constexpr int64_t MAX_N        = (int64_t)MAX_PARTIALS * CHUNK;
constexpr int     MAX_DEVICES  = 64;
static_assert(MAX_N == MGB_NOALLOC_DOT_MAX_N, "include/mgb200.h states the limit");

// [0, MAX_PARTIALS): partial sums; [MAX_PARTIALS]: ticket counter (a double-sized slot used as unsigned int)
static __device__ __align__(256) double g_noalloc_dot_scratch[MAX_PARTIALS + 32];

namespace
{
  template <typename T>
  __global__ void
  __launch_bounds__(256)
    k_noalloc_dot(const int64_t n, const T *__restrict__ x, const T *__restrict__ y, double *scratch, double *result)
  {
    const int64_t begin = (int64_t)blockIdx.x * CHUNK;
    double        acc   = 0;
#pragma unroll 4
    for (int k = 0; k < CHUNK / 256; ++k)
      {
        const int64_t i = begin + k * 256 + threadIdx.x;
        if (i < n)
          acc += (double)x[i] * (double)y[i];
      }
    __shared__ double red[8];
    __shared__ bool   last;
    for (int o = 16; o > 0; o >>= 1)
      acc += __shfl_down_sync(0xffffffffu, acc, o);
    if ((threadIdx.x & 31) == 0)
      red[threadIdx.x >> 5] = acc;
    __syncthreads();
    unsigned int *ticket = reinterpret_cast<unsigned int *>(scratch + MAX_PARTIALS);
    if (threadIdx.x == 0)
      {
        double v = 0;
        for (int w = 0; w < 8; ++w)
          v += red[w];
        scratch[blockIdx.x] = v;
        __threadfence();
        last = atomicAdd(ticket, 1u) == gridDim.x - 1;
      }
    __syncthreads();
    if (!last)
      return;
    // the last block to finish adds the partial sums in index order: thread t takes a contiguous range
    __threadfence();
    const int     nb  = gridDim.x;
    const int     per = (nb + 255) / 256;
    double        s   = 0;
    const int     b0 = threadIdx.x * per, b1 = min(nb, b0 + per);
    for (int b = b0; b < b1; ++b)
      s += reinterpret_cast<volatile double *>(scratch)[b];
    __shared__ double all[256];
    all[threadIdx.x] = s;
    __syncthreads();
    if (threadIdx.x == 0)
      {
        double v = 0;
        for (int t = 0; t < 256; ++t)
          v += all[t];
        *result = v;
        *ticket = 0; // leave the counter clean for the next call (the next call clears it as well)
      }
  }

  std::mutex g_mutex;
  double    *g_scratch_address[MAX_DEVICES] = {};

  int
  fail(const std::string &msg)
  {
    mgb_set_last_error(msg.c_str());
    return 1;
  }
} // namespace

extern "C" int
mgb_noalloc_dot(int precision, int64_t n, const void *x_dev, const void *y_dev, double *result_dev, void *cuda_stream)
{
  if (precision != MGB_F32 && precision != MGB_F64)
    return fail("mgb_noalloc_dot: precision must be MGB_F32 or MGB_F64");
  if (n < 0)
    return fail("mgb_noalloc_dot: negative length");
  if (n > MAX_N)
    return fail("mgb_noalloc_dot: " + std::to_string(n) + " entries exceed the maximum of " + std::to_string(MAX_N) +
                " the static scratch array is sized for");
  int device = -1;
  if (cudaGetDevice(&device) != cudaSuccess || device < 0 || device >= MAX_DEVICES)
    return fail("mgb_noalloc_dot: no current CUDA device (the call has no CPU fallback)");
  double *scratch = nullptr;
  {
    // set-up, once per device, with that device current: the device address of the static array
    std::lock_guard<std::mutex> lock(g_mutex);
    if (!g_scratch_address[device])
      {
        void             *p = nullptr;
        const cudaError_t e = cudaGetSymbolAddress(&p, g_noalloc_dot_scratch);
        if (e != cudaSuccess || !p)
          return fail(std::string("mgb_noalloc_dot: cudaGetSymbolAddress failed: ") + cudaGetErrorString(e));
        g_scratch_address[device] = static_cast<double *>(p);
      }
    scratch = g_scratch_address[device];
  }
  if (n == 0)
    return 0; // set-up only
  if (!x_dev || !y_dev || !result_dev)
    return fail("mgb_noalloc_dot: null vector or result pointer");
  cudaStream_t   stream = static_cast<cudaStream_t>(cuda_stream);
  const unsigned blocks = (unsigned)((n + CHUNK - 1) / CHUNK);
  cudaError_t    e      = cudaMemsetAsync(scratch + MAX_PARTIALS, 0, sizeof(double), stream); // the ticket counter
  if (e == cudaSuccess)
    {
      if (precision == MGB_F64)
        k_noalloc_dot<double><<<blocks, 256, 0, stream>>>(n, static_cast<const double *>(x_dev),
                                                           static_cast<const double *>(y_dev), scratch, result_dev);
      else
        k_noalloc_dot<float><<<blocks, 256, 0, stream>>>(n, static_cast<const float *>(x_dev), static_cast<const float *>(y_dev),
                                                          scratch, result_dev);
      e = cudaGetLastError();
    }
  if (e != cudaSuccess)
    return fail(std::string("mgb_noalloc_dot: CUDA error: ") + cudaGetErrorString(e));
  return 0;
}
