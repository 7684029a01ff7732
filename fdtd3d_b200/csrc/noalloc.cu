/*
 * fdtd_sumsq_noalloc -- the |field|^2 reduction for callers that must not allocate device memory.
 *
 * Who needs this.  fdtd_field_sumsq works on a solver handle and uses partial-sum scratch that the handle
 * obtained from cudaMalloc when it was created.  Some users of this library cannot go through an
 * allocating path at the moment they want the number:
 *   - a monitoring hook that runs between chunks of time steps while the next chunk is already queued:
 *     cudaMalloc / cudaFree may synchronise the whole device, which would stall the queued steps;
 *   - a host application that has handed the GPU's memory to its own pool (a framework's caching
 *     allocator, a fixed arena) and forbids any other allocation, or whose big run has simply filled
 *     HBM: a raw cudaMalloc there fails or fragments the pool;
 *   - code that holds a field in its own device buffer and has no solver handle at all.
 * For them this entry point takes the input and the output as plain device pointers and keeps its only
 * scratch -- one partial sum per thread block -- in a fixed array that is part of the loaded library image.
 *
 * The price of a fixed array is a size limit: at most FDTD_SUMSQ_NOALLOC_MAX_COUNT scalars per call (a 512^3
 * real double field is half of that).  Larger inputs are refused with a message; split them and add the
 * results.  Every GPU that loads the library carries the array (256 KiB) until the process exits.
 */
#include <cstdio>
#include <mutex>

#include <cuda_runtime.h>

#include "../../include/fdtd3d_b200.h"

namespace
{

constexpr int THREADS = 256;
constexpr int PER_THREAD = 32;
constexpr long long CHUNK = (long long) THREADS * PER_THREAD; /* scalars summed by one thread block */
constexpr int MAX_PARTIALS = 32768;                           /* 256 KiB of doubles */
constexpr int MAX_DEVICES = 64;

static_assert (CHUNK * MAX_PARTIALS == FDTD_SUMSQ_NOALLOC_MAX_COUNT, "header and kernel disagree on the limit");

/* The scratch: one partial sum per thread block of the first pass.  File scope, so it lives in the module's
 * global segment on every device that loads the library; no allocation call ever creates or frees it.
 * In host code this name is only a handle for the runtime, not a device address (see scratch_for_device). */
static __device__ __align__ (256) double g_partials[MAX_PARTIALS];

template <typename T>
__global__ void __launch_bounds__ (THREADS)
sumsq_chunks (const T *in, long long count, double *partials)
{
  __shared__ double red[THREADS / 32];
  const long long base = (long long) blockIdx.x * CHUNK;
  double acc = 0.0;
#pragma unroll 4
  for (int r = 0; r < PER_THREAD; ++r)
  {
    const long long i = base + (long long) r * THREADS + threadIdx.x;
    if (i < count)
    {
      const double v = (double) in[i];
      acc += v * v;
    }
  }
#pragma unroll
  for (int off = 16; off > 0; off >>= 1) acc += __shfl_down_sync (0xffffffffu, acc, off);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = acc;
  __syncthreads ();
  if (threadIdx.x == 0)
  {
    double t = 0.0;
    for (int w = 0; w < THREADS / 32; ++w) t += red[w];
    partials[blockIdx.x] = t; /* plain store: every slot the second pass reads is rewritten by this call */
  }
}

__global__ void __launch_bounds__ (THREADS)
sumsq_combine (const double *partials, int n, double *out)
{
  __shared__ double red[THREADS];
  double acc = 0.0;
  for (int i = threadIdx.x; i < n; i += THREADS) acc += partials[i];
  red[threadIdx.x] = acc;
  __syncthreads ();
  for (int s = THREADS / 2; s > 0; s >>= 1)
  {
    if (threadIdx.x < s) red[threadIdx.x] += red[threadIdx.x + s];
    __syncthreads ();
  }
  if (threadIdx.x == 0) *out = red[0];
}

/* Device address of g_partials on the CURRENT device, looked up once per device and remembered.  The name
 * g_partials cannot be handed to a kernel from host code: there it stands for the host-side shadow of the
 * symbol, so a kernel would fault on it or, on systems where the GPU can reach host memory, quietly use
 * the wrong storage.  cudaGetSymbolAddress translates it for the device that is current when it is called.
 * The remembered address belongs to the device's current primary context: cudaDeviceReset (or anything else that
 * tears that context down) unloads the module, and the address is then stale -- this entry point must not be used
 * across such a reset (include/fdtd3d_b200.h says so). */
cudaError_t scratch_for_device (int device, double **out)
{
  static std::mutex lock;
  static double *known[MAX_DEVICES] = {};
  std::lock_guard<std::mutex> guard (lock);
  if (!known[device])
  {
    double *ptr = NULL;
    cudaError_t e = cudaGetSymbolAddress ((void **) &ptr, g_partials);
    if (e != cudaSuccess) return e;
    known[device] = ptr;
  }
  *out = known[device];
  return cudaSuccess;
}

int refuse (char *errbuf, int64_t errbuf_len, int code, const char *fmt, long long a, long long b)
{
  if (errbuf && errbuf_len > 0) snprintf (errbuf, (size_t) errbuf_len, fmt, a, b);
  return code;
}

} /* namespace */

extern "C" int fdtd_sumsq_noalloc (const void *dev_in, int value_type, int64_t count, double *dev_out, void *cuda_stream,
                                   char *errbuf, int64_t errbuf_len)
{
  if (errbuf && errbuf_len > 0) errbuf[0] = 0;
  if (!dev_in || !dev_out) return refuse (errbuf, errbuf_len, FDTD_INVALID_ARGUMENT, "fdtd_sumsq_noalloc: NULL device pointer", 0, 0);
  if (value_type != 4 && value_type != 8)
    return refuse (errbuf, errbuf_len, FDTD_INVALID_ARGUMENT, "fdtd_sumsq_noalloc: value_type %lld is neither 4 (float) nor 8 (double)", value_type, 0);
  if (count < 0)
    return refuse (errbuf, errbuf_len, FDTD_INVALID_ARGUMENT, "fdtd_sumsq_noalloc: negative count %lld", count, 0);
  if (count > FDTD_SUMSQ_NOALLOC_MAX_COUNT)
    return refuse (errbuf, errbuf_len, FDTD_INVALID_ARGUMENT,
                   "fdtd_sumsq_noalloc: %lld scalars exceed the fixed scratch of this entry point (limit %lld per call); "
                   "split the array and add the partial results, or use fdtd_field_sumsq on a solver handle",
                   count, (long long) FDTD_SUMSQ_NOALLOC_MAX_COUNT);

  int device = -1;
  cudaError_t e = cudaGetDevice (&device);
  if (e != cudaSuccess || device < 0)
    return refuse (errbuf, errbuf_len, FDTD_NO_DEVICE, "fdtd_sumsq_noalloc: no current CUDA device (cudaGetDevice: %lld)", (long long) e, 0);
  if (device >= MAX_DEVICES)
    return refuse (errbuf, errbuf_len, FDTD_INVALID_ARGUMENT, "fdtd_sumsq_noalloc: device ordinal %lld beyond the %lld this entry point tracks", device, MAX_DEVICES);
  double *partials = NULL;
  e = scratch_for_device (device, &partials);
  if (e != cudaSuccess)
  {
    if (errbuf && errbuf_len > 0) snprintf (errbuf, (size_t) errbuf_len, "fdtd_sumsq_noalloc: cudaGetSymbolAddress failed: %s", cudaGetErrorString (e));
    return FDTD_ERROR;
  }

  /* cudaLaunchKernel hands back the status of THIS launch, so an older, unrelated error that the application has
   * not yet collected with cudaGetLastError is neither blamed on this call nor cleared by it, and a failure of the
   * first pass is seen before the second pass is queued. */
  cudaStream_t stream = (cudaStream_t) cuda_stream;
  int nblocks = (int) ((count + CHUNK - 1) / CHUNK);
  long long n = count;
  if (nblocks > 0)
  {
    void *args[] = { (void *) &dev_in, (void *) &n, (void *) &partials };
    const void *fn = value_type == 4 ? (const void *) sumsq_chunks<float> : (const void *) sumsq_chunks<double>;
    e = cudaLaunchKernel (fn, dim3 (nblocks), dim3 (THREADS), args, 0, stream);
    if (e != cudaSuccess)
    {
      if (errbuf && errbuf_len > 0) snprintf (errbuf, (size_t) errbuf_len, "fdtd_sumsq_noalloc: launch of the partial-sum pass failed: %s", cudaGetErrorString (e));
      return FDTD_ERROR;
    }
  }
  {
    const double *cpartials = partials;
    void *args[] = { (void *) &cpartials, (void *) &nblocks, (void *) &dev_out };
    e = cudaLaunchKernel ((const void *) sumsq_combine, dim3 (1), dim3 (THREADS), args, 0, stream);
    if (e != cudaSuccess)
    {
      if (errbuf && errbuf_len > 0) snprintf (errbuf, (size_t) errbuf_len, "fdtd_sumsq_noalloc: launch of the combining pass failed: %s", cudaGetErrorString (e));
      return FDTD_ERROR;
    }
  }
  return FDTD_OK;
}
