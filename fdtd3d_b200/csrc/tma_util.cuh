/*
 * TMA / mbarrier PTX wrappers and the host-side tensor-map encoder shared by the TMA-staged kernels
 * (kernels3d_tma.cu: the two half-step kernels; kernels3d_fused.cu: the one-pass core kernel).
 */
#ifndef FDTD_TMA_UTIL_CUH
#define FDTD_TMA_UTIL_CUH

#include <cuda.h>
#include <cuda_runtime.h>

#include <string>

#include "fdtd_device.cuh"

namespace fdtd
{
namespace tma
{

__host__ __device__ constexpr int align128 (int x) { return (x + 127) & ~127; }

/* ------------------------------------------------------------------ PTX wrappers */

__device__ __forceinline__ unsigned smem_u32 (const void *p) { return (unsigned) __cvta_generic_to_shared (p); }

__device__ __forceinline__ void mbar_init (unsigned long long *bar, int count)
{
  asm volatile ("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32 (bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx (unsigned long long *bar, unsigned bytes)
{
  asm volatile ("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32 (bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait (unsigned long long *bar, unsigned parity)
{
  unsigned ok;
  asm volatile ("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                : "=r"(ok) : "r"(smem_u32 (bar)), "r"(parity) : "memory");
  return ok != 0;
}
/* bounded wait: a mis-programmed pipeline traps (reported as a CUDA error) instead of hanging the GPU */
__device__ __forceinline__ void mbar_wait (unsigned long long *bar, unsigned parity)
{
  unsigned tries = 0;
  while (!mbar_try_wait (bar, parity))
  {
    if (++tries > (1u << 26)) __trap ();
  }
}
__device__ __forceinline__ void tma_load_3d (void *dst, const CUtensorMap *map, unsigned long long *bar, int x, int y, int z)
{
  asm volatile ("cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];"
                ::"r"(smem_u32 (dst)), "l"(map), "r"(smem_u32 (bar)), "r"(x), "r"(y), "r"(z) : "memory");
}

template <typename T, int W>
__device__ __forceinline__ Vec<T, W> lds (const unsigned char *p)
{
  typedef typename RawVec<sizeof (T) * W>::type R;
  Vec<T, W> r;
  *reinterpret_cast<R *> (&r) = *reinterpret_cast<const R *> (p);
  return r;
}

/* Shared-space loads / stores by 32-bit shared address (LDS / STS): the generic-pointer form above compiles to generic
 * LD, which costs 64-bit address arithmetic and the slower generic path.  volatile keeps them in program order with the
 * mbarrier waits and barriers around them. */
template <typename T, int W>
__device__ __forceinline__ Vec<T, W> lds_s (unsigned addr)
{
  Vec<T, W> r;
  constexpr int BYTES = (int) sizeof (T) * W;
  static_assert (BYTES == 4 || BYTES == 8 || BYTES == 16, "vector of 4, 8 or 16 bytes");
  unsigned *u = reinterpret_cast<unsigned *> (&r);
  if (BYTES == 16) asm volatile ("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(u[0]), "=r"(u[1]), "=r"(u[2]), "=r"(u[BYTES == 16 ? 3 : 0]) : "r"(addr));
  else if (BYTES == 8) asm volatile ("ld.shared.v2.b32 {%0, %1}, [%2];" : "=r"(u[0]), "=r"(u[BYTES >= 8 ? 1 : 0]) : "r"(addr));
  else asm volatile ("ld.shared.b32 %0, [%1];" : "=r"(u[0]) : "r"(addr));
  return r;
}
template <typename T, int W>
__device__ __forceinline__ void sts_s (unsigned addr, const Vec<T, W> &r)
{
  constexpr int BYTES = (int) sizeof (T) * W;
  const unsigned *u = reinterpret_cast<const unsigned *> (&r);
  if (BYTES == 16) asm volatile ("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "r"(u[0]), "r"(u[1]), "r"(u[2]), "r"(u[BYTES == 16 ? 3 : 0]) : "memory");
  else if (BYTES == 8) asm volatile ("st.shared.v2.b32 [%0], {%1, %2};" ::"r"(addr), "r"(u[0]), "r"(u[BYTES >= 8 ? 1 : 0]) : "memory");
  else asm volatile ("st.shared.b32 [%0], %1;" ::"r"(addr), "r"(u[0]) : "memory");
}

typedef CUresult (*EncodeTiledFn) (CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *,
                                   const cuuint64_t *, const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave,
                                   CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

inline EncodeTiledFn encode_fn (std::string *err)
{
  static EncodeTiledFn fn = NULL;
  static bool tried = false;
  if (!tried)
  {
    tried = true;
    void *ptr = NULL;
    cudaDriverEntryPointQueryResult q;
    cudaError_t e = cudaGetDriverEntryPoint ("cuTensorMapEncodeTiled", &ptr, cudaEnableDefault, &q);
    if (e == cudaSuccess && q == cudaDriverEntryPointSuccess) fn = (EncodeTiledFn) ptr;
  }
  if (!fn && err) *err = "cuTensorMapEncodeTiled is not available from the CUDA driver";
  return fn;
}

/* 3-D tensor of `scalar_bytes` scalars: x extent xs scalars (row pitch), ny rows, nplanes planes; box bx x by x 1 */
inline bool encode (CUtensorMap *out, void *base, int scalar_bytes, long long xs, long long ny, long long nplanes, int bx, int by,
             std::string *err)
{
  EncodeTiledFn fn = encode_fn (err);
  if (!fn) return false;
  cuuint64_t gdim[3] = { (cuuint64_t) xs, (cuuint64_t) ny, (cuuint64_t) nplanes };
  cuuint64_t gstride[2] = { (cuuint64_t) xs * scalar_bytes, (cuuint64_t) xs * ny * scalar_bytes };
  cuuint32_t box[3] = { (cuuint32_t) bx, (cuuint32_t) by, 1 };
  cuuint32_t estr[3] = { 1, 1, 1 };
  CUresult r = fn (out, scalar_bytes == 4 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32 : CU_TENSOR_MAP_DATA_TYPE_FLOAT64, 3, base, gdim,
                   gstride, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE,
                   CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS)
  {
    if (err) *err = "cuTensorMapEncodeTiled failed with CUresult " + std::to_string ((int) r);
    return false;
  }
  return true;
}


} /* namespace tma */
} /* namespace fdtd */

#endif /* FDTD_TMA_UTIL_CUH */
