/*
 * Per-cell arithmetic and the TF/SF interpolation shared by the 2D TMz kernels (kernels2d.cu: the two half-step
 * kernels; kernels2d_fused.cu: the one-pass core kernel).  Reference: Source/Scheme/SchemeTMz.cpp.
 */
#ifndef FDTD_KERNELS2D_COMMON_CUH
#define FDTD_KERNELS2D_COMMON_CUH

#include "fdtd_device.cuh"

namespace fdtd
{
namespace k2d
{

#define IN_RANGE(x, lo, hi) ((x) >= (lo) && (x) <= (hi))

/* inline interpolation of SchemeTMz.cpp:232-248 (H: `- 0.5`) and :645-663 (E: no offset term) */
template <typename T, int NC>
__device__ __noinline__ void incident_tmz (const StepParams<T> &p, bool isH, int gi, int gj, T offx, T offy, T *out)
{
  T x = ((T) gi + offx) - p.zero_inc[0];
  T y = ((T) gj + offy) - p.zero_inc[1];
  double dd = (double) x * p.cphi + (double) y * p.sphi;
  if (isH) dd = dd - 0.5;
  T d = (T) dd;
  unsigned long long n = (unsigned long long) d;
  T c1 = (T) n;
  T prop2 = d - c1;
  T prop1 = (T) 1 - prop2;
  const T *F = isH ? p.hinc : p.einc;
#pragma unroll
  for (int c = 0; c < NC; ++c) out[c] = prop1 * F[n * NC + c] + prop2 * F[(n + 1) * NC + c];
}

/* v -= s * w  with w a double (cos / sin of the incident angle): `prevHy1 -= -diffX * cos (...)`,
 * SchemeTMz.cpp:275-288 */
template <typename T, int NC, bool EXACT>
__device__ __forceinline__ T sub_scaled (T v, T s, double w)
{
  if (EXACT && sizeof (T) == 4 && NC == 1) return (T) ((double) v - (double) s * w);
  if (sizeof (T) == 8) return (T) ((double) v - (double) s * w);
  return v - s * (T) w;
}

template <typename T, bool EXACT>
__device__ __forceinline__ void sweep3_coefs (T pn, T mn, T q, T mod, T &cb, T &cc)
{
  if (EXACT) { cb = (pn / mod) / q; cc = pn == mn ? cb : (mn / mod) / q; }
  else { T r = q * rcp_t<T> (mod); cb = pn * r; cc = mn * r; }
}

template <typename T, bool EXACT>
__device__ __forceinline__ T plain_coef (T m, const StepParams<T> &p)
{
  if (EXACT) return p.dt / (m * p.dx);
  return p.dt_dx * rcp_t<T> (m);
}

/* B update of a TMz H component.  `neg` selects the Hx form (0.0 - 0.0 - Ez1 + Ez2) or the Hy form
 * (Ez1 - Ez2 - 0.0 + 0.0) of Kernels.h:38-42; the double zeros promote the curl of a real float build. */
template <typename T, int NC, bool EXACT>
__device__ __forceinline__ T tmz_first (bool plain, bool hx_form, T old, T ez1, T ez2, T ca, T cb)
{
  if (EXACT && sizeof (T) == 4 && NC == 1)
  {
    /* Hx: ((0.0 - 0.0) - Ez1) + Ez2 is double throughout; Hy: Ez1 - Ez2 is a float subtraction (rounded to float)
     * before the double zeros promote it */
    const double curl = hx_form ? (0.0 - 0.0 - (double) ez1 + (double) ez2) : ((double) (ez1 - ez2) - 0.0 + 0.0);
    const double lhs = plain ? 1.0 * (double) old : (double) (ca * old);
    return (T) (lhs + (double) cb * curl);
  }
  const T curl = hx_form ? ((T) 0 - (T) 0 - ez1 + ez2) : (ez1 - ez2 - (T) 0 + (T) 0);
  if (EXACT) return (plain ? old : ca * old) + cb * curl;
  return fma_t<T> (cb, curl, plain ? old : ca * old);
}

} /* namespace k2d */
} /* namespace fdtd */

#endif /* FDTD_KERNELS2D_COMMON_CUH */
