/*
 * Device-side parameter block and small helpers shared by the kernels of fdtd3d_b200.
 *
 * Device data layout (DESIGN.md "Data layout in HBM"):
 *   - x is the fastest axis, then y, then z, so that a z-plane (the unit the slab halo exchange moves)
 *     is contiguous and a warp reads 512 contiguous bytes of one x-row with 128-bit accesses.  The
 *     reference's host order is z-fastest (Source/Grid/Grid.cpp:138); the transposition happens in the
 *     upload / download kernels, outside the time loop.
 *   - a "cell" is one FieldValue: NC = 1 (real) or 2 (interleaved re,im) scalars of T.
 *   - field arrays: (nzl + 2) planes x Ny rows x pitch cells; plane p holds local z = p - 1, so plane 0
 *     is the low halo (z = -1) and plane nzl+1 the high halo; rows are padded to 128 bytes.
 *   - material arrays (eps, mu; real T): (nzl + 3) planes x (Ny+1) rows x pitchM values, plane p holds
 *     local material z = p - 1.
 */
#ifndef FDTD_DEVICE_CUH
#define FDTD_DEVICE_CUH

#include <cuda_runtime.h>
#include <stdint.h>

namespace fdtd
{

enum { KIND_AT = 0, KIND_BTW = 1 };

/*
 * UPML coefficient tables along one axis for one sampling kind (SURVEY.md §7 "Coefficients"):
 * s = sigma[n] (AT: all material samples of the component share index n on this axis) or
 * (sigma[n] + sigma[n+1]) / 2 (BTW: the samples straddle n and n+1).  Built on the host with the
 * reference's own expressions (Scheme3D.cpp:310-311, 401-403), so they are the reference's values:
 *   ca   = (2 eps0 - s dt) / (2 eps0 + s dt)            Ca of either sweep
 *   cbd  = (2 eps0 dt / dx) / (2 eps0 + s dt)           Cb of the D/B sweep
 *   p    = 2 eps0 + s dt                                numerator of Cb / divisor of Cb, Cc (sweep 3)
 *   m    = 2 eps0 - s dt                                numerator of Cc
 *   invp = 1 / p                                        fast mode only
 */
template <typename T>
struct CoefTab
{
  const T *ca, *cbd, *p, *m, *invp;
};

template <typename T>
struct StepParams
{
  /* fields: Ex Ey Ez Hx Hy Hz Dx Dy Dz Bx By Bz (scalar pointer; cell stride NC) */
  T *f[12];
  /* where the updated Ex..Hz go: the same arrays (in place) or, when the one-pass core kernel is in use, the other
   * buffer of the E / H ping-pong pair (kernels3d_fused.cu); D / B are always updated in place */
  T *fo[6];
  /* Sub-boxes a half-step launch covers when it does not cover the whole slab (the shell around the one-pass core):
   * nreg == 0: planes [kb, ke) of all tile rows, from the launch arguments; else region r covers tile rows
   * [jt0, jt0 + njt) and planes [kb, ke) in chunks of kc planes, and owns the CTAs [cta0, cta0 of region r + 1) of the
   * one-dimensional grid. */
  int nreg;
  struct Region { int jt0, njt, kb, ke, kc, cta0; } reg[5];
  const T *eps, *mu;
  T mu_const;           /* the value of mu when the whole array holds one value (UNIMU variants of the H kernels) */
  T eps_const;          /* likewise for eps (TMz kernels) */
  CoefTab<T> tab[3][2]; /* [axis][kind] */
  T *einc, *hinc;       /* 1-D incident line, cells */
  T *face_tab;          /* TF/SF incident values of the face cells, 24 tables (kernels3d_common.cuh, tfsf_curl_tab) */
  long long face_stride; /* cells per table */
  const T *src;         /* source samples for the steps of the current run, one cell per step */
  long long step;       /* index into src of the step being computed (set per launch by the time loop) */
  long long *step_ptr;  /* when not NULL the index is read from here instead (and advanced by the H half-step kernel): every
                           step's launches then have identical parameters, so one captured step can be replayed as a
                           CUDA graph (launch-bound grids; solver.cu, run_graph) */
  long long ninc;       /* cells in einc / hinc */

  int nx, ny;           /* global x/y size */
  int nzl;              /* planes owned by this rank */
  int kz0;              /* global z of local plane 0 */
  int nzg;              /* global z size */
  int pitch, pitchM;    /* row pitch of field arrays (cells) and material arrays (values) */
  long long ps, psM;    /* plane strides (cells / values) */
  int kchunk;           /* planes marched by one CTA */

  int core_lo[3], core_hi[3]; /* sigma-free core (global cells, inclusive; empty if lo > hi), see kernels3d_tma.cu */
  int L[3], R[3];       /* TF/SF box: leftBorderTotalField, rightBorderTotalField */
  int srcpos[3];        /* point source cell (global) */

  T dt, dx, eps0, mu0;
  T dt_dx;              /* dt / dx (fast mode) */
  T inc_ce, inc_ch;     /* dt / (relV eps0 dx), dt / (relV mu0 dx): Scheme3D.cpp:39,76 */

  /* incident-wave sampling, Scheme3D.cpp:100-123: zero coordinate (FPValue), direction cosines as
   * the doubles sin/cos return, projections of YeeGridLayout.cpp:811-845 as FPValue */
  T zero_inc[3];
  double s1, c1, s2, c2;
  T proj_e[3], proj_h[3]; /* proj_h[2] carries the leading minus of getHzFromIncidentH */
  double cphi, sphi;      /* TMz: cos / sin of incidentWaveAngle */
};

/* the index of the step being computed (StepParams::step / step_ptr) */
template <typename T>
__device__ __forceinline__ long long current_step (const StepParams<T> &p)
{
  /* written as a separate global load: as `cond ? *ptr : p.step` the compiler selected between the two ADDRESSES and
   * loaded p.step through a generic pointer into the kernel's parameter space, which faults */
  long long s = p.step;
  if (p.step_ptr) asm volatile ("ld.volatile.global.s64 %0, [%1];" : "=l"(s) : "l"(p.step_ptr));
  return s;
}
/* called by one thread of the H half-step kernel: the next step's kernels start after this kernel has ended */
template <typename T>
__device__ __forceinline__ void advance_step (const StepParams<T> &p)
{
  if (p.step_ptr) *p.step_ptr += 1;
}

/* vector of W scalars moved with one 4-, 8- or 16-byte access */
template <typename T, int W>
struct alignas (sizeof (T) * W) Vec
{
  T v[W];
};

template <int BYTES> struct RawVec;
template <> struct RawVec<4> { typedef int type; };
template <> struct RawVec<8> { typedef int2 type; };
template <> struct RawVec<16> { typedef int4 type; };

/* read-only path (arrays this kernel does not write) */
template <typename T, int W>
__device__ __forceinline__ Vec<T, W> ldv (const T *p)
{
  typedef typename RawVec<sizeof (T) * W>::type R;
  Vec<T, W> r;
  *reinterpret_cast<R *> (&r) = __ldg (reinterpret_cast<const R *> (p));
  return r;
}

/* plain (coherent) load for arrays this kernel also writes */
template <typename T, int W>
__device__ __forceinline__ Vec<T, W> ldv_rw (const T *p)
{
  typedef typename RawVec<sizeof (T) * W>::type R;
  Vec<T, W> r;
  *reinterpret_cast<R *> (&r) = *reinterpret_cast<const R *> (p);
  return r;
}

template <typename T, int W>
__device__ __forceinline__ void stv (T *p, const Vec<T, W> &r)
{
  typedef typename RawVec<sizeof (T) * W>::type R;
  *reinterpret_cast<R *> (p) = *reinterpret_cast<const R *> (&r);
}

template <typename T> __device__ __forceinline__ T fma_t (T a, T b, T c);
template <> __device__ __forceinline__ float fma_t<float> (float a, float b, float c) { return __fmaf_rn (a, b, c); }
template <> __device__ __forceinline__ double fma_t<double> (double a, double b, double c) { return __fma_rn (a, b, c); }

/* Reciprocal for fast mode (exact mode divides).  double: hardware seed (rcp.approx.ftz.f64, ~20 bits) plus two
 * Newton steps -- branch-free, relative error about 1 ulp, no denormal / special-case handling (the argument is
 * eps*eps0 or mu*mu0, about 1e-11 .. 1e-5).  float: MUFU seed (~1 ulp) plus one Newton step, also branch-free --
 * the correctly rounded __frcp_rn (seed + fix-up + a special-case branch, three per cell) was 20 % of all
 * instructions of the float kernels (profiles/r1_core_c5_ncu_summary.md). */
template <typename T> __device__ __forceinline__ T rcp_t (T a);
template <> __device__ __forceinline__ float rcp_t<float> (float a)
{
  float r;
  asm ("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(a));
  const float e = __fmaf_rn (-a, r, 1.0f);
  return __fmaf_rn (r, e, r);
}
template <> __device__ __forceinline__ double rcp_t<double> (double a)
{
  double r;
  asm ("rcp.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(a));
  double e = __fma_rn (-a, r, 1.0);
  r = __fma_rn (r, e, r);
  e = __fma_rn (-a, r, 1.0);
  r = __fma_rn (r, e, r);
  return r;
}

/*
 * Projected incident field at the real coordinate of a component's grid position
 * (Scheme3D::approximateIncidentWave, Scheme3D.cpp:100-123 + YeeGridLayout.cpp:811-845).
 * gpos = global integer position; off = the component's Yee offset (YeeGridLayout.h:416-424);
 * ddiff = 0 for E, 0.5 for H; F = einc / hinc; coef = projection.  Rare path (TF/SF faces only).
 */
template <typename T, int NC>
__device__ __noinline__ void incident_sample (const StepParams<T> &p, const T *F, T ddiff, T coef,
                                              int gi, int gj, int gk, T offx, T offy, T offz, T *out)
{
  T x = ((T) gi + offx) - p.zero_inc[0];
  T y = ((T) gj + offy) - p.zero_inc[1];
  T z = ((T) gk + offz) - p.zero_inc[2];
  /* evaluated in double and rounded once, as the reference does for float builds (sin/cos are the
   * C double functions there); for T = double this is the reference's expression verbatim */
  double dd = (double) x * p.s1 * p.c2 + (double) y * p.s1 * p.s2 + (double) z * p.c1 - (double) ddiff;
  T d = (T) dd;
  unsigned long long n = (unsigned long long) d;
  T c1 = (T) n;
  T prop2 = d - c1;
  T prop1 = (T) 1 - prop2;
#pragma unroll
  for (int c = 0; c < NC; ++c)
  {
    T v = prop1 * F[n * NC + c] + prop2 * F[(n + 1) * NC + c];
    out[c] = v * coef;
  }
}

} /* namespace fdtd */

#endif /* FDTD_DEVICE_CUH */
