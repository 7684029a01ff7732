/*
 * Per-cell arithmetic and the TF/SF rare path shared by the 3D half-step kernels
 * (kernels3d.cu: direct global-memory version; kernels3d_tma.cu: TMA-staged version).
 */
#ifndef FDTD_KERNELS3D_COMMON_CUH
#define FDTD_KERNELS3D_COMMON_CUH

#include "fdtd_device.cuh"

namespace fdtd
{
namespace k3d
{

#define IN_RANGE(x, lo, hi) ((x) >= (lo) && (x) <= (hi))

/* tuning knobs (DESIGN.md "Kernel tuning"): CTA shape, resident CTAs per SM the register allocator must
 * allow, and bytes per thread per array access */
#ifndef FDTD_TILE_X
#define FDTD_TILE_X 32
#endif
#ifndef FDTD_TILE_Y
#define FDTD_TILE_Y 8
#endif
#ifndef FDTD_MINB
#define FDTD_MINB 1
#endif
#ifndef FDTD_VECB
#define FDTD_VECB 16
#endif
constexpr int BX = FDTD_TILE_X, BY = FDTD_TILE_Y;

/* cells per thread: FDTD_VECB bytes per access, but at least one cell */
template <typename T, int NC> struct CellsPerThread
{
  static constexpr int value = FDTD_VECB / (NC * (int) sizeof (T)) > 0 ? FDTD_VECB / (NC * (int) sizeof (T)) : 1;
};


/* ------------------------------------------------------------------ per-cell arithmetic */

template <typename T, bool EXACT>
__device__ __forceinline__ void sweep3_coefs (T pn, T mn, T q, T mod, T &cb, T &cc)
{
  if (EXACT)
  {
    /* ((2 eps0 k + sigma dt) / modifier) / (2 eps0 k + sigma' dt), Scheme3D.cpp:402-403.  Outside the PML
     * sigma = 0, so both numerators are the same number 2 eps0 and so are the two quotients: computing one
     * of them is the identical arithmetic (warp-uniform branch almost everywhere). */
    cb = (pn / mod) / q;
    cc = pn == mn ? cb : (mn / mod) / q;
  }
  else
  {
    T r = q * rcp_t<T> (mod); /* q holds invp */
    cb = pn * r;
    cc = mn * r;
  }
}

/* D (or B) update followed by the E-from-D (H-from-B) update of the same cell:
 * Kernels.h:13-29 then Kernels.h:80-96 */
template <typename T, bool EXACT>
__device__ __forceinline__ void upml_cell (T &d, T &e, T curl, T ca1, T cb1, T ca3, T cb3, T cc3)
{
  T dold = d;
  T dn, en;
  if (EXACT)
  {
    dn = ca1 * dold + cb1 * curl;
    en = ca3 * e + cb3 * dn - cc3 * dold;
  }
  else
  {
    dn = fma_t<T> (cb1, curl, ca1 * dold);
    en = fma_t<T> (ca3, e, fma_t<T> (cb3, dn, -(cc3 * dold)));
  }
  d = dn;
  e = en;
}

/* non-PML update: Kernels.h:45-61, `1.0 * old + (dt / (m * dx)) * curl`, m = eps*eps0 or mu*mu0 */
template <typename T, bool EXACT>
__device__ __forceinline__ T plain_coef (T m, const StepParams<T> &p)
{
  if (EXACT) return p.dt / (m * p.dx);
  return p.dt_dx * rcp_t<T> (m);
}
template <typename T, int NC, bool EXACT>
__device__ __forceinline__ T plain_cell (T old, T curl, T cb)
{
  if (EXACT)
  {
    /* real float builds: the double literal 1.0 promotes the sum, one rounding to float at the end */
    if (sizeof (T) == 4 && NC == 1) return (T) ((double) old + (double) (cb * curl));
    return old + cb * curl;
  }
  return fma_t<T> (cb, curl, old);
}

/* ------------------------------------------------------------------ TF/SF (rare path) */

/*
 * Face membership of a cell, as integer boxes equivalent to YeeGridLayout::doNeedTFSFUpdate??Border
 * (YeeGridLayout.cpp:327-809; derivation in SURVEY.md a8 and DESIGN.md).  With a = component axis,
 * b = a+1, c = a+2 (mod 3), L/R the TF/SF box, the result has bit 0: low face normal to b, bit 1: high
 * face normal to b, bit 2: low face normal to c, bit 3: high face normal to c.
 *   E_a: own axis in [L-1, R-1]; normal: low = L-1, high = R; remaining axis in [L, R-1]
 *   H_a: own axis in [L, R-1];   normal: low = L-1, high = R-1; remaining axis in [L-1, R-1]
 */
template <typename T>
__host__ __device__ __forceinline__ int tfsf_e_faces (const StepParams<T> &p, int a, int i, int j, int k)
{
  const int pos[3] = { i, j, k };
  const int b = (a + 1) % 3, c = (a + 2) % 3;
  if (!IN_RANGE (pos[a], p.L[a] - 1, p.R[a] - 1)) return 0;
  int f = 0;
  if (IN_RANGE (pos[c], p.L[c], p.R[c] - 1)) f |= pos[b] == p.L[b] - 1 ? 1 : (pos[b] == p.R[b] ? 2 : 0);
  if (IN_RANGE (pos[b], p.L[b], p.R[b] - 1)) f |= pos[c] == p.L[c] - 1 ? 4 : (pos[c] == p.R[c] ? 8 : 0);
  return f;
}
template <typename T>
__host__ __device__ __forceinline__ int tfsf_h_faces (const StepParams<T> &p, int a, int i, int j, int k)
{
  const int pos[3] = { i, j, k };
  const int b = (a + 1) % 3, c = (a + 2) % 3;
  if (!IN_RANGE (pos[a], p.L[a], p.R[a] - 1)) return 0;
  int f = 0;
  if (IN_RANGE (pos[c], p.L[c] - 1, p.R[c] - 1)) f |= pos[b] == p.L[b] - 1 ? 1 : (pos[b] == p.R[b] - 1 ? 2 : 0);
  if (IN_RANGE (pos[b], p.L[b] - 1, p.R[b] - 1)) f |= pos[c] == p.L[c] - 1 ? 4 : (pos[c] == p.R[c] - 1 ? 8 : 0);
  return f;
}

/* Cheap necessary condition for a cell to lie on any TF/SF face of the E (or H) updates: on every face at
 * least one coordinate equals L-1 or R (E) / L-1 or R-1 (H) along the face normal.  The j and k parts are
 * hoisted by the callers (per thread / per plane); only the i part is evaluated per cell. */
template <typename T>
__device__ __forceinline__ bool tfsf_axis_hit (const StepParams<T> &p, int axis, int x, bool e_phase)
{
  return x == p.L[axis] - 1 || x == (e_phase ? p.R[axis] : p.R[axis] - 1);
}

/*
 * Curl of a face cell with the TF/SF correction applied to the operand that lies across the interface
 * (Scheme3D::calculate{Ex..Hz}TFSF, Scheme3D.cpp:138-208, 435-505, 732-802, 1029-1099, 1325-1395,
 * 1621-1691), scalar `comp` (0 = re, 1 = im) of the cell.
 *
 * E_a (IS_E):  curl = P1 - P2 - Q1 + Q2 with P = H_c differenced along b (P1 at pos, P2 at pos - e_b),
 *              Q = H_b differenced along c; low face of b: P1 -= inc, high: P2 -= inc; low face of c:
 *              Q1 -= inc, high: Q2 -= inc; inc = projected incident H at the operand's own position.
 * H_a (!IS_E): P = E_b differenced along c (P1 at pos + e_c, P2 at pos), Q = E_c differenced along b
 *              (Q1 at pos + e_b, Q2 at pos); low face of b: Q2 += inc, high: Q1 += inc; low face of c:
 *              P2 += inc, high: P1 += inc.
 * Yee offsets (YeeGridLayout.h:416-424): E_f is at +1 along f and +0.5 elsewhere, H_f at +0.5 along f
 * and +1 elsewhere.
 */
template <typename T, int NC, bool IS_E>
__device__ __noinline__ T tfsf_curl (const StepParams<T> &p, int a, int faces, int comp, int i, int j, int k,
                                     T P1, T P2, T Q1, T Q2)
{
  const int b = (a + 1) % 3, c = (a + 2) % 3;
  T inc[NC];
#pragma unroll 1
  for (int pass = 0; pass < 2; ++pass)
  {
    const int bits = pass == 0 ? (faces & 3) : ((faces >> 2) & 3);
    if (!bits) continue;
    const bool low = (bits & 1) != 0;
    const int n = pass == 0 ? b : c;  /* face normal */
    const int fc = pass == 0 ? c : b; /* component of the other family that is corrected */
    int pos[3] = { i, j, k };
    if (!low) pos[n] += IS_E ? -1 : 1;
    T off[3];
    for (int ax = 0; ax < 3; ++ax) off[ax] = IS_E ? (ax == fc ? (T) 0.5 : (T) 1) : (ax == fc ? (T) 1 : (T) 0.5);
    incident_sample<T, NC> (p, IS_E ? p.hinc : p.einc, IS_E ? (T) 0.5 : (T) 0, IS_E ? p.proj_h[fc] : p.proj_e[fc],
                            pos[0], pos[1], pos[2], off[0], off[1], off[2], inc);
    const T d = inc[comp];
    if (IS_E)
    {
      if (pass == 0) { if (low) P1 -= d; else P2 -= d; }
      else { if (low) Q1 -= d; else Q2 -= d; }
    }
    else
    {
      if (pass == 0) { if (low) Q2 += d; else Q1 += d; }
      else { if (low) P2 += d; else P1 += d; }
    }
  }
  return P1 - P2 - Q1 + Q2;
}

/*
 * The same correction with the incident values read from the per-step face tables (aux_kernels.cu,
 * face_tables_kernel): table ((a * 2 + pass) * 2 + side) of the phase, `pass` 0 = faces normal to b, 1 = normal to c,
 * side 0 = low, 1 = high, holds out[comp] of the incident_sample call tfsf_curl makes for that face cell, indexed by
 * the two in-face coordinates (normal z: j * nx + i; normal y: kl * nx + i; normal x: kl * ny + j; kl = local plane).
 * Same operations in the same order as tfsf_curl, so the results are the same bits -- but a face cell costs a load
 * and two additions instead of ~300 instructions, which matters because every warp that crosses an x face has a
 * lane on it in every plane (34 % of all instructions of the 512^3 run were spent there, profiles/r1_core_*).
 */
template <typename T, int NC, bool IS_E>
__device__ __forceinline__ T tfsf_curl_tab (const StepParams<T> &p, int a, int faces, int comp, int i, int j, int kl,
                                            T P1, T P2, T Q1, T Q2)
{
  const int b = (a + 1) % 3, c = (a + 2) % 3;
#pragma unroll
  for (int pass = 0; pass < 2; ++pass)
  {
    const int bits = pass == 0 ? (faces & 3) : ((faces >> 2) & 3);
    if (!bits) continue;
    const bool low = (bits & 1) != 0;
    const int n = pass == 0 ? b : c;
    const long long idx = n == 2 ? (long long) j * p.nx + i : (n == 1 ? (long long) kl * p.nx + i : (long long) kl * p.ny + j);
    const int t = (IS_E ? 0 : 12) + (a * 2 + pass) * 2 + (low ? 0 : 1);
    const T *src = p.face_tab + (t * p.face_stride + idx) * NC + comp;
    const T d = __ldg (src);
    /* the same lane needs the next plane's entry one plane later: have it in L1 by then */
    if (n != 2 && kl + 1 < p.nzl)
      asm volatile ("prefetch.global.L1 [%0];" ::"l"(src + (long long) (n == 1 ? p.nx : p.ny) * NC));
    if (IS_E)
    {
      if (pass == 0) { if (low) P1 -= d; else P2 -= d; }
      else { if (low) Q1 -= d; else Q2 -= d; }
    }
    else
    {
      if (pass == 0) { if (low) Q2 += d; else Q1 += d; }
      else { if (low) P2 += d; else P1 += d; }
    }
  }
  return P1 - P2 - Q1 + Q2;
}

} /* namespace k3d */
} /* namespace fdtd */

#endif /* FDTD_KERNELS3D_COMMON_CUH */
