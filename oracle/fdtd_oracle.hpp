/*
 * TEST INFRASTRUCTURE ONLY -- Tier-2 oracle.  Nothing under fdtd3d_b200/ may include, link or call
 * this file; only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
 * legs use it, and only as the checker / the timed CPU baseline.
 *
 * A flat-array CPU restatement of the reference's Yee leapfrog step, written from the reference's
 * algorithm (file:line cited at each function, paths relative to /root/reference/Source), with the
 * same expressions, operand order, literal types and FPValue/FieldValue promotions so that it is
 * bit-for-bit equal to the reference CPU solver.  Parity is PINNED: tests/test_oracle_golden.py
 * compares it with fixtures dumped from the unmodified reference solver (oracle/_ref, built by
 * oracle/build_ref.sh; fixtures + generating script in tests/golden/), and
 * tests/test_oracle_vs_ref.py compares it live against oracle/_ref when that binary is present.
 *
 * Differences from the reference, none of which change a rounded result:
 *   - storage is one flat array per field (cur only).  The reference's cur/prev pair collapses
 *     because every sweep reads only its own previous value and *other* fields;
 *     the E-from-D update (Kernels.h:80-96) is done right after the cell's D update with D_prev
 *     held in a local, which is the same arithmetic because that sweep has no neighbour reads;
 *   - the sigma grids are stored as their 1-D profiles (sigma_x depends on i only,
 *     Scheme3D.cpp:3667-3716) but are *sampled* through the reference's own 2-/4-point averaging
 *     expressions (YeeGridLayout.h:1005-1263, Approximation.cpp:15-26);
 *   - the six per-component functions are one generic function over the axis permutation
 *     (a, b, c) = (component axis, next, next-next);
 *   - sweeps are OpenMP-parallel over i (each sweep is data-parallel, so the result is identical
 *     for any thread count).  The reference itself is single-threaded (no OpenMP anywhere).
 *
 * Math-library calls: in the reference's translation units an unqualified sin/cos/pow/log/sqrt
 * on a float resolves to the C `double` function (only <cmath> is included, so the global
 * namespace holds just ::sin(double); probed with g++ 13.3), i.e. float builds evaluate every
 * transcendental in double and round once on assignment.  The d*() helpers below do that.
 */
#ifndef FDTD_ORACLE_HPP
#define FDTD_ORACLE_HPP

#include <cmath>
#include <complex>
#include <cstdint>
#include <cstring>
#include <vector>

namespace oracle
{

static inline double dsin (double x) { return ::sin (x); }
static inline double dcos (double x) { return ::cos (x); }
static inline double dasin (double x) { return ::asin (x); }
static inline double dsqrt (double x) { return ::sqrt (x); }
static inline double dpow (double x, double y) { return ::pow (x, y); }
static inline double dlog (double x) { return ::log (x); }

/* Physics/PhysicsConst.h:9-13 -- constants are *stored as FPValue*. */
template <class FP>
struct PhysConst
{
  static FP SpeedOfLight () { return (FP) 2.99792458e+8; }
  static FP Eps0 () { return (FP) 0.0000000000088541878176203892; }
  static FP Mu0 () { return (FP) 0.0000012566370614359173; }
  static FP Pi () { return (FP) M_PI; }
};

#define ORACLE_SQR(x) ((x) * (x))

/* Layout/Approximation.cpp:15-26 */
template <class FP>
static inline FP approximateMaterial (FP v1, FP v2) { return (v1 + v2) / 2.0; }
template <class FP>
static inline FP approximateMaterial (FP v1, FP v2, FP v3, FP v4)
{
  return (approximateMaterial (v1, v2) + approximateMaterial (v3, v4)) / 2.0;
}

/* Layout/Approximation.cpp:229-286, including the `=`-for-`==` quirk at :237 and :250. */
template <class FP>
static FP phaseVelocityIncidentWave3D (FP delta, FP freeSpaceWaveLentgh, FP courantNum, FP N_lambda,
                                       FP incidentWaveAngle1, FP incidentWaveAngle2)
{
  typedef PhysConst<FP> PC;
  if ((incidentWaveAngle1 = (PC::Pi () / 2
      && (incidentWaveAngle2 == 0
          || incidentWaveAngle2 == PC::Pi () / 2
          || incidentWaveAngle2 == PC::Pi ()
          || incidentWaveAngle2 == 3 * PC::Pi () / 2))))
  {
    return PC::SpeedOfLight () * PC::Pi () /
           (N_lambda * dasin (dsin (PC::Pi () * courantNum / N_lambda) / courantNum));
  }
  if ((incidentWaveAngle1 = (PC::Pi () / 2
      && (incidentWaveAngle2 == PC::Pi () / 4
          || incidentWaveAngle2 == 3 * PC::Pi () / 4
          || incidentWaveAngle2 == 5 * PC::Pi () / 4
          || incidentWaveAngle2 == 7 * PC::Pi () / 4))))
  {
    return PC::SpeedOfLight () * PC::Pi () /
           (N_lambda * dsqrt (2.0) * dasin (dsin (PC::Pi () * courantNum / N_lambda) / (courantNum * dsqrt (2.0))));
  }

  FP accuracy = 0.0000001; /* Approximation::accuracy, Approximation.cpp:7 */
  FP k = 2 * PC::Pi ();
  FP k_prev = k + accuracy;
  FP normalized_delta = delta / freeSpaceWaveLentgh;
  FP A = normalized_delta * dsin (incidentWaveAngle1) * dcos (incidentWaveAngle2) / 2;
  FP B = normalized_delta * dsin (incidentWaveAngle1) * dsin (incidentWaveAngle2) / 2;
  FP C = normalized_delta * dcos (incidentWaveAngle1) / 2;
  FP D = ORACLE_SQR (dsin (PC::Pi () * courantNum / N_lambda)) / ORACLE_SQR (courantNum);
  while (ORACLE_SQR (k_prev - k) >= accuracy)
  {
    k_prev = k;
    FP diff1 = ORACLE_SQR (dsin (A * k)) + ORACLE_SQR (dsin (B * k)) + ORACLE_SQR (dsin (C * k)) - D;
    FP diff2 = A * dsin (2 * A * k) + B * dsin (2 * B * k) + C * dsin (2 * C * k);
    FP diff = diff1 / diff2;
    k -= diff;
  }
  return PC::SpeedOfLight () * 2 * PC::Pi () / k;
}

/* Value-type helpers: complex arithmetic with *real* coefficients is component-wise in
 * std::complex (operator*(T, complex<T>)), which is what the reference relies on. */
template <class FP, bool CPLX> struct FVT;
template <class FP> struct FVT<FP, false>
{
  typedef FP type;
  static type make (FP re, FP) { return re; }
  /* `1.0 * old` of Kernels.h:45-61: the double literal promotes a float field value */
  static double one_times (FP old) { return 1.0 * old; }
  /* FieldValue * double (TMz TFSF, SchemeTMz.cpp:275-288) */
  static double times_d (FP v, double s) { return v * s; }
  static double zero () { return 0.0; }
};
template <> struct FVT<double, true>
{
  typedef std::complex<double> type;
  static type make (double re, double im) { return type (re, im); }
  static type one_times (type old) { return 1.0 * old; }
  static type times_d (type v, double s) { return v * s; }
  static double zero () { return 0.0; }
};
/* complex float: the reference does not compile (Kernels.h mixes `double * std::complex<float>`),
 * so there is nothing to be faithful to; arithmetic stays in float.  "parity unpinned" for this
 * one instantiation. */
template <> struct FVT<float, true>
{
  typedef std::complex<float> type;
  static type make (float re, float im) { return type (re, im); }
  static type one_times (type old) { return old; }
  static type times_d (type v, double s) { return v * (float) s; }
  static float zero () { return 0.0f; }
};

struct Params
{
  int dims;                 /* 3 or 2 (TMz) */
  bool usePML, useTFSF;
  int64_t N[3];             /* grid size; N[2] = 1 for 2D */
  int64_t L[3];             /* leftBorderTotalField */
  int64_t src[3];           /* point-source cell */
  double dx, dt, freq, relv;
  double a1, a2, a3;        /* incident angles teta, phi, psi (radians) */
  int64_t t0;
};

template <class FP, bool CPLX>
class Scheme
{
public:
  typedef typename FVT<FP, CPLX>::type FV;
  typedef PhysConst<FP> PC;

  Params p;
  int64_t Nx, Ny, Nz, Mx, My, Mz; /* field and material grid sizes */
  FP gridStep, gridTimeStep, sourceFrequency, relPhaseVelocity;
  FP angle1, angle2, angle3;
  FP zeroInc[3];            /* YeeGridLayout::zeroIncCoordFP, YeeGridLayout.h:438-440 */

  /* E[0..2] = Ex,Ey,Ez ; H ; D ; B in reference linear order */
  std::vector<FV> E[3], H[3], D[3], B[3];
  std::vector<FP> Eps, Mu;  /* (N+1)^d, real part only (Approximation.cpp:219-227) */
  std::vector<FP> Sigma[3]; /* 1-D profiles, N[axis]+1 */
  std::vector<FV> EInc, HInc;
  int64_t t;

  explicit Scheme (const Params &prm) : p (prm)
  {
    Nx = p.N[0]; Ny = p.N[1]; Nz = p.dims == 3 ? p.N[2] : 1;
    Mx = Nx + 1; My = Ny + 1; Mz = p.dims == 3 ? Nz + 1 : 1;
    gridStep = (FP) p.dx; gridTimeStep = (FP) p.dt; sourceFrequency = (FP) p.freq;
    relPhaseVelocity = (FP) p.relv;
    angle1 = (FP) p.a1; angle2 = (FP) p.a2; angle3 = (FP) p.a3;
    /* YeeGridLayout.h:438-440: grid_coord - double*double*double, then stored as FPValue */
    zeroInc[0] = p.L[0] - 2.5 * dsin (angle1) * dcos (angle2);
    zeroInc[1] = p.L[1] - 2.5 * dsin (angle1) * dsin (angle2);
    zeroInc[2] = p.L[2] - 2.5 * dcos (angle1);
    size_t n = (size_t) Nx * Ny * Nz;
    for (int a = 0; a < 3; ++a)
    {
      E[a].assign (n, FV ()); H[a].assign (n, FV ());
      if (p.usePML) { D[a].assign (n, FV ()); B[a].assign (n, FV ()); }
      Sigma[a].assign ((size_t) p.N[a] + 1, (FP) 0);
    }
    Eps.assign ((size_t) Mx * My * Mz, (FP) 1);
    Mu.assign ((size_t) Mx * My * Mz, (FP) 1);
    if (p.useTFSF)
    {
      /* Scheme3D.h:278-279, SchemeTMz.h:234-235 */
      size_t ni = (size_t) 100 * (Nx + Ny + (p.dims == 3 ? Nz : 0));
      EInc.assign (ni, FV ()); HInc.assign (ni, FV ());
    }
    t = p.t0;
  }

  /* Grid.cpp:138 / Grid.cpp:82 */
  inline size_t idx (int64_t i, int64_t j, int64_t k) const { return ((size_t) i * Ny + j) * Nz + k; }
  inline size_t midx (int64_t i, int64_t j, int64_t k) const { return ((size_t) i * My + j) * Mz + k; }
  inline size_t idxv (const int64_t c[3]) const { return idx (c[0], c[1], c[2]); }

  /* value of the 3-D sigma_axis grid at material cell m[] (Scheme3D.cpp:3667-3818: depends on m[axis] only) */
  inline FP sigmaAt (int axis, const int64_t m[3]) const { return Sigma[axis][m[axis]]; }
  inline FP matAt (const std::vector<FP> &M, const int64_t m[3]) const { return M[midx (m[0], m[1], m[2])]; }

  /* YeeGridLayout::getMaterial, E/D-type (YeeGridLayout.h:1024-1110, 1246): two samples along the
   * component axis a at material cells c and c + e_a. */
  template <class F> inline FP materialE (int a, const int64_t c[3], F sample) const
  {
    int64_t c2[3] = { c[0], c[1], c[2] };
    c2[a] += 1;
    return approximateMaterial<FP> (sample (c), sample (c2));
  }
  /* H/B-type (YeeGridLayout.h:1111-1197, 1255): four samples over the two transverse axes u < v:
   * 11 = c, 12 = c + e_v, 21 = c + e_u, 22 = c + e_u + e_v. */
  template <class F> inline FP materialH (int a, const int64_t c[3], F sample) const
  {
    int u = (a == 0) ? 1 : 0;
    int v = (a == 2) ? 1 : 2;
    int64_t c12[3] = { c[0], c[1], c[2] }, c21[3] = { c[0], c[1], c[2] }, c22[3] = { c[0], c[1], c[2] };
    c12[v] += 1; c21[u] += 1; c22[u] += 1; c22[v] += 1;
    return approximateMaterial<FP> (sample (c), sample (c12), sample (c21), sample (c22));
  }

  /* Scheme3D::approximateIncidentWave, Scheme3D.cpp:100-123 (real coordinate r[], dDiff 0.0 / 0.5) */
  FV approximateIncidentWave (const FP r[3], FP dDiff, const std::vector<FV> &FieldInc) const
  {
    FP x = r[0] - zeroInc[0];
    FP y = r[1] - zeroInc[1];
    FP z = r[2] - zeroInc[2];
    FP d = x * dsin (angle1) * dcos (angle2)
           + y * dsin (angle1) * dsin (angle2)
           + z * dcos (angle1) - dDiff;
    FP coordD1 = (FP) ((uint64_t) d);
    FP coordD2 = coordD1 + 1;
    FP proportionD2 = d - coordD1;
    FP proportionD1 = 1 - proportionD2;
    const FV &val1 = FieldInc[(size_t) coordD1];
    const FV &val2 = FieldInc[(size_t) coordD2];
    return proportionD1 * val1 + proportionD2 * val2;
  }

  /* YeeGridLayout::get{Ex,Ey,Ez}FromIncidentE / get{Hx,Hy,Hz}FromIncidentH, YeeGridLayout.cpp:811-845 */
  FV fromIncidentE (int a, FV valE) const
  {
    switch (a)
    {
      case 0: return valE * (FP) (dcos (angle3) * dsin (angle2) - dsin (angle3) * dcos (angle1) * dcos (angle2));
      case 1: return valE * (FP) (-dcos (angle3) * dcos (angle2) - dsin (angle3) * dcos (angle1) * dsin (angle2));
      default: return valE * (FP) (dsin (angle3) * dsin (angle1));
    }
  }
  FV fromIncidentH (int a, FV valH) const
  {
    switch (a)
    {
      case 0: return valH * (FP) (dsin (angle3) * dsin (angle2) + dcos (angle3) * dcos (angle1) * dcos (angle2));
      case 1: return valH * (FP) (-dsin (angle3) * dcos (angle2) + dcos (angle3) * dcos (angle1) * dsin (angle2));
      default: return -valH * (FP) (dcos (angle3) * dsin (angle1));
    }
  }

  /*
   * YeeGridLayout::doNeedTFSFUpdate{Ex..Hz}Border, YeeGridLayout.cpp:327-809, as FP comparisons on
   * the component's real coordinate.  All 24 cases follow one pattern (a = component axis,
   * n = face normal, t = remaining axis):
   *   E_a:  own axis in (L-0.5, R+0.5); normal: low (L-1, L), high (R, R+1); t in (L, R)
   *   H_a:  own axis in (L, R); normal: low (L-0.5, L+0.5), high (R-0.5, R+0.5); t in (L-0.5, R+0.5)
   * In 2D mode (is3Dim false) the z condition is skipped (e.g. YeeGridLayout.cpp:348-349).
   */
  bool needTFSF (bool isE, int a, const int64_t pos[3], int n, bool high) const
  {
    bool is3 = p.dims == 3;
    for (int ax = 0; ax < 3; ++ax)
    {
      if (ax == 2 && !is3) continue;
      FP off = isE ? (ax == a ? (FP) 1.0 : (FP) 0.5) : (ax == a ? (FP) 0.5 : (FP) 1.0);
      FP c = (FP) pos[ax] + off;
      FP Lf = (FP) 0.0 + (FP) p.L[ax];
      FP Rf = (FP) 0.0 + (FP) (p.N[ax] - p.L[ax]);
      bool ok;
      if (ax == n)
      {
        if (isE) ok = high ? (c > Rf && c < Rf + 1) : (c > Lf - 1 && c < Lf);
        else ok = high ? (c > Rf - 0.5 && c < Rf + 0.5) : (c > Lf - 0.5 && c < Lf + 0.5);
      }
      else if (ax == a)
      {
        if (isE) ok = c > Lf - 0.5 && c < Rf + 0.5;
        else ok = c > Lf && c < Rf;
      }
      else
      {
        if (isE) ok = c > Lf && c < Rf;
        else ok = c > Lf - 0.5 && c < Rf + 0.5;
      }
      if (!ok) return false;
    }
    return true;
  }

  /* real coordinate of component `comp` of an E (isE) or H grid at integer position c
   * (YeeGridLayout.h:416-424 offsets, :275-298 conversions) */
  void realCoord (bool isE, int comp, const int64_t c[3], FP r[3]) const
  {
    for (int ax = 0; ax < 3; ++ax)
    {
      FP off = isE ? (ax == comp ? (FP) 1.0 : (FP) 0.5) : (ax == comp ? (FP) 0.5 : (FP) 1.0);
      r[ax] = (FP) c[ax] + off;
    }
  }

  /* ---------------------------------------------------------------- 1-D incident line */

  /* Scheme3D::performPlaneWaveESteps, Scheme3D.cpp:24-59 / SchemeTMz.cpp:20-67 */
  void performPlaneWaveESteps (int tt)
  {
    FP rel = relPhaseVelocity; /* 2D: relPhi, SchemeTMz.cpp:35-49, passed in by the caller */
    size_t n = EInc.size ();
    for (size_t i = 1; i < n; ++i)
    {
      EInc[i] = EInc[i] + (gridTimeStep / (rel * PC::Eps0 () * gridStep)) * (HInc[i - 1] - HInc[i]);
    }
    EInc[0] = sourceValue (tt);
  }
  /* Scheme3D::performPlaneWaveHSteps, Scheme3D.cpp:61-82 / SchemeTMz.cpp:69-106 */
  void performPlaneWaveHSteps (int)
  {
    FP rel = relPhaseVelocity;
    size_t n = HInc.size ();
    for (size_t i = 0; i + 1 < n; ++i)
    {
      HInc[i] = HInc[i] + (gridTimeStep / (rel * PC::Mu0 () * gridStep)) * (EInc[i] - EInc[i + 1]);
    }
  }
  /* Scheme3D.cpp:47-52, 2016-2021: sin (gridTimeStep * t * 2 * Pi * sourceFrequency), t an int */
  FV sourceValue (int tt) const
  {
    FP arg = gridTimeStep * tt * 2 * PC::Pi () * sourceFrequency;
    if (CPLX) return FVT<FP, CPLX>::make ((FP) dsin (arg), (FP) dcos (arg));
    return FVT<FP, CPLX>::make ((FP) dsin (arg), 0);
  }

  /* ---------------------------------------------------------------- 3D E phase */

  /*
   * calculate{Ex,Ey,Ez}Step / ...StepPML, Scheme3D.cpp:211-416, 508-713, 805-1010, generic in
   * (a,b,c) = (0,1,2), (1,2,0), (2,0,1).  Stencil (YeeGridLayout.cpp:3-129):
   *   P = H_c differenced along b:  P1 = H_c[pos], P2 = H_c[pos - e_b]
   *   Q = H_b differenced along c:  Q1 = H_b[pos], Q2 = H_b[pos - e_c]
   *   val = Ca*old + Cb*(P1 - P2 - Q1 + Q2)                         (Kernels.h:13-20)
   * TFSF (calculateE?TFSF, Scheme3D.cpp:138-208 etc.): a low face normal to b corrects P1, a high
   * one P2; faces normal to c correct Q1 / Q2; all with `-=` of the projected incident H sampled at
   * the corrected operand's own real coordinate.
   * Compute range: start e_b + e_c, end N - e_a (YeeGridLayout.h:131-170).
   */
  void stepE (int a)
  {
    const int b = (a + 1) % 3, c = (a + 2) % 3;
    const FP eps0 = PC::Eps0 ();
    int64_t lo[3] = { 0, 0, 0 }, hi[3] = { Nx, Ny, Nz };
    lo[b] = 1; lo[c] = 1; hi[a] -= 1;
    if (p.dims == 2) { lo[2] = 0; hi[2] = 1; }
    std::vector<FV> &Dst = p.usePML ? D[a] : E[a];

    /* The reference runs the D sweep over the whole range and then the E-from-D sweep over the same
     * range (Scheme3D.cpp:270-324, 366-415).  The second sweep reads only the cell's own D_cur,
     * D_prev and E_prev, so doing it right after the cell's D update is the same arithmetic. */
#pragma omp parallel for schedule(static)
    for (int64_t i = lo[0]; i < hi[0]; ++i)
      for (int64_t j = lo[1]; j < hi[1]; ++j)
        for (int64_t k = lo[2]; k < hi[2]; ++k)
        {
          const int64_t pos[3] = { i, j, k };
          int64_t pb[3] = { i, j, k }, pc[3] = { i, j, k };
          pb[b] -= 1; pc[c] -= 1;
          FV P1 = H[c][idxv (pos)], P2 = H[c][idxv (pb)];
          FV Q1 = H[b][idxv (pos)], Q2 = H[b][idxv (pc)];
          if (p.useTFSF) tfsfE (a, b, c, pos, pb, pc, P1, P2, Q1, Q2);

          const size_t id = idxv (pos);
          FV old = Dst[id];
          if (p.usePML)
          {
            /* Scheme3D.cpp:281,310-319: sigma_b sampled E-type at the D_a position */
            FP sigma = materialE (a, pos, [&] (const int64_t *m) { return sigmaAt (b, m); });
            FP kk = 1;
            FP Ca = (2 * eps0 * kk - sigma * gridTimeStep) / (2 * eps0 * kk + sigma * gridTimeStep);
            FP Cb = (2 * eps0 * gridTimeStep / gridStep) / (2 * eps0 * kk + sigma * gridTimeStep);
            FV val = Ca * old + Cb * (P1 - P2 - Q1 + Q2);
            Dst[id] = val;

            /* third sweep, Scheme3D.cpp:366-415: E_a from D_a; Ca and the divisor from sigma_c,
             * numerators from sigma_a, modifier = eps*eps0 */
            FP eps = materialE (a, pos, [&] (const int64_t *m) { return matAt (Eps, m); });
            FP sigmaN = materialE (a, pos, [&] (const int64_t *m) { return sigmaAt (a, m); });
            FP sigmaD = materialE (a, pos, [&] (const int64_t *m) { return sigmaAt (c, m); });
            FP modifier = eps * eps0;
            FP k_n = 1, k_d = 1;
            FP Ca3 = (2 * eps0 * k_d - sigmaD * gridTimeStep) / (2 * eps0 * k_d + sigmaD * gridTimeStep);
            FP Cb3 = ((2 * eps0 * k_n + sigmaN * gridTimeStep) / (modifier)) / (2 * eps0 * k_d + sigmaD * gridTimeStep);
            FP Cc3 = ((2 * eps0 * k_n - sigmaN * gridTimeStep) / (modifier)) / (2 * eps0 * k_d + sigmaD * gridTimeStep);
            E[a][id] = Ca3 * E[a][id] + Cb3 * val - Cc3 * old;
          }
          else
          {
            /* Scheme3D.cpp:226,250-257 + Kernels.h:45-52: Ca = 1.0 (double literal), Cb = dt / ((eps*eps0) * dx) */
            FP eps = materialE (a, pos, [&] (const int64_t *m) { return matAt (Eps, m); });
            Dst[id] = FVT<FP, CPLX>::one_times (old) + ((gridTimeStep) / ((eps * eps0) * (gridStep))) * (P1 - P2 - Q1 + Q2);
          }
        }
  }

  void tfsfE (int a, int b, int c, const int64_t pos[3], const int64_t pb[3], const int64_t pc[3],
              FV &P1, FV &P2, FV &Q1, FV &Q2) const
  {
    bool lowB = needTFSF (true, a, pos, b, false), highB = needTFSF (true, a, pos, b, true);
    bool lowC = needTFSF (true, a, pos, c, false), highC = needTFSF (true, a, pos, c, true);
    FV diffB = FV (), diffC = FV ();
    if (lowB || highB)
    {
      /* low face -> operand 1 (H_c[pos]); else high face -> operand 2 (H_c[pos - e_b]) */
      FP r[3]; realCoord (false, c, lowB ? pos : pb, r);
      diffB = fromIncidentH (c, approximateIncidentWave (r, (FP) 0.5, HInc));
    }
    if (lowC || highC)
    {
      FP r[3]; realCoord (false, b, lowC ? pos : pc, r);
      diffC = fromIncidentH (b, approximateIncidentWave (r, (FP) 0.5, HInc));
    }
    if (lowB) P1 -= diffB; else if (highB) P2 -= diffB;
    if (lowC) Q1 -= diffC; else if (highC) Q2 -= diffC;
  }

  /* ---------------------------------------------------------------- 3D H phase */

  /*
   * calculate{Hx,Hy,Hz}Step / ...StepPML, Scheme3D.cpp:1102-1306, 1398-1602, 1693-1897.
   *   P = E_b differenced along c:  P1 = E_b[pos + e_c], P2 = E_b[pos]
   *   Q = E_c differenced along b:  Q1 = E_c[pos + e_b], Q2 = E_c[pos]
   *   val = Da*old + Db*(P1 - P2 - Q1 + Q2)                         (Kernels.h:22-29)
   * TFSF (calculateH?TFSF, Scheme3D.cpp:1029-1099 etc.): a low face normal to b corrects Q2, a high
   * one Q1; faces normal to c correct P2 / P1; all with `+=` of the projected incident E.
   * Compute range: start e_a, end N - e_b - e_c (YeeGridLayout.h:143-182).  eps0 is used in the B
   * equations too (Scheme3D.cpp:1200-1201).
   */
  void stepH (int a)
  {
    const int b = (a + 1) % 3, c = (a + 2) % 3;
    const FP eps0 = PC::Eps0 ();
    const FP mu0 = PC::Mu0 ();
    int64_t lo[3] = { 0, 0, 0 }, hi[3] = { Nx, Ny, Nz };
    lo[a] = 1; hi[b] -= 1; hi[c] -= 1;
    std::vector<FV> &Dst = p.usePML ? B[a] : H[a];

#pragma omp parallel for schedule(static)
    for (int64_t i = lo[0]; i < hi[0]; ++i)
      for (int64_t j = lo[1]; j < hi[1]; ++j)
        for (int64_t k = lo[2]; k < hi[2]; ++k)
        {
          const int64_t pos[3] = { i, j, k };
          int64_t pb[3] = { i, j, k }, pc[3] = { i, j, k };
          pb[b] += 1; pc[c] += 1;
          FV P1 = E[b][idxv (pc)], P2 = E[b][idxv (pos)];
          FV Q1 = E[c][idxv (pb)], Q2 = E[c][idxv (pos)];
          if (p.useTFSF) tfsfH (a, b, c, pos, pb, pc, P1, P2, Q1, Q2);

          const size_t id = idxv (pos);
          FV old = Dst[id];
          if (p.usePML)
          {
            FP sigma = materialH (a, pos, [&] (const int64_t *m) { return sigmaAt (b, m); });
            FP kk = 1;
            FP Ca = (2 * eps0 * kk - sigma * gridTimeStep) / (2 * eps0 * kk + sigma * gridTimeStep);
            FP Cb = (2 * eps0 * gridTimeStep / gridStep) / (2 * eps0 * kk + sigma * gridTimeStep);
            FV val = Ca * old + Cb * (P1 - P2 - Q1 + Q2);
            Dst[id] = val;

            /* Scheme3D.cpp:1256-1305 (same-cell dependency only, see stepE) */
            FP mu = materialH (a, pos, [&] (const int64_t *m) { return matAt (Mu, m); });
            FP sigmaN = materialH (a, pos, [&] (const int64_t *m) { return sigmaAt (a, m); });
            FP sigmaD = materialH (a, pos, [&] (const int64_t *m) { return sigmaAt (c, m); });
            FP modifier = mu * mu0;
            FP k_n = 1, k_d = 1;
            FP Ca3 = (2 * eps0 * k_d - sigmaD * gridTimeStep) / (2 * eps0 * k_d + sigmaD * gridTimeStep);
            FP Cb3 = ((2 * eps0 * k_n + sigmaN * gridTimeStep) / (modifier)) / (2 * eps0 * k_d + sigmaD * gridTimeStep);
            FP Cc3 = ((2 * eps0 * k_n - sigmaN * gridTimeStep) / (modifier)) / (2 * eps0 * k_d + sigmaD * gridTimeStep);
            H[a][id] = Ca3 * H[a][id] + Cb3 * val - Cc3 * old;
          }
          else
          {
            FP mu = materialH (a, pos, [&] (const int64_t *m) { return matAt (Mu, m); });
            Dst[id] = FVT<FP, CPLX>::one_times (old) + ((gridTimeStep) / ((mu * mu0) * (gridStep))) * (P1 - P2 - Q1 + Q2);
          }
        }
  }

  void tfsfH (int a, int b, int c, const int64_t pos[3], const int64_t pb[3], const int64_t pc[3],
              FV &P1, FV &P2, FV &Q1, FV &Q2) const
  {
    bool lowB = needTFSF (false, a, pos, b, false), highB = needTFSF (false, a, pos, b, true);
    bool lowC = needTFSF (false, a, pos, c, false), highC = needTFSF (false, a, pos, c, true);
    FV diffB = FV (), diffC = FV ();
    if (lowB || highB)
    {
      /* low face -> operand 2 (E_c[pos]); else high face -> operand 1 (E_c[pos + e_b]) */
      FP r[3]; realCoord (true, c, lowB ? pos : pb, r);
      diffB = fromIncidentE (c, approximateIncidentWave (r, (FP) 0.0, EInc));
    }
    if (lowC || highC)
    {
      FP r[3]; realCoord (true, b, lowC ? pos : pc, r);
      diffC = fromIncidentE (b, approximateIncidentWave (r, (FP) 0.0, EInc));
    }
    if (lowB) Q2 += diffB; else if (highB) Q1 += diffB;
    if (lowC) P2 += diffC; else if (highC) P1 += diffC;
  }

  /* ---------------------------------------------------------------- 2D TMz */

  /* the inline incident-wave interpolation of SchemeTMz.cpp:232-248 (dDiff 0.5 written as `- 0.5`)
   * and :645-663 (no dDiff term at all for E) */
  FV tmzInc (FP rx, FP ry, bool isH) const
  {
    FP x = rx - zeroInc[0];
    FP y = ry - zeroInc[1];
    FP d;
    if (isH) d = x * dcos (angle2) + y * dsin (angle2) - 0.5;
    else d = x * dcos (angle2) + y * dsin (angle2);
    FP coordD1 = (FP) ((uint64_t) d);
    FP coordD2 = coordD1 + 1;
    FP proportionD2 = d - coordD1;
    FP proportionD1 = 1 - proportionD2;
    const std::vector<FV> &F = isH ? HInc : EInc;
    return proportionD1 * F[(size_t) coordD1] + proportionD2 * F[(size_t) coordD2];
  }

  /* SchemeTMz::calculateEzStep / calculateEzStepPML, SchemeTMz.cpp:162-304, 307-584.
   * Materials are single-cell samples M[i,j] (cpp:181,326,528-530); sigma_z == 0 (cpp:2304-2316). */
  void stepEzTMz ()
  {
    const FP eps0 = PC::Eps0 ();
    std::vector<FV> &Dst = p.usePML ? D[2] : E[2];
#pragma omp parallel for schedule(static)
    for (int64_t i = 1; i < Nx; ++i)
      for (int64_t j = 1; j < Ny; ++j)
      {
        const int64_t pos[3] = { i, j, 0 };
        FV Hy1 = H[1][idx (i, j, 0)], Hy2 = H[1][idx (i - 1, j, 0)];
        FV Hx1 = H[0][idx (i, j, 0)], Hx2 = H[0][idx (i, j - 1, 0)];
        if (p.useTFSF)
        {
          bool left = needTFSF (true, 2, pos, 0, false), right = needTFSF (true, 2, pos, 0, true);
          bool down = needTFSF (true, 2, pos, 1, false), up = needTFSF (true, 2, pos, 1, true);
          FV diffX = FV (), diffY = FV ();
          if (left || right)
          {
            /* Hy at posRight = (i,j) for left, posLeft = (i-1,j) for right; Hy real = (i+1, j+0.5) */
            int64_t ii = left ? i : i - 1;
            diffX = tmzInc ((FP) ii + (FP) 1.0, (FP) j + (FP) 0.5, true);
          }
          if (down || up)
          {
            /* Hx at posUp = (i,j) for down, posDown = (i,j-1) for up; Hx real = (i+0.5, j+1) */
            int64_t jj = down ? j : j - 1;
            diffY = tmzInc ((FP) i + (FP) 0.5, (FP) jj + (FP) 1.0, true);
          }
          /* SchemeTMz.cpp:273-289 */
          if (left) Hy1 -= FVT<FP, CPLX>::times_d (-diffX, dcos (angle2));
          else if (right) Hy2 -= FVT<FP, CPLX>::times_d (-diffX, dcos (angle2));
          if (down) Hx1 -= FVT<FP, CPLX>::times_d (diffY, dsin (angle2));
          else if (up) Hx2 -= FVT<FP, CPLX>::times_d (diffY, dsin (angle2));
        }
        size_t id = idx (i, j, 0);
        FV old = Dst[id];
        if (p.usePML)
        {
          FP sigmaX = Sigma[0][i];
          FP k_x = 1;
          FP Ca = (2 * eps0 * k_x - sigmaX * gridTimeStep) / (2 * eps0 * k_x + sigmaX * gridTimeStep);
          FP Cb = (2 * eps0 * gridTimeStep / gridStep) / (2 * eps0 * k_x + sigmaX * gridTimeStep);
          FV val = Ca * old + Cb * (Hy1 - Hy2 - Hx1 + Hx2);
          Dst[id] = val;
          /* SchemeTMz.cpp:508-583 (same-cell dependency only, see stepE) */
          FP eps = Eps[midx (i, j, 0)];
          FP sigmaY = Sigma[1][j];
          FP sigmaZ = 0;
          FP modifier = eps * eps0;
          FP k_y = 1, k_z = 1;
          FP Ca3 = (2 * eps0 * k_y - sigmaY * gridTimeStep) / (2 * eps0 * k_y + sigmaY * gridTimeStep);
          FP Cb3 = ((2 * eps0 * k_z + sigmaZ * gridTimeStep) / (modifier)) / (2 * eps0 * k_y + sigmaY * gridTimeStep);
          FP Cc3 = ((2 * eps0 * k_z - sigmaZ * gridTimeStep) / (modifier)) / (2 * eps0 * k_y + sigmaY * gridTimeStep);
          E[2][id] = Ca3 * E[2][id] + Cb3 * val - Cc3 * old;
        }
        else
        {
          FP eps = Eps[midx (i, j, 0)];
          Dst[id] = FVT<FP, CPLX>::one_times (old) + ((gridTimeStep) / ((eps * eps0) * (gridStep))) * (Hy1 - Hy2 - Hx1 + Hx2);
        }
      }
  }

  /* SchemeTMz::calculateHxStep / ...PML (cpp:589-955) for a == 0 and calculateHyStep / ...PML
   * (cpp:958-1308) for a == 1.  Materials: 2-point averages `(M1 + M2) / 2` (cpp:606-612, 975-981). */
  void stepHTMz (int a)
  {
    const FP eps0 = PC::Eps0 ();
    const FP mu0 = PC::Mu0 ();
    const int64_t i0 = a == 0 ? 1 : 0, i1 = a == 0 ? Nx : Nx - 1;
    const int64_t j0 = a == 0 ? 0 : 1, j1 = a == 0 ? Ny - 1 : Ny;
    std::vector<FV> &Dst = p.usePML ? B[a] : H[a];
#pragma omp parallel for schedule(static)
    for (int64_t i = i0; i < i1; ++i)
      for (int64_t j = j0; j < j1; ++j)
      {
        const int64_t pos[3] = { i, j, 0 };
        /* Hx: Ez1 = Ez[posUp] = Ez[i,j+1], Ez2 = Ez[i,j]; Hy: Ez1 = Ez[posRight] = Ez[i+1,j], Ez2 = Ez[i,j] */
        int64_t i2 = a == 0 ? i : i + 1, j2 = a == 0 ? j + 1 : j;
        FV Ez1 = E[2][idx (i2, j2, 0)], Ez2 = E[2][idx (i, j, 0)];
        if (p.useTFSF)
        {
          int n = a == 0 ? 1 : 0;
          bool low = needTFSF (false, a, pos, n, false), high = needTFSF (false, a, pos, n, true);
          if (low || high)
          {
            /* Ez real = (i+0.5, j+0.5); low -> operand 2 at (i,j), high -> operand 1 */
            int64_t ii = low ? i : i2, jj = low ? j : j2;
            FV diff = tmzInc ((FP) ii + (FP) 0.5, (FP) jj + (FP) 0.5, false);
            if (low) Ez2 += diff; else Ez1 += diff;
          }
        }
        /* material samples: Hx real (i+0.5, j+1): y+0.5 -> M[i,j+1], y-0.5 -> M[i,j];
         * Hy real (i+1, j+0.5): x-0.5 -> M[i,j], x+0.5 -> M[i+1,j] */
        size_t m1 = a == 0 ? midx (i, j + 1, 0) : midx (i, j, 0);
        size_t m2 = a == 0 ? midx (i, j, 0) : midx (i + 1, j, 0);
        size_t id = idx (i, j, 0);
        FV old = Dst[id];
        if (p.usePML)
        {
          /* Bx <- sigma_y (cpp:770-771), By <- sigma_z == 0 (cpp:1139-1140) */
          FP sigma;
          if (a == 0) sigma = (Sigma[1][j + 1] + Sigma[1][j]) / 2;
          else sigma = ((FP) 0 + (FP) 0) / 2;
          FP kk = 1;
          FP Ca = (2 * eps0 * kk - sigma * gridTimeStep) / (2 * eps0 * kk + sigma * gridTimeStep);
          FP Cb = (2 * eps0 * gridTimeStep / gridStep) / (2 * eps0 * kk + sigma * gridTimeStep);
          /* Kernels.h:38-42: Hx: (0.0 - 0.0 - Ez1 + Ez2), Hy: (Ez1 - Ez2 - 0.0 + 0.0) */
          FV val;
          if (a == 0) val = Ca * old + Cb * (FVT<FP, CPLX>::zero () - FVT<FP, CPLX>::zero () - Ez1 + Ez2);
          else val = Ca * old + Cb * (Ez1 - Ez2 - FVT<FP, CPLX>::zero () + FVT<FP, CPLX>::zero ());
          Dst[id] = val;

          /* SchemeTMz.cpp:886-953, 1255-1307 (same-cell dependency only, see stepE) */
          FP mu = (Mu[m1] + Mu[m2]) / 2;
          FP sigmaN, sigmaD;
          if (a == 0)
          {
            /* Hx <- (sigma_z, sigma_x): cpp:925-927; both samples share i */
            sigmaN = (Sigma[0][i] + Sigma[0][i]) / 2;
            sigmaD = ((FP) 0 + (FP) 0) / 2;
          }
          else
          {
            /* Hy <- (sigma_x, sigma_y): cpp:1294-1296; samples at i and i+1, same j */
            sigmaN = (Sigma[1][j] + Sigma[1][j]) / 2;
            sigmaD = (Sigma[0][i] + Sigma[0][i + 1]) / 2;
          }
          FP modifier = mu * mu0;
          FP k_n = 1, k_d = 1;
          FP Ca3 = (2 * eps0 * k_d - sigmaD * gridTimeStep) / (2 * eps0 * k_d + sigmaD * gridTimeStep);
          FP Cb3 = ((2 * eps0 * k_n + sigmaN * gridTimeStep) / (modifier)) / (2 * eps0 * k_d + sigmaD * gridTimeStep);
          FP Cc3 = ((2 * eps0 * k_n - sigmaN * gridTimeStep) / (modifier)) / (2 * eps0 * k_d + sigmaD * gridTimeStep);
          H[a][id] = Ca3 * H[a][id] + Cb3 * val - Cc3 * old;
        }
        else
        {
          FP mu = (Mu[m1] + Mu[m2]) / 2;
          FP Cbv = (gridTimeStep) / ((mu * mu0) * (gridStep));
          if (a == 0) Dst[id] = FVT<FP, CPLX>::one_times (old) + Cbv * (FVT<FP, CPLX>::zero () - FVT<FP, CPLX>::zero () - Ez1 + Ez2);
          else Dst[id] = FVT<FP, CPLX>::one_times (old) + Cbv * (Ez1 - Ez2 - FVT<FP, CPLX>::zero () + FVT<FP, CPLX>::zero ());
        }
      }
  }

  /* ---------------------------------------------------------------- scattered field at save time */

  /* The block Scheme3D::performNSteps runs before saving results when --save-res is given
   * (Scheme3D.cpp:2767-2921, serial build): for each of the six fields, every cell whose real coordinate lies
   * inside the TF/SF box [leftBorderTotalField, rightBorderTotalField] (inclusive) gets the projected incident
   * field subtracted, in place. */
  void subtractIncident ()
  {
    for (int isE = 1; isE >= 0; --isE)
      for (int a = 0; a < 3; ++a)
      {
        std::vector<FV> &Fld = isE ? E[a] : H[a];
#pragma omp parallel for schedule(static)
        for (int64_t i = 0; i < Nx; ++i)
          for (int64_t j = 0; j < Ny; ++j)
            for (int64_t k = 0; k < Nz; ++k)
            {
              const int64_t pos[3] = { i, j, k };
              FP r[3];
              realCoord (isE != 0, a, pos, r);
              bool inside = true;
              for (int ax = 0; ax < 3; ++ax)
              {
                FP l = (FP) p.L[ax], rr = (FP) (p.N[ax] - p.L[ax]);
                if (r[ax] < l || r[ax] > rr) inside = false;
              }
              if (!inside) continue;
              FV incVal = isE ? fromIncidentE (a, approximateIncidentWave (r, (FP) 0.0, EInc))
                              : fromIncidentH (a, approximateIncidentWave (r, (FP) 0.5, HInc));
              Fld[idxv (pos)] = Fld[idxv (pos)] - incVal;
            }
      }
  }

  /* ---------------------------------------------------------------- time loop */

  /* Scheme3D::performNSteps, Scheme3D.cpp:1966-2069; SchemeTMz::performNSteps, SchemeTMz.cpp:1324-1427 */
  void performNSteps (int64_t nsteps)
  {
    for (int64_t s = 0; s < nsteps; ++s, ++t)
    {
      int tt = (int) t;
      if (p.useTFSF) performPlaneWaveESteps (tt);
      if (p.dims == 3) { stepE (0); stepE (1); stepE (2); }
      else stepEzTMz ();
      if (!p.useTFSF)
      {
        E[2][idx (p.src[0], p.src[1], p.dims == 3 ? p.src[2] : 0)] = sourceValue (tt);
      }
      if (p.useTFSF) performPlaneWaveHSteps (tt);
      if (p.dims == 3) { stepH (0); stepH (1); stepH (2); }
      else { stepHTMz (0); stepHTMz (1); }
    }
  }
};

/* ------------------------------------------------------------------ initialisation restated */

/* Scheme3D::initScheme, Scheme3D.cpp:3387-3402 (dims 3); SchemeTMz::initScheme, SchemeTMz.cpp:1948-1956
 * plus relPhi of SchemeTMz.cpp:35-49 (dims 2). */
template <class FP>
static void initScheme (int dims, FP dx, FP sourceFreq, FP a1, FP a2, FP &dt, FP &relv)
{
  typedef PhysConst<FP> PC;
  FP sourceFrequency = sourceFreq;
  FP sourceWaveLength = PC::SpeedOfLight () / sourceFrequency;
  FP gridStep = dx;
  if (dims == 3)
  {
    FP courantNum = 1.0 / 2.0;
    dt = gridStep * courantNum / PC::SpeedOfLight ();
    FP N_lambda = sourceWaveLength / gridStep;
    FP phaseVelocity0 = phaseVelocityIncidentWave3D<FP> (gridStep, sourceWaveLength, courantNum, N_lambda, PC::Pi () / 2, 0);
    FP phaseVelocity = phaseVelocityIncidentWave3D<FP> (gridStep, sourceWaveLength, courantNum, N_lambda, a1, a2);
    relv = phaseVelocity0 / phaseVelocity;
  }
  else
  {
    FP courantNum = 2.0;
    dt = gridStep / (courantNum * PC::SpeedOfLight ());
    FP S = 1 / courantNum;
    FP stepWaveLength = PC::SpeedOfLight () / (sourceFrequency * gridStep);
    FP arg = PC::Pi () * S / stepWaveLength;
    if (a2 == PC::Pi () / 4)
      relv = dsqrt (2) * dasin (dsin (arg) / (S * dsqrt (2))) / dasin (dsin (arg) / S);
    else
      relv = 1;
  }
}

/* PML conductivity profile along one axis, Scheme3D.cpp:3656-3716 (SchemeTMz.cpp:2212-2262): material
 * cell m has real coordinate m + 0.5 (getEpsCoordFP), the grid's "size" coordinate is (N+1) + 0.5;
 * `boundary` always uses the x PML size (Scheme3D.cpp:3661). */
template <class FP>
static void initSigmaProfile (int64_t N, int64_t pmlAxis, int64_t pmlX, FP gridStep, FP *out)
{
  typedef PhysConst<FP> PC;
  FP eps0 = PC::Eps0 ();
  FP mu0 = PC::Mu0 ();
  FP boundary = (uint64_t) pmlX * gridStep;
  uint32_t exponent = 6;
  FP R_err = 1e-16;
  FP sigma_max_1 = -dlog (R_err) * (exponent + 1.0) / (2.0 * dsqrt (mu0 / eps0) * boundary);
  FP boundaryFactor = sigma_max_1 / (gridStep * (dpow (boundary, exponent)) * (exponent + 1));
  FP size = (FP) (N + 1) + (FP) 0.5;
  uint64_t P = (uint64_t) pmlAxis;
  for (int64_t m = 0; m <= N; ++m)
  {
    FP posAbs = (FP) m + (FP) 0.5;
    FP val = 0;
    if (posAbs < P)
    {
      uint32_t dist = P - posAbs;
      FP x1 = (dist + 1) * gridStep;
      FP x2 = dist * gridStep;
      val = boundaryFactor * (dpow (x1, (exponent + 1)) - dpow (x2, (exponent + 1)));
    }
    else if (posAbs >= size - P)
    {
      uint32_t dist = posAbs - (size - P);
      FP x1 = (dist + 1) * gridStep;
      FP x2 = dist * gridStep;
      val = boundaryFactor * (dpow (x1, (exponent + 1)) - dpow (x2, (exponent + 1)));
    }
    out[m] = val;
  }
}

/* Approximation::approximateSphere (Approximation.cpp:303-331) over the Eps grid as in
 * Scheme3D::initGrids (Scheme3D.cpp:3413-3442); real part only. */
template <class FP>
static void initEpsSphere (int64_t Mx, int64_t My, int64_t Mz, FP cx, FP cy, FP cz, FP radius, FP epsIn, FP *out)
{
  for (int64_t i = 0; i < Mx; ++i)
    for (int64_t j = 0; j < My; ++j)
      for (int64_t k = 0; k < Mz; ++k)
      {
        FP px = (FP) i + (FP) 0.5, py = (FP) j + (FP) 0.5, pz = (FP) k + (FP) 0.5;
        FP d = dsqrt (ORACLE_SQR (px - cx) + ORACLE_SQR (py - cy) + ORACLE_SQR (pz - cz));
        FP diff = d - radius;
        FP eps_vacuum = 1;
        FP v;
        if (diff < -0.5) v = epsIn;
        else if (diff > 0.5) v = eps_vacuum;
        else
        {
          FP proportion = 0.5 - diff;
          v = proportion * epsIn + (1 - proportion) * eps_vacuum;
        }
        out[((size_t) i * My + j) * Mz + k] = v;
      }
}

} /* namespace oracle */

#endif /* FDTD_ORACLE_HPP */
