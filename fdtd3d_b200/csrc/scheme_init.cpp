/*
 * Host-side mirror of the parts of the reference's Scheme set-up that produce inputs of the time step:
 * Scheme3D::initScheme / SchemeTMz::initScheme (dt, relative phase velocity), the angle and frequency
 * conversions of main(), and the PML conductivity profile of initGrids.  Plain host C++ (no CUDA);
 * part of the C ABI so that callers in any language get the reference's numbers.
 *
 * FPValue semantics: every quantity is computed in the build's FPValue type (float or double) with the
 * reference's operand order; transcendental functions are the C double functions, as they are in the
 * reference's translation units (which only include <cmath>, so unqualified sin/pow/log/sqrt/asin
 * resolve to ::sin(double) etc. even in float builds).
 */
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <string>

#include "../../include/fdtd3d_b200.h"

namespace
{

template <typename T> struct PC
{
  /* Source/Physics/PhysicsConst.h:9-13 */
  static T c () { return (T) 2.99792458e+8; }
  static T eps0 () { return (T) 0.0000000000088541878176203892; }
  static T mu0 () { return (T) 0.0000012566370614359173; }
  static T pi () { return (T) M_PI; }
};

template <typename T> T sqr (T x) { return x * x; }

/* Approximation::phaseVelocityIncidentWave3D, Source/Layout/Approximation.cpp:229-286.  The reference
 * writes `incidentWaveAngle1 = PhysicsConst::Pi / 2 && (...)` (assignment, not comparison) at :237 and
 * :250; that behaviour is kept: teta becomes 1 or 0 and the two special cases depend on phi only. */
template <typename T>
T phase_velocity (T delta, T wavelength, T courant, T n_lambda, T teta, T phi)
{
  const T pi = PC<T>::pi ();
  const bool axis_aligned = phi == 0 || phi == pi / 2 || phi == pi || phi == 3 * pi / 2;
  teta = (T) (pi / 2 && axis_aligned);
  if (teta)
    return PC<T>::c () * pi / (n_lambda * ::asin (::sin (pi * courant / n_lambda) / courant));
  const bool diagonal = phi == pi / 4 || phi == 3 * pi / 4 || phi == 5 * pi / 4 || phi == 7 * pi / 4;
  teta = (T) (pi / 2 && diagonal);
  if (teta)
    return PC<T>::c () * pi / (n_lambda * ::sqrt (2.0) * ::asin (::sin (pi * courant / n_lambda) / (courant * ::sqrt (2.0))));

  /* general direction: Newton iteration on the numerical dispersion relation, with teta == 0 by now */
  const T accuracy = 0.0000001; /* Approximation.cpp:7 */
  T k = 2 * pi;
  T k_prev = k + accuracy;
  const T nd = delta / wavelength;
  const T A = nd * ::sin (teta) * ::cos (phi) / 2;
  const T B = nd * ::sin (teta) * ::sin (phi) / 2;
  const T C = nd * ::cos (teta) / 2;
  const T D = sqr (::sin (pi * courant / n_lambda)) / sqr (courant);
  while (sqr (k_prev - k) >= accuracy)
  {
    k_prev = k;
    const T diff1 = sqr (::sin (A * k)) + sqr (::sin (B * k)) + sqr (::sin (C * k)) - D;
    const T diff2 = A * ::sin (2 * A * k) + B * ::sin (2 * B * k) + C * ::sin (2 * C * k);
    const T diff = diff1 / diff2;
    k -= diff;
  }
  return PC<T>::c () * 2 * pi / k;
}

template <typename T>
void init_scheme (fdtd_config *cfg, double wavelength, double teta_deg, double phi_deg, double psi_deg)
{
  /* Source/main.cpp:79-85: angle * PhysicsConst::Pi / 180.0 */
  const T a1 = (T) teta_deg * PC<T>::pi () / 180.0;
  const T a2 = (T) phi_deg * PC<T>::pi () / 180.0;
  const T a3 = (T) psi_deg * PC<T>::pi () / 180.0;
  /* Source/main.cpp:148: PhysicsConst::SpeedOfLight / sourceWaveLength */
  const T freq = PC<T>::c () / (T) wavelength;
  const T dx = (T) cfg->dx;
  const T lambda = PC<T>::c () / freq; /* Scheme3D.cpp:3391 */
  T dt, relv;
  if (cfg->dims == 3)
  {
    /* Scheme3D::initScheme, Source/Scheme/Scheme3D.cpp:3387-3402 */
    const T courant = 1.0 / 2.0;
    dt = dx * courant / PC<T>::c ();
    const T n_lambda = lambda / dx;
    const T v0 = phase_velocity<T> (dx, lambda, courant, n_lambda, PC<T>::pi () / 2, 0);
    const T v = phase_velocity<T> (dx, lambda, courant, n_lambda, a1, a2);
    relv = v0 / v;
  }
  else
  {
    /* SchemeTMz::initScheme, Source/Scheme/SchemeTMz.cpp:1948-1956, and the relPhi the 1-D incident
     * line recomputes every cell, SchemeTMz.cpp:35-49 (only phi = 0 and pi/4 are accepted there) */
    const T courant = 2.0;
    dt = dx / (courant * PC<T>::c ());
    const T S = 1 / courant;
    const T step_wavelength = PC<T>::c () / (freq * dx);
    const T arg = PC<T>::pi () * S / step_wavelength;
    if (a2 == PC<T>::pi () / 4)
      relv = ::sqrt (2) * ::asin (::sin (arg) / (S * ::sqrt (2))) / ::asin (::sin (arg) / S);
    else
      relv = 1;
  }
  cfg->dx = dx;
  cfg->dt = dt;
  cfg->source_frequency = freq;
  cfg->rel_phase_velocity = relv;
  cfg->angle_teta = a1;
  cfg->angle_phi = a2;
  cfg->angle_psi = a3;
}

/* Scheme3D::initGrids, Source/Scheme/Scheme3D.cpp:3656-3716 (same code for y, z at :3718-3818 and in
 * SchemeTMz.cpp:2212-2290): the sigma value of material cell m along one axis.  Material cell m sits at
 * real coordinate m + 0.5; the grid's far corner is at (N + 1) + 0.5; `boundary` uses the x PML size for
 * every axis (Scheme3D.cpp:3661). */
template <typename T>
void pml_profile (int64_t n, int64_t pml_axis, int64_t pml_x, T dx, T *out)
{
  const T eps0 = PC<T>::eps0 (), mu0 = PC<T>::mu0 ();
  const T boundary = (uint64_t) pml_x * dx;
  const uint32_t exponent = 6;
  const T r_err = 1e-16;
  const T sigma_max = -::log (r_err) * (exponent + 1.0) / (2.0 * ::sqrt (mu0 / eps0) * boundary);
  const T factor = sigma_max / (dx * (::pow (boundary, exponent)) * (exponent + 1));
  const T far = (T) (n + 1) + (T) 0.5;
  const uint64_t P = (uint64_t) pml_axis;
  for (int64_t m = 0; m <= n; ++m)
  {
    const T pos = (T) m + (T) 0.5;
    T val = 0;
    if (pos < P)
    {
      const uint32_t dist = P - pos;
      const T x1 = (dist + 1) * dx, x2 = dist * dx;
      val = factor * (::pow (x1, (exponent + 1)) - ::pow (x2, (exponent + 1)));
    }
    else if (pos >= far - P)
    {
      const uint32_t dist = pos - (far - P);
      const T x1 = (dist + 1) * dx, x2 = dist * dx;
      val = factor * (::pow (x1, (exponent + 1)) - ::pow (x2, (exponent + 1)));
    }
    out[m] = val;
  }
}

} /* namespace */

extern "C"
{

int fdtd_init_scheme (fdtd_config *cfg, double wavelength, double teta_deg, double phi_deg, double psi_deg)
{
  if (!cfg || (cfg->dims != 2 && cfg->dims != 3) || (cfg->value_type != 4 && cfg->value_type != 8) || !(cfg->dx > 0) || !(wavelength > 0))
    return FDTD_INVALID_ARGUMENT;
  if (cfg->value_type == 4) init_scheme<float> (cfg, wavelength, teta_deg, phi_deg, psi_deg);
  else init_scheme<double> (cfg, wavelength, teta_deg, phi_deg, psi_deg);
  /* point source position: Scheme3D.cpp:2011-2013 (grid centre), SchemeTMz.cpp:1347 ((70, Ny/2)) */
  if (cfg->dims == 3)
    for (int a = 0; a < 3; ++a) cfg->source_pos[a] = cfg->size[a] / 2;
  else
  {
    cfg->source_pos[0] = 70;
    cfg->source_pos[1] = cfg->size[1] / 2;
    cfg->source_pos[2] = 0;
  }
  return FDTD_OK;
}

int fdtd_pml_sigma_profile (const fdtd_config *cfg, int axis, void *out, int64_t count)
{
  if (!cfg || !out || axis < 0 || axis >= cfg->dims || count != cfg->size[axis] + 1) return FDTD_INVALID_ARGUMENT;
  if (cfg->pml_size[0] <= 0) return FDTD_INVALID_ARGUMENT;
  if (cfg->value_type == 4) pml_profile<float> (cfg->size[axis], cfg->pml_size[axis], cfg->pml_size[0], (float) cfg->dx, (float *) out);
  else pml_profile<double> (cfg->size[axis], cfg->pml_size[axis], cfg->pml_size[0], cfg->dx, (double *) out);
  return FDTD_OK;
}

} /* extern "C" */
