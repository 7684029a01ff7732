/*
 * TEST INFRASTRUCTURE ONLY -- C entry points of the Tier-2 oracle (oracle/fdtd_oracle.hpp) for
 * ctypes (tests/, __graft_entry__.smoke(), bench.py cpu_baseline / --impl reference).
 * The only thing shared with the product is the POD config struct of include/fdtd3d_b200.h.
 */
#include <cstdio>
#include <cstring>
#include <new>
#ifdef _OPENMP
#include <omp.h>
#endif

#include "../include/fdtd3d_b200.h"
#include "fdtd_oracle.hpp"

using namespace oracle;

namespace
{

Params to_params (const fdtd_config *c)
{
  Params p;
  p.dims = c->dims;
  p.usePML = c->use_pml != 0;
  p.useTFSF = c->use_tfsf != 0;
  for (int a = 0; a < 3; ++a)
  {
    p.N[a] = c->size[a];
    p.L[a] = c->tfsf_size[a];
    p.src[a] = c->source_pos[a];
  }
  if (p.dims == 2) { p.N[2] = 1; p.L[2] = 0; p.src[2] = 0; }
  p.dx = c->dx; p.dt = c->dt; p.freq = c->source_frequency; p.relv = c->rel_phase_velocity;
  p.a1 = c->angle_teta; p.a2 = c->angle_phi; p.a3 = c->angle_psi;
  p.t0 = c->time_step_start;
  return p;
}

template <class FP, bool CPLX>
int run_t (const fdtd_config *c, fdtd_host_arrays *arr, int64_t nsteps, bool scattered)
{
  typedef Scheme<FP, CPLX> S;
  typedef typename S::FV FV;
  S s (to_params (c));
  const int ncomp = c->dims == 3 ? 3 : 3;
  size_t n = (size_t) s.Nx * s.Ny * s.Nz;
  for (int a = 0; a < ncomp; ++a)
  {
    if (arr->field[FDTD_EX + a]) memcpy (s.E[a].data (), arr->field[FDTD_EX + a], n * sizeof (FV));
    if (arr->field[FDTD_HX + a]) memcpy (s.H[a].data (), arr->field[FDTD_HX + a], n * sizeof (FV));
    if (c->use_pml)
    {
      if (arr->field[FDTD_DX + a]) memcpy (s.D[a].data (), arr->field[FDTD_DX + a], n * sizeof (FV));
      if (arr->field[FDTD_BX + a]) memcpy (s.B[a].data (), arr->field[FDTD_BX + a], n * sizeof (FV));
    }
    if (arr->sigma[a]) memcpy (s.Sigma[a].data (), arr->sigma[a], s.Sigma[a].size () * sizeof (FP));
  }
  if (arr->eps) memcpy (s.Eps.data (), arr->eps, s.Eps.size () * sizeof (FP));
  if (arr->mu) memcpy (s.Mu.data (), arr->mu, s.Mu.size () * sizeof (FP));
  if (c->use_tfsf)
  {
    if (arr->einc) memcpy (s.EInc.data (), arr->einc, s.EInc.size () * sizeof (FV));
    if (arr->hinc) memcpy (s.HInc.data (), arr->hinc, s.HInc.size () * sizeof (FV));
  }

  s.performNSteps (nsteps);
  if (scattered && c->dims == 3 && c->use_tfsf) s.subtractIncident ();

  for (int a = 0; a < ncomp; ++a)
  {
    if (arr->field[FDTD_EX + a]) memcpy (arr->field[FDTD_EX + a], s.E[a].data (), n * sizeof (FV));
    if (arr->field[FDTD_HX + a]) memcpy (arr->field[FDTD_HX + a], s.H[a].data (), n * sizeof (FV));
    if (c->use_pml)
    {
      if (arr->field[FDTD_DX + a]) memcpy (arr->field[FDTD_DX + a], s.D[a].data (), n * sizeof (FV));
      if (arr->field[FDTD_BX + a]) memcpy (arr->field[FDTD_BX + a], s.B[a].data (), n * sizeof (FV));
    }
  }
  if (c->use_tfsf)
  {
    if (arr->einc) memcpy (arr->einc, s.EInc.data (), s.EInc.size () * sizeof (FV));
    if (arr->hinc) memcpy (arr->hinc, s.HInc.data (), s.HInc.size () * sizeof (FV));
  }
  return 0;
}

} /* namespace */

extern "C"
{

/* Source/main.cpp:79-85: angle_deg * PhysicsConst::Pi / 180.0 evaluated in FPValue / double as the
 * reference does. */
double oracle_angle_from_degrees (int value_type, double deg)
{
  if (value_type == 4)
  {
    float a = (float) deg * PhysConst<float>::Pi () / 180.0;
    return a;
  }
  double a = deg * PhysConst<double>::Pi () / 180.0;
  return a;
}

/* Source/main.cpp:148: PhysicsConst::SpeedOfLight / solverSettings.getSourceWaveLength () */
double oracle_source_frequency (int value_type, double wavelength)
{
  if (value_type == 4) return (float) (PhysConst<float>::SpeedOfLight () / (float) wavelength);
  return PhysConst<double>::SpeedOfLight () / wavelength;
}

/* Fills cfg->dt and cfg->rel_phase_velocity from dims, value_type, dx, source_frequency, angles. */
int oracle_init_scheme (fdtd_config *cfg)
{
  if (cfg->value_type == 4)
  {
    float dt, relv;
    initScheme<float> (cfg->dims, (float) cfg->dx, (float) cfg->source_frequency, (float) cfg->angle_teta, (float) cfg->angle_phi, dt, relv);
    cfg->dt = dt; cfg->rel_phase_velocity = relv;
  }
  else
  {
    double dt, relv;
    initScheme<double> (cfg->dims, cfg->dx, cfg->source_frequency, cfg->angle_teta, cfg->angle_phi, dt, relv);
    cfg->dt = dt; cfg->rel_phase_velocity = relv;
  }
  return 0;
}

/* 1-D sigma profile along `axis` (size[axis]+1 FPValues). */
int oracle_init_sigma (const fdtd_config *cfg, int axis, void *out)
{
  if (cfg->value_type == 4)
    initSigmaProfile<float> (cfg->size[axis], cfg->pml_size[axis], cfg->pml_size[0], (float) cfg->dx, (float *) out);
  else
    initSigmaProfile<double> (cfg->size[axis], cfg->pml_size[axis], cfg->pml_size[0], cfg->dx, (double *) out);
  return 0;
}

int oracle_init_eps_sphere (const fdtd_config *cfg, void *out, double cx, double cy, double cz, double radius, double eps_in)
{
  int64_t Mx = cfg->size[0] + 1, My = cfg->size[1] + 1, Mz = cfg->dims == 3 ? cfg->size[2] + 1 : 1;
  if (cfg->value_type == 4)
    initEpsSphere<float> (Mx, My, Mz, (float) cx, (float) cy, (float) cz, (float) radius, (float) eps_in, (float *) out);
  else
    initEpsSphere<double> (Mx, My, Mz, cx, cy, cz, radius, eps_in, (double *) out);
  return 0;
}

static int oracle_run_impl (const fdtd_config *cfg, fdtd_host_arrays *arr, int64_t nsteps, int nthreads, bool scattered);

/* Runs nsteps of the restated performNSteps on host arrays (reference order, in/out). */
int oracle_run (const fdtd_config *cfg, fdtd_host_arrays *arr, int64_t nsteps, int nthreads)
{
  return oracle_run_impl (cfg, arr, nsteps, nthreads, false);
}

/* Same, followed by the --save-res post-processing (total -> scattered field inside the TF/SF box). */
int oracle_run_scattered (const fdtd_config *cfg, fdtd_host_arrays *arr, int64_t nsteps, int nthreads)
{
  return oracle_run_impl (cfg, arr, nsteps, nthreads, true);
}

static int oracle_run_impl (const fdtd_config *cfg, fdtd_host_arrays *arr, int64_t nsteps, int nthreads, bool scattered)
{
#ifdef _OPENMP
  if (nthreads > 0) omp_set_num_threads (nthreads);
#endif
  try
  {
    if (cfg->value_type == 8 && !cfg->complex_values) return run_t<double, false> (cfg, arr, nsteps, scattered);
    if (cfg->value_type == 8 && cfg->complex_values) return run_t<double, true> (cfg, arr, nsteps, scattered);
    if (cfg->value_type == 4 && !cfg->complex_values) return run_t<float, false> (cfg, arr, nsteps, scattered);
    if (cfg->value_type == 4 && cfg->complex_values) return run_t<float, true> (cfg, arr, nsteps, scattered);
  }
  catch (const std::bad_alloc &)
  {
    fprintf (stderr, "oracle_run: out of memory\n");
    return 1;
  }
  return 2;
}

int oracle_max_threads (void)
{
#ifdef _OPENMP
  return omp_get_max_threads ();
#else
  return 1;
#endif
}

} /* extern "C" */
