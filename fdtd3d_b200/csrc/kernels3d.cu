/*
 * 3D Yee leapfrog kernels for sm_100a: one fused kernel per half step.
 *
 *   e_phase_3d : Scheme3D::calculate{Ex,Ey,Ez}Step[PML] + calculate{Ex,Ey,Ez}TFSF + the point source
 *                (reference: Source/Scheme/Scheme3D.cpp:138-1010, 1995-2024)
 *   h_phase_3d : Scheme3D::calculate{Hx,Hy,Hz}Step[PML] + calculate{Hx,Hy,Hz}TFSF
 *                (reference: Source/Scheme/Scheme3D.cpp:1029-1897)
 *
 * The reference makes three passes per component over arrays of pointers and then copies cur->prev;
 * here all three components, their auxiliary D/B update, the E-from-D update, the TF/SF correction and
 * the source are done in ONE pass per half step, in place (each update reads only its own old value
 * and the *other* field family), so every array is read and written once per half step:
 * 15 FieldValue + 1 FPValue of traffic per cell per half step.
 *
 * Memory-bound stencil design:
 *   - thread = one 16-byte x-vector of cells, marching along z over `kchunk` planes; a warp covers 512
 *     contiguous bytes of a row; CTA = 32 x 8 threads = an (x, y) tile;
 *   - the z neighbour (k-1 for E, k+1 for H) is carried in registers from one plane to the next, so it is
 *     read from memory once; the y neighbour row and the single x-neighbour cell are re-read by the
 *     adjacent thread/row of the same CTA and hit L1;
 *   - UPML coefficients come from 1-D tables: the x-indexed ones live in registers for the whole
 *     march, the y-indexed ones are per-thread scalars, the z-indexed ones are warp-uniform loads;
 *   - cells outside a component's compute range (YeeGridLayout.h:131-182) keep their value.
 *
 * exact mode (EXACT = true): reference operand order, IEEE division per cell, no FMA contraction (the
 * file is compiled with -fmad=false; fast mode uses explicit fma intrinsics).
 */
#include "fdtd_device.cuh"
#include "kernels3d_common.cuh"
#include "launch.h"

namespace fdtd
{

namespace
{

using namespace k3d;

/* ------------------------------------------------------------------ E half step */

template <typename T, int NC, bool PML, bool TFSF, bool EXACT>
__global__ void __launch_bounds__ (BX *BY, FDTD_MINB)
e_phase_3d (const __grid_constant__ StepParams<T> p, int kb, int ke)
{
  constexpr int V = CellsPerThread<T, NC>::value; /* cells per thread */
  constexpr int W = V * NC;                        /* scalars per field vector */
  typedef Vec<T, W> FVec;
  typedef Vec<T, V> MVec;

  const int i0 = (blockIdx.x * BX + threadIdx.x) * V;
  const int j = blockIdx.y * BY + threadIdx.y;
  if (i0 >= p.nx || j >= p.ny) return;
  const int k0 = kb + blockIdx.z * p.kchunk;
  const int k1 = min (k0 + p.kchunk, ke);
  if (k0 >= k1) return;

  long long o = (((long long) (k0 + 1) * p.ny + j) * p.pitch + i0) * NC;
  const long long row = (long long) p.pitch * NC;
  const long long pl = p.ps * NC;
  long long om = ((long long) (k0 + 1) * (p.ny + 1) + j) * p.pitchM + i0;

  T *Ex = p.f[0], *Ey = p.f[1], *Ez = p.f[2];
  const T *Hx = p.f[3], *Hy = p.f[4], *Hz = p.f[5];
  T *Dx = p.f[6], *Dy = p.f[7], *Dz = p.f[8];

  /* compute-range masks that do not depend on k (YeeGridLayout.h:131-170) */
  const bool jlo = j >= 1, jhi = j < p.ny - 1;
  const bool j_face = TFSF && tfsf_axis_hit (p, 1, j, true);

  /* per-thread UPML coefficients */
  T ca_y_at = 0, cbd_y_at = 0, p_y_btw = 0, m_y_btw = 0, q_y_at = 0;
  MVec p_x_btw, m_x_btw, ca_x_at, cbd_x_at, q_x_at;
  if (PML)
  {
    ca_y_at = p.tab[1][KIND_AT].ca[j];   /* Dx sweep 1 and Ez sweep 3 */
    cbd_y_at = p.tab[1][KIND_AT].cbd[j]; /* Dx */
    p_y_btw = p.tab[1][KIND_BTW].p[j];   /* Ey numerators */
    m_y_btw = p.tab[1][KIND_BTW].m[j];
    q_y_at = EXACT ? p.tab[1][KIND_AT].p[j] : p.tab[1][KIND_AT].invp[j]; /* Ez divisor */
    p_x_btw = ldv<T, V> (p.tab[0][KIND_BTW].p + i0); /* Ex numerators */
    m_x_btw = ldv<T, V> (p.tab[0][KIND_BTW].m + i0);
    ca_x_at = ldv<T, V> (p.tab[0][KIND_AT].ca + i0);   /* Dz sweep 1 and Ey sweep 3 */
    cbd_x_at = ldv<T, V> (p.tab[0][KIND_AT].cbd + i0); /* Dz */
    q_x_at = ldv<T, V> ((EXACT ? p.tab[0][KIND_AT].p : p.tab[0][KIND_AT].invp) + i0); /* Ey divisor */
  }

  /* z-carried values: H at k-1 and eps at k */
  FVec hx_km = ldv<T, W> (Hx + o - pl);
  FVec hy_km = ldv<T, W> (Hy + o - pl);
  MVec eps_c = ldv<T, V> (p.eps + om);

  long long step = 0;
  if (!TFSF) step = p.step;

  for (int k = k0; k < k1; ++k, o += pl, om += p.psM)
  {
    const int gk = p.kz0 + k;
    const bool klo = gk >= 1, khi = gk < p.nzg - 1;
    const bool jk_face = j_face || (TFSF && tfsf_axis_hit (p, 2, gk, true));

    FVec hx = ldv<T, W> (Hx + o), hy = ldv<T, W> (Hy + o), hz = ldv<T, W> (Hz + o);
    FVec hz_jm = ldv<T, W> (Hz + o - row), hx_jm = ldv<T, W> (Hx + o - row);
    T hz_im[NC], hy_im[NC];
#pragma unroll
    for (int c = 0; c < NC; ++c)
    {
      hz_im[c] = __ldg (Hz + o - NC + c);
      hy_im[c] = __ldg (Hy + o - NC + c);
    }
    MVec eps_kp = ldv<T, V> (p.eps + om + p.psM);
    MVec eps_jp = ldv<T, V> (p.eps + om + p.pitchM);
    T eps_ip = __ldg (p.eps + om + V);

    FVec ex = ldv_rw<T, W> (Ex + o), ey = ldv_rw<T, W> (Ey + o), ez = ldv_rw<T, W> (Ez + o);
    FVec dx, dy, dz;
    if (PML)
    {
      dx = ldv_rw<T, W> (Dx + o); dy = ldv_rw<T, W> (Dy + o); dz = ldv_rw<T, W> (Dz + o);
    }

    T ca_z_at = 0, cbd_z_at = 0, q_z_at = 0, p_z_btw = 0, m_z_btw = 0;
    if (PML)
    {
      ca_z_at = p.tab[2][KIND_AT].ca[gk];   /* Dy sweep 1 and Ex sweep 3 */
      cbd_z_at = p.tab[2][KIND_AT].cbd[gk]; /* Dy */
      q_z_at = EXACT ? p.tab[2][KIND_AT].p[gk] : p.tab[2][KIND_AT].invp[gk]; /* Ex divisor */
      p_z_btw = p.tab[2][KIND_BTW].p[gk];   /* Ez numerators */
      m_z_btw = p.tab[2][KIND_BTW].m[gk];
    }

#pragma unroll
    for (int v = 0; v < V; ++v)
    {
      const int i = i0 + v;
      const bool ilo = i >= 1, ihi = i < p.nx - 1;
      const bool do_ex = ihi && jlo && klo;
      const bool do_ey = ilo && jhi && klo;
      const bool do_ez = ilo && jlo && khi;

      /* curls, operand "1" = higher coordinate (Kernels.h:6):
       * Ex: Hz[j] - Hz[j-1] - Hy[k] + Hy[k-1];  Ey: Hx[k] - Hx[k-1] - Hz[i] + Hz[i-1];
       * Ez: Hy[i] - Hy[i-1] - Hx[j] + Hx[j-1] */
      T cex[NC], cey[NC], cez[NC];
#pragma unroll
      for (int c = 0; c < NC; ++c)
      {
        const int s = v * NC + c;
        const T hz_m = v > 0 ? hz.v[v > 0 ? s - NC : s] : hz_im[c];
        const T hy_m = v > 0 ? hy.v[v > 0 ? s - NC : s] : hy_im[c];
        cex[c] = hz.v[s] - hz_jm.v[s] - hy.v[s] + hy_km.v[s];
        cey[c] = hx.v[s] - hx_km.v[s] - hz.v[s] + hz_m;
        cez[c] = hy.v[s] - hy_m - hx.v[s] + hx_jm.v[s];
      }

      if (TFSF && (jk_face || tfsf_axis_hit (p, 0, i, true)))
      {
        /* cells on a TF/SF face (O(N^2) of them) redo their curl with corrected operands */
        const int fx = tfsf_e_faces (p, 0, i, j, gk), fy = tfsf_e_faces (p, 1, i, j, gk), fz = tfsf_e_faces (p, 2, i, j, gk);
        if (fx | fy | fz)
        {
#pragma unroll
          for (int c = 0; c < NC; ++c)
          {
            const int s = v * NC + c;
            const T hz_m = v > 0 ? hz.v[v > 0 ? s - NC : s] : hz_im[c];
            const T hy_m = v > 0 ? hy.v[v > 0 ? s - NC : s] : hy_im[c];
            if (fx) cex[c] = tfsf_curl<T, NC, true> (p, 0, fx, c, i, j, gk, hz.v[s], hz_jm.v[s], hy.v[s], hy_km.v[s]);
            if (fy) cey[c] = tfsf_curl<T, NC, true> (p, 1, fy, c, i, j, gk, hx.v[s], hx_km.v[s], hz.v[s], hz_m);
            if (fz) cez[c] = tfsf_curl<T, NC, true> (p, 2, fz, c, i, j, gk, hy.v[s], hy_m, hx.v[s], hx_jm.v[s]);
          }
        }
      }

      /* material at the three E positions: 2-point average along the component axis
       * (YeeGridLayout.h:1048-1106, Approximation.cpp:15-19) */
      const T e_c = eps_c.v[v];
      const T e_xp = v + 1 < V ? eps_c.v[v + 1 < V ? v + 1 : v] : eps_ip;
      const T mod_x = ((e_c + e_xp) / (T) 2) * p.eps0;
      const T mod_y = ((e_c + eps_jp.v[v]) / (T) 2) * p.eps0;
      const T mod_z = ((e_c + eps_kp.v[v]) / (T) 2) * p.eps0;

      if (PML)
      {
        T cbx, ccx, cby, ccy, cbz, ccz;
        sweep3_coefs<T, EXACT> (p_x_btw.v[v], m_x_btw.v[v], q_z_at, mod_x, cbx, ccx);
        sweep3_coefs<T, EXACT> (p_y_btw, m_y_btw, q_x_at.v[v], mod_y, cby, ccy);
        sweep3_coefs<T, EXACT> (p_z_btw, m_z_btw, q_y_at, mod_z, cbz, ccz);
#pragma unroll
        for (int c = 0; c < NC; ++c)
        {
          const int s = v * NC + c;
          if (do_ex)
            upml_cell<T, EXACT> (dx.v[s], ex.v[s], cex[c],
                                 ca_y_at, cbd_y_at, ca_z_at, cbx, ccx);
          if (do_ey)
            upml_cell<T, EXACT> (dy.v[s], ey.v[s], cey[c],
                                 ca_z_at, cbd_z_at, ca_x_at.v[v], cby, ccy);
          if (do_ez)
            upml_cell<T, EXACT> (dz.v[s], ez.v[s], cez[c],
                                 ca_x_at.v[v], cbd_x_at.v[v], ca_y_at, cbz, ccz);
        }
      }
      else
      {
        const T cbx = plain_coef<T, EXACT> (mod_x, p);
        const T cby = plain_coef<T, EXACT> (mod_y, p);
        const T cbz = plain_coef<T, EXACT> (mod_z, p);
#pragma unroll
        for (int c = 0; c < NC; ++c)
        {
          const int s = v * NC + c;
          if (do_ex) ex.v[s] = plain_cell<T, NC, EXACT> (ex.v[s], cex[c], cbx);
          if (do_ey) ey.v[s] = plain_cell<T, NC, EXACT> (ey.v[s], cey[c], cby);
          if (do_ez) ez.v[s] = plain_cell<T, NC, EXACT> (ez.v[s], cez[c], cbz);
        }
      }

      if (!TFSF)
      {
        /* hard point source written after the E sweeps (Scheme3D.cpp:1995-2024); D is not touched */
        if (gk == p.srcpos[2] && j == p.srcpos[1] && i == p.srcpos[0])
        {
#pragma unroll
          for (int c = 0; c < NC; ++c) ez.v[v * NC + c] = p.src[step * NC + c];
        }
      }
    }

    stv<T, W> (Ex + o, ex); stv<T, W> (Ey + o, ey); stv<T, W> (Ez + o, ez);
    if (PML)
    {
      stv<T, W> (Dx + o, dx); stv<T, W> (Dy + o, dy); stv<T, W> (Dz + o, dz);
    }

    hx_km = hx; hy_km = hy; eps_c = eps_kp;
  }
}

/* ------------------------------------------------------------------ H half step */

template <typename T, int NC, bool PML, bool TFSF, bool EXACT>
__global__ void __launch_bounds__ (BX *BY, FDTD_MINB)
h_phase_3d (const __grid_constant__ StepParams<T> p, int kb, int ke)
{
  constexpr int V = CellsPerThread<T, NC>::value;
  constexpr int W = V * NC;
  typedef Vec<T, W> FVec;
  typedef Vec<T, V> MVec;

  const int i0 = (blockIdx.x * BX + threadIdx.x) * V;
  const int j = blockIdx.y * BY + threadIdx.y;
  if (i0 >= p.nx || j >= p.ny) return;
  const int k0 = kb + blockIdx.z * p.kchunk;
  const int k1 = min (k0 + p.kchunk, ke);
  if (k0 >= k1) return;

  long long o = (((long long) (k0 + 1) * p.ny + j) * p.pitch + i0) * NC;
  const long long row = (long long) p.pitch * NC;
  const long long pl = p.ps * NC;
  long long om = ((long long) (k0 + 1) * (p.ny + 1) + j) * p.pitchM + i0;

  const T *Ex = p.f[0], *Ey = p.f[1], *Ez = p.f[2];
  T *Hx = p.f[3], *Hy = p.f[4], *Hz = p.f[5];
  T *Bx = p.f[9], *By = p.f[10], *Bz = p.f[11];

  const bool jlo = j >= 1, jhi = j < p.ny - 1;
  const bool j_face = TFSF && tfsf_axis_hit (p, 1, j, false);

  T ca_y_btw = 0, cbd_y_btw = 0, p_y_at = 0, m_y_at = 0, q_y_btw = 0;
  MVec p_x_at, m_x_at, ca_x_btw, cbd_x_btw, q_x_btw;
  if (PML)
  {
    ca_y_btw = p.tab[1][KIND_BTW].ca[j];   /* Bx sweep 1 and Hz sweep 3 */
    cbd_y_btw = p.tab[1][KIND_BTW].cbd[j]; /* Bx */
    p_y_at = p.tab[1][KIND_AT].p[j];       /* Hy numerators */
    m_y_at = p.tab[1][KIND_AT].m[j];
    q_y_btw = EXACT ? p.tab[1][KIND_BTW].p[j] : p.tab[1][KIND_BTW].invp[j]; /* Hz divisor */
    p_x_at = ldv<T, V> (p.tab[0][KIND_AT].p + i0); /* Hx numerators */
    m_x_at = ldv<T, V> (p.tab[0][KIND_AT].m + i0);
    ca_x_btw = ldv<T, V> (p.tab[0][KIND_BTW].ca + i0);   /* Bz sweep 1 and Hy sweep 3 */
    cbd_x_btw = ldv<T, V> (p.tab[0][KIND_BTW].cbd + i0); /* Bz */
    q_x_btw = ldv<T, V> ((EXACT ? p.tab[0][KIND_BTW].p : p.tab[0][KIND_BTW].invp) + i0); /* Hy divisor */
  }

  /* z-carried values: E at k (from the previous iteration's k+1 loads) and mu at k */
  FVec ex = ldv<T, W> (Ex + o), ey = ldv<T, W> (Ey + o);
  MVec mu_c = ldv<T, V> (p.mu + om), mu_jp = ldv<T, V> (p.mu + om + p.pitchM);
  T mu_ip = __ldg (p.mu + om + V), mu_ip_jp = __ldg (p.mu + om + p.pitchM + V);

  for (int k = k0; k < k1; ++k, o += pl, om += p.psM)
  {
    const int gk = p.kz0 + k;
    const bool klo = gk >= 1, khi = gk < p.nzg - 1;
    const bool jk_face = j_face || (TFSF && tfsf_axis_hit (p, 2, gk, false));

    FVec ex_kp = ldv<T, W> (Ex + o + pl), ey_kp = ldv<T, W> (Ey + o + pl);
    FVec ez = ldv<T, W> (Ez + o);
    FVec ex_jp = ldv<T, W> (Ex + o + row), ez_jp = ldv<T, W> (Ez + o + row);
    T ey_ip[NC], ez_ip[NC];
#pragma unroll
    for (int c = 0; c < NC; ++c)
    {
      ey_ip[c] = __ldg (Ey + o + W + c);
      ez_ip[c] = __ldg (Ez + o + W + c);
    }
    MVec mu_kp = ldv<T, V> (p.mu + om + p.psM), mu_jp_kp = ldv<T, V> (p.mu + om + p.psM + p.pitchM);
    T mu_ip_kp = __ldg (p.mu + om + p.psM + V), mu_ip_jp_kp = __ldg (p.mu + om + p.psM + p.pitchM + V);

    FVec hx = ldv_rw<T, W> (Hx + o), hy = ldv_rw<T, W> (Hy + o), hz = ldv_rw<T, W> (Hz + o);
    FVec bx, by, bz;
    if (PML)
    {
      bx = ldv_rw<T, W> (Bx + o); by = ldv_rw<T, W> (By + o); bz = ldv_rw<T, W> (Bz + o);
    }

    T ca_z_btw = 0, cbd_z_btw = 0, q_z_btw = 0, p_z_at = 0, m_z_at = 0;
    if (PML)
    {
      ca_z_btw = p.tab[2][KIND_BTW].ca[gk];   /* By sweep 1 and Hx sweep 3 */
      cbd_z_btw = p.tab[2][KIND_BTW].cbd[gk]; /* By */
      q_z_btw = EXACT ? p.tab[2][KIND_BTW].p[gk] : p.tab[2][KIND_BTW].invp[gk]; /* Hx divisor */
      p_z_at = p.tab[2][KIND_AT].p[gk];       /* Hz numerators */
      m_z_at = p.tab[2][KIND_AT].m[gk];
    }

#pragma unroll
    for (int v = 0; v < V; ++v)
    {
      const int i = i0 + v;
      const bool ilo = i >= 1, ihi = i < p.nx - 1;
      const bool do_hx = ilo && jhi && khi;
      const bool do_hy = ihi && jlo && khi;
      const bool do_hz = ihi && jhi && klo;

      /* curls: Hx: Ey[k+1] - Ey[k] - Ez[j+1] + Ez[j];  Hy: Ez[i+1] - Ez[i] - Ex[k+1] + Ex[k];
       * Hz: Ex[j+1] - Ex[j] - Ey[i+1] + Ey[i] */
      T chx[NC], chy[NC], chz[NC];
#pragma unroll
      for (int c = 0; c < NC; ++c)
      {
        const int s = v * NC + c;
        const T ez_p = v + 1 < V ? ez.v[v + 1 < V ? s + NC : s] : ez_ip[c];
        const T ey_p = v + 1 < V ? ey.v[v + 1 < V ? s + NC : s] : ey_ip[c];
        chx[c] = ey_kp.v[s] - ey.v[s] - ez_jp.v[s] + ez.v[s];
        chy[c] = ez_p - ez.v[s] - ex_kp.v[s] + ex.v[s];
        chz[c] = ex_jp.v[s] - ex.v[s] - ey_p + ey.v[s];
      }

      if (TFSF && (jk_face || tfsf_axis_hit (p, 0, i, false)))
      {
        const int fx = tfsf_h_faces (p, 0, i, j, gk), fy = tfsf_h_faces (p, 1, i, j, gk), fz = tfsf_h_faces (p, 2, i, j, gk);
        if (fx | fy | fz)
        {
#pragma unroll
          for (int c = 0; c < NC; ++c)
          {
            const int s = v * NC + c;
            const T ez_p = v + 1 < V ? ez.v[v + 1 < V ? s + NC : s] : ez_ip[c];
            const T ey_p = v + 1 < V ? ey.v[v + 1 < V ? s + NC : s] : ey_ip[c];
            if (fx) chx[c] = tfsf_curl<T, NC, false> (p, 0, fx, c, i, j, gk, ey_kp.v[s], ey.v[s], ez_jp.v[s], ez.v[s]);
            if (fy) chy[c] = tfsf_curl<T, NC, false> (p, 1, fy, c, i, j, gk, ez_p, ez.v[s], ex_kp.v[s], ex.v[s]);
            if (fz) chz[c] = tfsf_curl<T, NC, false> (p, 2, fz, c, i, j, gk, ex_jp.v[s], ex.v[s], ey_p, ey.v[s]);
          }
        }
      }

      /* material at the three H positions: ((m11 + m12)/2 + (m21 + m22)/2)/2 over the two transverse
       * axes, sample order of YeeGridLayout.h:1132-1135, 1161-1164, 1190-1193 */
      const T m000 = mu_c.v[v], m001 = mu_kp.v[v], m010 = mu_jp.v[v], m011 = mu_jp_kp.v[v];
      const T m100 = v + 1 < V ? mu_c.v[v + 1 < V ? v + 1 : v] : mu_ip;
      const T m101 = v + 1 < V ? mu_kp.v[v + 1 < V ? v + 1 : v] : mu_ip_kp;
      const T m110 = v + 1 < V ? mu_jp.v[v + 1 < V ? v + 1 : v] : mu_ip_jp;
      const T two = (T) 2;
      const T mod_x = (((m000 + m001) / two + (m010 + m011) / two) / two) * p.mu0; /* (j,k) (j,k+1) (j+1,k) (j+1,k+1) */
      const T mod_y = (((m000 + m001) / two + (m100 + m101) / two) / two) * p.mu0; /* (i,k) (i,k+1) (i+1,k) (i+1,k+1) */
      const T mod_z = (((m000 + m010) / two + (m100 + m110) / two) / two) * p.mu0; /* (i,j) (i,j+1) (i+1,j) (i+1,j+1) */

      if (PML)
      {
        T cbx, ccx, cby, ccy, cbz, ccz;
        sweep3_coefs<T, EXACT> (p_x_at.v[v], m_x_at.v[v], q_z_btw, mod_x, cbx, ccx);
        sweep3_coefs<T, EXACT> (p_y_at, m_y_at, q_x_btw.v[v], mod_y, cby, ccy);
        sweep3_coefs<T, EXACT> (p_z_at, m_z_at, q_y_btw, mod_z, cbz, ccz);
#pragma unroll
        for (int c = 0; c < NC; ++c)
        {
          const int s = v * NC + c;
          if (do_hx)
            upml_cell<T, EXACT> (bx.v[s], hx.v[s], chx[c],
                                 ca_y_btw, cbd_y_btw, ca_z_btw, cbx, ccx);
          if (do_hy)
            upml_cell<T, EXACT> (by.v[s], hy.v[s], chy[c],
                                 ca_z_btw, cbd_z_btw, ca_x_btw.v[v], cby, ccy);
          if (do_hz)
            upml_cell<T, EXACT> (bz.v[s], hz.v[s], chz[c],
                                 ca_x_btw.v[v], cbd_x_btw.v[v], ca_y_btw, cbz, ccz);
        }
      }
      else
      {
        const T cbx = plain_coef<T, EXACT> (mod_x, p);
        const T cby = plain_coef<T, EXACT> (mod_y, p);
        const T cbz = plain_coef<T, EXACT> (mod_z, p);
#pragma unroll
        for (int c = 0; c < NC; ++c)
        {
          const int s = v * NC + c;
          if (do_hx) hx.v[s] = plain_cell<T, NC, EXACT> (hx.v[s], chx[c], cbx);
          if (do_hy) hy.v[s] = plain_cell<T, NC, EXACT> (hy.v[s], chy[c], cby);
          if (do_hz) hz.v[s] = plain_cell<T, NC, EXACT> (hz.v[s], chz[c], cbz);
        }
      }
    }

    stv<T, W> (Hx + o, hx); stv<T, W> (Hy + o, hy); stv<T, W> (Hz + o, hz);
    if (PML)
    {
      stv<T, W> (Bx + o, bx); stv<T, W> (By + o, by); stv<T, W> (Bz + o, bz);
    }

    ex = ex_kp; ey = ey_kp;
    mu_c = mu_kp; mu_jp = mu_jp_kp; mu_ip = mu_ip_kp; mu_ip_jp = mu_ip_jp_kp;
  }
}

/* ------------------------------------------------------------------ dispatch */

template <typename T, int NC, bool PML, bool TFSF, bool EXACT>
cudaError_t launch_t (bool e_phase, const StepParams<T> &p, int kb, int ke, cudaStream_t stream)
{
  constexpr int V = CellsPerThread<T, NC>::value;
  if (ke <= kb) return cudaSuccess;
  dim3 block (BX, BY);
  dim3 grid ((p.nx + BX * V - 1) / (BX * V), (p.ny + BY - 1) / BY, (ke - kb + p.kchunk - 1) / p.kchunk);
  if (e_phase)
    e_phase_3d<T, NC, PML, TFSF, EXACT><<<grid, block, 0, stream>>> (p, kb, ke);
  else
    h_phase_3d<T, NC, PML, TFSF, EXACT><<<grid, block, 0, stream>>> (p, kb, ke);
  return cudaGetLastError ();
}

template <typename T, int NC>
cudaError_t launch_nc (bool e_phase, const StepParams<T> &p, bool pml, bool tfsf, bool exact, int kb, int ke,
                       cudaStream_t s)
{
  const int key = (pml ? 4 : 0) | (tfsf ? 2 : 0) | (exact ? 1 : 0);
  switch (key)
  {
    case 0: return launch_t<T, NC, false, false, false> (e_phase, p, kb, ke, s);
    case 1: return launch_t<T, NC, false, false, true> (e_phase, p, kb, ke, s);
    case 2: return launch_t<T, NC, false, true, false> (e_phase, p, kb, ke, s);
    case 3: return launch_t<T, NC, false, true, true> (e_phase, p, kb, ke, s);
    case 4: return launch_t<T, NC, true, false, false> (e_phase, p, kb, ke, s);
    case 5: return launch_t<T, NC, true, false, true> (e_phase, p, kb, ke, s);
    case 6: return launch_t<T, NC, true, true, false> (e_phase, p, kb, ke, s);
    default: return launch_t<T, NC, true, true, true> (e_phase, p, kb, ke, s);
  }
}

} /* namespace */

template <typename T>
cudaError_t launch_phase_3d (bool e_phase, const StepParams<T> &p, bool cplx, bool pml, bool tfsf, bool exact,
                             int kb, int ke, cudaStream_t stream)
{
  if (cplx) return launch_nc<T, 2> (e_phase, p, pml, tfsf, exact, kb, ke, stream);
  return launch_nc<T, 1> (e_phase, p, pml, tfsf, exact, kb, ke, stream);
}

template cudaError_t launch_phase_3d<float> (bool, const StepParams<float> &, bool, bool, bool, bool, int, int, cudaStream_t);
template cudaError_t launch_phase_3d<double> (bool, const StepParams<double> &, bool, bool, bool, bool, int, int, cudaStream_t);

} /* namespace fdtd */
