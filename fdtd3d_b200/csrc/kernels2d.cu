/*
 * 2D TMz (Ez, Hx, Hy) half-step kernels for sm_100a.
 *
 *   ez_phase_tmz : SchemeTMz::calculateEzStep[PML] incl. TF/SF and the point source
 *                  (reference: Source/Scheme/SchemeTMz.cpp:162-584, 1341-1355)
 *   h_phase_tmz  : SchemeTMz::calculateHxStep[PML] + calculateHyStep[PML] incl. TF/SF
 *                  (reference: Source/Scheme/SchemeTMz.cpp:589-1308)
 *
 * Same fusion as the 3D kernels (D/B update + E-from-D / H-from-B + TF/SF + source in one pass, in
 * place).  A 2D grid is stored like a 3D one with a single z plane (plane index 1); thread = one
 * 16-byte x-vector marching along y over `kchunk` rows with the y neighbour carried in registers.
 * Material sampling differs from 3D: Ez uses the single cell M[i,j] (cpp:181,326,528-530), Hx the
 * pair (M[i,j+1] + M[i,j])/2 (cpp:606-612), Hy (M[i,j] + M[i+1,j])/2 (cpp:975-981); sigma_z == 0.
 *
 * exact mode reproduces the reference's promotions for real float builds: the `0.0` literals of
 * Kernels.h:38-42 and the `1.0` of Kernels.h:45-61 make parts of the expression double.
 *
 * CORE variants (fast mode with UPML): vectors whose cells all have sigma_x = sigma_y = 0 neither load nor store
 * Dz / Bx / By and take the direct update E += (dt/dx) / (eps eps0) curl, as in 3D (kernels3d_tma.cu); the solver
 * folds / unfolds the D/B arrays of exactly those cells (p.core_lo/hi, x aligned to whole vectors).
 */
#include "fdtd_device.cuh"
#include "kernels2d_common.cuh"
#include "launch.h"

namespace fdtd
{

namespace
{

using namespace k2d;

constexpr int TX = 128;

/* rows [j0, j1) this CTA marches over: a chunk of the whole grid, or of one of the row ranges of p.reg (kb, ke = rows;
 * cta0 = first blockIdx.y of the range) when the launch covers only the rows around the one-pass kernel's box */
template <typename T>
__device__ __forceinline__ bool tmz_rows (const StepParams<T> &p, int &j0, int &j1)
{
  if (p.nreg == 0)
  {
    j0 = blockIdx.y * p.kchunk;
    j1 = min (j0 + p.kchunk, p.ny);
    return j0 < j1;
  }
  int r = 0;
  for (int q = 1; q < p.nreg; ++q)
    if ((int) blockIdx.y >= p.reg[q].cta0) r = q;
  j0 = p.reg[r].kb + ((int) blockIdx.y - p.reg[r].cta0) * p.kchunk;
  j1 = min (j0 + p.kchunk, p.reg[r].ke);
  return j0 < j1;
}

/* ------------------------------------------------------------------ Ez half step */

/* UNI: the material array of the phase (eps for Ez, mu for Hx/Hy) holds one value (p.eps_const / p.mu_const, set with
 * fdtd_fill): it is not loaded; (m + m) / 2 == m exactly, so exact mode is unaffected. */
template <typename T, int NC, bool PML, bool TFSF, bool EXACT, bool CORE, bool UNI>
__global__ void __launch_bounds__ (TX)
ez_phase_tmz (const __grid_constant__ StepParams<T> p)
{
  static_assert (!CORE || (PML && !EXACT), "the sigma-free core is a fast-mode UPML optimisation");
  constexpr int V = 16 / (NC * (int) sizeof (T));
  constexpr int W = V * NC;
  typedef Vec<T, W> FVec;
  typedef Vec<T, V> MVec;

  const int i0 = (blockIdx.x * TX + threadIdx.x) * V;
  if (i0 >= p.nx) return;
  int j0, j1;
  if (!tmz_rows (p, j0, j1)) return;

  const long long row = (long long) p.pitch * NC;
  long long o = (p.ps + (long long) j0 * p.pitch + i0) * NC; /* plane 1 */
  long long om = p.psM + (long long) j0 * p.pitchM + i0;

  const T *Ez = p.f[2];
  T *EzOut = p.fo[2], *Dz = p.f[8];
  const T *Hx = p.f[3], *Hy = p.f[4];

  MVec ca_x_at, cbd_x_at;
  T p_z = 0, m_z = 0;
  if (PML)
  {
    ca_x_at = ldv<T, V> (p.tab[0][KIND_AT].ca + i0);   /* Dz <- sigma_x, cpp:442-443 */
    cbd_x_at = ldv<T, V> (p.tab[0][KIND_AT].cbd + i0);
    p_z = p.tab[2][KIND_AT].p[0];                      /* Ez numerators <- sigma_z == 0, cpp:556-558 */
    m_z = p.tab[2][KIND_AT].m[0];
  }

  FVec hx_jm = ldv<T, W> (Hx + o - row);
  long long step = 0;
  if (!TFSF) step = current_step (p);
  const bool x_core = CORE && i0 >= p.core_lo[0] && i0 + V - 1 <= p.core_hi[0];

  for (int j = j0; j < j1; ++j, o += row, om += p.pitchM)
  {
    const bool jlo = j >= 1;
    const bool shell = !(x_core && j >= p.core_lo[1] && j <= p.core_hi[1]); /* false: no Dz needed for this vector */
    FVec hx = ldv<T, W> (Hx + o), hy = ldv<T, W> (Hy + o);
    T hy_im[NC];
#pragma unroll
    for (int c = 0; c < NC; ++c) hy_im[c] = __ldg (Hy + o - NC + c);
    MVec eps;
    if (UNI)
    {
#pragma unroll
      for (int v = 0; v < V; ++v) eps.v[v] = p.eps_const;
    }
    else eps = ldv<T, V> (p.eps + om);
    FVec ez = ldv_rw<T, W> (Ez + o);
    FVec dz;
    if (PML && shell) dz = ldv_rw<T, W> (Dz + o);

    T ca_y_at = 0, q_y_at = 0;
    if (PML)
    {
      ca_y_at = p.tab[1][KIND_AT].ca[j]; /* Ez <- sigma_y */
      q_y_at = EXACT ? p.tab[1][KIND_AT].p[j] : p.tab[1][KIND_AT].invp[j];
    }

#pragma unroll
    for (int v = 0; v < V; ++v)
    {
      const int i = i0 + v;
      const bool do_ez = i >= 1 && jlo;

      T hy1[NC], hy2[NC], hx1[NC], hx2[NC];
#pragma unroll
      for (int c = 0; c < NC; ++c)
      {
        const int s = v * NC + c;
        hy1[c] = hy.v[s];
        hy2[c] = v > 0 ? hy.v[v > 0 ? s - NC : s] : hy_im[c];
        hx1[c] = hx.v[s];
        hx2[c] = hx_jm.v[s];
      }

      if (TFSF)
      {
        const int Lx = p.L[0], Ly = p.L[1], Rx = p.R[0], Ry = p.R[1];
        T inc[NC];
        if (IN_RANGE (j, Ly, Ry - 1) && (i == Lx - 1 || i == Rx))
        {
          /* Hy at (i,j) [left] / (i-1,j) [right], real (i+1, j+0.5); Hy_inc = -H_inc cos(phi) */
          const bool low = i == Lx - 1;
          incident_tmz<T, NC> (p, true, low ? i : i - 1, j, (T) 1, (T) 0.5, inc);
#pragma unroll
          for (int c = 0; c < NC; ++c)
          {
            if (low) hy1[c] = sub_scaled<T, NC, EXACT> (hy1[c], -inc[c], p.cphi);
            else hy2[c] = sub_scaled<T, NC, EXACT> (hy2[c], -inc[c], p.cphi);
          }
        }
        if (IN_RANGE (i, Lx, Rx - 1) && (j == Ly - 1 || j == Ry))
        {
          /* Hx at (i,j) [down] / (i,j-1) [up], real (i+0.5, j+1); Hx_inc = H_inc sin(phi) */
          const bool low = j == Ly - 1;
          incident_tmz<T, NC> (p, true, i, low ? j : j - 1, (T) 0.5, (T) 1, inc);
#pragma unroll
          for (int c = 0; c < NC; ++c)
          {
            if (low) hx1[c] = sub_scaled<T, NC, EXACT> (hx1[c], inc[c], p.sphi);
            else hx2[c] = sub_scaled<T, NC, EXACT> (hx2[c], inc[c], p.sphi);
          }
        }
      }

      const T mod = eps.v[v] * p.eps0;
      if (do_ez)
      {
        if (PML && shell)
        {
          T cb3, cc3;
          sweep3_coefs<T, EXACT> (p_z, m_z, q_y_at, mod, cb3, cc3);
#pragma unroll
          for (int c = 0; c < NC; ++c)
          {
            const int s = v * NC + c;
            const T curl = hy1[c] - hy2[c] - hx1[c] + hx2[c];
            const T dold = dz.v[s];
            T dn, en;
            if (EXACT)
            {
              dn = ca_x_at.v[v] * dold + cbd_x_at.v[v] * curl;
              en = ca_y_at * ez.v[s] + cb3 * dn - cc3 * dold;
            }
            else
            {
              dn = fma_t<T> (cbd_x_at.v[v], curl, ca_x_at.v[v] * dold);
              en = fma_t<T> (ca_y_at, ez.v[s], fma_t<T> (cb3, dn, -(cc3 * dold)));
            }
            dz.v[s] = dn;
            ez.v[s] = en;
          }
        }
        else
        {
          const T cb = plain_coef<T, EXACT> (mod, p);
#pragma unroll
          for (int c = 0; c < NC; ++c)
          {
            const int s = v * NC + c;
            const T curl = hy1[c] - hy2[c] - hx1[c] + hx2[c];
            if (EXACT && sizeof (T) == 4 && NC == 1) ez.v[s] = (T) ((double) ez.v[s] + (double) (cb * curl));
            else if (EXACT) ez.v[s] = ez.v[s] + cb * curl;
            else ez.v[s] = fma_t<T> (cb, curl, ez.v[s]);
          }
        }
      }

      if (!TFSF)
      {
        if (j == p.srcpos[1] && i == p.srcpos[0])
        {
          /* the source overwrites Ez, not Dz (SchemeTMz.cpp:1341-1355): in the core this cell keeps its true Dz in
           * memory (not folded) and advances it here, Dz += Cb curl with Ca = 1 */
          if (CORE && !shell && do_ez)
          {
#pragma unroll
            for (int c = 0; c < NC; ++c)
              Dz[o + v * NC + c] = fma_t<T> (cbd_x_at.v[v], hy1[c] - hy2[c] - hx1[c] + hx2[c], Dz[o + v * NC + c]);
          }
#pragma unroll
          for (int c = 0; c < NC; ++c) ez.v[v * NC + c] = p.src[step * NC + c];
        }
      }
    }

    stv<T, W> (EzOut + o, ez);
    if (PML && shell) stv<T, W> (Dz + o, dz);
    hx_jm = hx;
  }
}

/* ------------------------------------------------------------------ H half step */

template <typename T, int NC, bool PML, bool TFSF, bool EXACT, bool CORE, bool UNI>
__global__ void __launch_bounds__ (TX)
h_phase_tmz (const __grid_constant__ StepParams<T> p)
{
  static_assert (!CORE || (PML && !EXACT), "the sigma-free core is a fast-mode UPML optimisation");
  constexpr int V = 16 / (NC * (int) sizeof (T));
  constexpr int W = V * NC;
  typedef Vec<T, W> FVec;
  typedef Vec<T, V> MVec;

  if (blockIdx.x == 0 && blockIdx.y == 0 && threadIdx.x == 0) advance_step (p); /* nobody reads it in this half step */
  const int i0 = (blockIdx.x * TX + threadIdx.x) * V;
  if (i0 >= p.nx) return;
  int j0, j1;
  if (!tmz_rows (p, j0, j1)) return;

  const long long row = (long long) p.pitch * NC;
  long long o = (p.ps + (long long) j0 * p.pitch + i0) * NC;
  long long om = p.psM + (long long) j0 * p.pitchM + i0;

  const T *Ez = p.f[2];
  const T *Hx = p.f[3], *Hy = p.f[4];
  T *HxOut = p.fo[3], *HyOut = p.fo[4], *Bx = p.f[9], *By = p.f[10];

  MVec p_x_at, m_x_at, ca_x_btw, q_x_btw;
  T ca_z = 0, cbd_z = 0, q_z = 0;
  if (PML)
  {
    p_x_at = ldv<T, V> (p.tab[0][KIND_AT].p + i0);     /* Hx numerators <- sigma_x at i, cpp:925-927 */
    m_x_at = ldv<T, V> (p.tab[0][KIND_AT].m + i0);
    ca_x_btw = ldv<T, V> (p.tab[0][KIND_BTW].ca + i0); /* Hy Ca / divisor <- sigma_x between i, i+1, cpp:1294-1296 */
    q_x_btw = ldv<T, V> ((EXACT ? p.tab[0][KIND_BTW].p : p.tab[0][KIND_BTW].invp) + i0);
    ca_z = p.tab[2][KIND_BTW].ca[0];                   /* By sweep 1 and Hx sweep 3 <- sigma_z == 0 */
    cbd_z = p.tab[2][KIND_BTW].cbd[0];
    q_z = EXACT ? p.tab[2][KIND_BTW].p[0] : p.tab[2][KIND_BTW].invp[0];
  }

  FVec ez = ldv<T, W> (Ez + o);
  MVec mu_c;
  if (UNI)
  {
#pragma unroll
    for (int v = 0; v < V; ++v) mu_c.v[v] = p.mu_const;
  }
  else mu_c = ldv<T, V> (p.mu + om);
  const bool x_core = CORE && i0 >= p.core_lo[0] && i0 + V - 1 <= p.core_hi[0];

  for (int j = j0; j < j1; ++j, o += row, om += p.pitchM)
  {
    const bool jlo = j >= 1, jhi = j < p.ny - 1;
    const bool shell = !(x_core && j >= p.core_lo[1] && j <= p.core_hi[1]);
    FVec ez_jp = ldv<T, W> (Ez + o + row);
    T ez_ip[NC];
#pragma unroll
    for (int c = 0; c < NC; ++c) ez_ip[c] = __ldg (Ez + o + W + c);
    MVec mu_jp = mu_c;
    T mu_ip = p.mu_const;
    if (!UNI)
    {
      mu_jp = ldv<T, V> (p.mu + om + p.pitchM);
      mu_ip = __ldg (p.mu + om + V);
    }

    FVec hx = ldv_rw<T, W> (Hx + o), hy = ldv_rw<T, W> (Hy + o);
    FVec bx, by;
    if (PML && shell) { bx = ldv_rw<T, W> (Bx + o); by = ldv_rw<T, W> (By + o); }

    T ca_y_btw = 0, cbd_y_btw = 0, p_y_at = 0, m_y_at = 0;
    if (PML)
    {
      ca_y_btw = p.tab[1][KIND_BTW].ca[j];   /* Bx <- sigma_y between j, j+1, cpp:770-771 */
      cbd_y_btw = p.tab[1][KIND_BTW].cbd[j];
      p_y_at = p.tab[1][KIND_AT].p[j];       /* Hy numerators <- sigma_y at j */
      m_y_at = p.tab[1][KIND_AT].m[j];
    }

#pragma unroll
    for (int v = 0; v < V; ++v)
    {
      const int i = i0 + v;
      const bool do_hx = i >= 1 && jhi;
      const bool do_hy = i < p.nx - 1 && jlo;

      T xE1[NC], xE2[NC], yE1[NC], yE2[NC]; /* Hx: Ez[j+1], Ez[j]; Hy: Ez[i+1], Ez[i] */
#pragma unroll
      for (int c = 0; c < NC; ++c)
      {
        const int s = v * NC + c;
        xE1[c] = ez_jp.v[s]; xE2[c] = ez.v[s];
        yE1[c] = v + 1 < V ? ez.v[v + 1 < V ? s + NC : s] : ez_ip[c];
        yE2[c] = ez.v[s];
      }

      if (TFSF)
      {
        const int Lx = p.L[0], Ly = p.L[1], Rx = p.R[0], Ry = p.R[1];
        T inc[NC];
        if (IN_RANGE (i, Lx, Rx - 1) && (j == Ly - 1 || j == Ry - 1))
        {
          const bool low = j == Ly - 1;
          incident_tmz<T, NC> (p, false, i, low ? j : j + 1, (T) 0.5, (T) 0.5, inc);
#pragma unroll
          for (int c = 0; c < NC; ++c) { if (low) xE2[c] += inc[c]; else xE1[c] += inc[c]; }
        }
        if (IN_RANGE (j, Ly, Ry - 1) && (i == Lx - 1 || i == Rx - 1))
        {
          const bool low = i == Lx - 1;
          incident_tmz<T, NC> (p, false, low ? i : i + 1, j, (T) 0.5, (T) 0.5, inc);
#pragma unroll
          for (int c = 0; c < NC; ++c) { if (low) yE2[c] += inc[c]; else yE1[c] += inc[c]; }
        }
      }

      const T m00 = mu_c.v[v];
      const T m10 = v + 1 < V ? mu_c.v[v + 1 < V ? v + 1 : v] : mu_ip;
      const T mod_x = ((mu_jp.v[v] + m00) / (T) 2) * p.mu0;
      const T mod_y = ((m00 + m10) / (T) 2) * p.mu0;

      if (PML && shell)
      {
        T cbx, ccx, cby, ccy;
        sweep3_coefs<T, EXACT> (p_x_at.v[v], m_x_at.v[v], q_z, mod_x, cbx, ccx);
        sweep3_coefs<T, EXACT> (p_y_at, m_y_at, q_x_btw.v[v], mod_y, cby, ccy);
#pragma unroll
        for (int c = 0; c < NC; ++c)
        {
          const int s = v * NC + c;
          if (do_hx)
          {
            const T bold = bx.v[s];
            const T bn = tmz_first<T, NC, EXACT> (false, true, bold, xE1[c], xE2[c], ca_y_btw, cbd_y_btw);
            hx.v[s] = EXACT ? ca_z * hx.v[s] + cbx * bn - ccx * bold
                            : fma_t<T> (ca_z, hx.v[s], fma_t<T> (cbx, bn, -(ccx * bold)));
            bx.v[s] = bn;
          }
          if (do_hy)
          {
            const T bold = by.v[s];
            const T bn = tmz_first<T, NC, EXACT> (false, false, bold, yE1[c], yE2[c], ca_z, cbd_z);
            hy.v[s] = EXACT ? ca_x_btw.v[v] * hy.v[s] + cby * bn - ccy * bold
                            : fma_t<T> (ca_x_btw.v[v], hy.v[s], fma_t<T> (cby, bn, -(ccy * bold)));
            by.v[s] = bn;
          }
        }
      }
      else
      {
        const T cbx = plain_coef<T, EXACT> (mod_x, p);
        const T cby = plain_coef<T, EXACT> (mod_y, p);
#pragma unroll
        for (int c = 0; c < NC; ++c)
        {
          const int s = v * NC + c;
          if (do_hx) hx.v[s] = tmz_first<T, NC, EXACT> (true, true, hx.v[s], xE1[c], xE2[c], (T) 1, cbx);
          if (do_hy) hy.v[s] = tmz_first<T, NC, EXACT> (true, false, hy.v[s], yE1[c], yE2[c], (T) 1, cby);
        }
      }
    }

    stv<T, W> (HxOut + o, hx); stv<T, W> (HyOut + o, hy);
    if (PML && shell) { stv<T, W> (Bx + o, bx); stv<T, W> (By + o, by); }
    ez = ez_jp;
    mu_c = mu_jp;
  }
}

template <typename T, int NC, bool PML, bool TFSF, bool EXACT, bool CORE = false>
cudaError_t launch_t (bool e_phase, bool uni, const StepParams<T> &p, cudaStream_t stream)
{
  constexpr int V = 16 / (NC * (int) sizeof (T));
  dim3 block (TX);
  dim3 grid ((p.nx + TX * V - 1) / (TX * V), (p.ny + p.kchunk - 1) / p.kchunk);
  StepParams<T> pr = p;
  if (p.nreg > 0)
  {
    int total = 0;
    for (int r = 0; r < p.nreg; ++r)
    {
      pr.reg[r].cta0 = total;
      if (p.reg[r].ke > p.reg[r].kb) total += (p.reg[r].ke - p.reg[r].kb + p.kchunk - 1) / p.kchunk;
    }
    if (total == 0) return cudaSuccess;
    grid.y = total;
  }
  if (e_phase)
  {
    if (uni) ez_phase_tmz<T, NC, PML, TFSF, EXACT, CORE, true><<<grid, block, 0, stream>>> (pr);
    else ez_phase_tmz<T, NC, PML, TFSF, EXACT, CORE, false><<<grid, block, 0, stream>>> (pr);
  }
  else
  {
    if (uni) h_phase_tmz<T, NC, PML, TFSF, EXACT, CORE, true><<<grid, block, 0, stream>>> (pr);
    else h_phase_tmz<T, NC, PML, TFSF, EXACT, CORE, false><<<grid, block, 0, stream>>> (pr);
  }
  return cudaGetLastError ();
}

template <typename T, int NC>
cudaError_t launch_nc (bool e_phase, bool uni, const StepParams<T> &p, bool pml, bool tfsf, bool exact, bool core, cudaStream_t s)
{
  if (core && pml && !exact)
    return tfsf ? launch_t<T, NC, true, true, false, true> (e_phase, uni, p, s) : launch_t<T, NC, true, false, false, true> (e_phase, uni, p, s);
  const int key = (pml ? 4 : 0) | (tfsf ? 2 : 0) | (exact ? 1 : 0);
  switch (key)
  {
    case 0: return launch_t<T, NC, false, false, false> (e_phase, uni, p, s);
    case 1: return launch_t<T, NC, false, false, true> (e_phase, uni, p, s);
    case 2: return launch_t<T, NC, false, true, false> (e_phase, uni, p, s);
    case 3: return launch_t<T, NC, false, true, true> (e_phase, uni, p, s);
    case 4: return launch_t<T, NC, true, false, false> (e_phase, uni, p, s);
    case 5: return launch_t<T, NC, true, false, true> (e_phase, uni, p, s);
    case 6: return launch_t<T, NC, true, true, false> (e_phase, uni, p, s);
    default: return launch_t<T, NC, true, true, true> (e_phase, uni, p, s);
  }
}

} /* namespace */

template <typename T>
cudaError_t launch_phase_tmz (bool e_phase, const StepParams<T> &p, bool cplx, bool pml, bool tfsf, bool exact, bool core,
                              bool uniform_material, cudaStream_t stream)
{
  if (cplx) return launch_nc<T, 2> (e_phase, uniform_material, p, pml, tfsf, exact, core, stream);
  return launch_nc<T, 1> (e_phase, uniform_material, p, pml, tfsf, exact, core, stream);
}

template cudaError_t launch_phase_tmz<float> (bool, const StepParams<float> &, bool, bool, bool, bool, bool, bool, cudaStream_t);
template cudaError_t launch_phase_tmz<double> (bool, const StepParams<double> &, bool, bool, bool, bool, bool, bool, cudaStream_t);

} /* namespace fdtd */
