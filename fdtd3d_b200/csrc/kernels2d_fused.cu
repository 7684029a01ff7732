/*
 * 2D TMz: one pass per time step inside the sigma-free core, sm_100a -- the TMz counterpart of kernels3d_fused.cu.
 *
 * The half-step kernels (kernels2d.cu) move 9 field values per core cell and step (Ez pass: read Hx, Hy, read + write
 * Ez; H pass: read Ez, read + write Hx, Hy); this kernel reads Ez, Hx, Hy once and writes them once: 6 field values.
 * Reference semantics (SchemeTMz::performNSteps, SchemeTMz.cpp:1336-1427): Ez of step t from H of step t-1, then
 * Hx, Hy of step t from that Ez; per-cell arithmetic = the half-step kernels' (kernels2d_common.cuh).
 *
 * A CTA works on (x, y) tiles of TXC x TY cells, one below the other, staged by TMA (cp.async.bulk.tensor, a ring of
 * stages, one mbarrier each: boxes of Hy with an x halo, Hx with the row below, Ez, eps).  It covers one more 16-byte vector per row and
 * one more row than the tile (a warp per row + one warp for the rim column): every thread computes Ez' at its position, the rim threads for the first vector /
 * row of the neighbour tiles; Ez' goes through a double-buffered shared-memory tile, then the tile's own threads
 * compute Hx', Hy' from it.  Ez, Hx, Hy are ping-pong pairs (read `cur`, write `nxt`), because a neighbour CTA may
 * still need the old values this CTA would overwrite.  The rows outside the box (the y absorber) are advanced by the
 * half-step kernels on row ranges: Ez shell -> this kernel -> H shell; where the rim row lies outside the box its Ez'
 * is read from `nxt`.  The box spans the whole x range: lanes in the x absorber take the UPML path with Dz / Bx / By
 * (quarter-row boxes).  mu must be homogeneous (the reference's own set-up); eps may vary (UNIEPS = false).
 */
#include <cuda.h>

#include <string>

#include "fdtd_device.cuh"
#include "kernels2d_common.cuh"
#include "launch.h"
#include "tma_util.cuh"

namespace fdtd
{

namespace
{

using namespace k2d;
using namespace tma;

#ifndef FDTD_FUSED2D_TY
#define FDTD_FUSED2D_TY 8
#endif
#ifndef FDTD_FUSED2D_STAGES
#define FDTD_FUSED2D_STAGES 3
#endif

/* maps: 0 Hy (A) 1 Hx (C) 2 Ez 3 eps 4 Dz quarter rows 5 Bx 6 By quarter rows */
struct Fused2DSet
{
  CUtensorMap m[7];
};

template <typename T, int NC>
struct FTile2
{
  static constexpr int TY = FDTD_FUSED2D_TY;
  static constexpr int LX = 33;
  static constexpr int NTHREADS = 32 * (TY + 2);          /* a warp per tile row and rim row + one warp for the x rim column */
  static constexpr int CELL = NC * (int) sizeof (T);
  static constexpr int V = 16 / CELL;
  static constexpr int TXC = 32 * V;
  static constexpr int HM = 16 / (int) sizeof (T);
  static constexpr int QX = 8 * V;
  static constexpr int RW_A = (TXC + 2 * V) * CELL;   /* Hy rows: one vector of x halo on either side */
  static constexpr int RW_C = (TXC + V) * CELL;       /* Hx, Ez rows: the rim vector on the high side */
  static constexpr int MW = (TXC + V + HM - 1) / HM * HM;
  static constexpr int RW_M = MW * (int) sizeof (T);
  static constexpr int SZ_A = RW_A * (TY + 1);
  static constexpr int SZ_C = RW_C * (TY + 2);        /* rows j0 - 1 .. j0 + TY */
  static constexpr int SZ_E = RW_C * (TY + 1);
  static constexpr int SZ_M = RW_M * (TY + 1);
  static constexpr int SZ_QD = 128 * (TY + 1);
  static constexpr int SZ_QB = 128 * TY;
  static constexpr int OFF_A = 0;
  static constexpr int OFF_C = OFF_A + align128 (SZ_A);
  static constexpr int OFF_E = OFF_C + align128 (SZ_C);
  static constexpr int OFF_M = OFF_E + align128 (SZ_E);
  static constexpr int OFF_Q = OFF_M + align128 (SZ_M);
  static constexpr int X_BYTES = align128 (SZ_E);
  static constexpr int STAGES = FDTD_FUSED2D_STAGES;
  __host__ __device__ static constexpr int stage_bytes (int nq) { return OFF_Q + nq * (SZ_QD + 2 * SZ_QB); }
  __host__ __device__ static constexpr int smem_bytes (int nq) { return STAGES * stage_bytes (nq) + 2 * X_BYTES + 128; }
};

/* ------------------------------------------------------------------ the kernel */

template <typename T, int NC, bool TFSF, bool UNIEPS>
__global__ void __launch_bounds__ ((FTile2<T, NC>::NTHREADS))
step_core_tmz_fused (const __grid_constant__ StepParams<T> p, const __grid_constant__ Fused2DSet maps, const FusedBox fb)
{
  typedef FTile2<T, NC> L;
  constexpr int V = L::V, W = V * NC, S = L::STAGES, TY = L::TY;
  typedef Vec<T, W> FVec;
  typedef Vec<T, V> MVec;

  extern __shared__ unsigned char smem_raw[];
  __shared__ __align__ (8) unsigned long long full[S];
  unsigned char *smem = reinterpret_cast<unsigned char *> ((reinterpret_cast<uintptr_t> (smem_raw) + 127) & ~(uintptr_t) 127);
  const int stage_sz = L::stage_bytes (fb.nq);
  unsigned char *xbuf = smem + S * stage_sz;

  /* warps 0 .. TY: lane = x vector of row `warp` (row TY is the y rim); warp TY + 1: the x rim column, lane = row.
   * Row warps read, write and store whole contiguous rows; the rim warp touches one vector per row. */
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const bool valid = warp <= TY || lane <= TY;
  const int lx = warp <= TY ? lane : 32, ly = warp <= TY ? warp : (lane <= TY ? lane : 0);
  const bool leader = threadIdx.x == 0;
  const int i0t = blockIdx.x * L::TXC;
  const int ntiles = (fb.fy1 - fb.fy0 + 1) / TY;
  const int t0 = blockIdx.y * fb.kc, t1 = min (t0 + fb.kc, ntiles);
  if (t0 >= t1) return;
  const int i0 = i0t + lx * V;

  int qmask = 0;
#pragma unroll
  for (int q = 0; q < 4; ++q)
  {
    const int qa = i0t + q * L::QX, qb = qa + L::QX - 1;
    if (qa < p.nx && !(qa >= p.core_lo[0] && qb <= p.core_hi[0])) qmask |= 1 << q;
  }
  const int myq = lx >> 3;
  const bool shell = lx < 32 && ((qmask >> myq) & 1);
  const int qrank = __popc (qmask & ((1 << myq) - 1));
  /* compact quarter index in the Dz ping-pong arrays (launch.h, FusedBox): a rim thread needs the OLD Dz of a cell
   * the tile below it may already have updated, so Dz of the absorber quarters is a read-cur / write-nxt pair too */
  auto compact_q = [&] (int q) -> int
  {
    const int gq = i0t / L::QX + q;
    return gq < fb.nlo ? gq : fb.nlo + gq - fb.hq0;
  };

  auto issue = [&] (int slot, int t)
  {
    const int j0t = fb.fy0 + t * TY;
    unsigned char *base = smem + slot * stage_sz;
    unsigned long long *bar = &full[slot];
    mbar_expect_tx (bar, L::SZ_A + L::SZ_C + L::SZ_E + (UNIEPS ? 0 : L::SZ_M) + __popc (qmask) * (L::SZ_QD + 2 * L::SZ_QB));
    tma_load_3d (base + L::OFF_A, &maps.m[0], bar, (i0t - V) * NC, j0t, 1);
    tma_load_3d (base + L::OFF_C, &maps.m[1], bar, i0t * NC, j0t - 1, 1);
    tma_load_3d (base + L::OFF_E, &maps.m[2], bar, i0t * NC, j0t, 1);
    if (!UNIEPS) tma_load_3d (base + L::OFF_M, &maps.m[3], bar, i0t, j0t, 1);
    if (qmask)
    {
      int r = 0;
#pragma unroll
      for (int q = 0; q < 4; ++q)
        if ((qmask >> q) & 1)
        {
          tma_load_3d (base + L::OFF_Q + r * L::SZ_QD, &maps.m[4], bar, compact_q (q) * L::QX * NC, j0t, 1);
          tma_load_3d (base + L::OFF_Q + fb.nq * L::SZ_QD + r * L::SZ_QB, &maps.m[5], bar, (i0t + q * L::QX) * NC, j0t, 1);
          tma_load_3d (base + L::OFF_Q + fb.nq * (L::SZ_QD + L::SZ_QB) + r * L::SZ_QB, &maps.m[6], bar, (i0t + q * L::QX) * NC, j0t, 1);
          ++r;
        }
    }
  };

  if (leader)
  {
    for (int s = 0; s < S; ++s) mbar_init (&full[s], 1);
    asm volatile ("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads ();
  if (leader)
  {
    for (int s = 0; s < S && t0 + s < t1; ++s) issue (s, t0 + s);
  }

  const T mod_h = p.mu_const * p.mu0; /* homogeneous mu: (m + m) / 2 is m exactly */
  const T cb_h = plain_coef<T, false> (mod_h, p);
  long long step = 0;
  if (!TFSF) step = current_step (p);

  const int xo = lx * 16;
  const int oA = L::OFF_A + ly * L::RW_A + 16 + xo;
  const int oC_own = L::OFF_C + (ly + 1) * L::RW_C + xo, oC_jm = L::OFF_C + ly * L::RW_C + xo;
  const int oE = L::OFF_E + ly * L::RW_C + xo;
  const int oM = L::OFF_M + ly * L::RW_M + lx * V * (int) sizeof (T);
  const int oQ = ly * 128 + (lx & 7) * 16;
  const int oX = ly * L::RW_C + xo;

  for (int t = t0; t < t1; ++t)
  {
    const int n = t - t0;
    const int slot = n % S;
    const unsigned char *st = smem + slot * stage_sz;
    const int j0t = fb.fy0 + t * TY;
    const int j = j0t + ly;
    const bool ingrid = valid && i0 < p.nx && j < p.ny;
    const bool own = lx < 32 && ly < TY && ingrid;
    const bool rim_from_nxt = ly == TY && j0t + TY > fb.fy1; /* a shell row: Ez' as the Ez shell pass wrote it */
    const long long o = (p.ps + (long long) j * p.pitch + i0) * NC;
    mbar_wait (&full[slot], (n / S) & 1);

    /* ---------------- Ez' at this thread's position */
    const FVec hy = lds<T, W> (st + oA);
    const FVec hx = lds<T, W> (st + oC_own), hx_jm = lds<T, W> (st + oC_jm);
    T hy_im[NC];
#pragma unroll
    for (int c = 0; c < NC; ++c) hy_im[c] = *reinterpret_cast<const T *> (st + oA - L::CELL + c * sizeof (T));
    FVec ez = lds<T, W> (st + oE);
    MVec eps;
    if (UNIEPS)
    {
#pragma unroll
      for (int v = 0; v < V; ++v) eps.v[v] = p.eps_const;
    }
    else eps = lds<T, V> (st + oM);
    FVec dz, bx, by;
    if (shell)
    {
      dz = lds<T, W> (st + L::OFF_Q + qrank * L::SZ_QD + oQ);
      if (ly < TY)
      {
        bx = lds<T, W> (st + L::OFF_Q + fb.nq * L::SZ_QD + qrank * L::SZ_QB + oQ);
        by = lds<T, W> (st + L::OFF_Q + fb.nq * (L::SZ_QD + L::SZ_QB) + qrank * L::SZ_QB + oQ);
      }
    }
    const bool jlo = j >= 1, jhi = j < p.ny - 1;

    if (ingrid && !rim_from_nxt)
    {
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
            const bool low = i == Lx - 1;
            incident_tmz<T, NC> (p, true, low ? i : i - 1, j, (T) 1, (T) 0.5, inc);
#pragma unroll
            for (int c = 0; c < NC; ++c)
            {
              if (low) hy1[c] = sub_scaled<T, NC, false> (hy1[c], -inc[c], p.cphi);
              else hy2[c] = sub_scaled<T, NC, false> (hy2[c], -inc[c], p.cphi);
            }
          }
          if (IN_RANGE (i, Lx, Rx - 1) && (j == Ly - 1 || j == Ry))
          {
            const bool low = j == Ly - 1;
            incident_tmz<T, NC> (p, true, i, low ? j : j - 1, (T) 0.5, (T) 1, inc);
#pragma unroll
            for (int c = 0; c < NC; ++c)
            {
              if (low) hx1[c] = sub_scaled<T, NC, false> (hx1[c], inc[c], p.sphi);
              else hx2[c] = sub_scaled<T, NC, false> (hx2[c], inc[c], p.sphi);
            }
          }
        }
        const T mod = eps.v[v] * p.eps0;
        if (do_ez)
        {
          if (shell)
          {
            T cb3, cc3;
            sweep3_coefs<T, false> (p.tab[2][KIND_AT].p[0], p.tab[2][KIND_AT].m[0], p.tab[1][KIND_AT].invp[j], mod, cb3, cc3);
            const T ca_x = p.tab[0][KIND_AT].ca[i], cbd_x = p.tab[0][KIND_AT].cbd[i], ca_y = p.tab[1][KIND_AT].ca[j];
#pragma unroll
            for (int c = 0; c < NC; ++c)
            {
              const int s = v * NC + c;
              const T curl = hy1[c] - hy2[c] - hx1[c] + hx2[c];
              const T dold = dz.v[s];
              const T dn = fma_t<T> (cbd_x, curl, ca_x * dold);
              ez.v[s] = fma_t<T> (ca_y, ez.v[s], fma_t<T> (cb3, dn, -(cc3 * dold)));
              dz.v[s] = dn;
            }
          }
          else
          {
            const T cb = plain_coef<T, false> (mod, p);
#pragma unroll
            for (int c = 0; c < NC; ++c)
            {
              const int s = v * NC + c;
              ez.v[s] = fma_t<T> (cb, hy1[c] - hy2[c] - hx1[c] + hx2[c], ez.v[s]);
            }
          }
        }
        if (!TFSF)
        {
          if (j == p.srcpos[1] && i == p.srcpos[0])
          {
            /* the source overwrites Ez, not Dz (SchemeTMz.cpp:1341-1355); see ez_phase_tmz */
            if (!shell && do_ez && own)
            {
#pragma unroll
              for (int c = 0; c < NC; ++c)
                p.f[8][o + v * NC + c] = fma_t<T> (p.tab[0][KIND_AT].cbd[i], hy1[c] - hy2[c] - hx1[c] + hx2[c], p.f[8][o + v * NC + c]);
            }
#pragma unroll
            for (int c = 0; c < NC; ++c) ez.v[v * NC + c] = p.src[step * NC + c];
          }
        }
      }
      if (own)
      {
        stv<T, W> (p.fo[2] + o, ez);
        if (shell)
        {
          const long long od = ((((long long) p.ny + j) * fb.dq_pitch) + compact_q (myq) * L::QX + (lx & 7) * V) * NC; /* plane 1 */
          stv<T, W> ((T *) fb.dq_out[0] + od, dz);
        }
      }
    }
    else if (ingrid) ez = ldv_rw<T, W> (p.fo[2] + o);

    unsigned char *xw = xbuf + (t & 1) * L::X_BYTES;
    if (valid) *reinterpret_cast<FVec *> (xw + oX) = ez;

    /* Ez' of the whole tile is visible; every thread has read the stage: refill it with the tile S further down */
    __syncthreads ();
    if (leader && t + S < t1) issue (slot, t + S);

    /* ---------------- Hx', Hy' of the tile's own cells */
    if (own)
    {
      const FVec ez_jp = *reinterpret_cast<const FVec *> (xw + oX + L::RW_C);
      T ez_ip[NC];
#pragma unroll
      for (int c = 0; c < NC; ++c) ez_ip[c] = *reinterpret_cast<const T *> (xw + oX + 16 + c * sizeof (T));
      FVec hxn = hx, hyn = hy;
#pragma unroll
      for (int v = 0; v < V; ++v)
      {
        const int i = i0 + v;
        const bool do_hx = i >= 1 && jhi;
        const bool do_hy = i < p.nx - 1 && jlo;
        T xE1[NC], xE2[NC], yE1[NC], yE2[NC];
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
        if (shell)
        {
          T cbx, ccx, cby, ccy;
          sweep3_coefs<T, false> (p.tab[0][KIND_AT].p[i], p.tab[0][KIND_AT].m[i], p.tab[2][KIND_BTW].invp[0], mod_h, cbx, ccx);
          sweep3_coefs<T, false> (p.tab[1][KIND_AT].p[j], p.tab[1][KIND_AT].m[j], p.tab[0][KIND_BTW].invp[i], mod_h, cby, ccy);
          const T ca_z = p.tab[2][KIND_BTW].ca[0], cbd_z = p.tab[2][KIND_BTW].cbd[0];
          const T ca_y = p.tab[1][KIND_BTW].ca[j], cbd_y = p.tab[1][KIND_BTW].cbd[j], ca_x = p.tab[0][KIND_BTW].ca[i];
#pragma unroll
          for (int c = 0; c < NC; ++c)
          {
            const int s = v * NC + c;
            if (do_hx)
            {
              const T bold = bx.v[s];
              const T bn = tmz_first<T, NC, false> (false, true, bold, xE1[c], xE2[c], ca_y, cbd_y);
              hxn.v[s] = fma_t<T> (ca_z, hxn.v[s], fma_t<T> (cbx, bn, -(ccx * bold)));
              bx.v[s] = bn;
            }
            if (do_hy)
            {
              const T bold = by.v[s];
              const T bn = tmz_first<T, NC, false> (false, false, bold, yE1[c], yE2[c], ca_z, cbd_z);
              hyn.v[s] = fma_t<T> (ca_x, hyn.v[s], fma_t<T> (cby, bn, -(ccy * bold)));
              by.v[s] = bn;
            }
          }
        }
        else
        {
#pragma unroll
          for (int c = 0; c < NC; ++c)
          {
            const int s = v * NC + c;
            if (do_hx) hxn.v[s] = tmz_first<T, NC, false> (true, true, hxn.v[s], xE1[c], xE2[c], (T) 1, cb_h);
            if (do_hy) hyn.v[s] = tmz_first<T, NC, false> (true, false, hyn.v[s], yE1[c], yE2[c], (T) 1, cb_h);
          }
        }
      }
      stv<T, W> (p.fo[3] + o, hxn); stv<T, W> (p.fo[4] + o, hyn);
      if (shell) { stv<T, W> (p.f[9] + o, bx); stv<T, W> (p.f[10] + o, by); }
    }
  }
}

/* ------------------------------------------------------------------ host side */

template <typename T, int NC>
bool prepare_fused2_t (FusedMaps3D *out, void *ez, void *hx, void *hy, void *dz, long long dq_pitch, void *bx, void *by, void *eps,
                       int pitch, int ny, int pitchM, std::string *err)
{
  typedef FTile2<T, NC> L;
  static_assert (sizeof (Fused2DSet) <= sizeof (out->set), "FusedMaps3D too small");
  Fused2DSet *m = reinterpret_cast<Fused2DSet *> (out->set);
  const int sb = (int) sizeof (T);
  const long long xs = (long long) pitch * NC;
  /* a TMz array is stored like a 3D one with a single owned plane (plane 1 of 3) */
  bool ok = true;
  ok = ok && encode (&m->m[0], hy, sb, xs, ny, 3, (L::TXC + 2 * L::V) * NC, L::TY + 1, err);
  ok = ok && encode (&m->m[1], hx, sb, xs, ny, 3, (L::TXC + L::V) * NC, L::TY + 2, err);
  ok = ok && encode (&m->m[2], ez, sb, xs, ny, 3, (L::TXC + L::V) * NC, L::TY + 1, err);
  ok = ok && encode (&m->m[3], eps, sb, pitchM, ny + 1, 3, L::MW, L::TY + 1, err);
  ok = ok && encode (&m->m[4], dz, sb, dq_pitch * NC, ny, 3, L::QX * NC, L::TY + 1, err); /* compact Dz pair member */
  ok = ok && encode (&m->m[5], bx, sb, xs, ny, 3, L::QX * NC, L::TY, err);
  ok = ok && encode (&m->m[6], by, sb, xs, ny, 3, L::QX * NC, L::TY, err);
  out->valid = ok;
  return ok;
}

template <typename T, int NC, bool TFSF, bool UNIEPS>
cudaError_t launch_fused2_t (const StepParams<T> &p, const Fused2DSet &maps, const FusedBox &fb, cudaStream_t stream)
{
  typedef FTile2<T, NC> L;
  const int smem = L::smem_bytes (fb.nq);
  static int attr_dev[64] = {};
  int dev = 0;
  cudaError_t de = cudaGetDevice (&dev);
  if (de != cudaSuccess) return de;
  if (attr_dev[dev & 63] < smem)
  {
    cudaError_t e = cudaFuncSetAttribute (step_core_tmz_fused<T, NC, TFSF, UNIEPS>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    if (e != cudaSuccess) return e;
    attr_dev[dev & 63] = smem;
  }
  const int ntiles = (fb.fy1 - fb.fy0 + 1) / L::TY;
  if (ntiles <= 0) return cudaSuccess;
  dim3 block (L::NTHREADS);
  dim3 grid ((p.nx + L::TXC - 1) / L::TXC, (ntiles + fb.kc - 1) / fb.kc);
  step_core_tmz_fused<T, NC, TFSF, UNIEPS><<<grid, block, smem, stream>>> (p, maps, fb);
  return cudaGetLastError ();
}

template <typename T, int NC>
cudaError_t launch_fused2_nc (const StepParams<T> &p, const Fused2DSet &set, const FusedBox &fb, bool tfsf, bool uni_eps, cudaStream_t s)
{
  if (tfsf) return uni_eps ? launch_fused2_t<T, NC, true, true> (p, set, fb, s) : launch_fused2_t<T, NC, true, false> (p, set, fb, s);
  return uni_eps ? launch_fused2_t<T, NC, false, true> (p, set, fb, s) : launch_fused2_t<T, NC, false, false> (p, set, fb, s);
}

} /* namespace */

void fused_tile_geometry_tmz (int value_type, int nc, int *tile_cells_x, int *tile_rows, int *smem_per_quarter, int *smem_base)
{
#define GEOM(T, NC) do { typedef FTile2<T, NC> L; if (tile_cells_x) *tile_cells_x = L::TXC; if (tile_rows) *tile_rows = L::TY; \
    if (smem_per_quarter) *smem_per_quarter = L::smem_bytes (1) - L::smem_bytes (0); if (smem_base) *smem_base = L::smem_bytes (0); } while (0)
  if (value_type == 4 && nc == 1) GEOM (float, 1);
  else if (value_type == 4) GEOM (float, 2);
  else if (nc == 1) GEOM (double, 1);
  else GEOM (double, 2);
#undef GEOM
}

bool fused_prepare_tmz (FusedMaps3D *out, int value_type, int nc, void *ez, void *hx, void *hy, void *dz, long long dq_pitch,
                        void *bx, void *by, void *eps, int pitch, int ny, int pitchM, std::string *err)
{
  out->valid = false;
  if (value_type == 4)
    return nc == 2 ? prepare_fused2_t<float, 2> (out, ez, hx, hy, dz, dq_pitch, bx, by, eps, pitch, ny, pitchM, err)
                   : prepare_fused2_t<float, 1> (out, ez, hx, hy, dz, dq_pitch, bx, by, eps, pitch, ny, pitchM, err);
  return nc == 2 ? prepare_fused2_t<double, 2> (out, ez, hx, hy, dz, dq_pitch, bx, by, eps, pitch, ny, pitchM, err)
                 : prepare_fused2_t<double, 1> (out, ez, hx, hy, dz, dq_pitch, bx, by, eps, pitch, ny, pitchM, err);
}

template <typename T>
cudaError_t launch_step_core_tmz_fused (const StepParams<T> &p, const FusedMaps3D &maps, const FusedBox &fb, bool cplx, bool tfsf,
                                        bool uniform_eps, cudaStream_t stream)
{
  const Fused2DSet &set = *reinterpret_cast<const Fused2DSet *> (maps.set);
  if (cplx) return launch_fused2_nc<T, 2> (p, set, fb, tfsf, uniform_eps, stream);
  return launch_fused2_nc<T, 1> (p, set, fb, tfsf, uniform_eps, stream);
}

template cudaError_t launch_step_core_tmz_fused<float> (const StepParams<float> &, const FusedMaps3D &, const FusedBox &, bool, bool, bool, cudaStream_t);
template cudaError_t launch_step_core_tmz_fused<double> (const StepParams<double> &, const FusedMaps3D &, const FusedBox &, bool, bool, bool, cudaStream_t);

} /* namespace fdtd */
