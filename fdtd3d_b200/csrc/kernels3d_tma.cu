/*
 * 3D Yee leapfrog half-step kernels for sm_100a, TMA-staged version (the production path).
 *
 * Same arithmetic and fusion as kernels3d.cu (see there for the reference citations of each formula):
 * one pass per half step updates the three components, their auxiliary D/B fields, E-from-D / H-from-B,
 * the TF/SF correction and the point source, in place.  What differs is how the data reaches the SM.
 *
 * The direct version keeps every in-flight byte in registers (19 x 16-byte loads per thread, > 200
 * registers, one 8-warp CTA per SM) and tops out at about two thirds of the HBM roofline because too few
 * bytes are in flight (profiles/r1_direct_*).  Here a CTA owns an (x, y) tile of 32*V x TY cells and marches
 * along z; for each plane ONE elected thread issues one TMA box copy (cp.async.bulk.tensor.3d) per array
 * into a ring of shared-memory stages, each stage completing on its own mbarrier.  The boxes of the arrays
 * whose neighbours the stencil needs carry a one-vector x halo and / or a one-row y halo, so the y and x
 * neighbours are read from shared memory and no value is fetched twice by a CTA; the z neighbour is carried
 * in registers from plane to plane.  All stages but the one being consumed are in flight, independent of
 * register pressure.  Out-of-range parts of a box (grid edges) are zero-filled by the TMA unit; results are
 * written back with coalesced 128-bit stores.  Tile height TY, ring depth and CTAs per SM depend on the
 * variant (TileRows / Tile below): 8 rows x 4 stages x 1 CTA for the HBM-bound plain fast variants, 4 rows x
 * 3 stages x 3 CTAs for the CORE and EXACT variants, 16 rows x 2 stages for real float.
 *
 * Tile geometry (V = cells per 16 bytes = 16 / sizeof(FieldValue)), written for TY = 8:
 *   thread (tx, ty) owns cells [i0t + tx*V, +V) of row j0t + ty;  TXC = 32*V cells = 512 bytes per row.
 *   E phase stage for plane k:  Hz  (TXC+V) x 9 from (i0t-V, j0t-1)   Hy  (TXC+V) x 8 from (i0t-V, j0t)
 *                               Hx  TXC x 9 from (i0t, j0t-1)         D*, E*  TXC x 8
 *                               eps (TXC+HM) x 9 of material plane k+1 from (i0t, j0t)
 *   H phase stage for plane k:  Ex  TXC x 9 of plane k+1              Ey  (TXC+V) x 8 of plane k+1
 *                               Ez  (TXC+V) x 9 of plane k            B*, H*  TXC x 8
 *                               mu  (TXC+HM) x 9 of material plane k+1
 *
 * CORE variants (fast mode with UPML): where all three sigma are zero -- everything but the absorbing shell -- the
 * D/B recurrences collapse to E += (dt/dx) / (eps eps0) * curl, so those cells neither read nor write D/B
 * (DESIGN.md "sigma-free core"): 18 instead of 30 field values per cell-update.  The decision is made per plane
 * and per quarter of a tile row (8 lanes = 128 bytes): D/B travel as quarter-width boxes and only the quarters
 * that touch the shell are requested; the solver keeps D - eps eps0 E of the core cells in the D arrays between
 * runs (fold / unfold kernels in aux_kernels.cu), so D/B downloads are unaffected.
 */
#include <cuda.h>

#include <string>

#include "fdtd_device.cuh"
#include "kernels3d_common.cuh"
#include "launch.h"
#include "tma_util.cuh"

namespace fdtd
{

namespace
{

using namespace k3d;
using namespace tma;

/* tuning knobs: rows per CTA tile (0 = by value type, see Tile), ring depth (0 = as many as fit in 200 KB, at
 * most 4), CTAs per SM the register allocator must allow */
#ifndef FDTD_TMA_TY
#define FDTD_TMA_TY 0
#endif
#ifndef FDTD_TMA_STAGES
#define FDTD_TMA_STAGES 0
#endif
#ifndef FDTD_TMA_MINB
#define FDTD_TMA_MINB 1
#endif
/* the same three for the CORE and EXACT variants ("ALT" configuration; 0 = the defaults of TileRows / Tile) */
#ifndef FDTD_TMA_ALT_TY
#define FDTD_TMA_ALT_TY 0
#endif
#ifndef FDTD_TMA_ALT_STAGES
#define FDTD_TMA_ALT_STAGES 0
#endif
#ifndef FDTD_TMA_ALT_MINB
#define FDTD_TMA_ALT_MINB 0
#endif
constexpr int N_MAPS = 13; /* 10 full boxes + 3 quarter-width boxes of D / B (CORE variants) */

struct TmaSet
{
  CUtensorMap m[N_MAPS];
};


/*
 * Rows per CTA tile = warps per CTA, and with it the occupancy recipe.  Two configurations per value type:
 *
 *  - the plain fast variants (D/B through every cell) are HBM-bound with ONE CTA per SM and a deep ring: 8 rows x 4
 *    stages; real float, which packs four cells into a thread's 16 bytes (twice the arithmetic per byte) and was
 *    latency-bound with 8 warps (0.88 of the HBM peak), 16 rows x 2 stages = 512 threads at <= 128 registers (0.99;
 *    profiles/r1_tma_variants.log, r1_tma_variants_float.log);
 *  - the CORE variants move a third fewer bytes per plane with the same per-plane latency chain, and the EXACT variants
 *    are latency-bound by their IEEE divisions: both want more CTAs per SM.  ALT = 4 rows x 3 stages x THREE CTAs
 *    per SM (384 threads at up to 168 registers: no spills, 134 KB of ring per SM in flight) -- 512^3 double:
 *    29.0 (8 rows, 1 CTA) -> 31.4 (8 rows, 2 CTAs x 2 stages, 128 registers with spills) -> 33.9 Gcell/s
 *    (profiles/r1_core_variants.log).  Real float keeps its 16-row tiles (no difference measured).
 *
 * The tensor maps are encoded for the configuration in use (tma_prepare_3d (..., alt)).
 */
template <typename T, int NC, bool ALT>
struct TileRows
{
  static constexpr bool F1 = sizeof (T) == 4 && NC == 1;
  static constexpr int BASE = FDTD_TMA_TY > 0 ? FDTD_TMA_TY : (F1 ? 16 : 8);
  static constexpr int ALTV = FDTD_TMA_ALT_TY > 0 ? FDTD_TMA_ALT_TY : (F1 ? BASE : 4);
  static constexpr int value = ALT ? ALTV : BASE;
};

template <typename T, int NC, bool PML, bool ALT>
struct Tile
{
  static constexpr int TY = TileRows<T, NC, ALT>::value;
  static constexpr int NTHREADS = 32 * TY;
  static constexpr int CELL = NC * (int) sizeof (T);
  static constexpr int V = 16 / CELL;                 /* cells per thread = x halo cells */
  static constexpr int TXC = 32 * V;                  /* tile cells along x */
  static constexpr int HM = 16 / (int) sizeof (T);    /* material halo values */
  static constexpr int ROW = TXC * CELL;              /* 512 bytes */
  static constexpr int ROW_H = (TXC + V) * CELL;      /* row with x halo: 528 bytes */
  static constexpr int ROW_M = (TXC + HM) * (int) sizeof (T);
  static constexpr int N_OWN = PML ? 6 : 3;
  /* box bytes */
  static constexpr int SZ_A = ROW_H * (TY + 1);       /* x and y halo */
  static constexpr int SZ_B = ROW_H * TY;             /* x halo */
  static constexpr int SZ_C = ROW * (TY + 1);         /* y halo */
  static constexpr int SZ_OWN = ROW * TY;
  static constexpr int SZ_M = ROW_M * (TY + 1);
  static constexpr int OWN_STRIDE = (SZ_OWN + 127) & ~127; /* between the own-cell tiles of a stage */
  /* stage layout */
  static constexpr int OFF_A = 0;
  static constexpr int OFF_B = OFF_A + align128 (SZ_A);
  static constexpr int OFF_C = OFF_B + align128 (SZ_B);
  static constexpr int OFF_OWN = OFF_C + align128 (SZ_C);
  static constexpr int OFF_M = OFF_OWN + N_OWN * align128 (SZ_OWN);
  static constexpr int STAGE = OFF_M + align128 (SZ_M);
  static constexpr int TX_BYTES = SZ_A + SZ_B + SZ_C + N_OWN * SZ_OWN + SZ_M;
  /* CORE variants: D / B staged as four quarter-width boxes (8 lanes x 16 bytes = 128-byte rows) per array */
  static constexpr int QX = 8 * V;                    /* cells per quarter row */
  static constexpr int SZ_Q = 128 * TY;
  static constexpr int TX_BASE = SZ_A + SZ_B + SZ_C + 3 * SZ_OWN + SZ_M;
  static constexpr int BASE_STAGES = FDTD_TMA_STAGES > 0 ? FDTD_TMA_STAGES : ((4 * STAGE <= 200 * 1024) ? 4 : ((3 * STAGE <= 225 * 1024) ? 3 : 2));
  static constexpr int ALT_STAGES = FDTD_TMA_ALT_STAGES > 0 ? FDTD_TMA_ALT_STAGES : (TY <= 4 ? 3 : (TY == 8 ? 2 : BASE_STAGES));
  static constexpr int ALT_MINB = FDTD_TMA_ALT_MINB > 0 ? FDTD_TMA_ALT_MINB : (TY <= 4 ? 3 : (TY == 8 ? 2 : FDTD_TMA_MINB));
  static constexpr int STAGES = ALT ? ALT_STAGES : BASE_STAGES;
  static constexpr int MINB = ALT ? ALT_MINB : FDTD_TMA_MINB;
  static constexpr int SMEM = STAGES * STAGE + 128;
};

/*
 * CORE variants: which quarters (8 lanes) of a tile row need their D / B values.  `full`: every quarter that holds
 * grid cells (planes outside the z core, tiles outside the y core); `core`: in planes of the z core, the quarters
 * that are not entirely inside the x-y core.  p.core_lo/hi are aligned to quarters (x) and tile rows (y) by the
 * solver, which folds exactly the cells skipped here.
 */
template <typename L, typename T>
__device__ __forceinline__ void quarter_masks (const StepParams<T> &p, int i0t, int j0t, int &full, int &core)
{
  const bool y_core = j0t >= p.core_lo[1] && j0t + L::TY - 1 <= p.core_hi[1];
  full = 0;
  core = 0;
#pragma unroll
  for (int q = 0; q < 4; ++q)
  {
    const int qa = i0t + q * L::QX, qb = qa + L::QX - 1;
    if (qa >= p.nx) continue;
    full |= 1 << q;
    if (!(y_core && qa >= p.core_lo[0] && qb <= p.core_hi[0])) core |= 1 << q;
  }
}

/*
 * Which tile and which planes this CTA works on: the whole slab in a 3-D grid (x tiles, tile rows, plane chunks), or
 * one of the sub-boxes of p.reg in a flat grid (StepParams::reg).  Returns false if the CTA has nothing to do.
 */
template <typename L, typename T>
__device__ __forceinline__ bool cta_work (const StepParams<T> &p, int kb, int ke, int &i0t, int &j0t, int &k0, int &k1)
{
  if (p.nreg == 0)
  {
    i0t = blockIdx.x * L::TXC;
    j0t = blockIdx.y * L::TY;
    k0 = kb + blockIdx.z * p.kchunk;
    k1 = min (k0 + p.kchunk, ke);
    return k0 < k1;
  }
  int r = 0;
#pragma unroll
  for (int q = 1; q < 5; ++q)
    if (q < p.nreg && (int) blockIdx.x >= p.reg[q].cta0) r = q;
  const int ntx = (p.nx + L::TXC - 1) / L::TXC;
  int c = blockIdx.x - p.reg[r].cta0;
  const int bx = c % ntx;
  c /= ntx;
  const int by = c % p.reg[r].njt, bz = c / p.reg[r].njt;
  i0t = bx * L::TXC;
  j0t = (p.reg[r].jt0 + by) * L::TY;
  k0 = p.reg[r].kb + bz * p.reg[r].kc;
  k1 = min (k0 + p.reg[r].kc, p.reg[r].ke);
  return k0 < k1;
}

/* ------------------------------------------------------------------ E half step */

template <typename T, int NC, bool PML, bool TFSF, bool EXACT, bool CORE>
__global__ void __launch_bounds__ ((Tile<T, NC, PML, (CORE || EXACT)>::NTHREADS), (Tile<T, NC, PML, (CORE || EXACT)>::MINB))
e_phase_3d_tma (const __grid_constant__ StepParams<T> p, const __grid_constant__ TmaSet maps, int kb, int ke)
{
  static_assert (!CORE || (PML && !EXACT), "the sigma-free core is a fast-mode UPML optimisation");
  typedef Tile<T, NC, PML, (CORE || EXACT)> L;
  constexpr int V = L::V, W = V * NC, S = L::STAGES;
  typedef Vec<T, W> FVec;
  typedef Vec<T, V> MVec;

  extern __shared__ unsigned char smem_raw[];
  __shared__ __align__ (8) unsigned long long full[S];
  unsigned char *smem = reinterpret_cast<unsigned char *> ((reinterpret_cast<uintptr_t> (smem_raw) + 127) & ~(uintptr_t) 127);

  const int tx = threadIdx.x, ty = threadIdx.y;
  const bool leader = tx == 0 && ty == 0;
  int i0t, j0t, k0, k1;
  if (!cta_work<L> (p, kb, ke, i0t, j0t, k0, k1)) return; /* uniform over the CTA */
  const int i0 = i0t + tx * V, j = j0t + ty;
  const bool active = i0 < p.nx && j < p.ny;

  /* CORE: quarters of the tile row whose D values are needed -- all that hold cells in planes outside the z core
   * or in tiles outside the y core, else only those that stick out of the x core */
  int qmask_full = 0, qmask_core = 0;
  if (CORE) quarter_masks<L> (p, i0t, j0t, qmask_full, qmask_core);
  auto d_mask = [&] (int k) -> int
  {
    const int g = p.kz0 + k;
    return (g >= p.core_lo[2] && g <= p.core_hi[2]) ? qmask_core : qmask_full;
  };

  /* maps: 0 Hz (A), 1 Hy (B), 2 Hx (C), 3..8 Ex Ey Ez [Dx Dy Dz], 9 eps (M), 10..12 Dx Dy Dz as quarter boxes */
  auto issue = [&] (int slot, int k)
  {
    unsigned char *base = smem + slot * L::STAGE;
    unsigned long long *bar = &full[slot];
    const int m = CORE ? d_mask (k) : 0;
    mbar_expect_tx (bar, CORE ? L::TX_BASE + __popc (m) * 3 * L::SZ_Q : L::TX_BYTES);
    tma_load_3d (base + L::OFF_A, &maps.m[0], bar, (i0t - V) * NC, j0t - 1, k + 1);
    tma_load_3d (base + L::OFF_B, &maps.m[1], bar, (i0t - V) * NC, j0t, k + 1);
    tma_load_3d (base + L::OFF_C, &maps.m[2], bar, i0t * NC, j0t - 1, k + 1);
#pragma unroll
    for (int a = 0; a < (CORE ? 3 : L::N_OWN); ++a)
      tma_load_3d (base + L::OFF_OWN + a * L::OWN_STRIDE, &maps.m[3 + a], bar, i0t * NC, j0t, k + 1);
    if (CORE)
    {
#pragma unroll
      for (int a = 0; a < 3; ++a)
#pragma unroll
        for (int q = 0; q < 4; ++q)
          if ((m >> q) & 1)
            tma_load_3d (base + L::OFF_OWN + (3 + a) * L::OWN_STRIDE + q * L::SZ_Q, &maps.m[10 + a], bar,
                         (i0t + q * L::QX) * NC, j0t, k + 1);
    }
    tma_load_3d (base + L::OFF_M, &maps.m[9], bar, i0t, j0t, k + 2);
  };

  if (leader)
  {
    for (int s = 0; s < S; ++s) mbar_init (&full[s], 1);
    asm volatile ("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads ();
  if (leader)
  {
    for (int s = 0; s < S && k0 + s < k1; ++s) issue (s, k0 + s);
  }

  long long o = (((long long) (k0 + 1) * p.ny + j) * p.pitch + i0) * NC;
  const long long pl = p.ps * NC;
  const long long om = ((long long) (k0 + 1) * (p.ny + 1) + j) * p.pitchM + i0;

  T *Ex = p.fo[0], *Ey = p.fo[1], *Ez = p.fo[2]; /* stores only: the operands come from the stage */
  T *Dx = p.f[6], *Dy = p.f[7], *Dz = p.f[8];

  const bool jlo = j >= 1, jhi = j < p.ny - 1;
  const bool j_face = TFSF && tfsf_axis_hit (p, 1, j, true);

  T ca_y_at = 0, cbd_y_at = 0, p_y_btw = 0, m_y_btw = 0, q_y_at = 0;
  MVec p_x_btw = {}, m_x_btw = {}, ca_x_at = {}, cbd_x_at = {}, q_x_at = {};
  FVec hx_km = {}, hy_km = {};
  MVec eps_c = {}, eps_jp = {};
  T eps_ip = 0;
  if (active)
  {
    if (PML)
    {
      ca_y_at = p.tab[1][KIND_AT].ca[j];
      cbd_y_at = p.tab[1][KIND_AT].cbd[j];
      p_y_btw = p.tab[1][KIND_BTW].p[j];
      m_y_btw = p.tab[1][KIND_BTW].m[j];
      q_y_at = EXACT ? p.tab[1][KIND_AT].p[j] : p.tab[1][KIND_AT].invp[j];
      p_x_btw = ldv<T, V> (p.tab[0][KIND_BTW].p + i0);
      m_x_btw = ldv<T, V> (p.tab[0][KIND_BTW].m + i0);
      ca_x_at = ldv<T, V> (p.tab[0][KIND_AT].ca + i0);
      cbd_x_at = ldv<T, V> (p.tab[0][KIND_AT].cbd + i0);
      q_x_at = ldv<T, V> ((EXACT ? p.tab[0][KIND_AT].p : p.tab[0][KIND_AT].invp) + i0);
    }
    /* z-carried values for the first plane of the chunk, straight from global memory */
    hx_km = ldv<T, W> (p.f[3] + o - pl);
    hy_km = ldv<T, W> (p.f[4] + o - pl);
    eps_c = ldv<T, V> (p.eps + om);
    eps_jp = ldv<T, V> (p.eps + om + p.pitchM);
    eps_ip = __ldg (p.eps + om + V);
  }

  long long step = 0;
  if (!TFSF) step = current_step (p);

  /* per-thread byte offsets inside a stage */
  const int xo = tx * 16;
  const unsigned char *sA_own = smem + L::OFF_A + (ty + 1) * L::ROW_H + 16 + xo; /* Hz[j]   */
  const unsigned char *sA_jm = smem + L::OFF_A + ty * L::ROW_H + 16 + xo;        /* Hz[j-1] */
  const unsigned char *sB_own = smem + L::OFF_B + ty * L::ROW_H + 16 + xo;       /* Hy[j]   */
  const unsigned char *sC_own = smem + L::OFF_C + (ty + 1) * L::ROW + xo;        /* Hx[j]   */
  const unsigned char *sC_jm = smem + L::OFF_C + ty * L::ROW + xo;               /* Hx[j-1] */
  const unsigned char *sOwn = smem + L::OFF_OWN + ty * L::ROW + xo;
  /* D tiles: one full-width box, or (CORE) four quarter boxes with 128-byte rows */
  const unsigned char *sAux = CORE ? smem + L::OFF_OWN + (tx >> 3) * L::SZ_Q + ty * 128 + (tx & 7) * 16 : sOwn;
  const unsigned char *sM = smem + L::OFF_M + ty * L::ROW_M + tx * V * (int) sizeof (T);

  for (int k = k0; k < k1; ++k, o += pl)
  {
    const int n = k - k0;
    const int slot = n % S;
    const int so = slot * L::STAGE;
    mbar_wait (&full[slot], (n / S) & 1);

    const int gk = p.kz0 + k;
    const bool klo = gk >= 1, khi = gk < p.nzg - 1;
    const bool jk_face = j_face || (TFSF && tfsf_axis_hit (p, 2, gk, true));
    /* false: this thread's cells are in the sigma-free core in this plane */
    const bool shell = !CORE || ((d_mask (k) >> (tx >> 3)) & 1);

    const FVec hz = lds<T, W> (sA_own + so), hz_jm = lds<T, W> (sA_jm + so);
    const FVec hy = lds<T, W> (sB_own + so);
    const FVec hx = lds<T, W> (sC_own + so), hx_jm = lds<T, W> (sC_jm + so);
    T hz_im[NC], hy_im[NC];
#pragma unroll
    for (int c = 0; c < NC; ++c)
    {
      hz_im[c] = *reinterpret_cast<const T *> (sA_own + so - L::CELL + c * sizeof (T));
      hy_im[c] = *reinterpret_cast<const T *> (sB_own + so - L::CELL + c * sizeof (T));
    }
    FVec ex = lds<T, W> (sOwn + so + 0 * L::OWN_STRIDE);
    FVec ey = lds<T, W> (sOwn + so + 1 * L::OWN_STRIDE);
    FVec ez = lds<T, W> (sOwn + so + 2 * L::OWN_STRIDE);
    FVec dx, dy, dz;
    if (PML && shell)
    {
      dx = lds<T, W> (sAux + so + 3 * L::OWN_STRIDE);
      dy = lds<T, W> (sAux + so + 4 * L::OWN_STRIDE);
      dz = lds<T, W> (sAux + so + 5 * L::OWN_STRIDE);
    }
    const MVec eps_kp = lds<T, V> (sM + so);
    const MVec eps_jp_next = lds<T, V> (sM + so + L::ROW_M);
    const T eps_ip_next = *reinterpret_cast<const T *> (sM + so + V * sizeof (T));

    if (active)
    {
      T ca_z_at = 0, cbd_z_at = 0, q_z_at = 0, p_z_btw = 0, m_z_btw = 0;
      if (PML)
      {
        ca_z_at = p.tab[2][KIND_AT].ca[gk];
        cbd_z_at = p.tab[2][KIND_AT].cbd[gk];
        q_z_at = EXACT ? p.tab[2][KIND_AT].p[gk] : p.tab[2][KIND_AT].invp[gk];
        p_z_btw = p.tab[2][KIND_BTW].p[gk];
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
          const int fx = tfsf_e_faces (p, 0, i, j, gk), fy = tfsf_e_faces (p, 1, i, j, gk), fz = tfsf_e_faces (p, 2, i, j, gk);
          if (fx | fy | fz)
          {
#pragma unroll
            for (int c = 0; c < NC; ++c)
            {
              const int s = v * NC + c;
              const T hz_m = v > 0 ? hz.v[v > 0 ? s - NC : s] : hz_im[c];
              const T hy_m = v > 0 ? hy.v[v > 0 ? s - NC : s] : hy_im[c];
              if (fx) cex[c] = tfsf_curl_tab<T, NC, true> (p, 0, fx, c, i, j, k, hz.v[s], hz_jm.v[s], hy.v[s], hy_km.v[s]);
              if (fy) cey[c] = tfsf_curl_tab<T, NC, true> (p, 1, fy, c, i, j, k, hx.v[s], hx_km.v[s], hz.v[s], hz_m);
              if (fz) cez[c] = tfsf_curl_tab<T, NC, true> (p, 2, fz, c, i, j, k, hy.v[s], hy_m, hx.v[s], hx_jm.v[s]);
            }
          }
        }

        const T e_c = eps_c.v[v];
        const T e_xp = v + 1 < V ? eps_c.v[v + 1 < V ? v + 1 : v] : eps_ip;
        const T mod_x = ((e_c + e_xp) / (T) 2) * p.eps0;
        const T mod_y = ((e_c + eps_jp.v[v]) / (T) 2) * p.eps0;
        const T mod_z = ((e_c + eps_kp.v[v]) / (T) 2) * p.eps0;

        if (PML && shell)
        {
          T cbx, ccx, cby, ccy, cbz, ccz;
          sweep3_coefs<T, EXACT> (p_x_btw.v[v], m_x_btw.v[v], q_z_at, mod_x, cbx, ccx);
          sweep3_coefs<T, EXACT> (p_y_btw, m_y_btw, q_x_at.v[v], mod_y, cby, ccy);
          sweep3_coefs<T, EXACT> (p_z_btw, m_z_btw, q_y_at, mod_z, cbz, ccz);
#pragma unroll
          for (int c = 0; c < NC; ++c)
          {
            const int s = v * NC + c;
            if (do_ex) upml_cell<T, EXACT> (dx.v[s], ex.v[s], cex[c], ca_y_at, cbd_y_at, ca_z_at, cbx, ccx);
            if (do_ey) upml_cell<T, EXACT> (dy.v[s], ey.v[s], cey[c], ca_z_at, cbd_z_at, ca_x_at.v[v], cby, ccy);
            if (do_ez) upml_cell<T, EXACT> (dz.v[s], ez.v[s], cez[c], ca_x_at.v[v], cbd_x_at.v[v], ca_y_at, cbz, ccz);
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
          if (gk == p.srcpos[2] && j == p.srcpos[1] && i == p.srcpos[0])
          {
            /* the source overwrites Ez but not Dz (Scheme3D.cpp:1995-2024): in the core this one cell keeps its
             * true Dz in memory (the fold kernels skip it) and advances it here, Dz += Cb curl with Ca = 1 */
            if (CORE && !shell && do_ez)
            {
#pragma unroll
              for (int c = 0; c < NC; ++c) Dz[o + v * NC + c] = fma_t<T> (cbd_x_at.v[v], cez[c], Dz[o + v * NC + c]);
            }
#pragma unroll
            for (int c = 0; c < NC; ++c) ez.v[v * NC + c] = p.src[step * NC + c];
          }
        }
      }

      stv<T, W> (Ex + o, ex); stv<T, W> (Ey + o, ey); stv<T, W> (Ez + o, ez);
      if (PML && shell)
      {
        stv<T, W> (Dx + o, dx); stv<T, W> (Dy + o, dy); stv<T, W> (Dz + o, dz);
      }
    }

    hx_km = hx; hy_km = hy;
    eps_c = eps_kp; eps_jp = eps_jp_next; eps_ip = eps_ip_next;

    /* every thread has read the stage: hand the slot back to the TMA unit for plane k + S */
    __syncthreads ();
    if (leader && k + S < k1) issue (slot, k + S);
  }
}

/* ------------------------------------------------------------------ H half step */

/*
 * UNIMU: mu holds one value everywhere (p.mu_const; the solver knows because the array was last set with fdtd_fill --
 * the reference itself never initialises mu to anything but 1).  The material box is not loaded (FP bytes less per
 * cell) and the averaged mu -- ((m + m) / 2 + (m + m) / 2) / 2 is m exactly, so also in exact mode -- and everything
 * derived from it alone become loop invariants.
 */
template <typename T, int NC, bool PML, bool TFSF, bool EXACT, bool CORE, bool UNIMU>
__global__ void __launch_bounds__ ((Tile<T, NC, PML, (CORE || EXACT)>::NTHREADS), (Tile<T, NC, PML, (CORE || EXACT)>::MINB))
h_phase_3d_tma (const __grid_constant__ StepParams<T> p, const __grid_constant__ TmaSet maps, int kb, int ke)
{
  static_assert (!CORE || (PML && !EXACT), "the sigma-free core is a fast-mode UPML optimisation");
  typedef Tile<T, NC, PML, (CORE || EXACT)> L;
  constexpr int V = L::V, W = V * NC, S = L::STAGES;
  typedef Vec<T, W> FVec;
  typedef Vec<T, V> MVec;

  extern __shared__ unsigned char smem_raw[];
  __shared__ __align__ (8) unsigned long long full[S];
  unsigned char *smem = reinterpret_cast<unsigned char *> ((reinterpret_cast<uintptr_t> (smem_raw) + 127) & ~(uintptr_t) 127);

  const int tx = threadIdx.x, ty = threadIdx.y;
  const bool leader = tx == 0 && ty == 0;
  int i0t, j0t, k0, k1;
  if (!cta_work<L> (p, kb, ke, i0t, j0t, k0, k1)) return; /* uniform over the CTA */
  const int i0 = i0t + tx * V, j = j0t + ty;
  const bool active = i0 < p.nx && j < p.ny;

  if (blockIdx.x == 0 && blockIdx.y == 0 && blockIdx.z == 0 && leader) advance_step (p); /* nobody reads it in this half step */

  int qmask_full = 0, qmask_core = 0; /* see e_phase_3d_tma */
  if (CORE) quarter_masks<L> (p, i0t, j0t, qmask_full, qmask_core);
  auto d_mask = [&] (int k) -> int
  {
    const int g = p.kz0 + k;
    return (g >= p.core_lo[2] && g <= p.core_hi[2]) ? qmask_core : qmask_full;
  };

  /* maps: 0 Ez (A: x+y halo, plane k), 1 Ey (B: x halo, plane k+1), 2 Ex (C: y halo, plane k+1),
   *       3..8 Hx Hy Hz [Bx By Bz], 9 mu (M, material plane k+1), 10..12 Bx By Bz as quarter boxes */
  auto issue = [&] (int slot, int k)
  {
    unsigned char *base = smem + slot * L::STAGE;
    unsigned long long *bar = &full[slot];
    const int m = CORE ? d_mask (k) : 0;
    mbar_expect_tx (bar, (CORE ? L::TX_BASE + __popc (m) * 3 * L::SZ_Q : L::TX_BYTES) - (UNIMU ? L::SZ_M : 0));
    tma_load_3d (base + L::OFF_A, &maps.m[0], bar, i0t * NC, j0t, k + 1);
    tma_load_3d (base + L::OFF_B, &maps.m[1], bar, i0t * NC, j0t, k + 2);
    tma_load_3d (base + L::OFF_C, &maps.m[2], bar, i0t * NC, j0t, k + 2);
#pragma unroll
    for (int a = 0; a < (CORE ? 3 : L::N_OWN); ++a)
      tma_load_3d (base + L::OFF_OWN + a * L::OWN_STRIDE, &maps.m[3 + a], bar, i0t * NC, j0t, k + 1);
    if (CORE)
    {
#pragma unroll
      for (int a = 0; a < 3; ++a)
#pragma unroll
        for (int q = 0; q < 4; ++q)
          if ((m >> q) & 1)
            tma_load_3d (base + L::OFF_OWN + (3 + a) * L::OWN_STRIDE + q * L::SZ_Q, &maps.m[10 + a], bar,
                         (i0t + q * L::QX) * NC, j0t, k + 1);
    }
    if (!UNIMU) tma_load_3d (base + L::OFF_M, &maps.m[9], bar, i0t, j0t, k + 2);
  };

  if (leader)
  {
    for (int s = 0; s < S; ++s) mbar_init (&full[s], 1);
    asm volatile ("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads ();
  if (leader)
  {
    for (int s = 0; s < S && k0 + s < k1; ++s) issue (s, k0 + s);
  }

  long long o = (((long long) (k0 + 1) * p.ny + j) * p.pitch + i0) * NC;
  const long long row = (long long) p.pitch * NC;
  const long long pl = p.ps * NC;
  const long long om = ((long long) (k0 + 1) * (p.ny + 1) + j) * p.pitchM + i0;

  T *Hx = p.fo[3], *Hy = p.fo[4], *Hz = p.fo[5]; /* stores only */
  T *Bx = p.f[9], *By = p.f[10], *Bz = p.f[11];

  const bool jlo = j >= 1, jhi = j < p.ny - 1;
  const bool j_face = TFSF && tfsf_axis_hit (p, 1, j, false);

  T ca_y_btw = 0, cbd_y_btw = 0, p_y_at = 0, m_y_at = 0, q_y_btw = 0;
  MVec p_x_at = {}, m_x_at = {}, ca_x_btw = {}, cbd_x_btw = {}, q_x_btw = {};
  FVec ex = {}, ex_jp = {}, ey = {};
  T ey_ip[NC] = {};
  MVec mu_c = {}, mu_jp = {};
  T mu_ip = 0, mu_ip_jp = 0;
  if (active)
  {
    if (PML)
    {
      ca_y_btw = p.tab[1][KIND_BTW].ca[j];
      cbd_y_btw = p.tab[1][KIND_BTW].cbd[j];
      p_y_at = p.tab[1][KIND_AT].p[j];
      m_y_at = p.tab[1][KIND_AT].m[j];
      q_y_btw = EXACT ? p.tab[1][KIND_BTW].p[j] : p.tab[1][KIND_BTW].invp[j];
      p_x_at = ldv<T, V> (p.tab[0][KIND_AT].p + i0);
      m_x_at = ldv<T, V> (p.tab[0][KIND_AT].m + i0);
      ca_x_btw = ldv<T, V> (p.tab[0][KIND_BTW].ca + i0);
      cbd_x_btw = ldv<T, V> (p.tab[0][KIND_BTW].cbd + i0);
      q_x_btw = ldv<T, V> ((EXACT ? p.tab[0][KIND_BTW].p : p.tab[0][KIND_BTW].invp) + i0);
    }
    /* z-carried values of plane k0, straight from global memory */
    ex = ldv<T, W> (p.f[0] + o);
    ex_jp = ldv<T, W> (p.f[0] + o + row);
    ey = ldv<T, W> (p.f[1] + o);
#pragma unroll
    for (int c = 0; c < NC; ++c) ey_ip[c] = __ldg (p.f[1] + o + W + c);
    if (!UNIMU)
    {
      mu_c = ldv<T, V> (p.mu + om);
      mu_jp = ldv<T, V> (p.mu + om + p.pitchM);
      mu_ip = __ldg (p.mu + om + V);
      mu_ip_jp = __ldg (p.mu + om + p.pitchM + V);
    }
  }

  const int xo = tx * 16;
  const unsigned char *sA_own = smem + L::OFF_A + ty * L::ROW_H + xo;       /* Ez[j]      */
  const unsigned char *sA_jp = smem + L::OFF_A + (ty + 1) * L::ROW_H + xo;  /* Ez[j+1]    */
  const unsigned char *sB_own = smem + L::OFF_B + ty * L::ROW_H + xo;       /* Ey[k+1]    */
  const unsigned char *sC_own = smem + L::OFF_C + ty * L::ROW + xo;         /* Ex[k+1]    */
  const unsigned char *sC_jp = smem + L::OFF_C + (ty + 1) * L::ROW + xo;    /* Ex[k+1][j+1] */
  const unsigned char *sOwn = smem + L::OFF_OWN + ty * L::ROW + xo;
  const unsigned char *sAux = CORE ? smem + L::OFF_OWN + (tx >> 3) * L::SZ_Q + ty * 128 + (tx & 7) * 16 : sOwn;
  const unsigned char *sM = smem + L::OFF_M + ty * L::ROW_M + tx * V * (int) sizeof (T);

  for (int k = k0; k < k1; ++k, o += pl)
  {
    const int n = k - k0;
    const int slot = n % S;
    const int so = slot * L::STAGE;
    mbar_wait (&full[slot], (n / S) & 1);

    const int gk = p.kz0 + k;
    const bool klo = gk >= 1, khi = gk < p.nzg - 1;
    const bool jk_face = j_face || (TFSF && tfsf_axis_hit (p, 2, gk, false));
    const bool shell = !CORE || ((d_mask (k) >> (tx >> 3)) & 1);

    const FVec ez = lds<T, W> (sA_own + so), ez_jp = lds<T, W> (sA_jp + so);
    const FVec ey_kp = lds<T, W> (sB_own + so);
    const FVec ex_kp = lds<T, W> (sC_own + so), ex_kp_jp = lds<T, W> (sC_jp + so);
    T ez_ip[NC], ey_kp_ip[NC];
#pragma unroll
    for (int c = 0; c < NC; ++c)
    {
      ez_ip[c] = *reinterpret_cast<const T *> (sA_own + so + 16 + c * sizeof (T));
      ey_kp_ip[c] = *reinterpret_cast<const T *> (sB_own + so + 16 + c * sizeof (T));
    }
    FVec hx = lds<T, W> (sOwn + so + 0 * L::OWN_STRIDE);
    FVec hy = lds<T, W> (sOwn + so + 1 * L::OWN_STRIDE);
    FVec hz = lds<T, W> (sOwn + so + 2 * L::OWN_STRIDE);
    FVec bx, by, bz;
    if (PML && shell)
    {
      bx = lds<T, W> (sAux + so + 3 * L::OWN_STRIDE);
      by = lds<T, W> (sAux + so + 4 * L::OWN_STRIDE);
      bz = lds<T, W> (sAux + so + 5 * L::OWN_STRIDE);
    }
    MVec mu_kp, mu_jp_kp;
    T mu_ip_kp, mu_ip_jp_kp;
    if (UNIMU)
    {
#pragma unroll
      for (int v = 0; v < V; ++v) mu_kp.v[v] = mu_jp_kp.v[v] = mu_c.v[v] = mu_jp.v[v] = p.mu_const;
      mu_ip_kp = mu_ip_jp_kp = mu_ip = mu_ip_jp = p.mu_const;
    }
    else
    {
      mu_kp = lds<T, V> (sM + so);
      mu_jp_kp = lds<T, V> (sM + so + L::ROW_M);
      mu_ip_kp = *reinterpret_cast<const T *> (sM + so + V * sizeof (T));
      mu_ip_jp_kp = *reinterpret_cast<const T *> (sM + so + L::ROW_M + V * sizeof (T));
    }

    if (active)
    {
      T ca_z_btw = 0, cbd_z_btw = 0, q_z_btw = 0, p_z_at = 0, m_z_at = 0;
      if (PML)
      {
        ca_z_btw = p.tab[2][KIND_BTW].ca[gk];
        cbd_z_btw = p.tab[2][KIND_BTW].cbd[gk];
        q_z_btw = EXACT ? p.tab[2][KIND_BTW].p[gk] : p.tab[2][KIND_BTW].invp[gk];
        p_z_at = p.tab[2][KIND_AT].p[gk];
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
              if (fx) chx[c] = tfsf_curl_tab<T, NC, false> (p, 0, fx, c, i, j, k, ey_kp.v[s], ey.v[s], ez_jp.v[s], ez.v[s]);
              if (fy) chy[c] = tfsf_curl_tab<T, NC, false> (p, 1, fy, c, i, j, k, ez_p, ez.v[s], ex_kp.v[s], ex.v[s]);
              if (fz) chz[c] = tfsf_curl_tab<T, NC, false> (p, 2, fz, c, i, j, k, ex_jp.v[s], ex.v[s], ey_p, ey.v[s]);
            }
          }
        }

        const T m000 = mu_c.v[v], m001 = mu_kp.v[v], m010 = mu_jp.v[v], m011 = mu_jp_kp.v[v];
        const T m100 = v + 1 < V ? mu_c.v[v + 1 < V ? v + 1 : v] : mu_ip;
        const T m101 = v + 1 < V ? mu_kp.v[v + 1 < V ? v + 1 : v] : mu_ip_kp;
        const T m110 = v + 1 < V ? mu_jp.v[v + 1 < V ? v + 1 : v] : mu_ip_jp;
        const T two = (T) 2;
        const T mod_x = (((m000 + m001) / two + (m010 + m011) / two) / two) * p.mu0;
        const T mod_y = (((m000 + m001) / two + (m100 + m101) / two) / two) * p.mu0;
        const T mod_z = (((m000 + m010) / two + (m100 + m110) / two) / two) * p.mu0;

        if (PML && shell)
        {
          T cbx, ccx, cby, ccy, cbz, ccz;
          sweep3_coefs<T, EXACT> (p_x_at.v[v], m_x_at.v[v], q_z_btw, mod_x, cbx, ccx);
          sweep3_coefs<T, EXACT> (p_y_at, m_y_at, q_x_btw.v[v], mod_y, cby, ccy);
          sweep3_coefs<T, EXACT> (p_z_at, m_z_at, q_y_btw, mod_z, cbz, ccz);
#pragma unroll
          for (int c = 0; c < NC; ++c)
          {
            const int s = v * NC + c;
            if (do_hx) upml_cell<T, EXACT> (bx.v[s], hx.v[s], chx[c], ca_y_btw, cbd_y_btw, ca_z_btw, cbx, ccx);
            if (do_hy) upml_cell<T, EXACT> (by.v[s], hy.v[s], chy[c], ca_z_btw, cbd_z_btw, ca_x_btw.v[v], cby, ccy);
            if (do_hz) upml_cell<T, EXACT> (bz.v[s], hz.v[s], chz[c], ca_x_btw.v[v], cbd_x_btw.v[v], ca_y_btw, cbz, ccz);
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
      if (PML && shell)
      {
        stv<T, W> (Bx + o, bx); stv<T, W> (By + o, by); stv<T, W> (Bz + o, bz);
      }
    }

    ex = ex_kp; ex_jp = ex_kp_jp; ey = ey_kp;
#pragma unroll
    for (int c = 0; c < NC; ++c) ey_ip[c] = ey_kp_ip[c];
    mu_c = mu_kp; mu_jp = mu_jp_kp; mu_ip = mu_ip_kp; mu_ip_jp = mu_ip_jp_kp;

    __syncthreads ();
    if (leader && k + S < k1) issue (slot, k + S);
  }
}

/* ------------------------------------------------------------------ dispatch */

template <typename T, int NC, bool PML, bool TFSF, bool EXACT, bool CORE, bool UNIMU>
cudaError_t launch_h (const StepParams<T> &p, const TmaSet &maps, dim3 grid, dim3 block, int smem, int kb, int ke, cudaStream_t stream)
{
  /* the opt-in to > 48 KB of dynamic shared memory is a per-device function attribute */
  static bool attr_dev[64] = {};
  int dev = 0;
  cudaError_t de = cudaGetDevice (&dev);
  if (de != cudaSuccess) return de;
  if (!attr_dev[dev & 63])
  {
    cudaError_t e = cudaFuncSetAttribute (h_phase_3d_tma<T, NC, PML, TFSF, EXACT, CORE, UNIMU>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    if (e != cudaSuccess) return e;
    attr_dev[dev & 63] = true;
  }
  h_phase_3d_tma<T, NC, PML, TFSF, EXACT, CORE, UNIMU><<<grid, block, smem, stream>>> (p, maps, kb, ke);
  return cudaGetLastError ();
}

template <typename T, int NC, bool PML, bool TFSF, bool EXACT, bool CORE = false>
cudaError_t launch_t (bool e_phase, bool unimu, const StepParams<T> &p, const TmaSet &maps, int kb, int ke, cudaStream_t stream)
{
  typedef Tile<T, NC, PML, (CORE || EXACT)> L;
  constexpr int smem = L::SMEM;
  dim3 block (32, L::TY);
  dim3 grid ((p.nx + L::TXC - 1) / L::TXC, (p.ny + L::TY - 1) / L::TY, (ke - kb + p.kchunk - 1) / p.kchunk);
  StepParams<T> pr = p;
  if (p.nreg > 0)
  {
    /* flat grid over the sub-boxes: region r owns CTAs [cta0 (r), cta0 (r + 1)) */
    int total = 0;
    for (int r = 0; r < p.nreg; ++r)
    {
      pr.reg[r].cta0 = total;
      const int planes = p.reg[r].ke - p.reg[r].kb;
      if (planes > 0 && p.reg[r].njt > 0) total += (int) grid.x * p.reg[r].njt * ((planes + p.reg[r].kc - 1) / p.reg[r].kc);
    }
    if (total == 0) return cudaSuccess;
    grid = dim3 (total, 1, 1);
  }
  else if (ke <= kb) return cudaSuccess;
  if (!e_phase)
    return unimu ? launch_h<T, NC, PML, TFSF, EXACT, CORE, true> (pr, maps, grid, block, smem, kb, ke, stream)
                 : launch_h<T, NC, PML, TFSF, EXACT, CORE, false> (pr, maps, grid, block, smem, kb, ke, stream);
  /* the opt-in to > 48 KB of dynamic shared memory is a per-device function attribute */
  static bool attr_e_dev[64] = {};
  int dev = 0;
  cudaError_t de = cudaGetDevice (&dev);
  if (de != cudaSuccess) return de;
  if (!attr_e_dev[dev & 63])
  {
    cudaError_t e = cudaFuncSetAttribute (e_phase_3d_tma<T, NC, PML, TFSF, EXACT, CORE>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    if (e != cudaSuccess) return e;
    attr_e_dev[dev & 63] = true;
  }
  e_phase_3d_tma<T, NC, PML, TFSF, EXACT, CORE><<<grid, block, smem, stream>>> (pr, maps, kb, ke);
  return cudaGetLastError ();
}

template <typename T, int NC>
cudaError_t launch_nc (bool e_phase, bool unimu, const StepParams<T> &p, const TmaSet &maps, bool pml, bool tfsf, bool exact,
                       bool core, int kb, int ke, cudaStream_t s)
{
  if (core && pml && !exact)
    return tfsf ? launch_t<T, NC, true, true, false, true> (e_phase, unimu, p, maps, kb, ke, s)
                : launch_t<T, NC, true, false, false, true> (e_phase, unimu, p, maps, kb, ke, s);
  const int key = (pml ? 4 : 0) | (tfsf ? 2 : 0) | (exact ? 1 : 0);
  switch (key)
  {
    case 0: return launch_t<T, NC, false, false, false> (e_phase, unimu, p, maps, kb, ke, s);
    case 1: return launch_t<T, NC, false, false, true> (e_phase, unimu, p, maps, kb, ke, s);
    case 2: return launch_t<T, NC, false, true, false> (e_phase, unimu, p, maps, kb, ke, s);
    case 3: return launch_t<T, NC, false, true, true> (e_phase, unimu, p, maps, kb, ke, s);
    case 4: return launch_t<T, NC, true, false, false> (e_phase, unimu, p, maps, kb, ke, s);
    case 5: return launch_t<T, NC, true, false, true> (e_phase, unimu, p, maps, kb, ke, s);
    case 6: return launch_t<T, NC, true, true, false> (e_phase, unimu, p, maps, kb, ke, s);
    default: return launch_t<T, NC, true, true, true> (e_phase, unimu, p, maps, kb, ke, s);
  }
}

/* ------------------------------------------------------------------ tensor maps (host) */

template <typename T, int NC, bool ALT>
bool prepare_tt (TmaMaps3D *out, bool pml, void *const f[12], void *eps, void *mu, int pitch, int ny, int planes, int pitchM,
                int planesM, std::string *err)
{
  typedef Tile<T, NC, true, ALT> L; /* box shapes do not depend on PML */
  static_assert (sizeof (TmaSet) <= sizeof (out->e), "TmaMaps3D too small");
  TmaSet *e = reinterpret_cast<TmaSet *> (out->e), *h = reinterpret_cast<TmaSet *> (out->h);
  const int sb = (int) sizeof (T);
  const long long xs = (long long) pitch * NC;
  const int bxh = (L::TXC + L::V) * NC, bx = L::TXC * NC, bxm = L::TXC + L::HM;
  bool ok = true;
  /* E phase: Hz (A), Hy (B), Hx (C), Ex Ey Ez Dx Dy Dz, eps */
  ok = ok && encode (&e->m[0], f[5], sb, xs, ny, planes, bxh, L::TY + 1, err);
  ok = ok && encode (&e->m[1], f[4], sb, xs, ny, planes, bxh, L::TY, err);
  ok = ok && encode (&e->m[2], f[3], sb, xs, ny, planes, bx, L::TY + 1, err);
  for (int a = 0; a < 3 && ok; ++a) ok = encode (&e->m[3 + a], f[0 + a], sb, xs, ny, planes, bx, L::TY, err);
  for (int a = 0; a < 3 && ok && pml; ++a) ok = encode (&e->m[6 + a], f[6 + a], sb, xs, ny, planes, bx, L::TY, err);
  for (int a = 0; a < 3 && ok && pml; ++a) ok = encode (&e->m[10 + a], f[6 + a], sb, xs, ny, planes, bx / 4, L::TY, err);
  ok = ok && encode (&e->m[9], eps, sb, pitchM, ny + 1, planesM, bxm, L::TY + 1, err);
  /* H phase: Ez (A), Ey (B), Ex (C), Hx Hy Hz Bx By Bz, mu */
  ok = ok && encode (&h->m[0], f[2], sb, xs, ny, planes, bxh, L::TY + 1, err);
  ok = ok && encode (&h->m[1], f[1], sb, xs, ny, planes, bxh, L::TY, err);
  ok = ok && encode (&h->m[2], f[0], sb, xs, ny, planes, bx, L::TY + 1, err);
  for (int a = 0; a < 3 && ok; ++a) ok = encode (&h->m[3 + a], f[3 + a], sb, xs, ny, planes, bx, L::TY, err);
  for (int a = 0; a < 3 && ok && pml; ++a) ok = encode (&h->m[6 + a], f[9 + a], sb, xs, ny, planes, bx, L::TY, err);
  for (int a = 0; a < 3 && ok && pml; ++a) ok = encode (&h->m[10 + a], f[9 + a], sb, xs, ny, planes, bx / 4, L::TY, err);
  ok = ok && encode (&h->m[9], mu, sb, pitchM, ny + 1, planesM, bxm, L::TY + 1, err);
  out->valid = ok;
  return ok;
}

template <typename T, int NC>
bool prepare_t (TmaMaps3D *out, bool pml, bool alt, void *const f[12], void *eps, void *mu, int pitch, int ny, int planes,
                int pitchM, int planesM, std::string *err)
{
  return alt ? prepare_tt<T, NC, true> (out, pml, f, eps, mu, pitch, ny, planes, pitchM, planesM, err)
             : prepare_tt<T, NC, false> (out, pml, f, eps, mu, pitch, ny, planes, pitchM, planesM, err);
}

template <typename T, int NC>
void geometry_t (bool alt, int &txc, int &ty, int &qx, int &minb)
{
  if (alt) { typedef Tile<T, NC, true, true> L; txc = L::TXC; ty = L::TY; qx = L::QX; minb = L::MINB; }
  else { typedef Tile<T, NC, true, false> L; txc = L::TXC; ty = L::TY; qx = L::QX; minb = L::MINB; }
}

} /* namespace */

bool tma_prepare_3d (TmaMaps3D *out, int value_type, int nc, bool pml, bool alt, void *const f[12], void *eps, void *mu,
                     int pitch, int ny, int planes, int pitchM, int planesM, std::string *err)
{
  out->valid = false;
  if (value_type == 4)
    return nc == 2 ? prepare_t<float, 2> (out, pml, alt, f, eps, mu, pitch, ny, planes, pitchM, planesM, err)
                   : prepare_t<float, 1> (out, pml, alt, f, eps, mu, pitch, ny, planes, pitchM, planesM, err);
  return nc == 2 ? prepare_t<double, 2> (out, pml, alt, f, eps, mu, pitch, ny, planes, pitchM, planesM, err)
                 : prepare_t<double, 1> (out, pml, alt, f, eps, mu, pitch, ny, planes, pitchM, planesM, err);
}

void tma_tile_geometry (int value_type, int nc, bool alt, int *tile_cells_x, int *tile_rows, int *quarter_cells_x,
                        int *ctas_per_sm)
{
  int txc, ty, qx, minb;
  if (value_type == 4 && nc == 1) geometry_t<float, 1> (alt, txc, ty, qx, minb);
  else if (value_type == 4) geometry_t<float, 2> (alt, txc, ty, qx, minb);
  else if (nc == 1) geometry_t<double, 1> (alt, txc, ty, qx, minb);
  else geometry_t<double, 2> (alt, txc, ty, qx, minb);
  if (tile_cells_x) *tile_cells_x = txc;
  if (tile_rows) *tile_rows = ty;
  if (quarter_cells_x) *quarter_cells_x = qx;
  if (ctas_per_sm) *ctas_per_sm = minb;
}

template <typename T>
cudaError_t launch_phase_3d_tma (bool e_phase, const StepParams<T> &p, const TmaMaps3D &maps, bool cplx, bool pml, bool tfsf,
                                 bool exact, bool core, bool uniform_mu, int kb, int ke, cudaStream_t stream)
{
  const TmaSet &set = *reinterpret_cast<const TmaSet *> (e_phase ? maps.e : maps.h);
  if (cplx) return launch_nc<T, 2> (e_phase, uniform_mu, p, set, pml, tfsf, exact, core, kb, ke, stream);
  return launch_nc<T, 1> (e_phase, uniform_mu, p, set, pml, tfsf, exact, core, kb, ke, stream);
}

template cudaError_t launch_phase_3d_tma<float> (bool, const StepParams<float> &, const TmaMaps3D &, bool, bool, bool, bool, bool, bool, int, int, cudaStream_t);
template cudaError_t launch_phase_3d_tma<double> (bool, const StepParams<double> &, const TmaMaps3D &, bool, bool, bool, bool, bool, bool, int, int, cudaStream_t);

} /* namespace fdtd */
