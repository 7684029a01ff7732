/*
 * One pass per time step inside the sigma-free core ("temporal blocking" of the two half steps), sm_100a.
 *
 * The two half-step kernels (kernels3d_tma.cu) move 18 field values per core cell and step: the E pass reads H and
 * reads + writes E, the H pass reads E and reads + writes H.  This kernel reads E and H once and writes E' and H' once:
 * 12 field values (+ eps) per cell and step.  Reference semantics preserved (Scheme3D.cpp:1986-2069): E of step t is
 * computed from H of step t-1, then H of step t from that E; the per-cell arithmetic is the half-step kernels' own
 * (kernels3d_common.cuh), only the schedule differs, so the tolerance of fast mode is unchanged.
 *
 * How the dependency H'(i,j,k) <- E'(i+1,.,.), E'(.,j+1,.), E'(.,.,k+1) is met without a grid-wide barrier:
 *   - a CTA owns an (x, y) tile of TXC x TY cells and marches along z; in iteration k it computes E'(k) and then H'(k-1),
 *     so the z neighbour is simply the plane it has just produced (its own E'(k-1) is carried in registers);
 *   - the x and y neighbours that belong to the next tile are RECOMPUTED: the CTA covers one more 16-byte vector per row
 *     and one more row than its tile, and those rim threads evaluate E' for the first vector / row of the neighbour
 *     tiles from the same staged data (boxes one vector wider and one row higher; thread layout: a warp per row plus
 *     one warp for the rim column).  E' travels between the threads of
 *     a CTA through a double-buffered shared-memory tile; only the tile's own cells are stored;
 *   - a neighbour CTA may run earlier or later, so nobody may overwrite what a rim thread still has to read: E and H
 *     are ping-pong pairs -- this kernel (and, while it is in use, the half-step kernels of the shell) reads buffer
 *     `cur` and writes buffer `nxt` (StepParams::f / fo), and the solver swaps them after every step;
 *   - the planes, rows and cells outside this kernel's box (the absorbing shell in y and z, slab boundary planes of a
 *     multi-GPU run) are advanced by the half-step kernels on sub-boxes: E shell -> this kernel -> H shell.  Where a
 *     rim row or the plane above the last one lies outside the box, E' is read from `nxt` (the E shell pass has
 *     already written it) instead of being recomputed.
 *
 * The box spans the whole x range: the few lanes of the first and last x tile that lie in the x absorber take the
 * UPML path with their D / B values (quarter-row boxes, as in the CORE half-step kernels; their D values form a third
 * ping-pong pair in compact arrays, because a rim thread needs the neighbour cell's old D); everything else is
 * E' = E + (dt/dx)/(eps eps0) curl H, H' = H + (dt/dx)/(mu mu0) curl E'.  mu must be homogeneous (set with fdtd_fill):
 * the reference never initialises it to anything but 1; otherwise the solver keeps the two-pass kernels.
 *
 * Data movement as in kernels3d_tma.cu: per plane one elected thread issues the TMA box copies of a stage
 * (cp.async.bulk.tensor.3d, one mbarrier per stage, ring of S stages); results leave with 128-bit stores.
 *
 * Measured on B200 (profiles/r2_one_pass.md): the kernel moves 0.65 x the bytes of the two half-step kernels but runs
 * the instruction streams of both at the same 1.55 IPC per SM, which makes it latency-bound: 6 % faster than the two
 * passes for complex double (where the solver uses it by default), slower for real double, equal for float (there it
 * is used only on request, fdtd_set_tuning).
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

#ifndef FDTD_FUSED_TY
#define FDTD_FUSED_TY 0
#endif
#ifndef FDTD_FUSED_STAGES
#define FDTD_FUSED_STAGES 0
#endif
#ifndef FDTD_FUSED_UNROLL
#define FDTD_FUSED_UNROLL 1
#endif
#ifndef FDTD_FUSED_MINB
#define FDTD_FUSED_MINB 0
#endif

/* maps: 0 Hz (A) 1 Hy (B) 2 Hx (C) 3..5 Ex Ey Ez 6 eps 7..9 Dx Dy Dz quarter rows (compact pair member being read) 10..12 Bx By Bz quarter rows */
struct FusedSet
{
  CUtensorMap m[13];
};

template <typename T, int NC>
struct FTile
{
  static constexpr bool F1 = sizeof (T) == 4 && NC == 1;
  /* measured on 512^3 double (profiles/r2_one_pass.md): 8 rows x 3 stages x 1 CTA per SM 3.51 ms, 4 rows x 2 CTAs 3.85 ms,
   * 16 rows (register-capped, spills) 4.67 ms, 4 rows x 3 CTAs (spills) 6.8 ms */
  static constexpr int TY = FDTD_FUSED_TY > 0 ? FDTD_FUSED_TY : 8;
  static constexpr int NTHREADS = 32 * (TY + 2);          /* a warp per tile row and rim row + one warp for the x rim column */
  static constexpr int CELL = NC * (int) sizeof (T);
  static constexpr int V = 16 / CELL;
  static constexpr int TXC = 32 * V;
  static constexpr int HM = 16 / (int) sizeof (T);
  static constexpr int QX = 8 * V;
  static constexpr int RW_A = (TXC + 2 * V) * CELL;   /* Hz, Hy rows: one vector of x halo on either side */
  static constexpr int RW_C = (TXC + V) * CELL;       /* Hx, E rows: the rim vector on the high side */
  static constexpr int MW = (TXC + V + 1 + HM - 1) / HM * HM; /* eps values per row */
  static constexpr int RW_M = MW * (int) sizeof (T);
  static constexpr int SZ_A = RW_A * (TY + 2);
  static constexpr int SZ_B = RW_A * (TY + 1);
  static constexpr int SZ_C = RW_C * (TY + 2);
  static constexpr int SZ_E = RW_C * (TY + 1);
  static constexpr int SZ_M = RW_M * (TY + 2);
  static constexpr int SZ_QD = 128 * (TY + 1);        /* D of a quarter row, rim row included */
  static constexpr int SZ_QB = 128 * TY;              /* B of a quarter row */
  static constexpr int E_STRIDE = align128 (SZ_E);
  static constexpr int OFF_A = 0;
  static constexpr int OFF_B = OFF_A + align128 (SZ_A);
  static constexpr int OFF_C = OFF_B + align128 (SZ_B);
  static constexpr int OFF_E = OFF_C + align128 (SZ_C);
  static constexpr int OFF_M = OFF_E + 3 * E_STRIDE;
  static constexpr int OFF_Q = OFF_M + align128 (SZ_M);
  static constexpr int TX_BASE = SZ_A + SZ_B + SZ_C + 3 * SZ_E + SZ_M;
  static constexpr int X_BYTES = 3 * E_STRIDE;        /* one E' exchange buffer */
  static constexpr int STAGES = FDTD_FUSED_STAGES > 0 ? FDTD_FUSED_STAGES : 3;
  static constexpr int MINB = FDTD_FUSED_MINB > 0 ? FDTD_FUSED_MINB : 1;
  __host__ __device__ static constexpr int stage_bytes (int nq) { return OFF_Q + nq * 3 * (SZ_QD + SZ_QB); }
  __host__ __device__ static constexpr int smem_bytes (int nq) { return STAGES * stage_bytes (nq) + 2 * X_BYTES + 128; }
};

/* ------------------------------------------------------------------ the kernel */

/* one cell's scalars (re, im), and the three curls of a cell: passed by value so that nothing has its address taken */
template <typename T, int NC> struct Cell { T v[NC]; };
template <typename T, int NC> struct Curl3 { T x[NC], y[NC], z[NC]; };

/* Rare path, kept out of the main loop: curls of a cell on a TF/SF face (E update) */
/* inlined: as real calls they cost the whole kernel 25 % (ABI register constraints), measured */
#define FDTD_TFSF_FIX_ATTR __forceinline__
template <typename T, int NC>
__device__ FDTD_TFSF_FIX_ATTR Curl3<T, NC> tfsf_fix_e (const StepParams<T> &p, int i, int j, int gk, int k, Cell<T, NC> hz, Cell<T, NC> hz_jm,
                                                 Cell<T, NC> hy, Cell<T, NC> hy_km, Cell<T, NC> hx, Cell<T, NC> hx_km, Cell<T, NC> hx_jm,
                                                 Cell<T, NC> hz_m, Cell<T, NC> hy_m, Curl3<T, NC> cu)
{
  const int fx = tfsf_e_faces (p, 0, i, j, gk), fy = tfsf_e_faces (p, 1, i, j, gk), fz = tfsf_e_faces (p, 2, i, j, gk);
  if (!(fx | fy | fz)) return cu;
#pragma unroll
  for (int c = 0; c < NC; ++c)
  {
    if (fx) cu.x[c] = tfsf_curl_tab<T, NC, true> (p, 0, fx, c, i, j, k, hz.v[c], hz_jm.v[c], hy.v[c], hy_km.v[c]);
    if (fy) cu.y[c] = tfsf_curl_tab<T, NC, true> (p, 1, fy, c, i, j, k, hx.v[c], hx_km.v[c], hz.v[c], hz_m.v[c]);
    if (fz) cu.z[c] = tfsf_curl_tab<T, NC, true> (p, 2, fz, c, i, j, k, hy.v[c], hy_m.v[c], hx.v[c], hx_jm.v[c]);
  }
  return cu;
}

/* ... and of the H update (m = local plane, gm = global plane) */
template <typename T, int NC>
__device__ FDTD_TFSF_FIX_ATTR Curl3<T, NC> tfsf_fix_h (const StepParams<T> &p, int i, int j, int gm, int m, Cell<T, NC> ey_kp, Cell<T, NC> ey,
                                                 Cell<T, NC> ez_jp, Cell<T, NC> ez, Cell<T, NC> ez_p, Cell<T, NC> ex_kp, Cell<T, NC> ex,
                                                 Cell<T, NC> ex_jp, Cell<T, NC> ey_p, Curl3<T, NC> cu)
{
  const int fx = tfsf_h_faces (p, 0, i, j, gm), fy = tfsf_h_faces (p, 1, i, j, gm), fz = tfsf_h_faces (p, 2, i, j, gm);
  if (!(fx | fy | fz)) return cu;
#pragma unroll
  for (int c = 0; c < NC; ++c)
  {
    if (fx) cu.x[c] = tfsf_curl_tab<T, NC, false> (p, 0, fx, c, i, j, m, ey_kp.v[c], ey.v[c], ez_jp.v[c], ez.v[c]);
    if (fy) cu.y[c] = tfsf_curl_tab<T, NC, false> (p, 1, fy, c, i, j, m, ez_p.v[c], ez.v[c], ex_kp.v[c], ex.v[c]);
    if (fz) cu.z[c] = tfsf_curl_tab<T, NC, false> (p, 2, fz, c, i, j, m, ex_jp.v[c], ex.v[c], ey_p.v[c], ey.v[c]);
  }
  return cu;
}

template <typename T, int NC, bool TFSF>
__global__ void __launch_bounds__ ((FTile<T, NC>::NTHREADS), (FTile<T, NC>::MINB))
step_core_3d_fused (const __grid_constant__ StepParams<T> p, const __grid_constant__ FusedSet maps, const FusedBox fb)
{
  typedef FTile<T, NC> L;
  constexpr int V = L::V, W = V * NC, S = L::STAGES, TY = L::TY;
  typedef Vec<T, W> FVec;
  typedef Vec<T, V> MVec;

  /* all shared-memory accesses index smem_raw with integer offsets, so that they compile to shared-space loads /
   * stores with 32-bit addresses; `sm0` makes the first stage 128-byte aligned (TMA destination) */
  extern __shared__ unsigned char smem_raw[];
  __shared__ __align__ (8) unsigned long long full[S];
  const unsigned sbase = smem_u32 (smem_raw);
  const int sm0 = (int) ((128u - (sbase & 127u)) & 127u);
  const int stage_sz = L::stage_bytes (fb.nq);
  const int xbuf0 = sm0 + S * stage_sz; /* two E' exchange buffers */

  /* warps 0 .. TY: lane = x vector of row `warp` (row TY is the y rim); warp TY + 1: the x rim column, lane = row.
   * Row warps read, write and store whole contiguous rows; the rim warp touches one vector per row. */
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const bool valid = warp <= TY || lane <= TY;
  const int lx = warp <= TY ? lane : 32, ly = warp <= TY ? warp : (lane <= TY ? lane : 0);
  const bool leader = threadIdx.x == 0;
  const int i0t = blockIdx.x * L::TXC, j0t = fb.fy0 + blockIdx.y * TY;
  const int k0 = fb.fz0 + blockIdx.z * fb.kc;
  const int k1 = min (k0 + fb.kc, fb.fz1 + 1);   /* E' and H' of planes [k0, k1); iteration k1 only supplies E'(k1) */
  if (k0 >= k1) return;
  const int i0 = i0t + lx * V, j = j0t + ly;
  const bool ingrid = valid && i0 < p.nx && j < p.ny;
  const bool own = lx < 32 && ly < TY && ingrid;  /* a cell of this CTA's tile: stored */
  /* rim row above the last tile row of the box: a shell row, its E' is read from the E shell pass's output */
  const bool rim_from_nxt = ly == TY && j0t + TY > fb.fy1;
  /* the plane above the last one of the box likewise */
  const bool top_from_nxt = k1 > fb.fz1;
  const bool compute_e = ingrid && !rim_from_nxt;

  /* quarters (8 lanes) of the tile row that lie in the x absorber: they carry D / B */
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
  /* compact quarter index of a quarter of this tile in the D ping-pong arrays (launch.h, FusedBox) */
  auto compact_q = [&] (int q) -> int
  {
    const int gq = i0t / L::QX + q;
    return gq < fb.nlo ? gq : fb.nlo + gq - fb.hq0;
  };

  auto issue = [&] (int slot, int k)
  {
    unsigned char *base = smem_raw + sm0 + slot * stage_sz;
    unsigned long long *bar = &full[slot];
    mbar_expect_tx (bar, L::TX_BASE + __popc (qmask) * 3 * (L::SZ_QD + L::SZ_QB));
    tma_load_3d (base + L::OFF_A, &maps.m[0], bar, (i0t - V) * NC, j0t - 1, k + 1);
    tma_load_3d (base + L::OFF_B, &maps.m[1], bar, (i0t - V) * NC, j0t, k + 1);
    tma_load_3d (base + L::OFF_C, &maps.m[2], bar, i0t * NC, j0t - 1, k + 1);
#pragma unroll
    for (int a = 0; a < 3; ++a) tma_load_3d (base + L::OFF_E + a * L::E_STRIDE, &maps.m[3 + a], bar, i0t * NC, j0t, k + 1);
    tma_load_3d (base + L::OFF_M, &maps.m[6], bar, i0t, j0t, k + 2);
    if (qmask)
    {
      int r = 0;
#pragma unroll
      for (int q = 0; q < 4; ++q)
        if ((qmask >> q) & 1)
        {
#pragma unroll
          for (int a = 0; a < 3; ++a)
          {
            tma_load_3d (base + L::OFF_Q + (a * fb.nq + r) * L::SZ_QD, &maps.m[7 + a], bar, compact_q (q) * L::QX * NC, j0t, k + 1);
            /* B of plane k - 1: H'(k-1) is computed in iteration k */
            tma_load_3d (base + L::OFF_Q + 3 * fb.nq * L::SZ_QD + (a * fb.nq + r) * L::SZ_QB, &maps.m[10 + a], bar,
                         (i0t + q * L::QX) * NC, j0t, k);
          }
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
    for (int s = 0; s < S && k0 + s <= k1; ++s) issue (s, k0 + s);
  }

  long long o = (((long long) (k0 + 1) * p.ny + j) * p.pitch + i0) * NC;
  const long long pl = p.ps * NC;
  const long long om = ((long long) (k0 + 1) * (p.ny + 1) + j) * p.pitchM + i0;

  /* what does not change along the march, once: update ranges (YeeGridLayout.h:131-182) and TF/SF face candidates of
   * this thread's cells as bit masks over v; rows likewise */
  int ilo_m = 0, ihi_m = 0, ife_m = 0, ifh_m = 0;
#pragma unroll
  for (int v = 0; v < V; ++v)
  {
    const int i = i0 + v;
    if (i >= 1) ilo_m |= 1 << v;
    if (i < p.nx - 1) ihi_m |= 1 << v;
    if (TFSF && tfsf_axis_hit (p, 0, i, true)) ife_m |= 1 << v;
    if (TFSF && tfsf_axis_hit (p, 0, i, false)) ifh_m |= 1 << v;
  }
  const bool j_face_e = TFSF && tfsf_axis_hit (p, 1, j, true);
  const bool j_face_h = TFSF && tfsf_axis_hit (p, 1, j, false);
  /* the rows and planes the kernel computes lie strictly inside the grid (configure_fused), so of the update ranges
   * E: x needs ihi & jlo & klo, y: ilo & jhi & klo, z: ilo & jlo & khi; H: x: ilo & jhi & khi, y: ihi & jlo & khi,
   * z: ihi & jhi & klo only the x parts can be false, and only in the absorber lanes of the first / last x tile */
  const T eps0 = p.eps0, dt_dx = p.dt_dx;

  /* carried from plane to plane: H of the plane below (operands of E', then the own values H' starts from), E' of the
   * plane below, the material of this plane */
  FVec hx_km = {}, hy_km = {}, hz_km = {};
  FVec exm = {}, eym = {}, ezm = {};
  MVec eps_c = {}, eps_jp = {};
  T eps_ip = 0;
  if (ingrid)
  {
    hx_km = ldv<T, W> (p.f[3] + o - pl);
    hy_km = ldv<T, W> (p.f[4] + o - pl);
    eps_c = ldv<T, V> (p.eps + om);
    eps_jp = ldv<T, V> (p.eps + om + p.pitchM);
    eps_ip = __ldg (p.eps + om + V);
  }
  const T cb_h = dt_dx * rcp_t<T> (p.mu_const * p.mu0); /* homogeneous mu: (m + m) / 2 is m exactly */

  long long step = 0;
  if (!TFSF) step = current_step (p);
  const bool src_row = !TFSF && j == p.srcpos[1] && p.srcpos[0] >= i0 && p.srcpos[0] < i0 + V;

  /* byte offsets of this thread's operands inside a stage / an exchange buffer */
  const int xo = lx * 16;
  const int oA_own = L::OFF_A + (ly + 1) * L::RW_A + 16 + xo, oA_jm = L::OFF_A + ly * L::RW_A + 16 + xo;
  const int oB_own = L::OFF_B + ly * L::RW_A + 16 + xo;
  const int oC_own = L::OFF_C + (ly + 1) * L::RW_C + xo, oC_jm = L::OFF_C + ly * L::RW_C + xo;
  const int oE = L::OFF_E + ly * L::RW_C + xo;
  const int oM = L::OFF_M + ly * L::RW_M + lx * V * (int) sizeof (T);
  const int oQD = L::OFF_Q + qrank * L::SZ_QD + ly * 128 + (lx & 7) * 16;
  const int oQB = L::OFF_Q + 3 * fb.nq * L::SZ_QD + qrank * L::SZ_QB + ly * 128 + (lx & 7) * 16;
  const int oX = ly * L::RW_C + xo;

  int slot = 0, st = sm0;       /* ring slot of iteration k and its byte offset */
  unsigned phase = 0;           /* mbarrier parity of that slot */
  int xw = xbuf0, xr = xbuf0 + L::X_BYTES; /* exchange buffer written in this iteration / written in the previous one */

#ifdef FDTD_FUSED_GENERIC_LD
#define SMV(W_, off) lds<T, W_> (smem_raw + (off))
#define SM1(off) (*reinterpret_cast<const T *> (smem_raw + (off)))
#else
#define SMV(W_, off) lds_s<T, W_> (sbase + (unsigned) (off))
#define SM1(off) (lds_s<T, 1> (sbase + (unsigned) (off)).v[0])
#endif

  constexpr int UNROLL_K = FDTD_FUSED_UNROLL; /* 2: the plane-carried values alternate between two register sets instead of being moved */
#pragma unroll UNROLL_K
  for (int k = k0; k <= k1; ++k, o += pl)
  {
    mbar_wait (&full[slot], phase);

    const int gk = p.kz0 + k;
    /* ---------------- E'(k) at this thread's position (tile cell or rim) */
    const FVec hz = SMV (W, st + oA_own), hz_jm = SMV (W, st + oA_jm);
    const FVec hy = SMV (W, st + oB_own);
    const FVec hx = SMV (W, st + oC_own), hx_jm = SMV (W, st + oC_jm);
    T hz_im[NC], hy_im[NC];
#pragma unroll
    for (int c = 0; c < NC; ++c)
    {
      hz_im[c] = SM1 (st + oA_own - L::CELL + c * (int) sizeof (T));
      hy_im[c] = SM1 (st + oB_own - L::CELL + c * (int) sizeof (T));
    }
    FVec ex = SMV (W, st + oE + 0 * L::E_STRIDE);
    FVec ey = SMV (W, st + oE + 1 * L::E_STRIDE);
    FVec ez = SMV (W, st + oE + 2 * L::E_STRIDE);
    const MVec eps_kp = SMV (V, st + oM);
    const MVec eps_jp_next = SMV (V, st + oM + L::RW_M);
    const T eps_ip_next = SM1 (st + oM + V * (int) sizeof (T));

    if (compute_e && !(k == k1 && top_from_nxt))
    {
      /* cells that may lie on a TF/SF face in this plane (necessary condition; the exact test is in tfsf_fix_e) */
      const int face_m = !TFSF ? 0 : ((j_face_e || tfsf_axis_hit (p, 2, gk, true)) ? (1 << V) - 1 : ife_m);
      FVec dx, dy, dz;
      if (shell)
      {
        dx = SMV (W, st + oQD + (0 * fb.nq) * L::SZ_QD);
        dy = SMV (W, st + oQD + (1 * fb.nq) * L::SZ_QD);
        dz = SMV (W, st + oQD + (2 * fb.nq) * L::SZ_QD);
      }
#pragma unroll
      for (int v = 0; v < V; ++v)
      {
        const int i = i0 + v;
        T cex[NC], cey[NC], cez[NC];
        T hz_m_[NC], hy_m_[NC];
#pragma unroll
        for (int c = 0; c < NC; ++c)
        {
          const int s = v * NC + c;
          hz_m_[c] = v > 0 ? hz.v[v > 0 ? s - NC : s] : hz_im[c];
          hy_m_[c] = v > 0 ? hy.v[v > 0 ? s - NC : s] : hy_im[c];
          cex[c] = hz.v[s] - hz_jm.v[s] - hy.v[s] + hy_km.v[s];
          cey[c] = hx.v[s] - hx_km.v[s] - hz.v[s] + hz_m_[c];
          cez[c] = hy.v[s] - hy_m_[c] - hx.v[s] + hx_jm.v[s];
        }
        if (TFSF && ((face_m >> v) & 1))
        {
          auto cell = [&] (const FVec &f) { Cell<T, NC> r; for (int c = 0; c < NC; ++c) r.v[c] = f.v[v * NC + c]; return r; };
          Cell<T, NC> a_, b_;
          Curl3<T, NC> cu;
          for (int c = 0; c < NC; ++c) { a_.v[c] = hz_m_[c]; b_.v[c] = hy_m_[c]; cu.x[c] = cex[c]; cu.y[c] = cey[c]; cu.z[c] = cez[c]; }
          cu = tfsf_fix_e<T, NC> (p, i, j, gk, k, cell (hz), cell (hz_jm), cell (hy), cell (hy_km), cell (hx), cell (hx_km), cell (hx_jm), a_, b_, cu);
          for (int c = 0; c < NC; ++c) { cex[c] = cu.x[c]; cey[c] = cu.y[c]; cez[c] = cu.z[c]; }
        }

        const T e_c = eps_c.v[v];
        const T e_xp = v + 1 < V ? eps_c.v[v + 1 < V ? v + 1 : v] : eps_ip;

        if (shell)
        {
          /* x absorber lanes: the UPML form, coefficients straight from the tables (few lanes, L1-resident); these
           * lanes include the cells at the x ends of the grid, so they keep the update-range predicates */
          const T mod_x = ((e_c + e_xp) / (T) 2) * eps0;
          const T mod_y = ((e_c + eps_jp.v[v]) / (T) 2) * eps0;
          const T mod_z = ((e_c + eps_kp.v[v]) / (T) 2) * eps0;
          const bool do_ex = (ihi_m >> v) & 1, do_ey = (ilo_m >> v) & 1, do_ez = do_ey;
          T cbx, ccx, cby, ccy, cbz, ccz;
          sweep3_coefs<T, false> (p.tab[0][KIND_BTW].p[i], p.tab[0][KIND_BTW].m[i], p.tab[2][KIND_AT].invp[gk], mod_x, cbx, ccx);
          sweep3_coefs<T, false> (p.tab[1][KIND_BTW].p[j], p.tab[1][KIND_BTW].m[j], p.tab[0][KIND_AT].invp[i], mod_y, cby, ccy);
          sweep3_coefs<T, false> (p.tab[2][KIND_BTW].p[gk], p.tab[2][KIND_BTW].m[gk], p.tab[1][KIND_AT].invp[j], mod_z, cbz, ccz);
          const T ca_x = p.tab[0][KIND_AT].ca[i], cbd_x = p.tab[0][KIND_AT].cbd[i];
          const T ca_y = p.tab[1][KIND_AT].ca[j], cbd_y = p.tab[1][KIND_AT].cbd[j];
          const T ca_z = p.tab[2][KIND_AT].ca[gk], cbd_z = p.tab[2][KIND_AT].cbd[gk];
#pragma unroll
          for (int c = 0; c < NC; ++c)
          {
            const int s = v * NC + c;
            if (do_ex) upml_cell<T, false> (dx.v[s], ex.v[s], cex[c], ca_y, cbd_y, ca_z, cbx, ccx);
            if (do_ey) upml_cell<T, false> (dy.v[s], ey.v[s], cey[c], ca_z, cbd_z, ca_x, cby, ccy);
            if (do_ez) upml_cell<T, false> (dz.v[s], ez.v[s], cez[c], ca_x, cbd_x, ca_y, cbz, ccz);
          }
        }
        else
        {
          /* Core lanes.  Every cell computed here lies strictly inside the grid (configure_fused keeps the box off the
           * grid's faces), so all three components are inside their update ranges: no predicates.  Where the four
           * material samples of the cell coincide -- everywhere but at material interfaces -- ((e + e) / 2) eps0 is
           * e eps0 exactly and one reciprocal serves the three components (same bits as three). */
          T cbx, cby, cbz;
          if (e_xp == e_c && eps_jp.v[v] == e_c && eps_kp.v[v] == e_c) cbx = cby = cbz = dt_dx * rcp_t<T> (e_c * eps0);
          else
          {
            cbx = dt_dx * rcp_t<T> (((e_c + e_xp) / (T) 2) * eps0);
            cby = dt_dx * rcp_t<T> (((e_c + eps_jp.v[v]) / (T) 2) * eps0);
            cbz = dt_dx * rcp_t<T> (((e_c + eps_kp.v[v]) / (T) 2) * eps0);
          }
#pragma unroll
          for (int c = 0; c < NC; ++c)
          {
            const int s = v * NC + c;
            ex.v[s] = fma_t<T> (cbx, cex[c], ex.v[s]);
            ey.v[s] = fma_t<T> (cby, cey[c], ey.v[s]);
            ez.v[s] = fma_t<T> (cbz, cez[c], ez.v[s]);
          }
        }

        if (!TFSF && src_row && gk == p.srcpos[2] && i == p.srcpos[0])
        {
          /* the source overwrites Ez but not Dz (Scheme3D.cpp:1995-2024); see e_phase_3d_tma */
          if (!shell && own && k < k1)
          {
#pragma unroll
            for (int c = 0; c < NC; ++c) p.f[8][o + v * NC + c] = fma_t<T> (p.tab[0][KIND_AT].cbd[i], cez[c], p.f[8][o + v * NC + c]);
          }
#pragma unroll
          for (int c = 0; c < NC; ++c) ez.v[v * NC + c] = p.src[step * NC + c];
        }
      }
      if (own && k < k1)
      {
        stv<T, W> (p.fo[0] + o, ex); stv<T, W> (p.fo[1] + o, ey); stv<T, W> (p.fo[2] + o, ez);
        if (shell)
        {
          /* the other member of the D pair (compact layout) */
          const long long od = ((((long long) (k + 1) * p.ny + j) * fb.dq_pitch) + compact_q (myq) * L::QX + (lx & 7) * V) * NC;
          stv<T, W> ((T *) fb.dq_out[0] + od, dx); stv<T, W> ((T *) fb.dq_out[1] + od, dy); stv<T, W> ((T *) fb.dq_out[2] + od, dz);
        }
      }
    }
    else if (ingrid)
    {
      /* outside the box: E' as the E shell pass wrote it */
      ex = ldv_rw<T, W> (p.fo[0] + o);
      ey = ldv_rw<T, W> (p.fo[1] + o);
      ez = ldv_rw<T, W> (p.fo[2] + o);
    }

    /* ---------------- H'(k-1) of the tile's own cells: E'(k-1) of the neighbours from the exchange buffer written in
     * the previous iteration, E'(k-1) of this cell and H(k-1) from registers, E'x, E'y (k) just computed */
    if (k > k0 && own)
    {
      const int m = k - 1, gm = gk - 1;
      const FVec ex_jp = SMV (W, xr + oX + 0 * L::E_STRIDE + L::RW_C);
      const FVec ez_jp = SMV (W, xr + oX + 2 * L::E_STRIDE + L::RW_C);
      T ey_ip[NC], ez_ip[NC];
#pragma unroll
      for (int c = 0; c < NC; ++c)
      {
        ey_ip[c] = SM1 (xr + oX + 1 * L::E_STRIDE + 16 + c * (int) sizeof (T));
        ez_ip[c] = SM1 (xr + oX + 2 * L::E_STRIDE + 16 + c * (int) sizeof (T));
      }
      const int face_m = !TFSF ? 0 : ((j_face_h || tfsf_axis_hit (p, 2, gm, false)) ? (1 << V) - 1 : ifh_m);
      FVec bx, by, bz;
      if (shell)
      {
        bx = SMV (W, st + oQB + (0 * fb.nq) * L::SZ_QB);
        by = SMV (W, st + oQB + (1 * fb.nq) * L::SZ_QB);
        bz = SMV (W, st + oQB + (2 * fb.nq) * L::SZ_QB);
      }
      FVec hxn = hx_km, hyn = hy_km, hzn = hz_km;
#pragma unroll
      for (int v = 0; v < V; ++v)
      {
        const int i = i0 + v;
        T chx[NC], chy[NC], chz[NC], ez_p[NC], ey_p[NC];
#pragma unroll
        for (int c = 0; c < NC; ++c)
        {
          const int s = v * NC + c;
          ez_p[c] = v + 1 < V ? ezm.v[v + 1 < V ? s + NC : s] : ez_ip[c];
          ey_p[c] = v + 1 < V ? eym.v[v + 1 < V ? s + NC : s] : ey_ip[c];
          chx[c] = ey.v[s] - eym.v[s] - ez_jp.v[s] + ezm.v[s];
          chy[c] = ez_p[c] - ezm.v[s] - ex.v[s] + exm.v[s];
          chz[c] = ex_jp.v[s] - exm.v[s] - ey_p[c] + eym.v[s];
        }
        if (TFSF && ((face_m >> v) & 1))
        {
          auto cell = [&] (const FVec &f) { Cell<T, NC> r; for (int c = 0; c < NC; ++c) r.v[c] = f.v[v * NC + c]; return r; };
          Cell<T, NC> a_, b_;
          Curl3<T, NC> cu;
          for (int c = 0; c < NC; ++c) { a_.v[c] = ez_p[c]; b_.v[c] = ey_p[c]; cu.x[c] = chx[c]; cu.y[c] = chy[c]; cu.z[c] = chz[c]; }
          cu = tfsf_fix_h<T, NC> (p, i, j, gm, m, cell (ey), cell (eym), cell (ez_jp), cell (ezm), a_, cell (ex), cell (exm), cell (ex_jp), b_, cu);
          for (int c = 0; c < NC; ++c) { chx[c] = cu.x[c]; chy[c] = cu.y[c]; chz[c] = cu.z[c]; }
        }
        if (shell)
        {
          const T mod = p.mu_const * p.mu0;
          const bool do_hx = (ilo_m >> v) & 1, do_hy = (ihi_m >> v) & 1, do_hz = do_hy;
          T cbx, ccx, cby, ccy, cbz, ccz;
          sweep3_coefs<T, false> (p.tab[0][KIND_AT].p[i], p.tab[0][KIND_AT].m[i], p.tab[2][KIND_BTW].invp[gm], mod, cbx, ccx);
          sweep3_coefs<T, false> (p.tab[1][KIND_AT].p[j], p.tab[1][KIND_AT].m[j], p.tab[0][KIND_BTW].invp[i], mod, cby, ccy);
          sweep3_coefs<T, false> (p.tab[2][KIND_AT].p[gm], p.tab[2][KIND_AT].m[gm], p.tab[1][KIND_BTW].invp[j], mod, cbz, ccz);
          const T ca_x = p.tab[0][KIND_BTW].ca[i], cbd_x = p.tab[0][KIND_BTW].cbd[i];
          const T ca_y = p.tab[1][KIND_BTW].ca[j], cbd_y = p.tab[1][KIND_BTW].cbd[j];
          const T ca_z = p.tab[2][KIND_BTW].ca[gm], cbd_z = p.tab[2][KIND_BTW].cbd[gm];
#pragma unroll
          for (int c = 0; c < NC; ++c)
          {
            const int s = v * NC + c;
            if (do_hx) upml_cell<T, false> (bx.v[s], hxn.v[s], chx[c], ca_y, cbd_y, ca_z, cbx, ccx);
            if (do_hy) upml_cell<T, false> (by.v[s], hyn.v[s], chy[c], ca_z, cbd_z, ca_x, cby, ccy);
            if (do_hz) upml_cell<T, false> (bz.v[s], hzn.v[s], chz[c], ca_x, cbd_x, ca_y, cbz, ccz);
          }
        }
        else
        {
          /* core lanes: inside every update range (see above) */
#pragma unroll
          for (int c = 0; c < NC; ++c)
          {
            const int s = v * NC + c;
            hxn.v[s] = fma_t<T> (cb_h, chx[c], hxn.v[s]);
            hyn.v[s] = fma_t<T> (cb_h, chy[c], hyn.v[s]);
            hzn.v[s] = fma_t<T> (cb_h, chz[c], hzn.v[s]);
          }
        }
      }
      stv<T, W> (p.fo[3] + o - pl, hxn); stv<T, W> (p.fo[4] + o - pl, hyn); stv<T, W> (p.fo[5] + o - pl, hzn);
      if (shell)
      {
        stv<T, W> (p.f[9] + o - pl, bx); stv<T, W> (p.f[10] + o - pl, by); stv<T, W> (p.f[11] + o - pl, bz);
      }
    }

    /* ---------------- hand E'(k) to the neighbours in the CTA, shift the plane-carried values */
    if (k < k1 && valid)
    {
      sts_s<T, W> (sbase + (unsigned) (xw + oX + 0 * L::E_STRIDE), ex);
      sts_s<T, W> (sbase + (unsigned) (xw + oX + 1 * L::E_STRIDE), ey);
      sts_s<T, W> (sbase + (unsigned) (xw + oX + 2 * L::E_STRIDE), ez);
    }
    exm = ex; eym = ey; ezm = ez;
    hx_km = hx; hy_km = hy; hz_km = hz;
    eps_c = eps_kp; eps_jp = eps_jp_next; eps_ip = eps_ip_next;

    /* every thread has read the stage and the older exchange buffer: refill the slot with plane k + S */
    __syncthreads ();
    if (leader && k + S <= k1) issue (slot, k + S);
    if (++slot == S) { slot = 0; st = sm0; phase ^= 1u; }
    else st += stage_sz;
    { const int t_ = xw; xw = xr; xr = t_; }
  }
#undef SMV
#undef SM1
}

/* ------------------------------------------------------------------ host side */

template <typename T, int NC>
bool prepare_fused_t (FusedMaps3D *out, void *const e[3], void *const h[3], void *const d[3], long long dq_pitch, void *const b[3],
                      void *eps, int pitch, int ny, int planes, int pitchM, int planesM, std::string *err)
{
  typedef FTile<T, NC> L;
  static_assert (sizeof (FusedSet) <= sizeof (out->set), "FusedMaps3D too small");
  FusedSet *m = reinterpret_cast<FusedSet *> (out->set);
  const int sb = (int) sizeof (T);
  const long long xs = (long long) pitch * NC;
  bool ok = true;
  ok = ok && encode (&m->m[0], h[2], sb, xs, ny, planes, (L::TXC + 2 * L::V) * NC, L::TY + 2, err);
  ok = ok && encode (&m->m[1], h[1], sb, xs, ny, planes, (L::TXC + 2 * L::V) * NC, L::TY + 1, err);
  ok = ok && encode (&m->m[2], h[0], sb, xs, ny, planes, (L::TXC + L::V) * NC, L::TY + 2, err);
  for (int a = 0; a < 3 && ok; ++a) ok = encode (&m->m[3 + a], e[a], sb, xs, ny, planes, (L::TXC + L::V) * NC, L::TY + 1, err);
  ok = ok && encode (&m->m[6], eps, sb, pitchM, ny + 1, planesM, L::MW, L::TY + 2, err);
  /* d[] are the compact arrays: x extent dq_pitch cells */
  for (int a = 0; a < 3 && ok; ++a) ok = encode (&m->m[7 + a], d[a], sb, dq_pitch * NC, ny, planes, L::QX * NC, L::TY + 1, err);
  for (int a = 0; a < 3 && ok; ++a) ok = encode (&m->m[10 + a], b[a], sb, xs, ny, planes, L::QX * NC, L::TY, err);
  out->valid = ok;
  return ok;
}

template <typename T, int NC, bool TFSF>
cudaError_t launch_fused_t (const StepParams<T> &p, const FusedSet &maps, const FusedBox &fb, cudaStream_t stream)
{
  typedef FTile<T, NC> L;
  const int smem = L::smem_bytes (fb.nq);
  static int attr_dev[64] = {};
  int dev = 0;
  cudaError_t de = cudaGetDevice (&dev);
  if (de != cudaSuccess) return de;
  if (attr_dev[dev & 63] < smem)
  {
    cudaError_t e = cudaFuncSetAttribute (step_core_3d_fused<T, NC, TFSF>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    if (e != cudaSuccess) return e;
    attr_dev[dev & 63] = smem;
  }
  const int rows = fb.fy1 - fb.fy0 + 1, planes = fb.fz1 - fb.fz0 + 1;
  if (rows <= 0 || planes <= 0) return cudaSuccess;
  dim3 block (L::NTHREADS);
  dim3 grid ((p.nx + L::TXC - 1) / L::TXC, rows / L::TY, (planes + fb.kc - 1) / fb.kc);
  step_core_3d_fused<T, NC, TFSF><<<grid, block, smem, stream>>> (p, maps, fb);
  return cudaGetLastError ();
}

} /* namespace */

void fused_tile_geometry (int value_type, int nc, int *tile_cells_x, int *tile_rows, int *ctas_per_sm, int *smem_per_quarter,
                          int *smem_base)
{
#define GEOM(T, NC) do { typedef FTile<T, NC> L; if (tile_cells_x) *tile_cells_x = L::TXC; if (tile_rows) *tile_rows = L::TY; \
    if (ctas_per_sm) *ctas_per_sm = L::MINB; if (smem_per_quarter) *smem_per_quarter = L::smem_bytes (1) - L::smem_bytes (0); \
    if (smem_base) *smem_base = L::smem_bytes (0); } while (0)
  if (value_type == 4 && nc == 1) GEOM (float, 1);
  else if (value_type == 4) GEOM (float, 2);
  else if (nc == 1) GEOM (double, 1);
  else GEOM (double, 2);
#undef GEOM
}

bool fused_prepare_3d (FusedMaps3D *out, int value_type, int nc, void *const e[3], void *const h[3], void *const d[3],
                       long long dq_pitch, void *const b[3], void *eps, int pitch, int ny, int planes, int pitchM, int planesM,
                       std::string *err)
{
  out->valid = false;
  if (value_type == 4)
    return nc == 2 ? prepare_fused_t<float, 2> (out, e, h, d, dq_pitch, b, eps, pitch, ny, planes, pitchM, planesM, err)
                   : prepare_fused_t<float, 1> (out, e, h, d, dq_pitch, b, eps, pitch, ny, planes, pitchM, planesM, err);
  return nc == 2 ? prepare_fused_t<double, 2> (out, e, h, d, dq_pitch, b, eps, pitch, ny, planes, pitchM, planesM, err)
                 : prepare_fused_t<double, 1> (out, e, h, d, dq_pitch, b, eps, pitch, ny, planes, pitchM, planesM, err);
}

template <typename T>
cudaError_t launch_step_core_fused (const StepParams<T> &p, const FusedMaps3D &maps, const FusedBox &fb, bool cplx, bool tfsf,
                                    cudaStream_t stream)
{
  const FusedSet &set = *reinterpret_cast<const FusedSet *> (maps.set);
  if (cplx)
    return tfsf ? launch_fused_t<T, 2, true> (p, set, fb, stream) : launch_fused_t<T, 2, false> (p, set, fb, stream);
  return tfsf ? launch_fused_t<T, 1, true> (p, set, fb, stream) : launch_fused_t<T, 1, false> (p, set, fb, stream);
}

template cudaError_t launch_step_core_fused<float> (const StepParams<float> &, const FusedMaps3D &, const FusedBox &, bool, bool, cudaStream_t);
template cudaError_t launch_step_core_fused<double> (const StepParams<double> &, const FusedMaps3D &, const FusedBox &, bool, bool, cudaStream_t);

} /* namespace fdtd */
