/*
 * Host side of the C ABI declared in include/fdtd3d_b200.h: device-resident solver state, UPML
 * coefficient tables, the time loop (kernel launches + z-slab halo exchange) and host<->device
 * transfers.  There is no CPU compute path: every entry point that needs a device fails with
 * FDTD_NO_DEVICE if there is none.
 *
 * Reference being replaced: Source/Cuda/CudaInterface.cu (cudaExecute3DSteps :356-932,
 * cudaExecute2DTMzSteps :20-339, cudaInit :934, cudaInfo :939-959) and, for the exchange,
 * Source/Grid/ParallelGrid.cpp:1817-1823 (share) -- see DESIGN.md.
 */
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include <cuda_runtime.h>

#include "../../include/fdtd3d_b200.h"
#include "fdtd_device.cuh"
#include "kernels3d_common.cuh"
#include "launch.h"
#include "nccl_dl.h"

namespace fdtd
{

static thread_local std::string g_last_error;

static int fail (int code, const char *fmt, ...)
{
  char buf[1024];
  va_list ap;
  va_start (ap, fmt);
  vsnprintf (buf, sizeof (buf), fmt, ap);
  va_end (ap);
  g_last_error = buf;
  return code;
}

#define CUDA_TRY(expr) \
  do { cudaError_t e_ = (expr); if (e_ != cudaSuccess) return fail (FDTD_ERROR, "%s failed: %s (%s:%d)", #expr, cudaGetErrorString (e_), __FILE__, __LINE__); } while (0)

#define NCCL_TRY(expr) \
  do { int r_ = (expr); if (r_ != 0) return fail (FDTD_COMM_ERROR, "%s failed: %s (%s:%d)", #expr, nccl->GetErrorString (r_), __FILE__, __LINE__); } while (0)

static inline long long round_up (long long x, long long m) { return (x + m - 1) / m * m; }

/* Physics/PhysicsConst.h:9-13: constants stored as FPValue */
template <typename T> struct PC
{
  static T c () { return (T) 2.99792458e+8; }
  static T eps0 () { return (T) 0.0000000000088541878176203892; }
  static T mu0 () { return (T) 0.0000012566370614359173; }
  static T pi () { return (T) M_PI; }
};

} /* namespace fdtd */

using namespace fdtd;

struct fdtd_solver
{
  fdtd_config cfg;
  int nc;            /* scalars per cell */
  int fp;            /* sizeof FPValue */
  int cell;          /* bytes per cell */
  int nx, ny, nzg;   /* global sizes (nzg = 1 in 2D) */
  int nzl, kz0;      /* local planes and their global offset */
  int nzm;           /* local material planes (nzl+1 in 3D, 1 in 2D) */
  int pitch, pitchM;
  long long ps, psM;
  long long field_cells, mat_vals; /* allocation sizes */
  long long ninc;

  void *f[12];
  void *eps, *mu;
  void *einc, *hinc;
  void *face_tab;    /* TF/SF face tables of the TMA kernels (24 x face_stride cells) */
  long long face_stride;
  void *tab_mem;     /* all coefficient tables, one allocation */
  size_t tab_stride; /* bytes between tables */
  void *src;         /* source samples of the run being executed (device) */
  long long src_capacity;
  char *src_ring;    /* pinned host ring the samples of queued runs are copied from (fdtd_run_async never waits for it
                        unless more samples are queued than it holds) */
  size_t src_ring_bytes, src_ring_head, src_ring_used;
  void *staging;     /* layout-conversion scratch */
  size_t staging_bytes;
  double *red_partials, *red_out; /* reduction scratch */

  std::vector<double> sigma[3]; /* profiles as given (FPValue-valued) */

  cudaStream_t s_compute, s_comm;
  cudaEvent_t ev_eb, ev_ex, ev_hb, ev_hx;
  /* 3D TF/SF with face tables: the incident line and the tables depend on no 3-D field, so they are advanced on a
   * stream of their own, one half step ahead of the half-step kernels that read the tables (enqueue_steps) */
  cudaStream_t s_inc = NULL;
  cudaEvent_t ev_i0 = NULL, ev_i1 = NULL, ev_fe = NULL, ev_fh = NULL, ev_ed = NULL, ev_hd = NULL;
  std::vector<cudaEvent_t> run_ev; /* a (start, stop) pair per run queued since the last synchronisation */
  int runs_pending;
  NcclComm comm;
  bool comm_failed;   /* a send / recv of the halo exchange failed: the handle refuses further multi-rank runs */

  TmaMaps3D tma;        /* tensor maps of the TMA-staged 3D kernels */

  /* One pass per step inside the core (kernels3d_fused.cu): E and H become ping-pong pairs.  fA is the allocation
   * fdtd_create made, fB the second one; f[0..5] always point at the buffers that hold the current fields (parity 0:
   * fA), so everything outside the time loop is unaware of the pair. */
  void *fA[6], *fB[6];
  int parity;
  bool fused_ok;        /* geometry allows it and the second buffers exist */
  bool fused_run;       /* the run being queued uses it (fused_ok and mu homogeneous) */
  FusedBox fbox;
  int shell_ty;         /* rows per tile of the half-step kernels that advance the shell */
  FusedMaps3D fmaps[2][2]; /* [E/H parity][D parity]: reads the current buffers */
  /* D of the absorber quarters inside the box: compact ping-pong pair (launch.h, FusedBox); dq_live: the member dq_par
   * holds the authoritative values (gathered from the D arrays at the start of a one-pass run, scattered back before
   * anything else reads or writes D) */
  void *dq[2][3];
  size_t dq_bytes;
  int dq_par;
  bool dq_live;
  TmaMaps3D tma_pp[2];  /* per parity, half-step kernels of the shell: E phase reads cur, H phase reads E of nxt */
  int fused_kc;         /* planes marched per CTA of the fused kernel; 0 = default rule */
  int kchunk_forced;    /* fdtd_set_tuning: planes marched per CTA of the half-step kernels; 0 = default rule */
  /* launch-bound grids: one step captured as a CUDA graph and replayed (run_graph); the step index then lives in
   * device memory (StepParams::step_ptr).  Any change of what a step launches invalidates the captured graph. */
  long long *step_dev;
  cudaGraphExec_t step_graph;
  void *step_graph_src;   /* the source table the captured launches point at */
  bool step_graph_prof;
  int step_graph_launches;
  int fused_policy;     /* fdtd_set_tuning: 0 never, 1 wherever the geometry allows, -1 (default) where it was measured to
                           pay: complex double */
  float prof_f_ms;
  /* multi-rank runs with profiling on: nine events per step around the eight parts of a step (fdtd_step_breakdown) */
  std::vector<cudaEvent_t> mprof_ev;
  long long mprof_steps;
  float mprof_ms[8];
  bool use_tma;
  int kchunk;

  /* sigma-free core (kernels3d_tma.cu, CORE variants): box of cells whose D/B the time loop skips, and whether the
   * D/B arrays currently hold the folded representation D - eps eps0 E there (aux_kernels.cu, fold_aux_kernel) */
  bool mu_uniform;      /* the mu array was last set with fdtd_fill: one value everywhere (UNIMU H kernels) */
  double mu_value;
  bool eps_uniform;     /* likewise eps (TMz E kernel) */
  double eps_value;
  bool core_allowed, core_on, folded;
  bool tma_alt;         /* tile configuration the tensor maps and kchunk were set up for (exact || core_on) */
  int core_lo[3], core_hi[3];

  bool profiling;
  std::vector<cudaEvent_t> prof_ev; /* 4 per step: E begin/end, H begin/end */
  long long prof_steps;
  float prof_e_ms, prof_h_ms;

  long long t;          /* next time step index */
  long long launches;
  float last_ms;
};

namespace fdtd
{

static const int N_TABLES = 3 * 2 * 5;

template <typename T>
static StepParams<T> make_params (const fdtd_solver *s)
{
  StepParams<T> p;
  memset (&p, 0, sizeof (p));
  const fdtd_config &c = s->cfg;
  for (int i = 0; i < 12; ++i) p.f[i] = (T *) s->f[i];
  for (int i = 0; i < 6; ++i) p.fo[i] = (T *) s->f[i]; /* in place unless the time loop says otherwise */
  p.nreg = 0;
  p.eps = (const T *) s->eps;
  p.mu = (const T *) s->mu;
  p.mu_const = (T) s->mu_value;
  p.eps_const = (T) s->eps_value;
  for (int a = 0; a < 3; ++a)
    for (int k = 0; k < 2; ++k)
    {
      const char *base = (const char *) s->tab_mem + (size_t) ((a * 2 + k) * 5) * s->tab_stride;
      p.tab[a][k].ca = (const T *) (base + 0 * s->tab_stride);
      p.tab[a][k].cbd = (const T *) (base + 1 * s->tab_stride);
      p.tab[a][k].p = (const T *) (base + 2 * s->tab_stride);
      p.tab[a][k].m = (const T *) (base + 3 * s->tab_stride);
      p.tab[a][k].invp = (const T *) (base + 4 * s->tab_stride);
    }
  p.einc = (T *) s->einc;
  p.hinc = (T *) s->hinc;
  p.face_tab = (T *) s->face_tab;
  p.face_stride = s->face_stride;
  p.src = (const T *) s->src;
  p.step = 0;
  p.step_ptr = NULL;
  p.ninc = s->ninc;
  p.nx = s->nx; p.ny = s->ny; p.nzl = s->nzl; p.kz0 = s->kz0; p.nzg = s->nzg;
  p.pitch = s->pitch; p.pitchM = s->pitchM; p.ps = s->ps; p.psM = s->psM;
  p.kchunk = s->kchunk;
  for (int a = 0; a < 3; ++a)
  {
    p.core_lo[a] = s->core_on ? s->core_lo[a] : 0;
    p.core_hi[a] = s->core_on ? s->core_hi[a] : -1;
    p.L[a] = (int) c.tfsf_size[a];
    p.R[a] = (int) (c.size[a] - c.tfsf_size[a]);
    p.srcpos[a] = (int) c.source_pos[a];
  }
  const T dt = (T) c.dt, dx = (T) c.dx, relv = (T) c.rel_phase_velocity;
  p.dt = dt; p.dx = dx; p.eps0 = PC<T>::eps0 (); p.mu0 = PC<T>::mu0 ();
  p.dt_dx = dt / dx;
  /* Scheme3D.cpp:39,76: gridTimeStep / (relPhaseVelocity * Eps0 * gridStep) */
  p.inc_ce = dt / (relv * PC<T>::eps0 () * dx);
  p.inc_ch = dt / (relv * PC<T>::mu0 () * dx);
  const T a1 = (T) c.angle_teta, a2 = (T) c.angle_phi, a3 = (T) c.angle_psi;
  /* YeeGridLayout.h:438-440 */
  p.zero_inc[0] = (T) ((unsigned long long) c.tfsf_size[0] - 2.5 * sin ((double) a1) * cos ((double) a2));
  p.zero_inc[1] = (T) ((unsigned long long) c.tfsf_size[1] - 2.5 * sin ((double) a1) * sin ((double) a2));
  p.zero_inc[2] = (T) ((unsigned long long) c.tfsf_size[2] - 2.5 * cos ((double) a1));
  p.s1 = sin ((double) a1); p.c1 = cos ((double) a1);
  p.s2 = sin ((double) a2); p.c2 = cos ((double) a2);
  const double s3 = sin ((double) a3), c3 = cos ((double) a3);
  /* YeeGridLayout.cpp:811-845 */
  p.proj_e[0] = (T) (c3 * p.s2 - s3 * p.c1 * p.c2);
  p.proj_e[1] = (T) (-c3 * p.c2 - s3 * p.c1 * p.s2);
  p.proj_e[2] = (T) (s3 * p.s1);
  p.proj_h[0] = (T) (s3 * p.s2 + c3 * p.c1 * p.c2);
  p.proj_h[1] = (T) (-s3 * p.c2 + c3 * p.c1 * p.s2);
  p.proj_h[2] = -(T) (c3 * p.s1);
  p.cphi = p.c2; p.sphi = p.s2;
  return p;
}

/* UPML coefficient tables from the 1-D sigma profiles (fdtd_device.cuh, struct CoefTab), with the
 * reference's expressions Scheme3D.cpp:310-311, 401-403 evaluated in FPValue arithmetic. */
template <typename T>
static int build_tables (fdtd_solver *s)
{
  const fdtd_config &c = s->cfg;
  const size_t len = s->tab_stride / sizeof (T);
  std::vector<T> host ((size_t) N_TABLES * len);
  const T eps0 = PC<T>::eps0 ();
  const T dt = (T) c.dt, dx = (T) c.dx;
  for (int a = 0; a < 3; ++a)
  {
    const std::vector<double> &sg = s->sigma[a];
    const long long n = (long long) sg.size (); /* N_a + 1 */
    for (int kind = 0; kind < 2; ++kind)
    {
      T *ca = &host[(size_t) ((a * 2 + kind) * 5 + 0) * len];
      T *cbd = &host[(size_t) ((a * 2 + kind) * 5 + 1) * len];
      T *pp = &host[(size_t) ((a * 2 + kind) * 5 + 2) * len];
      T *mm = &host[(size_t) ((a * 2 + kind) * 5 + 3) * len];
      T *ip = &host[(size_t) ((a * 2 + kind) * 5 + 4) * len];
      for (size_t i = 0; i < len; ++i)
      {
        T sig = 0;
        if ((long long) i < n)
        {
          if (kind == KIND_AT) sig = (T) sg[i];
          else
          {
            /* Approximation.cpp:15-19: (val1 + val2) / 2.0 */
            T v1 = (T) sg[i], v2 = (long long) i + 1 < n ? (T) sg[i + 1] : (T) sg[i];
            sig = (v1 + v2) / 2.0;
          }
        }
        T k = 1;
        ca[i] = (2 * eps0 * k - sig * dt) / (2 * eps0 * k + sig * dt);
        cbd[i] = (2 * eps0 * dt / dx) / (2 * eps0 * k + sig * dt);
        pp[i] = 2 * eps0 * k + sig * dt;
        mm[i] = 2 * eps0 * k - sig * dt;
        ip[i] = (T) 1 / pp[i];
      }
    }
  }
  CUDA_TRY (cudaMemcpyAsync (s->tab_mem, host.data (), host.size () * sizeof (T), cudaMemcpyHostToDevice, s->s_compute));
  CUDA_TRY (cudaStreamSynchronize (s->s_compute));
  return FDTD_OK;
}

static int build_tables_any (fdtd_solver *s)
{
  return s->fp == 4 ? build_tables<float> (s) : build_tables<double> (s);
}

/* source samples for steps [t, t+n): sin (gridTimeStep * t * 2 * Pi * sourceFrequency), complex builds
 * (sin, cos) -- Scheme3D.cpp:47-52, 2016-2021 -- evaluated on the host with the C double sin/cos as
 * the reference does, so the waveform is the reference's bit for bit. */
template <typename T>
static int upload_source (fdtd_solver *s, long long n)
{
  const fdtd_config &c = s->cfg;
  const size_t bytes = (size_t) n * s->cell;
  /* The samples are written into a pinned ring and copied from there in stream order, so queueing a run does not
   * wait for the runs before it.  Only when the ring cannot take the new samples beside those of the runs still
   * queued, or the device table has to grow, the stream is drained first. */
  if (n > s->src_capacity || bytes > s->src_ring_bytes - s->src_ring_used)
  {
    CUDA_TRY (cudaStreamSynchronize (s->s_compute));
    s->src_ring_head = s->src_ring_used = 0;
  }
  if (n > s->src_capacity)
  {
    if (s->src) cudaFree (s->src);
    s->src = NULL;
    s->src_capacity = 0;
    CUDA_TRY (cudaMalloc (&s->src, bytes));
    s->src_capacity = n;
  }
  if (bytes > s->src_ring_bytes)
  {
    if (s->src_ring) cudaFreeHost (s->src_ring);
    s->src_ring = NULL;
    s->src_ring_bytes = 0;
    CUDA_TRY (cudaHostAlloc ((void **) &s->src_ring, bytes, cudaHostAllocDefault));
    s->src_ring_bytes = bytes;
  }
  if (s->src_ring_head + bytes > s->src_ring_bytes)
  {
    /* no wrap inside one run: skip the tail (it counts as used until the next drain) */
    s->src_ring_used += s->src_ring_bytes - s->src_ring_head;
    s->src_ring_head = 0;
    if (bytes > s->src_ring_bytes - s->src_ring_used)
    {
      CUDA_TRY (cudaStreamSynchronize (s->s_compute));
      s->src_ring_used = 0;
    }
  }
  T *host = reinterpret_cast<T *> (s->src_ring + s->src_ring_head);
  const T dt = (T) c.dt, freq = (T) c.source_frequency;
  for (long long i = 0; i < n; ++i)
  {
    const int t = (int) (s->t + i);
    const T arg = dt * t * 2 * PC<T>::pi () * freq;
    host[(size_t) i * s->nc] = (T) ::sin ((double) arg);
    if (s->nc == 2) host[(size_t) i * 2 + 1] = (T) ::cos ((double) arg);
  }
  CUDA_TRY (cudaMemcpyAsync (s->src, host, bytes, cudaMemcpyHostToDevice, s->s_compute));
  s->src_ring_head += bytes;
  s->src_ring_used += bytes;
  return FDTD_OK;
}

/* ------------------------------------------------------------------ halo exchange */

/* H phase -> E phase: top owned plane of Hx, Hy goes to the upper neighbour's low halo;
 * E phase -> H phase: bottom owned plane of Ex, Ey goes to the lower neighbour's high halo. */
static int exchange (fdtd_solver *s, bool h_fields, void *const *fld)
{
  const char *why = NULL;
  const NcclApi *nccl = nccl_api (&why);
  if (!nccl || !s->comm) return fail (FDTD_COMM_ERROR, "halo exchange without a communicator (%s)", why ? why : "");
  if (s->comm_failed) return fail (FDTD_COMM_ERROR, "an earlier halo exchange of this handle failed: destroy it");
  const size_t plane_bytes = (size_t) s->ps * s->cell;
  const int r = s->cfg.rank, n = s->cfg.nranks;
  NCCL_TRY (nccl->GroupStart ());
  /* a failure inside the group must still close it, or the group stays open on this thread */
  int rc = 0;
#define IN_GROUP(expr) do { if (!rc) rc = (expr); } while (0)
  if (h_fields)
  {
    for (int f = FDTD_HX; f <= FDTD_HY; ++f)
    {
      char *base = (char *) fld[f];
      if (r + 1 < n) IN_GROUP (nccl->Send (base + (size_t) s->nzl * plane_bytes, plane_bytes, NCCL_INT8, r + 1, s->comm, s->s_comm));
      if (r > 0) IN_GROUP (nccl->Recv (base, plane_bytes, NCCL_INT8, r - 1, s->comm, s->s_comm));
    }
  }
  else
  {
    for (int f = FDTD_EX; f <= FDTD_EY; ++f)
    {
      char *base = (char *) fld[f];
      if (r > 0) IN_GROUP (nccl->Send (base + plane_bytes, plane_bytes, NCCL_INT8, r - 1, s->comm, s->s_comm));
      if (r + 1 < n) IN_GROUP (nccl->Recv (base + (size_t) (s->nzl + 1) * plane_bytes, plane_bytes, NCCL_INT8, r + 1, s->comm, s->s_comm));
    }
  }
#undef IN_GROUP
  const int rc_end = nccl->GroupEnd ();
  if (!rc) rc = rc_end;
  if (rc)
  {
    s->comm_failed = true; /* later runs refuse cleanly instead of waiting on half-wired events */
    return fail (FDTD_COMM_ERROR, "halo exchange failed: %s", nccl->GetErrorString (rc));
  }
  return FDTD_OK;
}

/* ------------------------------------------------------------------ sigma-free core */

/*
 * The box of cells where all three sigma profiles vanish at both material samples a component can use along each
 * axis (sigma[i] and sigma[i+1]): per axis the longest run of zeros of the profile, shrunk in x to whole quarter
 * tile rows and in y to whole tile rows of the TMA kernels (they decide per quarter row), intersected over the axes.
 */
static bool compute_core_box (int dims, int fp, int nc, const std::vector<double> sigma[3], int lo_out[3], int hi_out[3])
{
  for (int a = 0; a < 3; ++a) { lo_out[a] = 0; hi_out[a] = -1; }
  int txc = 0, ty = 0, qx = 0;
  if (dims == 3) tma_tile_geometry (fp, nc, true, &txc, &ty, &qx, NULL); /* the CORE variants' tiles */
  else { qx = 8 * (16 / (fp * nc)); ty = 1; } /* TMz: per row; in x per quarter of a tile row (8 vectors), the unit in
                                                  which the one-pass kernel stages Dz / Bx / By (kernels2d_fused.cu) */
  int lo3[3], hi3[3];
  for (int a = 0; a < 3; ++a)
  {
    const std::vector<double> &sg = sigma[a];
    long long best0 = 0, best1 = -1, run0 = 0;
    for (long long i = 0; i <= (long long) sg.size (); ++i)
    {
      const bool zero = i < (long long) sg.size () && sg[(size_t) i] == 0.0;
      if (!zero)
      {
        if (i - 1 - run0 > best1 - best0) { best0 = run0; best1 = i - 1; }
        run0 = i + 1;
      }
    }
    long long lo = best0, hi = best1 - 1; /* cells whose samples i and i+1 are both in the run */
    if (a == 0) { lo = round_up (lo, qx); hi = (hi + 1) / qx * qx - 1; }
    if (a == 1) { lo = round_up (lo, ty); hi = (hi + 1) / ty * ty - 1; }
    if (hi < lo) return false;
    lo3[a] = (int) lo;
    hi3[a] = (int) hi;
  }
  for (int a = 0; a < 3; ++a) { lo_out[a] = lo3[a]; hi_out[a] = hi3[a]; }
  return true;
}

static void update_core (fdtd_solver *s)
{
  s->core_on = false;
  for (int a = 0; a < 3; ++a) { s->core_lo[a] = 0; s->core_hi[a] = -1; }
  if (!s->core_allowed) return;
  s->core_on = compute_core_box (s->cfg.dims, s->fp, s->nc, s->sigma, s->core_lo, s->core_hi);
}

/*
 * Tensor maps and planes marched per CTA for the tile configuration of the kernels that will run (3D only): the CORE and
 * EXACT variants use other tiles than the plain fast ones (kernels3d_tma.cu, TileRows), and whether the core is on
 * depends on the sigma profiles, so this runs at creation and again when fdtd_set_sigma changes the answer.
 */
static int configure_fused (fdtd_solver *s);
static int normalize_parity (fdtd_solver *s);
static int release_dq (fdtd_solver *s);

/* what a step launches has changed: a captured step is stale */
static void drop_step_graph (fdtd_solver *s)
{
  if (s->step_graph) cudaGraphExecDestroy (s->step_graph);
  s->step_graph = NULL;
}

static int configure_3d (fdtd_solver *s, bool first)
{
  drop_step_graph (s);
  if (s->cfg.dims != 3) return configure_fused (s); /* TMz: direct-load half-step kernels, no tensor maps of their own */
  const bool alt = s->cfg.exact_mode == 1 || s->core_on;
  int rc0 = normalize_parity (s); /* the maps below describe the first allocation */
  if (rc0) return rc0;
  if (!first && alt == s->tma_alt) return configure_fused (s); /* the core box may have moved all the same */
  s->tma_alt = alt;
  int txc = 32 * (16 / s->cell), tyr = 8, per_sm = 1; /* cells per CTA tile row, rows per CTA tile, resident CTAs per SM */
  if (s->use_tma) tma_tile_geometry (s->fp, s->nc, alt, &txc, &tyr, NULL, &per_sm);
  const long long tiles = (long long) ((s->nx + txc - 1) / txc) * ((s->ny + tyr - 1) / tyr);
  /* planes marched per CTA: measured best (profiles/r1_kchunk_sweep.log) when a launch has about 8192 CTAs per
   * resident CTA of an SM (~55 rounds) -- short enough marches to keep the SMs evenly loaded to the end, long enough to
   * amortise the pipeline fill; never more than 64 planes, and small grids still get at least two rounds */
  long long kc_ll = (long long) s->nzl * tiles / (8192LL * per_sm);
  const long long kc_small = (long long) s->nzl * tiles / (2LL * 148 * per_sm);
  if (kc_ll > kc_small) kc_ll = kc_small;
  int kc = (int) kc_ll;
  if (kc > 64) kc = 64;
  if (kc < 4) kc = 4;
  s->kchunk = s->kchunk_forced > 0 ? s->kchunk_forced : kc;
  if (s->use_tma)
  {
    std::string err;
    if (!tma_prepare_3d (&s->tma, s->fp, s->nc, s->cfg.use_pml != 0, alt, s->f, s->eps, s->mu, s->pitch, s->ny, s->nzl + 2,
                         s->pitchM, s->nzm + 2, &err))
      return fail (FDTD_ERROR, "TMA descriptor set-up failed: %s", err.c_str ());
  }
  return configure_fused (s);
}

/* ------------------------------------------------------------------ one pass per step in the core */

/* the current fields back into the first allocation (after an odd number of fused steps they live in the second) */
static int normalize_parity (fdtd_solver *s)
{
  if (s->parity == 0) return FDTD_OK;
  for (int i = 0; i < 6; ++i)
  {
    if (!s->fA[i]) continue; /* TMz has Ez, Hx, Hy only */
    CUDA_TRY (cudaMemcpyAsync (s->fA[i], s->fB[i], (size_t) s->field_cells * s->cell, cudaMemcpyDeviceToDevice, s->s_compute));
    s->f[i] = s->fA[i];
  }
  s->parity = 0;
  return FDTD_OK;
}

/*
 * Decides whether the one-pass kernel can run on this configuration and sets up its box, buffers and tensor maps
 * (3D, fast mode with a sigma-free core).  Requirements beyond a non-empty core: every x tile boundary inside the grid
 * lies in the x core (the rim lane recomputes E' there with the plain update), the D / B part of a stage fits in
 * shared memory, and the slab keeps one plane on either side that abuts another rank (its E' / H' travel mid-step).
 * Not being able to is not an error: the two half-step kernels then advance the whole slab as before.
 */
static int configure_fused (fdtd_solver *s)
{
  int rc = release_dq (s); /* with the box that gathered them */
  if (rc) return rc;
  s->fused_ok = false;
  const bool tmz = s->cfg.dims == 2;
  const bool wanted = s->fused_policy > 0 || (s->fused_policy < 0 && !tmz && s->nc == 2 && s->fp == 8);
  if (!s->core_on || (!tmz && !s->use_tma) || !wanted) return FDTD_OK;
  if ((rc = normalize_parity (s))) return rc;
  int txc = 0, fty = 0, smem_q = 0, smem_base = 0, sty = 1;
  if (tmz) fused_tile_geometry_tmz (s->fp, s->nc, &txc, &fty, &smem_q, &smem_base);
  else
  {
    fused_tile_geometry (s->fp, s->nc, &txc, &fty, NULL, &smem_q, &smem_base);
    tma_tile_geometry (s->fp, s->nc, true, NULL, &sty, NULL, NULL);
  }
  const int V = 16 / s->cell, qx = 8 * V;
  for (int b = txc; b < s->nx; b += txc)
    if (b < s->core_lo[0] || b + V - 1 > s->core_hi[0]) return FDTD_OK;
  FusedBox fb;
  fb.fy0 = s->core_lo[1];
  fb.fy1 = s->core_hi[1];
  if (tmz)
  {
    /* the half-step kernels of the shell work row by row: any whole number of tile rows will do */
    fb.fy1 = fb.fy0 + (fb.fy1 - fb.fy0 + 1) / fty * fty - 1;
    if (fb.fy1 < fb.fy0) return FDTD_OK;
    fb.fz0 = fb.fz1 = 0;
  }
  else
  {
    /* The kernel's core lanes carry no update-range predicates: every cell it computes must lie strictly inside the
     * grid.  With an absorber on every face this costs nothing; without one the box gives up the tile row / plane
     * on the face, and a core that reaches the x faces keeps the two-pass kernels. */
    if (s->core_lo[0] < 1 || s->core_hi[0] > s->nx - 2) return FDTD_OK;
    /* rows: whole tile rows of both the one-pass kernel and the half-step kernels that take the rest */
    const int al = sty % fty == 0 ? sty : (fty % sty == 0 ? fty : sty * fty);
    fb.fy0 = (fb.fy0 + al - 1) / al * al;
    if (fb.fy0 < 1) fb.fy0 = al;
    fb.fy1 = (fb.fy1 + 1) / al * al - 1;
    if (fb.fy1 > s->ny - 2) fb.fy1 = (s->ny - 1) / al * al - 1;
    if (fb.fy1 < fb.fy0) return FDTD_OK;
    int zlo = s->core_lo[2] - s->kz0, zhi = s->core_hi[2] - s->kz0;
    if (s->kz0 + zlo < 1) zlo = 1 - s->kz0;
    if (s->kz0 + zhi > s->nzg - 2) zhi = s->nzg - 2 - s->kz0;
    const int zmin = s->cfg.rank > 0 ? 1 : 0, zmax = s->cfg.rank + 1 < s->cfg.nranks ? s->nzl - 2 : s->nzl - 1;
    if (zlo < zmin) zlo = zmin;
    if (zhi > zmax) zhi = zmax;
    if (zhi < zlo) return FDTD_OK;
    fb.fz0 = zlo;
    fb.fz1 = zhi;
  }
  fb.nq = 0;
  for (int t = 0; t * txc < s->nx; ++t)
  {
    int n = 0;
    for (int q = 0; q < 4; ++q)
    {
      const int qa = t * txc + q * qx, qb = qa + qx - 1;
      if (qa < s->nx && !(qa >= s->core_lo[0] && qb <= s->core_hi[0])) ++n;
    }
    if (n > fb.nq) fb.nq = n;
  }
  if (smem_base + fb.nq * smem_q > 227 * 1024) return FDTD_OK;
  /* compact D pair of the absorber quarters: quarters below the x core, then those above it */
  fb.nlo = s->core_lo[0] / qx;
  fb.hq0 = (s->core_hi[0] + 1) / qx;
  fb.nqt = fb.nlo + ((s->nx + qx - 1) / qx - fb.hq0);
  fb.dq_pitch = (long long) (fb.nqt > 0 ? fb.nqt : 1) * qx;
  for (int a = 0; a < 3; ++a) fb.dq_out[a] = NULL;
  /* planes marched per CTA: one redundant plane per march (E' of the plane above the last one), so longer marches than
   * the half-step kernels'; still several rounds of CTAs per SM */
  const long long tiles = (long long) ((s->nx + txc - 1) / txc) * ((fb.fy1 - fb.fy0 + 1) / fty);
  if (tmz)
  {
    /* tiles marched per CTA (one below the other) */
    int kc = s->fused_kc > 0 ? s->fused_kc : 4;
    while (kc > 1 && tiles / kc < 8LL * 148) kc /= 2;
    fb.kc = kc;
  }
  else
  {
    int kc = s->fused_kc > 0 ? s->fused_kc : 32;
    const int planes = fb.fz1 - fb.fz0 + 1;
    while (kc > 8 && tiles * ((planes + kc - 1) / kc) < 8LL * 148) kc /= 2;
    if (kc > planes) kc = planes;
    fb.kc = kc;
  }

  if (!s->fB[2])
  {
    for (int i = 0; i < 6; ++i)
    {
      if (!s->fA[i]) continue;
      cudaError_t e = cudaMalloc (&s->fB[i], (size_t) s->field_cells * s->cell);
      if (e == cudaSuccess) e = cudaMemsetAsync (s->fB[i], 0, (size_t) s->field_cells * s->cell, s->s_compute);
      if (e != cudaSuccess)
      {
        /* no room for the second buffers: stay with the two-pass kernels */
        cudaGetLastError ();
        for (int k = 0; k <= i; ++k) { if (s->fB[k]) cudaFree (s->fB[k]); s->fB[k] = NULL; }
        return FDTD_OK;
      }
    }
  }
  {
    const size_t need = (size_t) (s->nzl + 2) * s->ny * fb.dq_pitch * s->cell;
    if (need != s->dq_bytes)
    {
      for (int d = 0; d < 2; ++d)
        for (int a = 0; a < 3; ++a) { if (s->dq[d][a]) cudaFree (s->dq[d][a]); s->dq[d][a] = NULL; }
      s->dq_bytes = 0;
      for (int d = 0; d < 2; ++d)
        for (int a = 0; a < (tmz ? 1 : 3); ++a)
        {
          cudaError_t e = cudaMalloc (&s->dq[d][a], need);
          if (e == cudaSuccess) e = cudaMemsetAsync (s->dq[d][a], 0, need, s->s_compute);
          if (e != cudaSuccess) { cudaGetLastError (); return FDTD_OK; } /* stay with the two-pass kernels */
        }
      s->dq_bytes = need;
    }
  }
  std::string err;
  for (int par = 0; par < 4 && tmz; ++par)
  {
    void **cur = (par >> 1) ? s->fB : s->fA;
    if (!fused_prepare_tmz (&s->fmaps[par >> 1][par & 1], s->fp, s->nc, cur[2], cur[3], cur[4], s->dq[par & 1][0], fb.dq_pitch, s->f[9],
                            s->f[10], s->eps, s->pitch, s->ny, s->pitchM, &err))
      return fail (FDTD_ERROR, "TMA descriptor set-up of the one-pass TMz kernel failed: %s", err.c_str ());
  }
  for (int par = 0; par < 2 && !tmz; ++par)
  {
    void **cur = par ? s->fB : s->fA, **nxt = par ? s->fA : s->fB;
    void *b[3] = { s->f[9], s->f[10], s->f[11] };
    for (int dpar = 0; dpar < 2; ++dpar)
      if (!fused_prepare_3d (&s->fmaps[par][dpar], s->fp, s->nc, cur, cur + 3, s->dq[dpar], fb.dq_pitch, b, s->eps, s->pitch, s->ny,
                             s->nzl + 2, s->pitchM, s->nzm + 2, &err))
        return fail (FDTD_ERROR, "TMA descriptor set-up of the one-pass kernel failed: %s", err.c_str ());
    /* half-step kernels of the shell: the E phase reads H and E of cur; the H phase reads E of nxt and H of cur */
    void *fe[12], *fh[12];
    for (int i = 0; i < 12; ++i) fe[i] = fh[i] = s->f[i];
    for (int i = 0; i < 6; ++i) fe[i] = fh[i] = cur[i];
    for (int i = 0; i < 3; ++i) fh[i] = nxt[i];
    TmaMaps3D te, th;
    if (!tma_prepare_3d (&te, s->fp, s->nc, true, true, fe, s->eps, s->mu, s->pitch, s->ny, s->nzl + 2, s->pitchM, s->nzm + 2, &err)
        || !tma_prepare_3d (&th, s->fp, s->nc, true, true, fh, s->eps, s->mu, s->pitch, s->ny, s->nzl + 2, s->pitchM, s->nzm + 2, &err))
      return fail (FDTD_ERROR, "TMA descriptor set-up failed: %s", err.c_str ());
    memcpy (s->tma_pp[par].e, te.e, sizeof (te.e));
    memcpy (s->tma_pp[par].h, th.h, sizeof (th.h));
    s->tma_pp[par].valid = true;
  }
  s->fbox = fb;
  s->shell_ty = sty;
  s->fused_ok = true;
  return FDTD_OK;
}

static int fold_core (fdtd_solver *s, bool fold)
{
  const bool cplx = s->cfg.complex_values != 0, tfsf = s->cfg.use_tfsf != 0;
  const bool tmz = s->cfg.dims == 2;
  if (s->fp == 4) CUDA_TRY (launch_fold_aux<float> (make_params<float> (s), tmz, cplx, tfsf, fold, s->s_compute));
  else CUDA_TRY (launch_fold_aux<double> (make_params<double> (s), tmz, cplx, tfsf, fold, s->s_compute));
  s->launches++;
  s->folded = fold;
  return FDTD_OK;
}

/* D of the absorber quarters back from the compact pair into the D arrays (stream-ordered) */
static int release_dq (fdtd_solver *s)
{
  if (!s->dq_live) return FDTD_OK;
  void *main_d[3] = { s->f[6], s->f[7], s->f[8] };
  if (s->cfg.dims == 2) main_d[0] = s->f[8];
  CUDA_TRY (launch_dq_copy (main_d, s->dq[s->dq_par], s->cfg.dims == 2 ? 1 : 3, false, s->fbox, s->cell, s->ny, s->pitch, s->ps, s->s_compute));
  s->launches += s->cfg.dims == 2 ? 1 : 3;
  s->dq_live = false;
  s->dq_par = 0;
  return FDTD_OK;
}

/* before anything but the time loop reads or changes D, B, E, H, eps, mu or sigma (stream-ordered, no sync) */
static int ensure_unfolded (fdtd_solver *s)
{
  int rc = release_dq (s);
  if (rc) return rc;
  if (!s->folded) return FDTD_OK;
  return fold_core (s, false);
}

/* ------------------------------------------------------------------ time loop */

template <typename T>
static cudaError_t phase_3d (fdtd_solver *s, bool e_phase, const StepParams<T> &p, int kb, int ke, cudaStream_t S)
{
  const fdtd_config &c = s->cfg;
  const bool cplx = c.complex_values != 0, pml = c.use_pml != 0, tfsf = c.use_tfsf != 0, exact = c.exact_mode == 1;
  return launch_phase_3d_tma<T> (e_phase, p, s->tma, cplx, pml, tfsf, exact, s->core_on, s->mu_uniform, kb, ke, S);
}

template <typename T>
static int enqueue_steps (fdtd_solver *s, long long nsteps)
{
  const fdtd_config &c = s->cfg;
  StepParams<T> p = make_params<T> (s);
  const bool cplx = c.complex_values != 0, pml = c.use_pml != 0, tfsf = c.use_tfsf != 0, exact = c.exact_mode == 1;
  const bool multi = c.nranks > 1;
  cudaStream_t S = s->s_compute, C = s->s_comm;
  const bool prof = s->profiling && !multi;
  s->prof_steps = 0;
  if (prof)
  {
    while ((long long) s->prof_ev.size () < 4 * nsteps)
    {
      cudaEvent_t e;
      CUDA_TRY (cudaEventCreate (&e));
      s->prof_ev.push_back (e);
    }
    s->prof_steps = nsteps;
  }

  const bool mprof = s->profiling && multi;
  s->mprof_steps = 0;
  if (mprof)
  {
    while ((long long) s->mprof_ev.size () < 9 * nsteps)
    {
      cudaEvent_t e;
      CUDA_TRY (cudaEventCreate (&e));
      s->mprof_ev.push_back (e);
    }
    s->mprof_steps = nsteps;
  }
#define MPROF(i) do { if (mprof) CUDA_TRY (cudaEventRecord (s->mprof_ev[9 * n + (i)], S)); } while (0)
  if (multi)
  {
    /* make the low H halo valid for the first E half step */
    CUDA_TRY (cudaEventRecord (s->ev_hb, S));
    CUDA_TRY (cudaStreamWaitEvent (C, s->ev_hb, 0));
    int rc = exchange (s, true, s->f);
    if (rc) return rc;
    CUDA_TRY (cudaEventRecord (s->ev_hx, C));
  }

  /* Small grids are launch-bound (c1: six launches of 5-25 us per step; TMz 4096^2: two of 50 us): there one step is
   * captured as a CUDA graph and replayed, which takes the per-launch overhead out of the stream.  All steps must then
   * launch identical parameters, so the step index is read from device memory and advanced by the H kernel. */
  const long long cells_rank = (long long) s->nx * s->ny * s->nzl;
  if (!multi && !prof && nsteps >= 8 && (c.dims == 2 || cells_rank <= (4LL << 20)) && s->step_dev)
  {
    p.step_ptr = s->step_dev;
    CUDA_TRY (cudaMemsetAsync (s->step_dev, 0, sizeof (long long), S));
    auto one_step = [&] () -> int
    {
      int nl = 0;
      if (tfsf) { CUDA_TRY (launch_incident<T> (true, p, cplx, exact, S)); ++nl; }
      if (tfsf && s->face_tab) { CUDA_TRY (launch_face_tables<T> (true, p, cplx, S)); ++nl; }
      if (c.dims == 2) CUDA_TRY (launch_phase_tmz<T> (true, p, cplx, pml, tfsf, exact, s->core_on, s->eps_uniform, S));
      else CUDA_TRY (phase_3d<T> (s, true, p, 0, s->nzl, S));
      if (tfsf) { CUDA_TRY (launch_incident<T> (false, p, cplx, exact, S)); ++nl; }
      if (tfsf && s->face_tab) { CUDA_TRY (launch_face_tables<T> (false, p, cplx, S)); ++nl; }
      if (c.dims == 2) CUDA_TRY (launch_phase_tmz<T> (false, p, cplx, pml, tfsf, exact, s->core_on, s->mu_uniform, S));
      else CUDA_TRY (phase_3d<T> (s, false, p, 0, s->nzl, S));
      s->step_graph_launches = nl + 2;
      return FDTD_OK;
    };
    /* the first step runs as plain launches (first-use set-up of the kernels happens outside a capture) */
    int rc = one_step ();
    if (rc) return rc;
    s->launches += s->step_graph_launches;
    if (!s->step_graph || s->step_graph_src != s->src)
    {
      if (s->step_graph) { cudaGraphExecDestroy (s->step_graph); s->step_graph = NULL; }
      cudaGraph_t g = NULL;
      CUDA_TRY (cudaStreamBeginCapture (S, cudaStreamCaptureModeThreadLocal));
      rc = one_step ();
      cudaError_t ce = cudaStreamEndCapture (S, &g);
      if (rc) { if (g) cudaGraphDestroy (g); return rc; }
      if (ce != cudaSuccess) return fail (FDTD_ERROR, "capturing a time step failed: %s", cudaGetErrorString (ce));
      ce = cudaGraphInstantiate (&s->step_graph, g, 0);
      cudaGraphDestroy (g);
      if (ce != cudaSuccess) { s->step_graph = NULL; return fail (FDTD_ERROR, "cudaGraphInstantiate failed: %s", cudaGetErrorString (ce)); }
      s->step_graph_src = s->src;
    }
    for (long long n = 1; n < nsteps; ++n) CUDA_TRY (cudaGraphLaunch (s->step_graph, S));
    s->launches += (nsteps - 1) * s->step_graph_launches;
    return FDTD_OK;
  }

  /* TF/SF with face tables: the incident line (EInc from HInc, HInc from EInc) and the tables made from it depend on no
   * 3-D field.  They run on their own stream A, up to one half step ahead: the E tables of step n are written while
   * the H kernel of step n-1 runs (they wait only for the E kernel of step n-1, their last reader), the H tables while
   * the E kernel of step n runs.  A has the LOWEST priority: its blocks fill the SMs the running half-step kernel leaves
   * idle in its tail (at the highest priority they displaced its blocks and the kernel lost what the overlap gained).
   * On one stream the two small launches sat between the half-step kernels, 0.03 ms per half step of a 3.6 ms step;
   * half of that is left (profiles/r2_scaling.md). */
  cudaStream_t A = s->s_inc;
  const bool side = tfsf && s->face_tab && A && c.dims == 3;
  if (side)
  {
    CUDA_TRY (cudaEventRecord (s->ev_i0, S));
    CUDA_TRY (cudaStreamWaitEvent (A, s->ev_i0, 0));
  }

  for (long long n = 0; n < nsteps; ++n)
  {
    /* Scheme3D::performNSteps, Scheme3D.cpp:1986-2069 (SchemeTMz.cpp:1336-1427) */
    p.step = n; /* the parameter block is passed by value with every launch */
    MPROF (0);
    if (side)
    {
      CUDA_TRY (launch_incident<T> (true, p, cplx, exact, A));
      if (n > 0) CUDA_TRY (cudaStreamWaitEvent (A, s->ev_ed, 0));
      CUDA_TRY (launch_face_tables<T> (true, p, cplx, A));
      CUDA_TRY (cudaEventRecord (s->ev_fe, A));
      CUDA_TRY (cudaStreamWaitEvent (S, s->ev_fe, 0));
      s->launches += 2;
    }
    else
    {
      if (tfsf) { CUDA_TRY (launch_incident<T> (true, p, cplx, exact, S)); s->launches++; }
      if (tfsf && s->face_tab) { CUDA_TRY (launch_face_tables<T> (true, p, cplx, S)); s->launches++; }
    }
    MPROF (1);
    if (prof) CUDA_TRY (cudaEventRecord (s->prof_ev[4 * n + 0], S));
    if (c.dims == 2)
    {
      CUDA_TRY (launch_phase_tmz<T> (true, p, cplx, pml, tfsf, exact, s->core_on, s->eps_uniform, S)); s->launches++;
    }
    else if (!multi)
    {
      CUDA_TRY (phase_3d<T> (s, true, p, 0, s->nzl, S)); s->launches++;
    }
    else
    {
      CUDA_TRY (cudaStreamWaitEvent (S, s->ev_hx, 0));
      MPROF (2);
      CUDA_TRY (phase_3d<T> (s, true, p, 0, 1, S)); s->launches++;
      MPROF (3);
      CUDA_TRY (cudaEventRecord (s->ev_eb, S));
      CUDA_TRY (cudaStreamWaitEvent (C, s->ev_eb, 0));
      int rc = exchange (s, false, s->f);
      if (rc) return rc;
      CUDA_TRY (cudaEventRecord (s->ev_ex, C));
      if (s->nzl > 1) { CUDA_TRY (phase_3d<T> (s, true, p, 1, s->nzl, S)); s->launches++; }
      MPROF (4);
    }

    if (prof) CUDA_TRY (cudaEventRecord (s->prof_ev[4 * n + 1], S));
    if (side)
    {
      CUDA_TRY (cudaEventRecord (s->ev_ed, S));
      CUDA_TRY (launch_incident<T> (false, p, cplx, exact, A));
      if (n > 0) CUDA_TRY (cudaStreamWaitEvent (A, s->ev_hd, 0));
      CUDA_TRY (launch_face_tables<T> (false, p, cplx, A));
      CUDA_TRY (cudaEventRecord (s->ev_fh, A));
      CUDA_TRY (cudaStreamWaitEvent (S, s->ev_fh, 0));
      s->launches += 2;
    }
    else
    {
      if (tfsf) { CUDA_TRY (launch_incident<T> (false, p, cplx, exact, S)); s->launches++; }
      if (tfsf && s->face_tab) { CUDA_TRY (launch_face_tables<T> (false, p, cplx, S)); s->launches++; }
    }
    if (prof) CUDA_TRY (cudaEventRecord (s->prof_ev[4 * n + 2], S));
    if (c.dims == 2)
    {
      CUDA_TRY (launch_phase_tmz<T> (false, p, cplx, pml, tfsf, exact, s->core_on, s->mu_uniform, S)); s->launches++;
    }
    else if (!multi)
    {
      CUDA_TRY (phase_3d<T> (s, false, p, 0, s->nzl, S)); s->launches++;
    }
    else
    {
      MPROF (5);
      CUDA_TRY (cudaStreamWaitEvent (S, s->ev_ex, 0));
      MPROF (6);
      CUDA_TRY (phase_3d<T> (s, false, p, s->nzl - 1, s->nzl, S)); s->launches++;
      MPROF (7);
      CUDA_TRY (cudaEventRecord (s->ev_hb, S));
      CUDA_TRY (cudaStreamWaitEvent (C, s->ev_hb, 0));
      int rc = exchange (s, true, s->f);
      if (rc) return rc;
      CUDA_TRY (cudaEventRecord (s->ev_hx, C));
      if (s->nzl > 1) { CUDA_TRY (phase_3d<T> (s, false, p, 0, s->nzl - 1, S)); s->launches++; }
      MPROF (8);
    }
    if (side) CUDA_TRY (cudaEventRecord (s->ev_hd, S));
    if (prof) CUDA_TRY (cudaEventRecord (s->prof_ev[4 * n + 3], S));
  }
#undef MPROF
  if (side)
  {
    CUDA_TRY (cudaEventRecord (s->ev_i1, A));
    CUDA_TRY (cudaStreamWaitEvent (S, s->ev_i1, 0));
  }
  if (multi) CUDA_TRY (cudaStreamWaitEvent (S, s->ev_hx, 0));
  return FDTD_OK;
}

/*
 * The sub-boxes of the slab that the one-pass kernel does not cover, as a region table for one half-step launch
 * (StepParams::reg): the planes below and above its box, and inside its plane range the rows below and above.
 * `part`: 0 = all of it; 1 = only the plane that has to leave for a neighbour rank first (E phase: local plane 0 when
 * there is a rank below; H phase: the top plane when there is a rank above); 2 = all but that plane.
 */
template <typename T>
static void shell_regions (const fdtd_solver *s, bool e_phase, int part, StepParams<T> &p)
{
  const FusedBox &fb = s->fbox;
  const int sty = s->shell_ty;
  const int njt_all = (s->ny + sty - 1) / sty;
  const bool has_b = e_phase ? s->cfg.rank > 0 : s->cfg.rank + 1 < s->cfg.nranks;
  const int bplane = e_phase ? 0 : s->nzl - 1;
  p.nreg = 0;
  auto add = [&] (int jt0, int njt, int kb, int ke)
  {
    if (njt <= 0 || ke <= kb) return;
    typename StepParams<T>::Region &r = p.reg[p.nreg++];
    r.jt0 = jt0; r.njt = njt; r.kb = kb; r.ke = ke; r.cta0 = 0;
    r.kc = ke - kb < 16 ? ke - kb : 16;
  };
  if (part == 1)
  {
    if (has_b) add (0, njt_all, bplane, bplane + 1);
    return;
  }
  int lo_b = 0, lo_e = fb.fz0, hi_b = fb.fz1 + 1, hi_e = s->nzl;
  if (part == 2 && has_b)
  {
    if (e_phase) lo_b = 1; else hi_e = s->nzl - 1;
  }
  add (0, njt_all, lo_b, lo_e);
  add (0, njt_all, hi_b, hi_e);
  add (0, fb.fy0 / sty, fb.fz0, fb.fz1 + 1);
  add ((fb.fy1 + 1) / sty, njt_all - (fb.fy1 + 1) / sty, fb.fz0, fb.fz1 + 1);
}

/*
 * nsteps leapfrog steps with the one-pass kernel in the core.  Per step, on the compute stream:
 *   incident line E and H updates and both sets of TF/SF face tables (they depend on the 1-D line only),
 *   E half step of the shell (cur -> nxt), one-pass kernel (cur -> nxt, E and H), H half step of the shell,
 * then the roles of the two buffer sets are swapped.  With several ranks the plane a neighbour waits for is computed
 * first and travels on the communication stream while the rest runs, as in the two-pass schedule.
 */
template <typename T>
static int enqueue_steps_fused (fdtd_solver *s, long long nsteps)
{
  const fdtd_config &c = s->cfg;
  const bool cplx = c.complex_values != 0, tfsf = c.use_tfsf != 0;
  const bool multi = c.nranks > 1;
  cudaStream_t S = s->s_compute, C = s->s_comm;
  const bool prof = s->profiling && !multi;
  s->prof_steps = 0;
  if (prof)
  {
    while ((long long) s->prof_ev.size () < 4 * nsteps)
    {
      cudaEvent_t e;
      CUDA_TRY (cudaEventCreate (&e));
      s->prof_ev.push_back (e);
    }
    s->prof_steps = nsteps;
  }
  if (multi)
  {
    CUDA_TRY (cudaEventRecord (s->ev_hb, S));
    CUDA_TRY (cudaStreamWaitEvent (C, s->ev_hb, 0));
    int rc = exchange (s, true, s->f);
    if (rc) return rc;
    CUDA_TRY (cudaEventRecord (s->ev_hx, C));
  }
  if (!s->dq_live)
  {
    void *main_d[3] = { s->f[6], s->f[7], s->f[8] };
    if (c.dims == 2) main_d[0] = s->f[8];
    s->dq_par = 0;
    CUDA_TRY (launch_dq_copy (main_d, s->dq[0], c.dims == 2 ? 1 : 3, true, s->fbox, s->cell, s->ny, s->pitch, s->ps, S));
    s->launches += c.dims == 2 ? 1 : 3;
    s->dq_live = true;
  }
  for (long long n = 0; n < nsteps; ++n)
  {
    StepParams<T> p = make_params<T> (s); /* s->f are the current buffers */
    void **nxt = s->parity ? s->fA : s->fB;
    FusedBox fb = s->fbox;
    for (int a = 0; a < 3; ++a) fb.dq_out[a] = s->dq[s->dq_par ^ 1][a];
    const FusedMaps3D &fmap = s->fmaps[s->parity][s->dq_par];
    for (int i = 0; i < 6; ++i) p.fo[i] = (T *) nxt[i];
    p.step = n;
    if (c.dims == 2)
    {
      /* TMz: Ez half step of the rows outside the box, one-pass kernel, H half step of those rows.  The TF/SF
       * correction samples the incident line in the kernels: the E updates need HInc of the previous step and the H
       * updates EInc of this one, so the line's E update runs before and its H update after them. */
      if (tfsf) { CUDA_TRY (launch_incident<T> (true, p, cplx, false, S)); s->launches++; }
      if (prof) CUDA_TRY (cudaEventRecord (s->prof_ev[4 * n + 0], S));
      StepParams<T> pe = p;
      pe.nreg = 2;
      pe.reg[0].kb = 0; pe.reg[0].ke = s->fbox.fy0;
      pe.reg[1].kb = s->fbox.fy1 + 1; pe.reg[1].ke = s->ny;
      CUDA_TRY (launch_phase_tmz<T> (true, pe, cplx, true, tfsf, false, true, s->eps_uniform, S)); s->launches++;
      if (prof) CUDA_TRY (cudaEventRecord (s->prof_ev[4 * n + 1], S));
      CUDA_TRY (launch_step_core_tmz_fused<T> (p, fmap, fb, cplx, tfsf, s->eps_uniform, S)); s->launches++;
      if (prof) CUDA_TRY (cudaEventRecord (s->prof_ev[4 * n + 2], S));
      StepParams<T> ph = pe;
      ph.f[2] = (T *) nxt[2];
      CUDA_TRY (launch_phase_tmz<T> (false, ph, cplx, true, tfsf, false, true, true, S)); s->launches++;
      if (prof) CUDA_TRY (cudaEventRecord (s->prof_ev[4 * n + 3], S));
      if (tfsf) { CUDA_TRY (launch_incident<T> (false, p, cplx, false, S)); s->launches++; }
      for (int i = 0; i < 6; ++i) if (nxt[i]) s->f[i] = nxt[i];
      s->parity ^= 1;
      s->dq_par ^= 1;
      continue;
    }
    const TmaMaps3D &shell_maps = s->tma_pp[s->parity];
    if (tfsf)
    {
      CUDA_TRY (launch_incident<T> (true, p, cplx, false, S));
      CUDA_TRY (launch_face_tables<T> (true, p, cplx, S));
      CUDA_TRY (launch_incident<T> (false, p, cplx, false, S));
      CUDA_TRY (launch_face_tables<T> (false, p, cplx, S));
      s->launches += 4;
    }
    if (prof) CUDA_TRY (cudaEventRecord (s->prof_ev[4 * n + 0], S));

    /* E half step of the shell */
    StepParams<T> pe = p;
    if (multi)
    {
      CUDA_TRY (cudaStreamWaitEvent (S, s->ev_hx, 0));
      shell_regions<T> (s, true, 1, pe);
      if (pe.nreg) { CUDA_TRY (launch_phase_3d_tma<T> (true, pe, shell_maps, cplx, true, tfsf, false, true, true, 0, 0, S)); s->launches++; }
      CUDA_TRY (cudaEventRecord (s->ev_eb, S));
      CUDA_TRY (cudaStreamWaitEvent (C, s->ev_eb, 0));
      int rc = exchange (s, false, nxt);
      if (rc) return rc;
      CUDA_TRY (cudaEventRecord (s->ev_ex, C));
    }
    shell_regions<T> (s, true, multi ? 2 : 0, pe);
    if (pe.nreg) { CUDA_TRY (launch_phase_3d_tma<T> (true, pe, shell_maps, cplx, true, tfsf, false, true, true, 0, 0, S)); s->launches++; }
    if (prof) CUDA_TRY (cudaEventRecord (s->prof_ev[4 * n + 1], S));

    /* the core: E and H in one pass */
    CUDA_TRY (launch_step_core_fused<T> (p, fmap, fb, cplx, tfsf, S));
    s->launches++;
    if (prof) CUDA_TRY (cudaEventRecord (s->prof_ev[4 * n + 2], S));

    /* H half step of the shell: reads E of nxt, H of cur */
    StepParams<T> ph = p;
    for (int i = 0; i < 3; ++i) ph.f[i] = (T *) nxt[i];
    if (multi)
    {
      CUDA_TRY (cudaStreamWaitEvent (S, s->ev_ex, 0));
      shell_regions<T> (s, false, 1, ph);
      if (ph.nreg) { CUDA_TRY (launch_phase_3d_tma<T> (false, ph, shell_maps, cplx, true, tfsf, false, true, true, 0, 0, S)); s->launches++; }
      CUDA_TRY (cudaEventRecord (s->ev_hb, S));
      CUDA_TRY (cudaStreamWaitEvent (C, s->ev_hb, 0));
      int rc = exchange (s, true, nxt);
      if (rc) return rc;
      CUDA_TRY (cudaEventRecord (s->ev_hx, C));
    }
    shell_regions<T> (s, false, multi ? 2 : 0, ph);
    if (ph.nreg) { CUDA_TRY (launch_phase_3d_tma<T> (false, ph, shell_maps, cplx, true, tfsf, false, true, true, 0, 0, S)); s->launches++; }
    if (prof) CUDA_TRY (cudaEventRecord (s->prof_ev[4 * n + 3], S));

    /* nxt holds the fields of the next step */
    for (int i = 0; i < 6; ++i) s->f[i] = nxt[i];
    s->parity ^= 1;
    s->dq_par ^= 1;
  }
  if (multi) CUDA_TRY (cudaStreamWaitEvent (S, s->ev_hx, 0));
  return FDTD_OK;
}

static int run_async (fdtd_solver *s, long long nsteps)
{
  if (nsteps < 0) return fail (FDTD_INVALID_ARGUMENT, "nsteps < 0");
  CUDA_TRY (cudaSetDevice (s->cfg.device));
  if (nsteps == 0) { if (!s->runs_pending) s->last_ms = 0; return FDTD_OK; }
  int rc = s->fp == 4 ? upload_source<float> (s, nsteps) : upload_source<double> (s, nsteps);
  if (rc) return rc;
  /* every queued run has its own (start, stop) events, so runs can be queued back to back */
  while ((int) s->run_ev.size () < 2 * (s->runs_pending + 1))
  {
    cudaEvent_t e;
    CUDA_TRY (cudaEventCreate (&e));
    s->run_ev.push_back (e);
  }
  CUDA_TRY (cudaEventRecord (s->run_ev[2 * s->runs_pending], s->s_compute));
  /* one pass per step in the core when the geometry allows it and mu is homogeneous (kernels3d_fused.cu) */
  s->fused_run = s->fused_ok && s->mu_uniform;
  if (!s->fused_run && ((rc = release_dq (s)) || (rc = normalize_parity (s)))) return rc; /* the two-pass tensor maps describe the first allocation */
  if (s->core_on && !s->folded && (rc = fold_core (s, true))) return rc;
  if (s->fused_run) rc = s->fp == 4 ? enqueue_steps_fused<float> (s, nsteps) : enqueue_steps_fused<double> (s, nsteps);
  else rc = s->fp == 4 ? enqueue_steps<float> (s, nsteps) : enqueue_steps<double> (s, nsteps);
  if (rc) return rc;
  CUDA_TRY (cudaEventRecord (s->run_ev[2 * s->runs_pending + 1], s->s_compute));
  s->t += nsteps;
  s->runs_pending++;
  return FDTD_OK;
}

static int sync_run (fdtd_solver *s)
{
  CUDA_TRY (cudaSetDevice (s->cfg.device));
  CUDA_TRY (cudaStreamSynchronize (s->s_compute));
  s->src_ring_head = s->src_ring_used = 0;
  if (s->runs_pending)
  {
    /* the time of the run queued last; with several runs queued, also what they took together */
    const int last = s->runs_pending - 1;
    CUDA_TRY (cudaEventElapsedTime (&s->last_ms, s->run_ev[2 * last], s->run_ev[2 * last + 1]));
    s->runs_pending = 0;
    for (int i = 0; i < 8; ++i) s->mprof_ms[i] = 0;
    for (long long n = 0; n < s->mprof_steps && !s->fused_run; ++n)
      for (int i = 0; i < 8; ++i)
      {
        float d = 0;
        CUDA_TRY (cudaEventElapsedTime (&d, s->mprof_ev[9 * n + i], s->mprof_ev[9 * n + i + 1]));
        s->mprof_ms[i] += d;
      }
    s->prof_e_ms = s->prof_h_ms = s->prof_f_ms = 0;
    for (long long n = 0; n < s->prof_steps; ++n)
    {
      /* two-pass: [0,1] E kernel, [2,3] H kernel; one-pass: [0,1] E shell, [1,2] core kernel, [2,3] H shell */
      float a = 0, b = 0, f = 0;
      CUDA_TRY (cudaEventElapsedTime (&a, s->prof_ev[4 * n + 0], s->prof_ev[4 * n + 1]));
      CUDA_TRY (cudaEventElapsedTime (&b, s->prof_ev[4 * n + 2], s->prof_ev[4 * n + 3]));
      if (s->fused_run) CUDA_TRY (cudaEventElapsedTime (&f, s->prof_ev[4 * n + 1], s->prof_ev[4 * n + 2]));
      s->prof_e_ms += a;
      s->prof_h_ms += b;
      s->prof_f_ms += f;
    }
  }
  return FDTD_OK;
}

/* ------------------------------------------------------------------ transfers */

struct FieldInfo
{
  void *dev;
  int cell;          /* bytes per cell */
  int nx, ny;        /* logical row length and rows per plane */
  int nzl_planes;    /* owned planes */
  int nzg_planes;    /* global planes */
  int pitch;
  long long ps;
  bool is_inc;
};

static int field_info (fdtd_solver *s, int field, FieldInfo *fi)
{
  memset (fi, 0, sizeof (*fi));
  const fdtd_config &c = s->cfg;
  if (field >= FDTD_EX && field <= FDTD_BZ)
  {
    if (!s->f[field]) return fail (FDTD_INVALID_ARGUMENT, "field %d is not part of this configuration", field);
    fi->dev = s->f[field]; fi->cell = s->cell; fi->nx = s->nx; fi->ny = s->ny;
    fi->nzl_planes = s->nzl; fi->nzg_planes = s->nzg; fi->pitch = s->pitch; fi->ps = s->ps;
    return FDTD_OK;
  }
  if (field == FDTD_EPS || field == FDTD_MU)
  {
    fi->dev = field == FDTD_EPS ? s->eps : s->mu; fi->cell = s->fp; fi->nx = s->nx + 1; fi->ny = s->ny + 1;
    fi->nzl_planes = s->nzm; fi->nzg_planes = c.dims == 3 ? s->nzg + 1 : 1; fi->pitch = s->pitchM; fi->ps = s->psM;
    return FDTD_OK;
  }
  if (field == FDTD_EINC || field == FDTD_HINC)
  {
    if (!s->einc) return fail (FDTD_INVALID_ARGUMENT, "incident line requested without use_tfsf");
    fi->dev = field == FDTD_EINC ? s->einc : s->hinc; fi->cell = s->cell; fi->is_inc = true;
    return FDTD_OK;
  }
  return fail (FDTD_INVALID_ARGUMENT, "unknown field id %d", field);
}

static int transfer (fdtd_solver *s, int field, void *host, int order, long long z0, long long nz, bool to_device)
{
  if (!host) return fail (FDTD_INVALID_ARGUMENT, "NULL host pointer");
  CUDA_TRY (cudaSetDevice (s->cfg.device));
  int rc = sync_run (s);
  if (rc) return rc;
  FieldInfo fi;
  rc = field_info (s, field, &fi);
  if (rc) return rc;
  if (to_device && field == FDTD_MU) s->mu_uniform = false;
  if (to_device && field == FDTD_EPS) s->eps_uniform = false;
  if (to_device && (field == FDTD_MU || field == FDTD_EPS)) drop_step_graph (s);
  /* uploads of anything the folded representation depends on, and reads of D / B, need the plain arrays */
  if (!fi.is_inc && (to_device || (field >= FDTD_DX && field <= FDTD_BZ)) && (rc = ensure_unfolded (s))) return rc;
  cudaStream_t S = s->s_compute;
  if (fi.is_inc)
  {
    const size_t bytes = (size_t) s->ninc * s->cell;
    if (to_device) CUDA_TRY (cudaMemcpyAsync (fi.dev, host, bytes, cudaMemcpyHostToDevice, S));
    else CUDA_TRY (cudaMemcpyAsync (host, fi.dev, bytes, cudaMemcpyDeviceToHost, S));
    CUDA_TRY (cudaStreamSynchronize (S));
    return FDTD_OK;
  }
  if (s->cfg.dims == 2) { z0 = 0; nz = 1; }
  if (z0 < 0 || nz <= 0 || z0 + nz > fi.nzg_planes)
    return fail (FDTD_INVALID_ARGUMENT, "plane range [%lld, %lld) outside [0, %d)", z0, z0 + nz, fi.nzg_planes);
  /* device plane index of buffer plane k: (z0 + k - kz0) + 1; owned planes are [1, nzl_planes + 1) */
  const int kofs = (int) (z0 - s->kz0) + 1;
  const int kmin = 1, kmax = fi.nzl_planes + 1;
  char *h = (char *) host;
  if (order == FDTD_ORDER_REFERENCE)
  {
    /* Only the planes of [z0, z0 + nz) that this rank owns move, in either direction: the host block may be a
     * whole-grid buffer that other ranks fill beside this one, so nothing outside the slab is read or written.
     * Buffer planes [ka, kb) are the owned ones; they are staged as a compact k-fastest block of kb - ka planes. */
    long long ka = s->kz0 - z0, kb = (long long) s->kz0 + fi.nzl_planes - z0;
    if (ka < 0) ka = 0;
    if (kb > nz) kb = nz;
    if (kb <= ka) return FDTD_OK;
    const long long nzo = kb - ka;
    const bool whole = nzo == nz;
    /* chunks of i-rows of the k-fastest host block */
    const size_t row_bytes = (size_t) fi.ny * nzo * fi.cell;   /* staged */
    const size_t host_row = (size_t) fi.ny * nz * fi.cell;     /* in the caller's block */
    long long rows_per_chunk = (long long) (s->staging_bytes / row_bytes);
    if (rows_per_chunk < 1) return fail (FDTD_INVALID_ARGUMENT, "one i-row (%zu bytes) exceeds the staging buffer", row_bytes);
    const int kofs_o = kofs + (int) ka;
    for (long long ia = 0; ia < fi.nx; ia += rows_per_chunk)
    {
      const long long ib = ia + rows_per_chunk < fi.nx ? ia + rows_per_chunk : fi.nx;
      const size_t bytes = (size_t) (ib - ia) * row_bytes;
      char *hp = h + (size_t) ia * host_row + (size_t) ka * fi.cell;
      const size_t w = (size_t) nzo * fi.cell, hpitch = (size_t) nz * fi.cell, lines = (size_t) (ib - ia) * fi.ny;
      if (to_device)
      {
        if (whole) CUDA_TRY (cudaMemcpyAsync (s->staging, hp, bytes, cudaMemcpyHostToDevice, S));
        else CUDA_TRY (cudaMemcpy2DAsync (s->staging, w, hp, hpitch, w, lines, cudaMemcpyHostToDevice, S));
        CUDA_TRY (launch_reorder (true, fi.dev, s->staging, fi.cell, (int) ia, (int) ib, fi.ny, (int) nzo, fi.pitch, fi.ps, kofs_o, kmin, kmax, S));
      }
      else
      {
        CUDA_TRY (launch_reorder (false, fi.dev, s->staging, fi.cell, (int) ia, (int) ib, fi.ny, (int) nzo, fi.pitch, fi.ps, kofs_o, kmin, kmax, S));
        if (whole) CUDA_TRY (cudaMemcpyAsync (hp, s->staging, bytes, cudaMemcpyDeviceToHost, S));
        else CUDA_TRY (cudaMemcpy2DAsync (hp, hpitch, s->staging, w, w, lines, cudaMemcpyDeviceToHost, S));
      }
      s->launches++;
      CUDA_TRY (cudaStreamSynchronize (S));
    }
  }
  else if (order == FDTD_ORDER_NATIVE)
  {
    /* x-fastest host block: plane by plane, rows of nx cells */
    for (long long k = 0; k < nz; ++k)
    {
      const int kd = (int) k + kofs;
      if (kd < kmin || kd >= kmax) continue;
      char *hp = h + (size_t) k * fi.ny * fi.nx * fi.cell;
      char *dp = (char *) fi.dev + (size_t) kd * fi.ps * fi.cell;
      if (to_device)
        CUDA_TRY (cudaMemcpy2DAsync (dp, (size_t) fi.pitch * fi.cell, hp, (size_t) fi.nx * fi.cell, (size_t) fi.nx * fi.cell, fi.ny, cudaMemcpyHostToDevice, S));
      else
        CUDA_TRY (cudaMemcpy2DAsync (hp, (size_t) fi.nx * fi.cell, dp, (size_t) fi.pitch * fi.cell, (size_t) fi.nx * fi.cell, fi.ny, cudaMemcpyDeviceToHost, S));
    }
    CUDA_TRY (cudaStreamSynchronize (S));
  }
  else
    return fail (FDTD_INVALID_ARGUMENT, "unknown order %d", order);
  if (to_device && (field == FDTD_MU || field == FDTD_EPS))
  {
    /* An uploaded material may still be homogeneous (the reference's own mu is 1 everywhere, and in TMz so is eps):
     * one scan of the owned planes decides whether the kernels may take the value from the parameter block. */
    const char *base = (const char *) fi.dev + (size_t) fi.ps * fi.cell;
    const long long nrows = (long long) fi.ny * fi.nzl_planes;
    if (s->fp == 4) CUDA_TRY (launch_uniform_check<float> ((const float *) base, fi.nx, nrows, fi.pitch, s->red_partials, S));
    else CUDA_TRY (launch_uniform_check<double> ((const double *) base, fi.nx, nrows, fi.pitch, s->red_partials, S));
    s->launches++;
    double res[2] = { 1.0, 0.0 };
    CUDA_TRY (cudaMemcpyAsync (res, s->red_partials, sizeof (res), cudaMemcpyDeviceToHost, S));
    CUDA_TRY (cudaStreamSynchronize (S));
    const bool uni = res[0] == 0.0;
    if (field == FDTD_MU) { s->mu_uniform = uni; s->mu_value = res[1]; }
    else { s->eps_uniform = uni; s->eps_value = res[1]; }
  }
  return FDTD_OK;
}

static int validate (const fdtd_config *c)
{
  if (!c) return fail (FDTD_INVALID_ARGUMENT, "NULL config");
  if (c->abi_version != FDTD3D_B200_ABI_VERSION) return fail (FDTD_INVALID_ARGUMENT, "abi_version %d, library is %d", c->abi_version, FDTD3D_B200_ABI_VERSION);
  if (c->dims != 2 && c->dims != 3) return fail (FDTD_INVALID_ARGUMENT, "dims must be 2 (TMz) or 3");
  if (c->exact_mode < 0 || c->exact_mode > 2) return fail (FDTD_INVALID_ARGUMENT, "exact_mode must be 0 (fast), 1 (exact) or 2 (fast, D/B kept everywhere)");
  if (c->value_type != 4 && c->value_type != 8) return fail (FDTD_INVALID_ARGUMENT, "value_type must be 4 (float) or 8 (double); long double has no device type");
  for (int a = 0; a < c->dims; ++a)
  {
    /* 65535: rows and planes index CUDA grid dimensions y / z in the layout-conversion and init kernels */
    if (c->size[a] < 2 || c->size[a] > 65534) return fail (FDTD_INVALID_ARGUMENT, "size[%d] = %lld outside [2, 65534]", a, (long long) c->size[a]);
    if (c->use_tfsf && (c->tfsf_size[a] < 1 || 2 * c->tfsf_size[a] >= c->size[a]))
      return fail (FDTD_INVALID_ARGUMENT, "tfsf_size[%d] = %lld does not leave a total-field region", a, (long long) c->tfsf_size[a]);
    if (!c->use_tfsf && (c->source_pos[a] < 0 || c->source_pos[a] >= c->size[a]))
      return fail (FDTD_INVALID_ARGUMENT, "source_pos[%d] = %lld outside the grid", a, (long long) c->source_pos[a]);
  }
  if (c->dims == 2 && c->size[2] != 1) return fail (FDTD_INVALID_ARGUMENT, "size[2] must be 1 for dims == 2");
  if (!(c->dx > 0) || !(c->dt > 0)) return fail (FDTD_INVALID_ARGUMENT, "dx and dt must be positive");
  if (c->use_tfsf && !(c->rel_phase_velocity > 0)) return fail (FDTD_INVALID_ARGUMENT, "rel_phase_velocity must be positive");
  if (c->nranks < 1 || c->rank < 0 || c->rank >= c->nranks) return fail (FDTD_INVALID_ARGUMENT, "bad rank %d of %d", c->rank, c->nranks);
  if (c->nranks > 1 && c->dims != 3) return fail (FDTD_INVALID_ARGUMENT, "z-slab decomposition needs dims == 3");
  if (c->nranks > 1 && c->size[2] / c->nranks < 1) return fail (FDTD_INVALID_ARGUMENT, "fewer z planes than ranks");
  return FDTD_OK;
}

} /* namespace fdtd */

/* ==================================================================== C ABI */

extern "C"
{

const char *fdtd_last_error (void) { return g_last_error.c_str (); }

int fdtd_device_info (int device, char *buf, int64_t buflen)
{
  if (!buf || buflen <= 0) return fail (FDTD_INVALID_ARGUMENT, "NULL buffer");
  int n = 0;
  if (cudaGetDeviceCount (&n) != cudaSuccess || n == 0) return fail (FDTD_NO_DEVICE, "no CUDA device");
  cudaDeviceProp prop;
  CUDA_TRY (cudaGetDeviceProperties (&prop, device));
  int memclk = 0, buswidth = 0;
  cudaDeviceGetAttribute (&memclk, cudaDevAttrMemoryClockRate, device);
  cudaDeviceGetAttribute (&buswidth, cudaDevAttrGlobalMemoryBusWidth, device);
  /* same "theoretical bandwidth" formula as cudaInfo, CudaInterface.cu:955-957 */
  snprintf (buf, (size_t) buflen, "%s: sm_%d%d, %d SMs, %.1f GiB, L2 %.0f MiB, theoretical bandwidth %.0f GB/s",
            prop.name, prop.major, prop.minor, prop.multiProcessorCount, prop.totalGlobalMem / 1073741824.0,
            prop.l2CacheSize / 1048576.0, 2.0 * memclk * (buswidth / 8) / 1.0e6);
  return FDTD_OK;
}

int fdtd_create (const fdtd_config *cfg, fdtd_handle *out)
{
  if (!out) return fail (FDTD_INVALID_ARGUMENT, "NULL output handle");
  *out = NULL;
  int rc = validate (cfg);
  if (rc) return rc;
  int ndev = 0;
  if (cudaGetDeviceCount (&ndev) != cudaSuccess || ndev == 0)
    return fail (FDTD_NO_DEVICE, "no CUDA device available: fdtd3d_b200 has no CPU fallback");
  if (cfg->device < 0 || cfg->device >= ndev) return fail (FDTD_INVALID_ARGUMENT, "device %d of %d", cfg->device, ndev);
  CUDA_TRY (cudaSetDevice (cfg->device));

  fdtd_solver *s = new fdtd_solver ();
  s->cfg = *cfg;
  s->nc = cfg->complex_values ? 2 : 1;
  s->fp = cfg->value_type;
  s->cell = s->fp * s->nc;
  s->nx = (int) cfg->size[0]; s->ny = (int) cfg->size[1]; s->nzg = cfg->dims == 3 ? (int) cfg->size[2] : 1;
  if (cfg->dims == 2) { s->cfg.size[2] = 1; s->cfg.tfsf_size[2] = 0; s->cfg.pml_size[2] = 0; s->cfg.source_pos[2] = 0; }
  /* slab partition: core = Nz / n, last rank takes the remainder (ParallelYeeGridLayout.cpp:24-36) */
  const int core = s->nzg / cfg->nranks;
  s->kz0 = cfg->rank * core;
  s->nzl = cfg->rank == cfg->nranks - 1 ? s->nzg - s->kz0 : core;
  s->nzm = cfg->dims == 3 ? s->nzl + 1 : 1;
  s->pitch = (int) round_up (s->nx, 128 / s->cell);
  s->pitchM = (int) round_up (s->nx + 1, 128 / s->fp);
  s->ps = (long long) s->pitch * s->ny;
  s->psM = (long long) s->pitchM * (s->ny + 1);
  s->field_cells = s->ps * (s->nzl + 2);
  s->mat_vals = s->psM * (s->nzm + 2);
  s->ninc = cfg->use_tfsf ? 100LL * (s->nx + s->ny + (cfg->dims == 3 ? s->nzg : 0)) : 0;
  s->t = cfg->time_step_start;
  s->fused_policy = -1;
  for (int a = 0; a < 3; ++a) s->sigma[a].assign ((size_t) (a < cfg->dims ? cfg->size[a] : 1) + 1, 0.0);

#define CREATE_TRY(expr) \
  do { cudaError_t e_ = (expr); if (e_ != cudaSuccess) { fail (FDTD_ERROR, "%s failed: %s", #expr, cudaGetErrorString (e_)); fdtd_destroy (s); return FDTD_ERROR; } } while (0)

  {
    /* The halo exchange runs at the highest stream priority: an interior launch fills every SM for more than a
     * millisecond, and the send/recv kernel of the plane just computed needs an SM of its own; at equal priority it was
     * dispatched only when the interior launch had no blocks left to place, and the next half step then waited for it
     * (profiles/r2_scaling.md, step breakdown at 8 ranks). */
    int prio_least = 0, prio_greatest = 0;
    CREATE_TRY (cudaDeviceGetStreamPriorityRange (&prio_least, &prio_greatest));
    /* three levels where the device has them: exchange > half-step kernels > incident line and face tables */
    const int prio_mid = prio_least - 1 > prio_greatest ? prio_least - 1 : prio_least;
    CREATE_TRY (cudaStreamCreateWithPriority (&s->s_compute, cudaStreamNonBlocking, prio_mid));
    CREATE_TRY (cudaStreamCreateWithPriority (&s->s_comm, cudaStreamNonBlocking, prio_greatest));
    if (cfg->use_tfsf && cfg->dims == 3)
    {
      CREATE_TRY (cudaStreamCreateWithPriority (&s->s_inc, cudaStreamNonBlocking, prio_least));
      cudaEvent_t *ie[] = { &s->ev_i0, &s->ev_i1, &s->ev_fe, &s->ev_fh, &s->ev_ed, &s->ev_hd };
      for (cudaEvent_t *e : ie) CREATE_TRY (cudaEventCreateWithFlags (e, cudaEventDisableTiming));
    }
  }
  s->src_ring_bytes = (size_t) 1 << 20; /* 65536 complex-double samples queued at a time */
  CREATE_TRY (cudaHostAlloc ((void **) &s->src_ring, s->src_ring_bytes, cudaHostAllocDefault));
  CREATE_TRY (cudaEventCreateWithFlags (&s->ev_eb, cudaEventDisableTiming));
  CREATE_TRY (cudaEventCreateWithFlags (&s->ev_ex, cudaEventDisableTiming));
  CREATE_TRY (cudaEventCreateWithFlags (&s->ev_hb, cudaEventDisableTiming));
  CREATE_TRY (cudaEventCreateWithFlags (&s->ev_hx, cudaEventDisableTiming));

  const bool is3 = cfg->dims == 3;
  for (int f = 0; f < 12; ++f)
  {
    const int comp = f % 3;
    const bool aux = f >= FDTD_DX;
    bool present = is3 ? true : ((f < FDTD_HX || (f >= FDTD_DX && f < FDTD_BX)) ? comp == 2 : comp < 2); /* TMz: Ez, Hx, Hy */
    if (aux && !cfg->use_pml) present = false;
    if (!present) continue;
    CREATE_TRY (cudaMalloc (&s->f[f], (size_t) s->field_cells * s->cell));
    CREATE_TRY (cudaMemsetAsync (s->f[f], 0, (size_t) s->field_cells * s->cell, s->s_compute));
  }
  CREATE_TRY (cudaMalloc (&s->eps, (size_t) s->mat_vals * s->fp));
  CREATE_TRY (cudaMalloc (&s->mu, (size_t) s->mat_vals * s->fp));
  if (cfg->use_tfsf)
  {
    CREATE_TRY (cudaMalloc (&s->einc, (size_t) s->ninc * s->cell));
    CREATE_TRY (cudaMalloc (&s->hinc, (size_t) s->ninc * s->cell));
    CREATE_TRY (cudaMemsetAsync (s->einc, 0, (size_t) s->ninc * s->cell, s->s_compute));
    CREATE_TRY (cudaMemsetAsync (s->hinc, 0, (size_t) s->ninc * s->cell, s->s_compute));
  }
  long long tlen = 0;
  for (int a = 0; a < 3; ++a) { long long l = round_up ((long long) s->sigma[a].size (), 128 / s->fp) + 128 / s->fp; if (l > tlen) tlen = l; }
  if (tlen < s->pitchM + 128 / s->fp) tlen = s->pitchM + 128 / s->fp;
  s->tab_stride = (size_t) tlen * s->fp;
  CREATE_TRY (cudaMalloc (&s->tab_mem, (size_t) N_TABLES * s->tab_stride));
  s->staging_bytes = (size_t) 64 << 20;
  {
    /* one i-row of the largest field must fit */
    const size_t need = (size_t) (s->ny + 1) * (s->nzg + 1) * (s->cell > s->fp ? s->cell : s->fp);
    if (need > s->staging_bytes) s->staging_bytes = need;
  }
  CREATE_TRY (cudaMalloc (&s->staging, s->staging_bytes));
  CREATE_TRY (cudaMalloc ((void **) &s->red_partials, sizeof (double) * (size_t) sumsq_partial_count ()));
  CREATE_TRY (cudaMalloc ((void **) &s->red_out, sizeof (double)));
  CREATE_TRY (cudaMalloc ((void **) &s->step_dev, sizeof (long long)));
#undef CREATE_TRY

  for (int i = 0; i < 6; ++i) s->fA[i] = s->f[i];
  /* kernel selection; the TF/SF face tables of the TMA kernels */
  s->use_tma = false;
  s->kchunk = 8; /* 2D: rows marched per thread (profiles/r1_kchunk_sweep.log) */
  if (is3)
  {
    s->use_tma = true;
    if (s->use_tma && cfg->use_tfsf)
    {
      long long m = (long long) s->nx * s->ny;
      if ((long long) s->nx * s->nzl > m) m = (long long) s->nx * s->nzl;
      if ((long long) s->ny * s->nzl > m) m = (long long) s->ny * s->nzl;
      s->face_stride = m;
      cudaError_t fe = cudaMalloc (&s->face_tab, (size_t) 24 * m * s->cell);
      if (fe == cudaSuccess) fe = cudaMemsetAsync (s->face_tab, 0, (size_t) 24 * m * s->cell, s->s_compute);
      if (fe != cudaSuccess)
      {
        fail (FDTD_ERROR, "allocating the TF/SF face tables failed: %s", cudaGetErrorString (fe));
        fdtd_destroy (s);
        return FDTD_ERROR;
      }
    }
  }

  /* fast mode with UPML skips D / B where sigma = 0; exact_mode 2 keeps the reference's formulation everywhere */
  s->core_allowed = cfg->use_pml && cfg->exact_mode == 0;
  update_core (s);
  if ((rc = configure_3d (s, true)))
  {
    fdtd_destroy (s);
    return rc;
  }

  *out = s;
  if ((rc = fdtd_fill (s, FDTD_EPS, 1.0, 0.0)) || (rc = fdtd_fill (s, FDTD_MU, 1.0, 0.0)) || (rc = build_tables_any (s)))
  {
    fdtd_destroy (s);
    *out = NULL;
    return rc;
  }
  return FDTD_OK;
}

int fdtd_destroy (fdtd_handle s)
{
  if (!s) return FDTD_OK;
  cudaSetDevice (s->cfg.device);
  if (s->s_compute) cudaStreamSynchronize (s->s_compute);
  if (s->s_comm) cudaStreamSynchronize (s->s_comm);
  if (s->comm)
  {
    const NcclApi *nccl = nccl_api (NULL);
    if (nccl) nccl->CommDestroy (s->comm);
  }
  for (int f = 0; f < 6; ++f) { if (s->fA[f]) s->f[f] = s->fA[f]; if (s->fB[f]) cudaFree (s->fB[f]); }
  for (int d = 0; d < 2; ++d) for (int a = 0; a < 3; ++a) if (s->dq[d][a]) cudaFree (s->dq[d][a]);
  for (int f = 0; f < 12; ++f) if (s->f[f]) cudaFree (s->f[f]);
  drop_step_graph (s);
  void *ptrs[] = { s->step_dev, s->face_tab, s->eps, s->mu, s->einc, s->hinc, s->tab_mem, s->src, s->staging, s->red_partials, s->red_out };
  for (void *p : ptrs) if (p) cudaFree (p);
  cudaEvent_t evs[] = { s->ev_eb, s->ev_ex, s->ev_hb, s->ev_hx, s->ev_i0, s->ev_i1, s->ev_fe, s->ev_fh, s->ev_ed, s->ev_hd };
  for (cudaEvent_t e : evs) if (e) cudaEventDestroy (e);
  for (cudaEvent_t e : s->run_ev) cudaEventDestroy (e);
  if (s->src_ring) cudaFreeHost (s->src_ring);
  for (cudaEvent_t e : s->prof_ev) cudaEventDestroy (e);
  for (cudaEvent_t e : s->mprof_ev) cudaEventDestroy (e);
  if (s->s_compute) cudaStreamDestroy (s->s_compute);
  if (s->s_comm) cudaStreamDestroy (s->s_comm);
  if (s->s_inc) cudaStreamDestroy (s->s_inc);
  delete s;
  return FDTD_OK;
}

int fdtd_set_sigma (fdtd_handle s, int axis, const void *profile, int64_t count)
{
  if (!s || !profile) return fail (FDTD_INVALID_ARGUMENT, "NULL argument");
  if (axis < 0 || axis >= s->cfg.dims) return fail (FDTD_INVALID_ARGUMENT, "axis %d", axis);
  if (count != (int64_t) s->sigma[axis].size ()) return fail (FDTD_INVALID_ARGUMENT, "sigma profile along axis %d needs %zu values, got %lld", axis, s->sigma[axis].size (), (long long) count);
  CUDA_TRY (cudaSetDevice (s->cfg.device));
  int rc = sync_run (s);
  if (rc) return rc;
  if ((rc = ensure_unfolded (s))) return rc; /* with the core box of the old profiles */
  for (int64_t i = 0; i < count; ++i)
    s->sigma[axis][(size_t) i] = s->fp == 4 ? (double) ((const float *) profile)[i] : ((const double *) profile)[i];
  update_core (s);
  if ((rc = configure_3d (s, false))) return rc;
  return build_tables_any (s);
}

int fdtd_upload (fdtd_handle s, int field, const void *host, int order, int64_t z0, int64_t nz)
{
  if (!s) return fail (FDTD_INVALID_ARGUMENT, "NULL handle");
  return transfer (s, field, (void *) host, order, z0, nz, true);
}

int fdtd_download (fdtd_handle s, int field, void *host, int order, int64_t z0, int64_t nz)
{
  if (!s) return fail (FDTD_INVALID_ARGUMENT, "NULL handle");
  return transfer (s, field, host, order, z0, nz, false);
}

int fdtd_fill (fdtd_handle s, int field, double re, double im)
{
  if (!s) return fail (FDTD_INVALID_ARGUMENT, "NULL handle");
  CUDA_TRY (cudaSetDevice (s->cfg.device));
  int rc = sync_run (s);
  if (rc) return rc;
  FieldInfo fi;
  rc = field_info (s, field, &fi);
  if (rc) return rc;
  const bool material = field == FDTD_EPS || field == FDTD_MU;
  if (!fi.is_inc && (rc = ensure_unfolded (s))) return rc;
  long long count = fi.is_inc ? s->ninc : (material ? s->mat_vals : s->field_cells);
  alignas (16) char pat[16];
  memset (pat, 0, sizeof (pat));
  if (s->fp == 4)
  {
    float v[2] = { (float) re, (float) im };
    memcpy (pat, v, material || s->nc == 1 ? 4 : 8);
  }
  else
  {
    double v[2] = { re, im };
    memcpy (pat, v, material || s->nc == 1 ? 8 : 16);
  }
  if (field == FDTD_MU || field == FDTD_EPS)
  {
    /* the value as the array holds it (FPValue) */
    const double v = s->fp == 4 ? (double) (float) re : re;
    if (field == FDTD_MU) { s->mu_uniform = true; s->mu_value = v; }
    else { s->eps_uniform = true; s->eps_value = v; }
    drop_step_graph (s);
  }
  CUDA_TRY (launch_fill (fi.dev, count, fi.cell, pat, s->s_compute));
  s->launches++;
  CUDA_TRY (cudaStreamSynchronize (s->s_compute));
  return FDTD_OK;
}

int fdtd_init_eps_sphere (fdtd_handle s, double cx, double cy, double cz, double radius, double eps_inside)
{
  if (!s) return fail (FDTD_INVALID_ARGUMENT, "NULL handle");
  if (s->cfg.dims != 3) return fail (FDTD_INVALID_ARGUMENT, "the eps sphere is a Scheme3D feature");
  CUDA_TRY (cudaSetDevice (s->cfg.device));
  int rc = sync_run (s);
  if (rc) return rc;
  if ((rc = ensure_unfolded (s))) return rc;
  s->eps_uniform = false;
  drop_step_graph (s);
  if (s->fp == 4)
    CUDA_TRY (launch_eps_sphere<float> ((float *) s->eps, s->nx + 1, s->ny + 1, s->nzm, s->kz0, s->pitchM, s->psM, (float) cx, (float) cy, (float) cz, (float) radius, (float) eps_inside, s->s_compute));
  else
    CUDA_TRY (launch_eps_sphere<double> ((double *) s->eps, s->nx + 1, s->ny + 1, s->nzm, s->kz0, s->pitchM, s->psM, cx, cy, cz, radius, eps_inside, s->s_compute));
  s->launches++;
  CUDA_TRY (cudaStreamSynchronize (s->s_compute));
  return FDTD_OK;
}

int fdtd_subtract_incident (fdtd_handle s)
{
  if (!s) return fail (FDTD_INVALID_ARGUMENT, "NULL handle");
  if (s->cfg.dims != 3 || !s->cfg.use_tfsf) return fail (FDTD_INVALID_ARGUMENT, "the scattered-field post-processing needs a 3D scheme with use_tfsf");
  CUDA_TRY (cudaSetDevice (s->cfg.device));
  int rc = sync_run (s);
  if (rc) return rc;
  if ((rc = ensure_unfolded (s))) return rc;
  for (int comp = 0; comp < 6; ++comp)
  {
    if (s->fp == 4) CUDA_TRY (launch_scattered<float> (make_params<float> (s), s->cfg.complex_values != 0, comp, s->s_compute));
    else CUDA_TRY (launch_scattered<double> (make_params<double> (s), s->cfg.complex_values != 0, comp, s->s_compute));
    s->launches++;
  }
  CUDA_TRY (cudaStreamSynchronize (s->s_compute));
  return FDTD_OK;
}

int fdtd_run_async (fdtd_handle s, int64_t nsteps)
{
  if (!s) return fail (FDTD_INVALID_ARGUMENT, "NULL handle");
  return run_async (s, nsteps);
}

int fdtd_busy (fdtd_handle s, int *busy)
{
  if (!s || !busy) return fail (FDTD_INVALID_ARGUMENT, "NULL argument");
  CUDA_TRY (cudaSetDevice (s->cfg.device));
  const cudaError_t q = cudaStreamQuery (s->s_compute);
  if (q != cudaSuccess && q != cudaErrorNotReady) return fail (FDTD_ERROR, "cudaStreamQuery failed: %s", cudaGetErrorString (q));
  *busy = q == cudaErrorNotReady;
  return FDTD_OK;
}

int fdtd_sync (fdtd_handle s)
{
  if (!s) return fail (FDTD_INVALID_ARGUMENT, "NULL handle");
  return sync_run (s);
}

int fdtd_run (fdtd_handle s, int64_t nsteps)
{
  if (!s) return fail (FDTD_INVALID_ARGUMENT, "NULL handle");
  int rc = run_async (s, nsteps);
  if (rc) return rc;
  return sync_run (s);
}

int fdtd_last_run_ms (fdtd_handle s, float *ms)
{
  if (!s || !ms) return fail (FDTD_INVALID_ARGUMENT, "NULL argument");
  *ms = s->last_ms;
  return FDTD_OK;
}

int fdtd_set_profiling (fdtd_handle s, int enable)
{
  if (!s) return fail (FDTD_INVALID_ARGUMENT, "NULL handle");
  int rc = sync_run (s);
  if (rc) return rc;
  s->profiling = enable != 0;
  if (!s->profiling) s->prof_steps = 0;
  return FDTD_OK;
}

int fdtd_phase_times (fdtd_handle s, float *e_ms, float *h_ms, int64_t *steps)
{
  if (!s || !e_ms || !h_ms || !steps) return fail (FDTD_INVALID_ARGUMENT, "NULL argument");
  int rc = sync_run (s);
  if (rc) return rc;
  *e_ms = s->prof_e_ms; *h_ms = s->prof_h_ms; *steps = s->prof_steps;
  return FDTD_OK;
}

int fdtd_step_breakdown (fdtd_handle s, float ms[8], int64_t *steps)
{
  if (!s || !ms || !steps) return fail (FDTD_INVALID_ARGUMENT, "NULL argument");
  int rc = sync_run (s);
  if (rc) return rc;
  for (int i = 0; i < 8; ++i) ms[i] = s->mprof_ms[i];
  *steps = s->fused_run ? 0 : s->mprof_steps;
  return FDTD_OK;
}

int fdtd_launch_count (fdtd_handle s, int64_t *count)
{
  if (!s || !count) return fail (FDTD_INVALID_ARGUMENT, "NULL argument");
  *count = s->launches;
  return FDTD_OK;
}

int fdtd_get_time_step (fdtd_handle s, int64_t *t)
{
  if (!s || !t) return fail (FDTD_INVALID_ARGUMENT, "NULL argument");
  *t = s->t;
  return FDTD_OK;
}

int fdtd_plan_core_box (const fdtd_config *cfg, const void *const sigma[3], int64_t lo[3], int64_t hi[3])
{
  if (!sigma || !lo || !hi) return fail (FDTD_INVALID_ARGUMENT, "NULL argument");
  int rc = validate (cfg);
  if (rc) return rc;
  std::vector<double> sg[3];
  for (int a = 0; a < 3; ++a)
  {
    const size_t n = (size_t) (a < cfg->dims ? cfg->size[a] : 1) + 1;
    sg[a].assign (n, 0.0);
    if (a < cfg->dims && sigma[a])
      for (size_t i = 0; i < n; ++i)
        sg[a][i] = cfg->value_type == 4 ? (double) ((const float *) sigma[a])[i] : ((const double *) sigma[a])[i];
  }
  int l[3], h[3];
  const bool on = cfg->use_pml && cfg->exact_mode == 0
                  && compute_core_box (cfg->dims, cfg->value_type, cfg->complex_values ? 2 : 1, sg, l, h);
  for (int a = 0; a < 3; ++a) { lo[a] = on ? l[a] : 0; hi[a] = on ? h[a] : -1; }
  return FDTD_OK;
}

int fdtd_tfsf_face_masks (const fdtd_config *cfg, int field, uint8_t *out, int64_t count)
{
  if (!out) return fail (FDTD_INVALID_ARGUMENT, "NULL output");
  int rc = validate (cfg);
  if (rc) return rc;
  if (cfg->dims != 3 || !cfg->use_tfsf) return fail (FDTD_INVALID_ARGUMENT, "the face boxes belong to a 3D scheme with use_tfsf");
  if (field < FDTD_EX || field > FDTD_HZ) return fail (FDTD_INVALID_ARGUMENT, "field %d is not one of Ex..Hz", field);
  const int nx = (int) cfg->size[0], ny = (int) cfg->size[1], nz = (int) cfg->size[2];
  if (count != (int64_t) nx * ny * nz) return fail (FDTD_INVALID_ARGUMENT, "count %lld, grid has %lld cells", (long long) count, (long long) nx * ny * nz);
  /* the very functions the half-step kernels evaluate (kernels3d_common.cuh), run on the host */
  StepParams<double> p;
  memset (&p, 0, sizeof (p));
  for (int a = 0; a < 3; ++a) { p.L[a] = (int) cfg->tfsf_size[a]; p.R[a] = (int) (cfg->size[a] - cfg->tfsf_size[a]); }
  const bool is_e = field <= FDTD_EZ;
  const int a = field % 3, b = (a + 1) % 3, c = (a + 2) % 3;
  for (int i = 0; i < nx; ++i)
    for (int j = 0; j < ny; ++j)
      for (int k = 0; k < nz; ++k)
      {
        const int f = is_e ? k3d::tfsf_e_faces (p, a, i, j, k) : k3d::tfsf_h_faces (p, a, i, j, k);
        /* relative bits (low / high face normal to b, then to c) -> LayoutDirection bits (2 * axis + high) */
        int m = 0;
        if (f & 1) m |= 1 << (2 * b);
        if (f & 2) m |= 1 << (2 * b + 1);
        if (f & 4) m |= 1 << (2 * c);
        if (f & 8) m |= 1 << (2 * c + 1);
        out[((size_t) i * ny + j) * nz + k] = (uint8_t) m;
      }
  return FDTD_OK;
}

int fdtd_get_fused_box (fdtd_handle s, int64_t lo[3], int64_t hi[3])
{
  if (!s || !lo || !hi) return fail (FDTD_INVALID_ARGUMENT, "NULL argument");
  const bool on = s->fused_ok && s->mu_uniform;
  lo[0] = 0; hi[0] = on ? s->nx - 1 : -1;
  lo[1] = on ? s->fbox.fy0 : 0; hi[1] = on ? s->fbox.fy1 : -1;
  lo[2] = on ? s->kz0 + s->fbox.fz0 : 0; hi[2] = on ? s->kz0 + s->fbox.fz1 : -1;
  return FDTD_OK;
}

int fdtd_fused_time (fdtd_handle s, float *ms)
{
  if (!s || !ms) return fail (FDTD_INVALID_ARGUMENT, "NULL argument");
  int rc = sync_run (s);
  if (rc) return rc;
  *ms = s->prof_f_ms;
  return FDTD_OK;
}

int fdtd_set_tuning (fdtd_handle s, int march_planes, int one_pass_march_planes, int allow_one_pass)
{
  if (!s) return fail (FDTD_INVALID_ARGUMENT, "NULL handle");
  if (march_planes < 0 || one_pass_march_planes < 0) return fail (FDTD_INVALID_ARGUMENT, "negative march length");
  CUDA_TRY (cudaSetDevice (s->cfg.device));
  int rc = sync_run (s);
  if (rc) return rc;
  s->kchunk_forced = march_planes;
  s->fused_kc = one_pass_march_planes;
  s->fused_policy = allow_one_pass;
  if (s->cfg.dims == 2) s->kchunk = march_planes > 0 ? march_planes : 8;
  return configure_3d (s, true);
}

int fdtd_get_core_box (fdtd_handle s, int64_t lo[3], int64_t hi[3])
{
  if (!s || !lo || !hi) return fail (FDTD_INVALID_ARGUMENT, "NULL argument");
  for (int a = 0; a < 3; ++a)
  {
    lo[a] = s->core_on ? s->core_lo[a] : 0;
    hi[a] = s->core_on ? s->core_hi[a] : -1;
  }
  return FDTD_OK;
}

int fdtd_field_sumsq (fdtd_handle s, int field, double *out)
{
  if (!s || !out) return fail (FDTD_INVALID_ARGUMENT, "NULL argument");
  CUDA_TRY (cudaSetDevice (s->cfg.device));
  int rc = sync_run (s);
  if (rc) return rc;
  FieldInfo fi;
  rc = field_info (s, field, &fi);
  if (rc) return rc;
  const bool material = field == FDTD_EPS || field == FDTD_MU;
  if (field >= FDTD_DX && field <= FDTD_BZ && (rc = ensure_unfolded (s))) return rc;
  const int scal_per_cell = material ? 1 : s->nc;
  cudaStream_t S = s->s_compute;
  if (fi.is_inc)
  {
    if (s->fp == 4) CUDA_TRY (launch_sumsq<float> ((const float *) fi.dev, (int) (s->ninc * s->nc), 1, 0, s->red_partials, s->red_out, S));
    else CUDA_TRY (launch_sumsq<double> ((const double *) fi.dev, (int) (s->ninc * s->nc), 1, 0, s->red_partials, s->red_out, S));
  }
  else
  {
    /* owned planes are contiguous rows: plane 1 .. nzl_planes */
    const long long nrows = (long long) fi.ny * fi.nzl_planes;
    const char *base = (const char *) fi.dev + (size_t) fi.ps * fi.cell;
    if (s->fp == 4) CUDA_TRY (launch_sumsq<float> ((const float *) base, fi.nx * scal_per_cell, nrows, (long long) fi.pitch * scal_per_cell, s->red_partials, s->red_out, S));
    else CUDA_TRY (launch_sumsq<double> ((const double *) base, fi.nx * scal_per_cell, nrows, (long long) fi.pitch * scal_per_cell, s->red_partials, s->red_out, S));
  }
  s->launches += 2;
  CUDA_TRY (cudaMemcpyAsync (out, s->red_out, sizeof (double), cudaMemcpyDeviceToHost, S));
  CUDA_TRY (cudaStreamSynchronize (S));
  return FDTD_OK;
}

int fdtd_comm_unique_id (void *id128)
{
  if (!id128) return fail (FDTD_INVALID_ARGUMENT, "NULL id buffer");
  const char *why = NULL;
  const NcclApi *nccl = nccl_api (&why);
  if (!nccl) return fail (FDTD_COMM_ERROR, "NCCL unavailable: %s", why);
  NcclUniqueId id;
  NCCL_TRY (nccl->GetUniqueId (&id));
  memcpy (id128, &id, sizeof (id));
  return FDTD_OK;
}

int fdtd_comm_init (fdtd_handle s, const void *id128)
{
  if (!s || !id128) return fail (FDTD_INVALID_ARGUMENT, "NULL argument");
  if (s->cfg.nranks < 2) return FDTD_OK;
  const char *why = NULL;
  const NcclApi *nccl = nccl_api (&why);
  if (!nccl) return fail (FDTD_COMM_ERROR, "NCCL unavailable: %s", why);
  CUDA_TRY (cudaSetDevice (s->cfg.device));
  NcclUniqueId id;
  memcpy (&id, id128, sizeof (id));
  NCCL_TRY (nccl->CommInitRank (&s->comm, s->cfg.nranks, id, s->cfg.rank));
  return FDTD_OK;
}

int fdtd_execute_steps (int *status, const fdtd_config *cfg, fdtd_host_arrays *arr, int64_t total_steps)
{
  if (status) *status = FDTD_ERROR;
  if (!arr) return fail (FDTD_INVALID_ARGUMENT, "NULL arrays");
  if (cfg && cfg->nranks > 1)
    return fail (FDTD_INVALID_ARGUMENT, "fdtd_execute_steps is the single-GPU one-shot call: it has no communicator for %d ranks "
                 "(use fdtd_create + fdtd_comm_init + fdtd_run per rank)", cfg->nranks);
  fdtd_handle h = NULL;
  int rc = fdtd_create (cfg, &h);
  if (rc) return rc;
  const int64_t nzf = cfg->dims == 3 ? cfg->size[2] : 1;
  const int64_t nzm = cfg->dims == 3 ? cfg->size[2] + 1 : 1;
  do
  {
    for (int a = 0; a < cfg->dims && !rc; ++a)
      if (arr->sigma[a]) rc = fdtd_set_sigma (h, a, arr->sigma[a], cfg->size[a] + 1);
    if (rc) break;
    if (arr->eps && (rc = fdtd_upload (h, FDTD_EPS, arr->eps, FDTD_ORDER_REFERENCE, 0, nzm))) break;
    if (arr->mu && (rc = fdtd_upload (h, FDTD_MU, arr->mu, FDTD_ORDER_REFERENCE, 0, nzm))) break;
    for (int f = 0; f < 12 && !rc; ++f)
      if (arr->field[f] && h->f[f]) rc = fdtd_upload (h, f, arr->field[f], FDTD_ORDER_REFERENCE, 0, nzf);
    if (rc) break;
    if (cfg->use_tfsf)
    {
      if (arr->einc && (rc = fdtd_upload (h, FDTD_EINC, arr->einc, FDTD_ORDER_REFERENCE, 0, 1))) break;
      if (arr->hinc && (rc = fdtd_upload (h, FDTD_HINC, arr->hinc, FDTD_ORDER_REFERENCE, 0, 1))) break;
    }
    if ((rc = fdtd_run (h, total_steps))) break;
    for (int f = 0; f < 12 && !rc; ++f)
      if (arr->field[f] && h->f[f]) rc = fdtd_download (h, f, arr->field[f], FDTD_ORDER_REFERENCE, 0, nzf);
    if (rc) break;
    if (cfg->use_tfsf)
    {
      if (arr->einc && (rc = fdtd_download (h, FDTD_EINC, arr->einc, FDTD_ORDER_REFERENCE, 0, 1))) break;
      if (arr->hinc && (rc = fdtd_download (h, FDTD_HINC, arr->hinc, FDTD_ORDER_REFERENCE, 0, 1))) break;
    }
  } while (0);
  std::string keep = g_last_error;
  fdtd_destroy (h);
  if (rc) { g_last_error = keep; return rc; }
  if (status) *status = FDTD_OK;
  return FDTD_OK;
}

} /* extern "C" */
