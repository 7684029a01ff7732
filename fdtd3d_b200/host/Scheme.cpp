#include "Scheme.h"

#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <thread>
#include <vector>

#include "Settings.h"

YeeGridLayout::YeeGridLayout (int d, const long long s[3], const long long p[3], const long long t[3], double teta,
                              double phi, double psi)
  : tetaDeg (teta), phiDeg (phi), psiDeg (psi), dims (d)
{
  for (int a = 0; a < 3; ++a)
  {
    size[a] = a < d ? s[a] : 1;
    pml[a] = a < d ? p[a] : 0;
    tfsf[a] = a < d ? t[a] : 0;
  }
  /* YeeGridLayout.h:446-447: ASSERT (angle >= 0 && angle <= Pi / 2) for teta and phi */
  if (teta < 0 || teta > 90.0 || phi < 0 || phi > 90.0)
  {
    printf ("ERROR: incident wave angles teta and phi must be within [0, 90] degrees.\n");
    exit (EXIT_ERROR);
  }
}

void DeviceScheme::fail (const char *what) const
{
  printf ("ERROR: %s: %s\n", what, fdtd_last_error ());
  exit (EXIT_ERROR);
}

DeviceScheme::DeviceScheme (YeeGridLayout *layout, long long tStep, bool doUsePML, bool doUseTFSF, int valueType,
                            bool complexValues, int exactMode, int device, int numGpus)
  : yeeLayout (layout), totalStep (tStep), lastRunMs (0)
{
  memset (&cfg, 0, sizeof (cfg));
  cfg.abi_version = FDTD3D_B200_ABI_VERSION;
  cfg.dims = layout->dims;
  cfg.value_type = valueType;
  cfg.complex_values = complexValues;
  cfg.use_pml = doUsePML;
  cfg.use_tfsf = doUseTFSF;
  cfg.exact_mode = exactMode;
  cfg.device = device;
  for (int a = 0; a < 3; ++a)
  {
    cfg.size[a] = layout->size[a];
    cfg.pml_size[a] = layout->pml[a];
    cfg.tfsf_size[a] = layout->tfsf[a];
  }
  cfg.nranks = numGpus < 1 ? 1 : numGpus;
  if (cfg.nranks > 1 && (cfg.dims != 3 || cfg.size[2] / cfg.nranks < 2))
  {
    printf ("ERROR: --num-cuda-gpus %d needs a 3D grid with at least two z planes per GPU.\n", cfg.nranks);
    exit (EXIT_ERROR);
  }
  /* Scheme3D.h:287-290: TF/SF needs a non-zero scattered zone, PML a non-zero PML size */
  if (doUseTFSF && cfg.tfsf_size[0] == 0 && cfg.tfsf_size[1] == 0 && cfg.tfsf_size[2] == 0)
  {
    printf ("ERROR: --use-tfsf needs a non-zero TF/SF size.\n");
    exit (EXIT_ERROR);
  }
  if (doUsePML && cfg.pml_size[0] == 0 && cfg.pml_size[1] == 0 && cfg.pml_size[2] == 0)
  {
    printf ("ERROR: --use-pml needs a non-zero PML size.\n");
    exit (EXIT_ERROR);
  }
}

DeviceScheme::~DeviceScheme ()
{
  for (fdtd_handle h : handles) if (h) fdtd_destroy (h);
}

bool DeviceScheme::forEachRank (const char *what, int (*fn) (DeviceScheme *, int))
{
  const int n = (int) handles.size ();
  std::vector<int> rc (n, 0);
  std::vector<std::string> msg (n);
  if (n == 1)
  {
    rc[0] = fn (this, 0);
    if (rc[0]) msg[0] = fdtd_last_error ();
  }
  else
  {
    std::vector<std::thread> th;
    for (int r = 0; r < n; ++r)
      th.emplace_back ([&, r] () { rc[r] = fn (this, r); if (rc[r]) msg[r] = fdtd_last_error (); /* per-thread message */ });
    for (std::thread &t : th) t.join ();
  }
  bool ok = true;
  for (int r = 0; r < n; ++r)
    if (rc[r]) { printf ("ERROR: %s (GPU %d): %s\n", what, (int) cfg.device + r, msg[r].c_str ()); ok = false; }
  return ok;
}

void DeviceScheme::initScheme (double dx, double wavelength)
{
  cfg.dx = dx;
  if (fdtd_init_scheme (&cfg, wavelength, yeeLayout->tetaDeg, yeeLayout->phiDeg, yeeLayout->psiDeg) != FDTD_OK)
  {
    printf ("ERROR: invalid scheme parameters (dx = %g, wavelength = %g).\n", dx, wavelength);
    exit (EXIT_ERROR);
  }
}

static unsigned char g_comm_id[128];

void DeviceScheme::initGrids ()
{
  const int n = cfg.nranks;
  handles.assign ((size_t) n, (fdtd_handle) NULL);
  std::vector<std::vector<char> > prof (3);
  if (cfg.use_pml)
    for (int a = 0; a < cfg.dims; ++a)
    {
      prof[a].resize ((size_t) (cfg.size[a] + 1) * cfg.value_type);
      if (fdtd_pml_sigma_profile (&cfg, a, prof[a].data (), cfg.size[a] + 1) != FDTD_OK)
      {
        printf ("ERROR: cannot build the PML profile along axis %d.\n", a);
        exit (EXIT_ERROR);
      }
    }
  for (int r = 0; r < n; ++r)
  {
    fdtd_config c = cfg;
    c.rank = r;
    c.device = cfg.device + r;
    if (fdtd_create (&c, &handles[r]) != FDTD_OK) fail ("cannot create the device scheme");
    for (int a = 0; cfg.use_pml && a < cfg.dims; ++a)
      if (fdtd_set_sigma (handles[r], a, prof[a].data (), cfg.size[a] + 1) != FDTD_OK) fail ("fdtd_set_sigma");
    /* Scheme3D.cpp:3437: eps = smoothed sphere, centre (40.5, 40.5, 40.5), radius 20, eps 2 (hard-coded there);
     * SchemeTMz.cpp:1966-2002: eps = 1.  mu = 1 in both (fdtd_create's default). */
    if (cfg.dims == 3 && fdtd_init_eps_sphere (handles[r], 40.5, 40.5, 40.5, 20.0, 2.0) != FDTD_OK) fail ("fdtd_init_eps_sphere");
  }
  if (n > 1)
  {
    /* one NCCL communicator over the GPUs of this process: the id is shared in memory, every rank joins from its own
     * thread (replaces ParallelGridCore's MPI set-up) */
    if (fdtd_comm_unique_id (g_comm_id) != FDTD_OK) fail ("fdtd_comm_unique_id");
    if (!forEachRank ("fdtd_comm_init", [] (DeviceScheme *s, int r) { return fdtd_comm_init (s->handles[r], g_comm_id); }))
      exit (EXIT_ERROR);
  }
}

void DeviceScheme::performSteps ()
{
  if (handles.empty ()) initGrids ();
  static long long steps;
  steps = totalStep;
  if (!forEachRank ("fdtd_run", [] (DeviceScheme *s, int r) { return fdtd_run (s->handles[r], steps); })) exit (EXIT_ERROR);
  lastRunMs = 0;
  for (fdtd_handle h : handles)
  {
    float ms = 0;
    fdtd_last_run_ms (h, &ms);
    if (ms > lastRunMs) lastRunMs = ms; /* the slowest rank */
  }
}

static const char *const kFieldNames[] = { "Ex", "Ey", "Ez", "Hx", "Hy", "Hz", "Dx", "Dy", "Dz", "Bx", "By", "Bz", "Eps", "Mu", "EInc", "HInc" };

/* fields that exist for this configuration: E/H (+D/B with PML) of the scheme's components, incident lines with TF/SF */
static bool fieldPresent (const fdtd_config &cfg, int f)
{
  if (f == FDTD_EINC || f == FDTD_HINC) return cfg.use_tfsf != 0;
  if (f >= FDTD_EPS) return false;
  const int comp = f % 3;
  const bool e_like = f < FDTD_HX || (f >= FDTD_DX && f < FDTD_BX);
  if (cfg.dims == 2 && (e_like ? comp != 2 : comp == 2)) return false;
  if (f >= FDTD_DX && !cfg.use_pml) return false;
  return true;
}

void DeviceScheme::loadDat (const std::string &dir, long long step)
{
  for (fdtd_handle h : handles) if (h) fdtd_destroy (h);
  handles.clear ();
  cfg.time_step_start = step;
  initGrids ();
  const long long cells = cfg.size[0] * cfg.size[1] * (cfg.dims == 3 ? cfg.size[2] : 1);
  const long long ninc = 100LL * (cfg.size[0] + cfg.size[1] + (cfg.dims == 3 ? cfg.size[2] : 0));
  const size_t cell = (size_t) cfg.value_type * (cfg.complex_values ? 2 : 1);
  for (int f = 0; f < FDTD_FIELD_COUNT; ++f)
  {
    if (!fieldPresent (cfg, f)) continue;
    const size_t bytes = (size_t) (f >= FDTD_EINC ? ninc : cells) * cell;
    std::vector<char> buf (bytes);
    const std::string path = dir + "/current[" + std::to_string (step) + "]_rank-0_" + kFieldNames[f] + ".dat";
    std::ifstream file (path.c_str (), std::ios::in | std::ios::binary);
    if (!file.is_open () || !file.read (buf.data (), (std::streamsize) bytes))
    {
      printf ("ERROR: cannot read %zu bytes from %s\n", bytes, path.c_str ());
      exit (EXIT_ERROR);
    }
    /* every rank takes its own planes out of the whole-grid block */
    for (fdtd_handle h : handles)
      if (fdtd_upload (h, f, buf.data (), FDTD_ORDER_REFERENCE, 0, cfg.dims == 3 ? cfg.size[2] : 1) != FDTD_OK) fail ("fdtd_upload");
  }
}

void DeviceScheme::makeScatteredField ()
{
  if (cfg.dims != 3 || !cfg.use_tfsf) return;
  for (fdtd_handle h : handles)
    if (fdtd_subtract_incident (h) != FDTD_OK) fail ("fdtd_subtract_incident");
}

void DeviceScheme::dumpDat (const std::string &dir, const std::string &prefix)
{
  const long long cells = cfg.size[0] * cfg.size[1] * (cfg.dims == 3 ? cfg.size[2] : 1);
  const long long ninc = 100LL * (cfg.size[0] + cfg.size[1] + (cfg.dims == 3 ? cfg.size[2] : 0));
  const size_t cell = (size_t) cfg.value_type * (cfg.complex_values ? 2 : 1);
  int64_t t = 0;
  fdtd_get_time_step (handles[0], &t);
  for (int f = 0; f < FDTD_FIELD_COUNT; ++f)
  {
    if (!fieldPresent (cfg, f)) continue;
    std::vector<char> buf ((size_t) (f >= FDTD_EINC ? ninc : cells) * cell);
    /* the slabs are gathered in place: each rank writes only the planes it owns into the whole-grid block (the
     * incident line is replicated: rank 0's copy) */
    for (size_t r = 0; r < (f >= FDTD_EINC ? (size_t) 1 : handles.size ()); ++r)
      if (fdtd_download (handles[r], f, buf.data (), FDTD_ORDER_REFERENCE, 0, cfg.dims == 3 ? cfg.size[2] : 1) != FDTD_OK) fail ("fdtd_download");
    const std::string path = dir + "/current[" + std::to_string ((long long) t) + "]_rank-0_" + prefix + kFieldNames[f] + ".dat";
    std::ofstream file (path.c_str (), std::ios::out | std::ios::binary);
    if (!file.is_open ())
    {
      printf ("ERROR: cannot write %s\n", path.c_str ());
      exit (EXIT_ERROR);
    }
    file.write (buf.data (), (std::streamsize) buf.size ());
  }
}

Scheme3D::Scheme3D (YeeGridLayout *layout, long long tStep, bool doUsePML, bool doUseTFSF, int valueType, bool complexValues,
                    int exactMode, int device, int numGpus)
  : DeviceScheme (layout, tStep, doUsePML, doUseTFSF, valueType, complexValues, exactMode, device, numGpus)
{
}

SchemeTMz::SchemeTMz (YeeGridLayout *layout, long long tStep, bool doUsePML, bool doUseTFSF, int valueType, bool complexValues,
                      int exactMode, int device)
  : DeviceScheme (layout, tStep, doUsePML, doUseTFSF, valueType, complexValues, exactMode, device)
{
  /* SchemeTMz.h:241-246: only phi = 0 and phi = 45 degrees are implemented for the TMz incident wave */
  if (doUseTFSF && layout->phiDeg != 0.0 && layout->phiDeg != 45.0)
  {
    printf ("ERROR: the TMz scheme supports TF/SF only for --angle-phi 0 or 45.\n");
    exit (EXIT_ERROR);
  }
}
