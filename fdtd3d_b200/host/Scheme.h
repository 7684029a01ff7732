/*
 * Host-side Scheme classes of the fdtd3d_b200 solver binary: the callers of the C ABI.
 *
 * They mirror the reference's driver objects -- YeeGridLayout (Source/Layout/YeeGridLayout.h:400-477),
 * Scheme (Source/Scheme/Scheme.h:4-8), Scheme3D (Source/Scheme/Scheme3D.h:224-297) and SchemeTMz
 * (Source/Scheme/SchemeTMz.h) -- with the same life cycle main() drives: construct, initScheme, initGrids,
 * performSteps.  The grids themselves live on the GPU behind an fdtd_handle; nothing is computed on the host.
 */
#ifndef FDTD3D_B200_HOST_SCHEME_H
#define FDTD3D_B200_HOST_SCHEME_H

#include <string>
#include <vector>

#include "../../include/fdtd3d_b200.h"

/* the part of YeeGridLayout the time step needs as input: sizes and incident-wave angles (degrees, as the CLI
 * gives them; the conversion `angle * Pi / 180.0` of Source/main.cpp:79-85 happens in fdtd_init_scheme) */
struct YeeGridLayout
{
  long long size[3], pml[3], tfsf[3];
  double tetaDeg, phiDeg, psiDeg;
  int dims;
  YeeGridLayout (int dims, const long long size[3], const long long pml[3], const long long tfsf[3], double tetaDeg,
                 double phiDeg, double psiDeg);
};

class Scheme
{
public:
  virtual void performSteps () = 0;
  virtual ~Scheme () {}
};

class DeviceScheme : public Scheme
{
protected:
  YeeGridLayout *yeeLayout;
  fdtd_config cfg;                    /* cfg.rank / nranks / device describe rank 0; rank r runs on device cfg.device + r */
  std::vector<fdtd_handle> handles;   /* one per GPU: --num-cuda-gpus z-slabs (Source/main.cpp:88-96, Settings.inc:81) */
  long long totalStep;
  double lastRunMs;

  /* runs fn (rank) for every rank on a thread of its own and waits: NCCL set-up and the time loop need all ranks of
   * the process inside the library at the same time; returns false if any rank failed (message printed) */
  bool forEachRank (const char *what, int (*fn) (DeviceScheme *, int));

  void fail (const char *what) const; /* prints the library's error and exits with EXIT_ERROR */

public:
  DeviceScheme (YeeGridLayout *layout, long long tStep, bool doUsePML, bool doUseTFSF, int valueType, bool complexValues,
                int exactMode, int device, int numGpus = 1);
  ~DeviceScheme ();
  int numRanks () const { return cfg.nranks; }

  /* Scheme3D::initScheme (dx, sourceFreq) / SchemeTMz::initScheme: dt, frequency, relative phase velocity */
  void initScheme (double dx, double wavelength);
  /* Scheme3D::initGrids / SchemeTMz::initGrids: eps (sphere in 3D, 1 in 2D), mu = 1, sigma profiles, fields = 0 */
  void initGrids ();
  void performSteps () override;

  /* DATDumper format (Source/File-Management/Dumper/DATDumper.h:84-118): raw FieldValue per cell in linear-index
   * order, file name current[<step>]_rank-<p>_<name>.dat (Source/File-Management/Commons.h:56-68) */
  void dumpDat (const std::string &dir, const std::string &prefix = "");

  /* resume: load every field (and the incident lines) from DAT dumps of step `step` and continue from there --
   * the counterpart of DATLoader (Source/File-Management/Loader/DATLoader.h), which the reference never wires up */
  void loadDat (const std::string &dir, long long step);

  /* the post-processing Scheme3D::performNSteps applies before saving results (Scheme3D.cpp:2767-2921):
   * total -> scattered field inside the TF/SF box, in place; no-op for 2D or without TF/SF */
  void makeScatteredField ();

  const fdtd_config &config () const { return cfg; }
  double deviceMilliseconds () const { return lastRunMs; }
};

class Scheme3D : public DeviceScheme
{
public:
  Scheme3D (YeeGridLayout *layout, long long tStep, bool doUsePML, bool doUseTFSF, int valueType, bool complexValues,
            int exactMode, int device, int numGpus = 1);
};

class SchemeTMz : public DeviceScheme
{
public:
  SchemeTMz (YeeGridLayout *layout, long long tStep, bool doUsePML, bool doUseTFSF, int valueType, bool complexValues,
             int exactMode, int device);
};

#endif /* FDTD3D_B200_HOST_SCHEME_H */
