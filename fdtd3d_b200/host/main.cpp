/*
 * fdtd3d (B200 build): the solver binary with the reference's command line.
 *
 * Mirrors Source/main.cpp:36-241: parse Settings, build the layout and one Scheme (Scheme3D for --3d, SchemeTMz for
 * --2d), initScheme, initGrids, time performSteps with gettimeofday and print the same summary.  Where the
 * reference's cmake chooses dimension / value type / complex values at compile time, this binary takes them from
 * --2d|--3d (default 3d), --value-type, --complex-values.
 */
#include <cstdio>
#include <cstdlib>
#include <sys/time.h>

#include "Scheme.h"
#include "Settings.h"

Settings solverSettings;

int main (int argc, char **argv)
{
  solverSettings.SetupFromCmd (argc, argv);
  Settings &s = solverSettings;
  if (s.printSettings)
  {
    s.print (stdout);
    return EXIT_OK;
  }

  const int dims = s.getDimension () == 2 ? 2 : 3;
  if (s.valueType != "d" && s.valueType != "f")
  {
    printf ("ERROR: --value-type must be f or d (long double has no device type).\n");
    return EXIT_ERROR;
  }
  if (s.doUseMetamaterials || s.doUseAmplitudeMode || s.doUseNTFF || s.doUseDoubleMaterialPrecision || s.doUseParallelGrid)
  {
    printf ("ERROR: metamaterials, amplitude mode, NTFF, double material precision and the MPI parallel grid are outside "
            "the time-step path this build accelerates.\n");
    return EXIT_ERROR;
  }

  const long long overallSize[3] = { s.sizeX, s.sizeY, s.sizeZ };
  const long long pmlSize[3] = { s.pmlSizeX, s.pmlSizeY, s.pmlSizeZ };
  const long long tfsfSize[3] = { s.tfsfSizeX, s.tfsfSizeY, s.tfsfSizeZ };
  YeeGridLayout yeeLayout (dims, overallSize, pmlSize, tfsfSize, s.incidentWaveAngle1, s.incidentWaveAngle2, s.incidentWaveAngle3);

  char info[512];
  if (fdtd_device_info ((int) s.cudaDevice, info, sizeof (info)) != FDTD_OK) /* cudaInfo (); cudaInit (...) of main.cpp:88-96 */
  {
    printf ("ERROR: %s\n", fdtd_last_error ());
    return EXIT_ERROR;
  }
  printf ("%s\n", info);

  const int vt = s.valueType == "f" ? 4 : 8;
  const int exact = (s.exactMode && !s.fastMode) ? 1 : (s.keepAux ? 2 : 0); /* fdtd_config.exact_mode */
  DeviceScheme *scheme;
  if (s.numCudaGPUs < 1 || (dims == 2 && s.numCudaGPUs != 1))
  {
    printf ("ERROR: --num-cuda-gpus must be at least 1, and 1 for --2d (the z-slab decomposition is 3D).\n");
    return EXIT_ERROR;
  }
  if (dims == 2) scheme = new SchemeTMz (&yeeLayout, s.numTimeSteps, s.doUsePML, s.doUseTFSF, vt, s.complexValues, exact, (int) s.cudaDevice);
  else scheme = new Scheme3D (&yeeLayout, s.numTimeSteps, s.doUsePML, s.doUseTFSF, vt, s.complexValues, exact, (int) s.cudaDevice,
                              (int) s.numCudaGPUs);

  scheme->initScheme (s.gridStep, s.sourceWaveLength);
  if (!s.loadDatDir.empty ()) scheme->loadDat (s.loadDatDir, s.startStep);
  else scheme->initGrids ();

  struct timeval tv1, tv2;
  gettimeofday (&tv1, NULL);
  scheme->performSteps ();
  gettimeofday (&tv2, NULL);

  if (s.doSaveRes)
  {
    /* --save-res as in Scheme3D::performNSteps (Scheme3D.cpp:2541-2940): inside the TF/SF box the incident field
     * is subtracted, then the grids "3D-in-time-Ex" ... are saved -- here as DAT files instead of BMP images */
    scheme->makeScatteredField ();
    scheme->dumpDat (s.dumpDatDir.empty () ? std::string (".") : s.dumpDatDir, dims == 3 ? "3D-in-time-" : "2D-TMz-in-time-");
  }
  else if (!s.dumpDatDir.empty ()) scheme->dumpDat (s.dumpDatDir);

  /* Source/main.cpp:173-233 */
  printf ("Total time = %f seconds\n", (double) (tv2.tv_usec - tv1.tv_usec) / 1000000 + (double) (tv2.tv_sec - tv1.tv_sec));
  printf ("Dimension: %d\n", s.getDimension ());
  if (dims == 2) printf ("Grid size: %lldx%lld\n", s.sizeX, s.sizeY);
  else printf ("Grid size: %lldx%lldx%lld\n", s.sizeX, s.sizeY, s.sizeZ);
  printf ("Number of time steps: %lld\n", s.numTimeSteps);
  printf ("\n");
  printf ("Value type: %s\n", vt == 4 ? "float" : "double");
  printf ("Number of time steps: 2\n");
  printf ("\n-------- Details --------\n");
  printf ("Parallel grid: %d\n", 0);
  printf ("GPUs: %d (z-slabs, NCCL halo exchange)\n", scheme->numRanks ());
  const double cells = (double) s.sizeX * s.sizeY * (dims == 3 ? s.sizeZ : 1);
  printf ("Device time: %f ms, %.1f Mcell-updates/s (%s arithmetic)\n", scheme->deviceMilliseconds (),
          scheme->deviceMilliseconds () > 0 ? cells * s.numTimeSteps / (scheme->deviceMilliseconds () * 1e-3) / 1e6 : 0.0,
          exact ? "exact" : "fast");
  delete scheme;
  return EXIT_OK;
}
