/*
 * TEST INFRASTRUCTURE ONLY.  Our probe around the reference's own Settings class (compiled from
 * /root/reference/Source/Settings/Settings.cpp by oracle/build_ref.sh): parses argv exactly as fdtd3d's
 * main() does (Source/main.cpp:38) and prints every setting, one `--option=value` per line, in the order of
 * Source/Settings/Settings.inc.  tests/golden/make_settings_golden.py turns its output into fixtures.
 */
#include <cstdio>
#include "Settings.h"

extern Settings solverSettings;

int main (int argc, char **argv)
{
  solverSettings.SetupFromCmd (argc, argv);
  Settings &s = solverSettings;
  printf ("dimension=%d\n", s.getDimension ());
  printf ("--log-level=%d\n", (int) s.getLogLevel ());
  printf ("--sizex=%lld\n--sizey=%lld\n--sizez=%lld\n", (long long) s.getSizeX (), (long long) s.getSizeY (), (long long) s.getSizeZ ());
  printf ("--pml-sizex=%lld\n--pml-sizey=%lld\n--pml-sizez=%lld\n", (long long) s.getPMLSizeX (), (long long) s.getPMLSizeY (), (long long) s.getPMLSizeZ ());
  printf ("--tfsf-sizex=%lld\n--tfsf-sizey=%lld\n--tfsf-sizez=%lld\n", (long long) s.getTFSFSizeX (), (long long) s.getTFSFSizeY (), (long long) s.getTFSFSizeZ ());
  printf ("--ntff-sizex=%lld\n--ntff-sizey=%lld\n--ntff-sizez=%lld\n", (long long) s.getNTFFSizeX (), (long long) s.getNTFFSizeY (), (long long) s.getNTFFSizeZ ());
  printf ("--time-steps=%lld\n--amplitude-time-steps=%lld\n", (long long) s.getNumTimeSteps (), (long long) s.getNumAmplitudeSteps ());
  printf ("--angle-teta=%.17g\n--angle-phi=%.17g\n--angle-psi=%.17g\n", (double) s.getIncidentWaveAngle1 (), (double) s.getIncidentWaveAngle2 (), (double) s.getIncidentWaveAngle3 ());
  printf ("--buffer-size=%lld\n--num-cuda-gpus=%lld\n", (long long) s.getBufferSize (), (long long) s.getNumCudaGPUs ());
  printf ("--parallel-grid=%d\n--optimal-topology=%d\n", (int) s.getDoUseParallelGrid (), (int) s.getDoUseOptimalVirtualTopology ());
  printf ("--available-topologies=%s\n", s.getFileWithAvailableTopologies ().c_str ());
  printf ("--topology-sizex=%lld\n--topology-sizey=%lld\n--topology-sizez=%lld\n", (long long) s.getTopologySizeX (), (long long) s.getTopologySizeY (), (long long) s.getTopologySizeZ ());
  printf ("--use-double-material-precision=%d\n--use-tfsf=%d\n--use-ntff=%d\n--use-pml=%d\n--use-metamaterials=%d\n--use-amp-mode=%d\n",
          (int) s.getDoUseDoubleMaterialPrecision (), (int) s.getDoUseTFSF (), (int) s.getDoUseNTFF (), (int) s.getDoUsePML (),
          (int) s.getDoUseMetamaterials (), (int) s.getDoUseAmplitudeMode ());
  printf ("--dx=%.17g\n--wavelength=%.17g\n", (double) s.getGridStep (), (double) s.getSourceWaveLength ());
  printf ("--save-res=%d\n--save-materials=%d\n--save-interm-res=%d\n--interm-save-step=%lld\n--save-scattered-field-res=%d\n--save-scattered--field-interm=%d\n",
          (int) s.getDoSaveRes (), (int) s.getDoSaveMaterials (), (int) s.getDoSaveIntermediateRes (), (long long) s.getIntermediateSaveStep (),
          (int) s.getDoSaveScatteredFieldRes (), (int) s.getDoSaveScatteredFieldIntermediate ());
  return 0;
}
