/*
 * Command-line settings of the fdtd3d_b200 solver binary.
 *
 * Mirrors the reference's Settings class (Source/Settings/Settings.{h,cpp,inc}): same option names, defaults,
 * argument types, parse order sensitivity and exit codes, so a command line written for fdtd3d works here
 * unchanged.  The reference generates everything from an X-macro table; here it is a run-time table.
 *
 * Kept quirks, because a drop-in must parse like the original:
 *   - float options are parsed with std::stof, i.e. in single precision, even in double builds
 *     (Source/Helpers/Assert.h:107-108): `--dx 0.0005` yields 0.00050000002374872565; the defaults are exact;
 *   - `--same-size*` copy the x value at the moment they are parsed (Settings.cpp:121-132);
 *   - `--same-size-ntff` sets ntffSizeZ and tfsfSizeY from tfsfSizeX (typo at Settings.cpp:133-136);
 *   - `--same-size-topology` is listed by --help but rejected as an unknown option;
 *   - bool options can only be switched on.
 *
 * Additions (the reference fixes these at compile time, CMakeLists.txt:58-199): --value-type f|d,
 * --complex-values, --exact / --fast, --cuda-device, --print-settings, --dump-dat, --load-dat + --start-step
 * (checkpoint / resume in DATDumper's format; the reference only has the unwired DATLoader class).
 */
#ifndef FDTD3D_B200_HOST_SETTINGS_H
#define FDTD3D_B200_HOST_SETTINGS_H

#include <string>
#include <vector>

#define SOLVER_VERSION "0.2.2"

/* Source/Helpers/Assert.h:114-117 */
#define EXIT_OK 0x0
#define EXIT_BREAK_ARG_PARSING 0x1
#define EXIT_ERROR 0xa
#define EXIT_UNKNOWN_OPTION 0xb

class Settings
{
public:
  enum ArgType { TYPE_NONE, TYPE_INT, TYPE_FLOAT, TYPE_STRING, TYPE_LOG_LEVEL, OPTION_NONE, OPTION_STRING };

  struct Entry
  {
    const char *cmdArg;
    ArgType type;
    const char *description;
    void *field;         /* bool* / long long* / double* / std::string* / int* (log level) */
    std::string defText; /* as --help prints it */
  };

private:
  int dimension;
  std::vector<Entry> table;

  int parseArg (int &index, int argc, char **argv, bool isCmd);
  int setFromCmd (int argc, char **argv, bool isCmd);
  int loadCmdFromFile (std::string fileName);
  int saveCmdToFile (int argc, char **argv, std::string fileName);
  void buildTable ();

public:
  /* fields, names as in Settings.inc */
  int logLevel;
  long long sizeX, sizeY, sizeZ;
  long long pmlSizeX, pmlSizeY, pmlSizeZ;
  long long tfsfSizeX, tfsfSizeY, tfsfSizeZ;
  long long ntffSizeX, ntffSizeY, ntffSizeZ;
  long long numTimeSteps, numAmplitudeTimeSteps;
  double incidentWaveAngle1, incidentWaveAngle2, incidentWaveAngle3;
  long long bufferSize, numCudaGPUs;
  bool doUseParallelGrid, doUseOptimalVirtualTopology;
  std::string fileWithAvailableTopologies;
  long long topologySizeX, topologySizeY, topologySizeZ;
  bool doUseDoubleMaterialPrecision, doUseTFSF, doUseNTFF, doUsePML, doUseMetamaterials, doUseAmplitudeMode;
  double gridStep, sourceWaveLength;
  bool doSaveRes, doSaveMaterials, doSaveIntermediateRes;
  long long intermediateSaveStep;
  bool doSaveScatteredFieldRes, doSaveScatteredFieldIntermediate;

  /* additions */
  std::string valueType; /* "d" or "f" */
  bool complexValues, exactMode, fastMode, keepAux, printSettings;
  long long cudaDevice;
  std::string dumpDatDir, loadDatDir;
  long long startStep;

  Settings ();

  /* Settings::SetupFromCmd, Settings.cpp:300-321: exits the process on --help/--version/errors */
  void SetupFromCmd (int argc, char **argv);
  /* same without exiting: returns the status code */
  int TrySetupFromCmd (int argc, char **argv) { return setFromCmd (argc, argv, true); }

  int getDimension () const { return dimension; }
  void print (FILE *out) const;
};

#endif /* FDTD3D_B200_HOST_SETTINGS_H */
