/* See Settings.h.  Reference: Source/Settings/Settings.cpp (parseArg :19-183, setFromCmd :190-221,
 * loadCmdFromFile :228-267, saveCmdToFile :274-295, SetupFromCmd :300-321). */
#include "Settings.h"

#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <stdexcept>

Settings::Settings ()
  : dimension (0), logLevel (0),
    sizeX (100), sizeY (100), sizeZ (100), pmlSizeX (10), pmlSizeY (10), pmlSizeZ (10),
    tfsfSizeX (20), tfsfSizeY (20), tfsfSizeZ (20), ntffSizeX (15), ntffSizeY (15), ntffSizeZ (15),
    numTimeSteps (100), numAmplitudeTimeSteps (10),
    incidentWaveAngle1 (90.0), incidentWaveAngle2 (0.0), incidentWaveAngle3 (90.0),
    bufferSize (1), numCudaGPUs (1), doUseParallelGrid (false), doUseOptimalVirtualTopology (false),
    fileWithAvailableTopologies ("nofile"), topologySizeX (1), topologySizeY (1), topologySizeZ (1),
    doUseDoubleMaterialPrecision (false), doUseTFSF (false), doUseNTFF (false), doUsePML (false),
    doUseMetamaterials (false), doUseAmplitudeMode (false), gridStep (0.0005), sourceWaveLength (0.02),
    doSaveRes (false), doSaveMaterials (false), doSaveIntermediateRes (false), intermediateSaveStep (100),
    doSaveScatteredFieldRes (false), doSaveScatteredFieldIntermediate (false),
    valueType ("d"), complexValues (false), exactMode (false), fastMode (false), keepAux (false), printSettings (false), cudaDevice (0), startStep (0)
{
  buildTable ();
}

/* The option table in the order of Source/Settings/Settings.inc:28-125 (that is the order of --help). */
void Settings::buildTable ()
{
  auto none = [&] (const char *a, const char *d) { table.push_back ({ a, OPTION_NONE, d, NULL, "" }); };
  auto flag = [&] (const char *a, bool *f, const char *d) { table.push_back ({ a, TYPE_NONE, d, f, "" }); };
  auto integer = [&] (const char *a, long long *f, const char *d) { table.push_back ({ a, TYPE_INT, d, f, std::to_string (*f) }); };
  auto real = [&] (const char *a, double *f, const char *d)
  {
    char buf[64];
    snprintf (buf, sizeof (buf), "%f", *f);
    table.push_back ({ a, TYPE_FLOAT, d, f, buf });
  };
  auto str = [&] (const char *a, std::string *f, const char *d) { table.push_back ({ a, TYPE_STRING, d, f, *f }); };

  none ("--2d", "2D mode computations");
  none ("--3d", "3D mode computations");
  table.push_back ({ "--log-level", TYPE_LOG_LEVEL, "Log level of fdtd3d (0,1,2,3)", &logLevel, "0" });
  integer ("--sizex", &sizeX, "Size of calculation area by x coordinate");
  integer ("--sizey", &sizeY, "Size of calculation area by y coordinate");
  integer ("--sizez", &sizeZ, "Size of calculation area by z coordinate");
  none ("--same-size", "Use size of calculation area by x coordinate for y and z coordinates too");
  integer ("--pml-sizex", &pmlSizeX, "Size of PML area by x coordinate. PML of this size will be applied to both left and right borders of area by x coordinate");
  integer ("--pml-sizey", &pmlSizeY, "Size of PML area by y coordinate. PML of this size will be applied to both left and right borders of area by y coordinate");
  integer ("--pml-sizez", &pmlSizeZ, "Size of PML area by z coordinate. PML of this size will be applied to both left and right borders of area by z coordinate");
  none ("--same-size-pml", "Use size of PML area by x coordinate for y and z coordinates too");
  integer ("--tfsf-sizex", &tfsfSizeX, "Size of TF/SF scattered area by x coordinate. Border of TF/SF will be placed at this distance from both left and right borders of area by x coordinate");
  integer ("--tfsf-sizey", &tfsfSizeY, "Size of TF/SF scattered area by y coordinate. Border of TF/SF will be placed at this distance from both left and right borders of area by y coordinate");
  integer ("--tfsf-sizez", &tfsfSizeZ, "Size of TF/SF scattered area by z coordinate. Border of TF/SF will be placed at this distance from both left and right borders of area by z coordinate");
  none ("--same-size-tfsf", "Use size of TF/SF scattered area by x coordinate for y and z coordinates too");
  integer ("--ntff-sizex", &ntffSizeX, "Size of NTFF area by x coordinate. Border of NTFF will be placed at this distance from both left and right borders of area by x coordinate");
  integer ("--ntff-sizey", &ntffSizeY, "Size of NTFF area by y coordinate. Border of NTFF will be placed at this distance from both left and right borders of area by y coordinate");
  integer ("--ntff-sizez", &ntffSizeZ, "Size of NTFF area by z coordinate. Border of NTFF will be placed at this distance from both left and right borders of area by z coordinate");
  none ("--same-size-ntff", "Use size of NTFF area by x coordinate for y and z coordinates too");
  integer ("--time-steps", &numTimeSteps, "Number of time steps for which to perform computations");
  integer ("--amplitude-time-steps", &numAmplitudeTimeSteps, "Number of time steps for which to perform amplitude computations");
  real ("--angle-teta", &incidentWaveAngle1, "Incident wave angle teta (degrees)");
  real ("--angle-phi", &incidentWaveAngle2, "Incident wave angle phi (degrees)");
  real ("--angle-psi", &incidentWaveAngle3, "Incident wave angle psi (degrees)");
  integer ("--buffer-size", &bufferSize, "Size of buffer for parallel grid");
  integer ("--num-cuda-gpus", &numCudaGPUs, "Number of GPUs to use in computations");
  flag ("--parallel-grid", &doUseParallelGrid, "Use parallel grid (if fdtd3d is built with it)");
  flag ("--optimal-topology", &doUseOptimalVirtualTopology, "Use optimal topology for parallel grid");
  str ("--available-topologies", &fileWithAvailableTopologies, "File with available topologies for current architecture");
  integer ("--topology-sizex", &topologySizeX, "Size by x coordinate of virtual topology");
  integer ("--topology-sizey", &topologySizeY, "Size by y coordinate of virtual topology");
  integer ("--topology-sizez", &topologySizeZ, "Size by z coordinate of virtual topology");
  none ("--same-size-topology", "Use size of topology by x coordinate for y and z coordinates too");
  flag ("--use-double-material-precision", &doUseDoubleMaterialPrecision, "Use double material precision");
  flag ("--use-tfsf", &doUseTFSF, "Use TF/SF");
  flag ("--use-ntff", &doUseNTFF, "Use NTFF");
  flag ("--use-pml", &doUsePML, "Use PML");
  flag ("--use-metamaterials", &doUseMetamaterials, "Use Metamaterials");
  flag ("--use-amp-mode", &doUseAmplitudeMode, "Use amplitude mode");
  real ("--dx", &gridStep, "Grid step (meters)");
  real ("--wavelength", &sourceWaveLength, "Wave length of source (meters)");
  flag ("--save-res", &doSaveRes, "Save results to files");
  flag ("--save-materials", &doSaveMaterials, "Save materials to files");
  flag ("--save-interm-res", &doSaveIntermediateRes, "Save intermediate results to files");
  integer ("--interm-save-step", &intermediateSaveStep, "Save step for intermediate save");
  flag ("--save-scattered-field-res", &doSaveScatteredFieldRes, "Save scattered field for result");
  flag ("--save-scattered--field-interm", &doSaveScatteredFieldIntermediate, "Save scattered field for intermediate");
  table.push_back ({ "--cmd-from-file", OPTION_STRING,
                     "Load command line from file. Cmd file has the next format:\n\t\t<cmd with arg>\n\t\t<value>\n\t\t<cmd with arg>\n"
                     "\t\t<value>\n\t\t<cmd without arg>\n\t\t<cmd without arg>", NULL, "" });
  table.push_back ({ "--save-cmd-to-file", OPTION_STRING, "Save command line to file", NULL, "" });
  /* additions of this build (compile-time switches of the reference made run-time) */
  str ("--value-type", &valueType, "[fdtd3d_b200] floating point type of values: f or d (VALUE_TYPE of the reference's cmake)");
  flag ("--complex-values", &complexValues, "[fdtd3d_b200] complex field values (COMPLEX_FIELD_VALUES of the reference's cmake)");
  flag ("--exact", &exactMode, "[fdtd3d_b200] reference operand order, IEEE division, no FMA: bit-for-bit with the CPU solver");
  flag ("--fast", &fastMode, "[fdtd3d_b200] reciprocal tables and FMA; D/B are not carried where all sigma are zero (default)");
  flag ("--keep-aux", &keepAux, "[fdtd3d_b200] with --fast: carry D/B through every cell as the reference does");
  integer ("--cuda-device", &cudaDevice, "[fdtd3d_b200] CUDA device ordinal");
  flag ("--print-settings", &printSettings, "[fdtd3d_b200] print the parsed settings and exit");
  str ("--load-dat", &loadDatDir, "[fdtd3d_b200] directory with DAT dumps (current[<start-step>]_rank-0_<name>.dat) to resume from");
  integer ("--start-step", &startStep, "[fdtd3d_b200] time step the loaded state belongs to; --time-steps more steps are run");
  str ("--dump-dat", &dumpDatDir, "[fdtd3d_b200] directory for DAT dumps of the final fields (DATDumper format)");
}

int Settings::parseArg (int &index, int argc, char **argv, bool isCmd)
{
  const char *arg = argv[index];
  if (strcmp (arg, "--help") == 0)
  {
    printf ("fdtd3d is an open source 1D [NYI], 2D, 3D FDTD electromagnetics solver with MPI, OpenMP [NYI] and CUDA [WIP] support.\n");
    printf ("Usage: fdtd3d [options]\n\n");
    printf ("Options:\n");
    for (const Entry &e : table)
    {
      switch (e.type)
      {
        case TYPE_NONE: case OPTION_NONE: printf ("  %s\n\t%s\n", e.cmdArg, e.description); break;
        case TYPE_INT: printf ("  %s <int> (default: %s)\n\t%s\n", e.cmdArg, e.defText.c_str (), e.description); break;
        case TYPE_FLOAT: printf ("  %s <float> (default: %s)\n\t%s\n", e.cmdArg, e.defText.c_str (), e.description); break;
        case TYPE_STRING: printf ("  %s <string> (default: %s)\n\t%s\n", e.cmdArg, e.defText.c_str (), e.description); break;
        case TYPE_LOG_LEVEL: printf ("  %s <int> (default: 0)\n\t%s\n", e.cmdArg, e.description); break;
        case OPTION_STRING: printf ("  %s <string>\n\t%s\n", e.cmdArg, e.description); break;
      }
    }
    return EXIT_BREAK_ARG_PARSING;
  }
  if (strcmp (arg, "--version") == 0)
  {
    printf ("Version: %s\n", SOLVER_VERSION);
    return EXIT_BREAK_ARG_PARSING;
  }

  for (const Entry &e : table)
  {
    if (e.type == OPTION_NONE || e.type == OPTION_STRING || strcmp (arg, e.cmdArg) != 0) continue;
    if (e.type == TYPE_NONE)
    {
      *(bool *) e.field = true;
      return EXIT_OK;
    }
    ++index;
    if (index >= argc)
    {
      /* the reference ASSERTs here (Settings.cpp:68) */
      printf ("ERROR: option %s needs a value.\n", e.cmdArg);
      return EXIT_ERROR;
    }
    try
    {
      switch (e.type)
      {
        case TYPE_INT: *(long long *) e.field = std::stoi (argv[index]); break;          /* STOI */
        case TYPE_FLOAT: *(double *) e.field = std::stof (argv[index]); break;           /* STOF: single precision */
        case TYPE_STRING: *(std::string *) e.field = std::string (argv[index]); break;
        case TYPE_LOG_LEVEL:
        {
          int v = std::stoi (argv[index]);
          if (v < 0 || v > 3) { printf ("ERROR: log level must be 0..3.\n"); return EXIT_ERROR; }
          *(int *) e.field = v;
          break;
        }
        default: break;
      }
    }
    catch (const std::exception &)
    {
      /* std::stoi / std::stof throw in the reference too (uncaught there) */
      printf ("ERROR: bad value [%s] for option %s.\n", argv[index], e.cmdArg);
      return EXIT_ERROR;
    }
    return EXIT_OK;
  }

  if (strcmp (arg, "--same-size") == 0) sizeZ = sizeY = sizeX;
  else if (strcmp (arg, "--same-size-pml") == 0) pmlSizeZ = pmlSizeY = pmlSizeX;
  else if (strcmp (arg, "--same-size-tfsf") == 0) tfsfSizeZ = tfsfSizeY = tfsfSizeX;
  else if (strcmp (arg, "--same-size-ntff") == 0) ntffSizeZ = tfsfSizeY = tfsfSizeX; /* sic: Settings.cpp:133-136 */
  else if (strcmp (arg, "--2d") == 0) dimension = 2;
  else if (strcmp (arg, "--3d") == 0) dimension = 3;
  else if (strcmp (arg, "--cmd-from-file") == 0)
  {
    if (!isCmd)
    {
      printf ("Command line files are not allowed in other command line files.\n");
      return EXIT_ERROR;
    }
    if (argc != 3)
    {
      printf ("Command line files are allowed only without other options.\n");
      return EXIT_ERROR;
    }
    ++index;
    int status = loadCmdFromFile (argv[index]);
    if (status == EXIT_ERROR) printf ("ERROR: Incorrect command line file.\n");
    return status;
  }
  else if (strcmp (arg, "--save-cmd-to-file") == 0)
  {
    ++index;
    if (index >= argc) { printf ("ERROR: option --save-cmd-to-file needs a value.\n"); return EXIT_ERROR; }
    saveCmdToFile (argc, argv, argv[index]);
  }
  else
  {
    printf ("Unknown option [%s]\n", arg);
    return EXIT_UNKNOWN_OPTION;
  }
  return EXIT_OK;
}

int Settings::setFromCmd (int argc, char **argv, bool isCmd)
{
  for (int i = (isCmd ? 1 : 0); i < argc; ++i)
  {
    int status = parseArg (i, argc, argv, isCmd);
    if (status != EXIT_OK) return status;
  }
  return EXIT_OK;
}

int Settings::loadCmdFromFile (std::string fileName)
{
  printf ("Loading command line from file %s\n", fileName.c_str ());
  std::ifstream infile (fileName);
  std::vector<std::string> words;
  std::string w;
  while (infile >> w) words.push_back (w);
  std::vector<char *> argv;
  for (std::string &s : words) argv.push_back (&s[0]);
  return setFromCmd ((int) argv.size (), argv.data (), false);
}

int Settings::saveCmdToFile (int argc, char **argv, std::string fileName)
{
  printf ("Saving command line to file %s\n", fileName.c_str ());
  std::ofstream outfile (fileName);
  for (int i = 1; i < argc; ++i)
  {
    if (strcmp (argv[i], "--save-cmd-to-file") == 0) ++i;
    else outfile << argv[i] << std::endl;
  }
  return EXIT_OK;
}

void Settings::SetupFromCmd (int argc, char **argv)
{
  int status = setFromCmd (argc, argv, true);
  switch (status)
  {
    case EXIT_OK: break;
    case EXIT_BREAK_ARG_PARSING: exit (EXIT_OK);
    default: exit (status);
  }
}

void Settings::print (FILE *out) const
{
  fprintf (out, "dimension=%d\n", dimension);
  for (const Entry &e : table)
  {
    switch (e.type)
    {
      case TYPE_NONE: fprintf (out, "%s=%d\n", e.cmdArg, (int) *(bool *) e.field); break;
      case TYPE_INT: fprintf (out, "%s=%lld\n", e.cmdArg, *(long long *) e.field); break;
      case TYPE_FLOAT: fprintf (out, "%s=%.17g\n", e.cmdArg, *(double *) e.field); break;
      case TYPE_STRING: fprintf (out, "%s=%s\n", e.cmdArg, ((std::string *) e.field)->c_str ()); break;
      case TYPE_LOG_LEVEL: fprintf (out, "%s=%d\n", e.cmdArg, *(int *) e.field); break;
      default: break;
    }
  }
}
