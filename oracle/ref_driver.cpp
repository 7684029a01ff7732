/*
 * TEST INFRASTRUCTURE ONLY (Tier-1 oracle driver).  Not part of the product path.
 *
 * This file is OUR code; it is compiled by oracle/build_ref.sh together with the reference's own,
 * unmodified solver sources (read in place from /root/reference/Source) and drives them exactly
 * the way the reference's main() does (Source/main.cpp:79-156): YeeGridLayout -> Scheme ->
 * initScheme -> initGrids -> performSteps.  It bypasses Settings' std::stof parsing (SURVEY.md
 * appendix) by taking exact doubles from argv via strtod, then dumps every grid's `cur` layer in
 * the reference's linear-index order (Grid.cpp:82,138) into a small record file that
 * tests/refio.py reads.  Golden fixtures under tests/golden/ are made from these dumps by
 * tests/golden/make_golden.py.
 *
 * argv: nx ny nz pml tfsf steps use_pml use_tfsf dx wavelength theta_deg phi_deg psi_deg out.bin [dump_res [dat]]
 * dat = 1 additionally saves Ez and Hy with the reference's own DATDumper (Source/File-Management/Dumper/DATDumper.h) into
 * the current directory -- current[<steps>]_rank-0_Ez.dat, ..._Hy.dat -- so that the DAT files the product writes can be
 * compared with files the reference wrote (tests/golden/make_golden.py keeps their bytes in the fixture).
 * dump_res = 1 constructs the scheme with doDumpRes (--save-res): the reference then turns the total field inside
 * the TF/SF box into the scattered field in place before saving (Scheme3D.cpp:2767-2921); BMP writes are stubbed.
 */
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <stdint.h>
#include <sys/time.h>
#include <vector>

#include "Grid.h"
#include "PhysicsConst.h"
#include "Settings.h"
#include "DATDumper.h"
#ifdef GRID_3D
#include "Scheme3D.h"
#endif
#ifdef GRID_2D
#include "SchemeTMz.h"
#endif

int cudaThreadsX = 8, cudaThreadsY = 8, cudaThreadsZ = 8;

static FILE *out;

static void
put_record (const char *name, const void *data, uint64_t count, uint32_t elem, uint32_t ncomp)
{
  char nm[16];
  memset (nm, 0, sizeof (nm));
  strncpy (nm, name, 15);
  fwrite (nm, 1, 16, out);
  fwrite (&count, 8, 1, out);
  fwrite (&elem, 4, 1, out);
  fwrite (&ncomp, 4, 1, out);
  fwrite (data, (size_t) elem * ncomp, count, out);
}

template <class G>
static void
put_grid (const char *name, G &g)
{
  uint64_t n = g.getSize ().calculateTotalCoord ();
  std::vector<FieldValue> buf (n);
  for (uint64_t i = 0; i < n; ++i)
  {
    buf[i] = g.getFieldPointValue (i)->getCurValue ();
  }
  put_record (name, buf.data (), n, sizeof (FPValue), sizeof (FieldValue) / sizeof (FPValue));
}

int
main (int argc, char **argv)
{
  if (argc < 15 || argc > 17)
  {
    fprintf (stderr, "usage: %s nx ny nz pml tfsf steps use_pml use_tfsf dx wavelength theta phi psi out.bin\n", argv[0]);
    return 2;
  }
  grid_coord nx = atoi (argv[1]), ny = atoi (argv[2]), nz = atoi (argv[3]);
  grid_coord pml = atoi (argv[4]), tfsf = atoi (argv[5]);
  time_step steps = atoi (argv[6]);
  bool usePML = atoi (argv[7]) != 0, useTFSF = atoi (argv[8]) != 0;
  FPValue dx = (FPValue) strtod (argv[9], NULL);
  FPValue wavelength = (FPValue) strtod (argv[10], NULL);
  FPValue theta = (FPValue) strtod (argv[11], NULL);
  FPValue phi = (FPValue) strtod (argv[12], NULL);
  FPValue psi = (FPValue) strtod (argv[13], NULL);
  bool dumpRes = argc >= 16 && atoi (argv[15]) != 0;
  bool dumpDat = argc >= 17 && atoi (argv[16]) != 0;
  out = fopen (argv[14], "wb");
  if (!out) { perror ("fopen"); return 2; }

  /* same expressions as Source/main.cpp:79-85 */
  FPValue a1 = theta * PhysicsConst::Pi / 180.0;
  FPValue a2 = phi * PhysicsConst::Pi / 180.0;
  FPValue a3 = psi * PhysicsConst::Pi / 180.0;

#ifdef GRID_3D
  GridCoordinate3D overallSize (nx, ny, nz);
  GridCoordinate3D pmlSize (pml, pml, pml);
  GridCoordinate3D tfsfSize (tfsf, tfsf, tfsf);
  YeeGridLayout yeeLayout (overallSize, pmlSize, tfsfSize, a1, a2, a3, false);
  Scheme3D scheme (&yeeLayout, overallSize, steps, false, 0, usePML, useTFSF, false, false, dumpRes);
#endif
#ifdef GRID_2D
  GridCoordinate2D overallSize (nx, ny);
  GridCoordinate2D pmlSize (pml, pml);
  GridCoordinate2D tfsfSize (tfsf, tfsf);
  YeeGridLayout yeeLayout (overallSize, pmlSize, tfsfSize, a1, a2, a3, false);
  SchemeTMz scheme (&yeeLayout, overallSize, steps, false, 0, usePML, useTFSF, a2, false, dumpRes);
#endif

  scheme.initScheme (dx, PhysicsConst::SpeedOfLight / wavelength);
  scheme.initGrids ();

  struct timeval tv1, tv2;
  gettimeofday (&tv1, NULL);
  scheme.performSteps ();
  gettimeofday (&tv2, NULL);
  double secs = (double) (tv2.tv_usec - tv1.tv_usec) / 1000000 + (double) (tv2.tv_sec - tv1.tv_sec);
  printf ("Total time = %f seconds\n", secs);

  if (dumpDat)
  {
#ifdef GRID_3D
    DATDumper<GridCoordinate3D> dat;
    dat.init (steps, CURRENT, 0, "Ez");
    dat.dumpGrid (scheme.Ez, GridCoordinate3D (0, 0, 0), scheme.Ez.getSize ());
    dat.init (steps, CURRENT, 0, "Hy");
    dat.dumpGrid (scheme.Hy, GridCoordinate3D (0, 0, 0), scheme.Hy.getSize ());
#else
    DATDumper<GridCoordinate2D> dat;
    dat.init (steps, CURRENT, 0, "Ez");
    dat.dumpGrid (scheme.Ez, GridCoordinate2D (0, 0), scheme.Ez.getSize ());
    dat.init (steps, CURRENT, 0, "Hy");
    dat.dumpGrid (scheme.Hy, GridCoordinate2D (0, 0), scheme.Hy.getSize ());
#endif
  }

  double scal[8];
  scal[0] = scheme.gridTimeStep;
  scal[1] = scheme.gridStep;
  scal[2] = scheme.sourceFrequency;
#ifdef GRID_3D
  scal[3] = scheme.relPhaseVelocity;
#else
  scal[3] = 0;
#endif
  scal[4] = a1; scal[5] = a2; scal[6] = a3; scal[7] = secs;
  put_record ("scalars", scal, 8, 8, 1);

#ifdef GRID_3D
  put_grid ("Ex", scheme.Ex); put_grid ("Ey", scheme.Ey); put_grid ("Ez", scheme.Ez);
  put_grid ("Hx", scheme.Hx); put_grid ("Hy", scheme.Hy); put_grid ("Hz", scheme.Hz);
  if (usePML)
  {
    put_grid ("Dx", scheme.Dx); put_grid ("Dy", scheme.Dy); put_grid ("Dz", scheme.Dz);
    put_grid ("Bx", scheme.Bx); put_grid ("By", scheme.By); put_grid ("Bz", scheme.Bz);
  }
#endif
#ifdef GRID_2D
  put_grid ("Ez", scheme.Ez); put_grid ("Hx", scheme.Hx); put_grid ("Hy", scheme.Hy);
  if (usePML)
  {
    put_grid ("Dz", scheme.Dz); put_grid ("Bx", scheme.Bx); put_grid ("By", scheme.By);
  }
#endif
  put_grid ("Eps", scheme.Eps); put_grid ("Mu", scheme.Mu);
  put_grid ("SigmaX", scheme.SigmaX); put_grid ("SigmaY", scheme.SigmaY); put_grid ("SigmaZ", scheme.SigmaZ);
  if (useTFSF)
  {
    put_grid ("EInc", scheme.EInc); put_grid ("HInc", scheme.HInc);
  }
  fclose (out);
  return 0;
}
