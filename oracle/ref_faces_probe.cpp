/*
 * TEST INFRASTRUCTURE ONLY.  Our probe around the reference's own YeeGridLayout (compiled from
 * /root/reference/Source/Layout/YeeGridLayout.cpp by oracle/build_ref.sh): for every cell of an nx x ny x nz grid
 * and each of the six components it evaluates the reference's floating-point TF/SF face predicates
 * doNeedTFSFUpdate{Ex..Hz}Border (YeeGridLayout.cpp:327-809) for the four directions Scheme3D asks about
 * (Scheme3D.cpp:148-152, 445-449, 742-746, 1039-1043, 1335-1339, 1631-1634) and writes one byte per cell and
 * component: bit d set <=> the predicate holds for LayoutDirection d (LEFT 0, RIGHT 1, DOWN 2, UP 3, BACK 4, FRONT 5).
 * Output: 6 * nx*ny*nz bytes, component-major, cells in the reference's linear order (i*ny + j)*nz + k.
 * tests/golden/make_faces_golden.py turns it into the fixture the integer face boxes of the product are tested against.
 *
 * argv: nx ny nz lx ly lz out.bin
 */
#include <cstdio>
#include <cstdlib>
#include <vector>

#include "YeeGridLayout.h"

int cudaThreadsX = 8, cudaThreadsY = 8, cudaThreadsZ = 8;

int main (int argc, char **argv)
{
  if (argc != 8) { fprintf (stderr, "usage: %s nx ny nz lx ly lz out.bin\n", argv[0]); return 2; }
  const int nx = atoi (argv[1]), ny = atoi (argv[2]), nz = atoi (argv[3]);
  GridCoordinate3D size (nx, ny, nz), pml (2, 2, 2), tfsf (atoi (argv[4]), atoi (argv[5]), atoi (argv[6]));
  YeeGridLayout lay (size, pml, tfsf, 0.3, 0.4, 0.5, false);
  typedef bool (YeeGridLayout::*Pred) (GridCoordinate3D, LayoutDirection, bool) const;
  const Pred pred[6] = { &YeeGridLayout::doNeedTFSFUpdateExBorder, &YeeGridLayout::doNeedTFSFUpdateEyBorder,
                         &YeeGridLayout::doNeedTFSFUpdateEzBorder, &YeeGridLayout::doNeedTFSFUpdateHxBorder,
                         &YeeGridLayout::doNeedTFSFUpdateHyBorder, &YeeGridLayout::doNeedTFSFUpdateHzBorder };
  const LayoutDirection dirs[6] = { LayoutDirection::LEFT, LayoutDirection::RIGHT, LayoutDirection::DOWN,
                                    LayoutDirection::UP, LayoutDirection::BACK, LayoutDirection::FRONT };
  std::vector<unsigned char> out ((size_t) 6 * nx * ny * nz, 0);
  for (int c = 0; c < 6; ++c)
    for (int i = 0; i < nx; ++i)
      for (int j = 0; j < ny; ++j)
        for (int k = 0; k < nz; ++k)
        {
          unsigned char m = 0;
          for (int d = 0; d < 6; ++d)
          {
            if (d / 2 == c % 3) continue; /* the component's own axis: the reference's switch is UNREACHABLE there */
            if ((lay.*pred[c]) (GridCoordinate3D (i, j, k), dirs[d], true)) m |= (unsigned char) (1 << d);
          }
          out[(((size_t) c * nx + i) * ny + j) * nz + k] = m;
        }
  FILE *f = fopen (argv[7], "wb");
  if (!f) { perror ("fopen"); return 2; }
  fwrite (out.data (), 1, out.size (), f);
  fclose (f);
  return 0;
}
