#!/usr/bin/env bash
# TEST INFRASTRUCTURE ONLY (build container: needs /root/reference and oracle/_ref from build_ref.sh).
# Settles SURVEY.md §8 row f4 on evidence: the unmodified reference solver, constructed with
# doUseMetamaterials = true (Scheme3D.h:231) and otherwise driven exactly like the Tier-1 oracle
# (initScheme -> initGrids -> performSteps), aborts in the first E half step -- there is no completing
# reference run of the Drude path to be on parity with.  Prints the reference's own assert line and backtrace.
# Usage: oracle/probe_metamaterials.sh [n]      (grid n^3, default 32; 100 = the reference's default size)
set -euo pipefail
HERE="$(cd "$(dirname "$0")" && pwd)"
S="${FDTD3D_REFERENCE:-/root/reference}/Source"
O="$HERE/_ref"
N="${1:-32}"
T="$(mktemp -d)"
sed 's/steps, false, 0, usePML, useTFSF, false, false, dumpRes);/steps, false, 0, usePML, useTFSF, true, false, dumpRes);/' "$HERE/ref_driver.cpp" > "$T/drv.cpp"
grep -q "useTFSF, true, false, dumpRes" "$T/drv.cpp"
INC="-I$S/Helpers -I$S/Physics -I$S/Kernels -I$S/Coordinate -I$S/Grid -I$S/Layout -I$S/File-Management -I$S/File-Management/Loader -I$S/File-Management/Dumper -I$S/Scheme -I$S/Settings"
g++ -std=c++11 -O1 -g -rdynamic -ffp-contract=off -include string -DTWO_TIME_STEPS= -DPRINT_MESSAGE=1 -DCXX11_ENABLED -w \
    -DDOUBLE_VALUES= -DGRID_3D= -I"$O/gen_3d_d" -I"$O/stub" $INC "$T/drv.cpp" "$O/gen_3d_d/Scheme3D.cpp" \
    "$S/Settings/Settings.cpp" "$S/Layout/YeeGridLayout.cpp" "$S/Layout/Approximation.cpp" "$S/Grid/Grid.cpp" \
    "$S/Kernels/FieldPoint.cpp" "$S/Coordinate/GridCoordinate.cpp" "$S/Helpers/Assert.cpp" "$S/File-Management/BMPHelper.cpp" \
    "$S/File-Management/Commons.cpp" "$S/File-Management/Dumper/BMPDumper.cpp" "$S/File-Management/Dumper/TXTDumper.cpp" \
    "$S/File-Management/Loader/BMPLoader.cpp" -o "$T/ref_meta"
set +e
"$T/ref_meta" "$N" "$N" "$N" 10 $((N > 60 ? 20 : 8)) 2 1 1 0.0005 0.02 90 0 90 "$T/out.bin" > "$T/log" 2>&1
echo "exit code $?"
grep -A 11 "^Assert at" "$T/log" | c++filt | sed 's/ \[0x[0-9a-f]*\]//; s/(GridCoordinate3DTemplate<unsigned long>[^)]*)//g'
rm -rf "$T"
