#!/usr/bin/env bash
# Tuning aid: device-resident throughput of one bench.py workload for several per-CTA march lengths.
# usage: tools/sweep_kchunk.sh <workload> <kchunk>...
w=$1; shift
for kc in "$@"; do
  FDTD3D_B200_KCHUNK=$kc python bench.py --workload $w --steps 20 --warmup 3 --no-e2e --no-cpu-baseline 2>/dev/null | tail -1 | python -c "
import json,sys
j=json.loads(sys.stdin.read()); r=j['roofline']
print('$w kchunk=$kc  %.0f Mcell/s  E %.4f ms  H %.4f ms  frac %.3f' % (j['value'], r.get('e_phase_ms',0), r.get('h_phase_ms',0), r['frac']))"
done
