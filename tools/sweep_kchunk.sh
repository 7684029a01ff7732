#!/bin/bash
# planes (3D) / rows (2D) marched per CTA: bench.py device-resident numbers for a list of march lengths
# usage: tools/sweep_kchunk.sh <workload> <one-pass: 0|1> k1 k2 ...   (run on the GPU box)
w=$1; op=$2; shift 2
for kc in "$@"; do
  if [ "$op" = 1 ]; then flag="--one-pass --one-pass-march $kc"; else flag="--march $kc"; fi
  python bench.py --workload $w --steps 20 --warmup 3 --no-e2e --no-cpu-baseline --no-parity --quick $flag 2>/dev/null | tail -1 | python -c "
import json,sys
d=json.loads(sys.stdin.read()); r=d['roofline']
print('$w kchunk=$kc one_pass=$op  %.0f Mcell/s  E %.4f ms  H %.4f ms  core %s ms  step frac %.3f' % (d['value'], r.get('e_phase_ms',0), r.get('h_phase_ms',0), r.get('one_pass_ms'), r['whole_step_frac']))"
done
