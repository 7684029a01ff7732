#!/bin/bash
# ncu visit (one GPU): launch list of a short default bench run, then ONE `--set full` capture of the dominant kernel.
# usage: tools/gpu_ncu.sh <prefix> <workload> <kernel regex> [launches] [extra bench flags]   e.g. tools/gpu_ncu.sh r2_c3 c3 phase_3d_tma 2
P=$1; W=$2; K=$3; N=${4:-1}; X=$5
B="python bench.py --workload $W --steps 3 --warmup 3 --no-e2e --no-cpu-baseline --no-parity --quick $X"
$B > gpurun_out/${P}_plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/${P}_plain.log; exit 1; }
ncu --metrics gpu__time_duration.sum --clock-control none -c 120 --csv --log-file gpurun_out/${P}_launches.csv $B > gpurun_out/${P}_ncu_launches.log 2>&1
echo "launch list rc=$?"
ncu --set full --clock-control none --import-source on --kernel-name regex:$K --launch-skip 4 --launch-count $N -f -o gpurun_out/${P}_full $B > gpurun_out/${P}_ncu_full.log 2>&1
echo "full capture rc=$?"
ls -la gpurun_out/${P}_*
