#!/bin/bash
# 8-GPU visit: multi-rank parity (2, 4, 8 ranks; logs kept), the solver binary on several GPUs, the scaling benches.
P=${1:-r2}
for n in 2 4 8; do
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2961$n tests/multi_rank_check.py > gpurun_out/${P}_multi_rank_check_$n.log 2>&1
  grep "multi_rank_check:" gpurun_out/${P}_multi_rank_check_$n.log
done
timeout 900 python -m pytest tests/test_gpu_cli.py -m gpu -q -k "num_cuda" 2>&1 | tail -2
for n in 1 2 4 8; do
  if [ $n = 1 ]; then L="python"; else L="python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2970$n"; fi
  timeout 900 $L bench.py --gpus $n --steps 20 --warmup 5 > gpurun_out/${P}_scale_c3_$n.json 2> gpurun_out/${P}_scale_c3_$n.err
  python - <<PY
import json
try:
    d = json.loads(open("gpurun_out/${P}_scale_c3_$n.json").read().strip().splitlines()[-1])
    print("c3 N=$n", "%.0f" % d["value"], "ms/step %.4f" % d["ms_per_step"], "parity", d["parity"]["ok"], {k: (round(v["value"]), round(v["ms_per_step"], 4)) for k, v in (d.get("other_configs") or {}).items() if "value" in v})
except Exception as e:
    print("c3 N=$n failed", e)
PY
done
