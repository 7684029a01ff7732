#!/bin/bash
# Final single-GPU validation of a build: the whole GPU suite and the default bench line.  Output: gpurun_out/${1}_*.
P=${1:-final}
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/${P}_pytest.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/${P}_pytest.log
tail -3 gpurun_out/${P}_pytest.log
timeout 900 python bench.py --steps 20 --warmup 5 > gpurun_out/${P}_bench.json 2> gpurun_out/${P}_bench.err; echo "bench rc=$?"
python - <<PY
import json
d = json.loads(open("gpurun_out/${P}_bench.json").read().strip().splitlines()[-1])
print("c3", round(d["value"]), d["ms_per_step"], "e2e", round(d["e2e"]["value"]), "parity", d["parity"]["ok"], "sustained", (d.get("sustained") or {}).get("value"))
print({k: (round(v["value"]), round(v["ms_per_step"], 4)) for k, v in (d.get("other_configs") or {}).items() if "value" in v})
print({k: round(m["value"]) for k, m in (d["roofline"].get("modes") or {}).items()})
PY
