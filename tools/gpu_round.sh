#!/bin/bash
# One visit to the GPU box: the one-pass kernels' tests first, then the whole suite, the bench line and A/B variants.
# Everything lands in gpurun_out/ (prefix $1).
P=${1:-r2}
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -q -k "one_pass" > gpurun_out/${P}_onepass.log 2>&1; echo "one_pass rc=$?" | tee -a gpurun_out/${P}_onepass.log
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/${P}_pytest.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/${P}_pytest.log
tail -3 gpurun_out/${P}_pytest.log
timeout 900 python bench.py --steps 20 --warmup 5 > gpurun_out/${P}_bench.json 2> gpurun_out/${P}_bench.err; echo "bench rc=$?"
for v in "--one-pass" "--one-pass --one-pass-march 16" "--one-pass --one-pass-march 64"; do
  timeout 300 python bench.py --steps 20 --warmup 3 --no-e2e --no-cpu-baseline --no-parity --quick $v 2>/dev/null | tail -1 >> gpurun_out/${P}_variants.jsonl
done
for w in c5 c2 c4s; do
  timeout 300 python bench.py --workload $w --steps 20 --warmup 3 --no-e2e --no-cpu-baseline --no-parity --quick 2>/dev/null | tail -1 >> gpurun_out/${P}_variants.jsonl
  timeout 300 python bench.py --workload $w --steps 20 --warmup 3 --no-e2e --no-cpu-baseline --no-parity --quick --one-pass 2>/dev/null | tail -1 >> gpurun_out/${P}_variants.jsonl
done
P=$P python - <<'PY'
import json, os
for l in open("gpurun_out/%s_variants.jsonl" % os.environ["P"]):
    try:
        d = json.loads(l); r = d["roofline"]
        print(d["config"]["name"], "%.0f Mcell/s" % d["value"], "ms/step %.3f" % d["ms_per_step"], "one_pass_ms", r.get("one_pass_ms"), "E", r.get("e_phase_ms"), "H", r.get("h_phase_ms"), "needed frac", r.get("whole_step_frac_of_needed_bytes"))
    except Exception as e:
        print("bad line", e)
PY
