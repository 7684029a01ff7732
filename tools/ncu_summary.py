#!/usr/bin/env python
"""Summarises an `ncu --set full` report (.ncu-rep) of the half-step kernels into profiles/.

usage: tools/ncu_summary.py <report.ncu-rep> <tag>      -> profiles/<tag>_ncu_summary.md
and merges the per-kernel DRAM bytes into profiles/ncu_summary.json (bench.py reads `traffic` from it).
Run in the build container (ncu can read reports without a GPU).
"""
import csv
import io
import json
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
METRICS = [
    ("gpu__time_duration.sum", "duration"),
    ("dram__bytes_read.sum", "DRAM bytes read"),
    ("dram__bytes_write.sum", "DRAM bytes written"),
    ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "DRAM throughput % of peak"),
    ("lts__t_sector_hit_rate.pct", "L2 hit rate %"),
    ("l1tex__t_sector_hit_rate.pct", "L1 hit rate %"),
    ("launch__registers_per_thread", "registers / thread"),
    ("launch__shared_mem_per_block_dynamic", "dynamic smem / block"),
    ("launch__grid_size", "grid size"),
    ("launch__block_size", "block size"),
    ("sm__warps_active.avg.pct_of_peak_sustained_active", "warps active % of peak"),
    ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "SM throughput %"),
    ("sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active", "FP64 pipe active %"),
    ("smsp__average_warp_latency_issue_stalled_long_scoreboard.ratio", "stall long scoreboard (ratio)"),
    ("smsp__average_warp_latency_issue_stalled_barrier.ratio", "stall barrier (ratio)"),
]
UNIT = {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1.0, "Tbyte": 1e12}


def main():
    rep, tag = sys.argv[1], sys.argv[2]
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr, units = rows[0], rows[1]
    lines = ["# ncu --set full summary: %s" % tag, "", "source report: `%s` (scratch, not committed); command: see profiles/README.md" % os.path.basename(rep), ""]
    js_path = os.path.join(ROOT, "profiles", "ncu_summary.json")
    js = json.load(open(js_path)) if os.path.exists(js_path) else {}
    for r in rows[2:]:
        name = r[hdr.index("Kernel Name")]
        short = re.search(r"([eh]z?_phase_\w+|step_core_\w+)", name)
        short = short.group(1) if short else name[:40]
        lines += ["## `%s`" % name, "", "| metric | value |", "|---|---|"]
        dram = 0.0
        for key, label in METRICS:
            if key in hdr:
                i = hdr.index(key)
                lines.append("| %s (`%s`) | %s %s |" % (label, key, r[i], units[i]))
                if key.startswith("dram__bytes"):
                    dram += float(r[i].replace(",", "")) * UNIT.get(units[i], 1.0)
        lines.append("| DRAM read+write per launch | %.3f GB |" % (dram / 1e9))
        lines.append("")
        base = short.replace("_tma", "")
        targs = re.search(r"<(.*?)>\(", name)
        targ = [a.strip() for a in targs.group(1).split(",")] if targs else []
        if "phase" in base and len(targ) >= 6 and targ[5] in ("1", "(bool)1", "true"):  # <T, NC, PML, TFSF, EXACT, CORE[, UNIMU]>
            base += "_core"  # the variants that skip D/B in the sigma-free core
        if len(sys.argv) > 3:
            base = sys.argv[3] + base  # e.g. "c4s_": a capture of another workload must not replace the headline's entry
        js[base] = {"dram_bytes_per_launch": dram, "tag": tag, "kernel": name}
    open(os.path.join(ROOT, "profiles", tag + "_ncu_summary.md"), "w").write("\n".join(lines))
    json.dump(js, open(js_path, "w"), indent=1)
    print("\n".join(lines))


if __name__ == "__main__":
    main()
