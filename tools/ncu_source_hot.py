#!/usr/bin/env python
"""Hot source lines of a kernel from an ncu report: instructions executed and stall samples per CUDA-C line.

usage: tools/ncu_source_hot.py <report.ncu-rep> <kernel-name-regex> [top-N]
"""
import collections
import csv
import io
import subprocess
import sys


def num(x):
    try:
        return int(float(x.replace(",", "")))
    except ValueError:
        return 0


def main():
    rep, kern = sys.argv[1], sys.argv[2]
    top_n = int(sys.argv[3]) if len(sys.argv) > 3 else 40
    out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass", "--kernel-name", "regex:" + kern],
                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    secs, cur = [], None
    for r in rows:
        if len(r) >= 2 and r[0] == "File Path":
            cur = {"file": r[1], "rows": [], "hdr": None}
            secs.append(cur)
        elif cur is not None and r and r[0] == "Line No":
            cur["hdr"] = r
        elif cur is not None and cur["hdr"] and len(r) == len(cur["hdr"]):
            cur["rows"].append(r)
    tot = tot_s = 0
    per = collections.defaultdict(lambda: [0, 0, "", collections.Counter()])
    ops = collections.Counter()
    for s in secs:
        h = s["hdr"]
        i_inst, i_samp = h.index("Instructions Executed"), h.index("# Samples")
        line, src = None, ""
        for r in s["rows"]:
            if r[0] != "":
                line, src = num(r[0]), r[1]
            if not r[2].startswith("0x"):  # source-line rows carry the sum of their SASS rows
                continue
            n, sm = num(r[i_inst]), num(r[i_samp])
            key = (s["file"].split("/")[-1], line)
            per[key][0] += n
            per[key][1] += sm
            per[key][2] = src
            toks = r[3].split()
            op = (toks[1] if toks[0].startswith("@") else toks[0]).split(".")[0]
            per[key][3][op] += n
            ops[op] += n
            tot += n
            tot_s += sm
    print("total warp instructions %d, samples %d" % (tot, tot_s))
    print("opcode mix:", ", ".join("%s %.1f%%" % (o, 100.0 * n / tot) for o, n in ops.most_common(18)))
    for k, v in sorted(per.items(), key=lambda kv: -kv[1][0])[:top_n]:
        print("%-22s %5d inst %5.2f%% samp %5.2f%%  %s | %s" % (k[0], k[1], 100.0 * v[0] / tot, 100.0 * v[1] / max(tot_s, 1),
                                                             v[2].strip()[:64], dict(v[3].most_common(4))))


if __name__ == "__main__":
    main()
