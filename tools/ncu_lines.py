#!/usr/bin/env python
"""Per-source-line instruction / stall-sample shares of one kernel from an ncu report.

ncu's CSV source page is SASS-only; this joins it with `nvdisasm -g` line info of the cubin
(same instruction order) so hot spots can be read against the .cu file without a GUI.

    python tools/ncu_lines.py <report.ncu-rep> <cubin> <ncu-kernel-regex> <mangled-regex> [min_pct]
"""
import csv
import re
import subprocess
import sys
from collections import defaultdict


def sass_lines(cubin, kregex):
    txt = subprocess.run(["nvdisasm", "-g", "-c", cubin], capture_output=True, text=True).stdout
    out, cur_fn, take, line = [], None, False, None
    for ln in txt.splitlines():
        m = re.match(r"\s*\.section\s+\.text\.(\S+?),", ln)
        if m:
            cur_fn = m.group(1)
            take = re.search(kregex, cur_fn) is not None
            continue
        if not take:
            continue
        m = re.search(r'//## File ".*?([^/"]+)", line (\d+)', ln)
        if m:
            line = (m.group(1), int(m.group(2)))
            continue
        if re.match(r"\s+/\*[0-9a-f]{4}\*/", ln):
            out.append(line)
    return out


def main():
    rep, cubin, kre, mre = sys.argv[1:5]      # kre: demangled (ncu), mre: mangled (cubin)
    min_pct = float(sys.argv[5]) if len(sys.argv) > 5 else 0.5
    csvtxt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "-k", f"regex:{kre}",
                             "-c", "1"], capture_output=True, text=True).stdout
    rows = list(csv.reader(csvtxt.splitlines()))
    hi = next(i for i, r in enumerate(rows) if "Instructions Executed" in r)
    hdr = rows[hi]
    ie, ss = hdr.index("Instructions Executed"), hdr.index("# Samples")
    body = [r for r in rows[hi + 1:] if len(r) > ie and r[ie].isdigit()]
    lines = sass_lines(cubin, mre)
    if len(body) > len(lines) and len(body) % len(lines) == 0:
        body = body[:len(lines)]                 # several launches in the report: first one
    if len(lines) != len(body):
        print(f"warning: {len(lines)} SASS lines in cubin vs {len(body)} in report", file=sys.stderr)
    inst, smp = defaultdict(int), defaultdict(int)
    for k, r in enumerate(body):
        key = lines[k] if k < len(lines) else None
        inst[key] += int(r[ie])
        smp[key] += int(r[ss])
    ti, ts = sum(inst.values()), sum(smp.values())
    print(f"total warp-instructions {ti}, samples {ts}")
    for key in sorted(inst, key=lambda k: (k is None, k)):
        pi, ps = inst[key] / ti * 100, smp[key] / max(ts, 1) * 100
        if pi >= min_pct or ps >= min_pct:
            print(f"{key}: {pi:5.1f}% inst  {ps:5.1f}% samples")


if __name__ == "__main__":
    main()
