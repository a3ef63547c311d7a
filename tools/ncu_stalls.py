#!/usr/bin/env python
"""Per-kernel stall breakdown and hot spots from an ncu report (source page, SASS view).

    python tools/ncu_stalls.py <report.ncu-rep> [top_n]
Prints, for every profiled launch: instructions executed, warp-stall samples by reason, the
code split at block barriers (instructions / samples per section) and the most-sampled
instructions. Needs the report to be taken with --set full --import-source on."""
import csv
import subprocess
import sys


def main(rep, top_n=25):
    txt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"],
                         capture_output=True, text=True).stdout
    rows = list(csv.reader(txt.splitlines()))
    i = 0
    while i < len(rows):
        if not rows[i] or rows[i][0] != "Kernel Name":
            i += 1
            continue
        name = rows[i][1].split("(")[0]
        hdr = rows[i + 1]
        body = []
        i += 2
        while i < len(rows) and rows[i] and rows[i][0] != "Kernel Name" and len(rows[i]) >= 10:
            body.append(rows[i])
            i += 1
        iA, iS = hdr.index("Address"), hdr.index("Source")
        iN, iW = hdr.index("Instructions Executed"), hdr.index("Warp Stall Sampling (All Samples)")
        stall = [k for k, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
        base = int(body[0][iA], 16)
        tot = sum(int(r[iN]) for r in body)
        tot_s = sum(int(r[iW]) for r in body) or 1
        print(f"==== {name}: {len(body)} SASS instructions ({len(body) * 16 / 1024:.1f} KB), "
              f"{tot} executed, {tot_s} samples")
        by = {hdr[k][6:]: sum(int(r[k]) for r in body) for k in stall}
        print("  stalls:", ", ".join(f"{k} {100 * v / tot_s:.1f}%" for k, v in
                                      sorted(by.items(), key=lambda kv: -kv[1]) if v))
        cur = [0, 0, 0, None]
        for r in body:
            off = int(r[iA], 16) - base
            if cur[3] is None:
                cur[3] = off
            cur[0] += int(r[iN]); cur[1] += int(r[iW]); cur[2] += 1
            if "BAR.SYNC" in r[iS]:
                print(f"  {cur[3]:#07x}-{off:#07x} n={cur[2]:5d} inst {100 * cur[0] / tot:5.1f}% "
                      f"samples {100 * cur[1] / tot_s:5.1f}%")
                cur = [0, 0, 0, None]
        print(f"  {cur[3]:#07x}-end     n={cur[2]:5d} inst {100 * cur[0] / tot:5.1f}% "
              f"samples {100 * cur[1] / tot_s:5.1f}%")
        for r in sorted(body, key=lambda r: -int(r[iW]))[:top_n]:
            st = {hdr[k][6:]: int(r[k]) for k in stall if int(r[k]) > 0}
            st = dict(sorted(st.items(), key=lambda kv: -kv[1])[:3])
            print(f"    {int(r[iA], 16) - base:#07x} {r[iS].strip()[:46]:46s} n={int(r[iN]):9d} "
                  f"s={int(r[iW]):5d} {st}")


if __name__ == "__main__":
    main(sys.argv[1], int(sys.argv[2]) if len(sys.argv) > 2 else 25)
