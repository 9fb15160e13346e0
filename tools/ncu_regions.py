#!/usr/bin/env python
"""Attribute ncu per-SASS-instruction counters of k_propagate to regions of propagate.cu.
Inlined math (mvs_math.cuh) is charged to the last propagate.cu line seen before it.

  tools/ncu_regions.py <source.csv> <nvdisasm -g listing> <kernel substring> <file> <line:name,...>
"""
import csv, re, sys
from collections import defaultdict

def main():
    src_csv, dis, kname, fname, spec = sys.argv[1:6]
    bounds = [(int(a.split(":")[0]), a.split(":")[1]) for a in spec.split(",")]
    lines = []; infn = False; last = 0
    for ln in open(dis, errors="replace"):
        if ln.startswith(".text.") or re.match(r"\s*\.section\s+\.text\.", ln):
            infn = kname in ln
        if not infn:
            continue
        m = re.search(r'//## File "([^"]+)", line (\d+)', ln)
        if m:
            if m.group(1).endswith(fname):
                last = int(m.group(2))
            continue
        if re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+\S", ln) and ".byte" not in ln:
            lines.append(last)
    rows = list(csv.reader(open(src_csv))); h = rows[1]
    ii, si, ti = h.index("Instructions Executed"), h.index("# Samples"), h.index("Thread Instructions Executed")
    data = [r for r in rows[2:] if len(r) > ii]
    reg = defaultdict(lambda: [0, 0, 0])
    def region(l):
        name = "?"
        for b, n in bounds:
            if l >= b: name = n
        return name
    for k, r in enumerate(data):
        rg = region(lines[k] if k < len(lines) else 0)
        reg[rg][0] += int(r[ii]); reg[rg][1] += int(r[si]); reg[rg][2] += int(r[ti])
    tot = sum(v[0] for v in reg.values()); ts = sum(v[1] for v in reg.values())
    print(f"total warp instructions {tot}, samples {ts}")
    for b, n in bounds:
        v = reg[n]
        print(f"{n:28s} inst {v[0]/tot*100:5.1f}%  samples {v[1]/max(ts,1)*100:5.1f}%  lanes {v[2]/max(v[0],1):.1f}")

if __name__ == "__main__":
    main()
