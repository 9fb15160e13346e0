#!/usr/bin/env python
"""Attribute ncu per-SASS-instruction counters to CUDA source lines.

  tools/ncu_lines.py <kernel.source.csv from `ncu --page source --csv`> <nvdisasm -g listing> <kernel name substring>

The ncu CLI prints metrics per SASS instruction only; nvdisasm -g gives the line of each
instruction (inlined callees report their own file/line).  Instructions are matched by order."""
import csv
import re
import sys
from collections import defaultdict


def main():
    src_csv, dis, kname = sys.argv[1:4]
    top = int(sys.argv[4]) if len(sys.argv) > 4 else 40
    lines = []          # (file, line) per instruction, in order
    cur = ("?", 0)
    infn = False
    for ln in open(dis, errors="replace"):
        if ln.startswith(".text.") or re.match(r"\s*\.section\s+\.text\.", ln):
            infn = kname in ln
        if not infn:
            continue
        m = re.search(r'//## File "([^"]+)", line (\d+)', ln)
        if m:
            cur = (m.group(1).split("/")[-1], int(m.group(2)))
            continue
        if re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+\S", ln) and ".byte" not in ln:
            lines.append(cur)
    rows = list(csv.reader(open(src_csv)))
    h = rows[1]
    ii, si, src = h.index("Instructions Executed"), h.index("# Samples"), h.index("Source")
    ti = h.index("Thread Instructions Executed")
    data = [r for r in rows[2:] if len(r) > ii]
    print(f"{len(lines)} instructions in listing, {len(data)} in ncu report")
    inst, smp, thr = defaultdict(int), defaultdict(int), defaultdict(int)
    for k, r in enumerate(data):
        key = lines[k] if k < len(lines) else ("?", 0)
        inst[key] += int(r[ii]); smp[key] += int(r[si]); thr[key] += int(r[ti])
    tot, tots = sum(inst.values()), sum(smp.values())
    print(f"total warp instructions {tot}, samples {tots}")
    for key, n in sorted(inst.items(), key=lambda kv: -kv[1])[:top]:
        print(f"{n / tot * 100:5.1f}% inst {smp[key] / max(tots, 1) * 100:5.1f}% smp  lanes {thr[key] / max(n, 1):4.1f}  {key[0]}:{key[1]}")


if __name__ == "__main__":
    main()
