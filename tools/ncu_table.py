#!/usr/bin/env python
"""Per-launch table from an `ncu --page raw --csv` export: time, DRAM traffic and rate against the
measured HBM peak, issue / L1 utilisation, cache hit rates, coalescing (sectors per request, useful bytes
per sector) for every kernel launch in the report.
  tools/ncu_table.py <raw.csv> [peak GB/s] > profiles/<name>_summary.md"""
import csv
import json
import os
import sys

UNIT = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12, "ns": 1e-9, "us": 1e-6, "ms": 1e-3, "s": 1,
        "usecond": 1e-6, "msecond": 1e-3, "nsecond": 1e-9, "second": 1}


def main():
    path = sys.argv[1]
    peak = float(sys.argv[2]) if len(sys.argv) > 2 else None
    if peak is None:
        try:
            mp = json.load(open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")))
            peak = float(mp.get("hbm_gbs") or 6556.8)
        except Exception:
            peak = 6556.8
    rows = list(csv.reader(open(path)))
    h, u = rows[0], rows[1]
    col = {n: i for i, n in enumerate(h)}

    def get(r, name, scale=True):
        i = col.get(name)
        if i is None or i >= len(r) or r[i] == "":
            return float("nan")
        v = float(r[i].replace(",", ""))
        return v * UNIT.get(u[i], 1) if scale else v

    print("| # | kernel | grid x block | regs | time us | DRAM rd MB | DRAM wr MB | DRAM GB/s | %% of %.0f | issue %% | L1 thr %% | "
          "L1 hit %% | L2 hit %% | ld sect/req | ld B/sector | st sect/req |" % peak)
    print("|" + "---|" * 16)
    for k, r in enumerate(rows[2:]):
        if len(r) < 10:
            continue
        name = r[col["Kernel Name"]].split("(")[0].replace("void ", "").replace("mvs::", "")
        t = get(r, "gpu__time_duration.sum")
        rd, wr = get(r, "dram__bytes_read.sum"), get(r, "dram__bytes_write.sum")
        gbps = (rd + wr) / t / 1e9
        lreq, lsec = get(r, "l1tex__t_requests_pipe_lsu_mem_global_op_ld.sum"), get(r, "l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum")
        sreq, ssec = get(r, "l1tex__t_requests_pipe_lsu_mem_global_op_st.sum"), get(r, "l1tex__t_sectors_pipe_lsu_mem_global_op_st.sum")
        print("| %d | %s | %s x %s | %d | %.1f | %.1f | %.1f | %.0f | %.1f | %.1f | %.1f | %.1f | %.1f | %.2f | %.1f | %.2f |" % (
            k, name, r[col["Grid Size"]], r[col["Block Size"]], get(r, "launch__registers_per_thread"), t * 1e6, rd / 1e6,
            wr / 1e6, gbps, 100 * gbps / peak, get(r, "sm__issue_active.avg.pct_of_peak_sustained_elapsed"),
            get(r, "l1tex__throughput.avg.pct_of_peak_sustained_elapsed"), get(r, "l1tex__t_sector_hit_rate.pct"),
            get(r, "lts__t_sector_hit_rate.pct"), lsec / lreq if lreq else float("nan"),
            get(r, "smsp__sass_average_data_bytes_per_sector_mem_global_op_ld.ratio"), ssec / sreq if sreq else float("nan")))


if __name__ == "__main__":
    main()
