#!/usr/bin/env python3
"""Summarise ncu captures for profiles/: per-kernel DRAM traffic and headline metrics from a
`--set full` report (.ncu-rep), written as JSON that bench.py reads for roofline.traffic.

    python tools/ncu_traffic.py <report.ncu-rep> <frames_per_launch> [out.json]
"""
import csv
import io
import json
import subprocess
import sys

WANT = {
    "dram__bytes_read.sum": "dram_read", "dram__bytes_write.sum": "dram_write",
    "gpu__time_duration.sum": "duration", "sm__inst_executed.sum": "warp_instructions", "smsp__inst_executed.sum": "warp_instructions",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed": "sm_throughput_pct",
    "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed": "dram_throughput_pct",
    "sm__warps_active.avg.pct_of_peak_sustained_active": "achieved_occupancy_pct",
    "smsp__thread_inst_executed_per_inst_executed.ratio": "active_threads_per_warp",
    "launch__registers_per_thread": "registers", "lts__t_bytes.sum": "l2_bytes",
    "sm__inst_issued.avg.pct_of_peak_sustained_active": "issue_slots_busy_pct",
}
UNIT = {"Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "byte": 1.0, "us": 1e-6, "ms": 1e-3, "ns": 1e-9, "s": 1.0,
        "usecond": 1e-6, "msecond": 1e-3, "nsecond": 1e-9, "second": 1.0}


def main():
    rep, frames = sys.argv[1], int(sys.argv[2])
    out = sys.argv[3] if len(sys.argv) > 3 else None
    txt = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True, check=True).stdout
    rows = list(csv.reader(io.StringIO(txt)))
    hdr, units = rows[0], rows[1]
    res = {}
    for r in rows[2:]:
        d = dict(zip(hdr, r))
        name = d["Kernel Name"].split("(")[0].replace("void ", "").split("<")[0]
        e = {"frames_per_launch": frames}
        for i, h in enumerate(hdr):
            if h in WANT and r[i] not in ("", "n/a"):
                v = float(r[i].replace(",", ""))
                e[WANT[h]] = v * UNIT.get(units[i], 1.0)
        if "dram_read" in e and "dram_write" in e:
            e["dram_bytes"] = e["dram_read"] + e["dram_write"]
        res[name] = e
    s = json.dumps(res, indent=1)
    print(s)
    if out:
        try:
            old = json.loads(open(out).read())
        except Exception:
            old = {}
        old.update(res)
        open(out, "w").write(json.dumps(old, indent=1) + "\n")


if __name__ == "__main__":
    main()
