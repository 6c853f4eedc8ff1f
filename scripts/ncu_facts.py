#!/usr/bin/env python
"""Distil `ncu --set full` raw pages (ncu -i x.ncu-rep --page raw --csv) into the few per-launch
facts bench.py quotes: warp instructions executed, DRAM bytes read + written, duration, issue
activity, registers.  Usage:
    python scripts/ncu_facts.py c4s=profiles/r02/ncu_full_c4s.raw.csv#134217728 c2=... > profiles/r02/ncu_facts.json
(`#units` = waypoint-samples the captured launch processed; `@text` picks a kernel by name.)
bench.py reads profiles/<round>/ncu_facts.json (never anything measured under the profiler for
its timings: only the instruction count and the DRAM traffic, which do not depend on clocks)."""
import csv
import json
import sys

UNIT = {"ns": 1e-9, "us": 1e-6, "ms": 1e-3, "s": 1.0, "byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "nsecond": 1e-9, "usecond": 1e-6,
        "msecond": 1e-3, "second": 1.0}


def facts(path, pick=None):
    rows = list(csv.reader(open(path)))
    hdr, units, data = rows[0], rows[1], rows[2:]
    if pick:
        data = [r for r in data if pick in r[hdr.index("Kernel Name")]] or data
    r = data[-1]

    def val(name, scale=True):
        if name not in hdr:
            return None
        i = hdr.index(name)
        try:
            v = float(r[i].replace(",", ""))
        except ValueError:
            return None
        return v * UNIT.get(units[i], 1.0) if scale else v

    rd, wr = val("dram__bytes_read.sum"), val("dram__bytes_write.sum")
    return {
        "kernel": r[hdr.index("Kernel Name")],
        "grid": r[hdr.index("Grid Size")], "block": r[hdr.index("Block Size")],
        "warp_inst_per_launch": val("smsp__inst_executed.sum", False),
        "dram_bytes_per_launch": None if rd is None else rd + (wr or 0.0),
        "duration_us_under_ncu": (val("gpu__time_duration.sum") or 0.0) * 1e6,
        "issue_active_pct": val("smsp__issue_active.avg.pct_of_peak_sustained_active", False),
        "registers": val("launch__registers_per_thread", False),
        "achieved_occupancy_pct": val("sm__warps_active.avg.pct_of_peak_sustained_active", False),
        "source": path,
    }


def main():
    out = {}
    for arg in sys.argv[1:]:
        key, path = arg.split("=", 1)
        pick, units = None, None
        if "#" in path:
            path, units = path.split("#", 1)
        if "@" in path:
            path, pick = path.split("@", 1)
        out[key] = facts(path, pick)
        if units:
            u = float(units)
            out[key]["units_per_launch"] = u
            out[key]["warp_inst_per_wps"] = out[key]["warp_inst_per_launch"] / u
            out[key]["thread_inst_per_wps"] = 32.0 * out[key]["warp_inst_per_launch"] / u
    json.dump(out, sys.stdout, indent=1)
    print()


if __name__ == "__main__":
    main()
