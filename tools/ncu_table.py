#!/usr/bin/env python
"""One markdown row per distinct kernel of an `ncu --page raw --csv` dump (the LAST captured instance of each,
i.e. a warm one): duration, DRAM bytes, achieved DRAM bandwidth as a fraction of the measured HBM peak
(MEASURED_PEAKS.json), registers, grid, issue-slot utilisation and the dominant stall reason.

    ncu -i prof.ncu-rep --page raw --csv > raw.csv && python tools/ncu_table.py raw.csv [more.csv ...]
"""
import csv
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def peak_gbs():
    try:
        return float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]), "measured"
    except Exception:  # noqa: BLE001
        return 6650.0, "fallback"


def rows_of(path):
    rows = list(csv.reader(open(path)))
    hdr, units = rows[0], rows[1]
    idx = {h: i for i, h in enumerate(hdr)}
    out = {}
    for r in rows[2:]:
        name = r[idx["Kernel Name"]].strip()

        def val(k, default=0.0):
            try:
                v = float(r[idx[k]].replace(",", ""))
            except (KeyError, ValueError):
                return default
            u = units[idx[k]].lower()
            if k.startswith("dram__bytes"):
                v *= {"byte": 1, "kbyte": 1e3, "mbyte": 1e6, "gbyte": 1e9}.get(u, 1)
            if k == "gpu__time_duration.sum":
                v *= {"ns": 1e-3, "us": 1.0, "usecond": 1.0, "ms": 1e3, "msecond": 1e3, "nsecond": 1e-3}.get(u, 1.0)
            return v
        stalls = []
        for i, h in enumerate(hdr):
            if "average_warps_issue_stalled" in h and "not_issued" not in h:
                try:
                    stalls.append((float(r[i]), h.replace("smsp__average_warps_issue_stalled_", "").replace("_per_issue_active.ratio", "")))
                except ValueError:
                    pass
        stalls = [s for s in sorted(stalls, reverse=True) if s[1] != "selected"]
        out[name] = {"us": val("gpu__time_duration.sum"), "rd": val("dram__bytes_read.sum"), "wr": val("dram__bytes_write.sum"),
                     "regs": int(val("launch__registers_per_thread")), "grid": int(val("launch__grid_size")),
                     "issue": val("smsp__issue_active.avg.pct_of_peak_sustained_active"),
                     "occ": val("sm__warps_active.avg.pct_of_peak_sustained_active"),
                     "stall": "%s %.1f" % (stalls[0][1], stalls[0][0]) if stalls else "-"}
    return out


def main(paths):
    peak, src = peak_gbs()
    print("| kernel | time µs | DRAM read / write MB | DRAM GB/s | of %s HBM (%.0f GB/s) | regs | grid | warps active %% | issue active %% | top stall (warps per issue) |" % (src, peak))
    print("|---|---|---|---|---|---|---|---|---|---|")
    for p in paths:
        for name, k in rows_of(p).items():
            gbs = (k["rd"] + k["wr"]) / (k["us"] * 1e-6) / 1e9 if k["us"] else 0.0
            short = name.replace("void ", "")
            short = short[: short.index("(")] if "(" in short else short
            print("| `%s` | %.1f | %.1f / %.1f | %.0f | %.2f | %d | %d | %.0f | %.0f | %s |" % (
                short, k["us"], k["rd"] / 1e6, k["wr"] / 1e6, gbs, gbs / peak, k["regs"], k["grid"], k["occ"], k["issue"], k["stall"]))


if __name__ == "__main__":
    main(sys.argv[1:])
