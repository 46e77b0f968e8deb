#!/usr/bin/env python
"""DRAM traffic and device time per kernel group of ONE bench step, from an ncu launch list taken with

    ncu --profile-from-start off --clock-control none --csv --log-file step.csv \\
        --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum \\
        python bench.py --steps 2 --warmup 1 --no-cpu-baseline --trigger-steps 0 --e2e-steps 0 --profile-step

(bench.py --profile-step brackets one step with cudaProfilerStart/Stop).  Writes the JSON bench.py reads
(profiles/r2_traffic.json): key = kernel group as in the bench line's roofline.kernels ("K1", "K5a", ...), value =
dram__bytes_read.sum + dram__bytes_write.sum summed over the group's launches of the step; plus per-group launch counts and
times under "_detail" (cold-cache, serialised: shares, not absolutes) and the command under "source".

    python tools/ncu_traffic.py step.csv [full_table.csv] > profiles/r2_traffic.json
The optional second list is the same capture of `--early-chunk 0` (one K1 launch over every frame pair): key "K1_full"."""
import csv
import json
import re
import sys
from collections import OrderedDict

GROUPS = [("K1", r"k1_trigger_tma_kernel|k1_trigger_kernel|k1_fill_bin0"), ("K2", r"k2_"), ("K5a", r"k1_genesis|k5_genesis"),
          ("K5b", r"k5_post"), ("K5c", r"k5_otsu|k5_flag_fill"), ("K6", r"k6_"), ("cub", r"cub::|DeviceRadixSort|DeviceScan|DeviceSelect"),
          ("K7a", r"k7_trace1"), ("K7b", r"k7_trace2"), ("K7c", r"k7_external|k7_emit"), ("K3", r"k3_"), ("K4", r"k4_"), ("K8", r"k8_")]


def read(path):
    with open(path) as f:
        lines = [l for l in f if l.startswith('"')]
    rd = csv.reader(lines)
    hdr = next(rd)
    ki, mi, vi, ui, ii = (hdr.index(c) for c in ("Kernel Name", "Metric Name", "Metric Value", "Metric Unit", "ID"))
    launches = OrderedDict()
    for r in rd:
        d = launches.setdefault(r[ii], {"name": r[ki]})
        v = float(r[vi].replace(",", ""))
        if r[mi] == "gpu__time_duration.sum":
            v *= {"ns": 1e-3, "us": 1.0, "ms": 1e3, "s": 1e6}.get(r[ui], 1.0)
        else:
            v *= {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(r[ui], 1.0)
        d[r[mi]] = v
    return list(launches.values())


def group(launches):
    out = OrderedDict()
    other = {"launches": 0, "us": 0.0, "dram_bytes": 0.0, "kernels": {}}
    for l in launches:
        g = next((k for k, pat in GROUPS if re.search(pat, l["name"])), None)
        d = out.setdefault(g, {"launches": 0, "us": 0.0, "dram_bytes": 0.0, "kernels": {}}) if g else other
        d["launches"] += 1
        d["us"] += l.get("gpu__time_duration.sum", 0.0)
        d["dram_bytes"] += l.get("dram__bytes_read.sum", 0.0) + l.get("dram__bytes_write.sum", 0.0)
        nm = re.sub(r"\(.*", "", l["name"]).replace("void ", "")[:60]
        d["kernels"][nm] = d["kernels"].get(nm, 0) + 1
    if other["launches"]:
        out["other"] = other
    return out


def main():
    step = group(read(sys.argv[1]))
    res = OrderedDict()
    for k, d in step.items():
        if k != "other":
            res[k] = d["dram_bytes"]
    tot = sum(d["us"] for d in step.values())
    for d in step.values():
        d["share_of_step_device_time"] = d["us"] / tot if tot else None
    res["_detail"] = step
    if len(sys.argv) > 2:
        full = group(read(sys.argv[2]))
        if "K1" in full:
            res["K1_full"] = full["K1"]["dram_bytes"]
            res["_detail_full_table"] = {"K1": full["K1"]}
    res["source"] = ("tools/ncu_traffic.py over ncu --profile-from-start off --metrics gpu__time_duration.sum,dram__bytes_read.sum,"
                     "dram__bytes_write.sum of one bench step (bench.py --profile-step); sums over the group's launches of the step")
    json.dump(res, sys.stdout, indent=1)
    print()


if __name__ == "__main__":
    main()
