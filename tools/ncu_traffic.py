#!/usr/bin/env python
"""DRAM traffic per kernel launch from an `ncu --set full` report -> JSON (bench.py's roofline.traffic reads it).

    python tools/ncu_traffic.py gpurun_out/x.ncu-rep "<command that was profiled>" > profiles/attn_instep_traffic_r01.json
"""
import csv
import json
import subprocess
import sys

rep, cmd = sys.argv[1], sys.argv[2]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
hdr, units = rows[0], rows[1]
ix = {h: i for i, h in enumerate(hdr)}


def to_bytes(v, unit):
    v = float(v.replace(",", ""))
    return v * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}[unit]


out = {"command": cmd, "kernels": []}
for r in rows[2:]:
    rd = to_bytes(r[ix["dram__bytes_read.sum"]], units[ix["dram__bytes_read.sum"]])
    wr = to_bytes(r[ix["dram__bytes_write.sum"]], units[ix["dram__bytes_write.sum"]])
    dur = float(r[ix["gpu__time_duration.sum"]].replace(",", ""))
    dunit = units[ix["gpu__time_duration.sum"]]
    dur_us = dur * {"ns": 1e-3, "us": 1.0, "ms": 1e3, "usecond": 1.0, "nsecond": 1e-3, "msecond": 1e3}.get(dunit, 1.0)
    out["kernels"].append({"name": r[ix["Kernel Name"]].split("(")[0], "grid": r[ix["Grid Size"]] if "Grid Size" in ix else None,
                           "dram_read_bytes": rd, "dram_write_bytes": wr, "dram_bytes": rd + wr, "duration_us": dur_us})
print(json.dumps(out, indent=1))
