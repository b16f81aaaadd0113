#!/usr/bin/env python
"""profiles/traffic.json from ncu --set full captures of the final build: DRAM bytes per launch of the dominant kernels.

    python profiles/make_traffic.py student=gpurun_out/prof_student.ncu-rep pool=gpurun_out/prof_pool.ncu-rep [note]
bench.py reads the file for `roofline.traffic` (ncu cannot run inside the timed bench)."""
import csv
import json
import os
import subprocess
import sys

out = {}
note = " ".join(a for a in sys.argv[1:] if "=" not in a)
for arg in sys.argv[1:]:
    if "=" not in arg:
        continue
    name, path = arg.split("=", 1)
    txt = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(txt.splitlines()))
    hdr = rows[0]
    body = rows[2:]
    rd, wr, du = hdr.index("dram__bytes_read.sum"), hdr.index("dram__bytes_write.sum"), hdr.index("gpu__time_duration.sum")
    units = rows[1]
    scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}

    def val(r, i):
        return float(r[i]) * scale[units[i]]
    tot = [val(r, rd) + val(r, wr) for r in body]
    tscale = {"ns": 1e-3, "us": 1.0, "ms": 1e3}[units[du]]
    out[name] = {"kernel": body[0][hdr.index("Kernel Name")].split("(")[0],
                 "dram_bytes_per_launch": sum(tot) / len(tot), "launches_captured": len(tot),
                 "ncu_duration_us": sum(float(r[du]) * tscale for r in body) / len(body),
                 "source": f"{os.path.basename(path)} (ncu --set full, N=150000 scene, dram__bytes_read.sum + "
                           f"dram__bytes_write.sum){' ' + note if note else ''}"}
json.dump(out, open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "traffic.json"), "w"), indent=1)
print(json.dumps(out, indent=1))
