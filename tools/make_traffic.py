#!/usr/bin/env python3
"""profiles/traffic.json from one `ncu --set full` capture of the benchmarked no-scale kernel.

    python tools/make_traffic.py gpurun_out/r2_convert_full.ncu-rep 32 "how it was captured"

Writes dram__bytes_read.sum + dram__bytes_write.sum of the captured launch, divided by the sessions of
that launch, together with the sha of the kernel source the library was built from: bench.py reports
the figure as `roofline.traffic` only while that sha still matches (a later kernel edit makes it stale)."""
import csv
import io
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
rep, sessions = sys.argv[1], int(sys.argv[2])
how = sys.argv[3] if len(sys.argv) > 3 else "ncu --set full --clock-control none"
out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True, check=True).stdout
rows = list(csv.reader(io.StringIO(out)))
hdr, units, vals = rows[0], rows[1], rows[2]


def metric(name):
    i = hdr.index(name)
    scale = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}[units[i]]
    return float(vals[i]) * scale


rd, wr = metric("dram__bytes_read.sum"), metric("dram__bytes_write.sum")
sys.path.insert(0, ROOT)
from bench import code_sha  # noqa: E402  (comments and white space do not count)
j = {"dram_bytes_per_frame": (rd + wr) / sessions,
     "kernel_sha": code_sha(os.path.join(ROOT, "selkies_gstreamer_b200", "csrc", "sk_kernels.cu")),
     "kernel": vals[hdr.index("Kernel Name")] if "Kernel Name" in hdr else "sk_convert_kernel",
     "sessions_per_launch": sessions, "dram_bytes_read": rd, "dram_bytes_write": wr,
     "source": f"{os.path.relpath(os.path.abspath(rep), ROOT)} ({how})"}
json.dump(j, open(os.path.join(ROOT, "profiles", "traffic.json"), "w"), indent=1)
print(json.dumps(j))
