#!/usr/bin/env python3
"""Turn what `tools/profile_run.sh <round>` brought back in gpurun_out/ into the tracked files under profiles/:

    python tools/collect_profiles.py r2

* profiles/traffic.json                     (tools/make_traffic.py on the benchmarked kernel's capture)
* profiles/<round>_launches_summary.md      + the launch list CSV
* profiles/<round>_bench_default.json       the default bench line of the same session
* profiles/<round>_noalloc_timing.md, <round>_scale_geometries.txt, <round>_convert_single_digest.txt
* profiles/sass_scale_tiled_kernel.txt      SASS excerpt of the config-3 instantiation of the built library
* prints the digest of the two full captures (the figures of <round>_ncu_summary.md are typed from it)
"""
import collections
import csv
import json
import os
import re
import shutil
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
R = sys.argv[1] if len(sys.argv) > 1 else "r2"
G, P = os.path.join(ROOT, "gpurun_out"), os.path.join(ROOT, "profiles")
SO = os.path.join(ROOT, "selkies_gstreamer_b200", "libselkies_preencode.so")


def run(*a):
    return subprocess.run(list(a), capture_output=True, text=True).stdout


# traffic.json
print(run(sys.executable, os.path.join(ROOT, "tools", "make_traffic.py"), os.path.join(G, f"{R}_convert_full.ncu-rep"), "32",
          "ncu --set full --clock-control none, bench.py --steps 20 --warmup 3, launch 73 of sk_convert_kernel")[:300])

# launch list
rows = [r for r in csv.reader(open(os.path.join(G, f"{R}_launches.csv"))) if len(r) > 5]
hdr = [i for i, r in enumerate(rows) if r and r[0] == "ID"][0]
H = rows[hdr]
ki, vi = H.index("Kernel Name"), H.index("Metric Value")
agg = collections.defaultdict(list)
for r in rows[hdr + 1:]:
    try:
        agg[r[ki].split("(")[0]].append(float(r[vi].replace(",", "")))
    except ValueError:
        pass
md = [f"# Round {R[1:]} — ncu launch list of `python bench.py --steps 20 --warmup 3 --no-configs --no-side --no-cpu-baseline`", "",
      "`ncu --metrics gpu__time_duration.sum --clock-control none -c 400` (serialised, cold-cache per launch; unit ns). "
      f"Full CSV: `{R}_launches_bench_steps20.csv`.", "", "| kernel | launches | total | mean | min | max |", "|---|---|---|---|---|---|"]
for k, v in sorted(agg.items(), key=lambda kv: -sum(kv[1])):
    md.append(f"| `{k}` | {len(v)} | {sum(v):.0f} | {sum(v) / len(v):.1f} | {min(v):.1f} | {max(v):.1f} |")
open(os.path.join(P, f"{R}_launches_summary.md"), "w").write("\n".join(md) + "\n")
shutil.copy(os.path.join(G, f"{R}_launches.csv"), os.path.join(P, f"{R}_launches_bench_steps20.csv"))

# bench line and small tables
line = json.loads(open(os.path.join(G, f"{R}_bench.log")).read().strip().splitlines()[-1])
json.dump(line, open(os.path.join(P, f"{R}_bench_default.json"), "w"), indent=1)
for name in (f"{R}_noalloc_timing.md", f"{R}_scale_geometries.txt", f"{R}_convert_single_digest.txt"):
    if os.path.exists(os.path.join(G, name)):
        shutil.copy(os.path.join(G, name), os.path.join(P, name))

# SASS excerpt of the config-3 scale kernel
names = [l for l in run("cuobjdump", "-sass", SO).splitlines() if "Function :" in l and "sk_scale_tiled_kernelILi1ELi1ELb1ELb1ELb1ELi7ELi32ELi192ELb0" in l]
if names:
    fn = names[0].split("Function :")[1].strip()
    ins = []
    for l in run("cuobjdump", "-sass", "-fun", fn, SO).splitlines():
        m = re.match(r"\s*/\*([0-9a-f]{4})\*/\s+(.*?);", l)
        if m:
            ins.append((m.group(1), re.sub(r"\s+", " ", m.group(2)).strip()))
    out = ["sk_scale_tiled_kernel<I420,h-cosited,DAMAGE,SH,SV,RPW=7,TR=32,RS=192,AFTER=0> — release build, cuobjdump -sass (sm_100a)",
           f"instructions: {len(ins)} (interior + edge path of one instantiation; 2592 at the start of round 2's last session)", "",
           "--- TMA / mbarrier instructions (the raw source window of a tile is ONE 2-D tensor-map copy; UTMAPF = the",
           "    successor tile's window asked into L2)"]
    out += [f"/*{a}*/ {t}" for a, t in ins if re.search(r"UTMALDG|UTMAPF|SYNCS|ELECT", t)]
    out += ["", "--- one pass-A row of the interior path: 8 LDS and 4 STS with immediate row offsets (compile-time stride 192 words),",
            "    no address arithmetic, 4 x (2 PRMT + 3 IDP.2A + 2 PRMT)"]
    idx = [i for i, (a, t) in enumerate(ins) if t.startswith("BRA.U UP")]
    start = idx[1] if len(idx) > 1 else 0
    out += [f"/*{a}*/ {t}" for a, t in ins[start:start + 44]]
    out += ["", "--- opcode histogram"]
    h = collections.Counter(re.sub(r"^@!?U?P\d+ ", "", t).split()[0].split(".")[0] for a, t in ins)
    out += [f"{k:<10} {v}" for k, v in h.most_common(14)]
    open(os.path.join(P, "sass_scale_tiled_kernel.txt"), "w").write("\n".join(out) + "\n")

for rep in (f"{R}_convert_full", f"{R}_scale_full"):
    print("==", rep)
    print("\n".join(run(sys.executable, os.path.join(ROOT, "tools", "ncu_digest.py"), os.path.join(G, rep + ".ncu-rep"), "2").splitlines()[:22]))
print({k: line[k] for k in ("value", "ms_per_step")}, line["roofline"]["frac"], line["roofline"]["traffic"], line["e2e"]["value"])
for k, v in line["configs"].items():
    print(k, round(v.get("kernel_us_per_launch", 0), 2), round(v["roofline"]["frac"], 3), round(v["frames_per_s"]))
