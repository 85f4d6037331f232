#!/usr/bin/env python3
"""Produce golden vectors from a REAL GStreamer for tests/test_gst_golden.py.

Where this project was built there is no GStreamer (no gi, no gst-launch-1.0, no network), so its
oracle is a restatement of GStreamer 1.24.12's generic converter and parity is "unpinned".  Run this
on any machine that has the version the reference pins (addons/gstreamer/Dockerfile:85-93: 1.24.12,
-Dorc=enabled, videoconvertscale enabled; docs/component.md:314 requires >= 1.22):

    python tools/make_gst_golden.py            # writes tests/golden/gst/*.npz (+ MANIFEST.json)
    python -m pytest tests/test_gst_golden.py  # oracle vs GStreamer on any host; add -m gpu on a B200

Each case writes the input frame as raw BGRx / RGBx, pushes it through

    gst-launch-1.0 filesrc ! rawvideoparse ! <stage> ! video/x-raw,format=..,width=..,height=..[,colorimetry=..,chroma-site=..] ! filesink

where <stage> is exactly what the reference instantiates — `videoconvert n-threads=N`
(src/selkies/legacy/gstwebrtc_app.py:611-613) — or, for the scaling case the reference has no call site
for, `videoconvertscale method=bilinear` (ONE converter: scale and matrix in the same chain, which is
what oracle/sk_oracle.c restates) and, separately labelled, `videoconvert ! videoscale` (TWO converters:
GStreamer is then free to scale after the conversion, in the YUV domain — a different computation).
The caps GStreamer negotiated (`-v`) are stored beside the pixels: they answer SURVEY.md A.8 R1/R2
(default colorimetry and chroma-site by height).  Cases without colorimetry/chroma-site in the caps
filter exercise those defaults.

Nothing here imports this project's kernels or oracle: the vectors are GStreamer's alone.
"""
from __future__ import annotations

import argparse
import json
import os
import re
import shutil
import subprocess
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from selkies_gstreamer_b200 import synth  # noqa: E402  (pure numpy frame generators, no GPU)

KAT_RGB = [(0, 0, 0), (255, 255, 255), (255, 0, 0), (0, 255, 0), (0, 0, 255), (255, 255, 0), (0, 255, 255),
           (255, 0, 255), (128, 128, 128), (1, 2, 3), (200, 100, 50)]


# {n} = n-threads, as the reference sets it (gstwebrtc_app.py:612)
CONVERT = "videoconvert n-threads={n}"
CONVERTSCALE = "videoconvertscale method=bilinear n-threads={n}"
CONVERT_THEN_SCALE = "videoconvert n-threads={n} ! videoscale method=bilinear"


def flat(rgb, w, h):
    f = np.empty((h, w, 4), np.uint8)
    f[..., 0], f[..., 1], f[..., 2], f[..., 3] = rgb[2], rgb[1], rgb[0], 0xFF
    return f


def cases():
    """(name, frame BGRx, in_fmt, out_w, out_h, out_fmt, colorimetry or None, chroma-site or None, stage)"""
    out = []
    for i, rgb in enumerate(KAT_RGB):                       # SURVEY A.7 known-answer colours, all matrices
        for col in ("bt601", "bt709", "bt2020"):
            out.append((f"kat{i:02d}_{col}", flat(rgb, 16, 4), "bgrx", 16, 4, "I420", col, "jpeg", CONVERT))
    for (w, h) in ((34, 18), (130, 34), (1920, 1080)):
        for col, site in (("bt601", "jpeg"), ("bt709", "mpeg2"), ("bt2020", "mpeg2")):
            for fmt in ("I420", "NV12"):
                out.append((f"noise_{w}x{h}_{col}_{site}_{fmt}", synth.noise(w, h, seed=1234), "bgrx", w, h, fmt, col,
                            site, CONVERT))
        out.append((f"bars_{w}x{h}", synth.bars(w, h), "bgrx", w, h, "I420", "bt709", "mpeg2", CONVERT))
        out.append((f"bars_{w}x{h}_jpeg", synth.bars(w, h), "bgrx", w, h, "I420", "bt601", "jpeg", CONVERT))
        # no colorimetry / chroma-site in the filter: GStreamer's defaults by height (R1, R2)
        out.append((f"defaults_{w}x{h}", synth.noise(w, h, seed=7), "bgrx", w, h, "I420", None, None, CONVERT))
    out.append(("defaults_3840x2160", synth.noise(3840, 2160, seed=8), "bgrx", 3840, 2160, "NV12", None, None, CONVERT))
    out.append(("edge_1920x1080_rgbx", synth.desktop(1920, 1080)[..., [2, 1, 0, 3]].copy(), "rgbx", 1920, 1080, "NV12",
                "bt709", "mpeg2", CONVERT))
    # scaling (BASELINE.json config 3 and small ragged ones), one converter and two converters
    # ... and enlarging ones (more output than input pixels: upstream then scales behind the matrix)
    for (w, h, ow, oh) in ((2560, 1440, 1920, 1080), (64, 48, 32, 24), (100, 60, 98, 34), (64, 48, 64, 24), (64, 48, 30, 48),
                           (1280, 720, 1920, 1080), (32, 24, 64, 48), (64, 24, 20, 96), (96, 20, 128, 18)):
        f = synth.noise(w, h, seed=w + oh)
        for stage in (CONVERTSCALE, CONVERT_THEN_SCALE):
            tag = "one" if stage is CONVERTSCALE else "two"
            out.append((f"scale_{w}x{h}_to_{ow}x{oh}_{tag}", f, "bgrx", ow, oh, "I420", "bt709", "mpeg2", stage))
    return out


def run_case(gst, tmp, name, frame, in_fmt, ow, oh, out_fmt, col, site, stage, n_threads):
    h, w = frame.shape[:2]
    src, dst = os.path.join(tmp, "in.raw"), os.path.join(tmp, "out.raw")
    frame.tofile(src)
    caps = f"video/x-raw,format={out_fmt},width={ow},height={oh}"
    if col:
        caps += f",colorimetry={col}"
    if site:
        caps += f",chroma-site={site}"
    stage = stage.format(n=n_threads)
    cmd = (f"{gst} -v filesrc location={src} blocksize={frame.nbytes} ! rawvideoparse format={in_fmt} width={w} height={h} "
           f"framerate=1/1 ! {stage} ! {caps} ! filesink location={dst}")
    p = subprocess.run(cmd, shell=True, capture_output=True, text=True, timeout=300)
    if p.returncode != 0 or not os.path.exists(dst):
        raise RuntimeError(f"{name}: gst-launch failed\n{cmd}\n{p.stderr[-2000:]}")
    out = np.fromfile(dst, np.uint8)
    negotiated = [l.strip() for l in p.stdout.splitlines() if "caps = video/x-raw" in l]
    os.remove(dst)
    return out, negotiated, cmd


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default=os.path.join(ROOT, "tests", "golden", "gst"))
    ap.add_argument("--gst-launch", default=shutil.which("gst-launch-1.0") or "gst-launch-1.0")
    ap.add_argument("--n-threads", type=int, default=1, help="videoconvert n-threads (results must not depend on it: R10)")
    args = ap.parse_args()
    try:
        ver = subprocess.run([args.gst_launch, "--version"], capture_output=True, text=True, timeout=30).stdout
    except OSError as e:
        raise SystemExit(f"{args.gst_launch} not found ({e}): run this where GStreamer >= 1.22 (ideally 1.24.12) is installed")
    m = re.search(r"GStreamer (\d+\.\d+\.\d+)", ver)
    version = m.group(1) if m else "unknown"
    os.makedirs(args.out, exist_ok=True)
    manifest = {"gstreamer": version, "n_threads": args.n_threads, "cases": []}
    with tempfile.TemporaryDirectory() as tmp:
        for (name, frame, in_fmt, ow, oh, out_fmt, col, site, stage) in cases():
            out, negotiated, cmd = run_case(args.gst_launch, tmp, name, frame, in_fmt, ow, oh, out_fmt, col, site, stage,
                                            args.n_threads)
            np.savez_compressed(os.path.join(args.out, name + ".npz"), input=frame, output=out,
                                meta=json.dumps({"in_fmt": in_fmt, "out_w": ow, "out_h": oh, "out_fmt": out_fmt,
                                                 "colorimetry": col, "chroma_site": site, "stage": stage,
                                                 "negotiated_caps": negotiated, "command": cmd, "gstreamer": version}))
            manifest["cases"].append(name)
            print(f"{name}: {out.nbytes} bytes", flush=True)
    json.dump(manifest, open(os.path.join(args.out, "MANIFEST.json"), "w"), indent=1)
    print(f"wrote {len(manifest['cases'])} cases from GStreamer {version} to {args.out}")


if __name__ == "__main__":
    main()
