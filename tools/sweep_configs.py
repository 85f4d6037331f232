#!/usr/bin/env python3
"""Device-resident throughput of every BASELINE.json geometry at several batch sizes.

    python tools/sweep_configs.py [--steps 200] > gpurun_out/sweep.md

Prints a markdown table: frames/s (CUDA events over all steps), mean kernel µs per launch,
algorithmic GB/s and the fraction of the measured HBM copy peak.  Same timing method as bench.py.
"""
import argparse
import json
import os
import statistics
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import selkies_gstreamer_b200 as sk  # noqa: E402
from selkies_gstreamer_b200 import synth  # noqa: E402

CONFIGS = [
    ("1080p BGRx->I420 bt709", 1920, 1080, 1920, 1080, "I420"),
    ("4K BGRx->NV12 bt709", 3840, 2160, 3840, 2160, "NV12"),
    ("1440p->1080p bilinear I420 bt709", 2560, 1440, 1920, 1080, "I420"),
    ("8K BGRx->I420 bt709", 7680, 4320, 7680, 4320, "I420"),
]


def run(name, w, h, ow, oh, fmt, batch, steps, peak):
    sess = [sk.PreencodeSession(w, h, out_width=ow, out_height=oh, out_format=fmt, colorimetry="bt709",
                                chroma_site="mpeg2", host_output=False, timing=(i == 0)) for i in range(batch)]
    try:
        base = synth.desktop(w, h)
        nxt = synth.mutate(base, 0.05, seed=1, aligned=False)
        for s in sess:
            s.upload(0, base); s.wait(s.submit(0))
            s.upload(1, nxt); s.wait(s.submit(1))
        b = sk.SessionBatch(sess)
        sets = [b.frame_set([s.device_input(k)[0] for s in sess]) for k in (0, 1)]
        stride = sess[0].device_input(0)[1]
        fid = 100
        for k in range(5):
            b.submit(sets[k & 1], stride, fid); fid += 1
        b.wait(fid - 1)
        kus = []
        sess[0].marker(0)
        for k in range(steps):
            b.submit(sets[k & 1], stride, fid)
            b.wait(fid - 1)
            kus.append(sess[0].last_kernel_us())
            fid += 1
        sess[0].marker(1)
        b.wait(fid - 1)
        us = sess[0].marker_elapsed_us()
        fps = batch * steps / (us * 1e-6)
        k_us = statistics.mean(kus[1:])
        bytes_launch = synth.algorithmic_bytes(w, h, ow, oh) * batch
        gbs = bytes_launch / (k_us * 1e-6) / 1e9
        ws = batch * (4 * w * h + 3 * ow * oh) / 1e6
        return f"| {name} | {batch} | {ws:.0f} | {fps:,.0f} | {k_us:.1f} | {gbs:,.0f} | {gbs / peak:.2f} |"
    finally:
        for s in sess:
            s.close()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--batches", default="1,4,16,32")
    args = ap.parse_args()
    try:
        peak = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
    except Exception:
        peak = 6650.0
    print("| config | sessions per launch | working set MB | frames/s | kernel µs/launch | algorithmic GB/s | frac of measured peak |")
    print("|---|---|---|---|---|---|---|")
    for (name, w, h, ow, oh, fmt) in CONFIGS:
        for batch in [int(x) for x in args.batches.split(",")]:
            if w >= 7680 and batch > 16:
                continue
            print(run(name, w, h, ow, oh, fmt, batch, args.steps, peak), flush=True)


if __name__ == "__main__":
    main()
