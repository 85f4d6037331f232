#!/usr/bin/env python3
"""Convert raw BGRx frames with the B200 stage and write what an encoder would receive.

    python tools/convert_raw.py --size 1920x1080 [--in frames.bgrx] [--out-size 1280x720] [--format NV12]
                                [--frames 3] --out out_prefix

Without --in, synthetic desktop frames are used.  Writes <prefix>.<format>.yuv (packed frames, GstVideoInfo
layout, playable with e.g. `ffplay -f rawvideo -pixel_format nv12 -video_size WxH`) and one
<prefix>.damage.<n>.pgm per frame (1 pixel per 16x16 macroblock, 255 = damaged).  Needs a CUDA device.
"""
import argparse
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from selkies_gstreamer_b200 import PreencodeSession, synth  # noqa: E402
from selkies_gstreamer_b200.damage_policy import DamagePolicy  # noqa: E402


def parse_size(s):
    w, h = s.lower().split("x")
    return int(w), int(h)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--size", required=True)
    ap.add_argument("--out-size")
    ap.add_argument("--in", dest="inp")
    ap.add_argument("--format", default="NV12", choices=["NV12", "I420"])
    ap.add_argument("--colorimetry")
    ap.add_argument("--frames", type=int, default=3)
    ap.add_argument("--out", required=True)
    a = ap.parse_args()
    w, h = parse_size(a.size)
    ow, oh = parse_size(a.out_size) if a.out_size else (w, h)
    if a.inp:
        raw = np.fromfile(a.inp, dtype=np.uint8)
        n = raw.size // (w * h * 4)
        frames = [raw[i * w * h * 4:(i + 1) * w * h * 4].reshape(h, w, 4) for i in range(min(n, a.frames))]
    else:
        frames = [synth.desktop(w, h)]
        for i in range(1, a.frames):
            frames.append(synth.mutate(frames[-1], 0.02, seed=i, aligned=False))
    pol = DamagePolicy(ow, oh)
    with PreencodeSession(w, h, out_width=ow, out_height=oh, out_format=a.format, colorimetry=a.colorimetry,
                          policy=True) as s, open(f"{a.out}.{a.format.lower()}.yuv", "wb") as fo:
        for i, f in enumerate(frames):
            r = s.convert(f, slot=i & 1)
            fo.write(r.frame.tobytes())
            d = pol.decide(r.damage_map, r.policy_map, r.damaged_count)
            with open(f"{a.out}.damage.{i}.pgm", "wb") as pg:
                pg.write(f"P5 {r.damage_map.shape[1]} {r.damage_map.shape[0]} 255\n".encode())
                pg.write((r.damage_map * 255).astype(np.uint8).tobytes())
            print(f"frame {i}: {r.damaged_count}/{r.damage_map.size} blocks damaged, {len(d.damage_regions)} regions, "
                  f"keyframe={d.force_keyframe}")


if __name__ == "__main__":
    main()
