"""Regenerates tests/golden/frozen_small.npz from the C oracle.

These are FROZEN outputs of this project's own oracle (regression fixtures and a
travel-proof copy for the GPU box), not outputs of GStreamer: the reference's
arithmetic cannot be run here (parity unpinned — see oracle/sk_oracle.h).
The only externally derived vectors are in kat_colours.json.

    python tests/golden/make_frozen.py
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import cpu_oracle as o  # noqa: E402

# (in_w, in_h, out_w, out_h, in_fmt, out_fmt, matrix, site, scale_variant, seed)
CASES = [
    (34, 18, 34, 18, 0, 0, 1, 1, 0, 1),
    (34, 18, 34, 18, 1, 1, 0, 0, 0, 2),
    (130, 34, 130, 34, 0, 1, 2, 1, 0, 3),
    (2, 2, 2, 2, 0, 0, 0, 1, 0, 4),
    (256, 40, 256, 40, 0, 0, 1, 0, 0, 5),
    (64, 48, 48, 36, 0, 1, 1, 1, 0, 6),
    (64, 48, 48, 36, 0, 0, 1, 1, 1, 7),
    (100, 60, 98, 34, 1, 0, 0, 0, 0, 8),
    (64, 48, 64, 24, 0, 0, 1, 1, 0, 9),
    (64, 48, 30, 48, 0, 1, 1, 1, 0, 10),
    # growing pictures: scaled behind the matrix
    (32, 24, 64, 48, 0, 1, 1, 1, 0, 11),
    (30, 20, 32, 48, 1, 0, 0, 0, 1, 12),
    (64, 24, 20, 96, 0, 0, 2, 1, 0, 13),
]


def main():
    out = {"cases": np.array(CASES, dtype=np.int32)}
    for i, (w, h, ow, oh, inf, outf, m, site, var, seed) in enumerate(CASES):
        frame = np.random.default_rng(seed).integers(0, 256, (h, w, 4), dtype=np.uint8)
        nxt = frame.copy()
        nxt[h // 3:h // 3 + 5, w // 2:w // 2 + 3, :3] ^= 0x55
        a = o.convert(frame, ow, oh, in_fmt=inf, out_fmt=outf, matrix=m, chroma_site=site, scale_variant=var)
        b = o.convert(nxt, ow, oh, in_fmt=inf, out_fmt=outf, matrix=m, chroma_site=site, scale_variant=var)
        dm, _ = o.damage(ow, oh, outf, b, a)
        out[f"in0_{i}"], out[f"in1_{i}"] = frame, nxt
        out[f"out0_{i}"], out[f"out1_{i}"], out[f"map1_{i}"] = a, b, dm
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "frozen_small.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
