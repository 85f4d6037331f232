"""Synthetic capture frames shared by tests/ and bench.py.

No capture device or videotestsrc exists here; these stand in for what
ximagesrc delivers (BGRx, src/selkies/legacy/gstwebrtc_app.py:210-255).
Pure numpy, host side only.
"""
from __future__ import annotations

import numpy as np

KAT_COLOURS = [(0, 0, 0), (255, 255, 255), (255, 0, 0), (0, 255, 0), (0, 0, 255), (255, 255, 0),
               (0, 255, 255), (255, 0, 255)]


def noise(width: int, height: int, seed: int = 1234) -> np.ndarray:
    """Uniform bytes in all four channels: worst case for damage, X byte random."""
    return np.random.default_rng(seed).integers(0, 256, (height, width, 4), dtype=np.uint8)


def desktop(width: int, height: int, seed: int = 42) -> np.ndarray:
    """Flat background, solid windows, 1-px 'text' runs; X byte 0xFF."""
    rng = np.random.default_rng(seed)
    f = np.empty((height, width, 4), dtype=np.uint8)
    f[...] = (0x36, 0x34, 0x2E, 0xFF)     # B, G, R, X of (0x2E,0x34,0x36)
    for _ in range(40):
        x0, y0 = int(rng.integers(0, width)), int(rng.integers(0, height))
        w, h = int(rng.integers(8, max(9, width // 3))), int(rng.integers(8, max(9, height // 3)))
        f[y0:y0 + h, x0:x0 + w, :3] = rng.integers(0, 256, 3, dtype=np.uint8)
    for _ in range(200):
        x0, y0 = int(rng.integers(0, width)), int(rng.integers(0, height))
        w, h = int(rng.integers(16, 160)), int(rng.integers(6, 14))
        blk = f[y0:y0 + h, x0:x0 + w, :3]
        pat = (rng.integers(0, 2, blk.shape[:2], dtype=np.uint8) * 255)[..., None]
        blk[...] = pat
    return f


def bars(width: int, height: int) -> np.ndarray:
    """Eight colour bars (the KAT colours) + a 1-px checkerboard strip."""
    f = np.empty((height, width, 4), dtype=np.uint8)
    f[..., 3] = 0xFF
    for i, (r, g, b) in enumerate(KAT_COLOURS):
        x0, x1 = width * i // 8, width * (i + 1) // 8
        f[:, x0:x1, 0], f[:, x0:x1, 1], f[:, x0:x1, 2] = b, g, r
    strip = max(2, height // 8)
    yy, xx = np.mgrid[0:strip, 0:width]
    f[height - strip:, :, :3] = (((yy + xx) & 1) * 255).astype(np.uint8)[..., None]
    return f


def mutate(frame: np.ndarray, fraction: float, seed: int = 7, aligned: bool = True) -> np.ndarray:
    """Next frame: re-randomise rectangles until about `fraction` of the 16x16 blocks changed."""
    rng = np.random.default_rng(seed)
    out = frame.copy()
    h, w = frame.shape[:2]
    mbw, mbh = (w + 15) // 16, (h + 15) // 16
    target = int(round(fraction * mbw * mbh))
    if fraction >= 1.0:
        out[..., :3] = rng.integers(0, 256, (h, w, 3), dtype=np.uint8)
        return out
    done = set()
    while len(done) < target:
        bx, by = int(rng.integers(0, mbw)), int(rng.integers(0, mbh))
        if (bx, by) in done:
            continue
        done.add((bx, by))
        if aligned:
            x0, y0, ww, hh = bx * 16, by * 16, 16, 16
        else:
            x0, y0 = bx * 16 + int(rng.integers(0, 12)), by * 16 + int(rng.integers(0, 12))
            ww = hh = 4
        blk = out[y0:y0 + hh, x0:x0 + ww, :3]
        blk[...] = rng.integers(0, 256, blk.shape, dtype=np.uint8)
    return out


def algorithmic_bytes(in_w: int, in_h: int, out_w: int, out_h: int, damage: bool = True) -> int:
    """bytes one frame must move, independent of caching."""
    b = 4 * in_w * in_h + (3 * out_w * out_h) // 2
    if damage:
        b += (3 * out_w * out_h) // 2 + ((out_w + 15) // 16) * ((out_h + 15) // 16)
    return b
