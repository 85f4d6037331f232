"""From the per-macroblock maps to what an encoder can use.

The reference disables XDamage in the GStreamer pipeline (`use-damage=0`,
src/selkies/legacy/gstwebrtc_app.py:241) and leaves region/quality policy to its capture
module, to which it only passes three knobs (src/selkies/selkies.py:2949-2951).  Here the
GPU produces, per frame, the damage map and a policy map (SK_POLICY_* bits, updated in the
convert kernel); this module turns them into rectangles and keyframe / paint-over decisions on
the host.  Pure Python/numpy on ~32 K bytes per 4K frame; no pixel work.
"""
from __future__ import annotations

from dataclasses import dataclass

import numpy as np

from . import _abi

MB = 16


@dataclass(frozen=True)
class Region:
    """Pixel rectangle, clipped to the frame; what GstVideoRegionOfInterestMeta carries."""
    x: int
    y: int
    w: int
    h: int
    kind: str = "damage"        # "damage" | "paint-over"


def regions_from_mask(mask: np.ndarray, width: int, height: int, kind: str = "damage",
                      max_regions: int = 64) -> list[Region]:
    """Cover the set macroblocks with rectangles: horizontal runs per macroblock row, runs with
    the same span in consecutive rows merged vertically.  If more than max_regions result, fall
    back to one bounding rectangle per band of macroblock rows (at most max_regions bands)."""
    mask = np.asarray(mask).astype(bool)
    mbh, mbw = mask.shape
    open_runs: dict[tuple[int, int], list[int]] = {}      # (x0, x1) -> [y0, y1)
    done: list[tuple[int, int, int, int]] = []
    for by in range(mbh):
        row = mask[by]
        runs = []
        if row.any():
            d = np.diff(np.concatenate(([0], row.astype(np.int8), [0])))
            runs = list(zip(np.flatnonzero(d == 1).tolist(), np.flatnonzero(d == -1).tolist()))
        nxt = {}
        for (x0, x1) in runs:
            if (x0, x1) in open_runs:
                y0, _ = open_runs.pop((x0, x1))
                nxt[(x0, x1)] = [y0, by + 1]
            else:
                nxt[(x0, x1)] = [by, by + 1]
        for (x0, x1), (y0, y1) in open_runs.items():
            done.append((x0, y0, x1, y1))
        open_runs = nxt
    for (x0, x1), (y0, y1) in open_runs.items():
        done.append((x0, y0, x1, y1))
    if len(done) > max_regions:
        # too fragmented for per-run rectangles: one bounding box per band of macroblock rows
        band = -(-mbh // max_regions)
        done = []
        for y0 in range(0, mbh, band):
            sub = mask[y0:y0 + band]
            if sub.any():
                ys, xs = np.nonzero(sub)
                done.append((int(xs.min()), y0 + int(ys.min()), int(xs.max()) + 1, y0 + int(ys.max()) + 1))
    out = []
    for (x0, y0, x1, y1) in sorted(done, key=lambda r: (r[1], r[0])):
        px, py = x0 * MB, y0 * MB
        out.append(Region(px, py, min(x1 * MB, width) - px, min(y1 * MB, height) - py, kind))
    return out


@dataclass
class FrameDecision:
    damaged_blocks: int
    total_blocks: int
    force_keyframe: bool          # send a downstream force-key-unit before this frame
    skip_frame: bool              # nothing changed and nothing to repaint: the encoder may repeat
    damage_regions: list[Region]
    paint_over_regions: list[Region]
    hot_fraction: float


class DamagePolicy:
    """Per-session decisions from the maps of each frame."""

    def __init__(self, width: int, height: int, keyframe_fraction: float = 0.6, max_regions: int = 64):
        if not 0.0 < keyframe_fraction <= 1.0:
            raise ValueError("keyframe_fraction must be in (0, 1]")
        self.width, self.height = width, height
        self.keyframe_fraction = keyframe_fraction
        self.max_regions = max_regions
        self._first = True

    def reset(self):
        self._first = True

    def decide(self, damage_map: np.ndarray, policy_map: np.ndarray | None, damaged_count: int | None = None) -> FrameDecision:
        dm = np.asarray(damage_map)
        total = dm.size
        n = int(dm.sum()) if damaged_count is None else int(damaged_count)
        paint = hot = None
        if policy_map is not None:
            pm = np.asarray(policy_map)
            paint = (pm & _abi.SK_POLICY_PAINT_OVER) != 0
            hot = (pm & _abi.SK_POLICY_HOT) != 0
        # the very first frame (or the one after a history reset) is all-damaged by definition:
        # the encoder produces a keyframe there anyway
        force = self._first or n >= self.keyframe_fraction * total
        self._first = False
        dr = [] if (n == 0 or n == total) else regions_from_mask(dm, self.width, self.height, "damage", self.max_regions)
        if n == total:
            dr = [Region(0, 0, self.width, self.height, "damage")]
        pr = regions_from_mask(paint, self.width, self.height, "paint-over", self.max_regions) if paint is not None and paint.any() else []
        return FrameDecision(n, total, bool(force), n == 0 and not pr, dr, pr,
                             float(hot.mean()) if hot is not None else 0.0)
