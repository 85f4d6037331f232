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
    done: list[tuple[int, int, int, int]] = []
    # all horizontal runs of the map at once: +1 where a run starts, -1 one past its end
    edge = np.diff(np.pad(mask.astype(np.int8), ((0, 0), (1, 1))), axis=1)
    rows, x0s = np.nonzero(edge == 1)
    _, x1s = np.nonzero(edge == -1)                 # same row-major order: run k is (rows[k], x0s[k], x1s[k])
    fragmented = rows.size > 8 * max_regions        # vertical merging cannot bring that under max_regions cheaply
    if not fragmented:
        open_runs: dict[tuple[int, int], list[int]] = {}      # (x0, x1) -> [y0, y1)
        k, nrun = 0, rows.size
        rows_l, x0_l, x1_l = rows.tolist(), x0s.tolist(), x1s.tolist()
        last_row = -1
        while k < nrun:
            by = rows_l[k]
            if by != last_row + 1:                  # a gap of empty rows closes everything that was open
                done.extend((x0, y0, x1, y1) for (x0, x1), (y0, y1) in open_runs.items())
                open_runs = {}
            nxt = {}
            while k < nrun and rows_l[k] == by:
                key = (x0_l[k], x1_l[k])
                if key in open_runs:
                    nxt[key] = [open_runs.pop(key)[0], by + 1]
                else:
                    nxt[key] = [by, by + 1]
                k += 1
            done.extend((x0, y0, x1, y1) for (x0, x1), (y0, y1) in open_runs.items())
            open_runs = nxt
            last_row = by
        done.extend((x0, y0, x1, y1) for (x0, x1), (y0, y1) in open_runs.items())
    if fragmented or len(done) > max_regions:
        # too fragmented for per-run rectangles: one bounding box per band of macroblock rows
        band = -(-mbh // max_regions)
        done = []
        any_row = mask.any(axis=1)
        any_col_lo = np.where(mask.any(axis=1), mask.argmax(axis=1), mbw)
        any_col_hi = np.where(mask.any(axis=1), mbw - mask[:, ::-1].argmax(axis=1), 0)
        for y0 in range(0, mbh, band):
            sub = any_row[y0:y0 + band]
            if sub.any():
                ys = np.flatnonzero(sub)
                done.append((int(any_col_lo[y0:y0 + band].min()), y0 + int(ys[0]),
                             int(any_col_hi[y0:y0 + band].max()), y0 + int(ys[-1]) + 1))
    out = []
    for (x0, y0, x1, y1) in sorted(done, key=lambda r: (r[1], r[0])):
        px, py = x0 * MB, y0 * MB
        out.append(Region(px, py, min(x1 * MB, width) - px, min(y1 * MB, height) - py, kind))
    return out


class FrameDecision:
    """What the maps of one frame mean for the encoder.  The rectangles are worked out when first
    asked for: a host that only wants the keyframe / skip decision does not pay for them."""

    def __init__(self, damaged_blocks: int, total_blocks: int, force_keyframe: bool, skip_frame: bool,
                 damage_regions: list[Region] | None = None, paint_over_regions: list[Region] | None = None,
                 hot_fraction: float = 0.0, _lazy=None):
        self.damaged_blocks = damaged_blocks
        self.total_blocks = total_blocks
        self.force_keyframe = force_keyframe      # send a downstream force-key-unit before this frame
        self.skip_frame = skip_frame              # nothing changed and nothing to repaint: the encoder may repeat
        self.hot_fraction = hot_fraction
        self._damage_regions = damage_regions
        self._paint_over_regions = paint_over_regions
        self._lazy = _lazy                        # (damage map, paint mask or None, width, height, max_regions)

    @property
    def damage_regions(self) -> list[Region]:
        if self._damage_regions is None:
            dm, _, w, h, mr = self._lazy
            n, total = self.damaged_blocks, self.total_blocks
            if n == 0:
                self._damage_regions = []
            elif n == total:
                self._damage_regions = [Region(0, 0, w, h, "damage")]
            else:
                self._damage_regions = regions_from_mask(dm, w, h, "damage", mr)
        return self._damage_regions

    @property
    def paint_over_regions(self) -> list[Region]:
        if self._paint_over_regions is None:
            _, paint, w, h, mr = self._lazy
            self._paint_over_regions = [] if paint is None else regions_from_mask(paint, w, h, "paint-over", mr)
        return self._paint_over_regions

    def __repr__(self):
        return (f"FrameDecision(damaged_blocks={self.damaged_blocks}, total_blocks={self.total_blocks}, "
                f"force_keyframe={self.force_keyframe}, skip_frame={self.skip_frame}, hot_fraction={self.hot_fraction:.4f})")


class DamagePolicy:
    """Per-session decisions from the maps of each frame.

    Key units are requested on the RISING EDGE of the damaged fraction only (the previous frame was
    below `keyframe_fraction`, this one is not) and never closer together than
    `min_keyframe_distance` frames.  Steady high-damage content — video, games, scrolling — is left
    to the encoder's rate control: the reference runs an infinite GOP under a tight VBV
    (`keyframe_distance=-1` -> gop-size -1 / key-int-max 1024, vbv = 1.5x, and its own
    `min_keyframe_frame_distance`, src/selkies/legacy/gstwebrtc_app.py:1227-1262), which an IDR per
    frame would overrun."""

    def __init__(self, width: int, height: int, keyframe_fraction: float = 0.6, min_keyframe_distance: int = 60,
                 max_regions: int = 64):
        if not 0.0 < keyframe_fraction <= 1.0:
            raise ValueError("keyframe_fraction must be in (0, 1]")
        if min_keyframe_distance < 0:
            raise ValueError("min_keyframe_distance must be >= 0")
        self.width, self.height = width, height
        self.keyframe_fraction = keyframe_fraction
        self.min_keyframe_distance = int(min_keyframe_distance)
        self.max_regions = max_regions
        self.reset()

    def reset(self):
        self._first = True
        self._was_high = False
        self._since_key = 0

    def decide(self, damage_map: np.ndarray, policy_map: np.ndarray | None, damaged_count: int | None = None) -> FrameDecision:
        dm = np.asarray(damage_map)
        total = dm.size
        n = int(dm.sum()) if damaged_count is None else int(damaged_count)
        paint = hot = None
        if policy_map is not None:
            pm = np.asarray(policy_map)
            paint = (pm & _abi.SK_POLICY_PAINT_OVER) != 0
            hot = (pm & _abi.SK_POLICY_HOT) != 0
        high = n >= self.keyframe_fraction * total
        # The very first frame (or the one after a history reset) is all-damaged by definition: the
        # encoder produces a keyframe there anyway.  After that: rising edge, and not too soon.
        force = self._first or (high and not self._was_high and self._since_key >= self.min_keyframe_distance)
        self._since_key = 0 if force else self._since_key + 1
        self._was_high = high
        self._first = False
        has_paint = paint is not None and bool(paint.any())
        return FrameDecision(n, total, bool(force), n == 0 and not has_paint, None, None if has_paint else [],
                             float(hot.mean()) if hot is not None else 0.0,
                             _lazy=(dm, paint if has_paint else None, self.width, self.height, self.max_regions))
