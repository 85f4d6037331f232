"""Frame geometry the stage can be asked for.

Where W x H comes from in the reference: the browser's requested size is clamped to the
display limit and snapped to even numbers before xrandr applies it
(src/selkies/legacy/resize.py:15-29 `fit_res`, limits at :62-69; even sizes are also forced at
src/selkies/legacy/webrtc_input.py:677 and src/selkies/selkies.py:3104-3105).  The stage relies on
that invariant (4:2:0 needs even sizes), so the same rule is available here for hosts that size
the encoder output themselves (capture size != encode size → the 2-tap shrink path).
"""
from __future__ import annotations

MAX_WIDTH, MAX_HEIGHT = 7680, 4320          # resize.py:66
MAX_WIDTH_DVI, MAX_HEIGHT_DVI = 2560, 1600  # resize.py:64


def fit_resolution(width: int, height: int, max_width: int = MAX_WIDTH, max_height: int = MAX_HEIGHT) -> tuple[int, int]:
    """Largest even size, same aspect, inside max_width x max_height (float steps of 0.01 % as the
    reference does, so results agree with it exactly — tests/golden/fit_res.json)."""
    if width < max_width and height < max_height:
        return width, height
    fw, fh = float(width), float(height)
    while fw > max_width or fh > max_height:
        fw, fh = fw * 0.9999, fh * 0.9999
    w, h = int(fw), int(fh)
    return w + (w & 1), h + (h & 1)


def is_valid_frame_size(width: int, height: int) -> bool:
    """What sk_session_create accepts: even, 2..8192."""
    return all(isinstance(v, int) and 2 <= v <= 8192 and v % 2 == 0 for v in (width, height))


def macroblocks(width: int, height: int) -> tuple[int, int]:
    return (width + 15) // 16, (height + 15) // 16
