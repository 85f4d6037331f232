"""Damaged stripes and the WebSocket-mode wire headers.

In its default (WebSocket) mode the reference does not run the GStreamer stage: its capture
module cuts the frame into horizontal stripes, encodes only the changed ones and hands them to
`selkies.py` (src/selkies/selkies.py:2865-2891), which relays them to the browser.  The browser
parses three packet kinds (addons/gst-web-core/selkies-core.js):

    0x00  full-frame H.264   4-byte header  :2812-2818   type, frame-type, u16 frame id
    0x03  JPEG stripe        6-byte header  :2908-2914   type, pad, u16 frame id, u16 y start
    0x04  H.264 stripe      10-byte header  :2925-2938   type, frame-type, u16 frame id, u16 y start,
                                                         u16 width, u16 height        (all big-endian)

This module lets the B200 stage sit under that mode: it turns the per-macroblock maps into the
list of stripes an encoder has to touch, gives zero-copy views of those stripes in the packed
NV12/I420 frame, and packs/parses the headers.  The encoders themselves stay out of scope.
"""
from __future__ import annotations

import struct
from dataclasses import dataclass

import numpy as np

from . import _abi

TYPE_H264_FULL, TYPE_JPEG_STRIPE, TYPE_H264_STRIPE = 0x00, 0x03, 0x04
FRAME_DELTA, FRAME_KEY = 0, 1


@dataclass(frozen=True)
class Stripe:
    index: int
    y_start: int
    height: int
    damaged: bool
    paint_over: bool      # went still long enough: resend at paint-over quality
    hot: bool             # keeps changing: streaming quality, no paint-over


def stripe_layout(height: int, n_stripes: int) -> list[tuple[int, int]]:
    """(y_start, height) of n horizontal stripes; boundaries on multiples of 16 rows so that a
    stripe owns whole macroblock rows (the last one takes the remainder)."""
    if n_stripes < 1:
        raise ValueError("n_stripes must be >= 1")
    mb_rows = (height + 15) // 16
    n = min(n_stripes, mb_rows)
    base, extra = divmod(mb_rows, n)
    out, row = [], 0
    for i in range(n):
        rows = base + (1 if i < extra else 0)
        y0 = row * 16
        out.append((y0, min(rows * 16, height - y0)))
        row += rows
    return out


def plan_stripes(damage_map: np.ndarray, policy_map: np.ndarray | None, height: int, n_stripes: int) -> list[Stripe]:
    """One Stripe per band, flagged from the macroblock rows it owns."""
    dm = np.asarray(damage_map)
    pm = None if policy_map is None else np.asarray(policy_map)
    out = []
    for i, (y0, h) in enumerate(stripe_layout(height, n_stripes)):
        r0, r1 = y0 // 16, (y0 + h + 15) // 16
        damaged = bool(dm[r0:r1].any())
        paint = bool(pm is not None and (pm[r0:r1] & _abi.SK_POLICY_PAINT_OVER).any())
        hot = bool(pm is not None and (pm[r0:r1] & _abi.SK_POLICY_HOT).any())
        out.append(Stripe(i, y0, h, damaged, paint, hot))
    return out


def stripes_to_send(stripes: list[Stripe]) -> list[Stripe]:
    return [s for s in stripes if s.damaged or s.paint_over]


def stripe_views(packed: np.ndarray, width: int, height: int, out_format: str, plane_offset, plane_stride,
                 y_start: int, stripe_height: int):
    """Zero-copy (Y, chroma...) row views of one stripe of a packed NV12/I420 frame.
    y_start and stripe_height must be even (they are: stripes start on multiples of 16)."""
    if y_start % 2 or (stripe_height % 2 and y_start + stripe_height != height):
        raise ValueError("stripes must start and end on even rows")
    h2 = (height + 1) & ~1
    y = packed[plane_offset[0]:plane_offset[0] + plane_stride[0] * h2].reshape(h2, plane_stride[0])
    ys = y[y_start:y_start + stripe_height, :width]
    c0, c1 = y_start // 2, (y_start + stripe_height + 1) // 2
    if out_format.upper() == "NV12":
        uv = packed[plane_offset[1]:plane_offset[1] + plane_stride[1] * (h2 // 2)].reshape(h2 // 2, plane_stride[1])
        return ys, uv[c0:c1, :2 * ((width + 1) // 2)]
    u = packed[plane_offset[1]:plane_offset[1] + plane_stride[1] * (h2 // 2)].reshape(h2 // 2, plane_stride[1])
    v = packed[plane_offset[2]:plane_offset[2] + plane_stride[2] * (h2 // 2)].reshape(h2 // 2, plane_stride[2])
    return ys, u[c0:c1, :(width + 1) // 2], v[c0:c1, :(width + 1) // 2]


# ---- headers (what the browser parses) ---------------------------------------------------
def pack_h264_full_header(frame_id: int, keyframe: bool) -> bytes:
    return struct.pack(">BBH", TYPE_H264_FULL, FRAME_KEY if keyframe else FRAME_DELTA, frame_id & 0xFFFF)


def pack_jpeg_stripe_header(frame_id: int, y_start: int) -> bytes:
    return struct.pack(">BBHH", TYPE_JPEG_STRIPE, 0, frame_id & 0xFFFF, y_start)


def pack_h264_stripe_header(frame_id: int, keyframe: bool, y_start: int, width: int, height: int) -> bytes:
    return struct.pack(">BBHHHH", TYPE_H264_STRIPE, FRAME_KEY if keyframe else FRAME_DELTA, frame_id & 0xFFFF,
                       y_start, width, height)


def parse_packet(packet: bytes) -> dict:
    """The browser's view of a packet (non-shared mode), for round-trip checks."""
    t = packet[0]
    if t == TYPE_H264_FULL:
        if len(packet) < 4:
            raise ValueError("short packet")
        _, ft, fid = struct.unpack_from(">BBH", packet)
        return {"type": t, "keyframe": ft == 1, "frame_id": fid, "payload": packet[4:]}
    if t == TYPE_JPEG_STRIPE:
        if len(packet) < 6:
            raise ValueError("short packet")
        _, _, fid, y0 = struct.unpack_from(">BBHH", packet)
        return {"type": t, "frame_id": fid, "y_start": y0, "payload": packet[6:]}
    if t == TYPE_H264_STRIPE:
        if len(packet) < 10:
            raise ValueError("short packet")
        _, ft, fid, y0, w, h = struct.unpack_from(">BBHHHH", packet)
        return {"type": t, "keyframe": ft == 1, "frame_id": fid, "y_start": y0, "width": w, "height": h,
                "payload": packet[10:]}
    raise ValueError(f"unknown packet type 0x{t:02x}")
