"""A `pixelflux`-shaped front for the B200 stage: the WebSocket mode's capture contract.

In its default mode the reference never builds the GStreamer pipeline; `selkies.py` drives a capture
module with exactly this surface (src/selkies/selkies.py:2846-2917):

    settings = CaptureSettings(); settings.capture_width = ...          # :2919-2964
    module = ScreenCapture()
    module.start_capture(settings, StripeCallback(cb))                  # :2897-2904
    # cb(result_ptr, user_data) on the module's thread; result = result_ptr.contents;
    # bytes(result.data[:result.size]), result.frame_id                 # :2865-2891
    module.stop_capture()

The module cuts each frame into horizontal stripes and delivers only those that changed (or that went
still long enough to be repainted at higher quality), each prefixed with the header the browser parses
(addons/gst-web-core/selkies-core.js:2812-2938; `stripes.py`).  This shim does that with the GPU stage:
convert + per-macroblock damage + temporal policy in one kernel, damage-gated read-back (only changed
macroblocks cross PCIe), stripes planned from the maps.  Two things the real module also does stay
outside the scope of this project and are therefore pluggable here: grabbing the screen (`frame_source`,
a callable returning a BGRx array) and entropy coding (`stripe_encoder`; the default hands over the
stripe's raw planes, which is what an encoder would be fed).
"""
from __future__ import annotations

import threading
import time

import numpy as np

from . import stripes as _st
from .session import PreencodeSession

OUTPUT_MODE_JPEG, OUTPUT_MODE_H264 = 0, 1


class CaptureSettings:
    """The fields `selkies.py` fills in (:2925-2963); unknown ones may be added freely, as with a ctypes struct."""

    def __init__(self):
        self.capture_width = 1920
        self.capture_height = 1080
        self.capture_x = 0
        self.capture_y = 0
        self.target_fps = 60.0
        self.capture_cursor = False
        self.debug_logging = False
        self.output_mode = OUTPUT_MODE_H264
        self.jpeg_quality = 40
        self.paint_over_jpeg_quality = 90
        self.h264_crf = 25
        self.h264_paintover_crf = 18
        self.h264_paintover_burst_frames = 5
        self.h264_fullcolor = False
        self.h264_streaming_mode = False
        self.h264_fullframe = False
        self.use_paint_over_quality = True
        self.paint_over_trigger_frames = 15      # selkies.py:2949
        self.damage_block_threshold = 10         # selkies.py:2950
        self.damage_block_duration = 20          # selkies.py:2951
        self.use_cpu = False
        self.vaapi_render_node_index = -1
        self.watermark_path = b""
        self.watermark_location_enum = 0
        # additions of this shim
        self.cuda_device_id = -1
        self.n_stripes = 0                       # 0: one stripe per 64 rows, at least 1


class StripeEncodeResult:
    """What the callback dereferences: `.contents` -> (`data`, `size`, `frame_id`) plus this shim's extras."""

    def __init__(self, data: bytes, frame_id: int, stripe: "_st.Stripe | None" = None, paint_over: bool = False):
        self.data = data
        self.size = len(data)
        self.frame_id = frame_id
        self.stripe_y_start = 0 if stripe is None else stripe.y_start
        self.stripe_height = 0 if stripe is None else stripe.height
        self.paint_over = paint_over

    @property
    def contents(self):
        return self


def StripeCallback(fn):
    """pixelflux wraps the Python callable in a CFUNCTYPE; here the callable is used as it is."""
    return fn


def raw_stripe_encoder(views, stripe, settings, paint_over: bool) -> bytes:
    """Stand-in for the entropy coder: the stripe's planes, row-packed (Y then chroma)."""
    return b"".join(np.ascontiguousarray(v).tobytes() for v in views)


class ScreenCapture:
    """start_capture(settings, callback) / stop_capture(), as `selkies.py` calls them."""

    def __init__(self, frame_source=None, stripe_encoder=raw_stripe_encoder):
        self.frame_source = frame_source
        self.stripe_encoder = stripe_encoder
        self._session: PreencodeSession | None = None
        self._settings: CaptureSettings | None = None
        self._callback = None
        self._thread = None
        self._stop = threading.Event()
        self._frame_id = 0
        self._slot = 0
        self.stats = {"frames": 0, "stripes_sent": 0, "stripes_total": 0, "readback_bytes": 0}

    # -- the pixelflux surface ------------------------------------------------------------
    def start_capture(self, settings: CaptureSettings, callback):
        if self._session is not None:
            raise RuntimeError("capture already started")
        w, h = int(settings.capture_width), int(settings.capture_height)
        if w % 2 or h % 2:
            raise ValueError("capture size must be even (selkies.py:3104-3105)")
        fmt = "I420" if settings.output_mode == OUTPUT_MODE_H264 and not settings.h264_fullcolor else "NV12"
        self._settings, self._callback = settings, callback
        self._n_stripes = int(settings.n_stripes) or max(1, h // 64)
        self._session = PreencodeSession(
            w, h, out_format=fmt, device_id=int(getattr(settings, "cuda_device_id", -1)), host_output=True,
            gated_readback=True, policy=True, damage_block_threshold=int(settings.damage_block_threshold),
            damage_block_duration=int(settings.damage_block_duration),
            paint_over_trigger_frames=int(settings.paint_over_trigger_frames))
        self._frame_id = 0
        self._stop.clear()
        if self.frame_source is not None:
            self._thread = threading.Thread(target=self._loop, name="b200-capture", daemon=True)
            self._thread.start()

    def stop_capture(self):
        self._stop.set()
        if self._thread is not None:
            self._thread.join(timeout=5)
            self._thread = None
        if self._session is not None:
            self._session.close()
            self._session = None

    # -- frames ---------------------------------------------------------------------------
    def _loop(self):
        period = 1.0 / max(1.0, float(self._settings.target_fps))
        nxt = time.perf_counter()
        while not self._stop.is_set():
            frame = self.frame_source()
            if frame is not None:
                self.push_frame(frame)
            nxt += period
            delay = nxt - time.perf_counter()
            if delay > 0:
                self._stop.wait(delay)
            else:
                nxt = time.perf_counter()

    def push_frame(self, frame: np.ndarray) -> list[StripeEncodeResult]:
        """Convert one captured BGRx frame and deliver its changed stripes through the callback
        (also returned, for callers that drive the shim synchronously)."""
        s, cs = self._session, self._settings
        if s is None:
            raise RuntimeError("capture not started")
        r = s.wait(s.submit_host(self._slot, np.ascontiguousarray(frame)))
        self._slot = (self._slot + 1) % s.n_inflight
        plan = _st.plan_stripes(r.damage_map, r.policy_map, s.out_height, self._n_stripes)
        if cs.h264_fullframe and any(p.damaged or p.paint_over for p in plan):
            # x264enc full-frame mode (selkies.py:2946): one packet for the whole frame when anything changed
            todo = [_st.Stripe(0, 0, s.out_height, True, any(p.paint_over for p in plan), any(p.hot for p in plan))]
        else:
            todo = _st.stripes_to_send(plan)
        out = []
        fid = self._frame_id & 0xFFFF
        for st in todo:
            paint = bool(st.paint_over and not st.damaged and cs.use_paint_over_quality)
            views = _st.stripe_views(r.frame, s.out_width, s.out_height, s.out_format, r.plane_offset, r.plane_stride,
                                     st.y_start, st.height)
            payload = self.stripe_encoder(views, st, cs, paint)
            key = self._frame_id == 0 or paint
            if cs.output_mode == OUTPUT_MODE_JPEG:
                # the server itself prepends 0x03 0x00 for JPEG (selkies.py:2874-2875): frame id + y start follow
                data = _st.pack_jpeg_stripe_header(fid, st.y_start)[2:] + payload
            elif cs.h264_fullframe:
                data = _st.pack_h264_full_header(fid, key) + payload
            else:
                data = _st.pack_h264_stripe_header(fid, key, st.y_start, s.out_width, st.height) + payload
            res = StripeEncodeResult(data, fid, st, paint)
            out.append(res)
            if self._callback is not None:
                self._callback(res, None)
        self.stats["frames"] += 1
        self.stats["stripes_sent"] += len(out)
        self.stats["stripes_total"] += 1 if cs.h264_fullframe else len(plan)
        self.stats["readback_bytes"] += r.readback_bytes
        self._frame_id += 1
        return out
