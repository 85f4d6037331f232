"""GstBase.BaseTransform adaptor: the drop-in for `videoconvert` / `cudaupload ! cudaconvert`
in the pipeline the reference builds (src/selkies/legacy/gstwebrtc_app.py:611-617, :263-284).

All logic lives in stage.PreEncodeStage; this file only translates between GStreamer
objects and it.  `gi` is imported lazily because it does not exist where the kernels are
developed and benchmarked; the reference itself subclasses a GObject type with do_*
vmethods in Python (PlayoutDelayExtension, gstwebrtc_app.py:1744-1780), which is the
precedent for registering an in-process Python element with Gst.Element.register.

    from selkies_gstreamer_b200 import gst_element
    gst_element.register()                      # after Gst.init()
    conv = Gst.ElementFactory.make("b200preencode")
"""
from __future__ import annotations

import ctypes

from . import stage as _stage

FACTORY_NAME = "b200preencode"
_element_class = None


class _MapError(RuntimeError):
    pass


class _GstMapInfo(ctypes.Structure):      # gstmemory.h: GstMapInfo
    _fields_ = [("memory", ctypes.c_void_p), ("flags", ctypes.c_int), ("data", ctypes.c_void_p),
                ("size", ctypes.c_size_t), ("maxsize", ctypes.c_size_t),
                ("user_data", ctypes.c_void_p * 4), ("_gst_reserved", ctypes.c_void_p * 4)]


_libgst = False


def _raw_gst():
    """libgstreamer through ctypes: PyGObject's Buffer.map() hands back *copies*, which would cost a
    33 MB Python-side memcpy per 4K frame; gst_buffer_map on the boxed pointer does not."""
    global _libgst
    if _libgst is False:
        try:
            lib = ctypes.CDLL("libgstreamer-1.0.so.0")
            lib.gst_buffer_map.argtypes = [ctypes.c_void_p, ctypes.POINTER(_GstMapInfo), ctypes.c_int]
            lib.gst_buffer_map.restype = ctypes.c_int
            lib.gst_buffer_unmap.argtypes = [ctypes.c_void_p, ctypes.POINTER(_GstMapInfo)]
            lib.gst_buffer_unmap.restype = None
            _libgst = lib
        except OSError:
            _libgst = None
    return _libgst


class _mapped:
    """Context manager giving (ptr, size) of a Gst.Buffer. Zero-copy through libgstreamer when it can
    be loaded; otherwise PyGObject's copying map, with the output written back by Buffer.fill()."""

    def __init__(self, Gst, buf, write: bool):
        self.Gst, self.buf, self.write = Gst, buf, write
        self.ptr = self.size = 0
        self._info = self._tmp = self._pymap = None

    def __enter__(self):
        lib = _raw_gst()
        if lib is not None:
            self._info = _GstMapInfo()
            flags = 2 if self.write else 1            # GST_MAP_WRITE / GST_MAP_READ
            if lib.gst_buffer_map(hash(self.buf), ctypes.byref(self._info), flags):
                self.ptr, self.size = self._info.data, self._info.size
                return self
            # e.g. a WRITE map refused because PyGObject holds an extra reference on outbuf inside
            # do_transform (the buffer is then not "writable"): take the copying path below instead
            # of failing the frame.
            self._info = None
        if self.write:
            self.size = self.buf.get_size()
            self._tmp = (ctypes.c_uint8 * self.size)()
        else:
            ok, self._pymap = self.buf.map(self.Gst.MapFlags.READ)
            if not ok:
                raise _MapError("Gst.Buffer.map failed")
            self.size = self._pymap.size
            self._tmp = (ctypes.c_uint8 * self.size).from_buffer_copy(self._pymap.data)
        self.ptr = ctypes.addressof(self._tmp)
        return self

    def __exit__(self, et, ev, tb):
        if self._info is not None:
            _raw_gst().gst_buffer_unmap(hash(self.buf), ctypes.byref(self._info))
        elif self.write:
            if et is None:
                self.buf.fill(0, bytes(self._tmp))
        elif self._pymap is not None:
            self.buf.unmap(self._pymap)
        return False


def _caps_from_gst(caps) -> list[dict]:
    return _stage.parse_caps(caps.to_string())


def _structure_to_dict(caps) -> dict:
    return _stage.parse_caps(caps.to_string())[0]


def element_class():
    """Build (once) and return the B200PreEncode GObject class. Needs PyGObject + GStreamer >= 1.22."""
    global _element_class
    if _element_class is not None:
        return _element_class
    import gi
    gi.require_version("Gst", "1.0")
    gi.require_version("GstBase", "1.0")
    from gi.repository import GObject, Gst, GstBase
    try:
        gi.require_version("GstVideo", "1.0")
        from gi.repository import GstVideo
    except (ValueError, ImportError):       # ROI metas / key-unit events need gst-plugins-base's GstVideo
        GstVideo = None

    class B200PreEncode(GstBase.BaseTransform):
        __gstmetadata__ = ("B200 pre-encode stage", "Filter/Converter/Video/Scaler",
                           "BGRx/RGBx to NV12/I420 (+bilinear shrink) with a 16x16 damage map, fused on sm_100a",
                           "selkies_gstreamer_b200")
        __gsttemplates__ = (
            Gst.PadTemplate.new("sink", Gst.PadDirection.SINK, Gst.PadPresence.ALWAYS,
                                Gst.Caps.from_string(_stage.SINK_TEMPLATE)),
            Gst.PadTemplate.new("src", Gst.PadDirection.SRC, Gst.PadPresence.ALWAYS,
                                Gst.Caps.from_string(_stage.SRC_TEMPLATE)),
        )
        # `qos` is inherited from GstBaseTransform (the reference sets it: gstwebrtc_app.py:613, :277).
        __gproperties__ = {
            "n-threads": (GObject.TYPE_UINT, "n-threads", "accepted for videoconvert compatibility (gstwebrtc_app.py:612)",
                          0, GObject.G_MAXUINT, 1, GObject.ParamFlags.READWRITE),
            "cuda-device-id": (GObject.TYPE_INT, "cuda-device-id", "GPU to run on, -1 = default (gstwebrtc_app.py:264-272)",
                               -1, GObject.G_MAXINT, -1, GObject.ParamFlags.READWRITE),
            "method": (GObject.TYPE_STRING, "method", "videoscale method; only 'bilinear'", "bilinear",
                       GObject.ParamFlags.READWRITE),
            "add-borders": (GObject.TYPE_BOOLEAN, "add-borders", "videoscale add-borders; only False", False,
                            GObject.ParamFlags.READWRITE),
            # stock videoconvert properties, accepted only at the values the arithmetic implements
            "dither": (GObject.TYPE_STRING, "dither", "inert for 8-bit in/out", "bayer", GObject.ParamFlags.READWRITE),
            "chroma-resampler": (GObject.TYPE_STRING, "chroma-resampler", "only 'linear'", "linear",
                                 GObject.ParamFlags.READWRITE),
            "chroma-mode": (GObject.TYPE_STRING, "chroma-mode", "only 'full'", "full", GObject.ParamFlags.READWRITE),
            "matrix-mode": (GObject.TYPE_STRING, "matrix-mode", "only 'full'", "full", GObject.ParamFlags.READWRITE),
            "gamma-mode": (GObject.TYPE_STRING, "gamma-mode", "only 'none'", "none", GObject.ParamFlags.READWRITE),
            "primaries-mode": (GObject.TYPE_STRING, "primaries-mode", "only 'none'", "none",
                               GObject.ParamFlags.READWRITE),
            "damage": (GObject.TYPE_BOOLEAN, "damage", "produce the per-16x16 damage map", True,
                       GObject.ParamFlags.READWRITE),
            "damage-block-threshold": (GObject.TYPE_UINT, "damage-block-threshold",
                                       "consecutive damaged frames that make a block hot (selkies.py:2950)",
                                       1, 255, 10, GObject.ParamFlags.READWRITE),
            "damage-block-duration": (GObject.TYPE_UINT, "damage-block-duration",
                                      "frames a hot block stays hot after its last change (selkies.py:2951)",
                                      1, 255, 20, GObject.ParamFlags.READWRITE),
            "paint-over-trigger-frames": (GObject.TYPE_UINT, "paint-over-trigger-frames",
                                          "still frames before a changed block is repainted (selkies.py:2949)",
                                          1, 255, 15, GObject.ParamFlags.READWRITE),
            "keyframe-damage-fraction": (GObject.TYPE_DOUBLE, "keyframe-damage-fraction",
                                         "damaged fraction of the frame that requests a key unit", 0.0, 1.0, 0.6,
                                         GObject.ParamFlags.READWRITE),
            "min-keyframe-distance": (GObject.TYPE_UINT, "min-keyframe-distance",
                                      "fewest frames between two key-unit requests (cf. the reference's "
                                      "min_keyframe_frame_distance, gstwebrtc_app.py:1227-1262)",
                                      0, GObject.G_MAXUINT, 60, GObject.ParamFlags.READWRITE),
            "pin-buffers": (GObject.TYPE_BOOLEAN, "pin-buffers", "page-lock the pool buffers the element is handed, so "
                            "both PCIe copies are DMA (turn off if upstream/downstream do not use buffer pools)", True,
                            GObject.ParamFlags.READWRITE),
            "coalesce": (GObject.TYPE_BOOLEAN, "coalesce", "share kernel launches with the other b200preencode "
                         "elements of this process that have the same caps and GPU", False, GObject.ParamFlags.READWRITE),
            "roi-meta": (GObject.TYPE_BOOLEAN, "roi-meta", "attach GstVideoRegionOfInterestMeta for damaged / "
                         "paint-over regions (x264enc honours them)", True, GObject.ParamFlags.READWRITE),
            "force-keyframes": (GObject.TYPE_BOOLEAN, "force-keyframes", "send downstream force-key-unit when the "
                                "damaged fraction RISES above keyframe-damage-fraction (never more often than "
                                "min-keyframe-distance; steady damage is left to the encoder's rate control)", True,
                                GObject.ParamFlags.READWRITE),
            "paint-over-delta-qp": (GObject.TYPE_INT, "paint-over-delta-qp", "delta-qp put on paint-over regions",
                                    -51, 51, -8, GObject.ParamFlags.READWRITE),
            "damaged-blocks": (GObject.TYPE_UINT, "damaged-blocks", "damaged macroblocks in the last frame",
                               0, GObject.G_MAXUINT, 0, GObject.ParamFlags.READABLE),
            "frames": (GObject.TYPE_UINT64, "frames", "frames converted since the last caps change",
                       0, GObject.G_MAXUINT64, 0, GObject.ParamFlags.READABLE),
        }

        def __init__(self):
            super().__init__()
            self.core = _stage.PreEncodeStage(self.get_name() or FACTORY_NAME)
            self.set_qos_enabled(True)      # videoconvert's default; late frames are dropped before transform
            self.set_passthrough(False)
            self.set_in_place(False)
            self._signalling = {"roi-meta": True, "force-keyframes": True, "paint-over-delta-qp": -8}
            self._keyunit_count = 0

        def do_get_property(self, prop):
            if prop.name in self._signalling:
                return self._signalling[prop.name]
            return self.core.get_property(prop.name)

        def do_set_property(self, prop, value):
            if prop.name in self._signalling:
                self._signalling[prop.name] = value
            else:
                self.core.set_property(prop.name, value)

        def apply_decision(self, outbuf, decision):
            """Turn one frame's decision into what encoders understand: ROI metas on the buffer
            and, when most of the frame changed, a downstream force-key-unit event."""
            if decision is None or GstVideo is None:
                return 0
            n = 0
            if self._signalling["roi-meta"]:
                for r in decision.damage_regions:
                    GstVideo.buffer_add_video_region_of_interest_meta(outbuf, "damage", r.x, r.y, r.w, r.h)
                    n += 1
                for r in decision.paint_over_regions:
                    meta = GstVideo.buffer_add_video_region_of_interest_meta(outbuf, "paint-over", r.x, r.y, r.w, r.h)
                    if meta is not None and hasattr(meta, "add_param"):
                        st = Gst.Structure.new_empty("roi/x264enc")
                        st.set_value("delta-qp", float(self._signalling["paint-over-delta-qp"]))
                        meta.add_param(st)
                    n += 1
            if self._signalling["force-keyframes"] and decision.force_keyframe and self.core.frames > 1:
                ev = GstVideo.video_event_new_downstream_force_key_unit(
                    outbuf.pts, Gst.CLOCK_TIME_NONE, Gst.CLOCK_TIME_NONE, True, self._keyunit_count)
                self._keyunit_count += 1
                self.srcpad.push_event(ev)
            return n

        def do_start(self):
            return self.core.start()

        def do_stop(self):
            return self.core.stop()

        def do_transform_caps(self, direction, caps, filter_caps):
            d = _stage.SINK if direction == Gst.PadDirection.SINK else _stage.SRC
            filt = _caps_from_gst(filter_caps) if filter_caps is not None else None
            out = self.core.transform_caps(d, _caps_from_gst(caps), filt)
            return Gst.Caps.from_string(_stage.caps_to_string(out)) if out else Gst.Caps.new_empty()

        def do_fixate_caps(self, direction, caps, othercaps):
            d = _stage.SINK if direction == Gst.PadDirection.SINK else _stage.SRC
            st = self.core.fixate_caps(d, _caps_from_gst(caps), _caps_from_gst(othercaps))
            return Gst.Caps.from_string(_stage.caps_to_string([st])).fixate()

        def do_set_caps(self, incaps, outcaps):
            ok = self.core.set_caps(_structure_to_dict(incaps), _structure_to_dict(outcaps))
            if not ok:
                Gst.error(f"{FACTORY_NAME}: {self.core.last_error}")
            return ok

        def do_get_unit_size(self, caps):
            try:
                return True, self.core.get_unit_size(_structure_to_dict(caps))
            except (KeyError, _stage.StageError):
                return False, 0

        def do_transform(self, inbuf, outbuf):
            try:
                with _mapped(Gst, inbuf, write=False) as src, _mapped(Gst, outbuf, write=True) as dst:
                    rc = self.core.transform(src.ptr, src.size, dst.ptr, dst.size)
            except _MapError as e:
                self.core.last_error = str(e)
                rc = _stage.FLOW_ERROR
            if rc == _stage.FLOW_OK:
                self.apply_decision(outbuf, self.core.last_decision)
                return Gst.FlowReturn.OK
            # Surfaces as a bus ERROR message, which the reference's bus_call handles (gstwebrtc_app.py:1654-1674).
            Gst.error(f"{FACTORY_NAME}: {self.core.last_error}")
            return Gst.FlowReturn.NOT_NEGOTIATED if rc == _stage.FLOW_NOT_NEGOTIATED else Gst.FlowReturn.ERROR

    _element_class = B200PreEncode
    return _element_class


def register(rank: int | None = None) -> bool:
    """Register the factory in-process (no plugin file needed; check_plugins() at
    gstwebrtc_app.py:1123-1172 therefore needs no new entry)."""
    import gi
    gi.require_version("Gst", "1.0")
    from gi.repository import GObject, Gst
    cls = element_class()
    gtype = GObject.type_register(cls) if hasattr(GObject, "type_register") else cls.__gtype__
    return Gst.Element.register(None, FACTORY_NAME, Gst.Rank.NONE if rank is None else rank, gtype)
