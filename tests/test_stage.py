"""Element surface without GStreamer: properties, caps negotiation, unit sizes, errors
(mirrors how src/selkies/legacy/gstwebrtc_app.py:611-617 / :263-284 drives the stock stage)."""
import sys
import types
from fractions import Fraction

import numpy as np
import pytest

from selkies_gstreamer_b200 import stage as st
from selkies_gstreamer_b200.stage import IntRange, PreEncodeStage, StageError


def test_reference_property_calls_are_accepted():
    s = PreEncodeStage()
    s.set_property("n-threads", 3)              # gstwebrtc_app.py:612
    s.set_property("qos", True)                 # :613
    s.set_property("cuda-device-id", 2)         # :264-272
    assert s.get_property("n-threads") == 3 and s.get_property("qos") is True and s.get_property("cuda-device-id") == 2
    assert s.get_property("method") == "bilinear"
    s.set_property("method", 1)
    assert s.get_property("method") == "bilinear"


def test_property_errors_like_gobject():
    s = PreEncodeStage()
    with pytest.raises(StageError, match="no property named"):
        s.set_property("brightness", 0)
    with pytest.raises(StageError, match="no property named"):
        s.get_property("nope")
    with pytest.raises(StageError, match="not writable"):
        s.set_property("frames", 3)
    with pytest.raises(StageError, match="out of range"):
        s.set_property("cuda-device-id", -2)
    with pytest.raises(StageError, match="out of range"):
        s.set_property("method", "lanczos")
    with pytest.raises(StageError, match="out of range"):
        s.set_property("add-borders", True)
    with pytest.raises(StageError, match="integer"):
        s.set_property("n-threads", "4")


def test_stock_videoconvert_properties_only_at_implemented_values():
    s = PreEncodeStage()
    for name, value in (("dither", "none"), ("chroma-resampler", "linear"), ("chroma-mode", "full"),
                        ("matrix-mode", "full"), ("gamma-mode", "none"), ("primaries-mode", "none"),
                        ("alpha-mode", "set"), ("alpha-value", 0.5), ("sharpness", 1.0), ("gamma-decode", False)):
        s.set_property(name, value)
        assert s.get_property(name) == value
    for name, value in (("chroma-resampler", "nearest"), ("chroma-mode", "none"), ("matrix-mode", "none"),
                        ("gamma-mode", "remap"), ("primaries-mode", "fast"), ("sharpness", 1.5), ("gamma-decode", True)):
        with pytest.raises(StageError, match="out of range"):
            s.set_property(name, value)
    with pytest.raises(StageError, match="string"):
        s.set_property("chroma-mode", 3)


def test_caps_parser_roundtrip():
    txt = ("video/x-raw, format=(string)BGRx, width=(int)1920, height=(int)1080, framerate=(fraction)60/1, "
           "pixel-aspect-ratio=(fraction)1/1")
    c = st.parse_caps(txt)
    assert c == [{"name": "video/x-raw", "format": "BGRx", "width": 1920, "height": 1080,
                  "framerate": Fraction(60, 1), "pixel-aspect-ratio": Fraction(1, 1)}]
    assert st.parse_caps(st.caps_to_string(c)) == c
    t = st.parse_caps(st.SINK_TEMPLATE)[0]
    assert t["format"] == ["BGRx", "BGRA", "RGBx", "RGBA"] and t["width"] == IntRange(2, 8192)
    cuda = st.parse_caps("video/x-raw(memory:CUDAMemory), format=(string)NV12")[0]
    assert cuda["features"] == "memory:CUDAMemory"


def test_transform_caps_sink_to_src():
    s = PreEncodeStage()
    sink = st.parse_caps("video/x-raw, format=(string)BGRx, width=(int)2560, height=(int)1440, framerate=(fraction)60/1")
    out = s.transform_caps(st.SINK, sink)
    assert len(out) == 1
    o = out[0]
    assert o["format"] == ["NV12", "I420"]
    assert o["width"] == IntRange(2, 8192) and o["height"] == IntRange(2, 8192)     # scales both ways
    assert o["framerate"] == Fraction(60, 1)
    assert "colorimetry" not in o
    # the downstream capsfilter (gstwebrtc_app.py:614-617) forces the format
    filt = st.parse_caps("video/x-raw, format=(string)NV12")
    out = s.transform_caps(st.SINK, sink, filt)
    assert out[0]["format"] == "NV12"


def test_transform_caps_src_to_sink_and_rejects_other_memory():
    s = PreEncodeStage()
    src = st.parse_caps("video/x-raw, format=(string)I420, width=(int)1920, height=(int)1080")
    out = s.transform_caps(st.SRC, src)
    assert out[0]["format"] == list(st.SINK_FORMATS)
    assert out[0]["width"] == IntRange(2, 8192)
    assert s.transform_caps(st.SINK, st.parse_caps("video/x-raw(memory:GLMemory), format=(string)BGRx")) == []
    assert s.transform_caps(st.SINK, st.parse_caps("audio/x-raw, format=(string)S16LE")) == []


def test_fixate_prefers_input_size_and_first_format():
    s = PreEncodeStage()
    sink = st.parse_caps("video/x-raw, format=(string)BGRx, width=(int)1920, height=(int)1080, framerate=(fraction)30/1")
    other = s.transform_caps(st.SINK, sink)
    fx = s.fixate_caps(st.SINK, sink, other)
    assert (fx["format"], fx["width"], fx["height"], fx["framerate"]) == ("NV12", 1920, 1080, Fraction(30, 1))
    assert st.is_fixed(fx)
    # a peer that wants 1280x720 I420 gets it
    peer = st.intersect_caps(other, st.parse_caps("video/x-raw, format=(string)I420, width=(int)1280, height=(int)720"))
    fx = s.fixate_caps(st.SINK, sink, peer)
    assert (fx["format"], fx["width"], fx["height"]) == ("I420", 1280, 720)
    with pytest.raises(StageError):
        s.fixate_caps(st.SINK, sink, [])


@pytest.mark.parametrize("fmt,w,h,size", [("BGRx", 1920, 1080, 8294400), ("NV12", 1920, 1080, 3110400),
                                          ("I420", 1920, 1080, 3110400), ("NV12", 3840, 2160, 12441600),
                                          ("I420", 7680, 4320, 49766400), ("I420", 1366, 768, 1368 * 768 + 2 * 684 * 384)])
def test_unit_sizes_match_survey_table(fmt, w, h, size):
    assert PreEncodeStage.get_unit_size({"format": fmt, "width": w, "height": h}) == size


def test_set_caps_rejects_bad_geometry_without_touching_the_gpu():
    s = PreEncodeStage()
    inc = {"name": "video/x-raw", "format": "BGRx", "width": 1921, "height": 1080}
    outc = {"name": "video/x-raw", "format": "NV12", "width": 1921, "height": 1080}
    assert s.set_caps(inc, outc) is False and "even" in s.last_error
    assert s.set_caps({"format": "NV12", "width": 2, "height": 2}, {"format": "NV12", "width": 2, "height": 2}) is False
    assert s.set_caps({"format": "BGRx", "width": IntRange(2, 4), "height": 2},
                      {"format": "NV12", "width": 2, "height": 2}) is False
    assert s.transform(0, 0, 0, 0) == st.FLOW_NOT_NEGOTIATED


# ---- the GstBase adaptor against a minimal fake `gi` (gi does not exist on this image) -------------
class _FakeCaps:
    def __init__(self, s):
        self.s = s

    @classmethod
    def from_string(cls, s):
        return cls(s)

    @classmethod
    def new_empty(cls):
        return cls("")

    def to_string(self):
        return self.s

    def fixate(self):
        return self


def _install_fake_gi(monkeypatch):
    gi = types.ModuleType("gi")
    gi.require_version = lambda *a: None
    repo = types.ModuleType("gi.repository")

    class _Enum:
        def __init__(self, **kw):
            self.__dict__.update(kw)

    Gst = types.ModuleType("Gst")
    Gst.Caps = _FakeCaps
    Gst.PadDirection = _Enum(SINK="sink", SRC="src")
    Gst.PadPresence = _Enum(ALWAYS=0)
    Gst.PadTemplate = types.SimpleNamespace(new=lambda *a: a)
    Gst.FlowReturn = _Enum(OK=0, ERROR=-5, NOT_NEGOTIATED=-4)
    Gst.MapFlags = _Enum(READ=1, WRITE=2)
    Gst.Rank = _Enum(NONE=0)
    Gst.errors = []
    Gst.error = Gst.errors.append
    registered = {}
    Gst.Element = types.SimpleNamespace(register=lambda plugin, name, rank, gtype: registered.setdefault(name, gtype) is gtype)
    Gst.registered = registered

    class BaseTransform:
        def __init__(self):
            self.qos = False

        def get_name(self):
            return "b200preencode0"

        def set_qos_enabled(self, v):
            self.qos = v

        def set_passthrough(self, v):
            self.passthrough = v

        def set_in_place(self, v):
            self.in_place = v

    GstBase = types.ModuleType("GstBase")
    GstBase.BaseTransform = BaseTransform
    GObject = types.ModuleType("GObject")
    for n in ("TYPE_UINT", "TYPE_INT", "TYPE_STRING", "TYPE_BOOLEAN", "TYPE_UINT64"):
        setattr(GObject, n, n)
    GObject.G_MAXUINT, GObject.G_MAXINT, GObject.G_MAXUINT64 = 2**32 - 1, 2**31 - 1, 2**64 - 1
    GObject.ParamFlags = _Enum(READWRITE=3, READABLE=1)
    GObject.type_register = lambda cls: cls
    Gst.CLOCK_TIME_NONE = 2**64 - 1

    class _Structure(dict):
        @classmethod
        def new_empty(cls, name):
            s = cls()
            s.name = name
            return s

        def set_value(self, k, v):
            self[k] = v

    Gst.Structure = _Structure

    class _Meta:
        def __init__(self, kind, rect):
            self.kind, self.rect, self.params = kind, rect, []

        def add_param(self, st):
            self.params.append(st)

    GstVideo = types.ModuleType("GstVideo")
    GstVideo.metas, GstVideo.events = [], []

    def add_roi(buf, kind, x, y, w, h):
        m = _Meta(kind, (x, y, w, h))
        GstVideo.metas.append(m)
        return m

    GstVideo.buffer_add_video_region_of_interest_meta = add_roi
    GstVideo.video_event_new_downstream_force_key_unit = lambda *a: ("force-key-unit",) + a
    GObject.TYPE_DOUBLE = "TYPE_DOUBLE"
    repo.Gst, repo.GstBase, repo.GObject, repo.GstVideo = Gst, GstBase, GObject, GstVideo
    gi.repository = repo
    for name, mod in (("gi", gi), ("gi.repository", repo)):
        monkeypatch.setitem(sys.modules, name, mod)
    return Gst


def test_gst_adaptor_negotiates_through_fake_gi(monkeypatch):
    Gst = _install_fake_gi(monkeypatch)
    from selkies_gstreamer_b200 import gst_element
    monkeypatch.setattr(gst_element, "_element_class", None)
    assert gst_element.register() is True
    assert gst_element.FACTORY_NAME in Gst.registered
    cls = gst_element.element_class()
    el = cls()
    assert el.qos is True and el.passthrough is False and el.in_place is False
    prop = types.SimpleNamespace
    el.do_set_property(prop(name="n-threads"), 4)
    el.do_set_property(prop(name="cuda-device-id"), 0)
    assert el.do_get_property(prop(name="n-threads")) == 4
    sink = Gst.Caps.from_string("video/x-raw, format=(string)BGRx, width=(int)1920, height=(int)1080, framerate=(fraction)60/1")
    filt = Gst.Caps.from_string("video/x-raw, format=(string)NV12")
    other = el.do_transform_caps(Gst.PadDirection.SINK, sink, filt)
    assert "NV12" in other.to_string() and "I420" not in other.to_string()
    fixed = el.do_fixate_caps(Gst.PadDirection.SINK, sink, other)
    parsed = st.parse_caps(fixed.to_string())[0]
    assert (parsed["format"], parsed["width"], parsed["height"]) == ("NV12", 1920, 1080)
    ok, size = el.do_get_unit_size(fixed)
    assert ok and size == 3110400
    ok, size = el.do_get_unit_size(sink)
    assert ok and size == 8294400
    assert el.do_start() is True and el.do_stop() is True
    # odd size is refused at set_caps and logged, as a failed negotiation would be
    bad = Gst.Caps.from_string("video/x-raw, format=(string)BGRx, width=(int)1919, height=(int)1080")
    assert el.do_set_caps(bad, fixed) is False and Gst.errors


def test_pipeline_patch_swaps_factory_names(monkeypatch):
    Gst = _install_fake_gi(monkeypatch)
    from selkies_gstreamer_b200 import gst_element, pipeline_patch
    monkeypatch.setattr(gst_element, "_element_class", None)
    made = []

    class FakeEl:
        def __init__(self, n):
            self.n, self.props = n, {}

        def set_property(self, k, v):
            self.props[k] = v

    def make(name, element_name=None):
        made.append(name)
        return FakeEl(name)

    Gst.ElementFactory = types.SimpleNamespace(make=make)
    with pipeline_patch.swapped_factories(Gst):
        a = Gst.ElementFactory.make("videoconvert")
        b = Gst.ElementFactory.make("cudaupload")
        c = Gst.ElementFactory.make("cudaconvert")
        d = Gst.ElementFactory.make("x264enc", "x264enc")
        b.set_property("cuda-device-id", 1)          # gstwebrtc_app.py:264-265 — swallowed on identity
        c.set_property("cuda-device-id", 1)
    assert made == ["b200preencode", "identity", "b200preencode", "x264enc"]
    assert b._e.props == {} and c.props == {"cuda-device-id": 1}
    assert Gst.ElementFactory.make is make


def test_gst_adaptor_signals_roi_and_keyframes(monkeypatch):
    Gst = _install_fake_gi(monkeypatch)
    from gi.repository import GstVideo
    from selkies_gstreamer_b200 import gst_element
    from selkies_gstreamer_b200.damage_policy import FrameDecision, Region
    monkeypatch.setattr(gst_element, "_element_class", None)
    el = gst_element.element_class()()
    pushed = []
    el.srcpad = types.SimpleNamespace(push_event=pushed.append)
    outbuf = types.SimpleNamespace(pts=1234)
    dec = FrameDecision(12, 920, True, False, [Region(80, 32, 64, 32)], [Region(0, 160, 16, 16, "paint-over")], 0.0)
    el.core.frames = 5
    assert el.apply_decision(outbuf, dec) == 2
    kinds = [(m.kind, m.rect) for m in GstVideo.metas]
    assert kinds == [("damage", (80, 32, 64, 32)), ("paint-over", (0, 160, 16, 16))]
    assert GstVideo.metas[1].params[0].name == "roi/x264enc" and GstVideo.metas[1].params[0]["delta-qp"] == -8.0
    assert pushed and pushed[0][0] == "force-key-unit" and pushed[0][1] == 1234
    # the first frame's keyframe is implicit; switches turn the signalling off
    el.core.frames = 1
    pushed.clear()
    el.apply_decision(outbuf, dec)
    assert not pushed
    prop = types.SimpleNamespace
    el.do_set_property(prop(name="roi-meta"), False)
    el.do_set_property(prop(name="force-keyframes"), False)
    el.core.frames = 9
    assert el.apply_decision(outbuf, dec) == 0 and not pushed
    el.do_set_property(prop(name="damage-block-threshold"), 4)
    assert el.do_get_property(prop(name="damage-block-threshold")) == 4
    assert el.do_get_property(prop(name="roi-meta")) is False


def test_cuda_memory_caps_are_offered_only_when_switched_on():
    """The NVENC branches ask for video/x-raw(memory:CUDAMemory),format=NV12 behind the stage
    (gstwebrtc_app.py:281-284): offered on the src pad with cuda-memory-output, refused without."""
    from selkies_gstreamer_b200 import stage as st
    sink = st.parse_caps("video/x-raw, format=(string)BGRx, width=(int)1920, height=(int)1080, framerate=(fraction)60/1")
    want = st.parse_caps("video/x-raw(memory:CUDAMemory), format=(string)NV12")
    s = st.PreEncodeStage()
    assert s.transform_caps(st.SINK, sink, want) == []                      # off: nothing to agree on
    assert not s.set_caps(sink[0], dict(want[0], width=1920, height=1080)) and "CUDAMemory" in s.last_error
    s.set_property("cuda-memory-output", True)
    got = s.transform_caps(st.SINK, sink, want)
    assert len(got) == 1 and got[0]["features"] == "memory:CUDAMemory" and got[0]["format"] == "NV12"
    fixed = s.fixate_caps(st.SINK, sink, got)
    assert fixed["features"] == "memory:CUDAMemory" and (fixed["width"], fixed["height"]) == (1920, 1080)
    both = s.transform_caps(st.SINK, sink)
    assert [c.get("features") for c in both] == [None, "memory:CUDAMemory"]
    # upstream direction: device-memory src caps map back to plain system-memory sink caps
    back = s.transform_caps(st.SRC, [dict(fixed)])
    assert back and all(c.get("features") is None for c in back) and "BGRx" in back[0]["format"]
    # the sink pad itself never takes device memory
    assert s.transform_caps(st.SINK, st.parse_caps("video/x-raw(memory:CUDAMemory), format=(string)BGRx, width=(int)64, height=(int)64")) == []
