"""Element logic without GStreamer: properties, caps negotiation, per-frame transform.

This is the host-side mirror of the element contract the reference relies on
when it instantiates and links the stock stage
(src/selkies/legacy/gstwebrtc_app.py:611-617 videoconvert, :263-284
cudaupload+cudaconvert, :941-994 add/link):

  * GObject properties it sets: `n-threads` (:612), `qos` (:613, :277),
    `cuda-device-id` (:264-272); videoscale's `method` / `add-borders` for the
    scale capability the stage also covers.
  * caps: sink video/x-raw {BGRx,BGRA,RGBx,RGBA}; src video/x-raw {NV12,I420},
    format forced by the downstream capsfilter (:614-617).
  * GstBaseTransform vmethods: transform_caps / fixate_caps / set_caps /
    get_unit_size / start / stop / transform.

`gst_element.py` adapts this class to a real GstBase.BaseTransform; keeping the
logic here lets it be tested where `gi` does not exist.  Pixel work is never
done here: transform() hands pointers to libselkies_preencode.so.
"""
from __future__ import annotations

import ctypes
import re
import threading
from fractions import Fraction

from . import _abi, colorimetry as _col
from .coalesce import LaunchCoalescer
from .damage_policy import DamagePolicy
from .hostbuf import PinCache
from .session import PreencodeError, PreencodeSession, frame_layout

SINK_FORMATS = ("BGRx", "BGRA", "RGBx", "RGBA")
SRC_FORMATS = ("NV12", "I420")
MIN_DIM, MAX_DIM = 2, 8192           # reference caps the display at 7680x4320 (resize.py:66)

FLOW_OK, FLOW_NOT_NEGOTIATED, FLOW_ERROR = 0, -4, -5      # Gst.FlowReturn values
SINK, SRC = "sink", "src"

SINK_TEMPLATE = ("video/x-raw, format=(string){ BGRx, BGRA, RGBx, RGBA }, "
                 "width=(int)[ 2, 8192 ], height=(int)[ 2, 8192 ], framerate=(fraction)[ 0/1, 2147483647/1 ]")
SRC_TEMPLATE = ("video/x-raw, format=(string){ NV12, I420 }, "
                "width=(int)[ 2, 8192 ], height=(int)[ 2, 8192 ], framerate=(fraction)[ 0/1, 2147483647/1 ]")
# The NVENC branches of the reference ask for this on the stage's src side (gstwebrtc_app.py:281-284, :377-380, :434-437)
CUDA_FEATURE = "memory:CUDAMemory"
SRC_TEMPLATE_CUDA = SRC_TEMPLATE.replace("video/x-raw,", f"video/x-raw({CUDA_FEATURE}),", 1)


class IntRange(tuple):
    """[lo, hi] inclusive."""
    def __new__(cls, lo, hi):
        return super().__new__(cls, (int(lo), int(hi)))


# ----------------------------------------------------------------- caps (subset of GstCaps syntax)
def parse_caps(text: str) -> list[dict]:
    """Parse the subset of Gst.Caps.to_string() this stage negotiates with."""
    out = []
    for part in _split_top(text, ";"):
        part = part.strip()
        if not part:
            continue
        fields = _split_top(part, ",")
        name = fields[0].strip()
        feat = None
        m = re.match(r"^([^()]+)\((.+)\)$", name)
        if m:
            name, feat = m.group(1).strip(), m.group(2).strip()
        st = {"name": name}
        if feat:
            st["features"] = feat
        for f in fields[1:]:
            if "=" not in f:
                continue
            k, v = f.split("=", 1)
            st[k.strip()] = _parse_value(v.strip())
        out.append(st)
    return out


def _split_top(s: str, sep: str) -> list[str]:
    parts, depth, cur = [], 0, []
    for ch in s:
        if ch in "{[(<":
            depth += 1
        elif ch in "}])>":
            depth -= 1
        if ch == sep and depth == 0:
            parts.append("".join(cur)); cur = []
        else:
            cur.append(ch)
    parts.append("".join(cur))
    return parts


def _parse_scalar(v: str):
    v = v.strip().strip('"')
    if re.fullmatch(r"-?\d+", v):
        return int(v)
    if re.fullmatch(r"-?\d+/\d+", v):
        return Fraction(v)
    return v


def _parse_value(v: str):
    v = re.sub(r"^\(\w+\)\s*", "", v)
    if v.startswith("{"):
        return [_parse_scalar(x) for x in _split_top(v[1:-1], ",")]
    if v.startswith("["):
        lo, hi = [_parse_scalar(x) for x in _split_top(v[1:-1], ",")][:2]
        if isinstance(lo, int) and isinstance(hi, int):
            return IntRange(lo, hi)
        return (lo, hi)
    return _parse_scalar(v)


def caps_to_string(caps: list[dict]) -> str:
    def val(v):
        if isinstance(v, IntRange):
            return f"(int)[ {v[0]}, {v[1]} ]"
        if isinstance(v, tuple):
            return f"(fraction)[ {v[0]}, {v[1]} ]" if isinstance(v[0], Fraction) else f"[ {v[0]}, {v[1]} ]"
        if isinstance(v, list):
            return "{ " + ", ".join(str(x) for x in v) + " }"
        if isinstance(v, Fraction):
            return f"(fraction){v.numerator}/{v.denominator}"
        if isinstance(v, int):
            return f"(int){v}"
        return f"(string){v}"
    out = []
    for st in caps:
        name = st["name"] + (f"({st['features']})" if st.get("features") else "")
        out.append(", ".join([name] + [f"{k}={val(v)}" for k, v in st.items() if k not in ("name", "features")]))
    return "; ".join(out)


def _as_list(v):
    return v if isinstance(v, list) else [v]


def _intersect_value(a, b):
    """Intersection of two field values (scalar / list / IntRange); None if empty."""
    if isinstance(a, IntRange) and isinstance(b, IntRange):
        lo, hi = max(a[0], b[0]), min(a[1], b[1])
        return None if lo > hi else (lo if lo == hi else IntRange(lo, hi))
    if isinstance(a, IntRange) or isinstance(b, IntRange):
        rng, other = (a, b) if isinstance(a, IntRange) else (b, a)
        keep = [x for x in _as_list(other) if isinstance(x, int) and rng[0] <= x <= rng[1]]
        return None if not keep else (keep[0] if len(keep) == 1 else keep)
    if isinstance(a, tuple) or isinstance(b, tuple):      # fraction ranges: accept the other side
        return b if isinstance(a, tuple) else a
    keep = [x for x in _as_list(a) if x in _as_list(b)]
    return None if not keep else (keep[0] if len(keep) == 1 else keep)


def intersect_caps(a: list[dict], b: list[dict]) -> list[dict]:
    out = []
    for sa in a:
        for sb in b:
            if sa["name"] != sb["name"] or sa.get("features") != sb.get("features"):
                continue
            st, ok = {"name": sa["name"]}, True
            if sa.get("features"):
                st["features"] = sa["features"]
            for k in list(sa.keys()) + [k for k in sb if k not in sa]:
                if k in ("name", "features"):
                    continue
                if k in sa and k in sb:
                    v = _intersect_value(sa[k], sb[k])
                    if v is None:
                        ok = False
                        break
                    st[k] = v
                else:
                    st[k] = sa.get(k, sb.get(k))
            if ok:
                out.append(st)
    return out


def is_fixed(st: dict) -> bool:
    return all(not isinstance(v, (list, tuple)) for k, v in st.items())


# ----------------------------------------------------------------- the stage
class StageError(RuntimeError):
    pass


class PreEncodeStage:
    """One convert[+scale]+damage stage instance (what one videoconvert element is)."""

    PROPERTIES = {
        # name: (type, default, validator)
        "qos": (bool, True, None),                                   # gstwebrtc_app.py:613 / :277
        "n-threads": (int, 1, lambda v: 0 <= v <= 0xFFFFFFFF),       # gstwebrtc_app.py:612 (accepted; the GPU path has no CPU threads)
        "cuda-device-id": (int, -1, lambda v: v >= -1),              # gstwebrtc_app.py:264-272
        "method": (str, "bilinear", lambda v: v in ("bilinear", 1)),  # videoscale method; only bilinear
        "add-borders": (bool, False, lambda v: v is False),
        # videoconvert / videoscale properties the reference leaves at their defaults: accepted at the values
        # the arithmetic implements, refused otherwise (so a mis-set pipeline fails loudly, not subtly)
        "dither": (str, "bayer", lambda v: v in ("none", "floyd-steinberg", "sierra-lite", "bayer", "verterr")),  # inert at 8 -> 8 bit
        "dither-quantization": (int, 1, lambda v: v >= 1),
        "chroma-resampler": (str, "linear", lambda v: v == "linear"),
        "chroma-mode": (str, "full", lambda v: v == "full"),
        "matrix-mode": (str, "full", lambda v: v == "full"),
        "gamma-mode": (str, "none", lambda v: v == "none"),
        "primaries-mode": (str, "none", lambda v: v == "none"),
        "alpha-mode": (str, "copy", lambda v: v in ("copy", "set", "mult")),      # no alpha on the output formats
        "alpha-value": (float, 1.0, lambda v: 0.0 <= v <= 1.0),
        "sharpness": (float, 1.0, lambda v: v == 1.0),
        "sharpen": (float, 0.0, lambda v: v == 0.0),
        "gamma-decode": (bool, False, lambda v: v is False),
        "damage": (bool, True, None),
        "n-inflight": (int, 2, lambda v: 2 <= v <= 8),
        "scale-variant": (int, _abi.SK_SCALE_VARIANT_GST_LQ, lambda v: v in (0, 1)),
        # temporal damage policy knobs (src/selkies/selkies.py:2949-2951)
        "damage-block-threshold": (int, 10, lambda v: 1 <= v <= 255),
        "damage-block-duration": (int, 20, lambda v: 1 <= v <= 255),
        "paint-over-trigger-frames": (int, 15, lambda v: 1 <= v <= 255),
        "keyframe-damage-fraction": (float, 0.6, lambda v: 0.0 < v <= 1.0),
        "min-keyframe-distance": (int, 60, lambda v: v >= 0),        # frames; cf. min_keyframe_frame_distance (gstwebrtc_app.py:1227-1262)
        # host side of the frame path
        "pin-buffers": (bool, True, None),           # page-lock the (recycled) buffers transform() is handed: DMA both ways
        "pin-cache-size": (int, 16, lambda v: 0 <= v <= 256),
        "coalesce": (bool, False, None),             # share launches with the other stages of this process (coalesce.py)
        "coalesce-window-us": (int, 200, lambda v: 0 <= v <= 100000),
        "spin-wait": (bool, False, None),            # busy-poll for completion instead of sleeping
        # offer video/x-raw(memory:CUDAMemory) on the src pad: the frame stays in HBM for a CUDA consumer (NVENC);
        # the Gst adaptor switches it on when the GstCuda typelib can be imported
        "cuda-memory-output": (bool, False, None),
    }
    READ_ONLY = ("frames", "damaged-blocks", "damaged-fraction", "total-blocks")

    def __init__(self, name: str = "b200preencode0"):
        self.name = name
        self._props = {k: d for k, (_, d, _) in self.PROPERTIES.items()}
        # set_caps / set_property may come from another thread than transform (§8b).  transform() holds the
        # lock only to submit; while it sleeps on the GPU the session is kept alive by `_busy`.
        self._lock = threading.Condition()
        self._busy = 0
        self._pins = PinCache(16)
        self._coalescer: LaunchCoalescer | None = None
        self._session: PreencodeSession | None = None
        self._in = self._out = None
        self._started = False
        self.frames = 0
        self.last_damage_map = None
        self.last_damaged = 0
        self.last_error = ""
        self.on_damage = None               # callable(stage, map, count): raw hook
        self.on_decision = None             # callable(stage, FrameDecision): ROI rectangles / keyframe / paint-over
        self.last_policy_map = None
        self.last_decision = None
        self._policy = None

    # -- properties ---------------------------------------------------------
    def set_property(self, name: str, value):
        if name in self.READ_ONLY:
            raise StageError(f"property '{name}' of '{self.name}' is not writable")
        if name not in self.PROPERTIES:
            raise StageError(f"object class 'B200PreEncode' has no property named '{name}'")
        typ, _, check = self.PROPERTIES[name]
        if typ is bool:
            if not isinstance(value, (bool, int)):
                raise StageError(f"property '{name}' expects a boolean")
            value = bool(value)
        elif typ is int:
            if isinstance(value, bool) or not isinstance(value, int):
                raise StageError(f"property '{name}' expects an integer")
        elif typ is float:
            if isinstance(value, bool) or not isinstance(value, (int, float)):
                raise StageError(f"property '{name}' expects a number")
            value = float(value)
        if name == "method" and value == 1:
            value = "bilinear"
        if typ is str and not isinstance(value, str):
            raise StageError(f"property '{name}' expects a string")
        if check and not check(value):
            raise StageError(f"value {value!r} out of range for property '{name}'")
        with self._lock:
            changed = self._props[name] != value
            self._props[name] = value
            needs_session = name in ("cuda-device-id", "damage", "n-inflight", "scale-variant", "spin-wait", "coalesce",
                                     "coalesce-window-us", "damage-block-threshold", "damage-block-duration",
                                     "paint-over-trigger-frames")
            if name == "keyframe-damage-fraction" and self._policy is not None:
                self._policy.keyframe_fraction = value
            if name == "min-keyframe-distance" and self._policy is not None:
                self._policy.min_keyframe_distance = value
            if name in ("pin-buffers", "pin-cache-size") and changed:
                self._pins.release()
                self._pins = PinCache(self._props["pin-cache-size"] if self._props["pin-buffers"] else 0)
            if changed and needs_session and self._session is not None:
                self._recreate_locked()

    def get_property(self, name: str):
        if name == "frames":
            return self.frames
        if name == "damaged-blocks":
            return self.last_damaged
        if name == "total-blocks":
            return 0 if not self._out else ((self._out["width"] + 15) // 16) * ((self._out["height"] + 15) // 16)
        if name == "damaged-fraction":
            t = self.get_property("total-blocks")
            return (self.last_damaged / t) if t else 0.0
        if name not in self._props:
            raise StageError(f"object class 'B200PreEncode' has no property named '{name}'")
        return self._props[name]

    # -- negotiation ----------------------------------------------------------
    def template(self, direction: str) -> list[dict]:
        if direction == SINK:
            return parse_caps(SINK_TEMPLATE)
        return parse_caps(SRC_TEMPLATE + ("; " + SRC_TEMPLATE_CUDA if self._props["cuda-memory-output"] else ""))

    def transform_caps(self, direction: str, caps: list[dict], filter_caps: list[dict] | None = None) -> list[dict]:
        """Caps possible on the OTHER pad given `caps` on the `direction` pad.

        Format becomes every format of the other side, size becomes the full range
        (the stage scales both ways; fixate_caps prefers the input size), colorimetry and
        chroma-site do not carry across an RGB↔YUV change; everything else
        (framerate, pixel-aspect-ratio, ...) is kept."""
        out = []
        other_formats = list(SRC_FORMATS if direction == SINK else SINK_FORMATS)
        cuda = self._props["cuda-memory-output"]
        for st in caps:
            feat = st.get("features")
            if st.get("name") != "video/x-raw":
                continue
            # the sink pad takes system memory only; the src pad may also hand out device memory
            if feat not in (None, "memory:SystemMemory") and not (direction == SRC and cuda and feat == CUDA_FEATURE):
                continue
            o = {k: v for k, v in st.items()
                 if k not in ("format", "colorimetry", "chroma-site", "width", "height", "features")}
            o["format"] = other_formats
            for dim in ("width", "height"):
                o[dim] = IntRange(MIN_DIM, MAX_DIM)       # like upstream videoconvertscale: any size either way
            out.append(o)
            if direction == SINK and cuda:
                out.append(dict(o, features=CUDA_FEATURE))
        out = intersect_caps(out, self.template(SRC if direction == SINK else SINK))
        if filter_caps is not None:
            out = intersect_caps(filter_caps, out)
        return out

    def fixate_caps(self, direction: str, caps: list[dict], othercaps: list[dict]) -> dict:
        """Pick one structure of `othercaps`: same size as the input if allowed, else the
        nearest allowed size; first listed format; framerate of the input."""
        if not othercaps:
            raise StageError("not negotiated: empty caps")
        src = caps[0]
        st = dict(othercaps[0])
        for dim in ("width", "height"):
            want = src.get(dim)
            v = st.get(dim)
            if isinstance(v, IntRange):
                pick = want if isinstance(want, int) else v[1]
                pick = min(max(pick, v[0]), v[1])
                st[dim] = pick & ~1 if pick > 1 else pick
            elif isinstance(v, list):
                st[dim] = min(v, key=lambda x: abs(x - (want if isinstance(want, int) else x)))
        if isinstance(st.get("format"), list):
            st["format"] = st["format"][0]
        fr = st.get("framerate")
        if isinstance(fr, (tuple, list)):
            st["framerate"] = src.get("framerate") if isinstance(src.get("framerate"), Fraction) else Fraction(60, 1)
        for k, v in list(st.items()):
            if isinstance(v, list):
                st[k] = v[0]
        return st

    @staticmethod
    def get_unit_size(caps: dict) -> int:
        """Bytes of one frame of fixed caps (GstVideoInfo default layout)."""
        w, h, fmt = caps["width"], caps["height"], caps["format"]
        if fmt in SINK_FORMATS:
            return 4 * w * h
        if fmt in SRC_FORMATS:
            return frame_layout(w, h, fmt)[0]
        raise StageError(f"unsupported format {fmt}")

    def start(self) -> bool:
        self._started = True
        return True

    def stop(self) -> bool:
        with self._lock:
            self._close_session_locked()
            self._pins.release()
            self._started = False
        return True

    def _close_session_locked(self):
        """Destroy the GPU session once no transform() is sleeping on it."""
        self._lock.wait_for(lambda: self._busy == 0)
        if self._session is not None:
            if self._coalescer is not None:
                self._coalescer.detach(self._session)
                self._coalescer = None
            self._session.close()
            self._session = None

    def set_caps(self, incaps: dict, outcaps: dict) -> bool:
        """(Re)create the GPU session for fixed caps; the next frame has no predecessor."""
        try:
            for c in (incaps, outcaps):
                if not is_fixed(c):
                    raise StageError("caps are not fixed")
            if incaps["format"] not in SINK_FORMATS or outcaps["format"] not in SRC_FORMATS:
                raise StageError(f"cannot convert {incaps['format']} to {outcaps['format']}")
            for c in (incaps, outcaps):
                if c["width"] % 2 or c["height"] % 2:
                    raise StageError("width and height must be even (the reference forces even sizes: "
                                     "webrtc_input.py:677, resize.py:28)")
            with self._lock:
                same = (self._in is not None and self._session is not None and
                        all(self._in.get(k) == incaps.get(k) for k in ("format", "width", "height")) and
                        all(self._out.get(k) == outcaps.get(k) for k in
                            ("format", "width", "height", "colorimetry", "chroma-site")))
                if outcaps.get("features") == CUDA_FEATURE and not self._props["cuda-memory-output"]:
                    raise StageError("memory:CUDAMemory output was not offered (cuda-memory-output is off)")
                same = same and (self._out.get("features") == outcaps.get("features"))
                self._in, self._out = dict(incaps), dict(outcaps)
                if same:
                    return True          # e.g. framerate-only renegotiation (set_framerate, gstwebrtc_app.py:1264-1266)
                self._recreate_locked()
            return True
        except (StageError, PreencodeError, ValueError, KeyError) as e:
            self.last_error = str(e)
            return False

    def _recreate_locked(self):
        self._close_session_locked()
        self._pins.release()            # new caps, new buffer pools
        i, o = self._in, self._out
        # The frame comes back into the buffer transform() is given (sk_submit_io), so the session
        # keeps no pinned output ring of its own.
        self._session = PreencodeSession(
            i["width"], i["height"], in_format=i["format"], out_format=o["format"],
            out_width=o["width"], out_height=o["height"], colorimetry=o.get("colorimetry"),
            chroma_site=o.get("chroma-site"), device_id=self._props["cuda-device-id"],
            n_inflight=self._props["n-inflight"], damage=self._props["damage"], host_output=False,
            scale_variant=self._props["scale-variant"], policy=self._props["damage"],
            damage_block_threshold=self._props["damage-block-threshold"],
            damage_block_duration=self._props["damage-block-duration"],
            paint_over_trigger_frames=self._props["paint-over-trigger-frames"],
            spin_wait=self._props["spin-wait"])
        if self._props["coalesce"]:
            self._coalescer = LaunchCoalescer.for_session(self._session, self._props["coalesce-window-us"])
            self._coalescer.attach()
        self._policy = DamagePolicy(o["width"], o["height"], self._props["keyframe-damage-fraction"],
                                    self._props["min-keyframe-distance"])
        self.frames = 0

    @property
    def session(self) -> PreencodeSession | None:
        return self._session

    @property
    def device_output(self) -> bool:
        """True when the negotiated src caps are video/x-raw(memory:CUDAMemory): use transform_device()."""
        return bool(self._out) and self._out.get("features") == CUDA_FEATURE

    def src_caps_fields(self) -> dict:
        """colorimetry / chroma-site actually used (what the src caps should advertise)."""
        s = self._session
        return {} if s is None else {"colorimetry": _col.MATRIX_NAMES[s.matrix], "chroma-site": _col.SITE_NAMES[s.site]}

    # -- per frame --------------------------------------------------------------
    def transform(self, in_ptr: int, in_size: int, out_ptr: int, out_size: int, in_stride: int = 0,
                  pin: bool | None = None) -> int:
        """One frame: raw host pointers in, raw host pointers out. Returns a FlowReturn value.

        H2D straight from the input buffer, D2H straight into the output buffer — no host-side copy.
        Buffers that come round again (buffer pools) are page-locked on first sight (`pin-buffers`;
        `pin=False` for one-off buffers: a pinned buffer must stay mapped until the stage stops or
        renegotiates, as pool memory does)."""
        with self._lock:
            s = self._session
            if s is None:
                self.last_error = "not negotiated"
                return FLOW_NOT_NEGOTIATED
            try:
                stride = in_stride or 4 * s.width
                need_in = stride * (s.height - 1) + 4 * s.width
                if in_size < need_in:
                    raise StageError(f"input buffer too small: {in_size} bytes")
                if out_size < s.frame_bytes:
                    raise StageError(f"output buffer too small: {out_size} < {s.frame_bytes}")
                if self._props["pin-buffers"] if pin is None else pin:
                    self._pins.ensure(in_ptr, in_size)
                    self._pins.ensure(out_ptr, out_size)
                co = self._coalescer if stride == s.in_stride else None
                fid = None
                if co is None:
                    # both buffers stay mapped until wait() returns
                    fid = s.submit_io(self.frames % s.n_inflight, in_ptr, stride, out_ptr)
                self._busy += 1
            except (StageError, PreencodeError) as e:
                self.last_error = str(e)
                return FLOW_ERROR
        # The GPU works; this thread sleeps without holding the lock (properties stay settable).
        try:
            r = co.convert(s, in_ptr, stride, out_ptr) if co is not None else s.wait(fid)
            self._after_frame(r)
            return FLOW_OK
        except (StageError, PreencodeError) as e:
            self.last_error = str(e)
            return FLOW_ERROR
        finally:
            with self._lock:
                self._busy -= 1
                self._lock.notify_all()

    def _after_frame(self, r):
        self.frames += 1
        # copies: the session's pinned map buffers are reused n_inflight frames later
        self.last_damage_map = None if r.damage_map is None else r.damage_map.copy()
        self.last_policy_map = None if r.policy_map is None else r.policy_map.copy()
        self.last_damaged = r.damaged_count
        if r.damage_map is not None:
            if self.on_damage is not None:
                self.on_damage(self, self.last_damage_map, r.damaged_count)
            self.last_decision = self._policy.decide(self.last_damage_map, self.last_policy_map, r.damaged_count)
            if self.on_decision is not None:
                self.on_decision(self, self.last_decision)

    def transform_device(self, in_ptr: int, in_size: int, consumer_stream: int = 0, in_stride: int = 0,
                         pin: bool | None = None):
        """One frame whose output STAYS IN HBM (src caps video/x-raw(memory:CUDAMemory), the reference's NVENC
        branches gstwebrtc_app.py:281-284): H2D from the input buffer, kernel, only the maps come back.
        Returns (FlowReturn, FrameResult or None); the result carries `dev_frame`, `plane_offset`, `plane_stride`
        and `frame_id`.  `consumer_stream` (a cudaStream_t, 0 = none) is made to wait for the kernel; call
        release_device(frame_id, stream) after queueing the consumer's read, or copy the surface before
        n-inflight more frames are converted."""
        with self._lock:
            s = self._session
            if s is None:
                self.last_error = "not negotiated"
                return FLOW_NOT_NEGOTIATED, None
            try:
                stride = in_stride or 4 * s.width
                if in_size < stride * (s.height - 1) + 4 * s.width:
                    raise StageError(f"input buffer too small: {in_size} bytes")
                if self._props["pin-buffers"] if pin is None else pin:
                    self._pins.ensure(in_ptr, in_size)
                fid = s.submit_io(self.frames % s.n_inflight, in_ptr, stride, None)
                if consumer_stream:
                    s.wait_device(fid, consumer_stream)
                self._busy += 1
            except (StageError, PreencodeError) as e:
                self.last_error = str(e)
                return FLOW_ERROR, None
        try:
            r = s.wait(fid)                 # maps and count only: the session has no host output
            self._after_frame(r)
            return FLOW_OK, r
        except (StageError, PreencodeError) as e:
            self.last_error = str(e)
            return FLOW_ERROR, None
        finally:
            with self._lock:
                self._busy -= 1
                self._lock.notify_all()

    def release_device(self, frame_id: int, consumer_stream: int):
        with self._lock:
            if self._session is not None:
                self._session.release_device(frame_id, consumer_stream)

    def transform_array(self, frame):
        """numpy convenience around transform(): returns (packed output, damage map, count)."""
        import numpy as np
        s = self._session
        if s is None:
            raise StageError("not negotiated")
        frame = np.ascontiguousarray(frame)
        out = np.empty(s.frame_bytes, np.uint8)
        rc = self.transform(frame.ctypes.data, frame.nbytes, out.ctypes.data, out.nbytes, pin=False)
        if rc != FLOW_OK:
            raise StageError(self.last_error)
        return out, (None if self.last_damage_map is None else self.last_damage_map.copy()), self.last_damaged

    def reset_history(self):
        with self._lock:
            if self._session is not None:
                self._session.reset_history()
            if self._policy is not None:
                self._policy.reset()
