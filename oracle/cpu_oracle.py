"""ctypes front end of the C oracle (oracle/sk_oracle.c).

TEST INFRASTRUCTURE ONLY — importable from tests/, __graft_entry__.smoke() and
bench.py's cpu_baseline / --impl reference legs; never from the product package.
PARITY UNPINNED (see sk_oracle.h): GStreamer 1.24.12 cannot be run here.
"""
from __future__ import annotations

import ctypes
import os
import subprocess
import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "libsk_oracle.so")

IN_BGRX, IN_RGBX = 0, 1
OUT_NV12, OUT_I420 = 0, 1
MATRIX_BT601, MATRIX_BT709, MATRIX_BT2020 = 0, 1, 2
SITE_NONE, SITE_H_COSITED = 0, 1
SCALE_GST_LQ, SCALE_PLAIN_LQ = 0, 1
# parity switches (sk_oracle.h SKO_PS_*; the product's SK_PS_* bits have the same values)
PS_MATRIX_ROUND, PS_COEFF_HALF_UP, PS_CS_EDGE_KEEP, PS_CHROMA_V_FIRST, PS_CHROMA_BOX4, PS_HINC_PLAIN, PS_V_FIRST_ON_TIE = (
    1, 2, 4, 8, 16, 32, 64)
PS_ALL = {"matrix-round": 1, "coeff-half-up": 2, "cs-edge-keep": 4, "chroma-v-first": 8, "chroma-box4": 16,
          "hinc-plain": 32, "v-first-on-tie": 64}


class SkoConfig(ctypes.Structure):
    _fields_ = [(n, ctypes.c_int) for n in (
        "in_w", "in_h", "in_stride", "in_fmt", "out_w", "out_h", "out_fmt",
        "matrix", "chroma_site", "scale_variant", "n_threads", "switches")]


def build(force: bool = False) -> str:
    """Compile libsk_oracle.so with gcc if missing or stale."""
    src = os.path.join(_HERE, "sk_oracle.c")
    hdr = os.path.join(_HERE, "sk_oracle.h")
    stale = (not os.path.exists(_LIB_PATH)
             or os.path.getmtime(_LIB_PATH) < max(os.path.getmtime(src), os.path.getmtime(hdr)))
    if force or stale:
        subprocess.check_call(["make", "-C", _HERE, "-B", "libsk_oracle.so"],
                              stdout=subprocess.DEVNULL)
    return _LIB_PATH


_lib = None


def lib() -> ctypes.CDLL:
    global _lib
    if _lib is None:
        build()
        L = ctypes.CDLL(_LIB_PATH)
        u8p = ctypes.POINTER(ctypes.c_uint8)
        L.sko_out_size.restype = ctypes.c_size_t
        L.sko_out_size.argtypes = [ctypes.c_int] * 3
        L.sko_out_layout.restype = None
        L.sko_out_layout.argtypes = [ctypes.c_int] * 3 + [
            ctypes.POINTER(ctypes.c_size_t), ctypes.POINTER(ctypes.c_int)]
        L.sko_matrix_coeffs.restype = None
        L.sko_matrix_coeffs.argtypes = [ctypes.c_int, ctypes.POINTER(ctypes.c_int)]
        L.sko_matrix_coeffs_sw.restype = None
        L.sko_matrix_coeffs_sw.argtypes = [ctypes.c_int, ctypes.c_int, ctypes.POINTER(ctypes.c_int)]
        L.sko_h_increment_sw.restype = ctypes.c_int32
        L.sko_h_increment_sw.argtypes = [ctypes.c_int] * 3
        L.sko_resampler_2tap.restype = None
        L.sko_resampler_2tap.argtypes = [ctypes.c_int] * 4 + [
            ctypes.POINTER(ctypes.c_int32), ctypes.POINTER(ctypes.c_int16)]
        L.sko_h_increment.restype = ctypes.c_int32
        L.sko_h_increment.argtypes = [ctypes.c_int, ctypes.c_int]
        L.sko_convert.restype = ctypes.c_int
        L.sko_convert.argtypes = [ctypes.POINTER(SkoConfig), u8p, u8p]
        L.sko_damage.restype = ctypes.c_int
        L.sko_damage.argtypes = [ctypes.c_int] * 3 + [u8p, u8p, u8p]
        for fn in (L.sko_chroma_down_h2, L.sko_chroma_down_h2_cs):
            fn.restype = None
            fn.argtypes = [u8p, ctypes.c_int]
        L.sko_chroma_down_v2.restype = None
        L.sko_chroma_down_v2.argtypes = [u8p, u8p, ctypes.c_int]
        L.sko_policy_update.restype = None
        L.sko_policy_update.argtypes = [ctypes.c_int, u8p, ctypes.POINTER(ctypes.c_uint32), u8p] + [ctypes.c_int] * 3
        _lib = L
    return _lib


def _u8p(a: np.ndarray):
    return a.ctypes.data_as(ctypes.POINTER(ctypes.c_uint8))


def out_size(w: int, h: int, fmt: int) -> int:
    return int(lib().sko_out_size(w, h, fmt))


def out_layout(w: int, h: int, fmt: int):
    off = (ctypes.c_size_t * 3)()
    st = (ctypes.c_int * 3)()
    lib().sko_out_layout(w, h, fmt, off, st)
    return list(off), list(st)


def matrix_coeffs(matrix: int, switches: int = 0) -> np.ndarray:
    c = (ctypes.c_int * 12)()
    lib().sko_matrix_coeffs_sw(matrix, switches, c)
    return np.array(list(c), dtype=np.int64).reshape(3, 4)


def resampler_2tap(in_size: int, out_size_: int, variant: int = SCALE_GST_LQ, precision: int = 6):
    off = np.zeros(out_size_, dtype=np.int32)
    tap = np.zeros(out_size_, dtype=np.int16)
    lib().sko_resampler_2tap(in_size, out_size_, variant, precision,
                             off.ctypes.data_as(ctypes.POINTER(ctypes.c_int32)),
                             tap.ctypes.data_as(ctypes.POINTER(ctypes.c_int16)))
    return off, tap


def h_increment(in_w: int, out_w: int, switches: int = 0) -> int:
    return int(lib().sko_h_increment_sw(in_w, out_w, switches))


def convert(frame: np.ndarray, out_w: int | None = None, out_h: int | None = None, *,
            in_fmt: int = IN_BGRX, out_fmt: int = OUT_NV12, matrix: int = MATRIX_BT709,
            chroma_site: int = SITE_H_COSITED, scale_variant: int = SCALE_GST_LQ,
            n_threads: int = 1, switches: int = 0) -> np.ndarray:
    """frame: uint8 [H, W, 4] (rows may be a strided view). Returns the packed output frame."""
    assert frame.dtype == np.uint8 and frame.ndim == 3 and frame.shape[2] == 4
    assert frame.strides[2] == 1 and frame.strides[1] == 4
    h, w = frame.shape[:2]
    out_w = out_w or w
    out_h = out_h or h
    cfg = SkoConfig(w, h, frame.strides[0], in_fmt, out_w, out_h, out_fmt, matrix,
                    chroma_site, scale_variant, n_threads, switches)
    out = np.zeros(out_size(out_w, out_h, out_fmt), dtype=np.uint8)
    rc = lib().sko_convert(ctypes.byref(cfg), _u8p(frame), _u8p(out))
    if rc != 0:
        raise ValueError(f"sko_convert failed: {rc}")
    return out


def damage(w: int, h: int, fmt: int, cur: np.ndarray, prev: np.ndarray | None):
    mbw, mbh = (w + 15) // 16, (h + 15) // 16
    m = np.zeros((mbh, mbw), dtype=np.uint8)
    n = lib().sko_damage(w, h, fmt, _u8p(cur), _u8p(prev) if prev is not None else None, _u8p(m))
    return m, int(n)


class PolicyTracker:
    """Host-side replay of the per-macroblock temporal policy (state lives here)."""

    def __init__(self, mb_w: int, mb_h: int, threshold: int = 10, duration: int = 20, trigger: int = 15):
        self.shape = (mb_h, mb_w)
        self.state = np.zeros(mb_w * mb_h, dtype=np.uint32)
        self.knobs = (threshold, duration, trigger)

    def update(self, damage_map: np.ndarray) -> np.ndarray:
        m = np.ascontiguousarray(damage_map, dtype=np.uint8).reshape(-1)
        pol = np.zeros(m.size, dtype=np.uint8)
        lib().sko_policy_update(m.size, _u8p(m), self.state.ctypes.data_as(ctypes.POINTER(ctypes.c_uint32)),
                                _u8p(pol), *self.knobs)
        return pol.reshape(self.shape)


def split_planes(buf: np.ndarray, w: int, h: int, fmt: int):
    """Views of Y and chroma planes of a packed output frame."""
    off, st = out_layout(w, h, fmt)
    h2 = (h + 1) & ~1
    y = buf[off[0]:off[0] + st[0] * h2].reshape(h2, st[0])[:h, :w]
    if fmt == OUT_I420:
        u = buf[off[1]:off[1] + st[1] * (h2 // 2)].reshape(h2 // 2, st[1])[:, :(w + 1) // 2]
        v = buf[off[2]:off[2] + st[2] * (h2 // 2)].reshape(h2 // 2, st[2])[:, :(w + 1) // 2]
        return y, u, v
    uv = buf[off[1]:off[1] + st[1] * (h2 // 2)].reshape(h2 // 2, st[1])[:, :2 * ((w + 1) // 2)]
    return y, uv[:, 0::2], uv[:, 1::2]
