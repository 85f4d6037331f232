"""Closed-form numpy restatement of SURVEY.md Appendix A — a second, independently
written statement of the same arithmetic, used only to cross-check the C oracle
(oracle/sk_oracle.c follows upstream's line-by-line, table-driven structure; this
file uses the whole-frame closed forms).

TEST INFRASTRUCTURE ONLY.  PARITY UNPINNED (GStreamer 1.24.12 is not runnable here).
"""
from __future__ import annotations

import numpy as np

# Appendix A.2 — integer coefficients, rows Y,U,V x columns R,G,B,offset.
COEFFS = {
    "bt601": np.array([[66, 129, 25, 4096], [-38, -74, 112, 32768], [112, -94, -18, 32768]], dtype=np.int64),
    "bt709": np.array([[47, 157, 16, 4096], [-26, -87, 112, 32768], [112, -102, -10, 32768]], dtype=np.int64),
    "bt2020": np.array([[58, 149, 13, 4096], [-31, -81, 112, 32768], [112, -103, -9, 32768]], dtype=np.int64),
}


def rgb_planes(frame: np.ndarray, in_fmt: str):
    f = frame.astype(np.int64)
    if in_fmt.upper() in ("BGRX", "BGRA"):
        return f[..., 2], f[..., 1], f[..., 0]
    return f[..., 0], f[..., 1], f[..., 2]


# Parity switches (sk_oracle.h SKO_PS_*): same bit values, independent implementation.
PS_MATRIX_ROUND, PS_COEFF_HALF_UP, PS_CS_EDGE_KEEP, PS_CHROMA_V_FIRST, PS_CHROMA_BOX4, PS_HINC_PLAIN, PS_V_FIRST_ON_TIE = (
    1, 2, 4, 8, 16, 32, 64)


def to_yuv444(r, g, b, matrix: str, switches: int = 0):
    c = COEFFS[matrix]          # none of the 27 coefficients sits near a tie: PS_COEFF_HALF_UP changes nothing
    rnd = 128 if switches & PS_MATRIX_ROUND else 0
    out = []
    for i in range(3):
        out.append((c[i, 0] * r + c[i, 1] * g + c[i, 2] * b + c[i, 3] + rnd) >> 8)
    return out


def avg(a, b):
    return (a + b + 1) >> 1


def _h_filter(c: np.ndarray, site: str, switches: int) -> np.ndarray:
    """[H, W] -> [H, W/2]: the horizontal half of the 4:2:0 down-sampling."""
    w = c.shape[1]
    if site == "none":
        return avg(c[:, 0::2], c[:, 1::2])
    ce = c[:, 0::2]            # even samples
    co = c[:, 1::2]            # odd samples (never modified upstream)
    left = np.empty_like(ce)
    left[:, 1:] = co[:, :-1]
    left[:, 0] = ce[:, 0]      # [3 1] opening tap == [1 2 1] with left := centre
    right = co.copy()
    if w > 2:                  # W == 2: only the opening tap runs upstream
        right[:, -1] = ce[:, -1]   # [1 3] closing tap == [1 2 1] with right := centre
    hh = (left + 2 * ce + right + 2) >> 2
    if switches & PS_CS_EDGE_KEEP:
        hh[:, 0] = ce[:, 0]
        if w > 2:
            hh[:, -1] = ce[:, -1]
    return hh


def chroma_420(c: np.ndarray, site: str, switches: int = 0) -> np.ndarray:
    """c: int64 [H, W] full-resolution chroma; returns [H/2, W/2]."""
    if site == "none" and switches & PS_CHROMA_BOX4:
        return (c[0::2, 0::2] + c[0::2, 1::2] + c[1::2, 0::2] + c[1::2, 1::2] + 2) >> 2
    if switches & PS_CHROMA_V_FIRST:
        return _h_filter(avg(c[0::2, :], c[1::2, :]), site, switches)
    hh = _h_filter(c, site, switches)
    return avg(hh[0::2, :], hh[1::2, :])


def scale_2tap_lq(rgb: np.ndarray, out_w: int, out_h: int, v_off, v_tap, h_inc: int, h_first: bool):
    """rgb: int64 [H, W, C]; taps come from the caller (tables are data, not code)."""
    def hs(a):
        if a.shape[1] == out_w:
            return a
        acc = np.arange(out_w, dtype=np.int64) * h_inc
        x0 = acc >> 16
        f = (acc >> 8) & 0xFF
        return (a[:, x0] * (256 - f)[None, :, None] + a[:, x0 + 1] * f[None, :, None]) >> 8

    def vs(a):
        if a.shape[0] == out_h:
            return a
        o = np.asarray(v_off, dtype=np.int64)
        p = np.asarray(v_tap, dtype=np.int64)[:, None, None]
        s1 = a[o]
        s2 = a[np.minimum(o + 1, a.shape[0] - 1)]
        return np.clip(s1 + (((s2 - s1) * p + 32) >> 6), 0, 255)

    return vs(hs(rgb)) if h_first else hs(vs(rgb))


def convert(frame: np.ndarray, *, in_fmt="BGRx", out_fmt="NV12", matrix="bt709", site="h-cosited",
            out_w=None, out_h=None, v_off=None, v_tap=None, h_inc=0, switches=0):
    """Returns (Y[H,W], U[H/2,W/2], V[H/2,W/2]) as uint8 — layout-free."""
    h, w = frame.shape[:2]
    out_w = out_w or w
    out_h = out_h or h
    r, g, b = rgb_planes(frame, in_fmt)
    scaling = (out_w, out_h) != (w, h)
    h_first = out_w * h <= w * out_h
    if switches & PS_V_FIRST_ON_TIE and out_w * h == w * out_h:
        h_first = False
    growing = out_w * out_h > w * h          # then the scaler runs behind the matrix, on Y/U/V
    if scaling and not growing:
        rgb = np.stack([r, g, b], axis=-1)
        rgb = scale_2tap_lq(rgb, out_w, out_h, v_off, v_tap, h_inc, h_first)
        r, g, b = rgb[..., 0], rgb[..., 1], rgb[..., 2]
    y, u, v = to_yuv444(r, g, b, matrix, switches)
    if scaling and growing:
        yuv = scale_2tap_lq(np.stack([y, u, v], axis=-1).astype(np.int64), out_w, out_h, v_off, v_tap, h_inc, h_first)
        y, u, v = yuv[..., 0], yuv[..., 1], yuv[..., 2]
    assert y.min() >= 0 and y.max() <= 255 and u.min() >= 0 and u.max() <= 255 and v.min() >= 0 and v.max() <= 255
    return (y.astype(np.uint8), chroma_420(u, site, switches).astype(np.uint8), chroma_420(v, site, switches).astype(np.uint8))


def damage(y, u, v, py, pu, pv):
    """Appendix A.6 on plane arrays; previous planes None → all ones."""
    h, w = y.shape
    mbh, mbw = (h + 15) // 16, (w + 15) // 16
    if py is None:
        return np.ones((mbh, mbw), dtype=np.uint8)
    m = np.zeros((mbh, mbw), dtype=np.uint8)
    dy = (y != py)
    dc = (u != pu) | (v != pv)
    for by in range(mbh):
        for bx in range(mbw):
            a = dy[by * 16:by * 16 + 16, bx * 16:bx * 16 + 16].any()
            c = dc[by * 8:by * 8 + 8, bx * 8:bx * 8 + 8].any()
            m[by, bx] = 1 if (a or c) else 0
    return m
