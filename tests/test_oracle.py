"""CPU tests of the oracle itself (no GPU): known answers, the second (numpy)
restatement, chroma filter vectors, layout, resampler properties, damage.

The oracle is a restatement of GStreamer 1.24.12's generic videoconvert path —
parity unpinned (GStreamer cannot run here); SURVEY.md A.7 vectors are the pin.
"""
import json
import os

import numpy as np
import pytest

from oracle import cpu_oracle as o
from oracle import np_restatement as npr

GOLDEN = os.path.join(os.path.dirname(__file__), "golden")
KAT = json.load(open(os.path.join(GOLDEN, "kat_colours.json")))
MATRIX_NAMES = ["bt601", "bt709", "bt2020"]
SITE_NAMES = ["none", "h-cosited"]


def flat(rgb, w=8, h=4, in_fmt=o.IN_BGRX, x=0xFF):
    f = np.empty((h, w, 4), np.uint8)
    r, g, b = rgb
    f[..., 3] = x
    if in_fmt == o.IN_BGRX:
        f[..., 0], f[..., 1], f[..., 2] = b, g, r
    else:
        f[..., 0], f[..., 1], f[..., 2] = r, g, b
    return f


def test_coefficients_match_survey_table():
    for m, name in enumerate(MATRIX_NAMES):
        assert o.matrix_coeffs(m).tolist() == KAT["coefficients"][name]
        assert npr.COEFFS[name].tolist() == KAT["coefficients"][name]


@pytest.mark.parametrize("vec", KAT["vectors"], ids=lambda v: "rgb" + "-".join(map(str, v["rgb"])))
def test_flat_colour_known_answers(vec):
    for m in range(3):
        for site in (o.SITE_NONE, o.SITE_H_COSITED):
            for out_fmt in (o.OUT_NV12, o.OUT_I420):
                for in_fmt in (o.IN_BGRX, o.IN_RGBX):
                    out = o.convert(flat(vec["rgb"], in_fmt=in_fmt), in_fmt=in_fmt, out_fmt=out_fmt,
                                    matrix=m, chroma_site=site)
                    y, u, v = o.split_planes(out, 8, 4, out_fmt)
                    ey, eu, ev = vec["yuv"][m]
                    assert (y == ey).all() and (u == eu).all() and (v == ev).all()


def test_bt709_white_fingerprint():
    # U row of BT.709 sums to -1, so white gives U = 127: a GStreamer fingerprint.
    out = o.convert(flat((255, 255, 255)), matrix=o.MATRIX_BT709, chroma_site=o.SITE_NONE)
    _, u, _ = o.split_planes(out, 8, 4, o.OUT_NV12)
    assert (u == 127).all()


def test_padding_byte_is_ignored():
    a = o.convert(flat((12, 200, 99), x=0x00))
    b = o.convert(flat((12, 200, 99), x=0xFF))
    assert (a == b).all()


def _ayuv(u_vals, v_vals=None):
    w = len(u_vals)
    line = np.zeros((w, 4), np.uint8)
    line[:, 2] = u_vals
    line[:, 3] = v_vals if v_vals is not None else u_vals
    return line


def test_chroma_site_none_kat():
    k = KAT["chroma"]["site_none_pair"]
    l0, l1 = _ayuv(k["rows"][0]), _ayuv(k["rows"][1])
    L = o.lib()
    L.sko_chroma_down_h2(o._u8p(l0), 2)
    L.sko_chroma_down_h2(o._u8p(l1), 2)
    assert [int(l0[0, 2]), int(l1[0, 2])] == k["h"]
    L.sko_chroma_down_v2(o._u8p(l0), o._u8p(l1), 2)
    assert int(l0[0, 2]) == k["v"] and int(l0[0, 3]) == k["v"]


def test_chroma_h_cosited_kat():
    k = KAT["chroma"]["h_cosited_w4"]
    l0 = _ayuv(k["row"])
    o.lib().sko_chroma_down_h2_cs(o._u8p(l0), 4)
    assert [int(l0[0, 2]), int(l0[2, 2])] == k["out_even"]
    assert [int(l0[1, 2]), int(l0[3, 2])] == [k["row"][1], k["row"][3]]   # odd samples untouched


def test_cascaded_rounding_is_not_box_filter():
    # (0,1 / 0,0): avg(avg(0,1), avg(0,0)) = avg(1,0) = 1, box (0+1+0+0+2)>>2 = 0  
    f = np.zeros((2, 2, 4), np.uint8)
    # choose two blues whose U differ by exactly 1 at BT.601: U = (112*B + 32768) >> 8
    b0, b1 = 0, 3      # U = 128, 129
    f[0, 0, 0], f[0, 1, 0] = b0, b1
    out = o.convert(f, matrix=o.MATRIX_BT601, chroma_site=o.SITE_NONE)
    _, u, _ = o.split_planes(out, 2, 2, o.OUT_NV12)
    assert int(u[0, 0]) == 129


@pytest.mark.parametrize("w,h", [(2, 2), (4, 2), (6, 4), (64, 34), (130, 18), (258, 50)])
def test_c_oracle_equals_numpy_restatement(w, h):
    f = np.random.default_rng(w * 1000 + h).integers(0, 256, (h, w, 4), dtype=np.uint8)
    for m, name in enumerate(MATRIX_NAMES):
        for site, sname in enumerate(SITE_NAMES):
            for out_fmt in (o.OUT_NV12, o.OUT_I420):
                for in_fmt, iname in ((o.IN_BGRX, "BGRx"), (o.IN_RGBX, "RGBx")):
                    out = o.convert(f, in_fmt=in_fmt, out_fmt=out_fmt, matrix=m, chroma_site=site, n_threads=3)
                    y, u, v = o.split_planes(out, w, h, out_fmt)
                    Y, U, V = npr.convert(f, in_fmt=iname, matrix=name, site=sname)
                    assert (y == Y).all() and (u == U).all() and (v == V).all()


def test_hd_frame_threads_agree_and_match_numpy():
    f = np.random.default_rng(5).integers(0, 256, (1080, 1920, 4), dtype=np.uint8)
    a = o.convert(f, n_threads=1)
    b = o.convert(f, n_threads=4)
    assert (a == b).all()
    y, u, v = o.split_planes(a, 1920, 1080, o.OUT_NV12)
    Y, U, V = npr.convert(f)
    assert (y == Y).all() and (u == U).all() and (v == V).all()


def test_strided_input_rows():
    big = np.random.default_rng(9).integers(0, 256, (20, 40, 4), dtype=np.uint8)
    view = big[2:18, 4:36]          # 32x16 sub-rectangle of a larger framebuffer
    assert (o.convert(view) == o.convert(np.ascontiguousarray(view))).all()


@pytest.mark.parametrize("w,h,fmt", [(1920, 1080, o.OUT_I420), (1920, 1080, o.OUT_NV12), (1366 + 0, 768, o.OUT_I420),
                                      (3840, 2160, o.OUT_NV12), (6, 2, o.OUT_I420)])
def test_output_layout_is_gstvideoinfo_default(w, h, fmt):
    off, st = o.out_layout(w, h, fmt)
    ru = lambda v, a: (v + a - 1) // a * a
    assert st[0] == ru(w, 4) and off[0] == 0 and off[1] == st[0] * ru(h, 2)
    if fmt == o.OUT_I420:
        assert st[1] == st[2] == ru(ru(w, 2) // 2, 4)
        assert off[2] == off[1] + st[1] * ru(h, 2) // 2
        assert o.out_size(w, h, fmt) == off[2] + st[2] * ru(h, 2) // 2
    else:
        assert st[1] == st[0]
        assert o.out_size(w, h, fmt) == off[1] + st[1] * ru(h, 2) // 2
    if w % 8 == 0:
        assert o.out_size(w, h, fmt) == w * h * 3 // 2


# ---- scale ---------------------------------------------------------------
@pytest.mark.parametrize("i,oo", [(1440, 1080), (48, 24), (60, 34), (2160, 1080), (100, 98)])
@pytest.mark.parametrize("variant", [o.SCALE_GST_LQ, o.SCALE_PLAIN_LQ])
def test_vertical_taps_properties(i, oo, variant):
    off, tap = o.resampler_2tap(i, oo, variant)
    assert (off >= 0).all() and (off <= i - 2).all()
    assert (np.diff(off) >= 0).all()
    assert (tap >= 0).all() and (tap <= 64).all()
    # symmetric resampling grid → taps mirror: tap[j] + tap[n-1-j] == 64 where offsets mirror too
    j = np.arange(oo)
    mirror = (off + off[::-1] == i - 2)
    assert ((tap + tap[::-1])[mirror] == 64).all()


def test_h_increment_is_edge_aligned():
    assert o.h_increment(2560, 1920) == ((2559 << 16) // 1919) - 1
    inc = o.h_increment(2560, 1920)
    assert ((1919 * inc) >> 16) + 1 <= 2559     # last tap stays inside the row


def test_half_size_exact_taps():
    off, tap = o.resampler_2tap(48, 24)
    assert (off == np.arange(24) * 2).all() and (tap == 32).all()


@pytest.mark.parametrize("dims", [(2560, 1440, 1920, 1080), (64, 48, 32, 24), (100, 60, 98, 34), (64, 48, 64, 24),
                                  (64, 48, 30, 48), (64, 64, 16, 60)])
def test_scale_c_oracle_equals_numpy_restatement(dims):
    w, h, ow, oh = dims
    f = np.random.default_rng(w + oh).integers(0, 256, (h, w, 4), dtype=np.uint8)
    for var in (o.SCALE_GST_LQ, o.SCALE_PLAIN_LQ):
        off, tap = o.resampler_2tap(h, oh, var)
        out = o.convert(f, ow, oh, scale_variant=var, n_threads=2, out_fmt=o.OUT_I420)
        y, u, v = o.split_planes(out, ow, oh, o.OUT_I420)
        Y, U, V = npr.convert(f, out_w=ow, out_h=oh, v_off=off, v_tap=tap, h_inc=o.h_increment(w, ow))
        assert (y == Y).all() and (u == U).all() and (v == V).all()


def test_scale_flat_colour_is_preserved():
    for vec in KAT["vectors"]:
        out = o.convert(flat(vec["rgb"], 64, 48), 48, 36, matrix=o.MATRIX_BT709)
        y, u, v = o.split_planes(out, 48, 36, o.OUT_NV12)
        ey, eu, ev = vec["yuv"][1]
        assert (y == ey).all() and (u == eu).all() and (v == ev).all()


# Enlarging: upstream's chain_scale() runs behind the matrix (on AYUV) once the picture grows.
ENLARGE_DIMS = [(32, 24, 64, 48), (64, 48, 98, 50), (30, 20, 32, 48), (64, 24, 20, 96), (96, 20, 128, 18),
                (1280, 720, 1920, 1080)]


@pytest.mark.parametrize("dims", ENLARGE_DIMS)
def test_enlarging_c_oracle_equals_numpy_restatement(dims):
    w, h, ow, oh = dims
    assert ow * oh > w * h
    f = np.random.default_rng(w + oh).integers(0, 256, (h, w, 4), dtype=np.uint8)
    for var in (o.SCALE_GST_LQ, o.SCALE_PLAIN_LQ):
        off, tap = o.resampler_2tap(h, oh, var)
        for fmt, site, sname in ((o.OUT_I420, o.SITE_H_COSITED, "h-cosited"), (o.OUT_NV12, o.SITE_NONE, "none")):
            out = o.convert(f, ow, oh, scale_variant=var, n_threads=3, out_fmt=fmt, chroma_site=site)
            y, u, v = o.split_planes(out, ow, oh, fmt)
            Y, U, V = npr.convert(f, out_w=ow, out_h=oh, v_off=off, v_tap=tap, h_inc=o.h_increment(w, ow), site=sname)
            assert (y == Y).all() and (u == U).all() and (v == V).all()


def test_enlarging_flat_colour_is_preserved_and_threads_do_not_matter():
    for vec in KAT["vectors"]:
        out = o.convert(flat(vec["rgb"], 32, 24), 80, 50, matrix=o.MATRIX_BT709)
        y, u, v = o.split_planes(out, 80, 50, o.OUT_NV12)
        ey, eu, ev = vec["yuv"][1]
        assert (y == ey).all() and (u == eu).all() and (v == ev).all()
    f = np.random.default_rng(5).integers(0, 256, (36, 50, 4), dtype=np.uint8)
    assert (o.convert(f, 100, 72, n_threads=1) == o.convert(f, 100, 72, n_threads=5)).all()


def test_enlarging_scales_converted_samples_not_rgb():
    """The two orders differ: a red/blue edge blended in RGB and then converted is not the blend of the two
    converted colours once the matrix rows have been floored separately."""
    f = np.zeros((8, 8, 4), np.uint8)
    f[:, :4] = (0, 0, 255, 255)        # BGRx red
    f[:, 4:] = (255, 0, 0, 255)        # BGRx blue
    off, tap = o.resampler_2tap(8, 16)
    Y, U, V = npr.convert(f, out_w=16, out_h=16, v_off=off, v_tap=tap, h_inc=o.h_increment(8, 16))
    out = o.convert(f, 16, 16)
    y, u, v = o.split_planes(out, 16, 16, o.OUT_NV12)
    assert (y == Y).all() and (u == U).all() and (v == V).all()
    yl, yr = int(y[0, 0]), int(y[0, -1])
    assert all(min(yl, yr) <= int(t) <= max(yl, yr) for t in y[0])       # a blend of the two converted lumas


# ---- damage ----------------------------------------------------------------
def test_damage_first_frame_all_ones_and_static_zero():
    f = np.random.default_rng(1).integers(0, 256, (40, 70, 4), dtype=np.uint8)
    out = o.convert(f)
    m, n = o.damage(70, 40, o.OUT_NV12, out, None)
    assert m.shape == (3, 5) and m.all() and n == 15
    m, n = o.damage(70, 40, o.OUT_NV12, out, out)
    assert not m.any() and n == 0


@pytest.mark.parametrize("fmt", [o.OUT_NV12, o.OUT_I420])
def test_damage_single_pixel_and_chroma_only(fmt):
    w, h = 64, 40      # last MB row is 8 px tall
    f = np.full((h, w, 4), 100, np.uint8)
    prev = o.convert(f, out_fmt=fmt)
    g = f.copy()
    g[39, 63, 2] = 180                       # bottom-right pixel, red channel
    cur = o.convert(g, out_fmt=fmt)
    m, n = o.damage(w, h, fmt, cur, prev)
    assert n == 1 and m[2, 3] == 1
    Y, U, V = o.split_planes(cur, w, h, fmt)
    pY, pU, pV = o.split_planes(prev, w, h, fmt)
    m2 = npr.damage(Y, U, V, pY, pU, pV)
    assert (m2 == m).all()
    # chroma-only change: craft planes directly
    cur2 = prev.copy()
    off, st = o.out_layout(w, h, fmt)
    cur2[off[1] + 9 * st[1] + 5] ^= 1        # chroma row 9 → MB row 1
    m3, n3 = o.damage(w, h, fmt, cur2, prev)
    assert n3 == 1 and m3[1].sum() == 1


def test_damage_matches_numpy_on_random_mutations():
    rng = np.random.default_rng(3)
    w, h = 130, 50
    f = rng.integers(0, 256, (h, w, 4), dtype=np.uint8)
    g = f.copy()
    for _ in range(6):
        x, y = int(rng.integers(0, w)), int(rng.integers(0, h))
        g[y, x, :3] ^= 0xFF
    for fmt in (o.OUT_NV12, o.OUT_I420):
        a, b = o.convert(f, out_fmt=fmt), o.convert(g, out_fmt=fmt)
        m, n = o.damage(w, h, fmt, b, a)
        m2 = npr.damage(*o.split_planes(b, w, h, fmt), *o.split_planes(a, w, h, fmt))
        assert (m == m2).all() and n == int(m2.sum()) and 0 < n <= 12


# ---- frozen fixtures ---------------------------------------------------------
def test_frozen_fixtures_still_reproduce():
    z = np.load(os.path.join(GOLDEN, "frozen_small.npz"))
    for i, (w, h, ow, oh, inf, outf, m, site, var, _seed) in enumerate(z["cases"].tolist()):
        a = o.convert(z[f"in0_{i}"], ow, oh, in_fmt=inf, out_fmt=outf, matrix=m, chroma_site=site, scale_variant=var)
        b = o.convert(z[f"in1_{i}"], ow, oh, in_fmt=inf, out_fmt=outf, matrix=m, chroma_site=site, scale_variant=var)
        assert (a == z[f"out0_{i}"]).all() and (b == z[f"out1_{i}"]).all()
        dm, _ = o.damage(ow, oh, outf, b, a)
        assert (dm == z[f"map1_{i}"]).all()


def test_opencv_sanity_within_two():
    # OpenCV adds a rounding term and box-filters chroma: only a +-2 luma sanity cross-check.
    cv2 = pytest.importorskip("cv2")
    f = np.random.default_rng(11).integers(0, 256, (32, 32, 4), dtype=np.uint8)
    y = o.split_planes(o.convert(f, matrix=o.MATRIX_BT601, chroma_site=o.SITE_NONE, out_fmt=o.OUT_I420), 32, 32, o.OUT_I420)[0]
    ref = cv2.cvtColor(f, cv2.COLOR_BGRA2YUV_I420)[:32, :32]
    assert np.abs(y.astype(int) - ref.astype(int)).max() <= 2


# ---- an anchor outside this repository: the standards' own equations --------------------------------
ITU = {o.MATRIX_BT601: (0.299, 0.114), o.MATRIX_BT709: (0.2126, 0.0722), o.MATRIX_BT2020: (0.2627, 0.0593)}


@pytest.mark.parametrize("matrix", [o.MATRIX_BT601, o.MATRIX_BT709, o.MATRIX_BT2020])
def test_matrix_stays_within_the_itu_equations(matrix):
    """E'Y = Kr R + Kg G + Kb B, Cb = (B - Y) / (2 (1 - Kb)), Cr = (R - Y) / (2 (1 - Kr)), quantised to 16 + 219 Y and
    128 + 224 C (BT.601-7 / BT.709-6 / BT.2020-2, 8 bit).  The integer chain (8 fractional bits per coefficient, sums
    floored, each coefficient off by up to half of 1/256) stays within 2.5 code values below and 1 above the real
    value, and on average half a code value low — a floor, not a round."""
    kr, kb = ITU[matrix]
    kg = 1.0 - kr - kb
    rng = np.random.default_rng(matrix)
    rgb = rng.integers(0, 256, (64, 64, 3)).astype(np.float64)
    rgb[0, :8] = [(0, 0, 0), (255, 255, 255), (255, 0, 0), (0, 255, 0), (0, 0, 255), (255, 255, 0), (0, 255, 255), (255, 0, 255)]
    f = np.empty((64, 64, 4), np.uint8)
    f[..., 0], f[..., 1], f[..., 2], f[..., 3] = rgb[..., 2], rgb[..., 1], rgb[..., 0], 255
    # 4:4:4 view of the oracle: every pixel doubled in both directions, so each chroma sample is one colour's
    big = np.repeat(np.repeat(f, 2, axis=0), 2, axis=1)
    out = o.convert(np.ascontiguousarray(big), matrix=matrix, chroma_site=o.SITE_NONE, out_fmt=o.OUT_I420)
    y, u, v = o.split_planes(out, 128, 128, o.OUT_I420)
    r, g, b = rgb[..., 0] / 255, rgb[..., 1] / 255, rgb[..., 2] / 255
    ey = kr * r + kg * g + kb * b
    exact = (16 + 219 * ey, 128 + 224 * (b - ey) / (2 * (1 - kb)), 128 + 224 * (r - ey) / (2 * (1 - kr)))
    for got, want in zip((y[::2, ::2], u, v), exact):
        d = got.astype(np.float64) - want
        assert d.max() <= 1.0 and d.min() > -2.5, (d.min(), d.max())
        # (BT.709's Cb row -26 -87 +112 sums to -1/256 after per-coefficient rounding: that plane sits a whole code
        #  value low on average, white comes out as U = 127 — SURVEY A.2/A.7, risk R8)
        assert -1.1 < d.mean() < -0.1, d.mean()
