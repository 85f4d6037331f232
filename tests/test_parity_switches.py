"""Every SURVEY.md A.8 risk is a switch carried by the oracle and by the product alike
(sk_oracle.h SKO_PS_*, selkies_preencode.h SK_PS_*, same bit values).

CPU part: each switch changes the C oracle's output on an input built to expose it, and the C
oracle with the switch equals the independent numpy restatement with the switch.
GPU part (-m gpu): the CUDA path with the switch equals the C oracle with the switch, bit for bit,
frames and damage maps — so flipping a default after real GStreamer vectors arrive
(tests/test_gst_golden.py) is a change of one constant on both sides.
"""
import numpy as np
import pytest

from oracle import cpu_oracle as o
from oracle import np_restatement as npr

MATRIX_NAMES = ["bt601", "bt709", "bt2020"]
SITE_NAMES = ["none", "h-cosited"]
SWITCHES = sorted(o.PS_ALL.items(), key=lambda kv: kv[1])


def planes(out, w, h, fmt=o.OUT_I420):
    return o.split_planes(out, w, h, fmt)


def noise(w, h, seed):
    return np.random.default_rng(seed).integers(0, 256, (h, w, 4), dtype=np.uint8)


def test_bits_agree_between_oracle_numpy_and_product():
    from selkies_gstreamer_b200 import _abi
    assert (o.PS_MATRIX_ROUND, o.PS_COEFF_HALF_UP, o.PS_CS_EDGE_KEEP, o.PS_CHROMA_V_FIRST, o.PS_CHROMA_BOX4,
            o.PS_HINC_PLAIN, o.PS_V_FIRST_ON_TIE) == (1, 2, 4, 8, 16, 32, 64)
    assert (_abi.SK_PS_MATRIX_ROUND, _abi.SK_PS_COEFF_HALF_UP, _abi.SK_PS_CS_EDGE_KEEP, _abi.SK_PS_CHROMA_V_FIRST,
            _abi.SK_PS_CHROMA_BOX4, _abi.SK_PS_HINC_PLAIN, _abi.SK_PS_V_FIRST_ON_TIE) == (1, 2, 4, 8, 16, 32, 64)
    assert (npr.PS_MATRIX_ROUND, npr.PS_CS_EDGE_KEEP, npr.PS_CHROMA_V_FIRST, npr.PS_CHROMA_BOX4) == (1, 4, 8, 16)
    hdr = open(__import__("os").path.join(__import__("os").path.dirname(__file__), "..", "include", "selkies_preencode.h")).read()
    for name in ("SK_PS_MATRIX_ROUND", "SK_PS_COEFF_HALF_UP", "SK_PS_CS_EDGE_KEEP", "SK_PS_CHROMA_V_FIRST",
                 "SK_PS_CHROMA_BOX4", "SK_PS_HINC_PLAIN", "SK_PS_V_FIRST_ON_TIE"):
        assert name in hdr


@pytest.mark.parametrize("name,bit", [kv for kv in SWITCHES if kv[1] in (1, 4, 8, 16)])
@pytest.mark.parametrize("site", [0, 1])
def test_pixel_switches_c_oracle_equals_numpy(name, bit, site):
    for (w, h) in [(2, 2), (4, 2), (6, 4), (66, 34), (130, 18)]:
        f = noise(w, h, w * 100 + h + bit)
        for m, mname in enumerate(MATRIX_NAMES):
            out = o.convert(f, out_fmt=o.OUT_I420, matrix=m, chroma_site=site, switches=bit, n_threads=2)
            y, u, v = planes(out, w, h)
            Y, U, V = npr.convert(f, matrix=mname, site=SITE_NAMES[site], switches=bit)
            assert (y == Y).all() and (u == U).all() and (v == V).all(), (name, w, h, mname)


def test_each_pixel_switch_changes_something_it_should():
    f = noise(66, 34, 7)
    base = {s: o.convert(f, out_fmt=o.OUT_I420, chroma_site=s) for s in (0, 1)}
    # R8 rounding term: luma and chroma move
    a = o.convert(f, out_fmt=o.OUT_I420, chroma_site=1, switches=o.PS_MATRIX_ROUND)
    y0, u0, _ = planes(base[1], 66, 34); y1, u1, _ = planes(a, 66, 34)
    assert (y1 >= y0).all() and (y1 != y0).any() and ((y1.astype(int) - y0) <= 1).all()
    # R3 edge taps: only the first and last chroma column may differ, luma never
    a = o.convert(f, out_fmt=o.OUT_I420, chroma_site=1, switches=o.PS_CS_EDGE_KEEP)
    y1, u1, v1 = planes(a, 66, 34)
    assert (y1 == y0).all() and (u1[:, 1:-1] == u0[:, 1:-1]).all() and ((u1[:, 0] != u0[:, 0]).any() or (u1[:, -1] != u0[:, -1]).any())
    # ... and it does nothing for site none
    assert (o.convert(f, out_fmt=o.OUT_I420, chroma_site=0, switches=o.PS_CS_EDGE_KEEP) == base[0]).all()
    # R4 order / single rounding: chroma differs by at most 1, luma never
    for bit, site in ((o.PS_CHROMA_V_FIRST, 0), (o.PS_CHROMA_V_FIRST, 1), (o.PS_CHROMA_BOX4, 0)):
        a = o.convert(f, out_fmt=o.OUT_I420, chroma_site=site, switches=bit)
        yb, ub, vb = planes(base[site], 66, 34); y1, u1, v1 = planes(a, 66, 34)
        d = np.abs(u1.astype(int) - ub)
        assert (y1 == yb).all() and d.max() == 1, (bit, site)
    # BOX4 is defined for site none only
    assert (o.convert(f, out_fmt=o.OUT_I420, chroma_site=1, switches=o.PS_CHROMA_BOX4) == base[1]).all()


def test_coefficient_rounding_switch_is_inert_for_the_three_matrices():
    """No coefficient of BT.601/709/2020 sits within 0.4 of a tie, so rint vs half-up cannot differ."""
    for m in range(3):
        assert (o.matrix_coeffs(m, o.PS_COEFF_HALF_UP) == o.matrix_coeffs(m)).all()
        c = o.matrix_coeffs(m, o.PS_MATRIX_ROUND)
        assert (c[:, 3] == o.matrix_coeffs(m)[:, 3] + 128).all() and (c[:, :3] == o.matrix_coeffs(m)[:, :3]).all()


def test_scale_switches_oracle_equals_numpy():
    # R6: increment without the "-1"
    assert o.h_increment(2560, 1920, o.PS_HINC_PLAIN) == o.h_increment(2560, 1920) + 1
    w, h, ow, oh = 100, 60, 70, 42
    f = noise(w, h, 3)
    off, tap = o.resampler_2tap(h, oh)
    out = o.convert(f, ow, oh, out_fmt=o.OUT_I420, switches=o.PS_HINC_PLAIN)
    Y, U, V = npr.convert(f, out_w=ow, out_h=oh, v_off=off, v_tap=tap, h_inc=o.h_increment(w, ow, o.PS_HINC_PLAIN))
    y, u, v = planes(out, ow, oh)
    assert (y == Y).all() and (u == U).all() and (v == V).all()
    assert (out != o.convert(f, ow, oh, out_fmt=o.OUT_I420)).any()
    # R7: 2560x1440 -> 1920x1080 costs the same either way; the switch picks vertical first
    w, h, ow, oh = 256, 144, 192, 108
    f = noise(w, h, 4)
    off, tap = o.resampler_2tap(h, oh)
    out = o.convert(f, ow, oh, out_fmt=o.OUT_I420, switches=o.PS_V_FIRST_ON_TIE)
    Y, U, V = npr.convert(f, out_w=ow, out_h=oh, v_off=off, v_tap=tap, h_inc=o.h_increment(w, ow), switches=npr.PS_V_FIRST_ON_TIE)
    y, u, v = planes(out, ow, oh)
    assert (y == Y).all() and (u == U).all() and (v == V).all()
    assert (out != o.convert(f, ow, oh, out_fmt=o.OUT_I420)).any()
    # no tie: the switch is inert
    assert (o.convert(f, 128, 108, switches=o.PS_V_FIRST_ON_TIE) == o.convert(f, 128, 108)).all()


def test_default_colorimetry_by_height_is_a_parameter():
    """R1: the >= 2160 -> bt2020 branch of upstream's default colorimetry is a host-side choice."""
    from selkies_gstreamer_b200 import colorimetry as col
    assert col.resolve(2160, None, None)[0] == 2
    assert col.resolve(2160, None, None, uhd_default="bt709")[0] == 1
    assert col.resolve(1080, None, None, uhd_default="bt709")[0] == 1
    assert col.resolve(480, None, None)[0] == 0


# ------------------------------------------------------------------------------ GPU
@pytest.mark.gpu
@pytest.mark.parametrize("name,bit", SWITCHES + [("all-pixel-switches", 1 | 4 | 8), ("round+box4", 1 | 16)])
def test_cuda_path_with_switch_equals_oracle_with_switch(name, bit):
    from selkies_gstreamer_b200 import PreencodeSession, synth
    cases = [(258, 50, None, None), (1280, 720, None, None), (256, 144, 192, 108), (200, 120, 140, 84)]
    for (w, h, ow, oh) in cases:
        f0 = noise(w, h, w + bit)
        f1 = synth.mutate(f0, 0.05, seed=bit, aligned=False)
        for fmt, site in (("NV12", 1), ("I420", 0)):
            ofmt = {"NV12": 0, "I420": 1}[fmt]
            with PreencodeSession(w, h, out_width=ow, out_height=oh, out_format=fmt, colorimetry=1, chroma_site=site,
                                  parity_switches=bit) as s:
                prev = None
                for i, f in enumerate((f0, f1)):
                    r = s.convert(f, slot=i)
                    e = o.convert(f, ow, oh, out_fmt=ofmt, matrix=1, chroma_site=site, switches=bit, n_threads=4)
                    assert (r.frame == e).all(), (name, w, h, ow, oh, fmt, i)
                    em, en = o.damage(ow or w, oh or h, ofmt, e, prev)
                    assert (r.damage_map == em).all() and r.damaged_count == en
                    prev = e
