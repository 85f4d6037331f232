"""GPU parity tests: the CUDA path, called through the C ABI (ctypes →
libselkies_preencode.so), must be bit-exact with the CPU oracle on the same
inputs, and exactly equal on the damage map.  Run on the B200 box:
    python -m pytest tests -m gpu -x -q
"""
import ctypes
import json
import os

import numpy as np
import pytest

from oracle import cpu_oracle as o
from selkies_gstreamer_b200 import (PreencodeError, PreencodeSession, SessionBatch, _abi, batch_submit, launch_count,
                                   synth)

pytestmark = pytest.mark.gpu

GOLDEN = os.path.join(os.path.dirname(__file__), "golden")
IN_NAMES = {0: "BGRx", 1: "RGBx"}
OUT_NAMES = {0: "NV12", 1: "I420"}


def oracle_frame(frame, sess_kw, out_w=None, out_h=None, variant=0):
    """The oracle's answer for a session created with `sess_kw` (kernel-selection keywords do not matter)."""
    return o.convert(frame, out_w, out_h,
                     in_fmt={"BGRx": 0, "RGBx": 1, "BGRA": 0, "RGBA": 1}[sess_kw.get("in_format", "BGRx")],
                     out_fmt={"NV12": 0, "I420": 1}[sess_kw.get("out_format", "NV12")],
                     matrix={"bt601": 0, "bt709": 1, "bt2020": 2}.get(sess_kw["colorimetry"], sess_kw["colorimetry"]),
                     chroma_site=sess_kw["chroma_site"],
                     scale_variant=variant, n_threads=8)


def check_sequence(frames, w, h, *, out_w=None, out_h=None, variant=0, n_inflight=2, **kw):
    """Run frames through one session; compare every output and map with the oracle."""
    kw.setdefault("colorimetry", 1)
    kw.setdefault("chroma_site", 1)
    ow, oh = out_w or w, out_h or h
    fmt = {"NV12": 0, "I420": 1}[kw.get("out_format", "NV12")]
    with PreencodeSession(w, h, out_width=out_w, out_height=out_h, n_inflight=n_inflight,
                          scale_variant=variant, **kw) as s:
        prev = None
        for i, f in enumerate(frames):
            r = s.convert(f, slot=i % n_inflight)
            exp = oracle_frame(f, kw, out_w, out_h, variant)
            assert r.frame.shape == exp.shape
            if not (r.frame == exp).all():
                bad = np.flatnonzero(r.frame != exp)
                raise AssertionError(f"frame {i}: {bad.size} bytes differ, first at {bad[:8]} "
                                     f"got {r.frame[bad[:8]]} want {exp[bad[:8]]}")
            em, en = o.damage(ow, oh, fmt, exp, prev)
            assert (r.damage_map == em).all(), f"frame {i}: damage map differs"
            assert r.damaged_count == en
            prev = exp


def rand_frame(w, h, seed):
    return np.random.default_rng(seed).integers(0, 256, (h, w, 4), dtype=np.uint8)


# --------------------------------------------------------------------------
def test_known_answer_colours_on_gpu():
    kat = json.load(open(os.path.join(GOLDEN, "kat_colours.json")))
    for m in range(3):
        for site in (0, 1):
            with PreencodeSession(16, 4, colorimetry=m, chroma_site=site, out_format="I420") as s:
                for vec in kat["vectors"]:
                    f = np.empty((4, 16, 4), np.uint8)
                    r, g, b = vec["rgb"]
                    f[..., 0], f[..., 1], f[..., 2], f[..., 3] = b, g, r, 0x5A
                    res = s.convert(f)
                    y, u, v = s.planes(res.frame)
                    ey, eu, ev = vec["yuv"][m]
                    assert (y == ey).all() and (u == eu).all() and (v == ev).all(), (vec, m, site)


def test_frozen_fixtures_on_gpu():
    z = np.load(os.path.join(GOLDEN, "frozen_small.npz"))
    for i, (w, h, ow, oh, inf, outf, m, site, var, _seed) in enumerate(z["cases"].tolist()):
        with PreencodeSession(w, h, out_width=ow, out_height=oh, in_format=IN_NAMES[inf], out_format=OUT_NAMES[outf],
                              colorimetry=m, chroma_site=site, scale_variant=var) as s:
            a = s.convert(z[f"in0_{i}"], 0)
            assert (a.frame == z[f"out0_{i}"]).all(), f"case {i} frame 0"
            assert a.damage_map.all()
            b = s.convert(z[f"in1_{i}"], 1)
            assert (b.frame == z[f"out1_{i}"]).all(), f"case {i} frame 1"
            assert (b.damage_map == z[f"map1_{i}"]).all(), f"case {i} map"
            assert b.damaged_count == int(z[f"map1_{i}"].sum())


@pytest.mark.parametrize("w,h", [(2, 2), (4, 2), (6, 4), (8, 16), (34, 18), (126, 30), (128, 16), (130, 34),
                                 (254, 24), (258, 50), (640, 360), (1366, 768)])
def test_small_and_ragged_sizes_all_modes(w, h):
    f0 = rand_frame(w, h, w * 7 + h)
    f1 = f0.copy()
    f1[h // 2, w // 2, :3] ^= 0x3C
    for m in range(3):
        for site in (0, 1):
            for out_format in ("NV12", "I420"):
                in_format = "RGBx" if (m + site) % 2 else "BGRx"
                check_sequence([f0, f1, f1], w, h, colorimetry=m, chroma_site=site, out_format=out_format,
                               in_format=in_format)


@pytest.mark.parametrize("variant", range(16))
def test_every_kernel_variant_is_bit_exact(variant):
    """Tuning build only (make -C selkies_gstreamer_b200/csrc tuning; SELKIES_PREENCODE_LIB=...tuning.so):
    the release library carries the default shape alone."""
    lib = _abi.load()
    if not hasattr(lib, "sk_tuning_set_variant"):
        if variant:
            pytest.skip("kernel-shape variants are compiled into the tuning build only")
        lib = None
    try:
        assert lib is None or lib.sk_tuning_set_variant(variant) == 0
        for (w, h) in [(258, 50), (1920, 1080), (1366, 768)]:
            f0 = rand_frame(w, h, 100 + variant)
            f1 = synth.mutate(f0, 0.05, seed=variant, aligned=False)
            for out_format in ("NV12", "I420"):
                for site in (0, 1):
                    check_sequence([f0, f1], w, h, out_format=out_format, chroma_site=site)
    finally:
        if lib is not None:
            lib.sk_tuning_set_variant(0)


@pytest.mark.parametrize("name,w,h,fmt", [("1080p-I420", 1920, 1080, "I420"), ("4K-NV12", 3840, 2160, "NV12"),
                                          ("8K-I420", 7680, 4320, "I420")])
def test_baseline_configs_full_size(name, w, h, fmt):
    f0 = synth.desktop(w, h)
    f1 = synth.mutate(f0, 0.005, seed=3, aligned=False)     # sparse damage (config 5)
    f2 = synth.noise(w, h)
    check_sequence([f0, f1, f1, f2], w, h, out_format=fmt)


def test_default_colorimetry_by_height():
    # upstream set_default_colorimetry: <=576 bt601/none, <2160 bt709/h-cosited, >=2160 bt2020/h-cosited
    for (w, h, m, site) in [(720, 576, 0, 0), (1280, 720, 1, 1), (3840, 2160, 2, 1)]:
        f = rand_frame(w, h, h)
        with PreencodeSession(w, h) as s:
            assert (s.matrix, s.site) == (m, site)
            r = s.convert(f)
            assert (r.frame == o.convert(f, matrix=m, chroma_site=site, n_threads=8)).all()


def test_bars_and_checkerboard_exercise_chroma_filters():
    f = synth.bars(1920, 1080)
    for site in (0, 1):
        check_sequence([f], 1920, 1080, chroma_site=site, out_format="I420")


def test_padding_byte_does_not_matter():
    f = rand_frame(256, 64, 1)
    g = f.copy()
    g[..., 3] = 0
    with PreencodeSession(256, 64) as s:
        assert (s.matrix, s.site) == (0, 0)      # SD default: bt601, chroma-site none
        a = s.convert(f, 0)
        fa = a.frame.copy()
        b = s.convert(g, 1)
        assert (fa == b.frame).all()
        assert b.damaged_count == 0 and not b.damage_map.any()


def test_strided_capture_rows():
    big = rand_frame(300, 80, 2)
    view = big[5:69, 10:266]                  # 256x64 region of a larger framebuffer
    with PreencodeSession(256, 64, in_stride=4 * 256 + 40) as s:     # staging stride not a multiple of 16
        st = s.staging(0)
        assert st.shape == (64, 4 * 256 + 40)
        s.upload(0, view)
        r = s.wait(s.submit(0))
        assert (r.frame == o.convert(np.ascontiguousarray(view), matrix=s.matrix, chroma_site=s.site, n_threads=2)).all()


def test_submit_host_pageable_strided_and_registered():
    from selkies_gstreamer_b200 import host_register, host_unregister
    big = rand_frame(700, 400, 21)
    view = big[20:380, 30:670]                                  # 640x360 window of a larger framebuffer (pageable)
    f2 = np.ascontiguousarray(synth.mutate(np.ascontiguousarray(view), 0.03, seed=5, aligned=False))
    with PreencodeSession(640, 360, colorimetry=1, chroma_site=1) as s:
        r0 = s.wait(s.submit_host(0, view))
        e0 = o.convert(np.ascontiguousarray(view), n_threads=2)
        assert (r0.frame == e0).all() and r0.damage_map.all()
        host_register(f2)
        try:
            r1 = s.wait(s.submit_host(1, f2))
            e1 = o.convert(f2, n_threads=2)
            em, en = o.damage(640, 360, 0, e1, e0)
            assert (r1.frame == e1).all() and (r1.damage_map == em).all() and r1.damaged_count == en
        finally:
            host_unregister(f2)
        with pytest.raises(ValueError):
            s.submit_host(0, view[:, :, ::-1])


def test_idempotence_and_reset_history():
    f = synth.desktop(1920, 1080)
    with PreencodeSession(1920, 1080, n_inflight=3) as s:
        r0 = s.convert(f, 0)
        assert r0.damage_map.all() and r0.damaged_count == 120 * 68
        r1 = s.convert(f, 1)
        assert r1.damaged_count == 0 and not r1.damage_map.any()
        s.reset_history()
        r2 = s.convert(f, 2)
        assert r2.damage_map.all() and r2.damaged_count == 120 * 68
        r3 = s.convert(f, 0)
        assert r3.damaged_count == 0


def test_frames_in_flight_ring():
    w, h, n = 640, 360, 4
    frames = [rand_frame(w, h, 50)]
    for i in range(7):
        frames.append(synth.mutate(frames[-1], 0.02 * (i + 1), seed=i, aligned=False))
    exp = [o.convert(f, n_threads=4) for f in frames]
    with PreencodeSession(w, h, n_inflight=n, colorimetry=1, chroma_site=1) as s:
        got = []
        for base in range(0, len(frames), n):
            ids = []
            for k in range(n):
                s.upload(k, frames[base + k])
                ids.append(s.submit(k))
            for fid in ids:
                r = s.wait(fid)
                got.append((r.frame.copy(), r.damage_map.copy(), r.damaged_count))
    prev = None
    for i, (fr, dm, cnt) in enumerate(got):
        assert (fr == exp[i]).all(), i
        em, en = o.damage(w, h, 0, exp[i], prev)
        assert (dm == em).all() and cnt == en, i
        prev = exp[i]


def test_convert_only_mode():
    f = rand_frame(640, 360, 8)
    with PreencodeSession(640, 360, damage=False, colorimetry=0, chroma_site=0) as s:
        r = s.convert(f)
        assert r.damage_map is None
        assert (r.frame == o.convert(f, matrix=0, chroma_site=0)).all()


# ---- scale -----------------------------------------------------------------
@pytest.mark.parametrize("dims", [(2560, 1440, 1920, 1080), (64, 48, 32, 24), (100, 60, 98, 34), (64, 48, 64, 24),
                                  (64, 48, 30, 48), (64, 64, 16, 60), (1920, 1080, 1280, 720), (3840, 2160, 1920, 1080)])
@pytest.mark.parametrize("variant", [0, 1])
@pytest.mark.parametrize("kernel", ["tiled", "generic"])
def test_scale_parity(dims, variant, kernel):
    """Both scale kernels (shared-memory tiled fast path, generic gather) against the oracle."""
    w, h, ow, oh = dims
    f0 = rand_frame(w, h, w + oh + variant)
    f1 = synth.mutate(f0, 0.03, seed=variant, aligned=False)
    for out_format, site in (("I420", 1), ("NV12", 0)):
        check_sequence([f0, f1], w, h, out_w=ow, out_h=oh, variant=variant, out_format=out_format, chroma_site=site,
                       scale_generic=(kernel == "generic"))


@pytest.mark.parametrize("dims", [(128, 64, 64, 96), (64, 128, 96, 64), (200, 100, 100, 150), (100, 200, 160, 100)])
@pytest.mark.parametrize("kernel", ["tiled", "generic"])
def test_scale_one_axis_enlarging_with_fewer_pixels(dims, kernel):
    """Fewer output than input pixels, but one axis grows: still the scale-before-matrix chain."""
    w, h, ow, oh = dims
    f0 = rand_frame(w, h, w + oh)
    f1 = synth.mutate(f0, 0.1, seed=3, aligned=False)
    for out_format, site in (("I420", 1), ("NV12", 0)):
        check_sequence([f0, f1], w, h, out_w=ow, out_h=oh, out_format=out_format, chroma_site=site,
                       scale_generic=(kernel == "generic"))


@pytest.mark.parametrize("layout", ["tile16", "separate-buffer", "tile16+separate-buffer"])
@pytest.mark.parametrize("dims", [(2560, 1440, 1920, 1080), (1366, 768, 1280, 720), (258, 130, 130, 66), (640, 360, 320, 360)])
def test_scale_kernel_layouts_that_the_default_choice_skips(dims, layout, monkeypatch):
    """The session picks 32-row tiles and in-place h-scaled rows when they fit; the 16-row tiles and the
    separate shared-memory buffer stay reachable (other geometries use them) and must be just as exact."""
    if "tile16" in layout:
        monkeypatch.setenv("SELKIES_PREENCODE_SCALE_TILE", "16")
    if "separate" in layout:
        monkeypatch.setenv("SELKIES_PREENCODE_SCALE_SEPARATE", "1")
    w, h, ow, oh = dims
    f0 = rand_frame(w, h, w + oh)
    f1 = synth.mutate(f0, 0.04, seed=5, aligned=False)
    for out_format, site in (("I420", 1), ("NV12", 0)):
        check_sequence([f0, f1], w, h, out_w=ow, out_h=oh, out_format=out_format, chroma_site=site)


@pytest.mark.parametrize("dims", [(1366, 768, 1280, 720), (258, 130, 130, 66), (2560, 1440, 1278, 720), (514, 96, 258, 90)])
def test_scale_ragged_edges(dims):
    w, h, ow, oh = dims
    f0 = rand_frame(w, h, w * 3 + oh)
    f1 = synth.mutate(f0, 0.05, seed=2, aligned=False)
    for out_format, site in (("NV12", 1), ("I420", 0)):
        check_sequence([f0, f1, f1], w, h, out_w=ow, out_h=oh, out_format=out_format, chroma_site=site)


def test_config3_desktop_content_bt709():
    f0 = synth.desktop(2560, 1440)
    f1 = synth.mutate(f0, 0.005, seed=1, aligned=False)
    check_sequence([f0, f1, f1], 2560, 1440, out_w=1920, out_h=1080, out_format="I420", colorimetry="bt709")


# Enlarging (more output than input pixels): upstream scales behind the matrix, on the converted samples.
@pytest.mark.parametrize("dims", [(32, 24, 64, 48), (64, 48, 98, 50), (30, 20, 32, 48), (64, 24, 20, 96),
                                  (96, 20, 128, 18), (130, 66, 258, 130), (640, 360, 1280, 720)])
@pytest.mark.parametrize("variant", [0, 1])
@pytest.mark.parametrize("kernel", ["tiled", "generic"])
def test_enlarging_parity(dims, variant, kernel):
    """tiled = whatever the session picks (the tiled kernel when both axes change and the window fits)."""
    w, h, ow, oh = dims
    assert ow * oh > w * h
    f0 = rand_frame(w, h, w + oh + variant)
    f1 = synth.mutate(f0, 0.05, seed=variant, aligned=False)
    for out_format, site, col in (("I420", 1, 1), ("NV12", 0, 0), ("NV12", 1, 2)):
        check_sequence([f0, f1, f1], w, h, out_w=ow, out_h=oh, variant=variant, out_format=out_format, chroma_site=site,
                       colorimetry=col, scale_generic=(kernel == "generic"))


@pytest.mark.parametrize("dims", [(1280, 720, 1920, 1080), (1366, 768, 1920, 1080), (258, 130, 386, 258), (640, 360, 642, 362)])
def test_enlarging_ragged_edges_tiled(dims):
    w, h, ow, oh = dims
    f0 = rand_frame(w, h, w * 3 + oh)
    f1 = synth.mutate(f0, 0.05, seed=2, aligned=False)
    for out_format, site in (("NV12", 1), ("I420", 0)):
        check_sequence([f0, f1, f1], w, h, out_w=ow, out_h=oh, out_format=out_format, chroma_site=site)


def test_enlarging_1080p_to_4k_rgbx_desktop_content():
    f0 = synth.desktop(1920, 1080)[..., [2, 1, 0, 3]].copy()
    f1 = synth.mutate(f0, 0.01, seed=4, aligned=False)
    check_sequence([f0, f1], 1920, 1080, out_w=3840, out_h=2160, out_format="NV12", in_format="RGBx")


def test_enlarging_through_the_switch_kernel_and_in_a_batch():
    """SK_PS_MATRIX_ROUND moves the matrix offsets: behind-the-matrix scaling must honour it per source pixel;
    and three sessions of one launch must each get their own frame."""
    from oracle import cpu_oracle as oo
    w, h, ow, oh = 96, 54, 160, 90
    f = rand_frame(w, h, 77)
    with PreencodeSession(w, h, out_width=ow, out_height=oh, out_format="I420", colorimetry=1, chroma_site=1,
                          parity_switches=oo.PS_MATRIX_ROUND) as s:
        exp = o.convert(f, ow, oh, out_fmt=1, matrix=1, chroma_site=1, switches=oo.PS_MATRIX_ROUND)
        assert (s.convert(f).frame == exp).all()
    frames = [rand_frame(w, h, 100 + i) for i in range(3)]
    nexts = [synth.mutate(fr, 0.1, seed=i, aligned=False) for i, fr in enumerate(frames)]
    sess = [PreencodeSession(w, h, out_width=ow, out_height=oh, colorimetry=1, chroma_site=1) for _ in frames]
    try:
        for s_, fr, nx in zip(sess, frames, nexts):          # a first pass each brings both inputs into HBM
            s_.upload(0, fr); s_.wait(s_.submit(0))
            s_.upload(1, nx); s_.wait(s_.submit(1))
            s_.reset_history()
        b = SessionBatch(sess)
        stride = sess[0].device_input(0)[1]
        prev = [None] * len(sess)
        for fid, (slot, ins) in enumerate(((0, frames), (1, nexts))):
            b.submit(b.frame_set([s_.device_input(slot)[0] for s_ in sess]), stride, 10 + fid)
            b.wait(10 + fid)
            for i, (s_, fr) in enumerate(zip(sess, ins)):
                r = s_.wait(10 + fid)
                exp = o.convert(fr, ow, oh, matrix=1, chroma_site=1)
                assert (r.frame == exp).all()
                em, en = o.damage(ow, oh, 0, exp, prev[i])
                assert (r.damage_map == em).all() and r.damaged_count == en
                prev[i] = exp
    finally:
        for s_ in sess:
            s_.close()


# ---- device-resident and batched submission -------------------------------------
def test_submit_device_and_alignment_check():
    w, h = 1280, 720
    f0, f1 = rand_frame(w, h, 1), rand_frame(w, h, 2)
    with PreencodeSession(w, h, n_inflight=2, colorimetry=1, chroma_site=1) as s:
        s.upload(0, f0); s.wait(s.submit(0))
        s.upload(1, f1); s.wait(s.submit(1))
        d0, st = s.device_input(0)
        d1, _ = s.device_input(1)
        before = launch_count()
        r = s.wait(s.submit_device(d0, st))
        assert launch_count() == before + 1
        e0, e1 = o.convert(f0, n_threads=4), o.convert(f1, n_threads=4)
        assert (r.frame == e0).all()
        em, en = o.damage(w, h, 0, e0, e1)
        assert (r.damage_map == em).all() and r.damaged_count == en
        with pytest.raises(PreencodeError) as e:
            s.submit_device(d1 + 4, st)
        assert e.value.code == _abi.SK_ERR_ALIGN


@pytest.mark.parametrize("fmt", ["NV12", "I420"])
def test_batch_submit_matches_oracle_per_session(fmt):
    w, h, n = 1920, 1080, 5
    sessions = [PreencodeSession(w, h, out_format=fmt, colorimetry=1, chroma_site=1) for _ in range(n)]
    try:
        f0 = [rand_frame(w, h, 10 + i) for i in range(n)]
        f1 = [synth.mutate(f, 0.01 * (i + 1), seed=i, aligned=False) for i, f in enumerate(f0)]
        for s, a, b in zip(sessions, f0, f1):
            s.upload(0, a); s.wait(s.submit(0))
            s.upload(1, b); s.wait(s.submit(1))
            s.reset_history()
        ptr0 = [s.device_input(0)[0] for s in sessions]
        ptr1 = [s.device_input(1)[0] for s in sessions]
        stride = sessions[0].device_input(0)[1]
        before = launch_count()
        batch_submit(sessions, ptr0, stride, frame_id=100)
        batch_submit(sessions, ptr1, stride, frame_id=101)
        assert launch_count() == before + 2
        ofmt = {"NV12": 0, "I420": 1}[fmt]
        for i, s in enumerate(sessions):
            e0 = o.convert(f0[i], out_fmt=ofmt, n_threads=4)
            e1 = o.convert(f1[i], out_fmt=ofmt, n_threads=4)
            r0 = s.wait(100)
            assert (r0.frame == e0).all() and r0.damage_map.all()
            r1 = s.wait(101)
            em, en = o.damage(w, h, ofmt, e1, e0)
            assert (r1.frame == e1).all() and (r1.damage_map == em).all() and r1.damaged_count == en
    finally:
        for s in sessions:
            s.close()


def test_batch_rejects_mixed_configuration():
    a = PreencodeSession(640, 360)
    b = PreencodeSession(640, 360, out_format="I420")
    try:
        a.upload(0, rand_frame(640, 360, 1)); a.wait(a.submit(0))
        b.upload(0, rand_frame(640, 360, 2)); b.wait(b.submit(0))
        with pytest.raises(PreencodeError) as e:
            batch_submit([a, b], [a.device_input(0)[0], b.device_input(0)[0]], 0, 5)
        assert e.value.code == _abi.SK_ERR_STATE
    finally:
        a.close(); b.close()


def test_policy_map_is_absent_unless_asked_for():
    with PreencodeSession(64, 64) as s:
        assert s.convert(rand_frame(64, 64, 1)).policy_map is None


def test_timing_fields_are_filled():
    f = rand_frame(1920, 1080, 4)
    with PreencodeSession(1920, 1080, timing=True) as s:
        s.convert(f, 0)
        r = s.convert(f, 1)
        assert r.kernel_us > 0 and r.h2d_us > 0 and r.d2h_us > 0
        assert s.last_kernel_us() == pytest.approx(r.kernel_us)


def test_unknown_frame_id():
    with PreencodeSession(64, 64) as s:
        with pytest.raises(PreencodeError) as e:
            s.wait(12345)
        assert e.value.code == _abi.SK_ERR_STATE


def test_linearity_property_on_luma_at_4k():
    # size-independent property: adding k to G of every pixel raises the Y accumulator by 157k (BT.709), so
    # Y changes by floor((acc+157k)/256)-floor(acc/256) ∈ {floor(157k/256), +1}; chroma of a grey ramp is 128/127.
    w, h = 3840, 2160
    f = np.zeros((h, w, 4), np.uint8)
    ramp = (np.arange(w) % 200).astype(np.uint8)
    f[..., 0] = f[..., 1] = f[..., 2] = ramp[None, :]
    g = f.copy()
    g[..., 1] += 32
    with PreencodeSession(w, h, colorimetry=1, chroma_site=1, host_output=True) as s:
        ya = s.planes(s.convert(f, 0).frame)[0].astype(int)
        rb = s.convert(g, 1)
        yb = s.planes(rb.frame)[0].astype(int)
        d = yb - ya
        assert d.min() >= (157 * 32) // 256 and d.max() <= (157 * 32) // 256 + 1
        assert rb.damaged_count == 240 * 135


# ---- the element logic on the GPU ---------------------------------------------------
def test_stage_transform_and_renegotiation():
    from selkies_gstreamer_b200 import stage as st
    seen = []
    s = st.PreEncodeStage()
    s.set_property("n-threads", 4)
    s.set_property("qos", True)
    s.on_damage = lambda stage, m, n: seen.append(n)
    sink = st.parse_caps("video/x-raw, format=(string)BGRx, width=(int)1280, height=(int)720, framerate=(fraction)60/1")
    other = s.transform_caps(st.SINK, sink, st.parse_caps("video/x-raw, format=(string)I420"))
    fixed = s.fixate_caps(st.SINK, sink, other)
    assert s.start() and s.set_caps(sink[0], fixed), s.last_error
    f0 = synth.desktop(1280, 720)
    f1 = synth.mutate(f0, 0.02, seed=4, aligned=False)
    out0, m0, n0 = s.transform_array(f0)
    out1, m1, n1 = s.transform_array(f1)
    e0 = o.convert(f0, out_fmt=o.OUT_I420, matrix=1, chroma_site=1)      # 720p default: bt709 + h-cosited
    e1 = o.convert(f1, out_fmt=o.OUT_I420, matrix=1, chroma_site=1)
    assert (out0 == e0).all() and (out1 == e1).all()
    em, en = o.damage(1280, 720, o.OUT_I420, e1, e0)
    assert m0.all() and (m1 == em).all() and n1 == en and seen == [n0, n1]
    assert s.get_property("damaged-blocks") == en and s.get_property("frames") == 2
    assert s.src_caps_fields() == {"colorimetry": "bt709", "chroma-site": "mpeg2"}
    # framerate-only renegotiation keeps the session and the history (set_framerate, gstwebrtc_app.py:1264-1266)
    sink2 = dict(sink[0], framerate=30)
    assert s.set_caps(sink2, fixed)
    _, _, n2 = s.transform_array(f1)
    assert n2 == 0
    # a resize recreates it: next frame has no predecessor
    small_in = dict(sink[0], width=640, height=360)
    small_out = dict(fixed, width=640, height=360)
    assert s.set_caps(small_in, small_out)
    _, m3, n3 = s.transform_array(synth.desktop(640, 360))
    assert m3.all() and n3 == 40 * 23
    # buffers that are too small are flow errors, not crashes
    assert s.transform(f0.ctypes.data, 16, out0.ctypes.data, out0.nbytes) == st.FLOW_ERROR
    assert s.stop()


def test_stage_scales_both_ways_when_the_downstream_caps_ask_for_it():
    """A capsfilter behind the stage that fixes another size (videoconvertscale's job): smaller, then larger."""
    from selkies_gstreamer_b200 import stage as st
    f = synth.desktop(1280, 720)
    sink = st.parse_caps("video/x-raw, format=(string)BGRx, width=(int)1280, height=(int)720, framerate=(fraction)60/1")
    for ow, oh in ((854, 480), (1920, 1080)):
        s = st.PreEncodeStage()
        filt = st.parse_caps(f"video/x-raw, format=(string)NV12, width=(int){ow}, height=(int){oh}")
        other = s.transform_caps(st.SINK, sink, filt)
        fixed = s.fixate_caps(st.SINK, sink, other)
        assert (fixed["width"], fixed["height"]) == (ow, oh)
        assert s.start() and s.set_caps(sink[0], fixed), s.last_error
        out, m, n = s.transform_array(f)
        site = s.session.site
        assert (out == o.convert(f, ow, oh, out_fmt=o.OUT_NV12, matrix=s.session.matrix, chroma_site=site)).all()
        assert m.all() and n == m.size
        assert s.stop()


# ---- temporal damage policy (in-kernel counters) ------------------------------------
@pytest.mark.parametrize("knobs", [(10, 20, 15), (2, 3, 4)])
@pytest.mark.parametrize("mode", ["convert", "scale"])
def test_policy_map_matches_oracle_over_a_long_sequence(knobs, mode):
    w, h = 640, 360
    ow, oh = (w, h) if mode == "convert" else (480, 270)
    mbw, mbh = (ow + 15) // 16, (oh + 15) // 16
    rng = np.random.default_rng(knobs[0])
    base = synth.desktop(w, h)
    tracker = o.PolicyTracker(mbw, mbh, *knobs)
    with PreencodeSession(w, h, out_width=ow, out_height=oh, colorimetry=1, chroma_site=1, policy=True,
                          damage_block_threshold=knobs[0], damage_block_duration=knobs[1],
                          paint_over_trigger_frames=knobs[2]) as s:
        cur = base.copy()
        seen_hot = seen_paint = False
        for t in range(70):
            if t < 14:
                cur[40:90, 100:260, :3] = rng.integers(0, 256, (50, 160, 3), dtype=np.uint8)   # a video region
            if t in (3, 30):
                cur[200:210, 400:420, :3] ^= 0x40                                               # one-off edits
            if 40 <= t < 44:
                cur[300:330, 20:60, :3] = rng.integers(0, 256, (30, 40, 3), dtype=np.uint8)
            r = s.convert(cur, slot=t & 1)
            want = tracker.update(r.damage_map)
            assert (r.policy_map == want).all(), f"frame {t}"
            assert ((r.policy_map & 1) == r.damage_map).all()
            seen_hot |= bool((r.policy_map & 2).any())
            seen_paint |= bool((r.policy_map & 4).any())
        assert seen_hot and seen_paint
        # a history reset makes everything damaged again but keeps the counters running
        s.reset_history()
        r = s.convert(cur, slot=0)
        assert (r.policy_map == tracker.update(r.damage_map)).all() and r.damage_map.all()


def test_policy_in_batched_launches():
    w, h, n = 640, 360, 3
    sessions = [PreencodeSession(w, h, colorimetry=1, chroma_site=1, policy=True, damage_block_threshold=2,
                                 damage_block_duration=3, paint_over_trigger_frames=2) for _ in range(n)]
    try:
        frames = [[rand_frame(w, h, 10 * i + k) if k < 2 else None for k in range(2)] for i in range(n)]
        for s, (a, b) in zip(sessions, frames):
            s.upload(0, a); s.wait(s.submit(0))
            s.upload(1, b); s.wait(s.submit(1))
            s.reset_history()
        trackers = [o.PolicyTracker(40, 23, 2, 3, 2) for _ in range(n)]
        for tr in trackers:                       # the two set-up frames already advanced the counters
            tr.update(np.ones((23, 40), np.uint8)); tr.update(np.ones((23, 40), np.uint8))
        batch = SessionBatch(sessions)
        sets = [batch.frame_set([s.device_input(k)[0] for s in sessions]) for k in (0, 1)]
        stride = sessions[0].device_input(0)[1]
        order = [0, 1, 1, 1, 1, 1, 0, 0, 0]
        for fid, k in enumerate(order):
            batch.submit(sets[k], stride, 500 + fid)
            batch.wait(500 + fid)
            for s, tr in zip(sessions, trackers):
                r = s.wait(500 + fid)
                assert (r.policy_map == tr.update(r.damage_map)).all(), (fid, k)
    finally:
        for s in sessions:
            s.close()
