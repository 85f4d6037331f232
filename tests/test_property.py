"""Property tests (hypothesis) over random geometry and modes.

CPU part: the C oracle against the closed-form numpy restatement.  GPU part (marked gpu): the
CUDA path through the C ABI against the oracle — ragged even sizes, both scale kernels,
every matrix / chroma-site / layout / input-order combination, damage maps included.
"""
import numpy as np
import pytest
from hypothesis import HealthCheck, given, settings, strategies as st

from oracle import cpu_oracle as o
from oracle import np_restatement as npr

even = lambda lo, hi: st.integers(lo // 2, hi // 2).map(lambda v: 2 * v)
MATRIX = ["bt601", "bt709", "bt2020"]
SITE = ["none", "h-cosited"]


@settings(max_examples=60, deadline=None, derandomize=True, suppress_health_check=[HealthCheck.too_slow])
@given(w=even(2, 200), h=even(2, 120), m=st.integers(0, 2), site=st.integers(0, 1), fmt=st.integers(0, 1),
       infmt=st.integers(0, 1), seed=st.integers(0, 2**31 - 1))
def test_oracle_vs_numpy_random_geometry(w, h, m, site, fmt, infmt, seed):
    f = np.random.default_rng(seed).integers(0, 256, (h, w, 4), dtype=np.uint8)
    out = o.convert(f, in_fmt=infmt, out_fmt=fmt, matrix=m, chroma_site=site, n_threads=2)
    y, u, v = o.split_planes(out, w, h, fmt)
    Y, U, V = npr.convert(f, in_fmt=["BGRx", "RGBx"][infmt], matrix=MATRIX[m], site=SITE[site])
    assert (y == Y).all() and (u == U).all() and (v == V).all()


@settings(max_examples=80, deadline=None, derandomize=True, suppress_health_check=[HealthCheck.too_slow])
@given(w=even(4, 160), h=even(4, 100), fx=st.floats(0.3, 2.5), fy=st.floats(0.3, 2.5), var=st.integers(0, 1),
       seed=st.integers(0, 2**31 - 1))
def test_oracle_scale_vs_numpy_random_geometry(w, h, fx, fy, var, seed):
    """Factors above 1 make the picture grow on one or both axes: scale behind the matrix when the pixel count grows."""
    ow, oh = max(2, int(w * fx) & ~1), max(2, int(h * fy) & ~1)
    f = np.random.default_rng(seed).integers(0, 256, (h, w, 4), dtype=np.uint8)
    off, tap = o.resampler_2tap(h, oh, var)
    out = o.convert(f, ow, oh, scale_variant=var, out_fmt=o.OUT_I420)
    y, u, v = o.split_planes(out, ow, oh, o.OUT_I420)
    Y, U, V = npr.convert(f, out_w=ow, out_h=oh, v_off=off, v_tap=tap, h_inc=o.h_increment(w, ow))
    assert (y == Y).all() and (u == U).all() and (v == V).all()


@pytest.mark.gpu
@settings(max_examples=40, deadline=None, derandomize=True, suppress_health_check=[HealthCheck.too_slow, HealthCheck.function_scoped_fixture])
@given(w=even(2, 700), h=even(2, 200), m=st.integers(0, 2), site=st.integers(0, 1), fmt=st.sampled_from(["NV12", "I420"]),
       infmt=st.sampled_from(["BGRx", "RGBx"]), variant=st.integers(0, 15), seed=st.integers(0, 2**31 - 1))
def test_gpu_convert_random_geometry(w, h, m, site, fmt, infmt, variant, seed):
    from selkies_gstreamer_b200 import PreencodeSession, _abi
    rng = np.random.default_rng(seed)
    f0 = rng.integers(0, 256, (h, w, 4), dtype=np.uint8)
    f1 = f0.copy()
    for _ in range(3):
        y, x = int(rng.integers(0, h)), int(rng.integers(0, w))
        f1[y, x, :3] ^= 0x5A
    ofmt, oin = {"NV12": 0, "I420": 1}[fmt], {"BGRx": 0, "RGBx": 1}[infmt]
    lib = _abi.load()
    tuning = hasattr(lib, "sk_tuning_set_variant")      # the release library has the default shape only
    try:
        if tuning:
            lib.sk_tuning_set_variant(variant)
        with PreencodeSession(w, h, in_format=infmt, out_format=fmt, colorimetry=m, chroma_site=site) as s:
            prev = None
            for i, f in enumerate((f0, f1, f1)):
                r = s.convert(f, slot=i & 1)
                e = o.convert(f, in_fmt=oin, out_fmt=ofmt, matrix=m, chroma_site=site)
                assert (r.frame == e).all()
                em, en = o.damage(w, h, ofmt, e, prev)
                assert (r.damage_map == em).all() and r.damaged_count == en
                prev = e
    finally:
        if tuning:
            lib.sk_tuning_set_variant(0)


@pytest.mark.gpu
@settings(max_examples=80, deadline=None, derandomize=True, suppress_health_check=[HealthCheck.too_slow, HealthCheck.function_scoped_fixture])
@given(w=even(4, 600), h=even(4, 200), fx=st.floats(0.2, 2.5), fy=st.floats(0.2, 2.5), var=st.integers(0, 1),
       site=st.integers(0, 1), fmt=st.sampled_from(["NV12", "I420"]), generic=st.booleans(), seed=st.integers(0, 2**31 - 1))
def test_gpu_scale_random_geometry(w, h, fx, fy, var, site, fmt, generic, seed):
    from selkies_gstreamer_b200 import PreencodeSession, _abi
    ow, oh = max(2, int(w * fx) & ~1), max(2, int(h * fy) & ~1)
    rng = np.random.default_rng(seed)
    f0 = rng.integers(0, 256, (h, w, 4), dtype=np.uint8)
    f1 = f0.copy()
    f1[h // 2, w // 2, :3] ^= 0xC3
    ofmt = {"NV12": 0, "I420": 1}[fmt]
    with PreencodeSession(w, h, out_width=ow, out_height=oh, out_format=fmt, colorimetry=1, chroma_site=site,
                          scale_variant=var, scale_generic=generic) as s:
        prev = None
        for i, f in enumerate((f0, f1)):
            r = s.convert(f, slot=i & 1)
            e = o.convert(f, ow, oh, out_fmt=ofmt, matrix=1, chroma_site=site, scale_variant=var)
            assert (r.frame == e).all()
            em, en = o.damage(ow, oh, ofmt, e, prev)
            assert (r.damage_map == em).all() and r.damaged_count == en
            prev = e


@pytest.mark.gpu
@settings(max_examples=30, deadline=None, derandomize=True, suppress_health_check=[HealthCheck.too_slow, HealthCheck.function_scoped_fixture])
@given(w=even(2, 700), h=even(2, 200), fmt=st.sampled_from(["NV12", "I420"]), n=st.integers(2, 4), site=st.integers(0, 1),
       seed=st.integers(0, 2**31 - 1))
def test_gpu_gated_readback_random_geometry(w, h, fmt, n, site, seed):
    """Damage-gated read-back: whatever the geometry, ring depth and damage pattern, the persistent host frame
    equals the oracle's whole frame after every conversion."""
    from selkies_gstreamer_b200 import PreencodeSession
    rng = np.random.default_rng(seed)
    ofmt = {"NV12": 0, "I420": 1}[fmt]
    cur = rng.integers(0, 256, (h, w, 4), dtype=np.uint8)
    with PreencodeSession(w, h, out_format=fmt, colorimetry=1, chroma_site=site, n_inflight=n, host_output=True,
                          gated_readback=True) as s:
        for k in range(2 * n + 3):
            if k and rng.random() < 0.8:                      # sometimes nothing changes at all
                cur = cur.copy()
                for _ in range(int(rng.integers(1, 4))):
                    y, x = int(rng.integers(0, h)), int(rng.integers(0, w))
                    cur[y:y + int(rng.integers(1, 20)), x:x + int(rng.integers(1, 40)), :3] ^= 0x3C
            r = s.convert(cur, slot=k % n)
            e = o.convert(cur, out_fmt=ofmt, matrix=1, chroma_site=site)
            assert (r.frame == e).all(), k
            assert r.readback_bytes % 384 == 0 and r.readback_bytes <= 384 * r.damage_map.size
