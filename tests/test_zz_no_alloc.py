"""sk_scale_frame_noalloc — the entry point that allocates nothing — against the CPU oracle and
against the session path.

This file is named so that pytest collects it last: if it fails under `-x`, every other
test has already run.

The scenario is the one the entry point exists for: the caller owns all GPU memory (here it
comes from torch's caching allocator, standing in for an encoder's buffer pool or any host
that forbids library-side cudaMalloc while other sessions are streaming) and asks the library
to shrink+convert a frame and diff it against the previous one on those buffers.  The library's
only scratch is the fixed array compiled into it.  The results must be byte-for-byte what the
oracle computes — and therefore also what the ordinary, allocating entry points
(sk_session_create → sk_submit → sk_wait) produce for the same frames; both are checked, with the
same exact-equality bar every other GPU test in this repository uses — on two calls in a row,
the second of which reads the first one's output as its previous frame.
"""
import ctypes

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

FMT = {"NV12": 0, "I420": 1}


def _oracle_reference(frames, w, h, ow, oh, fmt, matrix, site, in_fmt=0, damage=True):
    """Frames, maps and counts from the CPU oracle."""
    from oracle import cpu_oracle as o
    out, prev = [], None
    for f in frames:
        e = o.convert(f, ow, oh, in_fmt=in_fmt, out_fmt=FMT[fmt], matrix=matrix, chroma_site=site, n_threads=8)
        m, n = o.damage(ow, oh, FMT[fmt], e, prev) if damage else (None, None)
        out.append((e, m, n))
        prev = e
    return out


def _session_reference(frames, w, h, ow, oh, fmt, matrix, site):
    """The same from the existing (allocating) entry points."""
    from selkies_gstreamer_b200 import PreencodeSession
    out = []
    with PreencodeSession(w, h, out_width=ow, out_height=oh, out_format=fmt, colorimetry=matrix,
                          chroma_site=site) as s:
        for i, f in enumerate(frames):
            r = s.convert(f, slot=i % 2)
            out.append((r.frame.copy(), r.damage_map.copy(), r.damaged_count))
    return out


def _config(w, h, ow, oh, fmt, matrix, site, device_id=-1, in_fmt=0, flags=0):
    from selkies_gstreamer_b200 import _abi
    return _abi.SkConfig(ctypes.sizeof(_abi.SkConfig), device_id, w, h, 0, in_fmt, ow, oh, FMT[fmt], matrix, site,
                         _abi.SK_SCALE_BILINEAR, _abi.SK_SCALE_VARIANT_GST_LQ, 0, flags)


def _run_pair(lib, torch, frames, w, h, ow, oh, fmt, matrix, site, *, in_fmt=0, pad_px=0, damage=True, device_id=-1):
    """Two (or more) calls in a row on caller-owned buffers; returns [(frame, map, count), ...].
    pad_px > 0 puts the frames into wider device buffers: rows are then 4*(w+pad_px) bytes apart."""
    from selkies_gstreamer_b200 import _abi, frame_layout
    dev = torch.device("cuda:0")
    frame_bytes = frame_layout(ow, oh, fmt)[0]
    mbw, mbh = (ow + 15) // 16, (oh + 15) // 16
    d_in = []
    for f in frames:
        wide = np.full((h, w + pad_px, 4), 0x99, np.uint8)
        wide[:, :w] = f
        d_in.append(torch.from_numpy(wide).to(dev))
    stride = 4 * (w + pad_px) if pad_px else 0
    d_out = [torch.full((frame_bytes,), 0xEE, dtype=torch.uint8, device=dev) for _ in frames]
    d_map = torch.full((mbh * mbw,), 0x77, dtype=torch.uint8, device=dev)
    d_cnt = torch.full((1,), -1, dtype=torch.int32, device=dev)
    torch.cuda.synchronize()
    cfg = _config(w, h, ow, oh, fmt, matrix, site, device_id, in_fmt, 0 if damage else _abi.SK_FLAG_NO_DAMAGE)
    err = ctypes.create_string_buffer(256)
    stream = torch.cuda.current_stream().cuda_stream
    got = []
    for i in range(len(frames)):
        prev = d_out[i - 1].data_ptr() if i > 0 else None
        rc = lib.sk_scale_frame_noalloc(ctypes.byref(cfg), d_in[i].data_ptr(), stride, d_out[i].data_ptr(), prev,
                                        d_map.data_ptr() if damage else None, d_cnt.data_ptr() if damage else None,
                                        stream, err, len(err))
        assert rc == 0, err.value.decode()
        torch.cuda.synchronize()
        got.append((d_out[i].cpu().numpy(), d_map.cpu().numpy().reshape(mbh, mbw), int(d_cnt.cpu()[0])))
    return got


def _same_pixels(a, b, w, h, fmt):
    """Equal on every pixel of every plane.  Row-padding bytes (strides are rounded up to 4) are not
    compared: the kernel never writes them, and here the buffers are the caller's, pre-filled with 0xEE."""
    from oracle import cpu_oracle as o
    return all(np.array_equal(x, y) for x, y in zip(o.split_planes(a, w, h, FMT[fmt]), o.split_planes(b, w, h, FMT[fmt])))


def _need_gpu():
    torch = pytest.importorskip("torch")
    if not torch.cuda.is_available():
        pytest.skip("needs a CUDA device")
    return torch


@pytest.mark.parametrize("w,h,ow,oh,fmt", [(1280, 720, 960, 540, "I420"), (640, 360, 640, 360, "NV12")])
def test_noalloc_matches_oracle_and_session_path_twice_in_a_row(w, h, ow, oh, fmt):
    torch = _need_gpu()
    from selkies_gstreamer_b200 import _abi, synth
    lib = _abi.load()
    matrix, site = _abi.SK_MATRIX_BT709, _abi.SK_SITE_H_COSITED
    f0 = synth.desktop(w, h, seed=5)
    f1 = synth.mutate(f0, 0.04, seed=9, aligned=False)
    want = _oracle_reference([f0, f1], w, h, ow, oh, fmt, matrix, site)
    sess = _session_reference([f0, f1], w, h, ow, oh, fmt, matrix, site)
    got = _run_pair(lib, torch, [f0, f1], w, h, ow, oh, fmt, matrix, site)
    for i, ((gf, gm, gc), (wf, wm, wc), (sf, sm, sc)) in enumerate(zip(got, want, sess)):
        assert np.array_equal(gf, wf), f"call {i}: output frame differs from the oracle"
        assert np.array_equal(gm, wm), f"call {i}: damage map differs from the oracle"
        assert gc == wc, f"call {i}: damaged count {gc} != {wc}"
        assert np.array_equal(gf, sf) and np.array_equal(gm, sm) and gc == sc, f"call {i}: differs from the session path"
    assert got[0][1].all() and 0 < got[1][2] < got[1][1].size
    # A second pair right after, with another geometry: the shared array is rewritten per call.
    g0 = synth.noise(256, 144, seed=3)
    g1 = synth.mutate(g0, 0.1, seed=4, aligned=True)
    want = _oracle_reference([g0, g1], 256, 144, 128, 72, "NV12", _abi.SK_MATRIX_BT601, _abi.SK_SITE_NONE)
    got = _run_pair(lib, torch, [g0, g1], 256, 144, 128, 72, "NV12", _abi.SK_MATRIX_BT601, _abi.SK_SITE_NONE)
    for (gf, gm, gc), (wf, wm, wc) in zip(got, want):
        assert np.array_equal(gf, wf) and np.array_equal(gm, wm) and gc == wc


def test_noalloc_padded_rows_rgbx_and_convert_only():
    """Rows further apart than 4*width (a sub-rectangle of a larger framebuffer), RGBx byte order, and
    SK_FLAG_NO_DAMAGE (no map, no count, no previous frame) — each against the oracle, on the device
    named explicitly instead of 'the current one'."""
    torch = _need_gpu()
    from selkies_gstreamer_b200 import _abi, synth
    lib = _abi.load()
    w, h, ow, oh = 802, 450, 640, 360
    rng = np.random.default_rng(21)
    f0 = rng.integers(0, 256, (h, w, 4), dtype=np.uint8)
    f1 = synth.mutate(f0, 0.05, seed=2, aligned=False)
    want = _oracle_reference([f0, f1], w, h, ow, oh, "NV12", 0, 0, in_fmt=1)
    got = _run_pair(lib, torch, [f0, f1], w, h, ow, oh, "NV12", 0, 0, in_fmt=1, pad_px=37, device_id=0)
    for i, ((gf, gm, gc), (wf, wm, wc)) in enumerate(zip(got, want)):
        assert np.array_equal(gf, wf) and np.array_equal(gm, wm) and gc == wc, f"call {i}"
    # convert only: the map and count buffers may be NULL and stay untouched (802 wide: the I420 chroma
    # rows carry 3 bytes of padding each)
    want = _oracle_reference([f0, f1], w, h, w, h, "I420", 2, 1, damage=False)
    got = _run_pair(lib, torch, [f0, f1], w, h, w, h, "I420", 2, 1, damage=False, pad_px=3)
    for i, ((gf, gm, gc), (wf, _, _)) in enumerate(zip(got, want)):
        assert _same_pixels(gf, wf, w, h, "I420"), f"call {i}"
        assert (gm == 0x77).all() and gc == -1


def test_noalloc_accepts_the_tallest_output_that_fits():
    """out_height == SK_NOALLOC_MAX_OUT_HEIGHT uses the whole tap table and must still be exact."""
    torch = _need_gpu()
    from selkies_gstreamer_b200 import _abi
    lib = _abi.load()
    w, h, ow, oh = 64, 4400, 48, _abi.SK_NOALLOC_MAX_OUT_HEIGHT
    rng = np.random.default_rng(4)
    f0 = rng.integers(0, 256, (h, w, 4), dtype=np.uint8)
    f1 = f0.copy()
    f1[4399, 63, :3] ^= 0x11                   # the very last source row feeds the very last tap entry
    want = _oracle_reference([f0, f1], w, h, ow, oh, "NV12", 1, 1)
    got = _run_pair(lib, torch, [f0, f1], w, h, ow, oh, "NV12", 1, 1)
    for i, ((gf, gm, gc), (wf, wm, wc)) in enumerate(zip(got, want)):
        assert np.array_equal(gf, wf) and np.array_equal(gm, wm) and gc == wc, f"call {i}"
    assert got[1][2] >= 1 and got[1][1][-1, -1] == 1


def test_noalloc_rejects_what_does_not_fit_the_fixed_array():
    torch = _need_gpu()
    from selkies_gstreamer_b200 import _abi
    lib = _abi.load()
    assert _abi.SK_NOALLOC_MAX_OUT_HEIGHT == 4320
    buf = torch.zeros(64, dtype=torch.uint8, device="cuda:0")
    err = ctypes.create_string_buffer(256)
    cfg = _config(8192, 8192, 8192, 4322, "NV12", 1, 1)       # taller than the tap table
    rc = lib.sk_scale_frame_noalloc(ctypes.byref(cfg), buf.data_ptr(), 0, buf.data_ptr(), None, buf.data_ptr(),
                                    buf.data_ptr(), None, err, len(err))
    assert rc == _abi.SK_ERR_ARG
    assert b"SK_NOALLOC_MAX_OUT_HEIGHT" in err.value and b"4322" in err.value
    cfg = _config(640, 360, 1280, 720, "NV12", 1, 1)          # enlarging
    rc = lib.sk_scale_frame_noalloc(ctypes.byref(cfg), buf.data_ptr(), 0, buf.data_ptr(), None, buf.data_ptr(),
                                    buf.data_ptr(), None, err, len(err))
    assert rc == _abi.SK_ERR_UNSUPPORTED and err.value
    cfg = _config(640, 360, 640, 360, "NV12", 1, 1)           # missing map/count
    rc = lib.sk_scale_frame_noalloc(ctypes.byref(cfg), buf.data_ptr(), 0, buf.data_ptr(), None, None, None, None,
                                    err, len(err))
    assert rc == _abi.SK_ERR_ARG and b"d_map" in err.value
    cfg = _config(2, 8192, 2, 4096, "NV12", 1, 1)             # a pitch the kernel's 32-bit offsets cannot address
    rc = lib.sk_scale_frame_noalloc(ctypes.byref(cfg), buf.data_ptr(), 1 << 20, buf.data_ptr(), None, buf.data_ptr(),
                                    buf.data_ptr(), None, err, len(err))
    assert rc == _abi.SK_ERR_ARG and b"4 GiB" in err.value
