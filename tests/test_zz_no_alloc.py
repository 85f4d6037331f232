"""sk_scale_frame_noalloc — the entry point that allocates nothing — against the session path.

This file is named so that pytest collects it last: if it fails under `-x`, every other
test has already run.

The scenario is the one the entry point exists for: the caller owns all GPU memory (here it
comes from torch's caching allocator, standing in for an encoder's buffer pool or any host
that forbids library-side cudaMalloc while other sessions are streaming) and asks the library
to shrink+convert a frame and diff it against the previous one on those buffers.  The library's
only scratch is the fixed array compiled into it.  The results must be byte-for-byte what the
ordinary, allocating entry points (sk_session_create → sk_submit → sk_wait) produce for the
same frames — the same exact-equality bar every other GPU test in this repository uses — on
two calls in a row, the second of which reads the first one's output as its previous frame.
"""
import ctypes

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _session_reference(frames, w, h, ow, oh, fmt, matrix, site):
    """Frames, maps and counts from the existing (allocating) entry points."""
    from selkies_gstreamer_b200 import PreencodeSession
    out = []
    with PreencodeSession(w, h, out_width=ow, out_height=oh, out_format=fmt, colorimetry=matrix,
                          chroma_site=site) as s:
        for i, f in enumerate(frames):
            r = s.convert(f, slot=i % 2)
            out.append((r.frame.copy(), r.damage_map.copy(), r.damaged_count))
    return out


def _config(w, h, ow, oh, fmt, matrix, site, device_id=-1):
    from selkies_gstreamer_b200 import _abi
    return _abi.SkConfig(ctypes.sizeof(_abi.SkConfig), device_id, w, h, 0, _abi.SK_IN_BGRX, ow, oh,
                         {"NV12": _abi.SK_OUT_NV12, "I420": _abi.SK_OUT_I420}[fmt], matrix, site,
                         _abi.SK_SCALE_BILINEAR, _abi.SK_SCALE_VARIANT_GST_LQ, 0, 0)


def _run_pair(lib, torch, frames, w, h, ow, oh, fmt, matrix, site):
    """Two calls in a row on caller-owned buffers; returns [(frame, map, count), ...]."""
    from selkies_gstreamer_b200 import frame_layout
    dev = torch.device("cuda:0")
    frame_bytes = frame_layout(ow, oh, fmt)[0]
    mbw, mbh = (ow + 15) // 16, (oh + 15) // 16
    d_in = [torch.from_numpy(f).to(dev) for f in frames]
    d_out = [torch.full((frame_bytes,), 0xEE, dtype=torch.uint8, device=dev) for _ in frames]
    d_map = torch.full((mbh * mbw,), 0x77, dtype=torch.uint8, device=dev)
    d_cnt = torch.full((1,), -1, dtype=torch.int32, device=dev)
    torch.cuda.synchronize()
    cfg = _config(w, h, ow, oh, fmt, matrix, site)
    err = ctypes.create_string_buffer(256)
    stream = torch.cuda.current_stream().cuda_stream
    got = []
    for i in range(len(frames)):
        prev = d_out[i - 1].data_ptr() if i > 0 else None
        rc = lib.sk_scale_frame_noalloc(ctypes.byref(cfg), d_in[i].data_ptr(), 0, d_out[i].data_ptr(), prev,
                                        d_map.data_ptr(), d_cnt.data_ptr(), stream, err, len(err))
        assert rc == 0, err.value.decode()
        torch.cuda.synchronize()
        got.append((d_out[i].cpu().numpy(), d_map.cpu().numpy().reshape(mbh, mbw), int(d_cnt.cpu()[0])))
    return got


@pytest.mark.parametrize("w,h,ow,oh,fmt", [(1280, 720, 960, 540, "I420"), (640, 360, 640, 360, "NV12")])
def test_noalloc_matches_session_path_twice_in_a_row(w, h, ow, oh, fmt):
    torch = pytest.importorskip("torch")
    if not torch.cuda.is_available():
        pytest.skip("needs a CUDA device")
    from selkies_gstreamer_b200 import _abi, synth
    lib = _abi.load()
    matrix, site = _abi.SK_MATRIX_BT709, _abi.SK_SITE_H_COSITED
    f0 = synth.desktop(w, h, seed=5)
    f1 = synth.mutate(f0, 0.04, seed=9, aligned=False)
    want = _session_reference([f0, f1], w, h, ow, oh, fmt, matrix, site)
    got = _run_pair(lib, torch, [f0, f1], w, h, ow, oh, fmt, matrix, site)
    for i, ((gf, gm, gc), (wf, wm, wc)) in enumerate(zip(got, want)):
        assert np.array_equal(gf, wf), f"call {i}: output frame differs from the session path"
        assert np.array_equal(gm, wm), f"call {i}: damage map differs from the session path"
        assert gc == wc, f"call {i}: damaged count {gc} != {wc}"
    assert got[0][1].all() and 0 < got[1][2] < got[1][1].size
    # A second pair right after, with another geometry: the shared array is rewritten per call.
    g0 = synth.noise(256, 144, seed=3)
    g1 = synth.mutate(g0, 0.1, seed=4, aligned=True)
    want = _session_reference([g0, g1], 256, 144, 128, 72, "NV12", _abi.SK_MATRIX_BT601, _abi.SK_SITE_NONE)
    got = _run_pair(lib, torch, [g0, g1], 256, 144, 128, 72, "NV12", _abi.SK_MATRIX_BT601, _abi.SK_SITE_NONE)
    for (gf, gm, gc), (wf, wm, wc) in zip(got, want):
        assert np.array_equal(gf, wf) and np.array_equal(gm, wm) and gc == wc


def test_noalloc_rejects_what_does_not_fit_the_fixed_array():
    torch = pytest.importorskip("torch")
    if not torch.cuda.is_available():
        pytest.skip("needs a CUDA device")
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
