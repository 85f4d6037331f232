"""The pixelflux-shaped front (capture_shim.py): surface on CPU, stripes vs the oracle on the GPU."""
import numpy as np
import pytest

from selkies_gstreamer_b200 import capture_shim as cs
from selkies_gstreamer_b200 import stripes as st


def test_settings_carry_the_fields_the_reference_sets():
    """src/selkies/selkies.py:2925-2963 assigns these on a CaptureSettings()."""
    s = cs.CaptureSettings()
    for name in ("capture_width", "capture_height", "capture_x", "capture_y", "target_fps", "capture_cursor",
                 "debug_logging", "output_mode", "jpeg_quality", "paint_over_jpeg_quality", "h264_crf",
                 "h264_paintover_crf", "h264_paintover_burst_frames", "h264_fullcolor", "h264_streaming_mode",
                 "h264_fullframe", "use_paint_over_quality", "paint_over_trigger_frames", "damage_block_threshold",
                 "damage_block_duration", "use_cpu", "vaapi_render_node_index", "watermark_path",
                 "watermark_location_enum"):
        assert hasattr(s, name), name
    assert (s.paint_over_trigger_frames, s.damage_block_threshold, s.damage_block_duration) == (15, 10, 20)


def test_result_looks_like_what_the_callback_dereferences():
    """selkies.py:2870-2872: result = result_ptr.contents; bytes(result.data[:result.size]); result.frame_id."""
    r = cs.StripeEncodeResult(b"abcdef", 7)
    assert r.contents is r and bytes(r.contents.data[:r.contents.size]) == b"abcdef" and r.contents.frame_id == 7
    fn = lambda p, u: None  # noqa: E731
    assert cs.StripeCallback(fn) is fn


def test_start_needs_even_sizes_and_push_needs_start():
    m = cs.ScreenCapture()
    bad = cs.CaptureSettings()
    bad.capture_width = 1281
    with pytest.raises(ValueError):
        m.start_capture(bad, None)
    with pytest.raises(RuntimeError):
        m.push_frame(np.zeros((2, 2, 4), np.uint8))


@pytest.mark.gpu
@pytest.mark.parametrize("mode", ["h264-stripes", "jpeg", "h264-fullframe"])
def test_stripes_sent_cover_exactly_the_oracles_damaged_rows(mode):
    from oracle import cpu_oracle as o
    from selkies_gstreamer_b200 import synth
    w, h = 1280, 720
    s = cs.CaptureSettings()
    s.capture_width, s.capture_height = w, h
    s.output_mode = cs.OUTPUT_MODE_JPEG if mode == "jpeg" else cs.OUTPUT_MODE_H264
    s.h264_fullframe = mode == "h264-fullframe"
    s.paint_over_trigger_frames, s.damage_block_threshold, s.damage_block_duration = 3, 4, 5
    s.n_stripes = 9
    fmt_name = "NV12" if mode == "jpeg" else "I420"
    ofmt = {"NV12": o.OUT_NV12, "I420": o.OUT_I420}[fmt_name]
    got = []
    mod = cs.ScreenCapture()
    mod.start_capture(s, cs.StripeCallback(lambda ptr, user: got.append(ptr.contents)))
    try:
        layout = st.stripe_layout(h, 9)
        tracker = o.PolicyTracker((w + 15) // 16, (h + 15) // 16, 4, 5, 3)
        cur = synth.desktop(w, h, seed=11)
        prev_e = None
        for k in range(12):
            if k in (1, 2):
                cur = cur.copy()
                cur[100 + 200 * k:110 + 200 * k, 300:340, :3] ^= 0x55         # one small edit: one stripe
            elif k == 7:
                cur = synth.mutate(cur, 0.01, seed=k, aligned=False)             # scattered edits: several stripes
            got.clear()
            sent = mod.push_frame(cur)
            assert [r.data for r in sent] == [r.data for r in got]                # callback saw the same packets
            e = o.convert(cur, out_fmt=ofmt, matrix=1, chroma_site=1, n_threads=4)
            dm, _ = o.damage(w, h, ofmt, e, prev_e)
            pm = tracker.update(dm)
            want = [i for i, (y0, hh) in enumerate(layout)
                    if dm[y0 // 16:(y0 + hh + 15) // 16].any() or (pm[y0 // 16:(y0 + hh + 15) // 16] & 4).any()]
            if mode == "h264-fullframe":
                assert len(sent) == (1 if want else 0), k
                if sent:
                    pkt = st.parse_packet(sent[0].data)
                    assert pkt["type"] == st.TYPE_H264_FULL and pkt["frame_id"] == k
                    assert pkt["payload"] == b"".join(np.ascontiguousarray(p).tobytes() for p in o.split_planes(e, w, h, ofmt))
            else:
                assert [r.stripe_y_start for r in sent] == [layout[i][0] for i in want], k
                for r, i in zip(sent, want):
                    y0, hh = layout[i]
                    data = (b"\x03\x00" + r.data) if mode == "jpeg" else r.data     # the server prepends these for JPEG
                    pkt = st.parse_packet(data)
                    assert pkt["frame_id"] == k and pkt["y_start"] == y0
                    if mode != "jpeg":
                        assert (pkt["width"], pkt["height"]) == (w, hh)
                    # payload = the oracle's pixels of exactly that stripe
                    views = st.stripe_views(e, w, h, fmt_name, *o.out_layout(w, h, ofmt), y0, hh)
                    assert pkt["payload"] == b"".join(np.ascontiguousarray(v).tobytes() for v in views), (k, i)
            prev_e = e
        assert mod.stats["frames"] == 12 and mod.stats["stripes_sent"] < mod.stats["stripes_total"]
        assert mod.stats["readback_bytes"] < 12 * len(e) // 2                     # gated read-back: far less than 12 frames
    finally:
        mod.stop_capture()
