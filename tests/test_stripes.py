"""Stripe planning and the WebSocket-mode headers, against the byte offsets the browser reads
(addons/gst-web-core/selkies-core.js:2812-2818, :2908-2914, :2925-2938)."""
import numpy as np
import pytest

from selkies_gstreamer_b200 import _abi, frame_layout, stripes as st


def js_view(packet: bytes):
    """The DataView reads the JS client performs, transcribed offset by offset."""
    u8 = lambda o: packet[o]
    u16 = lambda o: (packet[o] << 8) | packet[o + 1]            # getUint16(o, false) = big-endian
    t = u8(0)
    if t == 0x00:
        return {"frameTypeFlag": u8(1), "frameId": u16(2), "payload": packet[4:]}
    if t == 0x03:
        return {"frameId": u16(2), "stripe_y_start": u16(4), "payload": packet[6:]}
    if t == 0x04:
        return {"video_frame_type_byte": u8(1), "vncFrameID": u16(2), "vncStripeYStart": u16(4),
                "stripeWidth": u16(6), "stripeHeight": u16(8), "payload": packet[10:]}
    raise AssertionError(t)


def test_headers_match_the_browser_offsets():
    p = st.pack_h264_stripe_header(0x1234, True, 1072, 3840, 64) + b"NAL"
    assert len(p) == 13
    assert js_view(p) == {"video_frame_type_byte": 1, "vncFrameID": 0x1234, "vncStripeYStart": 1072,
                          "stripeWidth": 3840, "stripeHeight": 64, "payload": b"NAL"}
    p = st.pack_jpeg_stripe_header(70000, 512) + b"\xff\xd8"
    assert js_view(p) == {"frameId": 70000 & 0xFFFF, "stripe_y_start": 512, "payload": b"\xff\xd8"}
    p = st.pack_h264_full_header(7, False) + b"x"
    assert js_view(p) == {"frameTypeFlag": 0, "frameId": 7, "payload": b"x"}
    for pk in (st.pack_h264_stripe_header(9, False, 16, 1920, 1080 - 16), st.pack_jpeg_stripe_header(1, 0),
               st.pack_h264_full_header(65535, True)):
        d = st.parse_packet(pk + b"zz")
        assert d["payload"] == b"zz" and d["frame_id"] in (9, 1, 65535)
    with pytest.raises(ValueError):
        st.parse_packet(b"\x04\x00")
    with pytest.raises(ValueError):
        st.parse_packet(b"\x09abcdefghij")


@pytest.mark.parametrize("height,n", [(1080, 8), (2160, 16), (4320, 32), (16, 4), (720, 7)])
def test_stripe_layout_tiles_the_frame_on_macroblock_rows(height, n):
    lay = st.stripe_layout(height, n)
    assert lay[0][0] == 0 and sum(h for _, h in lay) == height
    for (y0, h), (y1, _) in zip(lay, lay[1:]):
        assert y0 + h == y1 and y1 % 16 == 0
    assert len(lay) == min(n, (height + 15) // 16)
    sizes = [(h + 15) // 16 for _, h in lay]
    assert max(sizes) - min(sizes) <= 1


def test_plan_flags_follow_the_maps():
    mbh, mbw = 68, 120            # 1080p
    dm = np.zeros((mbh, mbw), np.uint8)
    pm = np.zeros((mbh, mbw), np.uint8)
    dm[0, 5] = 1; pm[0, 5] = _abi.SK_POLICY_DAMAGED
    dm[40, 100] = 1; pm[40, 100] = _abi.SK_POLICY_DAMAGED | _abi.SK_POLICY_HOT
    pm[67, 0] = _abi.SK_POLICY_PAINT_OVER
    plan = st.plan_stripes(dm, pm, 1080, 8)
    assert [s.damaged for s in plan] == [True, False, False, False, True, False, False, False]
    assert [s.hot for s in plan][4] and plan[7].paint_over and not plan[7].damaged
    assert [s.index for s in st.stripes_to_send(plan)] == [0, 4, 7]
    assert st.plan_stripes(dm, None, 1080, 8)[7].paint_over is False


@pytest.mark.parametrize("fmt", ["NV12", "I420"])
def test_stripe_views_are_zero_copy_rows_of_the_packed_frame(fmt):
    w, h = 1920, 1080
    size, off, stride = frame_layout(w, h, fmt)
    packed = np.arange(size, dtype=np.uint32).astype(np.uint8)
    y0, sh = st.stripe_layout(h, 8)[7]
    assert y0 + sh == h and sh % 2 == 0
    views = st.stripe_views(packed, w, h, fmt, off, stride, y0, sh)
    assert views[0].shape == (sh, w) and views[0].base is not None
    assert (views[0][0] == packed[y0 * stride[0]: y0 * stride[0] + w]).all()
    if fmt == "NV12":
        assert views[1].shape == (sh // 2, w)
        assert views[1][0, 0] == packed[off[1] + (y0 // 2) * stride[1]]
    else:
        assert views[1].shape == (sh // 2, w // 2) and views[2].shape == (sh // 2, w // 2)
        assert views[2][0, 0] == packed[off[2] + (y0 // 2) * stride[2]]
    views[0][0, 0] ^= 0xFF
    assert packed[y0 * stride[0]] == views[0][0, 0]              # a view, not a copy
    with pytest.raises(ValueError):
        st.stripe_views(packed, w, h, fmt, off, stride, 17, 16)
