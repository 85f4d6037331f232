#!/usr/bin/env python3
"""Smoke test of the element against a REAL GStreamer — for hosts that have PyGObject + GStreamer >= 1.22
and a B200 (neither the build container nor the GPU box of this project has `gi`, so the adaptor in
selkies_gstreamer_b200/gst_element.py has only ever met a hand-written fake; SURVEY rows a3/a10 can only
be closed by running this).

    python tools/gst_smoke.py                 # videotestsrc ! b200preencode ! NV12 ! fakesink, 60 buffers
    python tools/gst_smoke.py --reference /path/to/selkies/src/selkies/legacy
                                              # also: the reference's own build_video_pipeline with the
                                              # stock factories swapped (gstwebrtc_app.py:941-994)
Exit code 0 = every step passed; each step prints PASS/FAIL with the bus error if any.
"""
import argparse
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def run_pipeline(Gst, desc, seconds=10):
    pipe = Gst.parse_launch(desc)
    bus = pipe.get_bus()
    pipe.set_state(Gst.State.PLAYING)
    msg = bus.timed_pop_filtered(seconds * Gst.SECOND, Gst.MessageType.EOS | Gst.MessageType.ERROR)
    ok, why = False, "timeout"
    if msg is not None and msg.type == Gst.MessageType.EOS:
        ok, why = True, "EOS"
    elif msg is not None:
        err, dbg = msg.parse_error()
        why = f"{err.message} ({dbg})"
    pipe.set_state(Gst.State.NULL)
    return ok, why


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--reference", help="directory holding the reference's gstwebrtc_app.py")
    ap.add_argument("--encoder", default="x264enc")
    args = ap.parse_args()
    try:
        import gi
        gi.require_version("Gst", "1.0")
        from gi.repository import Gst
    except (ImportError, ValueError) as e:
        print(f"SKIP: PyGObject/GStreamer not importable here ({e})")
        return 2
    Gst.init(None)
    from selkies_gstreamer_b200 import gst_element, pipeline_patch
    assert gst_element.register(), "Gst.Element.register failed"
    failed = 0
    steps = [
        ("convert 1080p BGRx->NV12", "videotestsrc num-buffers=60 ! video/x-raw,format=BGRx,width=1920,height=1080,framerate=60/1 "
                                     "! b200preencode ! video/x-raw,format=NV12 ! fakesink"),
        ("convert + shrink 1440p->1080p I420", "videotestsrc num-buffers=30 ! video/x-raw,format=BGRx,width=2560,height=1440 "
                                               "! b200preencode ! video/x-raw,format=I420,width=1920,height=1080 ! fakesink"),
        ("convert + enlarge 720p->1080p NV12", "videotestsrc num-buffers=30 ! video/x-raw,format=BGRx,width=1280,height=720 "
                                               "! b200preencode ! video/x-raw,format=NV12,width=1920,height=1080 ! fakesink"),
        ("properties the reference sets", "videotestsrc num-buffers=10 ! video/x-raw,format=BGRx,width=640,height=360 "
                                          "! b200preencode n-threads=4 qos=true cuda-device-id=0 ! video/x-raw,format=NV12 ! fakesink"),
        ("ROI metas reach x264enc", "videotestsrc num-buffers=30 ! video/x-raw,format=BGRx,width=1280,height=720 "
                                    "! b200preencode ! video/x-raw,format=NV12 ! x264enc tune=zerolatency ! fakesink"),
    ]
    for name, desc in steps:
        ok, why = run_pipeline(Gst, desc)
        print(("PASS " if ok else "FAIL ") + name + ("" if ok else f": {why}"))
        failed += 0 if ok else 1
    if args.reference:
        sys.path.insert(0, args.reference)
        import asyncio
        from gstwebrtc_app import GSTWebRTCApp       # the reference's own builder, unmodified
        app = GSTWebRTCApp(asyncio.new_event_loop(), encoder=args.encoder)
        app.pipeline = Gst.Pipeline.new()
        try:
            with pipeline_patch.swapped_factories(Gst):
                app.build_webrtcbin_pipeline()
                app.build_video_pipeline()
            print("PASS reference build_video_pipeline links with b200preencode in place of the stock stage")
        except Exception as e:                        # noqa: BLE001
            print(f"FAIL reference build_video_pipeline: {e}")
            failed += 1
    return 1 if failed else 0


if __name__ == "__main__":
    sys.exit(main())
