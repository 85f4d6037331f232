"""B200-native replacement for the pre-encode stage of the Selkies GStreamer pipeline.

The stage (BGRx/RGBx → I420/NV12 with GStreamer's integer matrices, optional
bilinear shrink, per-16x16 damage map) runs as hand-written sm_100a CUDA inside
libselkies_preencode.so; this package is the ctypes host side and the
GstBase.BaseTransform element that drops into
src/selkies/legacy/gstwebrtc_app.py:611 / :263-270 of the reference.
"""
from . import _abi
from ._abi import PreencodeLibraryError
from .session import (FrameResult, PreencodeError, PreencodeSession, SessionBatch, batch_submit, device_count,
                      frame_layout, host_register, host_unregister, launch_count)

__all__ = ["PreencodeSession", "PreencodeError", "PreencodeLibraryError", "FrameResult", "SessionBatch", "batch_submit",
           "device_count", "frame_layout", "host_register", "host_unregister", "launch_count"]
