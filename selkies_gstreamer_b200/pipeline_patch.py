"""Swap the stock convert stage for b200preencode in the reference's pipeline builder
without editing it: wraps Gst.ElementFactory.make for the duration of build_video_pipeline
(src/selkies/legacy/gstwebrtc_app.py:200-1001).

  videoconvert              → b200preencode                         (:611, :643, :669, :686, :725, :742, :767)
  cudaconvert               → b200preencode (cuda-device-id kept)   (:270, :373, :430)
  cudaupload                → identity (the upload is part of b200preencode)   (:263, :370, :427)

For the NVENC branches the capsfilter after the stage asks for memory:CUDAMemory
(:281-284); until CUDAMemory output exists use the software-encoder
branches, or edit that capsfilter to system memory as INTEGRATION.md shows.
"""
from __future__ import annotations

import contextlib

from . import gst_element

SWAPS = {"videoconvert": gst_element.FACTORY_NAME, "cudaconvert": gst_element.FACTORY_NAME, "cudaupload": "identity"}
_IGNORED_ON_IDENTITY = {"cuda-device-id"}


class _PropertyFilter:
    """identity has no cuda-device-id: swallow what the reference sets on cudaupload (:264-265)."""

    def __init__(self, element):
        object.__setattr__(self, "_e", element)

    def set_property(self, name, value):
        if name in _IGNORED_ON_IDENTITY:
            return None
        return self._e.set_property(name, value)

    def __getattr__(self, name):
        return getattr(self._e, name)


@contextlib.contextmanager
def swapped_factories(Gst, swaps: dict | None = None):
    """with swapped_factories(Gst): app.build_video_pipeline()"""
    swaps = dict(SWAPS if swaps is None else swaps)
    gst_element.register()
    factory = Gst.ElementFactory
    original = factory.make

    def make(name, element_name=None):
        target = swaps.get(name, name)
        el = original(target, element_name) if element_name is not None else original(target)
        if name == "cudaupload" and target == "identity" and el is not None:
            return _PropertyFilter(el)
        return el

    factory.make = make
    try:
        yield
    finally:
        factory.make = original
