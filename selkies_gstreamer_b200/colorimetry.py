"""Caps-driven choice of colour matrix and chroma site.

The reference never sets colorimetry or chroma-site on the stage's src caps
(src/selkies/legacy/gstwebrtc_app.py:614-617 forces only `format`), and ximagesrc
emits none, so GStreamer's defaults by output height apply: upstream
video-info.c set_default_colorimetry().  A capsfilter that names them wins.
"""
from __future__ import annotations

from . import _abi

MATRIX_BY_NAME = {
    "bt601": _abi.SK_MATRIX_BT601, "smpte170m": _abi.SK_MATRIX_BT601,
    "bt709": _abi.SK_MATRIX_BT709,
    "bt2020": _abi.SK_MATRIX_BT2020, "bt2020-10": _abi.SK_MATRIX_BT2020,
}
# GStreamer serialises unknown combinations as range:matrix:transfer:primaries; matrix ids per
# GstVideoColorMatrix: 3 = bt709, 4 = bt601, 6 = bt2020.
_MATRIX_BY_ID = {"3": _abi.SK_MATRIX_BT709, "4": _abi.SK_MATRIX_BT601, "6": _abi.SK_MATRIX_BT2020}

SITE_BY_NAME = {
    "none": _abi.SK_SITE_NONE, "jpeg": _abi.SK_SITE_NONE, "center": _abi.SK_SITE_NONE,
    "h-cosited": _abi.SK_SITE_H_COSITED, "mpeg2": _abi.SK_SITE_H_COSITED, "left": _abi.SK_SITE_H_COSITED,
}

MATRIX_NAMES = {_abi.SK_MATRIX_BT601: "bt601", _abi.SK_MATRIX_BT709: "bt709", _abi.SK_MATRIX_BT2020: "bt2020"}
SITE_NAMES = {_abi.SK_SITE_NONE: "jpeg", _abi.SK_SITE_H_COSITED: "mpeg2"}


# upstream's set_default_colorimetry() is recalled as switching to bt2020 from 2160
# lines up.  If real GStreamer caps say otherwise, this is the one name to change (or pass uhd_default).
UHD_DEFAULT_MATRIX = "bt2020"


def default_for_height(out_height: int, uhd_default: str | None = None) -> tuple[int, int]:
    """(matrix, chroma_site) GStreamer assigns to YUV caps that carry neither."""
    if out_height >= 2160:
        return MATRIX_BY_NAME[uhd_default or UHD_DEFAULT_MATRIX], _abi.SK_SITE_H_COSITED
    if out_height > 576:
        return _abi.SK_MATRIX_BT709, _abi.SK_SITE_H_COSITED
    return _abi.SK_MATRIX_BT601, _abi.SK_SITE_NONE


def parse_matrix(colorimetry) -> int | None:
    if colorimetry is None:
        return None
    if isinstance(colorimetry, int):
        if colorimetry in MATRIX_NAMES:
            return colorimetry
        raise ValueError(f"unknown matrix id {colorimetry}")
    name = str(colorimetry).strip().lower()
    if name in MATRIX_BY_NAME:
        return MATRIX_BY_NAME[name]
    parts = name.split(":")
    if len(parts) == 4 and parts[1] in _MATRIX_BY_ID:
        if parts[0] not in ("2",):   # 2 = GST_VIDEO_COLOR_RANGE_16_235
            raise ValueError("only limited-range (16-235) output is supported")
        return _MATRIX_BY_ID[parts[1]]
    raise ValueError(f"unsupported colorimetry {colorimetry!r} (bt601, bt709, bt2020)")


def parse_site(chroma_site) -> int | None:
    if chroma_site is None:
        return None
    if isinstance(chroma_site, int):
        if chroma_site in SITE_NAMES:
            return chroma_site
        raise ValueError(f"unknown chroma-site id {chroma_site}")
    name = str(chroma_site).strip().lower()
    if name in SITE_BY_NAME:
        return SITE_BY_NAME[name]
    raise ValueError(f"unsupported chroma-site {chroma_site!r} (none/jpeg, h-cosited/mpeg2)")


def resolve(out_height: int, colorimetry=None, chroma_site=None, uhd_default: str | None = None) -> tuple[int, int]:
    m, s = default_for_height(out_height, uhd_default)
    pm, ps = parse_matrix(colorimetry), parse_site(chroma_site)
    return (m if pm is None else pm, s if ps is None else ps)
