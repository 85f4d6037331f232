"""geometry.fit_resolution against vectors generated from the reference's own fit_res
(src/selkies/legacy/resize.py:15-29; generator: tests/golden/make_fit_res.py)."""
import json
import os

from selkies_gstreamer_b200 import geometry

GOLDEN = os.path.join(os.path.dirname(__file__), "golden", "fit_res.json")


def test_fit_resolution_matches_reference_vectors():
    vectors = json.load(open(GOLDEN))["vectors"]
    assert len(vectors) >= 50
    for v in vectors:
        assert list(geometry.fit_resolution(*v["in"])) == v["out"], v


def test_results_that_shrink_are_even_and_inside_the_limit():
    for v in json.load(open(GOLDEN))["vectors"]:
        w, h, mw, mh = v["in"]
        ow, oh = geometry.fit_resolution(w, h, mw, mh)
        if (ow, oh) != (w, h):
            assert ow % 2 == 0 and oh % 2 == 0 and ow <= mw + 1 and oh <= mh + 1


def test_frame_size_rule_and_macroblocks():
    assert geometry.is_valid_frame_size(1920, 1080) and geometry.is_valid_frame_size(8192, 2)
    assert not geometry.is_valid_frame_size(1921, 1080) and not geometry.is_valid_frame_size(8194, 1080)
    assert geometry.macroblocks(1920, 1080) == (120, 68) and geometry.macroblocks(3840, 2160) == (240, 135)
    assert geometry.macroblocks(7680, 4320) == (480, 270)
