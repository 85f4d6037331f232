"""Temporal damage policy and host-side region/keyframe decisions.
The policy definition is this project's (the reference only passes knobs to its capture module,
src/selkies/selkies.py:2949-2951); the C oracle is cross-checked against an independent Python
statement written here, and the GPU kernel against the oracle in test_gpu_parity.py."""
import numpy as np
import pytest

from oracle import cpu_oracle as o
from selkies_gstreamer_b200 import _abi
from selkies_gstreamer_b200.damage_policy import DamagePolicy, Region, regions_from_mask

D, HOT, PAINT = _abi.SK_POLICY_DAMAGED, _abi.SK_POLICY_HOT, _abi.SK_POLICY_PAINT_OVER


class PyBlock:
    """Independent statement of one block's policy."""

    def __init__(self, thr, dur, trig):
        self.thr, self.dur, self.trig = thr, dur, trig
        self.streak = self.hot = self.still = 0
        self.dirty = False

    def step(self, damaged):
        flags = 0
        if damaged:
            self.streak = min(self.streak + 1, 255)
            self.still, self.dirty = 0, True
            if self.streak >= self.thr or self.hot > 0:
                self.hot = self.dur
            flags |= D
        else:
            self.streak = 0
            self.still = min(self.still + 1, 255)
            self.hot = max(self.hot - 1, 0)
            if self.dirty and self.hot == 0 and self.still >= self.trig:
                flags |= PAINT
                self.dirty = False
        if self.hot > 0:
            flags |= HOT
        return flags


def run_block(pattern, thr=10, dur=20, trig=15):
    t = o.PolicyTracker(1, 1, thr, dur, trig)
    return [int(t.update(np.array([[d]], np.uint8))[0, 0]) for d in pattern]


def test_single_change_is_repainted_once_after_trigger_still_frames():
    out = run_block([1] + [0] * 40)
    assert out[0] == D
    assert out[1:15] == [0] * 14            # still frames 1..14
    assert out[15] == PAINT                 # 15th still frame
    assert out[16:] == [0] * 25             # only once


def test_persistent_motion_turns_hot_then_cools_and_is_repainted():
    out = run_block([1] * 12 + [0] * 40)
    assert out[:9] == [D] * 9
    assert out[9:12] == [D | HOT] * 3       # 10th consecutive damaged frame onwards
    assert out[12:31] == [HOT] * 19         # cooling: 20 frames of hot after the last change
    assert out[31] == PAINT                 # hot expired and >= 15 still frames
    assert out[32:] == [0] * 20


def test_change_while_hot_refreshes_the_countdown():
    out = run_block([1] * 10 + [0] * 5 + [1] + [0] * 19 + [0, 0])
    assert out[15] == D | HOT
    assert out[16:35] == [HOT] * 19 and out[35] == PAINT


def test_never_damaged_block_is_never_flagged():
    assert run_block([0] * 50) == [0] * 50


@pytest.mark.parametrize("knobs", [(10, 20, 15), (1, 1, 1), (3, 40, 5), (255, 255, 255)])
def test_c_oracle_equals_python_statement_on_random_maps(knobs):
    rng = np.random.default_rng(sum(knobs))
    mbw, mbh = 7, 5
    t = o.PolicyTracker(mbw, mbh, *knobs)
    blocks = [PyBlock(*knobs) for _ in range(mbw * mbh)]
    p = rng.random(mbw * mbh) ** 2          # per-block activity
    for frame in range(300):
        m = (rng.random(mbw * mbh) < p * (0.2 + 0.8 * ((frame // 50) % 2))).astype(np.uint8)
        got = t.update(m.reshape(mbh, mbw)).reshape(-1)
        want = np.array([b.step(bool(d)) for b, d in zip(blocks, m)], np.uint8)
        assert (got == want).all(), frame


# ---- regions ------------------------------------------------------------------------
def paint(regions, mbw, mbh, w, h):
    img = np.zeros((h, w), np.int32)
    for r in regions:
        img[r.y:r.y + r.h, r.x:r.x + r.w] += 1
    return img


@pytest.mark.parametrize("seed", range(6))
def test_regions_cover_exactly_the_mask_without_overlap(seed):
    rng = np.random.default_rng(seed)
    w, h = 1366, 760            # ragged: last column 6 px, last row 8 px
    mbw, mbh = (w + 15) // 16, (h + 15) // 16
    mask = rng.random((mbh, mbw)) < 0.08
    mask[5:9, 10:20] = True
    mask[-1, -1] = True
    regs = regions_from_mask(mask, w, h, max_regions=10_000)
    img = paint(regs, mbw, mbh, w, h)
    assert img.max() == 1
    want = np.kron(mask.astype(np.int32), np.ones((16, 16), np.int32))[:h, :w]
    assert (img == want).all()
    assert all(r.x + r.w <= w and r.y + r.h <= h and r.w > 0 and r.h > 0 for r in regs)


def test_equal_runs_in_consecutive_rows_merge_vertically():
    mask = np.zeros((10, 12), bool)
    mask[2:6, 3:8] = True           # 4 rows x 5 blocks -> one rectangle
    mask[7, 0:2] = True
    assert regions_from_mask(mask, 192, 160) == [Region(48, 32, 80, 64), Region(0, 112, 32, 16)]


def test_too_many_regions_fall_back_to_row_band_boxes():
    mask = np.zeros((20, 30), bool)
    mask[2:18:2, 3:27:2] = True
    regs = regions_from_mask(mask, 480, 320, max_regions=16)
    assert 1 <= len(regs) <= 16
    img = paint(regs, 30, 20, 480, 320)
    assert img.max() == 1                                         # bands do not overlap
    want = np.kron(mask.astype(np.int32), np.ones((16, 16), np.int32))
    assert (img[want == 1] == 1).all()                            # every damaged block is covered
    assert all(r.x == 48 and r.w == 23 * 16 for r in regs)        # each band's box spans columns 3..25
    one = regions_from_mask(mask, 480, 320, max_regions=1)
    assert one == [Region(48, 32, 368, 240)]


def test_decisions():
    pol = DamagePolicy(640, 360, keyframe_fraction=0.5, min_keyframe_distance=0)
    mbh, mbw = 23, 40
    ones = np.ones((mbh, mbw), np.uint8)
    zeros = np.zeros((mbh, mbw), np.uint8)
    d = pol.decide(ones, ones * D)
    assert d.force_keyframe and not d.skip_frame and d.damage_regions == [Region(0, 0, 640, 360)]
    d = pol.decide(zeros, zeros)
    assert not d.force_keyframe and d.skip_frame and d.damage_regions == [] and d.paint_over_regions == []
    m = zeros.copy(); m[2:4, 5:9] = 1
    p = zeros.copy(); p[2:4, 5:9] = D | HOT; p[10, 0] = PAINT
    d = pol.decide(m, p)
    assert not d.force_keyframe and not d.skip_frame
    assert d.damage_regions == [Region(80, 32, 64, 32)]
    assert d.paint_over_regions == [Region(0, 160, 16, 16, "paint-over")]
    assert d.hot_fraction == pytest.approx(8 / 920)
    big = zeros.copy(); big[:12] = 1
    assert pol.decide(big, None).force_keyframe          # 480/920 >= 0.5
    pol.reset()
    assert pol.decide(zeros, None).force_keyframe        # first frame after a reset
    with pytest.raises(ValueError):
        DamagePolicy(64, 64, keyframe_fraction=0.0)


def test_keyframes_only_on_the_rising_edge_and_not_too_often():
    """Full-screen video / games / scrolling damage every frame: the reference runs an infinite GOP
    under a tight VBV (gstwebrtc_app.py:1227-1262), so steady damage must not turn every frame into an IDR."""
    mbh, mbw = 23, 40
    ones = np.ones((mbh, mbw), np.uint8)
    zeros = np.zeros((mbh, mbw), np.uint8)
    pol = DamagePolicy(640, 360, keyframe_fraction=0.6, min_keyframe_distance=30)
    forced = [pol.decide(ones, None).force_keyframe for _ in range(100)]
    assert forced[0] and sum(forced) == 1                  # only the implicit first frame
    # still for a moment, then a scene change: one request on the rising edge ...
    pol.decide(zeros, None)
    seq = [pol.decide(ones, None).force_keyframe for _ in range(5)]
    assert seq == [True, False, False, False, False]
    # ... but a second flip right after is too soon (min distance 30 frames)
    pol.decide(zeros, None)
    assert not pol.decide(ones, None).force_keyframe
    for _ in range(30):
        pol.decide(zeros, None)
    assert pol.decide(ones, None).force_keyframe
    with pytest.raises(ValueError):
        DamagePolicy(64, 64, min_keyframe_distance=-1)
