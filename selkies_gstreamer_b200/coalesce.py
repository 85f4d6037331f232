"""One kernel launch for the frames that several stage instances submit at the same moment.

The reference builds several pipelines in one process (src/selkies/legacy/webrtc.py:684-685 creates
two `GSTWebRTCApp`s; one per browser session in a multi-user host), each with its own convert stage
on its own streaming thread (gstwebrtc_app.py:611-617).  Alone, a 4K frame keeps the GPU busy for
~16 us of which ~7 us are launch latency and tail; converted together, N frames cost one launch
(`sk_batch_submit_host`).  This module is the meeting point: `transform()` calls of stages that
share a GPU and a configuration hand their buffers over, the first caller waits a short window
for the others, submits the batch, and every caller then waits for its own frame.

It is opt-in (stage property `coalesce`): a host whose stages are bound by PCIe gains nothing —
a batch's uploads all precede its one launch, so read-back of one frame no longer overlaps upload of
the next — while a host whose frames are already on the GPU, or are small, does.
"""
from __future__ import annotations

import ctypes
import threading
import time

from . import _abi
from .session import PreencodeError, PreencodeSession

_FRAME_ID_BASE = 1 << 40        # apart from the ids a session hands out itself (0, 1, 2, ...)


class _Batch:
    """Completion bookkeeping of one submitted batch (so a leaving member can wait for its peers)."""

    def __init__(self, n: int):
        self.remaining = n
        self.cv = threading.Condition()

    def finish_one(self):
        with self.cv:
            self.remaining -= 1
            if self.remaining <= 0:
                self.cv.notify_all()

    def wait_all(self, timeout: float = 5.0):
        with self.cv:
            self.cv.wait_for(lambda: self.remaining <= 0, timeout)


class _Request:
    __slots__ = ("session", "in_ptr", "in_stride", "out_ptr", "event", "lead", "frame_id", "error", "batch")

    def __init__(self, session, in_ptr, in_stride, out_ptr):
        self.session, self.in_ptr, self.in_stride, self.out_ptr = session, in_ptr, in_stride, out_ptr
        self.event = threading.Event()
        self.lead = False
        self.frame_id = None
        self.error = None
        self.batch = None


def batch_key(s: PreencodeSession):
    """Sessions with equal keys may share a launch (sk_batch_submit's own rule)."""
    f = s.flags & (_abi.SK_FLAG_NO_DAMAGE | _abi.SK_FLAG_POLICY | _abi.SK_FLAG_SCALE_GENERIC)
    return (s.device_index, s.width, s.height, s.in_stride, s.in_format.upper()[:3], s.out_width, s.out_height,
            s.out_format, s.matrix, s.site, s.scale_variant, s.n_inflight, f, s.policy_knobs, s.parity_switches)


class LaunchCoalescer:
    _registry: dict = {}
    _registry_lock = threading.Lock()

    @classmethod
    def for_session(cls, s: PreencodeSession, window_us: int = 200) -> "LaunchCoalescer":
        key = batch_key(s)
        with cls._registry_lock:
            c = cls._registry.get(key)
            if c is None:
                c = cls._registry[key] = LaunchCoalescer(key, window_us)
            c.window_s = window_us * 1e-6
            return c

    def __init__(self, key, window_us: int = 200):
        self.key = key
        self.window_s = window_us * 1e-6
        self._lib = _abi.load()
        self._cv = threading.Condition()
        self._pending: list[_Request] = []
        self._leader_active = False
        self._members = 0
        self._expected = 1
        self._next_id = _FRAME_ID_BASE
        self._last_batch: dict[int, _Batch] = {}          # id(session) -> last batch it took part in
        self.launches = 0
        self.frames = 0

    # -- membership ------------------------------------------------------------
    def attach(self):
        with self._cv:
            self._members += 1
            self._expected = self._members

    def detach(self, session: PreencodeSession):
        """Call before closing `session`: waits until every member of the last batch it took part in
        has collected its frame (they share the leader's streams and completion event)."""
        with self._cv:
            self._members = max(0, self._members - 1)
            self._expected = max(1, min(self._expected, self._members))
            b = self._last_batch.pop(id(session), None)
        if b is not None:
            b.wait_all()

    # -- per frame ---------------------------------------------------------------
    def convert(self, session: PreencodeSession, in_ptr: int, in_stride: int, out_ptr: int | None):
        """Hand one frame over; returns the FrameResult once it is done (blocking, like transform())."""
        req = _Request(session, in_ptr, in_stride, out_ptr)
        with self._cv:
            self._pending.append(req)
            req.lead = not self._leader_active
            if req.lead:
                self._leader_active = True
            else:
                self._cv.notify_all()           # a leader may be counting arrivals
        if req.lead:
            self._collect_window()
            self._lead_one_batch()
        else:
            req.event.wait()
            if req.lead:                        # promoted: frames were already waiting, no window
                self._lead_one_batch()
        if req.error is not None:
            raise req.error
        try:
            return session.wait(req.frame_id)
        finally:
            req.batch.finish_one()

    def _collect_window(self):
        deadline = time.perf_counter() + self.window_s
        with self._cv:
            while len(self._pending) < self._expected:
                left = deadline - time.perf_counter()
                if left <= 0:
                    break
                self._cv.wait(left)

    def _lead_one_batch(self):
        with self._cv:
            batch = self._pending[:_abi.SK_MAX_BATCH]
            del self._pending[:_abi.SK_MAX_BATCH]
            # how many to expect next time: what actually came together now
            self._expected = max(1, min(len(batch), self._members))
            fid = self._next_id
            self._next_id += 1
        n = len(batch)
        err = None
        done = _Batch(n)
        try:
            handles = (ctypes.c_void_p * n)(*[r.session.handle.value for r in batch])
            ins = (ctypes.c_void_p * n)(*[r.in_ptr for r in batch])
            outs = (ctypes.c_void_p * n)(*[r.out_ptr or 0 for r in batch])
            rc = self._lib.sk_batch_submit_host(handles, n, ins, batch[0].in_stride, outs, fid)
            if rc != _abi.SK_OK:
                err = PreencodeError(rc, (self._lib.sk_last_error(batch[0].session.handle) or b"").decode())
        except Exception as e:                  # noqa: BLE001 - handed to every waiting caller
            err = e
        with self._cv:
            self.launches += 1
            self.frames += n
            for r in batch:
                r.frame_id, r.error, r.batch = fid, err, done
                self._last_batch[id(r.session)] = done
            # hand leadership on if more frames arrived meanwhile, else step down
            nxt = self._pending[0] if self._pending else None
            if nxt is None:
                self._leader_active = False
        if err is not None:
            done.remaining = 0                  # nobody will collect a frame of this batch
        for r in batch:
            r.event.set()
        if nxt is not None:
            nxt.lead = True
            nxt.event.set()
