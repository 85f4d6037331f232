"""Host-side runtime logic that needs no GPU: the pin cache and the launch coalescer, against a fake library."""
import threading
import types

import pytest

from selkies_gstreamer_b200 import _abi, coalesce, hostbuf


class FakeLib:
    def __init__(self, fail_register=()):
        self.registered, self.calls, self.fail = {}, [], set(fail_register)
        self.batches = []
        self.lock = threading.Lock()

    def sk_host_register(self, ptr, n):
        self.calls.append(("reg", ptr, n))
        if ptr in self.fail:
            return _abi.SK_ERR_CUDA
        self.registered[ptr] = n
        return 0

    def sk_host_unregister(self, ptr):
        self.calls.append(("unreg", ptr))
        self.registered.pop(ptr, None)
        return 0

    def sk_batch_submit_host(self, handles, n, ins, stride, outs, fid):
        with self.lock:
            self.batches.append((n, fid, [handles[i] for i in range(n)]))
        return 0

    def sk_last_error(self, h):
        return b"fake"


@pytest.fixture
def fake(monkeypatch):
    lib = FakeLib()
    monkeypatch.setattr(_abi, "load", lambda: lib)
    return lib


def test_pin_cache_pins_once_and_evicts_least_recently_used(fake):
    c = hostbuf.PinCache(capacity=2)
    assert c.ensure(0x1000, 64) and c.ensure(0x1000, 64) and c.ensure(0x2000, 64)
    assert (c.hits, c.misses, len(c)) == (1, 2, 2)
    assert c.ensure(0x1000, 64)                      # touch: 0x2000 is now the oldest
    assert c.ensure(0x3000, 64)                      # evicts 0x2000
    assert set(fake.registered) == {0x1000, 0x3000} and ("unreg", 0x2000) in fake.calls
    assert c.ensure(0x1000, 128)                     # same address, larger buffer: pinned again
    assert fake.registered[0x1000] == 128
    c.release()
    assert not fake.registered and len(c) == 0


def test_pin_cache_gives_up_on_buffers_that_never_recur(fake):
    c = hostbuf.PinCache(capacity=4, give_up_after=10)
    for k in range(20):
        c.ensure(0x10000 + 0x1000 * k, 32)
    assert not c.enabled and not fake.registered     # switched itself off and unpinned everything
    assert c.ensure(0x999000, 32) is False


def test_pin_cache_survives_memory_that_cannot_be_pinned(monkeypatch):
    lib = FakeLib(fail_register={0x5000})
    monkeypatch.setattr(_abi, "load", lambda: lib)
    c = hostbuf.PinCache(4)
    assert c.ensure(0x5000, 16) is False and c.failures == 1 and c.enabled
    assert c.ensure(0x6000, 16) is True
    assert hostbuf.PinCache(0).ensure(0x7000, 16) is False      # capacity 0 = off


def _fake_session(handle, waited):
    s = types.SimpleNamespace(handle=types.SimpleNamespace(value=handle), device_index=0, width=64, height=64, in_stride=256,
                              in_format="BGRx", out_width=64, out_height=64, out_format="NV12", matrix=1, site=1,
                              scale_variant=0, n_inflight=2, flags=0, policy_knobs=(10, 20, 15), parity_switches=0)
    s.wait = lambda fid: waited.append((handle, fid)) or ("result", handle, fid)
    return s


def test_coalescer_shares_a_launch_between_threads_that_arrive_together(fake):
    waited = []
    sessions = [_fake_session(100 + i, waited) for i in range(4)]
    co = coalesce.LaunchCoalescer.for_session(sessions[0], window_us=200000)
    assert coalesce.LaunchCoalescer.for_session(sessions[1]) is co          # same key, same meeting point
    co.window_s = 0.2
    for _ in sessions:
        co.attach()
    results = {}
    gate = threading.Barrier(4)

    def work(s):
        gate.wait()
        results[s.handle.value] = co.convert(s, 0x1000 * s.handle.value, 256, 0x2000 * s.handle.value)

    ths = [threading.Thread(target=work, args=(s,)) for s in sessions]
    [t.start() for t in ths]
    [t.join(10) for t in ths]
    assert len(results) == 4 and sum(n for n, _, _ in fake.batches) == 4
    assert len(fake.batches) < 4, fake.batches                               # at least one launch was shared
    for h, (tag, hh, fid) in results.items():
        assert tag == "result" and hh == h and fid >= (1 << 40)
    # every member waited for the frame id of the batch it was in
    ids = {h: fid for (n, fid, hs) in fake.batches for h in hs}
    assert {h: f for h, f in waited} == ids
    for s in sessions:
        co.detach(s)                                                         # returns: nobody is still collecting


def test_coalescer_reports_a_failed_launch_to_every_member(monkeypatch):
    lib = FakeLib()
    lib.sk_batch_submit_host = lambda *a: _abi.SK_ERR_STATE
    monkeypatch.setattr(_abi, "load", lambda: lib)
    waited = []
    s = _fake_session(7, waited)
    s.width = 66                                      # another key: a fresh coalescer bound to this fake library
    co = coalesce.LaunchCoalescer.for_session(s, window_us=0)
    co.attach()
    with pytest.raises(coalesce.PreencodeError):
        co.convert(s, 1, 256, 2)
    assert not waited
    co.detach(s)
