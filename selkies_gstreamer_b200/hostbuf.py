"""Pinning of the host buffers a stage is handed, so that both PCIe copies are plain DMA.

In the reference's pipeline the stage's input buffers come from `ximagesrc`'s pool of XShm
images and its output buffers from the pool negotiated with the encoder
(src/selkies/legacy/gstwebrtc_app.py:210-255 source, :941-994 linking): a handful of long-lived
buffers that come round again every few frames.  Page-locking such a buffer once
(`sk_host_register`, milliseconds) turns every later `cudaMemcpyAsync` from / into it into an
asynchronous DMA at the full PCIe rate; an unpinned (pageable) buffer is staged by the driver at a
fraction of that (bench.py: `e2e.unpinned_fps`).

`PinCache` keeps the most recently seen buffers pinned, notices when buffers do NOT come round
again (every frame a new address: pinning would cost more than it saves) and then switches itself
off, and unpins everything when the stage stops.  Pinned pages stay resident: the cache is bounded.
"""
from __future__ import annotations

from collections import OrderedDict

from . import _abi


class PinCache:
    def __init__(self, capacity: int = 16, give_up_after: int = 64):
        self.capacity = int(capacity)
        self.give_up_after = int(give_up_after)
        self._pinned: OrderedDict[int, int] = OrderedDict()     # address -> bytes, least recently used first
        self.hits = self.misses = self.failures = 0
        self.enabled = self.capacity > 0
        self._lib = None

    def _load(self):
        if self._lib is None:
            self._lib = _abi.load()
        return self._lib

    def ensure(self, ptr: int, nbytes: int) -> bool:
        """True if [ptr, ptr+nbytes) is page-locked when this returns."""
        if not self.enabled or not ptr:
            return False
        have = self._pinned.get(ptr)
        if have is not None and have >= nbytes:
            self._pinned.move_to_end(ptr)
            self.hits += 1
            return True
        lib = self._load()
        if have is not None:                       # same address, larger buffer now: pin again
            lib.sk_host_unregister(ptr)
            del self._pinned[ptr]
        self.misses += 1
        # Buffers that never come round again are not a pool: stop paying for page-locking.
        if self.misses >= self.give_up_after and self.hits < self.misses:
            self.disable()
            return False
        while len(self._pinned) >= self.capacity:
            old, _ = self._pinned.popitem(last=False)
            lib.sk_host_unregister(old)
        if lib.sk_host_register(ptr, nbytes) != _abi.SK_OK:
            # e.g. a range that overlaps one pinned by somebody else, or memory that cannot be locked
            self.failures += 1
            if self.failures >= 8 and self.hits == 0:
                self.disable()
            return False
        self._pinned[ptr] = nbytes
        return True

    def disable(self):
        self.release()
        self.enabled = False

    def release(self):
        """Unpin everything (stage stop / caps change)."""
        if self._pinned:
            lib = self._load()
            for ptr in self._pinned:
                lib.sk_host_unregister(ptr)
            self._pinned.clear()

    def __len__(self):
        return len(self._pinned)
