"""Which input shapes get a CUDA graph, and how many graphs stay alive.

A captured step owns its activations (a private memory pool of 1-2 GB at 768x1024), so graphs cannot simply be kept
per input shape: the reference's data sets hold images of many sizes and are read one image per batch
(``train.py:104-113``), and a cache keyed by shape alone would capture -- and keep -- one graph per image.  Policy:

* a shape is captured only when it has been used in more than ``warm`` CONSECUTIVE calls (fixed-size workloads
  reach this on their third step; a stream of changing sizes never does and stays eager);
* at most ``capacity`` captured entries are kept, least recently used first out (``CAN_B200_GRAPH_CACHE``,
  default 4); dropping an entry releases its graph and static buffers;
* a change of ``signature`` (the caller's notion of "the captured constants are stale") drops everything and
  restarts the warm-up (the first calls after a change redo lazy set-up work that must not be captured).

Pure host logic, no CUDA calls: the entries are opaque to this class.
"""
from __future__ import annotations

import os
from collections import OrderedDict


class GraphCache:
    def __init__(self, capacity: int | None = None, warm: int = 2):
        if capacity is None:
            capacity = int(os.environ.get("CAN_B200_GRAPH_CACHE", "4"))
        self.capacity = max(1, capacity)
        self.warm = warm
        self._entries: OrderedDict = OrderedDict()
        self._signature = None
        self._last_key = None
        self._streak = 0

    def __len__(self) -> int:
        return len(self._entries)

    def lookup(self, key, signature=None):
        """Registers one call with input ``key``; returns its captured entry or ``None`` (run eagerly / capture)."""
        if signature != self._signature:
            self._entries.clear()
            self._signature = signature
            self._last_key, self._streak = None, 0   # warm up again: the first calls after a change redo lazy set-up work
        if key == self._last_key:
            self._streak += 1
        else:
            self._last_key, self._streak = key, 1
        ent = self._entries.get(key)
        if ent is not None:
            self._entries.move_to_end(key)
        return ent

    def should_capture(self) -> bool:
        """True when the key of the last ``lookup`` has now been used more than ``warm`` times in a row."""
        return self._streak > self.warm

    def insert(self, key, entry) -> None:
        self._entries.pop(key, None)
        while len(self._entries) >= self.capacity:
            self._entries.popitem(last=False)
        self._entries[key] = entry

    def clear(self) -> None:
        self._entries.clear()
        self._last_key, self._streak = None, 0
