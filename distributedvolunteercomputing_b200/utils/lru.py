"""Small LRU map with an eviction hook (used for the engine's per-(batch, frame size, requester) device state, which
would otherwise grow without bound on a long-lived worker that serves many requesters and video sizes)."""
from __future__ import annotations

import collections
from typing import Callable, Generic, Hashable, Iterator, Optional, TypeVar

V = TypeVar("V")


class LruCache(Generic[V]):
    def __init__(self, capacity: int, on_evict: Optional[Callable[[Hashable, V], None]] = None,
                 can_evict: Optional[Callable[[V], bool]] = None):
        assert capacity >= 1
        self.capacity = capacity
        self.on_evict = on_evict
        self.can_evict = can_evict or (lambda v: True)
        self._d: "collections.OrderedDict[Hashable, V]" = collections.OrderedDict()
        self.evictions = 0

    def get(self, key: Hashable) -> Optional[V]:
        v = self._d.get(key)
        if v is not None:
            self._d.move_to_end(key)
        return v

    def put(self, key: Hashable, value: V) -> None:
        self._d[key] = value
        self._d.move_to_end(key)
        if len(self._d) <= self.capacity:
            return
        for k in list(self._d):                       # oldest first; pinned entries are skipped, never the new one
            if len(self._d) <= self.capacity:
                break
            if k == key or not self.can_evict(self._d[k]):
                continue
            v = self._d.pop(k)
            self.evictions += 1
            if self.on_evict is not None:
                self.on_evict(k, v)

    def values(self) -> Iterator[V]:
        return iter(list(self._d.values()))

    def clear(self) -> None:
        self._d.clear()

    def __len__(self) -> int:
        return len(self._d)

    def __contains__(self, key: Hashable) -> bool:
        return key in self._d
