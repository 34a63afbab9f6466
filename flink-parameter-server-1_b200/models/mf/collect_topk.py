"""Parallelism-1 merge of per-worker partial top-Ks with a per-user "seen" filter.

Parity: ``utils/CollectTopKFromEachWorker.scala:23-74``.
"""
from __future__ import annotations

from collections import deque
from typing import Dict, List, Optional, Tuple

from ...entities import Either


class CollectTopKFromEachWorker:
    def __init__(self, K: int, memory: int, worker_parallelism: int):
        self.K, self.memory, self.W = K, memory, worker_parallelism
        self.outputs: Dict[float, List[Optional[list]]] = {}
        self.seen_set: Dict[int, set] = {}
        self.seen_list: Dict[int, deque] = {}

    def flat_map(self, value):
        """``value`` is one element of the PS job's output stream; PS outputs (``Right``) are
        dropped (``:57``).  Yields ``(user, item, timestamp, [(score, item), ...])``."""
        if value.is_right:
            return
        rich, topk = value.value
        all_topk = self.outputs.setdefault(rich.rating_id, [None] * self.W)
        all_topk[rich.target_worker] = list(topk)
        if all(t is not None for t in all_topk):
            seen = self.seen_set.setdefault(rich.user, set())
            merged = [x for part in all_topk for x in part if x[1] not in seen]
            merged.sort(key=lambda t: (-t[0], t[1]))
            yield (rich.user, rich.item, rich.timestamp, merged[: self.K])
            del self.outputs[rich.rating_id]
            seen.add(rich.item)
            lst = self.seen_list.setdefault(rich.user, deque())
            lst.append(rich.item)
            if self.memory > -1 and len(lst) > self.memory:
                seen.discard(lst.popleft())

    def process(self, stream) -> list:
        out = []
        for v in stream:
            out.extend(self.flat_map(v))
        return out
