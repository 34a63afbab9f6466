"""Manual back-pressure: sleep once per subtask before the first element goes through.

Parity: ``M/utils/FlinkSleepBlocker.scala:7-38``.
"""
from __future__ import annotations

import time

from ..parallel.streams import DataStream, as_stream


def block(stream, milliseconds: float) -> DataStream:
    s = stream if isinstance(stream, DataStream) else as_stream(stream, 1)

    def blocked(p):
        slept = False
        for rec in p:
            if not slept:
                time.sleep(milliseconds / 1000.0)
                slept = True
            yield rec

    return DataStream([blocked(p) for p in s.partitions()])
