"""Logging, counters and profiler ranges.

The reference only has slf4j ``log.debug`` per message (``M/FlinkParameterServer.scala:237``, ``:247``,
``:284``) and no metrics.  Here: a package logger (``FPS_LOG=DEBUG`` to enable per-event logging in
the host engine), a counters report that merges the host engine's ``EngineStats`` with the device
``FpsFlow`` block (pulls, pushes, remote fraction, limiter stalls, max in flight, NaNs), and NVTX
ranges so the fused launches show up named in Nsight timelines.
"""
from __future__ import annotations

import contextlib
import logging
import os
from typing import Dict, Optional

LOGGER = logging.getLogger("fps_b200")
if not LOGGER.handlers:
    _h = logging.StreamHandler()
    _h.setFormatter(logging.Formatter("%(asctime)s %(levelname)s %(name)s: %(message)s"))
    LOGGER.addHandler(_h)
    LOGGER.setLevel(getattr(logging, os.environ.get("FPS_LOG", "WARNING").upper(), logging.WARNING))


@contextlib.contextmanager
def nvtx_range(name: str):
    """``with nvtx_range("mf.step"):`` - a no-op without CUDA."""
    pushed = False
    try:
        import torch
        if torch.cuda.is_available():
            torch.cuda.nvtx.range_push(name)
            pushed = True
    except Exception:  # pragma: no cover
        pushed = False
    try:
        yield
    finally:
        if pushed:
            import torch
            torch.cuda.nvtx.range_pop()


def flow_report(stats: Dict[str, int], link_row_bytes: Optional[int] = None) -> Dict[str, float]:
    """Derived counters from a decoded ``FpsFlow`` (``ops.table.read_flow``)."""
    pulls = max(stats.get("pulls", 0), 1)
    physical = stats.get("combined_rows", 0) or stats.get("pulls", 0)
    out = {
        "records": stats.get("records", 0),
        "pulls": stats.get("pulls", 0),
        "pushes": stats.get("pushes", 0),
        "physical_rows_pulled": physical,
        "combine_ratio": stats.get("records", 0) / max(physical, 1),
        "remote_fraction": stats.get("remote_pulls", 0) / max(physical, 1),
        "limiter_stalls": stats.get("limiter_stalls", 0),
        "max_in_flight": stats.get("max_in_flight", 0),
        "nan_count": stats.get("nan_count", 0),
        "bad_ids": stats.get("bad_ids", 0),
        "mean_loss": stats.get("loss_sum", 0.0) / max(stats.get("records", 0), 1),
    }
    if link_row_bytes:
        out["nvlink_bytes_each_way"] = stats.get("remote_pulls", 0) * link_row_bytes
    return out


def engine_report(engine) -> Dict[str, int]:
    """Counters of a finished host-engine job (``transform(..., return_engine=True)``)."""
    d = engine.stats.as_dict()
    limiters = [lg for lg in engine.worker_logics if hasattr(lg, "max_in_flight")]
    if limiters:
        d["max_in_flight"] = max(lg.max_in_flight for lg in limiters)
        d["stalled_pulls"] = sum(getattr(lg, "stalled_pulls", 0) for lg in limiters)
    return d
