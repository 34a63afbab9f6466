#!/usr/bin/env python
"""Driver: implicit-feedback online MF on a LastFM-style log.

Parity: ``T/matrix/factorization/PSOnlineMatrixFactorizationImplicitTest.scala:31-98`` - reads lines
``"<timestamp> <user> <item>"``, turns each into an implicit rating 1.0, trains online MF with
negative sampling, and writes the user and item factors as ``id;value`` lines (one per coordinate).

    python drivers/online_mf_implicit.py INPUT USER_OUT ITEM_OUT [--engine host|device] [...]
"""
from __future__ import annotations

import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

from fps_b200 import last_write_wins, lefts, rights  # noqa: E402
from fps_b200.models.mf import Rating, ps_online_mf  # noqa: E402
from fps_b200.utils.checkpoint import write_id_value  # noqa: E402


def read_lastfm(path):
    with open(path) as f:
        for line in f:
            fields = line.split()
            if len(fields) >= 3:
                yield Rating(int(fields[1]), int(fields[2]), 1.0, int(float(fields[0])))


def main(argv=None):
    ap = argparse.ArgumentParser()
    ap.add_argument("input")
    ap.add_argument("user_out")
    ap.add_argument("item_out")
    ap.add_argument("--engine", default="host", choices=["host", "device"])
    ap.add_argument("--num-factors", type=int, default=10)
    ap.add_argument("--range-min", type=float, default=-0.01)
    ap.add_argument("--range-max", type=float, default=0.01)
    ap.add_argument("--learning-rate", type=float, default=0.4)
    ap.add_argument("--negative-sample-rate", type=int, default=9)
    ap.add_argument("--user-memory", type=int, default=128)
    ap.add_argument("--pull-limit", type=int, default=1500)
    ap.add_argument("--worker-parallelism", type=int, default=4)
    ap.add_argument("--ps-parallelism", type=int, default=4)
    ap.add_argument("--iteration-wait-time", type=int, default=10000)
    a = ap.parse_args(argv)
    ratings = list(read_lastfm(a.input))
    if a.engine == "device":
        from fps_b200.models.mf.device import ps_online_mf_device
        out = ps_online_mf_device(ratings, num_factors=a.num_factors, range_min=a.range_min,
                                  range_max=a.range_max, learning_rate=a.learning_rate,
                                  negative_sample_rate=a.negative_sample_rate, pull_limit=a.pull_limit)
    else:
        out = ps_online_mf(ratings, num_factors=a.num_factors, range_min=a.range_min, range_max=a.range_max,
                           learning_rate=a.learning_rate, negative_sample_rate=a.negative_sample_rate,
                           user_memory=a.user_memory, pull_limit=a.pull_limit,
                           worker_parallelism=a.worker_parallelism, ps_parallelism=a.ps_parallelism,
                           iteration_wait_time=a.iteration_wait_time)
    users, items = last_write_wins(lefts(out)), last_write_wins(rights(out))
    write_id_value(a.user_out, users)
    write_id_value(a.item_out, items)
    print(f"{len(ratings)} events, {len(users)} user vectors -> {a.user_out}, {len(items)} item vectors -> {a.item_out}")
    return 0


if __name__ == "__main__":
    sys.exit(main())
