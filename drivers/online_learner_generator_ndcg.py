#!/usr/bin/env python
"""Driver: prequential recommend-then-learn with a per-period nDCG report.

Parity: ``T/matrix/factorization/PSOnlineMatrixFactorizationAndTopKGeneratorTest.scala:27-52`` - reads
CSV lines ``timestamp,user,item``, runs the online learner + top-K generator and writes nDCG /
hit-rate per period (default one day = 86 400 s) as CSV (``nDCGSink.nDCGPeriodsToCsv``).

    python drivers/online_learner_generator_ndcg.py INPUT.csv OUT.csv [--engine host|device] [...]
"""
from __future__ import annotations

import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

from fps_b200.models.mf import LI, Rating, from_string, ps_online_learner_and_generator  # noqa: E402
from fps_b200.utils.evaluation import NDCGSink  # noqa: E402


def read_events(path):
    with open(path) as f:
        for line in f:
            fields = line.strip().split(",")
            if len(fields) >= 3:
                yield Rating(int(fields[1]), int(fields[2]), 1.0, int(float(fields[0])))


def main(argv=None):
    ap = argparse.ArgumentParser()
    ap.add_argument("input")
    ap.add_argument("output")
    ap.add_argument("--engine", default="host", choices=["host", "device"])
    ap.add_argument("--period", type=int, default=86400)
    ap.add_argument("--num-factors", type=int, default=10)
    ap.add_argument("--range-min", type=float, default=-0.01)
    ap.add_argument("--range-max", type=float, default=0.01)
    ap.add_argument("--learning-rate", type=float, default=0.4)
    ap.add_argument("--negative-sample-rate", type=int, default=9)
    ap.add_argument("--user-memory", type=int, default=128)
    ap.add_argument("--K", type=int, default=100)
    ap.add_argument("--worker-k", type=int, default=75)
    ap.add_argument("--bucket-size", type=int, default=100)
    ap.add_argument("--pruning", default="li:5:2.5", help="length | coord | incr:N | lc:T | li:N:T")
    ap.add_argument("--pull-limit", type=int, default=800)
    ap.add_argument("--worker-parallelism", type=int, default=4)
    ap.add_argument("--ps-parallelism", type=int, default=4)
    ap.add_argument("--batch-size", type=int, default=256, help="device engine: queries answered per step")
    a = ap.parse_args(argv)
    events = list(read_events(a.input))
    if a.engine == "device":
        from fps_b200.models.mf.device_topk import ps_online_learner_and_generator_device
        n_users = max(r.user for r in events) + 1
        n_items = max(r.item for r in events) + 1
        res, _ = ps_online_learner_and_generator_device(
            events, n_users, n_items, batch_size=a.batch_size, num_factors=a.num_factors,
            range_min=a.range_min, range_max=a.range_max, learning_rate=a.learning_rate,
            negative_sample_rate=a.negative_sample_rate, user_memory=a.user_memory, K=a.K,
            worker_k=a.worker_k, pull_limit=a.pull_limit)
    else:
        res = ps_online_learner_and_generator(
            events, num_factors=a.num_factors, range_min=a.range_min, range_max=a.range_max,
            learning_rate=a.learning_rate, negative_sample_rate=a.negative_sample_rate,
            user_memory=a.user_memory, K=a.K, worker_k=a.worker_k, bucket_size=a.bucket_size,
            pruning_algorithm=from_string(a.pruning), pull_limit=a.pull_limit,
            worker_parallelism=a.worker_parallelism, ps_parallelism=a.ps_parallelism, iteration_wait_time=100)
    sink = NDCGSink(period_length=a.period).consume(res)
    sink.write_csv(a.output)
    print(f"{len(events)} events: nDCG {sink.ndcg:.4f}, hit rate {sink.hit_rate:.4f} -> {a.output}")
    return 0


if __name__ == "__main__":
    sys.exit(main())
