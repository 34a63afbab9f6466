#!/usr/bin/env python
"""Script equivalents of the reference's two notebooks.

``split``  (``R/Notebooks/Data_Manipulation.ipynb``): split a session log ``"<ts> <user> <item>"``
           into the first N train days and the following M test days, print the dataset statistics
           the notebook prints, write ``<out>_train`` (``ts user item``) and ``<out>_test``
           (``user item`` sorted by user).
``eval``   (``R/Notebooks/Tester.ipynb``): load dumped ``id;value`` user / item vectors, recommend
           the top-5 items per active user by brute-force dot products (optionally excluding the
           user's training items), report recall and precision against the test events.

    python drivers/session_tools.py split data_session --train-days 31 --test-days 14 --out data_session
    python drivers/session_tools.py eval UserVector.map ItemVector.map data_session_train data_session_test
"""
from __future__ import annotations

import argparse
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

from fps_b200.utils.checkpoint import read_id_value  # noqa: E402

DAY = 86400


def read_log(path):
    rows = []
    with open(path) as f:
        for line in f:
            p = line.split()
            if len(p) >= 3:
                rows.append((int(float(p[0])), int(p[1]), int(p[2])))
    return np.asarray(rows, dtype=np.int64).reshape(-1, 3)


def split(log_path: str, out_prefix: str, train_days: int = 31, test_days: int = 14) -> dict:
    data = read_log(log_path)
    t0 = data[0, 0]
    t1 = t0 + DAY * train_days
    t2 = t1 + DAY * test_days
    data = data[data[:, 0] <= t2]
    train, test = data[data[:, 0] <= t1], data[data[:, 0] > t1]
    per_user = np.unique(test[:, 1], return_counts=True)[1] if len(test) else np.zeros(0)
    stats = {
        "train_events": int(len(train)), "train_users": int(len(np.unique(train[:, 1]))),
        "train_items": int(len(np.unique(train[:, 2]))), "max_item_id": int(train[:, 2].max()) if len(train) else -1,
        "test_users": int(len(np.unique(test[:, 1]))),
        "users_in_both": int(len(np.intersect1d(train[:, 1], test[:, 1]))),
        "test_events_per_user": {q: float(np.quantile(per_user, q)) if len(per_user) else 0.0 for q in (0.0, 0.5, 0.75, 0.85, 1.0)},
    }
    np.savetxt(out_prefix + "_train", train, fmt="%d", delimiter=" ")
    order = np.argsort(test[:, 1], kind="stable")
    np.savetxt(out_prefix + "_test", test[order][:, 1:], fmt="%d", delimiter=" ")
    return stats


def evaluate(user_vec_path: str, item_vec_path: str, train_path: str, test_path: str, k: int = 5,
             min_item_frequency: int = 0, exclude_train: bool = True) -> dict:
    users, items = read_id_value(user_vec_path), read_id_value(item_vec_path)
    train = np.loadtxt(train_path, dtype=np.int64).reshape(-1, 3)
    test = np.loadtxt(test_path, dtype=np.int64).reshape(-1, 2)
    if min_item_frequency > 0:
        ids, cnt = np.unique(train[:, 2], return_counts=True)
        train = train[np.isin(train[:, 2], ids[cnt > min_item_frequency])]
    train_by_user, test_by_user = {}, {}
    for _ts, u, i in train:
        train_by_user.setdefault(int(u), set()).add(int(i))
    for u, i in test:
        test_by_user.setdefault(int(u), []).append(int(i))
    active_users = sorted(set(train_by_user) & set(test_by_user) & set(users))
    active_items = np.asarray(sorted((set(test[:, 1].tolist()) & set(train[:, 2].tolist())) & set(items)), dtype=np.int64)
    if not active_users or not len(active_items):
        return {"active_users": 0, "recall": 0.0, "precision": 0.0}
    V = np.stack([items[int(i)] for i in active_items])
    hit = count = 0
    for u in active_users:
        scores = V @ users[u]
        if exclude_train:
            scores = np.where(np.isin(active_items, list(train_by_user[u])), -np.inf, scores)
        top = set(active_items[np.argsort(-scores)[:k]].tolist())
        hit += len(top & set(test_by_user[u]))
        count += len(test_by_user[u])
    return {"active_users": len(active_users), "test_events": count, "hits": hit,
            "recall": hit / max(count, 1), "precision": hit / (len(active_users) * float(k))}


def main(argv=None):
    ap = argparse.ArgumentParser()
    sub = ap.add_subparsers(dest="cmd", required=True)
    s = sub.add_parser("split")
    s.add_argument("log")
    s.add_argument("--out", required=True)
    s.add_argument("--train-days", type=int, default=31)
    s.add_argument("--test-days", type=int, default=14)
    e = sub.add_parser("eval")
    e.add_argument("user_vectors")
    e.add_argument("item_vectors")
    e.add_argument("train")
    e.add_argument("test")
    e.add_argument("--k", type=int, default=5)
    e.add_argument("--min-item-frequency", type=int, default=0)
    a = ap.parse_args(argv)
    if a.cmd == "split":
        print(split(a.log, a.out, a.train_days, a.test_days))
    else:
        print(evaluate(a.user_vectors, a.item_vectors, a.train, a.test, a.k, a.min_item_frequency))
    return 0


if __name__ == "__main__":
    sys.exit(main())
