#!/usr/bin/env python
"""Headline benchmark: large-scale SGD matrix factorisation on the parameter server
(BASELINE.json config 5: 100M users x 10M items, rank 128, pull/push throughput at 1/2/4/8 GPUs).

    python bench.py --gpus N --steps K --warmup W [--impl ours|nccl_a2a|cpu|reference]

One step = one fused pull-compute-push pass over ``--batch`` ratings per GPU (weak scaling: the
per-GPU batch is fixed, the 100M x 10M model is sharded over the N GPUs).  Every rating pulls
one item row from its owner shard, updates the worker-resident user row and pushes the item
delta back, so ratings/s == pull rows/s == push rows/s.  Prints ONE JSON line (rank 0).
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "mf_sgd_ratings_per_sec"


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "nccl_a2a", "cpu", "reference"])
    ap.add_argument("--batch", type=int, default=1 << 22, help="ratings per GPU per step")
    ap.add_argument("--users", type=int, default=100_000_000)
    ap.add_argument("--items", type=int, default=10_000_000)
    ap.add_argument("--rank", type=int, default=128, help="number of factors")
    ap.add_argument("--item-dist", default="movielens", choices=["movielens", "uniform"])
    ap.add_argument("--variant", default="auto", choices=["auto", "direct", "combine"],
                    help="auto = direct on one GPU, combine (hot-key pre-combining) when items are sharded over several")
    ap.add_argument("--pull-limit", type=int, default=0)
    ap.add_argument("--lr", type=float, default=0.01)
    ap.add_argument("--pool", type=int, default=4, help="distinct pre-generated batches rotated through")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--unroll", type=int, default=0, help="kernel tuning: ratings in flight per sub-warp")
    ap.add_argument("--blocks-per-sm", type=int, default=0, help="kernel tuning: persistent CTAs per SM")
    ap.add_argument("--min-blocks", type=int, default=0, help="kernel tuning: 8 = 32-register build of unroll 1")
    return ap.parse_args()


def reference_arm(a):
    """The reference is a Scala/sbt project on Flink 1.4 (no setup.py / pyproject, no JVM, sbt or
    jars in this image, no network) - it cannot be pip-installed or run; see DESIGN.md."""
    ref_dir = os.path.join(ROOT, "baseline", "_ref")
    why = ("reference is a Scala 2.11 / sbt / Apache Flink 1.4 JVM project: `pip install /root/reference` "
           "fails (no setup.py or pyproject.toml) and the image has no java/sbt/scala nor the Flink/Breeze jars")
    if os.path.isdir(ref_dir) and os.listdir(ref_dir):
        why += "; baseline/_ref holds no runnable artefact"
    print(json.dumps({"impl": "reference", "unavailable": why}))
    return 0


class ClockSampler:
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
         "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        self.gpu_index = gpu_index
        self.proc = None

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100",
                 "-i", str(self.gpu_index)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        except Exception:
            self.proc = None

    def stop(self) -> dict:
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            out, _ = self.proc.communicate(timeout=5)
        except Exception:
            self.proc.kill()
            out = ""
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in out.strip().splitlines():
            f = [x.strip() for x in line.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1])); mx.append(float(f[2]))
            except ValueError:
                continue
            for name, val in zip(names, f[5:9]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "samples": len(sm), "reasons": sorted(reasons)}


def make_batch(torch, a, rank, world, seed):
    """Synthetic MovieLens-shaped ratings for this rank's users (user % world == rank).
    Item popularity: id = floor(n_items * u^3) - the most popular item draws ~0.46 % of the
    ratings, like MovieLens' head; ``uniform`` for the unskewed sweep."""
    g = torch.Generator(device="cpu").manual_seed(seed)
    per_rank_users = (a.users + world - 1) // world
    slot = torch.randint(0, per_rank_users, (a.batch,), generator=g, dtype=torch.int64)
    users = (slot * world + rank).clamp_(max=a.users - 1)
    users = users - ((users - rank) % world)         # keep ownership after the clamp
    x = torch.rand(a.batch, generator=g, dtype=torch.float64)
    if a.item_dist == "movielens":
        x = x ** 3
    items = (x * a.items).long().clamp_(max=a.items - 1)
    ratings = torch.rand(a.batch, generator=g, dtype=torch.float32)
    return users.int(), items.int(), ratings


def main():
    a = parse()
    if a.impl == "reference":
        return reference_arm(a)
    import torch
    if a.impl == "cpu":
        return cpu_arm(a, torch)
    import torch.distributed as dist
    from fps_b200.parallel.device_context import init_context
    ctx = init_context()
    rank, world = ctx.rank, ctx.world_size
    assert world == a.gpus, f"--gpus {a.gpus} but WORLD_SIZE={world}; launch with torchrun for N>1"
    dev = ctx.device

    pool_h = [make_batch(torch, a, rank, world, 1234 + 97 * rank + k) for k in range(a.pool)]
    pool_h = [tuple(t.pin_memory() for t in b) for b in pool_h]
    pool_d = [tuple(t.to(dev, non_blocking=True) for t in b) for b in pool_h]

    if a.impl == "ours":
        from fps_b200.ops import _native
        if not _native.available():
            import build_native
            build_native.build()
        from fps_b200.models.mf.device import DeviceMF
        model = DeviceMF(a.users, a.items, a.rank, -0.01, 0.01, a.lr, pull_limit=a.pull_limit, seed=7,
                         lazy=False, variant=a.variant, ctx=ctx)
        model.params.unroll, model.params.blocks_per_sm = a.unroll, a.blocks_per_sm
        model.params.reserved = a.min_blocks
        step = lambda b: model.step(*b)                      # noqa: E731
        e2e_step = lambda b: model.step_from_host(*b)        # noqa: E731
        # direct: k_mf_sgd; combine: k_iota + k_gather_payload + k_mf_sgd_combine (the cub radix-sort
        # passes in between are library kernels and are not counted)
        launches_per_step = 3 if model.variant == "combine" else 1
    else:
        from fps_b200.parallel.nccl_baseline import AllToAllPS
        model = AllToAllPS(a.users, a.items, a.rank, a.lr, device=dev, seed=7)
        step = lambda b: model.step(*b)                      # noqa: E731

        def e2e_step(b):
            d = tuple(t.to(dev, non_blocking=True) for t in b)
            model.step(*d)
            return float(model.loss_sum.item()) / max(model.records, 1)
        launches_per_step = 0            # none of our kernels: torch ops + NCCL only

    def sync_all():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, batches):
        for k in range(a.warmup):
            fn(batches[k % len(batches)])
        sync_all()
        sampler = ClockSampler(dev.index if dev.index is not None else 0)
        if rank == 0:
            sampler.start()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for k in range(a.steps):
            fn(batches[(a.warmup + k) % len(batches)])
        e1.record()
        sync_all()
        clocks = sampler.stop() if rank == 0 else None
        ms = torch.tensor([e0.elapsed_time(e1)], device=dev, dtype=torch.float64)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms.item()), clocks

    ms_total, clocks = timed(step, pool_d)
    ms_per_step = ms_total / a.steps
    value = a.batch * world / (ms_per_step * 1e-3)

    e2e = None
    if not a.no_e2e:
        e2e_ms, _ = timed(e2e_step, pool_h)
        e2e = {"value": a.batch * world / (e2e_ms / a.steps * 1e-3), "unit": "ratings/s",
               "h2d_bytes_per_step": a.batch * 12, "d2h_bytes_per_step": 16 if a.impl == "ours" else 8,
               "ms_per_step": e2e_ms / a.steps}

    if a.impl == "ours":
        st = model.stats(check_nan=False)
        mse = st["loss_sum"] / max(st["records"], 1)
        # a diverged model makes the kernel skip its (non-finite) updates and look fast: say so in the line
        extra = {"nan_count": st["nan_count"], "training_valid": st["nan_count"] == 0 and st["bad_ids"] == 0,
                 "bad_ids": st["bad_ids"], "mse": mse,
                 "remote_pull_fraction": (st["remote_pulls"] / max(st["pulls"], 1)) if world > 1 else 0.0}
    else:
        extra = {"mse": float(model.loss_sum.item()) / max(model.records, 1)}

    # roofline bookkeeping: per rating, item row read + item row RMW by the atomic
    # + user row read + write = 5 row transfers locally; (G-1)/G of pulls and pushes cross NVLink.
    row_bytes = a.rank * 4
    hbm_bytes = 5 * row_bytes * a.batch
    link_bytes = 2 * row_bytes * a.batch * (world - 1) / world
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    hbm_peak = peaks.get("hbm_gbs", 6650.0)
    t_s = ms_per_step * 1e-3
    # "logical" = 5 row transfers per rating as if nothing hit L2; popular items do (the ncu capture in
    # profiles/r1_mf shows 7.8 GB of real DRAM traffic for 10.7 GB logical = 92 % of the copy bandwidth),
    # so the logical figure can exceed the measured peak
    roof = {"hbm_gbs_logical": hbm_bytes / t_s / 1e9,
            "hbm_logical_over_measured_copy_peak": hbm_bytes / t_s / 1e9 / hbm_peak,
            "nvlink_gbs_per_dir_achieved": link_bytes / t_s / 1e9,
            "nvlink_frac_of_770": link_bytes / t_s / 1e9 / 770.0 if world > 1 else None}

    if rank == 0:
        print(json.dumps({
            "metric": METRIC, "value": value, "unit": "ratings/s", "n_gpus": world, "steps": a.steps,
            "warmup": a.warmup, "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "fp32", "data": "synthetic", "impl": a.impl,
            "config": {"model": f"mf_sgd_rank{a.rank}_{a.users}x{a.items}",
                       "global_batch": a.batch * world, "seq_len": None,
                       "parallelism": f"w{world}xps{world}", "item_dist": a.item_dist,
                       "variant": model.variant if a.impl == "ours" else None, "pull_limit": a.pull_limit,
                       "l2": f"working set per step {hbm_bytes * 2 // 5 >> 20} MiB of random rows >> 126 MB L2; "
                             f"{a.pool} distinct batches rotated"},
            "clocks": clocks, "e2e": e2e, "gpu_launches": launches_per_step * a.steps,
            "pull_rows_per_sec": value, "push_rows_per_sec": value, "roofline": roof, **extra}))
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    return 0


def cpu_arm(a, torch):
    """BL2: the multi-threaded C++ restatement of the PS loop on the box's CPUs (stands in for
    the unbuildable Flink run; labelled "CPU restatement, not Flink")."""
    from fps_b200.parallel import cpu_baseline
    users = min(a.users, 2_000_000)
    items = min(a.items, 1_000_000)
    batch = min(a.batch, 1 << 18)
    threads = max(2, (os.cpu_count() or 8) // 2 * 2)
    res = cpu_baseline.bench_mf(users, items, a.rank, batch, a.steps, a.warmup, workers=threads // 2,
                                shards=threads // 2, lr=a.lr)
    print(json.dumps({"metric": METRIC, "value": res["ratings_per_sec"], "unit": "ratings/s", "n_gpus": 0,
                      "steps": a.steps, "warmup": a.warmup, "ms_per_step": res["ms_per_step"],
                      "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "fp64",
                      "data": "synthetic", "impl": "cpu_restatement_not_flink",
                      "config": {"model": f"mf_sgd_rank{a.rank}_{users}x{items}", "global_batch": batch,
                                 "parallelism": f"w{threads // 2}xps{threads // 2} threads",
                                 "note": "table shrunk to fit host RAM; per-record work identical"},
                      "gpu_launches": 0}))
    return 0


if __name__ == "__main__":
    sys.exit(main())
