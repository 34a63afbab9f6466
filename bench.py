#!/usr/bin/env python
"""Headline benchmark: large-scale SGD matrix factorisation on the parameter server
(BASELINE.json config 5: 100M users x 10M items, rank 128, pull/push throughput at 1/2/4/8 GPUs).

    python bench.py --gpus N --steps K --warmup W [--impl ours|nccl_a2a|cpu|reference]

One step = one fused pull-compute-push pass over ``--batch`` ratings per GPU (weak scaling: the
per-GPU batch is fixed, the 100M x 10M model is sharded over the N GPUs).  Every rating needs the
row of its item from the owner shard, updates the worker-resident user row and sends the item delta
back; the tile variant pulls / pushes one row per *run* of equal items, so rows/s is reported from
the kernel's own counters next to ratings/s.  Prints ONE JSON line (rank 0).

``vs_baseline``: BASELINE.md publishes no number and the reference (Scala / Flink) cannot run here,
so the ratio is taken against the NCCL all-to-all PS baseline (``--impl nccl_a2a``) measured in this
same process at the same users / items / rank / batch / N right before the timed arm.
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
DEFAULT_PULL_LIMIT = 262144
# measured on this pool (profiles/r1_2gpu/p2p_probe_2gpu.json): per-direction P2P rates of the two access
# types the fused kernels use; the NVLink roofline of a step is rows pulled / LOAD + rows pushed / RED
P2P_LOAD_GBS, P2P_RED_GBS = 759.0, 687.0
# measured with EVERY GPU pulling and pushing random 512-byte rows at once (scripts/p2p_symmetric_probe.py,
# profiles/r2_2gpu/p2p_symmetric_2gpu.json): payload per direction per GPU when requests / acknowledgements of
# one direction share the wire with the payload of the other - the pattern the fused kernels produce
P2P_SYMMETRIC_GBS = 638.6


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=100)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", default="ours", choices=["ours", "nccl_a2a", "cpu", "reference"])
    ap.add_argument("--batch", type=int, default=1 << 22, help="ratings per GPU per step")
    ap.add_argument("--users", type=int, default=100_000_000)
    ap.add_argument("--items", type=int, default=10_000_000)
    ap.add_argument("--rank", type=int, default=128, help="number of factors")
    ap.add_argument("--item-dist", default="movielens", choices=["movielens", "uniform"])
    ap.add_argument("--variant", default="auto", choices=["auto", "direct", "combine", "tiles"],
                    help="auto = DeviceMF's choice (tiles when items are sharded over several GPUs)")
    ap.add_argument("--pull-limit", type=int, default=DEFAULT_PULL_LIMIT,
                    help="bounded in-flight pulls per worker (device credit pool); 0 = unlimited")
    ap.add_argument("--baseline-steps", type=int, default=3,
                    help="timed steps of the in-process NCCL all-to-all baseline that fills vs_baseline (0 = skip)")
    ap.add_argument("--no-probe", action="store_true", help="skip the fp64 replay of a probe batch")
    ap.add_argument("--tile", type=int, default=0, help="tile variant: records per tile (512 / 1024 / 2048)")
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
    torch.cuda.synchronize()
    ready = torch.cuda.Event()
    ready.record()              # the resident batches are valid from here on

    def sync_all():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, batches, steps, warmup, before=None):
        for k in range(warmup):
            fn(batches[k % len(batches)])
        sync_all()
        if before is not None:
            before()
        sampler = ClockSampler(dev.index if dev.index is not None else 0)
        if rank == 0:
            sampler.start()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for k in range(steps):
            fn(batches[(warmup + k) % len(batches)])
        e1.record()
        sync_all()
        clocks = sampler.stop() if rank == 0 else None
        ms = torch.tensor([e0.elapsed_time(e1)], device=dev, dtype=torch.float64)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms.item()), clocks

    def a2a_arm(steps, warmup):
        from fps_b200.parallel.nccl_baseline import AllToAllPS
        m = AllToAllPS(a.users, a.items, a.rank, a.lr, device=dev, seed=7)
        ms, clk = timed(lambda b: m.step(*b), pool_d, steps, warmup)
        return m, ms, clk

    baseline = None
    if a.impl == "ours":
        from fps_b200.ops import _native
        if not _native.available():
            import build_native
            build_native.build()
        if a.baseline_steps > 0:
            m, ms, _ = a2a_arm(a.baseline_steps, 3)
            baseline = {"impl": "nccl_a2a", "value": a.batch * world / (ms / a.baseline_steps * 1e-3),
                        "ms_per_step": ms / a.baseline_steps, "steps": a.baseline_steps, "warmup": 3,
                        "same_config": True}
            del m
            torch.cuda.empty_cache()
            sync_all()
        from fps_b200.models.mf.device import DeviceMF
        model = DeviceMF(a.users, a.items, a.rank, -0.01, 0.01, a.lr, pull_limit=a.pull_limit, seed=7,
                         lazy=False, variant=a.variant, ctx=ctx)
        model.params.unroll, model.params.blocks_per_sm = a.unroll, a.blocks_per_sm
        model.params.reserved = a.tile if a.tile else a.min_blocks
        step = lambda b: model.step(*b, ready_event=ready)   # noqa: E731
        e2e_step = lambda b: model.step_from_host(*b)        # noqa: E731
        launches0 = [0]

        def before():
            model.reset_stats()
            launches0[0] = model.kernel_launches
        ms_total, clocks = timed(step, pool_d, a.steps, a.warmup, before)
        st = model.stats(check_nan=False)
        gpu_launches = model.kernel_launches - launches0[0]
    else:
        model, ms_total, clocks = a2a_arm(a.steps, a.warmup)
        step = lambda b: model.step(*b)                      # noqa: E731

        def e2e_step(b):
            d = tuple(t.to(dev, non_blocking=True) for t in b)
            model.step(*d)
            return float(model.loss_sum.item()) / max(model.records, 1)
        gpu_launches = 0                 # none of our kernels: torch ops + NCCL only
        st = None

    ms_per_step = ms_total / a.steps
    value = a.batch * world / (ms_per_step * 1e-3)

    e2e = None
    if not a.no_e2e:
        e2e_ms, _ = timed(e2e_step, pool_h, a.steps, a.warmup)
        e2e = {"value": a.batch * world / (e2e_ms / a.steps * 1e-3), "unit": "ratings/s",
               "h2d_bytes_per_step": a.batch * 12, "d2h_bytes_per_step": 16 if a.impl == "ours" else 8,
               "ms_per_step": e2e_ms / a.steps}
        if a.impl == "ours":
            e2e["last_mse"] = model.flush_host_steps()

    row_bytes = a.rank * 4
    t_s = ms_per_step * 1e-3
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    hbm_peak = peaks.get("hbm_gbs", 6650.0)
    if a.impl == "ours":
        # whole-job row counts from the kernels' own counters (timed region only)
        cnt = torch.tensor([st["pulls"], st["pushes"], st["remote_pulls"], st["remote_pushes"], st["records"],
                            st["nan_count"], st["bad_ids"]], device=dev, dtype=torch.float64)
        worst_link = torch.tensor([st["remote_pulls"] / P2P_LOAD_GBS + st["remote_pushes"] / P2P_RED_GBS],
                                  device=dev, dtype=torch.float64)
        if world > 1:
            dist.all_reduce(cnt)
            dist.all_reduce(worst_link, op=dist.ReduceOp.MAX)
        pulls, pushes, rpulls, rpushes, records, nans, bads = (float(x) for x in cnt.tolist())
        steps_t = a.steps * t_s
        # per GPU and direction: the rows it pulls come in, the deltas it pushes go out, and by symmetry it
        # serves the same amounts to its peers; roofline time = rows pulled / load rate + rows pushed / red.add rate
        link_bytes_dir = (rpulls + rpushes) * row_bytes / world / a.steps
        link_roof_s = float(worst_link.item()) * row_bytes / 1e9 / a.steps
        # local HBM traffic modelled from the counters: user row read + write per record, item row read per
        # local pull, read-modify-write per push that lands in this GPU's shard (all of them, job-wide)
        hbm_bytes = (2 * records + (pulls - rpulls) + 2 * pushes) * row_bytes / world / a.steps
        roof = {"rows_pulled_per_step_per_gpu": pulls / world / a.steps,
                "rows_pushed_per_step_per_gpu": pushes / world / a.steps,
                "combine_ratio_records_per_pull": records / max(pulls, 1.0),
                "remote_pull_fraction_of_rows": rpulls / max(pulls, 1.0),
                "nvlink_gbs_per_dir_actual": link_bytes_dir / t_s / 1e9 if world > 1 else 0.0,
                "nvlink_frac_of_measured_p2p": (link_roof_s / t_s) if world > 1 else None,
                "nvlink_frac_of_symmetric_probe": (link_bytes_dir / t_s / 1e9 / P2P_SYMMETRIC_GBS) if world > 1 else None,
                "nvlink_peaks_gbs": {"one_way_load": P2P_LOAD_GBS, "one_way_red_add": P2P_RED_GBS,
                                     "all_gpus_pull_and_push": P2P_SYMMETRIC_GBS,
                                     "source": "profiles/r1_2gpu/p2p_probe_2gpu.json, profiles/r2_2gpu/p2p_symmetric_2gpu.json"},
                "hbm_gbs_modelled_from_counters": hbm_bytes / t_s / 1e9,
                "hbm_frac_of_measured_copy_peak": hbm_bytes / t_s / 1e9 / hbm_peak,
                "logical_rows_per_sec_if_uncombined": value}
        extra = {"nan_count": int(nans), "bad_ids": int(bads), "mse": st["loss_sum"] / max(st["records"], 1),
                 "pull_rows_per_sec": pulls / steps_t, "push_rows_per_sec": pushes / steps_t,
                 "max_in_flight": st["max_in_flight"], "limiter_stalls": st["limiter_stalls"]}
        probe = None if a.no_probe else pushes_land_probe(torch, dist, model, a, rank, world, dev)
        if probe is not None:
            extra["probe"] = probe
        extra["training_valid"] = bool(nans == 0 and bads == 0 and (probe is None or probe["ok"]))
    else:
        roof = {"logical_rows_per_sec_if_uncombined": value}
        extra = {"mse": float(model.loss_sum.item()) / max(model.records, 1),
                 "pull_rows_per_sec": value, "push_rows_per_sec": value}

    if rank == 0:
        print(json.dumps({
            "metric": METRIC, "value": value, "unit": "ratings/s", "n_gpus": world, "steps": a.steps,
            "warmup": a.warmup, "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": (value / baseline["value"]) if baseline else None, "baseline": baseline,
            "dtype": "fp32", "data": "synthetic", "impl": a.impl,
            "config": {"model": f"mf_sgd_rank{a.rank}_{a.users}x{a.items}",
                       "global_batch": a.batch * world, "seq_len": None,
                       "parallelism": f"w{world}xps{world}", "item_dist": a.item_dist,
                       "variant": model.variant if a.impl == "ours" else None, "pull_limit": a.pull_limit,
                       "l2": f"working set per step {2 * row_bytes * a.batch >> 20} MiB of random rows >> 126 MB L2; "
                             f"{a.pool} distinct batches rotated"},
            "clocks": clocks, "e2e": e2e, "gpu_launches": gpu_launches, "roofline": roof, **extra}))
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    return 0


def pushes_land_probe(torch, dist, model, a, rank, world, dev, n=4096):
    """One extra step on a small conflict-free batch whose items all live on the NEXT rank's shard,
    replayed in fp64 from rows gathered before it: proves that pulls read the owner's row and that
    the remote red.add pushes land there (``training_valid`` is more than "no NaN")."""
    k = torch.arange(n, device=dev, dtype=torch.int64)
    users = k * world + rank
    items = k * world + (rank + 1) % world
    if int(users.max()) >= a.users or int(items.max()) >= a.items:
        return None
    r = torch.full((n,), 0.5, device=dev)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    U0 = model.users.gather(users).double()
    V0 = model.items.gather(items).double()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    model.step(users.int(), items.int(), r)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    U1 = model.users.gather(users).double()
    V1 = model.items.gather(items).double()
    e = (0.5 - (U0 * V0).sum(1))[:, None]
    err = max(float((U0 + a.lr * e * V0 - U1).abs().max()), float((V0 + a.lr * e * U0 - V1).abs().max()))
    moved = float((V1 - V0).abs().max())
    t = torch.tensor([err, -moved], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    err, moved = float(t[0]), -float(t[1])
    return {"max_abs_err_vs_fp64_replay": err, "min_over_ranks_max_item_row_change": moved,
            "items_owned_by": "next rank" if world > 1 else "local shard", "n": n,
            "ok": bool(err < 1e-6 and moved > 0.0)}


def cpu_arm(a, torch):
    """BL2: the multi-threaded C++ restatement of the PS loop on the box's CPUs (stands in for
    the unbuildable Flink run; labelled "CPU restatement, not Flink").  The tables are as large as
    the host's free memory allows (fp64 rows), the per-record work is identical."""
    from fps_b200.parallel import cpu_baseline
    try:
        import psutil
        avail = psutil.virtual_memory().available
    except Exception:
        avail = 64 << 30
    budget = min(int(avail * 0.5), 32 << 30)      # the single-threaded table fill takes ~1 s per GB
    row = a.rank * 8
    users = min(a.users, max(1_000_000, budget * 10 // 11 // row))
    items = min(a.items, max(100_000, budget // 11 // row))
    batch = min(a.batch, 1 << 20)
    threads = max(2, (os.cpu_count() or 8) // 2 * 2)
    res = cpu_baseline.bench_mf(users, items, a.rank, batch, a.steps, a.warmup, workers=threads // 2,
                                shards=threads // 2, lr=a.lr)
    print(json.dumps({"metric": METRIC, "value": res["ratings_per_sec"], "unit": "ratings/s", "n_gpus": 0,
                      "steps": a.steps, "warmup": a.warmup, "ms_per_step": res["ms_per_step"],
                      "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "fp64",
                      "data": "synthetic", "impl": "cpu_restatement_not_flink",
                      "config": {"model": f"mf_sgd_rank{a.rank}_{users}x{items}", "global_batch": batch,
                                 "parallelism": f"w{threads // 2}xps{threads // 2} threads",
                                 "host_cpus": os.cpu_count(), "host_mem_available_gb": round(avail / 2 ** 30, 1),
                                 "note": "tables sized to half of the host's free memory, at most 32 GB; per-record work identical"},
                      "gpu_launches": 0}))
    return 0


if __name__ == "__main__":
    sys.exit(main())
