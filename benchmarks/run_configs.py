#!/usr/bin/env python
"""The other BASELINE.json configs (bench.py covers config 5, the headline):

  1  offline SGD MF, rank 10, MovieLens-100K-shaped (943 users x 1682 items, 100K ratings), 10 epochs
     - device engine vs the C++ CPU restatement with workerParallelism = psParallelism = 4
  2  online (streaming) SGD MF, rank 64, MovieLens-20M-shaped stream (138K users x 27K items)
  3  binary passive-aggressive, RCV1-shaped (47 236 features, ~75 nnz), pull limiter on
  4  multiclass passive-aggressive (one-vs-all), 100 classes, 1M features, ~100 nnz
  6  LEMP top-K retrieval, rank 128, 1M items: tcgen05 kernel vs the CUDA-core twin vs torch matmul+topk

    python benchmarks/run_configs.py [--configs 1,2,3,4,6] [--out gpurun_out/configs.jsonl]
    torchrun --nproc-per-node N benchmarks/run_configs.py --configs 2,3,4       (multi-GPU)

Synthetic data of the named shapes, random-init parameters, CUDA-event timing (max over ranks),
3 warm-up + 10 timed launches unless stated.  One JSON line per config (rank 0).
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from fps_b200.parallel.device_context import init_context  # noqa: E402


def timed(fn, warmup=3, steps=10):
    import torch.distributed as dist
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    if dist.is_initialized():
        dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    ms = torch.tensor([e0.elapsed_time(e1) / steps], device="cuda", dtype=torch.float64)
    if dist.is_initialized():
        dist.barrier()
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    return float(ms.item())


def zipf_items(n, n_items, g, power=3.0):
    return (torch.rand(n, generator=g, dtype=torch.float64) ** power * n_items).long().clamp_(max=n_items - 1)


def config1(ctx):
    """Offline MF rank 10 on ML-100K-shaped data: whole job = 10 epochs over 100K ratings."""
    from fps_b200.models.mf.device import DeviceMF
    from fps_b200.parallel import cpu_baseline
    nu, ni, n, dim, epochs = 943, 1682, 100_000, 10, 10
    g = torch.Generator().manual_seed(1)
    u = torch.randint(0, nu, (n,), generator=g).int()
    i = zipf_items(n, ni, g, 2.0).int()
    truth_u, truth_i = torch.randn(nu, 4, generator=g), torch.randn(ni, 4, generator=g)
    r = ((truth_u[u.long()] * truth_i[i.long()]).sum(1) * 0.1 + 0.6).clamp(0.2, 1.0).float()   # stars / 5
    out = {"config": 1, "what": "offline MF rank 10, ML-100K-shaped, 10 epochs", "ratings_per_job": n * epochs}
    if ctx.world_size == 1:
        mf = DeviceMF(nu, ni, dim, -0.01, 0.01, 0.01, pull_limit=1600, seed=1, ctx=ctx, variant="auto")
        du, di, dr = u.cuda(), i.cuda(), r.cuda()

        def job():
            for _ in range(epochs):
                mf.step(du, di, dr)
        ms = timed(job, 2, 5)
        fresh = DeviceMF(nu, ni, dim, -0.01, 0.01, 0.01, pull_limit=1600, seed=1, ctx=ctx, variant="auto")
        for _ in range(epochs):
            fresh.step(du, di, dr)
        st = fresh.stats(check_nan=False)
        out.update(device_ms_per_job=ms, device_ratings_per_sec=n * epochs / ms * 1e3,
                   rmse_after_10_epochs=fresh.rmse(du, di, dr), nan_count=st["nan_count"], pull_limit=1600,
                   variant=fresh.variant, max_in_flight=st["max_in_flight"])
        # CPU restatement, W = PS = 4 (the reference's own parallelism for this config)
        W = S = 4
        per = n // W
        uw = np.zeros((W, per), np.int32)
        iw = np.zeros((W, per), np.int32)
        rw = np.zeros((W, per), np.float32)
        un, inn, rn = u.numpy(), i.numpy(), r.numpy()
        for w in range(W):
            idx = np.nonzero(un % W == w)[0][:per]
            uw[w, : len(idx)] = un[idx] // W
            iw[w, : len(idx)] = inn[idx]
            rw[w, : len(idx)] = rn[idx]
        ut = np.random.default_rng(0).uniform(-0.01, 0.01, (W, (nu + W - 1) // W, dim))
        it = np.random.default_rng(1).uniform(-0.01, 0.01, (S, (ni + S - 1) // S, dim))
        secs = 0.0
        for _ in range(epochs):
            s, _sse = cpu_baseline.run_mf(uw, iw, rw, ut, it, 0.01, 1600)
            secs += s
        out.update(cpu_restatement_ms_per_job=secs * 1e3, cpu_restatement_ratings_per_sec=per * W * epochs / secs,
                   cpu_note="C++ restatement of the PS loop (not Flink), 4 worker + 4 server threads, fp64")
    return out


def config2(ctx):
    """Online MF rank 64, ML-20M-shaped stream, micro-batches of 1M ratings per GPU."""
    from fps_b200.models.mf.device import DeviceMF
    nu, ni, dim, B = 138_493, 26_744, 64, int(os.environ.get("FPS_CONFIG2_BATCH", 1 << 20))
    W, rank = ctx.world_size, ctx.rank
    # lr 0.002: with 1M-rating launches per GPU on 26.7K items a hot item takes thousands of concurrent steps
    # from stale reads; at lr 0.01 that diverges on 2+ GPUs (nan_count > 0 - and a diverged run looks fast
    # because non-finite updates are skipped), so the learning rate is the one that trains at every N
    lr = float(os.environ.get("FPS_CONFIG2_LR", "0.002"))
    limit = int(os.environ.get("FPS_CONFIG2_PULL_LIMIT", "65536"))
    mf = DeviceMF(nu, ni, dim, -0.01, 0.01, lr, pull_limit=limit, seed=2, ctx=ctx,
                  variant=os.environ.get("FPS_CONFIG2_VARIANT", "auto"))    # FPS_MF_REPLICATE_ITEMS=1: replicated item table
    g = torch.Generator().manual_seed(2 + rank)
    batches = []
    for _ in range(4):
        u = (torch.randint(0, (nu + W - 1) // W, (B,), generator=g) * W + rank).clamp_(max=nu - 1)
        u = u - ((u - rank) % W)
        batches.append((u.int().cuda(), zipf_items(B, ni, g).int().cuda(), (torch.rand(B, generator=g) * 0.8 + 0.2).cuda()))   # stars / 5: the 1..5 scale diverges at lr 0.01 with 1M-rating launches
    k = [0]
    torch.cuda.synchronize()
    ready = torch.cuda.Event()
    ready.record()                      # the resident batches are valid from here on: the tile path may group
                                        # batch k + 1 on its side stream while step k runs

    def step():
        mf.step(*batches[k[0] % 4], ready_event=ready)
        k[0] += 1
    ms = timed(step, 3, 20)
    st = mf.stats(check_nan=False)      # a diverged model would skip its updates and look fast: report it
    return {"config": 2, "what": "online MF rank 64, ML-20M-shaped stream", "n_gpus": W, "variant": mf.variant,
            "batch_per_gpu": B, "ms_per_step": ms, "ratings_per_sec": B * W / ms * 1e3,
            "learning_rate": lr, "pull_limit": limit, "max_in_flight": st["max_in_flight"],
            "rows_pulled_per_record": st["pulls"] / max(st["records"], 1),
            "nan_count": st["nan_count"], "mse": st["loss_sum"] / max(st["records"], 1),
            "items_table": "replicated" if mf.items.replicated else "sharded"}


def _csr_batch(n_ex, n_feat, nnz, g, device):
    idx = torch.randint(0, n_feat, (n_ex, nnz), generator=g)
    idx = torch.sort(idx, dim=1).values
    val = torch.randn(n_ex, nnz, generator=g)
    indptr = torch.arange(0, (n_ex + 1) * nnz, nnz, dtype=torch.int64)
    return indptr.to(device), idx.reshape(-1).int().to(device), val.reshape(-1).float().to(device)


def config3(ctx):
    """Binary PA on RCV1-shaped vectors (47 236 features, 75 nnz), pull limiter 10 000."""
    from fps_b200.models.pa.device import DevicePA
    F, nnz, B = 47_236, 75, 1 << 17
    pa = DevicePA(F, algorithm="binary", variant="pa_i", aggressiveness=1.0, range_partitioning=True,
                  pull_limit=0, lazy=False, ctx=ctx)
    g = torch.Generator().manual_seed(3 + ctx.rank)
    indptr, idx, val = _csr_batch(B, F, nnz, g, "cuda")
    lab = (torch.randint(0, 2, (B,), generator=g) * 2 - 1).int().cuda()
    ms_unlimited = timed(lambda: pa.step(indptr, idx, val, lab), 3, 20)
    pa.params.pull_limit = 10_000            # the config's title says "with pull limiter": this is the headline
    ms = timed(lambda: pa.step(indptr, idx, val, lab), 3, 20)
    s = pa.stats()
    cpu = None
    if ctx.rank == 0:
        from fps_b200.parallel import cpu_baseline
        cpu = cpu_baseline.bench_pa_binary(F, nnz, 1 << 15, workers=4, shards=4, variant="pa_i", pull_limit=10_000)
    return {"config": 3, "cpu_restatement": cpu, "what": "binary PA-I, RCV1-shaped (47K features, 75 nnz)", "n_gpus": ctx.world_size,
            "examples_per_gpu": B, "ms_per_step": ms, "examples_per_sec": B * ctx.world_size / ms * 1e3,
            "pull_rows_per_sec": B * nnz * ctx.world_size / ms * 1e3,
            "table": ("replicated, deferred delta exchange" if pa.table.deferred else
                      "replicated, multimem.red pushes" if pa.table.multicast else "replicated, red.add per replica")
            if pa.table.replicated else "sharded",
            "pull_limit": 10_000, "max_in_flight": s["max_in_flight"],
            "without_limiter": {"ms_per_step": ms_unlimited, "examples_per_sec": B * ctx.world_size / ms_unlimited * 1e3}}


def config4(ctx):
    """Multiclass one-vs-all PA, 100 classes, 1M features, 100 nnz."""
    from fps_b200.models.pa.device import DevicePA
    F, L, nnz, B = 1_000_000, 100, 100, 1 << 15
    pa = DevicePA(F, label_count=L, algorithm="ova", variant="pa_i", aggressiveness=0.1,
                  range_partitioning=True, pull_limit=0, lazy=False, ctx=ctx)
    g = torch.Generator().manual_seed(4 + ctx.rank)
    indptr, idx, val = _csr_batch(B, F, nnz, g, "cuda")
    lab = torch.randint(0, L, (B,), generator=g).int().cuda()
    ms_unlimited = timed(lambda: pa.step(indptr, idx, val, lab), 3, 20)
    pa.params.pull_limit = 10_000
    ms = timed(lambda: pa.step(indptr, idx, val, lab), 3, 20)
    rows = B * nnz * ctx.world_size
    cpu = None
    if ctx.rank == 0:
        from fps_b200.parallel import cpu_baseline
        cpu = cpu_baseline.bench_pa_ova(F, L, nnz, 1 << 12, workers=4, shards=4, variant="pa_i", pull_limit=10_000)
    return {"config": 4, "cpu_restatement": cpu, "pull_limit": 10_000, "max_in_flight": pa.stats()["max_in_flight"],
            "without_limiter_ms_per_step": ms_unlimited, "what": "multiclass OVA PA-I, 100 classes, 1M features, 100 nnz", "n_gpus": ctx.world_size,
            "examples_per_gpu": B, "ms_per_step": ms, "examples_per_sec": B * ctx.world_size / ms * 1e3,
            "pull_rows_per_sec": rows / ms * 1e3, "push_rows_per_sec": rows / ms * 1e3,
            "row_bytes": L * 4, "gather_scatter_gbs": rows * L * 4 * 3 / ms / 1e6}


def config6(ctx):
    """LEMP top-K, rank 128: 4096 queries x 1M items, k = 100."""
    from fps_b200.ops.topk import lemp_topk, sort_items_by_length, topk_simt
    if ctx.world_size > 1:
        return {"config": 6, "skipped": "single-GPU kernel benchmark"}
    dim, n_q, n_items, k = 128, 4096, 1_000_000, 100
    g = torch.Generator(device="cuda").manual_seed(6)
    q = torch.randn(n_q, dim, device="cuda", generator=g)
    lens = torch.rand(n_items, 1, device="cuda", generator=g) ** 4 * 5 + 0.05       # long-tailed lengths
    items = torch.nn.functional.normalize(torch.randn(n_items, dim, device="cuda", generator=g), dim=1) * lens
    it_s, len_s, ids_s = sort_items_by_length(items)
    stats = torch.zeros(4, device="cuda", dtype=torch.int64)
    out = {"config": 6, "what": f"LEMP top-{k}, rank {dim}, {n_q} queries x {n_items} items"}
    ms_p = timed(lambda: lemp_topk(q, it_s, k, item_len=len_s, item_ids=ids_s, use_pruning=True, stats=stats), 2, 5)
    st = stats.tolist()
    ms_np = timed(lambda: lemp_topk(q, it_s, k, item_ids=ids_s, use_pruning=False), 1, 3)
    ms_simt = timed(lambda: topk_simt(q, it_s, k, item_ids=ids_s), 1, 2)
    from fps_b200.ops.topk import PackedItems, lemp_topk_packed
    t0 = time.time()
    packed = PackedItems(it_s, len_s, ids_s)
    torch.cuda.synchronize()
    pack_ms = (time.time() - t0) * 1e3
    stats.zero_()
    ms_pk_p = timed(lambda: lemp_topk_packed(q, packed, k, use_pruning=True, stats=stats), 2, 5)
    st2 = stats.tolist()
    ms_pk_np = timed(lambda: lemp_topk_packed(q, packed, k, use_pruning=False), 1, 3)
    out.update(packed_pipeline_all_tiles_ms=ms_pk_np, packed_pipeline_pruned_ms=ms_pk_p,
               packed_tiles_scored_fraction=st2[0] / max(st2[1], 1), pack_once_ms=pack_ms,
               packed_all_tiles_tflops_fp32_equiv=2.0 * n_q * n_items * dim / ms_pk_np / 1e9)

    def torch_ref():
        best_v = best_i = None
        for s in range(0, n_items, 1 << 17):
            sc = q @ it_s[s:s + (1 << 17)].T
            v, i = sc.topk(k, dim=1)
            i = i + s
            if best_v is None:
                best_v, best_i = v, i
            else:
                cv, ci = torch.cat([best_v, v], 1), torch.cat([best_i, i], 1)
                best_v, sel = cv.topk(k, dim=1)
                best_i = torch.gather(ci, 1, sel)
        return best_v
    torch.backends.cuda.matmul.allow_tf32 = False
    ms_torch = timed(torch_ref, 1, 3)
    flops = 2.0 * n_q * n_items * dim
    out.update(tcgen05_pruned_ms=ms_p, tiles_scored_fraction=st[0] / max(st[1], 1), tcgen05_all_tiles_ms=ms_np,
               tcgen05_all_tiles_tflops_fp32_equiv=flops / ms_np / 1e9, cuda_core_twin_ms=ms_simt,
               torch_fp32_matmul_topk_ms=ms_torch, queries_per_sec_pruned=n_q / ms_p * 1e3)
    # CPU arm: the same retrieval with the host's BLAS (numpy sgemm over item chunks + argpartition), on a slice
    # of the queries - a generous stand-in for the reference's per-query TreeSet / heap loop on the JVM
    n_cpu = 64
    q_h, it_h = q[:n_cpu].cpu().numpy(), it_s.cpu().numpy()
    t0 = time.time()
    best_v = np.full((n_cpu, k), -np.inf, dtype=np.float32)
    for s0 in range(0, n_items, 1 << 17):
        sc = q_h @ it_h[s0:s0 + (1 << 17)].T
        part = np.partition(sc, sc.shape[1] - k, axis=1)[:, -k:]
        best_v = np.partition(np.concatenate([best_v, part], axis=1), k, axis=1)[:, -k:]
    dt = time.time() - t0
    out["cpu_blas_topk"] = {"queries": n_cpu, "seconds": dt, "ms_per_4096_queries": dt / n_cpu * n_q * 1e3,
                            "threads": os.cpu_count(), "note": "numpy sgemm + partition on the GPU box's CPUs"}
    return out


def config7(ctx):
    """Prequential recommend-then-learn (psOnlineLearnerAndGenerator) on a stream shaped like the authors'
    session log (69K events, 12.5K users, 19K items; Notebooks/Data_Manipulation.ipynb): device engine
    vs the host engine (the exact-semantics Python event loop) on a prefix."""
    import random
    import time
    from fps_b200.models.mf import Rating, ps_online_learner_and_generator
    from fps_b200.models.mf.device_topk import ps_online_learner_and_generator_device
    from fps_b200.utils.evaluation import NDCGSink
    if ctx.world_size > 1:
        return {"config": 7, "skipped": "single-GPU job benchmark"}
    rnd = random.Random(7)
    n_users, n_items, n_events = 12_547, 19_048, 69_064
    pop = [int(n_items * rnd.random() ** 3) for _ in range(n_events)]          # long-tailed popularity
    stream = [Rating(rnd.randrange(n_users) if rnd.random() < 0.3 else (pop[t] * 7) % n_users, pop[t], 1.0, t)
              for t in range(n_events)]
    kw = dict(num_factors=10, range_min=-0.01, range_max=0.01, learning_rate=0.05, negative_sample_rate=9,
              K=100, worker_k=75, pull_limit=500)
    res = {"config": 7, "what": "prequential learner+generator, 69K events / 12.5K users / 19K items, rank 10, K=100, 9 negatives"}
    ps_online_learner_and_generator_device(stream[:4096], n_users=n_users, n_items=n_items, batch_size=2048,
                                           user_memory=128, **kw)          # warm-up (module load, allocator)
    for bs, graph in ((64, False), (64, True), (256, False), (256, True), (2048, False), (8192, False), (8192, True)):
        times = []
        for rep in range(4):                # the first job of a setting pays kernel loading and first allocations (~0.1 s)
            torch.cuda.synchronize()
            t0 = time.time()
            out, lg = ps_online_learner_and_generator_device(stream, n_users=n_users, n_items=n_items, batch_size=bs,
                                                             user_memory=128, cuda_graph=graph, **kw)
            torch.cuda.synchronize()
            if rep:
                times.append(time.time() - t0)
        dt = sorted(times)[1]               # median of three warm jobs (the loop around the steps is Python)
        s = lg.stats()
        res[f"device_batch{bs}" + ("_cuda_graph" if graph else "")] = {
            "events_per_sec": n_events / dt, "seconds": dt, "seconds_all": times, "ndcg_last20k": out.ndcg(n_events - 20000),
            "hit_rate_last20k": out.hit_rate(n_events - 20000), "sgd_records": s["records"],
            "nan_count": s["nan_count"], "max_in_flight": s["max_in_flight"], "cuda_graph": graph}
    # the same job fed with arrays instead of 69K Rating objects (unpacking them is half of the job's time)
    cols = (np.array([r.user for r in stream], dtype=np.int32), np.array([r.item for r in stream], dtype=np.int32),
            np.array([r.rating for r in stream], dtype=np.float32), np.array([r.timestamp for r in stream]))
    times = []
    for rep in range(4):
        torch.cuda.synchronize()
        t0 = time.time()
        out, lg = ps_online_learner_and_generator_device(cols, n_users=n_users, n_items=n_items, batch_size=8192,
                                                         user_memory=128, **kw)
        torch.cuda.synchronize()
        if rep:
            times.append(time.time() - t0)
    dt = sorted(times)[1]
    res["device_batch8192_columnar_input"] = {"events_per_sec": n_events / dt, "seconds": dt, "seconds_all": times,
                                              "ndcg_last20k": out.ndcg(n_events - 20000)}
    n_host = 3000
    t0 = time.time()
    out_h = ps_online_learner_and_generator(stream[:n_host], user_memory=128, worker_parallelism=4, ps_parallelism=4,
                                            iteration_wait_time=0, **kw)
    dt = time.time() - t0
    sink = NDCGSink().consume(out_h)
    res["host_engine_prefix"] = {"events": n_host, "events_per_sec": n_host / dt, "seconds": dt, "outputs": len(out_h),
                                 "ndcg": sink.ndcg, "hit_rate": sink.hit_rate}
    dev_prefix, _ = ps_online_learner_and_generator_device(stream[:n_host], n_users=n_users, n_items=n_items,
                                                           batch_size=64, user_memory=128, **kw)
    res["device_same_prefix_batch64"] = {"ndcg": dev_prefix.ndcg(), "hit_rate": dev_prefix.hit_rate()}
    return res


def config8(ctx):
    """Streaming ring (persistent consumer kernel; the host feeds a device-resident ring through the copy engine): config-2 shapes
    (rank 64, ML-20M-sized tables), records pushed by the host in 1M-record feeds."""
    import time
    from fps_b200.models.mf.device import DeviceMF
    from fps_b200.models.mf.stream import OnlineStream
    if ctx.world_size > 1:
        return {"config": 8, "skipped": "single-GPU streaming benchmark"}
    nu, ni, dim, n, feed = 138_493, 26_744, 64, 16_000_000, 1_000_000
    mf = DeviceMF(nu, ni, dim, -0.01, 0.01, 0.01, pull_limit=0, seed=2, ctx=ctx, variant="direct")
    g = torch.Generator().manual_seed(8)
    u = torch.randint(0, nu, (n,), generator=g, dtype=torch.int32).numpy()
    i = (torch.rand(n, generator=g) ** 3 * ni).int().numpy()
    r = torch.rand(n, generator=g).numpy()
    st = OnlineStream(mf, capacity=1 << 22, iteration_wait_ms=500, max_lifetime_ms=120_000)
    st.feed(u[:feed], i[:feed], r[:feed])          # starts the kernel; warm-up
    t0 = time.time()
    for k in range(feed, n, feed):
        st.feed(u[k:k + feed], i[k:k + feed], r[k:k + feed])
    while st._retired() < n:
        time.sleep(0.0005)
    dt = time.time() - t0
    reason = st.join()
    s = mf.stats()
    return {"config": 8, "what": "online MF rank 64 through the persistent ring consumer (host-fed device ring, copy-engine DMA)",
            "records": n - feed, "seconds": dt, "records_per_sec": (n - feed) / dt, "exit_reason": reason,
            "ring_done": s["ring_done"], "relaunches": st.relaunches, "nan_count": s["nan_count"]}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--configs", default="1,2,3,4,6")
    ap.add_argument("--out", default=os.path.join(ROOT, "gpurun_out", "configs.jsonl"))
    a = ap.parse_args()
    ctx = init_context()
    fns = {1: config1, 2: config2, 3: config3, 4: config4, 6: config6, 7: config7, 8: config8}
    os.makedirs(os.path.dirname(a.out), exist_ok=True)
    for c in [int(x) for x in a.configs.split(",")]:
        try:
            res = fns[c](ctx)
        except Exception as e:  # noqa: BLE001
            res = {"config": c, "error": repr(e)[:400]}
        if ctx.rank == 0:
            line = json.dumps(res)
            print(line, flush=True)
            with open(a.out, "a") as f:
                f.write(line + "\n")
        torch.cuda.synchronize()
        ctx.barrier()
    if ctx.distributed:
        import torch.distributed as dist
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
