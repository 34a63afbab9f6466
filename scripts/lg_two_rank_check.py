"""Learner+generator on several ranks with batches large enough for the packed in-place scorer: runs and learns
(hit rate over the last third above the overall one).  Whether the ranks return identical lists is printed, not
required: every rank gathers the user rows at its own time while the others keep pushing (asynchronous semantics).
Measured on 2 GPUs: packed scorer 0.070 -> 0.128, in-kernel scorer (FPS_TOPK_INPLACE_PACKED=0) 0.072 -> 0.132.  torchrun --nproc-per-node 2 scripts/lg_two_rank_check.py"""
import os, random, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch.distributed as dist
from fps_b200.models.mf import Rating
from fps_b200.models.mf.device_topk import ps_online_learner_and_generator_device
from fps_b200.parallel.device_context import init_context
ctx = init_context()
rnd = random.Random(3)
n_users, n_items = 400, 600
stream = [Rating(u := rnd.randrange(n_users), (u % 6) * 100 + rnd.randrange(100), 1.0, t) for t in range(24000)]
out, lg = ps_online_learner_and_generator_device(stream, n_users=n_users, n_items=n_items, batch_size=2048, num_factors=16,
                                                 range_min=-0.1, range_max=0.1, learning_rate=0.1, negative_sample_rate=3,
                                                 user_memory=8, K=20, worker_k=20, pull_limit=5000)
h0, h1 = out.hit_rate(0), out.hit_rate(16000)
ids = torch.from_numpy(out.ids[-2048:].copy()).cuda()
ref = ids.clone()
dist.broadcast(ref, 0)
same = bool(torch.equal(ids, ref))
print(f"[rank {ctx.rank}] scorer {'packed' if lg._scorer is not None else 'in-kernel'}; hit rate {h0:.3f} overall, {h1:.3f} after 16K events; same ids on all ranks: {same}", flush=True)
ok = torch.tensor([1 if h1 > h0 else 0], device="cuda")
dist.all_reduce(ok, op=dist.ReduceOp.MIN)
if ctx.rank == 0:
    print("LG_TWO_RANK", "PASS" if ok.item() == 1 else "FAIL", flush=True)
dist.destroy_process_group()
