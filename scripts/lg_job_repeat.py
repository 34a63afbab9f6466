"""Config-7 job at one batch size, repeated: separates one-time costs (kernel loading, first allocations) from the
steady rate.  python scripts/lg_job_repeat.py [batch] [repeats]"""
import os, random, sys, time, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from fps_b200.models.mf import Rating
from fps_b200.models.mf.device_topk import ps_online_learner_and_generator_device
from fps_b200.parallel.device_context import init_context
ctx = init_context()
bs = int(sys.argv[1]) if len(sys.argv) > 1 else 8192
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 4
rnd = random.Random(7)
n_users, n_items, n_events = 12_547, 19_048, 69_064
pop = [int(n_items * rnd.random() ** 3) for _ in range(n_events)]
stream = [Rating(rnd.randrange(n_users) if rnd.random() < 0.3 else (pop[t] * 7) % n_users, pop[t], 1.0, t) for t in range(n_events)]
kw = dict(num_factors=10, range_min=-0.01, range_max=0.01, learning_rate=0.05, negative_sample_rate=9, K=100, worker_k=75, pull_limit=500)
for r in range(reps):
    torch.cuda.synchronize(); t0 = time.time()
    out, lg = ps_online_learner_and_generator_device(stream, n_users=n_users, n_items=n_items, batch_size=bs, user_memory=128, **kw)
    torch.cuda.synchronize(); dt = time.time() - t0
    print(f"batch {bs} run {r}: {dt * 1e3:.1f} ms = {n_events / dt / 1e6:.3f} M events/s, ndcg {out.ndcg(n_events - 20000):.5f}", flush=True)
