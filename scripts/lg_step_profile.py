"""Where a learner+generator step spends its GPU time (config 7 shape): kernel table of torch.profiler
over 40 steps at a given batch size.  python scripts/lg_step_profile.py [batch]"""
import os, random, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from fps_b200.models.mf.device_topk import DeviceLearnerGenerator
from fps_b200.parallel.device_context import init_context
ctx = init_context()
bs = int(sys.argv[1]) if len(sys.argv) > 1 else 256
graph = (sys.argv[2] == "graph") if len(sys.argv) > 2 else False
rnd = random.Random(7)
n_users, n_items = 12_547, 19_048
lg = DeviceLearnerGenerator(n_users, n_items, num_factors=10, range_min=-0.01, range_max=0.01, learning_rate=0.05,
                            negative_sample_rate=9, K=100, worker_k=75, pull_limit=500, user_memory=128, cuda_graph=graph)
g = torch.Generator().manual_seed(1)
def batch():
    return (torch.randint(0, n_users, (bs,), generator=g).int().cuda(), (torch.rand(bs, generator=g) ** 3 * n_items).int().cuda(),
            torch.ones(bs).cuda())
for _ in range(10):
    lg.step_tensors(*batch())
torch.cuda.synchronize()
bb = [batch() for _ in range(40)]
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for b in bb:
    lg.step_tensors(*b)
e1.record(); torch.cuda.synchronize()
print("batch", bs, "graph", graph, "ms/step", e0.elapsed_time(e1) / 40)
from torch.profiler import profile, ProfilerActivity
with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
    for b in bb:
        lg.step_tensors(*b)
    torch.cuda.synchronize()
print(prof.key_averages().table(sort_by="cuda_time_total", row_limit=14, max_name_column_width=60))
