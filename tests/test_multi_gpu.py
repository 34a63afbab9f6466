"""Multi-GPU end-to-end check (needs >= 2 GPUs on the box; skipped otherwise): launches
``scripts/multi_gpu_check.py`` under torchrun - peer-mapped pulls, remote red.add pushes, the fused
MF kernel with items sharded over all ranks, identical reads on every rank."""
import os
import subprocess
import sys

import pytest

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.timeout(600)
def test_two_rank_peer_mapped_parameter_server():
    n = torch.cuda.device_count()
    if n < 2:
        pytest.skip("needs at least 2 GPUs")
    world = 2
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
           "--master-addr", "127.0.0.1", "--master-port", "29577", os.path.join(ROOT, "scripts", "multi_gpu_check.py")]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=500, cwd=ROOT)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert "MULTI_GPU_CHECK PASS" in r.stdout


@pytest.mark.timeout(600)
def test_two_ranks_on_one_gpu_cross_process_data_plane():
    """Runs on a one-GPU box too: two processes share GPU 0 (gloo control plane, CUDA-IPC peer mapping) and pull
    from / push to each other's shards, then train the fused MF kernel over the split item table."""
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2",
           "--master-addr", "127.0.0.1", "--master-port", "29579", os.path.join(ROOT, "scripts", "two_ranks_one_gpu_check.py")]
    env = dict(os.environ, CUDA_VISIBLE_DEVICES=os.environ.get("CUDA_VISIBLE_DEVICES", "0").split(",")[0])
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=500, cwd=ROOT, env=env)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert "TWO_RANKS_ONE_GPU PASS" in r.stdout
