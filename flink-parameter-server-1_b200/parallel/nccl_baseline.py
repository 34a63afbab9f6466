"""BL1 - the NCCL all-to-all parameter server (baseline, not the product).

The same MF step as the fused kernel, expressed the way a collective-based PS does it
(BASELINE.md §3): ids are bucketed by owner and exchanged with
``all_to_all_single`` (X1), owners gather the rows and send them back (X2), workers run the SGD
maths with torch ops, deltas go back with a third all-to-all and are applied with
``index_add_``.  Four collectives (counts, ids, rows, deltas) plus ~15 torch kernels per step - that cost is what
the fused P2P kernel removes.  Works on NCCL (GPU) and gloo (CPU tests).
"""
from __future__ import annotations

from typing import Optional

import torch
import torch.distributed as dist


class AllToAllPS:
    def __init__(self, n_users: int, n_items: int, dim: int, lr: float, *, init_range=(-0.01, 0.01),
                 seed: int = 0, device: Optional[torch.device] = None, lambda_: float = 0.0):
        self.rank = dist.get_rank() if dist.is_initialized() else 0
        self.world = dist.get_world_size() if dist.is_initialized() else 1
        self.device = device or torch.device("cpu")
        self.dim, self.lr, self.lambda_ = dim, lr, lambda_
        W = self.world
        g = torch.Generator(device="cpu").manual_seed(seed * 1000 + self.rank)
        lo, hi = init_range
        n_i = (n_items + W - 1) // W
        n_u = (n_users + W - 1) // W
        self.items = torch.empty(n_i, dim, device=self.device).uniform_(lo, hi)
        self.users = torch.empty(n_u, dim, device=self.device).uniform_(lo, hi)
        self.loss_sum = torch.zeros((), device=self.device, dtype=torch.float64)
        self.records = 0
        self.kernel_launches_per_step = 0

    def _a2a(self, send: torch.Tensor, send_counts, recv_counts) -> torch.Tensor:
        out = send.new_empty((int(sum(recv_counts)),) + tuple(send.shape[1:]))
        if self.world == 1:
            out.copy_(send)
            return out
        dist.all_to_all_single(out, send, output_split_sizes=list(recv_counts),
                               input_split_sizes=list(send_counts))
        return out

    def step(self, users: torch.Tensor, items: torch.Tensor, ratings: torch.Tensor) -> None:
        W = self.world
        items = items.long()
        owner = items % W
        order = torch.argsort(owner, stable=True)
        it_sorted = items[order]
        send_counts = torch.bincount(owner, minlength=W)
        if W > 1:
            recv_counts = torch.empty_like(send_counts)
            dist.all_to_all_single(recv_counts, send_counts)
        else:
            recv_counts = send_counts
        sc, rc = send_counts.tolist(), recv_counts.tolist()
        # pull: ids -> owners -> rows back
        req = self._a2a(it_sorted // W, sc, rc)
        rows = self.items[req]
        v = self._a2a(rows, rc, sc)
        # compute on the worker
        u_slot = users.long()[order] // W
        u = self.users[u_slot]
        r = ratings[order]
        e = r - (u * v).sum(dim=1)
        g = (self.lr * e)[:, None]
        du = g * v - self.lr * self.lambda_ * u
        dv = g * u - self.lr * self.lambda_ * v
        self.users.index_add_(0, u_slot, du)
        # push: deltas -> owners, applied with index_add_
        dv_recv = self._a2a(dv, sc, rc)
        self.items.index_add_(0, req, dv_recv)
        self.loss_sum += (e.double() ** 2).sum()
        self.records += int(users.numel())
