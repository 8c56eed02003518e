"""Data-parallel training: one process per GPU, replicated weights, ONE exchange step per update.

The reference places one graph tower per GPU inside a single process, gathers every tower's
gradients on the CPU and averages them variable by variable (Tacotron.py:484-501).  Here each rank
owns one tower (batch shard) and the exchange is a single ``all_reduce(SUM)`` of the flat fp32 gradient
buffer (20 425 491 elements, 81.7 MB) over NCCL / NVLink; the division by the world size, the L2 term,
the global-norm clip and the Adam update are one fused pass over the reduced buffer on every rank
(identical inputs -> identical weights, no broadcast needed).  Batch-norm moving statistics are
per-rank batch estimates; they are averaged across ranks in the same step (the reference applies every
tower's UPDATE_OPS to the same CPU variables, Tacotron.py:513 -- a momentum update per tower; the mean
of the per-rank updates is the order-independent equivalent).
"""
from __future__ import annotations

import torch

from .step import TrainStep


class DataParallelTrainer(object):
    def __init__(self, hp, weights, ops=None, device=0):
        if ops is None:
            from .ops import CudaOps
            ops = CudaOps(device)
        self.hp = hp
        self.ops = ops
        self.step_fn = TrainStep(hp, ops, weights)
        self.global_step = 0
        self.dist = torch.distributed if (torch.distributed.is_available() and torch.distributed.is_initialized()) else None
        self.world = self.dist.get_world_size() if self.dist else 1
        self.rank = self.dist.get_rank() if self.dist else 0

    def _all_reduce(self, t):
        self.dist.all_reduce(t, op=self.dist.ReduceOp.SUM)

    def step(self, ids, lens, mel_targets, stop_targets, seed=None):
        """One update on this rank's shard.  Inputs: device tensors (int32 ids/lens, fp32 targets).
        Returns the 5 losses [before, after, stop, regularization, total] of this rank's tower (device tensor)."""
        st = self.step_fn
        seed = (self.hp.tacotron_random_seed + 7919 * self.global_step + 104729 * self.rank) if seed is None else seed
        out = st.forward(ids, lens, mel_targets, stop_targets, seed)
        st.backward()
        st.apply(self.global_step, world=self.world, all_reduce=self._all_reduce if self.world > 1 else None)
        if self.world > 1 and st.store.s.numel():
            self._all_reduce(st.store.s)
            self.ops.ew("scale", st.store.s, st.store.s, None, 1.0 / self.world)
        self.global_step += 1
        ls = out["losses"]
        self.ops.ew("add", ls[4:5], ls[0:1], ls[1:2])
        self.ops.ew("add", ls[4:5], ls[4:5], ls[2:3])
        self.ops.ew("add", ls[4:5], ls[4:5], ls[3:4])
        return ls

    def export(self):
        return self.step_fn.store.export()
