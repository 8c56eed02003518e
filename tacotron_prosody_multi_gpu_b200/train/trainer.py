"""Data-parallel training: one process per GPU, replicated weights, ONE exchange step per update.

The reference places one graph tower per GPU inside a single process, gathers every tower's
gradients on the CPU and averages them variable by variable (Tacotron.py:484-501).  Here each rank
owns one tower (batch shard) and the exchange is a single ``all_reduce(SUM)`` of the flat fp32 gradient
buffer (20 425 491 elements, 81.7 MB) over NCCL / NVLink; the division by the world size, the L2 term,
the global-norm clip and the Adam update are one fused pass over the reduced buffer on every rank
(identical inputs -> identical weights, no broadcast needed).  Batch-norm moving statistics are
per-rank batch estimates; they travel in the SAME flat bucket ([statistics | gradients], train/params.py) and are
averaged across ranks (the reference applies every tower's UPDATE_OPS to the same CPU variables,
Tacotron.py:513 -- a momentum update per tower; the mean of the per-rank updates is the order-independent
equivalent).  The bucket is exchanged in two contiguous halves: the decoder + postnet gradients (70 % of the bytes)
are final when the decoder BPTT ends, so their all-reduce starts there on a side stream and runs under the style /
encoder backward; the rest follows when backward() returns.
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
        self.overlap = True
        self._early = None            # (work handle | None, side-stream event) of the early half started in backward()
        self._side = None
        if self.world > 1:
            self.step_fn.on_tail_ready = self._start_early_half

    # ---- the exchange step ----
    def _on_cuda(self):
        return self.step_fn.store.bucket.is_cuda

    def _start_early_half(self):
        """Called inside backward() once store.g[tail_start:] is final: all-reduce that half on a side stream."""
        if not self.overlap or self.world == 1:
            return
        st = self.step_fn.store
        early = st.bucket[st.ns + st.tail_start:]
        if not self._on_cuda():                       # gloo / CPU twin (tests): no streams, reduce in place right away
            self.dist.all_reduce(early, op=self.dist.ReduceOp.SUM)
            self._early = (None, None)
            return
        if self._side is None:
            self._side = torch.cuda.Stream(device=st.bucket.device)
        ready = torch.cuda.Event()
        ready.record()
        with torch.cuda.stream(self._side):
            self._side.wait_event(ready)
            self.dist.all_reduce(early, op=self.dist.ReduceOp.SUM)
            done = torch.cuda.Event()
            done.record()
        self._early = (None, done)

    def _all_reduce(self, t):
        """Reducer handed to TrainStep.apply: SUM of the flat bucket over the ranks.  If the early half is already in flight
        only the late half [statistics | encoder / style gradients] is reduced here, then both are joined."""
        st = self.step_fn.store
        if t is st.bucket and self._early is not None:
            _, done = self._early
            self._early = None
            self.dist.all_reduce(st.bucket[:st.ns + st.tail_start], op=self.dist.ReduceOp.SUM)
            if done is not None:
                torch.cuda.current_stream().wait_event(done)
            return
        self.dist.all_reduce(t, op=self.dist.ReduceOp.SUM)

    def step(self, ids, lens, mel_targets, stop_targets, seed=None):
        """One update on this rank's shard.  Inputs: device tensors (int32 ids/lens, fp32 targets).
        Returns the 5 losses [before, after, stop, regularization, total] of this rank's tower (device tensor)."""
        st = self.step_fn
        seed = (self.hp.tacotron_random_seed + 7919 * self.global_step + 104729 * self.rank) if seed is None else seed
        out = st.forward(ids, lens, mel_targets, stop_targets, seed)
        st.backward()
        st.apply(self.global_step, world=self.world, all_reduce=self._all_reduce if self.world > 1 else None)
        self.global_step += 1
        ls = out["losses"]
        self.ops.ew("add", ls[4:5], ls[0:1], ls[1:2])
        self.ops.ew("add", ls[4:5], ls[4:5], ls[2:3])
        self.ops.ew("add", ls[4:5], ls[4:5], ls[3:4])
        return ls

    def export(self):
        return self.step_fn.store.export()

    # ---- checkpoint / resume (the reference's tf.train.Saver keeps variables, Adam slots and global_step, train.py:60) ----
    def save(self, path):
        self.step_fn.store.save_state(path, self.global_step)

    def load(self, path):
        self.global_step = self.step_fn.store.load_state(path)
        return self.global_step
