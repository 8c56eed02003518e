"""Batched synthesis sharded over the GPUs of one box (BASELINE config C3).

Independent utterances, replicated weights, one process per GPU, NO data-path collective.  The plan is
a pure function of the utterance lengths, so every rank computes the same plan:

1. sort utterances by token length (stable), cut into batches of ``batch_size`` -- the batch composition
   (hence the padded T and the numerics of every utterance) does not depend on the number of ranks;
2. assign batches to ranks by greedy longest-processing-time on the cost estimate T * S_est, where
   S_est = frames_per_token * T / r (the reference stops a whole batch at the first utterance whose stop
   token fires, Helpers.py:51-52, so the batch cost is data dependent; measured S can be fed back).
"""
from __future__ import annotations

import numpy as np


def plan_batches(lengths, batch_size=64):
    """-> list of index arrays (into ``lengths``), each <= batch_size utterances of similar length."""
    order = np.argsort(np.asarray(lengths), kind="stable")
    return [order[i:i + batch_size] for i in range(0, len(order), batch_size)]


def batch_cost(lengths, idx, frames_per_token=13.0, r=2):
    T = int(np.max(np.asarray(lengths)[idx]))
    return float(T) * (frames_per_token * T / r) * len(idx)


def assign_batches(costs, world):
    """Greedy LPT: returns ``owner[b]`` in [0, world) minimising the heaviest rank."""
    owner = np.zeros(len(costs), np.int64)
    load = np.zeros(world)
    for b in sorted(range(len(costs)), key=lambda i: (-costs[i], i)):
        r = int(np.argmin(load))
        owner[b] = r
        load[r] += costs[b]
    return owner


def shard(lengths, world, batch_size=64, frames_per_token=13.0, r=2, measured_steps=None):
    """-> (batches, owner).  ``measured_steps[b]`` (optional) replaces the S estimate of batch b."""
    batches = plan_batches(lengths, batch_size)
    costs = []
    for b, idx in enumerate(batches):
        c = batch_cost(lengths, idx, frames_per_token, r)
        if measured_steps is not None and measured_steps[b] is not None:
            c = float(np.max(np.asarray(lengths)[idx])) * measured_steps[b] * len(idx)
        costs.append(c)
    return batches, assign_batches(costs, world)


class _Ticket(object):
    """Dynamic batch assignment: a shared counter in the job's rendezvous store (control plane only -- no tensor ever crosses
    ranks).  Every call hands out the next batch index of the common, cost-sorted order; the reference's batch-global stop makes
    the cost of a batch data dependent (Helpers.py:51-52), which a static plan cannot see."""

    def __init__(self, store, key):
        self.store, self.key = store, key

    def next(self):
        return int(self.store.add(self.key, 1)) - 1


def synthesize_batches(engine, hp, seqs, rank=0, world=1, batch_size=64, gst_weights=None, max_iters=None, philox_seed=None,
                       fetch=("mel_outputs", "stop"), store=None, store_key="taco_b200/next_batch"):
    """Runs this rank's share of ``seqs`` (list of int token arrays).  Returns {utterance index: dict(mel [len,80],
    steps)} for the utterances this rank owns; outputs are trimmed at the first rounded stop token and clipped like
    Synthesizer.synthesize.  ``engine`` is a TacoEngine bound to this rank's GPU, or an EnginePool on it (several batches
    in flight, same results).  ``store`` (a torch.distributed Store shared by the ranks, e.g. the default process-group store):
    batches are then claimed dynamically from one cost-sorted list instead of the static LPT plan -- same batch composition,
    same outputs, better balance when the stop-driven lengths differ."""
    from .Utils.Tacotron_feeder import _get_output_lengths, _prepare_inputs
    from .synthetic import gst_random_weights
    lengths = [len(s) for s in seqs]
    batches, owner = shard(lengths, world, batch_size, r=hp.outputs_per_step)
    gw = gst_random_weights(hp) if gst_weights is None else gst_weights
    seed = hp.tacotron_random_seed if philox_seed is None else philox_seed
    lo, hi = (-hp.max_abs_value, hp.max_abs_value) if hp.symmetric_mels else (0, hp.max_abs_value)
    mine = [(b, idx) for b, idx in enumerate(batches) if owner[b] == rank]
    # longest batches first: the library's workspaces are grow-only, so they are sized once by the first (largest) batch
    mine.sort(key=lambda job: -max(lengths[i] for i in job[1]))

    def run(eng, job):
        b, idx = job
        ids, _ = _prepare_inputs([np.asarray(seqs[i], np.int32) for i in idx])
        lens = np.asarray([lengths[i] for i in idx], np.int32)
        res = eng.forward_host(ids.astype(np.int32), lens, gst_weights=gw, max_iters=max_iters, philox_seed=seed, fetch=fetch)
        keep = _get_output_lengths(res["stop"])
        # ONE clip of the whole batch into a fresh array (the pinned buffers are reused by the engine's next batch); the per-utterance
        # results are views of it trimmed at the first rounded stop token -- the host cost per batch stays a single vector pass
        mel = np.clip(res["mel_outputs"], lo, hi)
        return {int(i): dict(mel=mel[j, :keep[j]], steps=res["steps"], batch=b) for j, i in enumerate(idx)}

    if store is not None and world > 1:
        order = sorted(range(len(batches)), key=lambda b: (-batch_cost(lengths, batches[b], r=hp.outputs_per_step), b))
        ticket = _Ticket(store, store_key)

        def claimed():
            while True:
                k = ticket.next()
                if k >= len(order):
                    return
                yield (order[k], batches[order[k]])

        import threading
        lock = threading.Lock()
        it = claimed()

        def run_claimed(eng, _):
            out_ = {}
            while True:
                with lock:
                    job = next(it, None)
                if job is None:
                    return out_
                out_.update(run(eng, job))

        n_workers = len(engine) if hasattr(engine, "map") else 1
        parts = engine.map(run_claimed, range(n_workers)) if hasattr(engine, "map") else [run_claimed(engine, 0)]
        out = {}
        for part in parts:
            out.update(part)
        return out
    # an EnginePool keeps several batches in flight on this GPU (engine.EnginePool); a single TacoEngine runs them in turn
    parts = engine.map(run, mine) if hasattr(engine, "map") else [run(engine, job) for job in mine]
    out = {}
    for part in parts:
        out.update(part)
    return out
