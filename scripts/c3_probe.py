#!/usr/bin/env python
"""C3 per-batch cost on one rank of an 8-rank sharding (run on one GPU): where the time of a small shard goes."""
import os, sys, time
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from tacotron_prosody_multi_gpu_b200 import sharding, synthetic, weights
from tacotron_prosody_multi_gpu_b200.engine import EnginePool
from tacotron_prosody_multi_gpu_b200.hparams import default_hparams

hp = default_hparams()
w = weights.scale_stop_projection(weights.init_weights(hp, perturb=0.05), 8.0, 0.104)
pool = EnginePool(hp, w, precision="bf16", in_flight=2)
ulens = synthetic.utterance_set(hp, 4096)
rng = np.random.Generator(np.random.PCG64(hp.tacotron_data_random_state + 7))
seqs = [np.concatenate([rng.integers(2, 66, size=int(n) - 1), [1]]).astype(np.int32) for n in ulens]
gw = synthetic.gst_random_weights(hp)
longest = sorted(range(4096), key=lambda i: -len(seqs[i]))[:128]
sharding.synthesize_batches(pool, hp, [seqs[i] for i in longest], 0, 1, 64, gw, 720, 45454)
torch.cuda.synchronize()
for world in (8, 1):
    for rank in range(min(world, 3)):
        t0 = time.perf_counter()
        res = sharding.synthesize_batches(pool, hp, seqs, rank, world, 64, gw, 720, 45454)
        torch.cuda.synchronize()
        dt = 1e3 * (time.perf_counter() - t0)
        steps = {}
        for v in res.values():
            steps[v["batch"]] = v["steps"]
        print("world %d rank %d: %d batches, %d decoder steps (%.1f ms of decoder at 16.2 us), %.1f ms wall" % (
            world, rank, len(steps), sum(steps.values()), sum(steps.values()) * 16.2e-3, dt))
