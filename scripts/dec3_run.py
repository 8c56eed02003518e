#!/usr/bin/env python
"""One free-running launch of the cluster decoder at the bench shape (for ncu captures): B, S from argv."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from tacotron_prosody_multi_gpu_b200 import synthetic, weights  # noqa: E402
from tacotron_prosody_multi_gpu_b200.engine import TacoEngine  # noqa: E402
from tacotron_prosody_multi_gpu_b200.hparams import default_hparams  # noqa: E402

B = int(sys.argv[1]) if len(sys.argv) > 1 else 64
S = int(sys.argv[2]) if len(sys.argv) > 2 else 100
hp = default_hparams()
w = weights.set_stop_bias(weights.init_weights(hp, perturb=0.05), -10.0)
eng = TacoEngine(hp, w, precision="bf16")
MAXLEN = int(sys.argv[3]) if len(sys.argv) > 3 else 1000
ids, lens = synthetic.token_batch(hp, [min(n, MAXLEN) for n in synthetic.cycle_lengths(B)])
mem = eng.encode(torch.from_numpy(ids), torch.from_numpy(lens), gst_weights=synthetic.gst_random_weights(hp))
steps = -1
for _ in range(3):
    try:
        d = eng.decode(mem, torch.from_numpy(lens), max_iters=S, philox_seed=1)
        steps = d["steps"]
    except Exception as e:   # ablation runs (TACO_DEC3_DBG) may abort by design; the stage timer is still valid
        steps = -1
print("B=%d T=%d S=%d steps=%d decoder stage %.3f ms -> %.2f us/step" % (B, ids.shape[1], S, steps, eng.stage_ms(1), 1e3 * eng.stage_ms(1) / S))
