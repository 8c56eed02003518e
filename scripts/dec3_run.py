#!/usr/bin/env python
"""Run the decoder kernel alone (for ncu captures): python scripts/dec3_run.py [B] [S]"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from tacotron_prosody_multi_gpu_b200 import synthetic, weights  # noqa: E402
from tacotron_prosody_multi_gpu_b200.engine import TacoEngine  # noqa: E402
from tacotron_prosody_multi_gpu_b200.hparams import default_hparams  # noqa: E402

B = int(sys.argv[1]) if len(sys.argv) > 1 else 64
S = int(sys.argv[2]) if len(sys.argv) > 2 else 200
hp = default_hparams()
w = weights.set_stop_bias(weights.init_weights(hp, perturb=0.05), -10.0)
eng = TacoEngine(hp, w, precision="bf16")
ids, lens = synthetic.token_batch(hp, synthetic.cycle_lengths(B))
gw = synthetic.gst_random_weights(hp)
mem = eng.encode(torch.from_numpy(ids), torch.from_numpy(lens), gst_weights=gw)
for _ in range(2):
    d = eng.decode(mem, torch.from_numpy(lens), max_iters=S, philox_seed=1)
print("B=%d T=%d S=%d steps=%d kernel %.3f ms -> %.2f us/step" % (B, ids.shape[1], S, d["steps"], eng.stage_ms(1), 1e3 * eng.stage_ms(1) / S))
