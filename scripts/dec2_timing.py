#!/usr/bin/env python
"""Per-stage time of the dataflow decoder kernel (TACO_DEC2_TIMING): stage nanoseconds per step, per CTA class."""
import ctypes as C
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
os.environ["TACO_DEC2_TIMING"] = "1"
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
buf = np.zeros((128, 16), np.int64)
eng._check(eng.L.taco_debug_read_workspace(eng._h, b"dec2_timing", buf.ctypes.data_as(C.c_void_p), buf.nbytes), "debug_read")
names = ["A.mma", "A.epi", "B.mma", "B.epi", "C.mma", "C.epi", "D.mma", "D.epi", "E.att", "E6.loc", "F.mma", "F.epi", "G.ctx", "G.h1z", "G.h2z", "-"]
us = buf / 1e3 / S
for cls, sel in (("cta 0", slice(0, 1)), ("ctas 0-39", slice(0, 40)), ("ctas 40-79 (x producers)", slice(40, 80)), ("cta 80 (stop)", slice(80, 81)),
                 ("ctas 81-127", slice(81, 128))):
    m = us[sel].mean(0)
    print("%-26s " % cls + "  ".join("%s %.2f" % (n.split()[0], v) for n, v in zip(names, m)) + "  | sum %.2f us" % m.sum())
