#!/usr/bin/env python
"""Per-stage time of the cluster decoder kernel (TACO_DEC2_TIMING): stage microseconds per step, per CTA class."""
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
steps = -1
for _ in range(2):
    try:
        steps = eng.decode(mem, torch.from_numpy(lens), max_iters=S, philox_seed=1)["steps"]
    except Exception:   # ablation runs (TACO_DEC3_DBG=513: nobody waits) abort by design; timers stay valid
        steps = -1
print("B=%d T=%d S=%d steps=%d kernel %.3f ms -> %.2f us/step" % (B, ids.shape[1], S, steps, eng.stage_ms(1), 1e3 * eng.stage_ms(1) / S))
buf = np.zeros((128, 24), np.int64)
eng._check(eng.L.taco_debug_read_workspace(eng._h, b"dec2_timing", buf.ctypes.data_as(C.c_void_p), buf.nbytes), "debug_read")
names = ["P1.mma", "P1.epi", "P2.mma", "P2.epi", "L1.mma", "L1.epi", "L2.mma", "L2.epi", "E1.qwait", "zshadow", "E3.softmax", "W1C", "stop", "-", "E6loc", "E5.mcast", "E2.energy", "E5.ctx", "sm.load", "sm.max", "sm.exp", "sm.sum", "sm.store", "-"]
us = buf / 1e3 / S
for cls, sel in (("cluster0 rank 0 (longest)", slice(0, 1)), ("cluster0 rank 5", slice(5, 6)), ("cluster0 rank 10 (stop)", slice(10, 11)), ("cluster0 rank 15", slice(15, 16)),
                 ("cluster0 mean", slice(0, 16)), ("cluster1 rank 0 (T=111)", slice(16, 17)), ("cluster1 rank 9", slice(25, 26)), ("cluster1 rank 12", slice(28, 29))):
    m = us[sel].mean(0)
    print("%-24s " % cls + "  ".join("%s %.2f" % (n, v) for n, v in zip(names, m) if n != "-") + "  | sum %.2f us" % m.sum())

# the straggler of every cluster: the CTA that waits least for the context rows
for c in range(8):
    blk = us[16 * c:16 * c + 16]
    if blk.sum() == 0:
        continue
    r = int(np.argmin(blk[:, 0]))
    m = blk[r]
    print("cluster%d straggler rank %2d  " % (c, r) + "  ".join("%s %.2f" % (n, v) for n, v in zip(names, m) if n != "-") + "  | sum %.2f us" % m.sum())
