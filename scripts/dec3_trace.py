#!/usr/bin/env python
"""One-step timeline of the cluster decoder (TACO_DEC3_TRACE=<step>): per CTA of a cluster, the time of every trace point of that
step relative to the cluster's earliest step start (globaltimer for the cross-CTA offset, clock64 deltas inside a CTA)."""
import ctypes as C
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
STEP = int(sys.argv[3]) if len(sys.argv) > 3 else 150
os.environ["TACO_DEC3_TRACE"] = str(STEP)
from tacotron_prosody_multi_gpu_b200 import synthetic, weights  # noqa: E402
from tacotron_prosody_multi_gpu_b200.engine import TacoEngine  # noqa: E402
from tacotron_prosody_multi_gpu_b200.hparams import default_hparams  # noqa: E402

B = int(sys.argv[1]) if len(sys.argv) > 1 else 64
S = int(sys.argv[2]) if len(sys.argv) > 2 else 200
CL = [int(x) for x in (sys.argv[4].split(",") if len(sys.argv) > 4 else ["1"])]
hp = default_hparams()
w = weights.set_stop_bias(weights.init_weights(hp, perturb=0.05), -10.0)
eng = TacoEngine(hp, w, precision="bf16")
ids, lens = synthetic.token_batch(hp, synthetic.cycle_lengths(B))
gw = synthetic.gst_random_weights(hp)
mem = eng.encode(torch.from_numpy(ids), torch.from_numpy(lens), gst_weights=gw)
for _ in range(2):
    d = eng.decode(mem, torch.from_numpy(lens), max_iters=S, philox_seed=1)
print("B=%d T=%d S=%d kernel %.3f ms -> %.2f us/step (trace mode adds pre-polls + barriers)" % (B, ids.shape[1], S, eng.stage_ms(1), 1e3 * eng.stage_ms(1) / S))
buf = np.zeros((128, 64, 2), np.int64)
eng._check(eng.L.taco_debug_read_workspace(eng._h, b"dec3_trace", buf.ctypes.data_as(C.c_void_p), buf.nbytes), "debug_read")
NAMES = {0: "step", 1: "P1.data", 2: "P1.mma", 3: "bar", 4: "P1.epi", 5: "bar", 6: "P1.sent", 7: "W1C", 8: "stop", 9: "bar+rearm", 10: "P2.data", 11: "P2.mma",
         12: "bar", 13: "P2.epi", 14: "bar", 15: "P2.sent", 16: "loc1", 17: "L1.data", 18: "L1.mma", 19: "bar", 20: "L1.epi", 21: "bar", 22: "L1.sent",
         23: "loc2", 24: "L2.data", 25: "L2.mma", 26: "bar", 27: "L2.epi", 28: "bar", 29: "q.mma", 30: "bar", 31: "L2.sent", 32: "Z.data", 33: "Z.mma",
         34: "q.recv", 35: "bar", 36: "energy", 37: "softmax", 38: "a.sent", 39: "cum+pairs", 40: "cs.bar", 41: "a.data", 42: "cs.fma", 43: "bar", 44: "cs.red",
         45: "bar", 46: "ctx.sent", 47: "end.bar"}
ghz = 1.965
for c in CL:
    blk = buf[16 * c:16 * c + 16]
    if blk[:, 0, 0].min() == 0:
        continue
    g0 = blk[:, 0, 1].min()
    print("cluster %d, step %d: microseconds since the earliest step start in the cluster; lens %s" % (c, STEP, lens[c * ((B + 6) // 7):(c + 1) * ((B + 6) // 7)].tolist()))
    for r in range(16):
        off = (blk[r, 0, 1] - g0) / 1e3
        out = []
        prev = 0.0
        for i in range(48):
            if blk[r, i, 0] == 0:
                continue
            t = off + (blk[r, i, 0] - blk[r, 0, 0]) / ghz / 1e3
            out.append("%s %.2f" % (NAMES[i], t))
        print("rank %2d: " % r + " | ".join(out))
