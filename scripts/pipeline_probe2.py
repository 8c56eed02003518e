#!/usr/bin/env python
"""Which part of the two-batches-in-flight pipeline bounds a batch: decoders only / + postnet / + encoder, two engines, two threads."""
import os, sys, threading, time
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from tacotron_prosody_multi_gpu_b200 import synthetic, weights
from tacotron_prosody_multi_gpu_b200.engine import EnginePool
from tacotron_prosody_multi_gpu_b200.hparams import default_hparams

hp = default_hparams()
w = weights.set_stop_bias(weights.init_weights(hp, perturb=0.05), -10.0)
pool = EnginePool(hp, w, precision="bf16", in_flight=2)
ids, lens = synthetic.token_batch(hp, synthetic.cycle_lengths(64))
gw = synthetic.gst_random_weights(hp)
dev = torch.device("cuda", 0)
ids_d, lens_d, gw_d = torch.from_numpy(ids).to(dev), torch.from_numpy(lens).to(dev), torch.from_numpy(gw).to(dev)
S = 720
mems = {}
for e in pool.engines:
    mems[id(e)] = e.encode(ids_d, lens_d, gst_weights=gw_d)
torch.cuda.synchronize()

def dec_only(e, _):
    return e.decode(mems[id(e)], lens_d, max_iters=S, philox_seed=1)["steps"]
def dec_post(e, _):
    d = e.decode(mems[id(e)], lens_d, max_iters=S, philox_seed=1)
    e.postnet(d["frames"], 2 * d["steps"])
    return d["steps"]
def enc_dec(e, _):
    m = e.encode(ids_d, lens_d, gst_weights=gw_d)
    return e.decode(m, lens_d, max_iters=S, philox_seed=1)["steps"]
def full(e, _):
    return e.forward(ids_d, lens_d, gst_weights=gw_d, max_iters=S, philox_seed=1)["steps"]

for name, fn in (("decoders only", dec_only), ("decoder + postnet", dec_post), ("encoder + decoder", enc_dec), ("full forward", full)):
    pool.map(fn, range(4))
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    n = 10
    pool.map(fn, range(n))
    torch.cuda.synchronize()
    print("%-20s %.2f ms per batch" % (name, 1e3 * (time.perf_counter() - t0) / n))
