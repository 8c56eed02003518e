"""Probe: do two batches in flight (two engines, two streams, two host threads) overlap one batch's encoder / postnet with
the other's persistent decoder?  Prints ms per batch for 1 and 2 batches in flight."""
import os
import sys
import threading

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from tacotron_prosody_multi_gpu_b200 import synthetic, weights  # noqa: E402
from tacotron_prosody_multi_gpu_b200.engine import TacoEngine  # noqa: E402
from tacotron_prosody_multi_gpu_b200.hparams import default_hparams  # noqa: E402


def main():
    K = int(sys.argv[1]) if len(sys.argv) > 1 else 12
    hp = default_hparams()
    w = weights.set_stop_bias(weights.init_weights(hp, perturb=0.05), -10.0)
    dev = torch.device("cuda", 0)
    lengths = synthetic.cycle_lengths(64)
    ids, lens = synthetic.token_batch(hp, lengths, seed=hp.tacotron_data_random_state)
    gw = synthetic.gst_random_weights(hp)
    ids_d, lens_d, gw_d = (torch.from_numpy(x).to(dev) for x in (ids, lens, gw))
    for n_flight in (1, 2):
        engs = [TacoEngine(hp, w, device=0, precision="bf16") for _ in range(n_flight)]
        streams = [torch.cuda.Stream(device=dev) for _ in range(n_flight)]

        def worker(i, n):
            with torch.cuda.stream(streams[i]):
                for _ in range(n):
                    engs[i].forward(ids_d, lens_d, gst_weights=gw_d, max_iters=720, philox_seed=1)
                streams[i].synchronize()

        for rep in range(2):   # first repetition = warm-up
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for s in streams:
                s.wait_event(e0)
            ts = [threading.Thread(target=worker, args=(i, K // n_flight)) for i in range(n_flight)]
            for t in ts:
                t.start()
            for t in ts:
                t.join()
            for s in streams:
                torch.cuda.current_stream().wait_stream(s)
            e1.record()
            torch.cuda.synchronize()
        print("in flight %d: %.3f ms per batch" % (n_flight, e0.elapsed_time(e1) / (K // n_flight * n_flight)))
        for e in engs:
            e.close()


if __name__ == "__main__":
    main()
