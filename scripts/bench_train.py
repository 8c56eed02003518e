#!/usr/bin/env python
"""Config C4 of BASELINE.json: teacher-forced training step (forward + backward + clip + Adam), batch 32 per GPU,
data-parallel over N GPUs with ONE NCCL all-reduce of the flat gradient buffer per step.

  python scripts/bench_train.py [--steps K] [--warmup W] [--batch 32] [--max-frames 1000]
  torchrun --nproc-per-node N scripts/bench_train.py ...

Prints one JSON line (rank 0): training steps/s and mel frames/s (whole job), per-phase device milliseconds
(CUDA events on the launching stream), kernel launches per step, and -- at N=1 -- the CPU restatement of the
reference (autograd through the oracle) on a bounded sample of the same workload."""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def workload(hp, batch, max_frames, rank):
    from tacotron_prosody_multi_gpu_b200 import synthetic
    lengths = synthetic.cycle_lengths(batch)
    ids, lens = synthetic.token_batch(hp, lengths, seed=hp.tacotron_data_random_state + 1000 * rank)
    mel, stop, flen = synthetic.mel_targets(hp, lens, max_frames=max_frames, seed=hp.tacotron_data_random_state + 1 + 1000 * rank)
    return ids, lens, mel, stop


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=2)
    ap.add_argument("--batch", type=int, default=32)
    ap.add_argument("--max-frames", type=int, default=1000)
    ap.add_argument("--precision", default="bf16", choices=["bf16", "fp32"],
                    help="bf16: large contractions on mma.sync tensor cores (bf16 operands, fp32 accumulate); fp32: SIMT parity mode")
    ap.add_argument("--cpu-sample", type=int, default=1, help="1: time the oracle on a bounded sample (N=1 only)")
    args = ap.parse_args()
    import torch
    import torch.distributed as dist
    from tacotron_prosody_multi_gpu_b200 import weights
    from tacotron_prosody_multi_gpu_b200.hparams import default_hparams
    from tacotron_prosody_multi_gpu_b200.train.trainer import DataParallelTrainer

    rank, local, world = (int(os.environ.get(k, d)) for k, d in (("RANK", 0), ("LOCAL_RANK", 0), ("WORLD_SIZE", 1)))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
    hp = default_hparams()
    w = weights.init_weights(hp)
    tr = DataParallelTrainer(hp, w, device=local)
    tr.ops.set_precision(args.precision)
    ids, lens, mel, stop = workload(hp, args.batch, args.max_frames, rank)
    o = tr.ops
    d_ids, d_lens, d_mel, d_stop = o.to_device(ids), o.to_device(lens), o.to_device(mel), o.to_device(stop)
    B, T = ids.shape
    F = mel.shape[1]
    st = tr.step_fn

    def ev():
        e = torch.cuda.Event(enable_timing=True)
        e.record()
        return e

    phases = {"forward_ms": 0.0, "backward_ms": 0.0, "exchange_apply_ms": 0.0}
    losses = []
    total_ms = 0.0
    launches = 0
    wall = 0.0
    for it in range(args.warmup + args.steps):
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        l0 = o.launches
        t0 = time.perf_counter()
        e0 = ev()
        out = st.forward(d_ids, d_lens, d_mel, d_stop, hp.tacotron_random_seed + it)
        e1 = ev()
        st.backward()
        e2 = ev()
        st.apply(tr.global_step, world=world, all_reduce=tr._all_reduce if world > 1 else None)
        e3 = ev()
        torch.cuda.synchronize()
        tr.global_step += 1
        if it >= args.warmup:
            wall += time.perf_counter() - t0
            phases["forward_ms"] += e0.elapsed_time(e1)
            phases["backward_ms"] += e1.elapsed_time(e2)
            phases["exchange_apply_ms"] += e2.elapsed_time(e3)
            total_ms += e0.elapsed_time(e3)
            launches = o.launches - l0
        losses.append([float(x) for x in out["losses"].cpu().numpy()[:4]])
    if world > 1:
        t = torch.tensor([total_ms], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        total_ms = float(t.item())
    if rank == 0:
        k = args.steps
        line = {"metric": "train_steps_per_sec_b%d_per_gpu" % args.batch, "value": k / (total_ms * 1e-3), "unit": "steps/s",
                "mel_frames_per_sec": world * B * F * k / (total_ms * 1e-3), "n_gpus": world, "steps": k, "warmup": args.warmup,
                "ms_per_step": total_ms / k, "host_wall_ms_per_step": 1e3 * wall / k, "scaling": "weak", "dtype": "bf16" if args.precision == "bf16" else "f32",
                "precision": "bf16 operands + fp32 accumulation in the conv / all-rows contractions (mma.sync); recurrent per-step contractions, "
                             "state, normalisation, losses, optimizer fp32" if args.precision == "bf16" else "fp32 everywhere",
                "data": "synthetic", "config": {"workload": "C4 teacher-forced training step fwd+bwd+clip+Adam, B=%d/GPU, T=%d, F=%d (S=%d decoder steps), "
                                                "random-init seed 45454, one all-reduce of %d fp32 gradients per step" % (B, T, F, F // 2, st.store.n)},
                "phases_ms": {kk: v / k for kk, v in phases.items()}, "gpu_launches": launches,
                "losses_first_last": [losses[0], losses[-1]]}
        if world == 1 and args.cpu_sample:
            from oracle import train_oracle
            torch.set_num_threads(os.cpu_count() or 1)
            Bs, Fs = 4, 100
            t0 = time.perf_counter()
            out, orc = train_oracle.train_forward(hp, w, ids[:Bs], lens[:Bs], mel[:Bs, :Fs], stop[:Bs, :Fs], 1)
            out["loss"].backward()
            dt = time.perf_counter() - t0
            line["cpu_baseline"] = {"value": Bs * Fs / dt, "unit": "mel frames/s (training)", "cores": os.cpu_count(), "kind": "port",
                                    "sample": "autograd through the oracle, B=%d, F=%d frames of the same batch, %.1f s" % (Bs, Fs, dt)}
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
