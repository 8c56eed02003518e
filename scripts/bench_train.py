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



def _claim_stdout():
    """Only the JSON line may reach stdout: libraries write to fd 1 behind Python's back (NCCL prints its version banner there).
    fd 1 is pointed at stderr for the rest of the process; the returned file is the real stdout."""
    sys.stdout.flush()
    saved = os.dup(1)
    os.dup2(2, 1)
    return os.fdopen(saved, "w")


OUT = None          # the real stdout (set by main)


def _emit(line):
    out = OUT if OUT is not None else sys.stdout
    print(json.dumps(line), file=out, flush=True)


def workload(hp, batch, max_frames, rank):
    from tacotron_prosody_multi_gpu_b200 import synthetic
    lengths = synthetic.cycle_lengths(batch)
    ids, lens = synthetic.token_batch(hp, lengths, seed=hp.tacotron_data_random_state + 1000 * rank)
    mel, stop, flen = synthetic.mel_targets(hp, lens, max_frames=max_frames, seed=hp.tacotron_data_random_state + 1 + 1000 * rank)
    return ids, lens, mel, stop


METRIC = "train_steps_per_sec_b%d_per_gpu"


def c4_config(B, T, F, n_grad):
    return {"workload": "C4 teacher-forced training step fwd+bwd+clip+Adam, B=%d/GPU, T=%d, F=%d (S=%d decoder steps), "
                        "random-init seed 45454, one flat bucket of %d fp32 gradients (+ BN statistics) all-reduced per step" % (B, T, F, F // 2, n_grad)}


def run_reference(args):
    """--impl reference for C4: the reference's training graph needs TF 1.9 -> autograd through the CPU restatement (oracle/),
    all host threads, on a bounded sample (B=4, F=100 of the same batch); frames/s scaled to steps/s of the full step."""
    if int(os.environ.get("RANK", "0")) != 0:
        return 0
    import torch
    from oracle import train_oracle
    from tacotron_prosody_multi_gpu_b200 import weights
    from tacotron_prosody_multi_gpu_b200.hparams import default_hparams
    hp = default_hparams()
    w = weights.init_weights(hp)
    ids, lens, mel, stop = workload(hp, args.batch, args.max_frames, 0)
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    Bs, Fs = 4, 100
    times = []
    for it in range(args.warmup + args.steps):
        t0 = time.perf_counter()
        out, orc = train_oracle.train_forward(hp, w, ids[:Bs], lens[:Bs], mel[:Bs, :Fs], stop[:Bs, :Fs], 1 + it)
        out["loss"].backward()
        if it >= args.warmup:
            times.append(time.perf_counter() - t0)
    fps = Bs * Fs * len(times) / float(np.sum(times))
    B, T, F = ids.shape[0], ids.shape[1], mel.shape[1]
    value = fps / (B * F)
    sample = "autograd through the oracle, B=%d, F=%d frames of the same batch per pass; steps/s = sample frames/s / (%d x %d)" % (Bs, Fs, B, F)
    n_grad = 20425491
    _emit({"impl": "reference", "metric": METRIC % args.batch, "value": value, "unit": "steps/s", "mel_frames_per_sec": fps,
                      "n_gpus": int(os.environ.get("WORLD_SIZE", "1")), "steps": args.steps, "warmup": args.warmup,
                      "ms_per_step": 1e3 / value, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
                      "data": "synthetic", "config": c4_config(B, T, F, n_grad),
                      "cpu_baseline": {"value": value, "unit": "steps/s", "cores": cores, "kind": "port", "sample": sample},
                      "e2e": {"value": value, "unit": "steps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}, "gpu_launches": 0})
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=2)
    ap.add_argument("--batch", type=int, default=32)
    ap.add_argument("--max-frames", type=int, default=1000)
    ap.add_argument("--precision", default="bf16", choices=["bf16", "fp32"],
                    help="bf16: large contractions on mma.sync tensor cores (bf16 operands, fp32 accumulate); fp32: SIMT parity mode")
    ap.add_argument("--cpu-sample", type=int, default=1, help="1: time the oracle on a bounded sample (N=1 only)")
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"],
                    help="reference: autograd through the CPU restatement (oracle/) on a bounded sample, rank 0 only")
    ap.add_argument("--host-inputs", type=int, default=1, help="1: every timed step copies its batch from pinned host memory (e2e form)")
    ap.add_argument("--phases", action="store_true", help="print time stamps of the phases of the last step to stderr (CUDA events)")
    ap.add_argument("--no-overlap", action="store_true", help="one all-reduce of the whole bucket after backward (no early half)")
    args = ap.parse_args()
    global OUT
    if OUT is None:
        OUT = _claim_stdout()
    if args.impl == "reference":
        return run_reference(args)
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
    tr.overlap = not args.no_overlap
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

    p_ids, p_lens, p_mel, p_stop = (torch.from_numpy(np.ascontiguousarray(a)).pin_memory() for a in (ids, lens, mel, stop))
    h2d_bytes = sum(int(t.numel() * t.element_size()) for t in (p_ids, p_lens, p_mel, p_stop))
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
        if args.host_inputs:      # the e2e form: this step's batch comes from pinned host memory
            d_ids.copy_(p_ids, non_blocking=True); d_lens.copy_(p_lens, non_blocking=True)
            d_mel.copy_(p_mel, non_blocking=True); d_stop.copy_(p_stop, non_blocking=True)
        st.marks = [] if args.phases else None
        out = st.forward(d_ids, d_lens, d_mel, d_stop, hp.tacotron_random_seed + it)
        e1 = ev()
        st.backward()
        e2 = ev()
        st.apply(tr.global_step, world=world, all_reduce=tr._all_reduce if world > 1 else None)
        if args.host_inputs:
            loss_host = out["losses"].cpu()      # the step's result read back inside the timed region (e2e form)
        e3 = ev()
        torch.cuda.synchronize()
        if args.phases and rank == 0 and it == args.warmup + args.steps - 1:
            m = st.marks
            for n1, b2 in m[1:]:          # time stamps relative to the first marker ([side] = recorded on a side stream)
                print("[phase] %-52s at %7.3f ms" % (n1, m[0][1].elapsed_time(b2)), file=sys.stderr)
            print("[phase] %-52s at %7.3f ms" % ("backward done", m[0][1].elapsed_time(e2)), file=sys.stderr)
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
    # ---- the exchange alone: K all-reduces of the whole flat bucket (statistics + gradients), CUDA events, max over ranks ----
    exch = None
    if world > 1:
        bucket = st.store.bucket
        scratch = torch.empty_like(bucket)
        for _ in range(3):
            dist.all_reduce(scratch)
        torch.cuda.synchronize()
        dist.barrier()
        a0 = ev()
        for _ in range(10):
            dist.all_reduce(scratch)
        a1 = ev()
        torch.cuda.synchronize()
        t = torch.tensor([a0.elapsed_time(a1) / 10], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ar_ms = float(t.item())
        nbytes = bucket.numel() * 4
        exch = {"allreduce_ms": ar_ms, "bucket_bytes": nbytes, "algbw_GBps": nbytes / (ar_ms * 1e-3) / 1e9,
                "busbw_GBps": 2.0 * (world - 1) / world * nbytes / (ar_ms * 1e-3) / 1e9,
                "overlap": "early half (decoder + postnet gradients, %d of %d floats) all-reduced on a side stream from the end of the "
                           "decoder BPTT; the rest after backward" % (st.store.n - st.store.tail_start, st.store.ns + st.store.n)
                           if tr.overlap else "none: one all-reduce after backward"}
        del scratch
    if rank == 0:
        k = args.steps
        val = world * k / (total_ms * 1e-3)
        line = {"metric": METRIC % args.batch, "value": val, "unit": "steps/s", "higher_is_better": True, "vs_baseline": None,
                "mel_frames_per_sec": world * B * F * k / (total_ms * 1e-3), "n_gpus": world, "steps": k, "warmup": args.warmup,
                "ms_per_step": total_ms / k, "host_wall_ms_per_step": 1e3 * wall / k, "scaling": "weak", "dtype": "bf16" if args.precision == "bf16" else "f32",
                "value_note": "whole-job training steps/s = ranks x (1 / step time): every rank completes one B=%d step per iteration" % B,
                "e2e": {"value": val, "unit": "steps/s", "h2d_bytes_per_step": h2d_bytes if args.host_inputs else 0, "d2h_bytes_per_step": 20,
                        "note": "the timed step copies its batch from pinned host memory and the 5 losses are read back" if args.host_inputs else "device-resident batch"},
                "exchange": exch,
                "precision": "bf16 operands + fp32 accumulation in the conv / all-rows contractions (mma.sync); recurrent per-step contractions, "
                             "state, normalisation, losses, optimizer fp32" if args.precision == "bf16" else "fp32 everywhere",
                "data": "synthetic", "config": c4_config(B, T, F, st.store.n_params),
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
        _emit(line)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
