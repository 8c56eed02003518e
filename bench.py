#!/usr/bin/env python
"""Headline benchmark: mel frames/s of the Tacotron-2 mel-generation hot path (BASELINE.json).

One "step" = one pass of the whole hot path (embedding -> encoder convs -> BiLSTM -> GST style ->
persistent attention decoder -> postnet) over ONE batch of 64 synthetic utterances in free-running
mode (config C2 of BASELINE.json / SURVEY.md 8d): token lengths = 4 x the 16 lengths of the
reference's sentences.txt (T = 111), S = 720 decoder steps fixed by max_iters (stop bias -10 so the
batch-global stop never fires), r = 2 -> 92 160 mel frames per step per GPU.

  python bench.py [--gpus N] [--steps K] [--warmup W]            (our arm, N=1 default)
  torchrun ... bench.py --gpus N ...                              (N>1: one rank per GPU, weak scaling)
  python bench.py --impl reference ...                            (CPU restatement of the reference)

`value`      device-resident inputs, CUDA-event time of the K steps (max over ranks), `--in-flight` (default 2)
             independent batches in flight per GPU (engine.EnginePool): the decoder kernel of one batch overlaps
             the encoder / postnet / copies of the next; `stages_ms.one_batch_in_flight_ms_per_step` is the
             latency of a single batch.
`e2e`        same metric through the reference-facing API (Tacotron.initialize on HOST arrays):
             pinned-host -> device copies of ids/lengths and device -> host reads of
             mel_outputs, alignments and stop tokens inside the timed region.
`roofline`   dominant kernel = the persistent decoder kernel, timed live by CUDA events recorded
             around its launch inside the library (taco_stage_ms), in a pass with ONE batch in flight.
`cpu_baseline` the oracle (CPU restatement of the reference; TF 1.9 is not installable) on the
             box's host cores, rank 0, N=1 only.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "mel_frames_per_sec_free_running_b64"
UNIT = "mel frames/s"
B_PER_GPU = 64
S_STEPS = 720


def load_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return float(d["hbm_gbs"]), "measured"
    return 6650.0, "fallback"


def load_traffic(kernel):
    """DRAM bytes per launch of the dominant kernel from the committed ncu --set full capture."""
    p = os.path.join(ROOT, "profiles", "decoder_traffic.json")
    try:
        d = json.load(open(p))
        if d.get("kernel") and d["kernel"] in kernel:
            return int(d["dram_bytes_per_launch"])
    except (OSError, ValueError, KeyError):
        pass
    return None


def workload(hp, rank=0, batch=B_PER_GPU):
    from tacotron_prosody_multi_gpu_b200 import synthetic
    lengths = synthetic.cycle_lengths(batch)
    ids, lens = synthetic.token_batch(hp, lengths, seed=hp.tacotron_data_random_state + 1000 * rank)
    gw = synthetic.gst_random_weights(hp)
    return ids, lens, gw


def decoder_step_bytes(lens, T, B, key_bytes=4, value_bytes=4):
    """Algorithmic bytes of ONE decoder step (SURVEY 8d / DESIGN.md): keys+values of the valid
    positions read once AT THEIR STORED ELEMENT SIZE (fp32 keys; values fp32, or bf16 in the
    bf16 precision mode), cumulative alignments read+write and alignment history write
    (3 x 4 bytes per padded position), frame + stop outputs."""
    valid = int(np.sum(lens))
    return valid * (128 * key_bytes + 768 * value_bytes) + 12 * B * T + B * (160 + 2) * 4


class ClockSampler(object):
    """nvidia-smi clocks / throttle reasons sampled every 200 ms during the timed region."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.f = tempfile.NamedTemporaryFile("w+", suffix=".csv", delete=False)
        self.p = None
        try:
            self.p = subprocess.Popen(["nvidia-smi", "-i", str(index), "--query-gpu=" + self.Q,
                                       "--format=csv,noheader,nounits", "-lms", "200"],
                                      stdout=self.f, stderr=subprocess.DEVNULL)
        except OSError:
            self.p = None

    def stop(self):
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": []}
        if self.p is None:
            return out
        self.p.terminate()
        try:
            self.p.wait(timeout=5)
        except Exception:
            self.p.kill()
        self.f.flush()
        self.f.seek(0)
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in self.f.read().splitlines():
            c = [x.strip() for x in line.split(",")]
            if len(c) < 7:
                continue
            try:
                sm.append(float(c[0]))
                mx.append(float(c[1]))
            except ValueError:
                continue
            for n, v in zip(names, c[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        self.f.close()
        try:
            os.unlink(self.f.name)
        except OSError:
            pass
        if sm:
            out["sm_mhz"] = float(np.median(sm))
            out["sm_max_mhz"] = float(max(mx))
            out["samples"] = len(sm)
        out["reasons"] = sorted(reasons)
        return out


def cpu_reference_pass(hp, w, ids, lens, gw, S, threads):
    """One pass of the CPU restatement of the reference over the same batch (oracle/ is test
    infrastructure: this is the cpu_baseline / --impl reference leg only).  Returns seconds."""
    import torch
    from oracle.tacotron_oracle import TacotronOracle
    torch.set_num_threads(threads)
    orc = TacotronOracle(hp, w)
    t0 = time.perf_counter()
    out = orc.forward(ids, lens, gst_weights=gw, max_iters=S, philox_seed=hp.tacotron_random_seed)
    dt = time.perf_counter() - t0
    assert out["steps"] == S
    return dt, int(out["mel_outputs"].shape[0] * out["mel_outputs"].shape[1])


def run_reference(args):
    """--impl reference: the reference's own implementation needs tensorflow-gpu==1.9 (absent, not
    installable) -> time its CPU restatement (oracle) with all host threads on the same config."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    from tacotron_prosody_multi_gpu_b200 import weights
    from tacotron_prosody_multi_gpu_b200.hparams import default_hparams
    hp = default_hparams()
    w = weights.set_stop_bias(weights.init_weights(hp, perturb=0.05), -10.0)
    ids, lens, gw = workload(hp)
    cores = os.cpu_count() or 1
    # bounded sample: the same batch (B=64, T=111) decoded for S_sample <= 720 steps, sized from a
    # short probe so that the whole --steps/--warmup run stays within ~args.ref_budget_s seconds
    S = args.ref_steps
    if S <= 0:
        cpu_reference_pass(hp, w, ids, lens, gw, 8, cores)  # warm the CPU kernels
        probe_s, _ = cpu_reference_pass(hp, w, ids, lens, gw, 48, cores)
        per_step = max(probe_s / 48.0, 1e-4)
        S = int(args.ref_budget_s / ((args.warmup + args.steps) * per_step))
        S = max(8, min(S_STEPS, S))
    times = []
    frames = 0
    for i in range(args.warmup + args.steps):
        dt, frames = cpu_reference_pass(hp, w, ids, lens, gw, S, cores)
        if i >= args.warmup:
            times.append(dt)
    total = float(np.sum(times))
    value = frames * len(times) / total
    sample = "B=64,T=111 batch, %d of %d decoder steps per pass (encoder + decoder + postnet)" % (S, S_STEPS)
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * total / len(times),
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": "C2 free-running B=64 T=111 S=720 (sentences.txt lengths x4), CPU restatement of the "
                               "reference (TF 1.9 unavailable)", "sample_decoder_steps": S},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line))
    return 0


def run_b200(args):
    import torch
    import torch.distributed as dist
    from tacotron_prosody_multi_gpu_b200 import _lib, weights
    from tacotron_prosody_multi_gpu_b200.engine import TacoEngine
    from tacotron_prosody_multi_gpu_b200.hparams import default_hparams

    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device -- the B200 path has no CPU fallback (use --impl reference)")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)

    hp = default_hparams()
    w = weights.set_stop_bias(weights.init_weights(hp, perturb=0.05), -10.0)
    from tacotron_prosody_multi_gpu_b200.engine import EnginePool
    pool = EnginePool(hp, w, device=local_rank, precision=args.precision, in_flight=args.in_flight)
    eng = pool.engines[0]
    ids, lens, gw = workload(hp, rank)
    B, T = ids.shape
    S = args.decoder_steps
    r = hp.outputs_per_step
    frames_per_step = B * S * r
    seed = hp.tacotron_random_seed
    torch.cuda.set_device(dev)
    ids_d = torch.from_numpy(ids).to(dev)
    lens_d = torch.from_numpy(lens).to(dev)
    gw_d = torch.from_numpy(gw).to(dev)
    # one 512 MB (> 126 MB L2) flush buffer per engine: written on the engine's stream before each of its steps
    flush = [torch.empty(512 * 1024 * 1024 // 4, dtype=torch.float32, device=dev) for _ in range(len(pool))]
    fidx = {id(e): i for i, e in enumerate(pool.engines)}

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def device_step(e, _):
        flush[fidx[id(e)]].add_(1.0)
        return e.forward(ids_d, lens_d, gst_weights=gw_d, max_iters=S, philox_seed=seed)["steps"]

    def e2e_step(e, _):
        flush[fidx[id(e)]].add_(1.0)
        o = e.forward_host(ids, lens, gst_weights=gw, max_iters=S, philox_seed=seed)
        return o["steps"], int(o["h2d_bytes"]), int(o["d2h_bytes"])

    def timed(fn, n):
        """EXACTLY n steps, `in_flight` of them at a time (one engine + stream + host thread each); ONE pair of CUDA events
        on the default stream brackets the region: every engine stream waits for the first, the second is recorded after all
        of them were joined.  The L2 flush of each step runs on the step's own stream, inside the timed region."""
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        res = pool.map(fn, range(n), after=e0)
        for st in pool.streams:
            torch.cuda.current_stream().wait_stream(st)
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1), res

    # ---- device-resident arm ---------------------------------------------------------------
    timed(device_step, max(args.warmup, len(pool)))
    barrier()
    sampler = ClockSampler(local_rank) if rank == 0 else None
    l0 = pool.launch_count
    ms_total, steps_seen = timed(device_step, args.steps)
    launches = pool.launch_count - l0
    barrier()
    assert all(x == S for x in steps_seen), "stop fired in a throughput run"
    ms_total = max_over_ranks(ms_total)

    # ---- end-to-end arm: host buffers in, host buffers out -----------------------------------
    timed(e2e_step, max(min(args.warmup, 2), len(pool)))
    barrier()
    e2e_ms, eo = timed(e2e_step, args.steps)
    barrier()
    e2e_ms = max_over_ranks(e2e_ms)
    h2d, d2h = eo[-1][1], eo[-1][2]

    # ---- one batch at a time: per-stage and decoder-kernel durations for the roofline (CUDA events around each kernel
    #      inside the library; with several batches in flight they would include the wait for free SMs) ----
    dec_ms = []
    seq_ms = 0.0
    for _ in range(3):
        flush[0].add_(1.0)
        torch.cuda.synchronize()
        q0, q1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        q0.record()
        eng.forward(ids_d, lens_d, gst_weights=gw_d, max_iters=S, philox_seed=seed)
        q1.record()
        torch.cuda.synchronize()
        seq_ms = q0.elapsed_time(q1)
        dec_ms.append(eng.stage_ms(_lib.TACO_STAGE_DECODER_KERNEL))
    dec_ms = dec_ms[1:]
    stage = {"encode_ms": eng.stage_ms(_lib.TACO_STAGE_ENCODE), "decoder_kernel_ms": float(np.mean(dec_ms)),
             "postnet_ms": eng.stage_ms(_lib.TACO_STAGE_POSTNET), "one_batch_in_flight_ms_per_step": seq_ms}
    clocks = sampler.stop() if sampler is not None else None

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return 0

    peak, peak_src = load_peaks()
    step_bytes = decoder_step_bytes(lens, T, B, 4, 2 if args.precision == "bf16" else 4)
    dec_s = float(np.mean(dec_ms)) * 1e-3
    achieved = step_bytes * S / dec_s / 1e9
    line = {
        "metric": METRIC, "value": world * frames_per_step * args.steps / (ms_total * 1e-3), "unit": UNIT,
        "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_total / args.steps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "bf16" if args.precision == "bf16" else "f32", "data": "synthetic",
        "precision": ("bf16 operands + fp32 accumulation in every contraction (tcgen05 in the postnet, mma.sync in the decoder step); "
                      "attention values streamed as bf16; recurrent state, softmax, location term and all outputs fp32") if args.precision == "bf16" else "fp32 everywhere",
        "config": {"workload": "C2 free-running inference, B=64/GPU, sentences.txt token lengths x4 (T=111), "
                               "S=%d decoder steps (max_iters, stop bias -10), r=2, random-init seed 45454" % S,
                   "batch_per_gpu": B, "T": T, "decoder_steps": S, "frames_counted": "all B*2S frames (no stop fires)",
                   "l2": "512 MB flush write before every step, on the step's stream, inside the timed region",
                   "in_flight": args.in_flight,
                   "parallelism": "independent batches per GPU, no collective; %d batches in flight per GPU (one engine handle, CUDA "
                                  "stream and host thread each): the encoder / postnet / copies of one batch overlap the persistent "
                                  "decoder kernel of another" % args.in_flight},
        "decoder_step_us": 1e6 * dec_s / S,
        "stages_ms": stage,
        "roofline": {"bound": "hbm", "kernel": "decoder3_kernel (cluster persistent decoder)" if args.precision == "bf16" else "decoder_persistent_kernel", "achieved": achieved, "peak": peak,
                     "peak_source": peak_src + " (burst copy bandwidth)", "unit": "GB/s", "frac": achieved / peak,
                     "algorithmic_bytes_per_launch": step_bytes * S, "launch_ms": dec_s * 1e3,
                     "traffic": load_traffic("decoder3_kernel" if args.precision == "bf16" else "decoder_persistent_kernel")},
        "e2e": {"value": world * frames_per_step * args.steps / (e2e_ms * 1e-3), "unit": UNIT,
                "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h, "ms_per_step": e2e_ms / args.steps},
        "gpu_launches": int(launches),
        "clocks": clocks,
    }
    if world == 1 and not args.no_cpu_baseline:
        cores = os.cpu_count() or 1
        Sc = args.cpu_steps
        dt, fr = cpu_reference_pass(hp, w, ids, lens, gw, Sc, cores)
        line["cpu_baseline"] = {"value": fr / dt, "unit": UNIT, "cores": cores, "kind": "port",
                                "sample": "one pass of the same B=64,T=111 batch, %d of %d decoder steps, torch-CPU fp32 "
                                          "restatement of the reference (TF 1.9 unavailable), %.1f s" % (Sc, S, dt)}
    print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--decoder-steps", type=int, default=S_STEPS)
    ap.add_argument("--cpu-steps", type=int, default=S_STEPS, help="decoder steps of the cpu_baseline sample")
    ap.add_argument("--ref-steps", type=int, default=0, help="--impl reference: decoder steps per pass (0 = auto)")
    ap.add_argument("--ref-budget-s", type=float, default=240.0)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--in-flight", type=int, default=2, help="independent batches in flight per GPU (engine.EnginePool)")
    ap.add_argument("--precision", default="bf16", choices=["bf16", "fp32"],
                    help="bf16: tensor-core (tcgen05) contractions with bf16 operands, fp32 accumulate; fp32: SIMT everywhere")
    args = ap.parse_args()
    if args.warmup < 3 and args.impl == "b200":
        args.warmup = 3
    if args.impl == "reference":
        return run_reference(args)
    return run_b200(args)


if __name__ == "__main__":
    sys.exit(main())
