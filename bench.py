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


def config_dict(S, in_flight, batch=B_PER_GPU, T=111):
    """The workload description BOTH arms print (the driver compares the two `config` objects)."""
    return {"workload": "C2 free-running inference, B=%d/GPU, sentences.txt token lengths x4 (T=%d), S=%d decoder steps "
                        "(max_iters, stop bias -10), r=2, random-init seed 45454" % (batch, T, S),
            "batch_per_gpu": batch, "T": T, "decoder_steps": S, "frames_counted": "all B*2S frames (no stop fires)",
            "l2": "512 MB flush write before every step, on the step's stream, inside the timed region",
            "in_flight": in_flight,
            "parallelism": "independent batches per GPU, no collective; %d batches in flight per GPU (one engine handle, CUDA "
                           "stream and host thread each): the encoder / postnet / copies of one batch overlap the persistent "
                           "decoder kernel of another" % in_flight}


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
        "config": config_dict(S_STEPS, args.in_flight),
        "reference_arm": "CPU restatement of the reference (oracle/; TensorFlow 1.9 is not installable here), %d host threads, "
                         "%d of %d decoder steps per pass" % (cores, S, S_STEPS),
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    _emit(line)
    return 0


C5_LENGTHS = [300, 287, 300, 251, 199, 300, 123, 64, 300, 18, 277, 300, 150, 290, 300, 42]


def _event_ms(fn, n):
    import torch
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        out = fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n, out


def c3_stop_bias(eng, hp, ids, lens, gw, S_cap, seed, scale=8.0):
    """C3 stop behaviour with random-init weights: the stop kernel is scaled x8 and the bias is placed, from ONE
    calibration batch decoded to S_cap steps without stopping (raw logits, bias -30), in the widest gap between the running
    maximum of the per-step max logit and a later step that exceeds it, in the second half of the sequence.  Every other
    batch then stops wherever ITS logits first cross that threshold (batch-global stop, Helpers.py:51-52)."""
    import torch
    mem = eng.encode(torch.from_numpy(ids), torch.from_numpy(lens), gst_weights=gw)
    d = eng.decode(mem, torch.from_numpy(lens), max_iters=S_cap, philox_seed=seed, logits_only=True)
    logit = d["stop"].double().cpu().numpy() + 30.0
    m = logit.reshape(logit.shape[0], S_cap, -1).max(axis=(0, 2))
    best = None
    run = m[: S_cap // 2].max()
    for s in range(S_cap // 2, S_cap - 1):
        if m[s] > run and (best is None or m[s] - run > best[0]):
            best = (m[s] - run, s, -(m[s] + run) / 2.0)
        run = max(run, m[s])
    if best is None:
        return -float(m.max()) + 1e-3, S_cap
    return float(best[2]), best[1] + 1


def extra_configs(args, hp, w, dev, rank, world, dist):
    """The other configurations of BASELINE.json beside the headline: C2 B=1 step latency, C5 long-form,
    C3 = 4096 utterances sharded over the ranks with stop-driven lengths (strong scaling, valid frames), fp32-mode C2."""
    import torch
    from tacotron_prosody_multi_gpu_b200 import _lib, sharding, synthetic, weights
    from tacotron_prosody_multi_gpu_b200.engine import EnginePool, TacoEngine
    local = dev.index
    out = {}
    gw = synthetic.gst_random_weights(hp)
    seed = hp.tacotron_random_seed
    if rank == 0:
        # ---- C2, batch 1 (synthesize.py --mode=inference runs one sentence at a time): decoder step latency ----
        c2 = {}
        for prec in ("bf16", "fp32"):
            eng = TacoEngine(hp, w, device=local, precision=prec)
            for T1, S1 in ((42, 277), (111, 720)):
                ids, lens = synthetic.token_batch(hp, [T1])
                a = (torch.from_numpy(ids).to(dev), torch.from_numpy(lens).to(dev))
                for _ in range(2):
                    eng.forward(*a, gst_weights=gw, max_iters=S1, philox_seed=seed)
                ms, o = _event_ms(lambda: eng.forward(*a, gst_weights=gw, max_iters=S1, philox_seed=seed), 3)
                assert o["steps"] == S1
                dk = eng.stage_ms(_lib.TACO_STAGE_DECODER_KERNEL)
                c2["%s_T%d_S%d" % (prec, T1, S1)] = {"decoder_step_us": 1e3 * dk / S1, "utterance_ms": ms, "mel_frames_per_sec": 2 * S1 / (ms * 1e-3)}
            if prec == "fp32":
                # ---- the headline batch in fp32 mode (exact-parity kernels), one batch in flight ----
                ids, lens, _ = workload(hp, rank)
                a = (torch.from_numpy(ids).to(dev), torch.from_numpy(lens).to(dev))
                S = args.decoder_steps
                eng.forward(*a, gst_weights=gw, max_iters=S, philox_seed=seed)
                ms, o = _event_ms(lambda: eng.forward(*a, gst_weights=gw, max_iters=S, philox_seed=seed), 2)
                dk = eng.stage_ms(_lib.TACO_STAGE_DECODER_KERNEL)
                out["c2_b64_fp32"] = {"value": ids.shape[0] * 2 * S / (ms * 1e-3), "unit": UNIT, "ms_per_step": ms, "decoder_step_us": 1e3 * dk / S,
                                      "dtype": "f32", "in_flight": 1}
            eng.close()
        out["c2_b1"] = c2
        # ---- C5 long form: B=16, T=300, reference encoder on [16, 2000, 80], 1000 free-running steps ----
        eng = TacoEngine(hp, w, device=local, precision="bf16")
        ids, lens = synthetic.token_batch(hp, C5_LENGTHS)
        rm = torch.from_numpy(synthetic.reference_mel(hp, len(C5_LENGTHS), 2000)).to(dev)
        a = (torch.from_numpy(ids).to(dev), torch.from_numpy(lens).to(dev))
        for _ in range(2):
            eng.forward(*a, reference_mel=rm, max_iters=1000, philox_seed=5)
        ms, o = _event_ms(lambda: eng.forward(*a, reference_mel=rm, max_iters=1000, philox_seed=5), 3)
        assert o["steps"] == 1000
        dk = eng.stage_ms(_lib.TACO_STAGE_DECODER_KERNEL)
        out["c5"] = {"workload": "B=16, T=300, reference mel [16,2000,80] through the reference encoder + GST, S=1000 free-running steps",
                     "mel_frames_per_sec": 16 * 2000 / (ms * 1e-3), "ms_per_batch": ms, "decoder_step_us": 1e3 * dk / 1000,
                     "encode_ms": eng.stage_ms(_lib.TACO_STAGE_ENCODE), "postnet_ms": eng.stage_ms(_lib.TACO_STAGE_POSTNET), "dtype": "bf16"}
        eng.close()
    # ---- C3: 4096 utterances, 64 batches of 64 sorted by length, LPT-assigned to the ranks, stop-driven S per batch ----
    n_utt, S_cap = args.c3_utterances, args.decoder_steps
    ulens = synthetic.utterance_set(hp, n_utt)
    rng = np.random.Generator(np.random.PCG64(hp.tacotron_data_random_state + 7))
    seqs = [np.concatenate([rng.integers(2, 66, size=int(n) - 1), [1]]).astype(np.int32) for n in ulens]
    bias_t = torch.zeros(1, dtype=torch.float64, device=dev)
    if rank == 0:
        cal = TacoEngine(hp, weights.scale_stop_projection(w, 8.0, -30.0), device=local, precision="bf16")
        batches = sharding.plan_batches([len(x) for x in seqs], 64)
        from tacotron_prosody_multi_gpu_b200.Utils.Tacotron_feeder import _prepare_inputs
        idx = batches[len(batches) // 2]
        cids, _ = _prepare_inputs([seqs[i] for i in idx])
        bias, s_cal = c3_stop_bias(cal, hp, cids.astype(np.int32), np.asarray([len(seqs[i]) for i in idx], np.int32), gw, S_cap, seed)
        cal.close()
        bias_t[0] = bias
    if world > 1:
        dist.broadcast(bias_t, 0)
    bias = float(bias_t.item())
    pool = EnginePool(hp, weights.scale_stop_projection(w, 8.0, bias), device=local, precision="bf16", in_flight=args.in_flight)
    longest = sorted(range(n_utt), key=lambda i: -len(seqs[i]))[:64 * max(2, args.in_flight)]
    sharding.synthesize_batches(pool, hp, [seqs[i] for i in longest], 0, 1, 64, gw, S_cap, seed)   # warm-up: the largest batches on every engine
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter()
    e0.record()
    store = None
    if world > 1 and not args.c3_static:
        store = dist.distributed_c10d._get_default_store()   # control plane only: a shared "next batch" counter
    res = sharding.synthesize_batches(pool, hp, seqs, rank, world, 64, gw, S_cap, seed, store=store, store_key="taco_b200/c3/%d" % args.steps)
    for st in pool.streams:
        torch.cuda.current_stream().wait_stream(st)
    e1.record()
    torch.cuda.synchronize()
    busy_ms = e0.elapsed_time(e1)
    wall_ms = 1e3 * (time.perf_counter() - t0)
    valid = sum(int(v["mel"].shape[0]) for v in res.values())
    steps_by_batch = {}
    for v in res.values():
        steps_by_batch[v["batch"]] = v["steps"]
    padded = sum(64 * 2 * sN for sN in steps_by_batch.values())
    stat = torch.tensor([busy_ms, float(valid), float(padded), float(len(res)), float(sum(steps_by_batch.values())), wall_ms], dtype=torch.float64, device=dev)
    if world > 1:
        allst = [torch.zeros_like(stat) for _ in range(world)]
        dist.all_gather(allst, stat)
    else:
        allst = [stat]
    pool.close()
    if rank == 0:
        a = np.stack([x.cpu().numpy() for x in allst])
        job_ms = float(a[:, 0].max())
        steps_list = sorted(steps_by_batch.values())
        plan = "static LPT plan by estimated cost" if (world == 1 or args.c3_static) else "claimed dynamically from a shared counter in the rendezvous store"
        out["c3"] = {"workload": "%d utterances (256 x the sentences.txt lengths, shuffled seed 1324), sorted into batches of 64, assigned to "
                                 "%d rank(s) (%s), no collective; stop kernel x8, bias %.3f calibrated on one batch: every batch stops at its own "
                                 "first any-row crossing (cap %d steps); host token arrays in, trimmed host mels out" % (n_utt, world, plan, bias, S_cap),
                     "scaling": "strong", "utterances": int(a[:, 3].sum()), "valid_mel_frames_per_sec": float(a[:, 1].sum()) / (job_ms * 1e-3),
                     "padded_mel_frames_per_sec": float(a[:, 2].sum()) / (job_ms * 1e-3), "job_ms": job_ms,
                     "rank_busy_ms": [float(x) for x in a[:, 0]], "rank_decoder_steps": [int(x) for x in a[:, 4]],
                     "imbalance_max_over_mean": float(a[:, 0].max() / a[:, 0].mean()),
                     "valid_frames": int(a[:, 1].sum()), "padded_frames": int(a[:, 2].sum()),
                     "rank0_batch_steps_min_median_max": [int(steps_list[0]), int(steps_list[len(steps_list) // 2]), int(steps_list[-1])],
                     "in_flight": args.in_flight}
    return out


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
    if os.environ.get("TACO_BLOCKING_SYNC", "1" if world > 1 else "0") == "1":
        # host threads that wait for the GPU sleep instead of spinning (N ranks x (in_flight + 1) threads share the host cores)
        import ctypes
        ctypes.CDLL("libcudart.so").cudaSetDeviceFlags(4)   # cudaDeviceScheduleBlockingSync, before the context exists
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
    pool.close()
    del flush
    torch.cuda.empty_cache()
    extras = None if args.no_extra else extra_configs(args, hp, w, dev, rank, world, dist)

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
        "config": config_dict(S, args.in_flight, B, T),
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
    if extras is not None:
        line["configs"] = extras
    if world == 1 and not args.no_cpu_baseline:
        cores = os.cpu_count() or 1
        Sc = args.cpu_steps
        dt, fr = cpu_reference_pass(hp, w, ids, lens, gw, Sc, cores)
        line["cpu_baseline"] = {"value": fr / dt, "unit": UNIT, "cores": cores, "kind": "port",
                                "sample": "one pass of the same B=64,T=111 batch, %d of %d decoder steps, torch-CPU fp32 "
                                          "restatement of the reference (TF 1.9 unavailable), %.1f s" % (Sc, S, dt)}
    _emit(line)
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
    ap.add_argument("--no-extra", action="store_true", help="skip the `configs` record (C2 B=1, C3, C5, fp32 mode)")
    ap.add_argument("--c3-utterances", type=int, default=4096)
    ap.add_argument("--c3-static", action="store_true", help="C3 at N > 1: static LPT plan instead of dynamic claiming")
    ap.add_argument("--workload", default="c2", choices=["c2", "c4"],
                    help="c2: the headline (free-running synthesis, B=64/GPU); c4: the data-parallel training step (scripts/bench_train.py)")
    ap.add_argument("--in-flight", type=int, default=2, help="independent batches in flight per GPU (engine.EnginePool)")
    ap.add_argument("--precision", default="bf16", choices=["bf16", "fp32"],
                    help="bf16: tensor-core (tcgen05) contractions with bf16 operands, fp32 accumulate; fp32: SIMT everywhere")
    args = ap.parse_args()
    global OUT
    OUT = _claim_stdout()
    if args.warmup < 3 and args.impl == "b200":
        args.warmup = 3
    if args.workload == "c4":
        sys.argv = [os.path.join(ROOT, "scripts", "bench_train.py"), "--steps", str(args.steps), "--warmup", str(args.warmup)] + (
            ["--impl", "reference"] if args.impl == "reference" else [])
        sys.path.insert(0, os.path.join(ROOT, "scripts"))
        import bench_train
        bench_train.OUT = OUT
        return bench_train.main()
    if args.impl == "reference":
        return run_reference(args)
    return run_b200(args)


if __name__ == "__main__":
    sys.exit(main())
