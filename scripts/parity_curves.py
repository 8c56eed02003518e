#!/usr/bin/env python
"""Error-vs-step curves of the CUDA decoder against the CPU oracle at BASELINE.json's full sizes, plus the
latency of the configurations bench.py does not time in its headline (C2 B=1, C5).  Run on a B200:

    python scripts/parity_curves.py [out.json] [cases,comma,separated]

Writes one JSON document (default gpurun_out/r02_parity_curves.json); tests/test_gpu_full_size.py asserts the
tolerances that were derived from it.  The oracle (oracle/) is the checker here, never the thing measured.
"""
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle.tacotron_oracle import TacotronOracle  # noqa: E402
from tacotron_prosody_multi_gpu_b200 import synthetic, weights  # noqa: E402
from tacotron_prosody_multi_gpu_b200.engine import TacoEngine  # noqa: E402
from tacotron_prosody_multi_gpu_b200.hparams import default_hparams  # noqa: E402

MARKS = (1, 16, 100, 277, 360, 720, 1000)


def curve(gpu, ref, axis_steps, r=1):
    """max |gpu - ref| per decoder step; ``axis_steps``: which axis indexes the step (frames: axis 1 in units of r)."""
    d = np.abs(np.asarray(gpu, np.float64) - np.asarray(ref, np.float64))
    d = np.moveaxis(d, axis_steps, 0)
    if r > 1:
        d = d.reshape(d.shape[0] // r, -1)
    return d.reshape(d.shape[0], -1).max(axis=1)


def marks(c):
    out = {}
    for m in MARKS:
        if m <= len(c):
            out[str(m)] = float(c[:m].max())
    return out


def decode_case(name, hp, w, lengths, S, teacher, precision, memory_seed=64, seed=21):
    ids, lens = synthetic.token_batch(hp, lengths)
    B, T = ids.shape
    rng = np.random.default_rng(memory_seed)
    memory = (rng.standard_normal((B, T, 768)) * 0.5).astype(np.float32)
    mel = None
    if teacher:
        mel, _, _ = synthetic.mel_targets(hp, np.full(B, T, np.int32), frames_per_token=13, max_frames=2 * S)
        mel = mel[:, :2 * S]
    bf = precision == "bf16"
    eng = TacoEngine(hp, w, precision=precision)
    mem_d = torch.from_numpy(memory).cuda()
    tgt = None if mel is None else torch.from_numpy(mel)
    d = eng.decode(mem_d, torch.from_numpy(lens), mel_targets=tgt, max_iters=S, philox_seed=seed)
    t0 = time.perf_counter()
    d = eng.decode(mem_d, torch.from_numpy(lens), mel_targets=tgt, max_iters=S, philox_seed=seed)
    torch.cuda.synchronize()
    wall = time.perf_counter() - t0
    kern_ms = eng.stage_ms(1)
    frames = d["frames"].cpu().numpy()
    stop = d["stop"].cpu().numpy()
    al = eng.alignments(d["align_hist"], d["steps"]).cpu().numpy()
    res = {"case": name, "B": B, "T": T, "S": S, "teacher_forced": teacher, "precision": precision, "gpu_steps": d["steps"],
           "decoder_kernel_ms": kern_ms, "decoder_step_us": 1e3 * kern_ms / max(d["steps"], 1), "wall_ms": 1e3 * wall}
    with torch.no_grad():
        refs = {"fp32": TacotronOracle(hp, w)}
        if bf:
            refs["mirror"] = TacotronOracle(hp, w, conv_bf16=True, dec_bf16=True)
        for k, orc in refs.items():
            t0 = time.perf_counter()
            ref = orc.decode(torch.from_numpy(memory), lens, mel_targets=mel, max_iters=S, philox_seed=seed)
            res["oracle_%s_s" % k] = time.perf_counter() - t0
            n = min(ref["steps"], d["steps"])
            cf = curve(frames[:, :2 * n], ref["frames_raw"].numpy()[:, :2 * n], 1, 2)
            ca = curve(al[:, :, :n], ref["alignments"].numpy()[:, :, :n], 2)
            cs = curve(stop[:, :2 * n], ref["stop"].numpy()[:, :2 * n], 1, 2)
            res["vs_" + k] = {"steps": ref["steps"], "frames_max_by_step": marks(cf), "align_max_by_step": marks(ca),
                              "stop_max_by_step": marks(cs), "frames_curve": [float(x) for x in cf[::8]],
                              "align_curve": [float(x) for x in ca[::8]],
                              "frames_rms": float(np.sqrt(np.mean((frames[:, :2 * n] - ref["frames_raw"].numpy()[:, :2 * n]) ** 2)))}
    eng.close()
    print(json.dumps({k: v for k, v in res.items() if not isinstance(v, dict)}), flush=True)
    for k in ("vs_fp32", "vs_mirror"):
        if k in res:
            print("   ", k, "frames", res[k]["frames_max_by_step"], "align", res[k]["align_max_by_step"], "rms %.2e" % res[k]["frames_rms"], flush=True)
    return res


def c5_case(hp, w):
    """C5 end to end (reference encoder + GST, T=300, 1000 free-running steps): timing only (parity: tests)."""
    lengths = [300, 287, 300, 251, 199, 300, 123, 64, 300, 18, 277, 300, 150, 290, 300, 42]
    ids, lens = synthetic.token_batch(hp, lengths)
    rm = synthetic.reference_mel(hp, len(lengths), 2000)
    res = {"case": "c5_e2e"}
    for precision in ("bf16", "fp32"):
        eng = TacoEngine(hp, w, precision=precision)
        args = (torch.from_numpy(ids).cuda(), torch.from_numpy(lens).cuda())
        rmd = torch.from_numpy(rm).cuda()
        for _ in range(2):
            eng.forward(*args, reference_mel=rmd, max_iters=1000, philox_seed=5)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        n = 3
        for _ in range(n):
            out = eng.forward(*args, reference_mel=rmd, max_iters=1000, philox_seed=5)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / n
        res[precision] = {"ms_per_batch": ms, "frames_per_s": 16 * 2000 / (ms * 1e-3), "steps": out["steps"],
                          "decoder_kernel_ms": eng.stage_ms(1), "decoder_step_us": eng.stage_ms(1), "encode_ms": eng.stage_ms(0),
                          "postnet_ms": eng.stage_ms(2)}
        eng.close()
    print(json.dumps(res), flush=True)
    return res


def main():
    out_path = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "gpurun_out", "r02_parity_curves.json")
    only = set(sys.argv[2].split(",")) if len(sys.argv) > 2 else None
    hp = default_hparams()
    w = weights.set_stop_bias(weights.init_weights(hp, perturb=0.05), -10.0)
    L64 = synthetic.cycle_lengths(64)
    cases = [
        ("c2_b64_tf_bf16", L64, 720, True, "bf16"),
        ("c2_b64_free_bf16", L64, 720, False, "bf16"),
        ("c2_b64_tf_fp32", L64, 720, True, "fp32"),
        ("c2_b1_t42_free_bf16", [42], 277, False, "bf16"),
        ("c2_b1_t42_free_fp32", [42], 277, False, "fp32"),
        ("c2_b1_t111_free_bf16", [111], 720, False, "bf16"),
        ("c2_b1_t111_free_fp32", [111], 720, False, "fp32"),
        ("c2_b1_t111_tf_bf16", [111], 720, True, "bf16"),
        ("c5_b16_t300_tf_bf16", [300, 287, 300, 251, 199, 300, 123, 64, 300, 18, 277, 300, 150, 290, 300, 42], 1000, True, "bf16"),
    ]
    results = []
    for name, lengths, S, teacher, prec in cases:
        if only and name not in only:
            continue
        results.append(decode_case(name, hp, w, lengths, S, teacher, prec))
    if not only or "c5_e2e" in only:
        results.append(c5_case(hp, w))
    os.makedirs(os.path.dirname(out_path), exist_ok=True)
    with open(out_path, "w") as f:
        json.dump(results, f, indent=1)
    print("wrote", out_path)


if __name__ == "__main__":
    main()
