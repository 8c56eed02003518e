"""Bench of the mel -> waveform step on one B200: C2-sized batch (B utterances x F frames,
60 Griffin-Lim iterations).  Prints one JSON line: frames/s device-resident (`value`), through the host-buffer C ABI
(`e2e`), the iteration kernel's HBM roofline, and the float64 oracle timed on a bounded sample (`cpu_baseline`)."""
import argparse
import json
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from tacotron_prosody_multi_gpu_b200.Utils.AudioProcessing import AudioPreprocess as AP  # noqa: E402
from tacotron_prosody_multi_gpu_b200.Utils.Hyperparams import hparams as hp  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=64)
    ap.add_argument("--frames", type=int, default=1440)
    ap.add_argument("--iters", type=int, default=60)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--cpu-frames", type=int, default=360, help="oracle sample: one utterance of this many frames (0 = skip)")
    a = ap.parse_args()
    B, F, it = a.batch, a.frames, a.iters
    plan = AP.GriffinLim(hp, 0)
    rng = np.random.default_rng(1324)
    mel_h = np.clip(rng.normal(size=(B, F, 80)) * 1.5, -4, 4).astype(np.float32)
    mel = torch.from_numpy(mel_h).cuda()
    audio = torch.empty((B, plan.hop * (F - 1)), dtype=torch.float32, device="cuda")

    def step():
        mag = plan.mel_to_mag(mel)
        plan.griffin_lim(mag, seed=45454, iters=it, out=audio)

    for _ in range(a.warmup):
        step()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(a.steps):
        step()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / a.steps
    # the iteration kernel alone
    mag = plan.mel_to_mag(mel)
    plan.griffin_lim(mag, seed=45454, iters=0, out=audio)
    torch.cuda.synchronize()
    e0.record()
    plan.griffin_lim(mag, seed=45454, iters=0, out=audio)
    e1.record()
    torch.cuda.synchronize()
    ms0 = e0.elapsed_time(e1)
    ms_iter = (ms - ms0) / it if it else 0.0
    # end to end through host buffers
    mels = [mel_h[b] for b in range(B)]
    plan.mel_to_audio(mels, seed=45454, iters=it)
    t = time.perf_counter()
    for _ in range(max(1, a.steps // 2)):
        wavs = plan.mel_to_audio(mels, seed=45454, iters=it)
    e2e_s = (time.perf_counter() - t) / max(1, a.steps // 2)
    bytes_iter = B * F * (1025 + 2 * plan.win) * 4          # magnitudes + frames read once + frames written once
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")))
    except Exception:
        pass
    peak = float(peaks.get("hbm_gbps_sustained", peaks.get("hbm_gbps", 6551.7))) if isinstance(peaks, dict) else 6551.7
    out = {
        "metric": "griffin_lim_mel_frames_per_sec", "value": B * F / (ms * 1e-3), "unit": "frames/s", "n_gpus": 1,
        "steps": a.steps, "warmup": a.warmup, "ms_per_step": ms, "higher_is_better": True, "dtype": "f32", "data": "synthetic",
        "config": {"workload": "mel -> waveform, B=%d utterances x F=%d frames, %d Griffin-Lim iterations, n_fft 2048 / hop 275 / win 1024"
                               % (B, F, it), "l2": "working set %d MB > L2" % (B * F * 3 * 1024 * 4 >> 20)},
        "audio_seconds_per_sec": B * plan.hop * (F - 1) / hp.sample_rate / (ms * 1e-3),
        "ms_per_iteration": ms_iter,
        "e2e": {"value": B * F / e2e_s, "unit": "frames/s", "h2d_bytes_per_step": int(mel_h.nbytes + 4 * B),
                "d2h_bytes_per_step": int(sum(w.nbytes for w in wavs))},
        "roofline": {"bound": "hbm", "achieved": bytes_iter / (ms_iter * 1e-3) / 1e9 if ms_iter else None, "peak": peak, "unit": "GB/s",
                     "frac": (bytes_iter / (ms_iter * 1e-3) / 1e9 / peak) if ms_iter else None, "traffic": None,
                     "note": "gl_kernel<ITER>: algorithmic bytes = B*F*(1025 + 2*win)*4 per launch; the kernel is bound by shared-memory "
                             "FFT passes, not HBM"},
        "gpu_launches": a.steps * (it + 3),
    }
    if a.cpu_frames:
        from oracle import griffin_lim as gl
        m = mel_h[0, :a.cpu_frames]
        t = time.perf_counter()
        gl.mel_to_audio(m, hp, 45454, iters=it)
        dt = time.perf_counter() - t
        out["cpu_baseline"] = {"value": a.cpu_frames / dt, "unit": "frames/s", "cores": 1, "kind": "port",
                               "sample": "one utterance of %d frames, %d iterations, float64 numpy oracle" % (a.cpu_frames, it)}
    print(json.dumps(out))


if __name__ == "__main__":
    main()
