"""Writes tests/golden/*.npz: small seeded input/output vectors of the hot path, produced by the CPU oracle.

The reference itself cannot produce them (TensorFlow 1.9 / librosa are not installable here), so these
are ORACLE outputs: they pin the oracle against drift (tests/test_golden_cpu.py recomputes them on CPU) and give the CUDA
path fixed vectors that travel to the GPU box (tests/test_golden_gpu.py).      python scripts/make_golden.py
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import griffin_lim as gl  # noqa: E402
from oracle.tacotron_oracle import TacotronOracle  # noqa: E402
from tacotron_prosody_multi_gpu_b200 import synthetic, weights  # noqa: E402
from tacotron_prosody_multi_gpu_b200.hparams import default_hparams  # noqa: E402

OUT = os.path.join(ROOT, "tests", "golden")
WEIGHT_SEED_NOTE = "weights = set_stop_bias(init_weights(default_hparams(), perturb=0.05), -10.0)  (seed 45454)"


def synthesis_case():
    """C1-like: teacher-forced (GTA) and free-running passes of a tiny ragged batch, fp32."""
    hp = default_hparams()
    w = weights.set_stop_bias(weights.init_weights(hp, perturb=0.05), -10.0)
    orc = TacotronOracle(hp, w)
    ids, lens = synthetic.token_batch(hp, [9, 5, 7], seed=4242)
    gw = synthetic.gst_random_weights(hp)
    tgt = synthetic.mel_targets(hp, [9, 5, 7], seed=4243, frames_per_token=2, max_frames=12)[0][:, :12].copy()
    tf = orc.forward(ids, lens, mel_targets=tgt, gst_weights=gw, philox_seed=77)
    fr = orc.forward(ids, lens, gst_weights=gw, max_iters=5, philox_seed=77)
    return dict(note=WEIGHT_SEED_NOTE, ids=ids, lens=lens, gst_weights=gw, mel_targets=tgt, philox_seed=77,
                tf_decoder_output=tf["decoder_output"].numpy(), tf_mel_outputs=tf["mel_outputs"].numpy(),
                tf_stop=tf["stop"].numpy(), tf_alignments=tf["alignments"].numpy(),
                encoder_outputs=tf["encoder_outputs"].numpy(),
                fr_mel_outputs=fr["mel_outputs"].numpy(), fr_stop=fr["stop"].numpy(), fr_steps=fr["steps"])


def griffin_lim_case():
    hp = default_hparams()
    rng = np.random.default_rng(99)
    mel = np.clip(rng.normal(size=(9, hp.num_mels)) * 1.5, -4, 4).astype(np.float32)
    mag = gl.mel_to_magnitude(mel, hp)
    wav = gl.mel_to_audio(mel, hp, seed=1234, iters=3)
    return dict(mel=mel, seed=1234, iters=3, magnitude=mag.T.astype(np.float32), waveform=wav.astype(np.float32))


def main():
    os.makedirs(OUT, exist_ok=True)
    np.savez_compressed(os.path.join(OUT, "synthesis_small.npz"), **synthesis_case())
    np.savez_compressed(os.path.join(OUT, "griffin_lim_small.npz"), **griffin_lim_case())
    for f in sorted(os.listdir(OUT)):
        print(f, os.path.getsize(os.path.join(OUT, f)))


if __name__ == "__main__":
    main()
