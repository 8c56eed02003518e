"""The committed golden vectors (tests/golden/, written by scripts/make_golden.py from the CPU oracle) are reproduced by
the oracle on this machine: guards the oracle against drift and pins the fixtures the GPU tests use."""
import os

import numpy as np

from oracle import griffin_lim as gl
from oracle.tacotron_oracle import TacotronOracle
from tacotron_prosody_multi_gpu_b200 import weights
from tacotron_prosody_multi_gpu_b200.hparams import default_hparams

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def test_synthesis_golden_reproduced_by_oracle():
    g = np.load(os.path.join(GOLDEN, "synthesis_small.npz"))
    hp = default_hparams()
    orc = TacotronOracle(hp, weights.set_stop_bias(weights.init_weights(hp, perturb=0.05), -10.0))
    seed = int(g["philox_seed"])
    tf = orc.forward(g["ids"], g["lens"], mel_targets=g["mel_targets"], gst_weights=g["gst_weights"], philox_seed=seed)
    np.testing.assert_allclose(tf["encoder_outputs"].numpy(), g["encoder_outputs"], atol=1e-5)
    np.testing.assert_allclose(tf["decoder_output"].numpy(), g["tf_decoder_output"], atol=2e-5)
    np.testing.assert_allclose(tf["mel_outputs"].numpy(), g["tf_mel_outputs"], atol=5e-5)
    np.testing.assert_allclose(tf["alignments"].numpy(), g["tf_alignments"], atol=1e-6)
    np.testing.assert_allclose(tf["stop"].numpy(), g["tf_stop"], atol=1e-6)
    fr = orc.forward(g["ids"], g["lens"], gst_weights=g["gst_weights"], max_iters=5, philox_seed=seed)
    assert fr["steps"] == int(g["fr_steps"])
    np.testing.assert_allclose(fr["mel_outputs"].numpy(), g["fr_mel_outputs"], atol=5e-5)
    # shapes of the reference's tower outputs (Tacotron.py:314-317): [B, F, 80], [B, F], [B, T, S]
    assert g["tf_mel_outputs"].shape == (3, 12, 80) and g["tf_stop"].shape == (3, 12) and g["tf_alignments"].shape == (3, 9, 6)


def test_griffin_lim_golden_reproduced_by_oracle():
    g = np.load(os.path.join(GOLDEN, "griffin_lim_small.npz"))
    hp = default_hparams()
    np.testing.assert_allclose(gl.mel_to_magnitude(g["mel"], hp).T, g["magnitude"], rtol=1e-5, atol=1e-7)
    wav = gl.mel_to_audio(g["mel"], hp, seed=int(g["seed"]), iters=int(g["iters"]))
    assert wav.shape == (275 * 8,)
    np.testing.assert_allclose(wav, g["waveform"], atol=1e-5 * float(np.abs(g["waveform"]).max()))
