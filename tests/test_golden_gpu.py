"""CUDA path (through the C ABI) against the committed golden vectors of tests/golden/ (oracle outputs, see
scripts/make_golden.py): nothing here needs /root/reference or recomputes the oracle."""
import os

import numpy as np
import pytest
import torch

from tacotron_prosody_multi_gpu_b200 import weights
from tacotron_prosody_multi_gpu_b200.hparams import default_hparams

pytestmark = pytest.mark.gpu
GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def test_synthesis_matches_golden_fp32():
    from tacotron_prosody_multi_gpu_b200.engine import TacoEngine
    g = np.load(os.path.join(GOLDEN, "synthesis_small.npz"))
    hp = default_hparams()
    eng = TacoEngine(hp, weights.set_stop_bias(weights.init_weights(hp, perturb=0.05), -10.0))
    try:
        seed = int(g["philox_seed"])
        ids, lens, gw = (torch.from_numpy(g[k]) for k in ("ids", "lens", "gst_weights"))
        tf = eng.forward(ids, lens, mel_targets=torch.from_numpy(g["mel_targets"]), gst_weights=gw, philox_seed=seed)
        np.testing.assert_allclose(tf["encoder_outputs"].cpu().numpy(), g["encoder_outputs"], atol=1e-4)
        np.testing.assert_allclose(tf["decoder_output"].cpu().numpy(), g["tf_decoder_output"], atol=2e-4)   # fp32 kernels
        np.testing.assert_allclose(tf["mel_outputs"].cpu().numpy(), g["tf_mel_outputs"], atol=5e-4)
        np.testing.assert_allclose(tf["alignments"].cpu().numpy(), g["tf_alignments"], atol=2e-5)
        np.testing.assert_allclose(tf["stop"].cpu().numpy(), g["tf_stop"], atol=2e-5)
        fr = eng.forward(ids, lens, gst_weights=gw, max_iters=5, philox_seed=seed)
        assert fr["steps"] == int(g["fr_steps"])
        np.testing.assert_allclose(fr["mel_outputs"].cpu().numpy(), g["fr_mel_outputs"], atol=5e-4)
    finally:
        eng.close()


def test_griffin_lim_matches_golden():
    from tacotron_prosody_multi_gpu_b200.Utils.AudioProcessing import AudioPreprocess as AP
    g = np.load(os.path.join(GOLDEN, "griffin_lim_small.npz"))
    hp = default_hparams()
    plan = AP.GriffinLim(hp, 0)
    try:
        mag = plan.mel_to_mag(torch.from_numpy(g["mel"][None]).cuda())
        assert float(np.abs(mag.cpu().numpy()[0] - g["magnitude"]).max() / g["magnitude"].max()) < 2e-5
        wav = plan.mel_to_audio([g["mel"]], seed=int(g["seed"]), iters=int(g["iters"]))[0]
        assert float(np.abs(wav - g["waveform"]).max() / np.abs(g["waveform"]).max()) < 1e-4
    finally:
        plan.close()
