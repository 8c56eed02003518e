"""CPU checks of the mel -> waveform oracle (oracle/griffin_lim.py) and of the host side of the audio step.

librosa is not installed, so the oracle's stft / istft are pinned against torch.stft / torch.istft, whose
documentation states librosa-compatible framing (center / reflect padding, window zero-padded to n_fft, window
envelope normalisation)."""
import numpy as np
import pytest
import torch

from oracle import griffin_lim as gl
from tacotron_prosody_multi_gpu_b200 import _lib
from tacotron_prosody_multi_gpu_b200.Utils.AudioProcessing import AudioPreprocess as AP
from tacotron_prosody_multi_gpu_b200.Utils.Hyperparams import hparams as hp

N_FFT, HOP, WIN = 2048, 275, 1024


def test_oracle_stft_istft_match_torch():
    rng = np.random.default_rng(3)
    y = rng.normal(size=HOP * 17)
    w = torch.hann_window(WIN, periodic=True, dtype=torch.float64)
    X = gl.stft(y, N_FFT, HOP, WIN)
    Xt = torch.stft(torch.from_numpy(y), N_FFT, HOP, WIN, window=w, center=True, pad_mode="reflect", return_complex=True).numpy()
    assert X.shape == Xt.shape == (1025, 18)
    np.testing.assert_allclose(X, Xt, atol=1e-10)
    S = X * np.exp(1j * rng.uniform(0, 6.28, size=X.shape))       # an inconsistent spectrogram, as in the iteration
    yi = gl.istft(S, HOP, WIN)
    yt = torch.istft(torch.from_numpy(S), N_FFT, HOP, WIN, window=w, center=True).numpy()
    assert yi.shape == yt.shape == (HOP * 17,)
    np.testing.assert_allclose(yi, yt, atol=1e-10)
    np.testing.assert_allclose(gl.istft(X, HOP, WIN), y, atol=1e-12)   # consistent spectrogram: exact round trip


def test_mel_filter_properties_and_host_copy():
    fb = gl.mel_filter(hp.sample_rate, hp.n_fft, hp.num_mels, hp.fmin, hp.fmax)
    assert fb.shape == (80, 1025) and (fb >= 0).all() and (fb.sum(1) > 0).all()
    freqs = np.linspace(0, hp.sample_rate / 2, 1025)
    support = freqs[(fb > 0).any(0)]
    assert support.min() >= hp.fmin and support.max() <= hp.fmax
    # Slaney normalisation: every triangle has unit area in Hz
    np.testing.assert_allclose((fb * (freqs[1] - freqs[0])).sum(1), 1.0, atol=0.12)
    peaks = fb.argmax(1)
    assert (np.diff(peaks) > 0).all()
    np.testing.assert_allclose(AP._build_mel_basis(hp), fb, atol=1e-12)
    # mel scale: linear below 1 kHz, log above, continuous at the knee
    np.testing.assert_allclose(gl.hz_to_mel([0, 500, 1000]), [0, 7.5, 15.0])
    np.testing.assert_allclose(gl.mel_to_hz(gl.hz_to_mel([30.0, 999.0, 1001.0, 7600.0])), [30.0, 999.0, 1001.0, 7600.0])


def test_denormalize_matches_reference_formula():
    x = np.array([-5.0, -4.0, 0.0, 4.0, 6.0])
    # symmetric, clipped: [-4,4] -> [-100, 0] dB   (AudioPreprocess.py:471-475)
    np.testing.assert_allclose(gl.denormalize(x, hp), [-100, -100, -50, 0, 0])
    np.testing.assert_allclose(AP._denormalize(x, hp), [-100, -100, -50, 0, 0])


def test_griffin_lim_oracle_converges_and_is_seeded():
    rng = np.random.default_rng(0)
    mel = np.clip(rng.normal(size=(14, 80)) * 1.5, -4, 4)
    mag = gl.mel_to_magnitude(mel, hp)
    assert mag.shape == (1025, 14) and (mag >= 1e-10 ** hp.power).all()
    t0 = gl.phase_init(45454, 0, *mag.shape)
    assert t0.min() >= 0 and t0.max() < 1 and abs(t0.mean() - 0.5) < 0.02
    assert not np.array_equal(t0, gl.phase_init(45454, 1, *mag.shape))
    trace = []
    y = gl.griffin_lim(mag, t0, 8, HOP, WIN, trace)
    assert y.shape == (HOP * 13,)
    assert all(b <= a + 1e-12 for a, b in zip(trace, trace[1:]))      # the projection never increases the distance
    y2 = gl.mel_to_audio(mel, hp, 45454, iters=8)
    np.testing.assert_array_equal(y, y2)


def test_inv_preemphasis_oracle_matches_scipy():
    from scipy import signal
    x = np.random.default_rng(1).normal(size=1000)
    np.testing.assert_allclose(gl.inv_preemphasis(x, 0.97), signal.lfilter([1], [1, -0.97], x), atol=1e-12)
    np.testing.assert_allclose(AP.inv_preemphasis(x, 0.97), signal.lfilter([1], [1, -0.97], x), atol=0)
    assert AP.inv_preemphasis(x, 0.97, False) is x


def test_audio_symbols_exported_and_no_gpu_failure():
    import ctypes
    L = ctypes.CDLL(_lib.LIB_PATH)
    header = open("include/taco_b200_audio.h").read()
    for name in AP.AUDIO_SYMBOLS:
        assert hasattr(L, name), name
        assert name + "(" in header, name
    if not torch.cuda.is_available():
        with pytest.raises(_lib.TacoError):
            AP.GriffinLim(hp)
        with pytest.raises(_lib.TacoError):
            AP.mel_to_audio_serie(np.zeros((80, 10), np.float32), hp)


def test_plan_argument_errors():
    class H:
        def __init__(self, **kw):
            self.__dict__.update(kw)
    with pytest.raises(ValueError):
        AP.GriffinLim(H(use_lws=True, n_fft=2048))
    with pytest.raises(ValueError):
        AP.GriffinLim(H(use_lws=False, n_fft=1024))
    assert AP.get_hop_size(H(hop_size=None, frame_shift_ms=12.5, sample_rate=22050)) == 275
