"""GPU parity of the mel -> waveform step (include/taco_b200_audio.h) against oracle/griffin_lim.py (float64).

Tolerances: the device path is fp32 (FFT in shared memory, fp32 twiddles); the oracle is float64.  One projection
step agrees to ~1e-5 of the signal's peak; the iteration is a non-expansive projection, so the error does not blow up,
but 60 iterations compound fp32 rounding, hence the looser bound written in that test."""
import numpy as np
import pytest
import torch

from oracle import griffin_lim as gl
from tacotron_prosody_multi_gpu_b200.Utils.AudioProcessing import AudioPreprocess as AP
from tacotron_prosody_multi_gpu_b200.Utils.Hyperparams import hparams as hp

pytestmark = pytest.mark.gpu
HOP, WIN = 275, 1024


@pytest.fixture(scope="module")
def plan():
    p = AP.GriffinLim(hp, 0)
    yield p
    p.close()


def _mels(lens, seed=0):
    rng = np.random.default_rng(seed)
    return [np.clip(rng.normal(size=(n, 80)) * 1.5, -4.2, 4.2).astype(np.float32) for n in lens]


def _batch(mels):
    F = max(m.shape[0] for m in mels)
    out = np.zeros((len(mels), F, 80), np.float32)
    for b, m in enumerate(mels):
        out[b, :m.shape[0]] = m
    return torch.from_numpy(out).cuda(), torch.tensor([m.shape[0] for m in mels], dtype=torch.int32).cuda()


def _rel(a, b):
    return float(np.abs(a - b).max() / np.abs(b).max())


def test_mel_to_magnitude_parity(plan):
    mels = _mels([13, 9, 16])
    mel_t, nf = _batch(mels)
    mag = plan.mel_to_mag(mel_t, nf).cpu().numpy()
    for b, m in enumerate(mels):
        ref = gl.mel_to_magnitude(m, hp).T                       # [F, bins]
        assert _rel(mag[b, :m.shape[0]], ref) < 2e-5
        assert (mag[b, m.shape[0]:] == 0).all()


@pytest.mark.parametrize("iters", [0, 1, 3])
def test_griffin_lim_explicit_phases_parity(plan, iters):
    rng = np.random.default_rng(5)
    F = 11                                                       # odd: the last CTA owns a single frame
    mag = np.abs(rng.normal(size=(2, F, 1025))).astype(np.float32) + 0.01
    turns = rng.uniform(size=(2, F, 1025)).astype(np.float32)
    y = plan.griffin_lim(torch.from_numpy(mag).cuda(), turns=torch.from_numpy(turns).cuda(), iters=iters, preemphasis=0.0).cpu().numpy()
    assert y.shape == (2, HOP * (F - 1))
    for b in range(2):
        ref = gl.griffin_lim(mag[b].T.astype(np.float64), turns[b].T.astype(np.float64), iters, HOP, WIN)
        assert _rel(y[b], ref) < 5e-5, (b, iters)


def test_ragged_batch_philox_phases_and_preemphasis(plan):
    mels = _mels([12, 7, 5, 20], seed=2)
    mel_t, nf = _batch(mels)
    mag = plan.mel_to_mag(mel_t, nf)
    y = plan.griffin_lim(mag, nf, seed=45454, utterance0=3, iters=2, preemphasis=0.97).cpu().numpy()
    for b, m in enumerate(mels):
        M = gl.mel_to_magnitude(m, hp)
        ref = gl.griffin_lim(M, gl.phase_init(45454, 3 + b, *M.shape), 2, HOP, WIN)
        ref = gl.inv_preemphasis(ref, 0.97)
        L = HOP * (m.shape[0] - 1)
        assert _rel(y[b, :L], ref) < 1e-4, b
        assert (y[b, L:] == 0).all()


def test_sixty_iterations_against_oracle(plan):
    mels = _mels([24], seed=7)
    mel_t, nf = _batch(mels)
    mag = plan.mel_to_mag(mel_t, nf)
    y = plan.griffin_lim(mag, nf, seed=9, iters=60).cpu().numpy()[0]
    trace = []
    M = gl.mel_to_magnitude(mels[0], hp)
    ref = gl.griffin_lim(M, gl.phase_init(9, 0, *M.shape), 60, HOP, WIN, trace)
    assert _rel(y, ref) < 2e-2            # 60 compounded fp32 projections vs float64
    d = float(plan.spectral_distance(torch.from_numpy(y[None]).cuda(), mag, nf).cpu()[0])
    d_ref = float(np.linalg.norm(np.abs(gl.stft(ref, 2048, HOP, WIN)) - M) / np.linalg.norm(M))
    assert abs(d - d_ref) < 2e-3 * max(d_ref, 1e-3) + 1e-4


def test_spectral_distance_parity_and_monotone_convergence_full_size(plan):
    # C2-sized batch: 8 utterances of 1440 frames (S = 720 decoder steps x r = 2)
    rng = np.random.default_rng(11)
    mel = torch.from_numpy(np.clip(rng.normal(size=(8, 1440, 80)) * 1.5, -4, 4).astype(np.float32)).cuda()
    mag = plan.mel_to_mag(mel)
    dist = []
    for it in (0, 5, 20, 60):
        y = plan.griffin_lim(mag, seed=1, iters=it)
        assert torch.isfinite(y).all()
        dist.append(plan.spectral_distance(y, mag).cpu().numpy())
    for a, b in zip(dist, dist[1:]):
        assert (b < a).all()
    # determinism: same seed, same bits
    y2 = plan.griffin_lim(mag, seed=1, iters=60)
    assert torch.equal(y, y2)
    # the measure itself against the oracle on one utterance
    y0 = y[0].cpu().numpy().astype(np.float64)
    M = mag[0].cpu().numpy().astype(np.float64).T
    d_ref = np.linalg.norm(np.abs(gl.stft(y0, 2048, HOP, WIN)) - M) / np.linalg.norm(M)
    assert abs(dist[-1][0] - d_ref) < 1e-3 * d_ref + 1e-5


def test_mel_to_audio_serie_host_api(plan):
    mels = _mels([15, 10], seed=4)
    wav = AP.mel_to_audio_serie(mels[0].T, hp, seed=77)
    assert wav.dtype == np.float32 and wav.shape == (HOP * 14,)
    ref = gl.mel_to_audio(mels[0], hp, 77, iters=hp.griffin_lim_iters)
    assert _rel(wav, ref) < 2e-2
    both = AP.mel_to_audio_serie([m.T for m in mels], hp, seed=77)
    np.testing.assert_array_equal(both[0], wav)                  # batching does not change an utterance
    assert both[1].shape == (HOP * 9,)
    with pytest.raises(Exception):
        AP.mel_to_audio_serie(mels[0][:4].T, hp)                 # too short for reflect padding
