"""CPU oracle for the mel -> waveform step.  TEST INFRASTRUCTURE ONLY: imported
by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg, never by the product package.

Restates, in float64 numpy:
  * ``mel_to_audio_serie``  (Utils/AudioProcessing/AudioPreprocess.py:152-186): de-normalise (:469-484), dB -> amplitude,
    pinv(mel basis) (:163-165), ``** power``, random phases, ``griffin_lim_iters`` x (stft -> unit phases -> istft),
    ``inv_preemphasis`` (:45-48);
  * the third-party functions that path calls.  ``librosa`` is an unpinned requirement of the reference
    (requirements.txt) and is NOT installed here, so its published algorithms are restated:
      - ``librosa.filters.mel(sr, n_fft, n_mels, fmin, fmax)`` (Slaney scale, area-normalised triangles),
      - ``librosa.stft(y, n_fft, hop_length, win_length)`` (center=True, reflect padding, periodic Hann of
        ``win_length`` zero-padded to ``n_fft``),
      - ``librosa.istft(S, hop_length, win_length)`` (windowed overlap-add, division by the window sum of squares
        where it exceeds float32 tiny, n_fft/2 trimmed from both ends).
    PARITY UNPINNED: neither librosa nor a reference-produced waveform fixture exists in this container; the FFT
    itself is numpy's.

The reference draws the initial phases from the unseeded global ``np.random.rand`` (:175); here they come from
Philox4x32-10 so that the CUDA path can be checked:  element idx = k * F + t (k = frequency bin, t = frame, the
C order of ``np.random.rand(*D.shape)`` with D [n_bins, F]) uses word ``idx & 3`` of
Philox(counter = (idx >> 2, utterance, idx >> 34, "GLIM"), key = seed);  u = (word >> 8) * 2**-24;  phase = 2*pi*u.
"""
import numpy as np

from .philox import philox4x32_10

GLIM_TAG = 0x474C494D
_MASK = np.uint64(0xFFFFFFFF)


def hz_to_mel(f):
    f = np.asarray(f, np.float64)
    f_sp = 200.0 / 3
    mels = f / f_sp
    min_log_hz = 1000.0
    min_log_mel = min_log_hz / f_sp
    logstep = np.log(6.4) / 27.0
    return np.where(f >= min_log_hz, min_log_mel + np.log(np.maximum(f, 1e-30) / min_log_hz) / logstep, mels)


def mel_to_hz(m):
    m = np.asarray(m, np.float64)
    f_sp = 200.0 / 3
    min_log_hz = 1000.0
    min_log_mel = min_log_hz / f_sp
    logstep = np.log(6.4) / 27.0
    return np.where(m >= min_log_mel, min_log_hz * np.exp(logstep * (m - min_log_mel)), f_sp * m)


def mel_filter(sr, n_fft, n_mels, fmin, fmax):
    """librosa.filters.mel (htk=False, norm=1/'slaney'): [n_mels, 1 + n_fft//2]."""
    n_bins = 1 + n_fft // 2
    fftfreqs = np.linspace(0.0, sr / 2.0, n_bins)
    mel_f = mel_to_hz(np.linspace(hz_to_mel(fmin), hz_to_mel(fmax), n_mels + 2))
    fdiff = np.diff(mel_f)
    ramps = mel_f[:, None] - fftfreqs[None, :]
    w = np.zeros((n_mels, n_bins))
    for i in range(n_mels):
        lower = -ramps[i] / fdiff[i]
        upper = ramps[i + 2] / fdiff[i + 1]
        w[i] = np.maximum(0.0, np.minimum(lower, upper))
    enorm = 2.0 / (mel_f[2:n_mels + 2] - mel_f[:n_mels])
    return w * enorm[:, None]


def denormalize(D, hp):
    """AudioPreprocess.py:469-484."""
    D = np.asarray(D, np.float64)
    m, lo = float(hp.max_abs_value), float(hp.min_level_db)
    if hp.allow_clipping_in_normalization:
        if hp.symmetric_mels:
            return (np.clip(D, -m, m) + m) * -lo / (2 * m) + lo
        return np.clip(D, 0, m) * -lo / m + lo
    if hp.symmetric_mels:
        return (D + m) * -lo / (2 * m) + lo
    return D * -lo / m + lo


def mel_to_magnitude(mel, hp, inv_basis=None):
    """mel [F, num_mels] -> D ** power, [n_bins, F]   (AudioPreprocess.py:157-165,174)."""
    S = denormalize(np.asarray(mel).T, hp) if hp.signal_normalization else np.asarray(mel, np.float64).T
    amp = np.power(10.0, (S + hp.ref_level_db) * 0.05)
    if inv_basis is None:
        inv_basis = np.linalg.pinv(mel_filter(hp.sample_rate, hp.n_fft, hp.num_mels, hp.fmin, hp.fmax))
    return np.maximum(1e-10, inv_basis @ amp) ** hp.power


def hann_padded(win, n_fft):
    """scipy get_window('hann', win, fftbins=True) centred in n_fft zeros (librosa.util.pad_center)."""
    w = 0.5 - 0.5 * np.cos(2.0 * np.pi * np.arange(win) / win)
    out = np.zeros(n_fft)
    lpad = (n_fft - win) // 2
    out[lpad:lpad + win] = w
    return out


def stft(y, n_fft, hop, win):
    w = hann_padded(win, n_fft)
    yp = np.pad(np.asarray(y, np.float64), n_fft // 2, mode="reflect")
    n_frames = 1 + (len(yp) - n_fft) // hop
    idx = np.arange(n_fft)[:, None] + hop * np.arange(n_frames)[None, :]
    return np.fft.rfft(w[:, None] * yp[idx], axis=0)          # [n_bins, n_frames]


def istft(S, hop, win):
    n_fft = 2 * (S.shape[0] - 1)
    n_frames = S.shape[1]
    w = hann_padded(win, n_fft)
    y = np.zeros(n_fft + hop * (n_frames - 1))
    wsum = np.zeros_like(y)
    frames = np.fft.irfft(S, n=n_fft, axis=0) * w[:, None]
    for i in range(n_frames):
        y[i * hop:i * hop + n_fft] += frames[:, i]
        wsum[i * hop:i * hop + n_fft] += w * w
    nz = wsum > np.finfo(np.float32).tiny
    y[nz] /= wsum[nz]
    return y[n_fft // 2:-(n_fft // 2)]


def phase_init(seed, utterance, n_bins, n_frames):
    """Philox phases in turns (u in [0,1)), [n_bins, n_frames]."""
    idx = np.arange(n_bins * n_frames, dtype=np.uint64)
    blk = idx >> np.uint64(2)
    o = philox4x32_10(blk & _MASK, np.uint64(utterance), blk >> np.uint64(32), np.uint64(GLIM_TAG),
                      seed & 0xFFFFFFFF, (seed >> 32) & 0xFFFFFFFF)
    lane = (idx & np.uint64(3)).astype(np.int64)
    word = np.where(lane == 0, o[0], np.where(lane == 1, o[1], np.where(lane == 2, o[2], o[3])))
    return ((word >> np.uint32(8)).astype(np.float64) * 2.0 ** -24).reshape(n_bins, n_frames)


def griffin_lim(mag, turns, iters, hop, win, trace=None):
    """mag, turns [n_bins, F] -> waveform [hop * (F - 1)]   (AudioPreprocess.py:175-185)."""
    n_fft = 2 * (mag.shape[0] - 1)
    y = istft(mag * np.exp(2j * np.pi * turns), hop, win)
    for _ in range(iters):
        X = stft(y, n_fft, hop, win)
        if trace is not None:
            trace.append(float(np.linalg.norm(np.abs(X) - mag) / np.linalg.norm(mag)))
        y = istft(mag * np.exp(1j * np.angle(X)), hop, win)
    return y


def inv_preemphasis(wav, k, on=True):
    """scipy.signal.lfilter([1], [1, -k], wav)   (AudioPreprocess.py:45-48)."""
    if not on:
        return wav
    out = np.empty_like(wav)
    acc = 0.0
    for n, x in enumerate(wav):
        acc = x + k * acc
        out[n] = acc
    return out


def mel_to_audio(mel, hp, seed, utterance=0, iters=None, turns=None):
    mag = mel_to_magnitude(mel, hp)
    if turns is None:
        turns = phase_init(seed, utterance, mag.shape[0], mag.shape[1])
    y = griffin_lim(mag, turns, hp.griffin_lim_iters if iters is None else iters, hp.hop_size, hp.win_size)
    return inv_preemphasis(y, hp.preemphasis, hp.preemphasize)
