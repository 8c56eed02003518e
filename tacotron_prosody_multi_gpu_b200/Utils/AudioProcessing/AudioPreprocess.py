"""Mel -> waveform, the step after the synthesis path.

Mirrors the names of the reference's Utils/AudioProcessing/AudioPreprocess.py that the synthesis / training
scripts import (Tacotron_synthesizer.py:6, train.py:124-127): ``mel_to_audio_serie``, ``inv_preemphasis``,
``save_wav``, ``get_hop_size``, ``_denormalize``, ``_db_to_amp``, ``_build_mel_basis``.  The spectral work
(de-normalise -> pinv(mel basis) -> Griffin-Lim -> inverse pre-emphasis) runs in libtaco_b200.so
(include/taco_b200_audio.h); there is no CPU path: without the CUDA library the calls raise ``TacoError``.

Host-side numpy here only builds constants (mel filter bank and its pseudo-inverse, once per hparams) and
does the int16 scaling of ``save_wav``.  Differences from the reference, all deliberate:
  * the initial phases are Philox-seeded (``seed`` argument, default hparams.tacotron_random_seed) instead of
    the unseeded global numpy RNG (:175), so that a run is reproducible and checkable;
  * ``use_lws=True`` (:167-171, needs the third-party ``lws`` package) is rejected;
  * ``mel_to_audio_serie`` also accepts a list of mels and synthesises them as one ragged batch (each utterance
    with its own length, exactly as if called one by one).
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from ... import _lib

N_FFT_SUPPORTED = 2048


class TacoMelNorm(C.Structure):
    _fields_ = [("signal_normalization", C.c_int32), ("allow_clipping_in_normalization", C.c_int32),
                ("symmetric_mels", C.c_int32), ("max_abs_value", C.c_float), ("min_level_db", C.c_float),
                ("ref_level_db", C.c_float), ("power", C.c_float)]


AUDIO_SYMBOLS = ("taco_gl_create", "taco_gl_destroy", "taco_gl_last_error", "taco_gl_mel_to_mag", "taco_gl_griffin_lim",
                 "taco_gl_spectral_distance", "taco_mel_to_audio")

_bound = False


def _L():
    global _bound
    L = _lib.lib()
    if not _bound:
        vp, i32, i64, u64, f32 = C.c_void_p, C.c_int, C.c_longlong, C.c_uint64, C.c_float
        L.taco_gl_create.argtypes = [i32, i32, i32, i32, i32, vp, C.POINTER(vp)]
        L.taco_gl_destroy.argtypes = [vp]
        L.taco_gl_destroy.restype = None
        L.taco_gl_last_error.argtypes = [vp]
        L.taco_gl_last_error.restype = C.c_char_p
        L.taco_gl_mel_to_mag.argtypes = [vp, vp, vp, i32, i32, C.POINTER(TacoMelNorm), vp, vp]
        L.taco_gl_griffin_lim.argtypes = [vp, vp, vp, vp, u64, i32, i32, i32, i32, f32, vp, i64, vp]
        L.taco_gl_spectral_distance.argtypes = [vp, vp, i64, vp, vp, i32, i32, vp, vp]
        L.taco_mel_to_audio.argtypes = [vp, vp, vp, i32, i32, C.POINTER(TacoMelNorm), u64, i32, i32, f32, vp, i64]
        for name in AUDIO_SYMBOLS:
            if name not in ("taco_gl_destroy", "taco_gl_last_error"):
                getattr(L, name).restype = i32
        _bound = True
    return L


def get_hop_size(hparams):
    """AudioPreprocess.py:352-356."""
    hop_size = hparams.hop_size
    if hop_size is None:
        hop_size = int(hparams.frame_shift_ms / 1000 * hparams.sample_rate)
    return hop_size


def _hz_to_mel(f):
    f = np.asarray(f, np.float64)
    return np.where(f >= 1000.0, 15.0 + np.log(np.maximum(f, 1e-30) / 1000.0) * (27.0 / np.log(6.4)), f * (3.0 / 200.0))


def _mel_to_hz(m):
    m = np.asarray(m, np.float64)
    return np.where(m >= 15.0, 1000.0 * np.exp((m - 15.0) * (np.log(6.4) / 27.0)), m * (200.0 / 3.0))


def _build_mel_basis(hparams):
    """librosa.filters.mel(sample_rate, n_fft, n_mels, fmin, fmax) as called at :436-439 (Slaney scale, unit-area
    triangles); librosa itself is not a dependency of this package."""
    n_bins = 1 + hparams.n_fft // 2
    freqs = np.arange(n_bins, dtype=np.float64) * (hparams.sample_rate / 2.0 / (n_bins - 1))
    edges = _mel_to_hz(np.linspace(_hz_to_mel(hparams.fmin), _hz_to_mel(hparams.fmax), hparams.num_mels + 2))
    rising = (freqs[None, :] - edges[:-2, None]) / (edges[1:-1] - edges[:-2])[:, None]
    falling = (edges[2:, None] - freqs[None, :]) / (edges[2:] - edges[1:-1])[:, None]
    tri = np.clip(np.minimum(rising, falling), 0.0, None)
    return tri * (2.0 / (edges[2:] - edges[:-2]))[:, None]


def _db_to_amp(x):
    return np.power(10.0, np.asarray(x) * 0.05)


def _denormalize(D, hparams):
    """AudioPreprocess.py:469-484 (host restatement for callers that import it; the synthesis path does this on
    the device)."""
    m, lo = hparams.max_abs_value, hparams.min_level_db
    if hparams.allow_clipping_in_normalization:
        D = np.clip(D, -m, m) if hparams.symmetric_mels else np.clip(D, 0, m)
    if hparams.symmetric_mels:
        return (D + m) * -lo / (2 * m) + lo
    return D * -lo / m + lo


def _mel_norm(hparams):
    return TacoMelNorm(int(bool(hparams.signal_normalization)), int(bool(hparams.allow_clipping_in_normalization)),
                       int(bool(hparams.symmetric_mels)), float(hparams.max_abs_value), float(hparams.min_level_db),
                       float(hparams.ref_level_db), float(hparams.power))


class GriffinLim:
    """Plan object over the C ABI (window, twiddles, pinv(mel basis), growable frame workspace) for one device."""

    def __init__(self, hparams, device=0):
        if getattr(hparams, "use_lws", False):
            raise ValueError("use_lws=True needs the third-party lws package; only the Griffin-Lim branch is built")
        if hparams.n_fft != N_FFT_SUPPORTED:
            raise ValueError("n_fft must be %d (the in-shared-memory FFT is built for hparams.n_fft)" % N_FFT_SUPPORTED)
        self.hp = hparams
        self.hop = get_hop_size(hparams)
        self.win = hparams.win_size or hparams.n_fft
        self.n_bins = hparams.n_fft // 2 + 1
        self.n_mels = hparams.num_mels
        self.device = device
        self.inv_basis = np.ascontiguousarray(np.linalg.pinv(_build_mel_basis(hparams)), np.float32)   # [n_bins, n_mels]
        self.norm = _mel_norm(hparams)
        self._h = C.c_void_p()
        L = _L()
        rc = L.taco_gl_create(device, hparams.n_fft, self.hop, self.win, self.n_mels, self.inv_basis.ctypes.data, C.byref(self._h))
        if rc != 0:
            self._h = C.c_void_p()
            raise _lib.TacoError("taco_gl_create failed (%d): is a CUDA device present?" % rc)

    def close(self):
        if self._h:
            _L().taco_gl_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _check(self, rc, what):
        if rc != 0:
            raise _lib.TacoError("%s failed (%d): %s" % (what, rc, _L().taco_gl_last_error(self._h).decode()))

    def min_frames(self):
        """Shortest utterance librosa's reflect padding (and this path) accepts."""
        return (self.hp.n_fft // 2) // self.hop + 2

    # ---- device-pointer forms (torch CUDA tensors: plumbing only) ----
    def mel_to_mag(self, mel, n_frames=None, stream=None):
        import torch
        B, F, _ = mel.shape
        mag = torch.empty((B, F, self.n_bins), dtype=torch.float32, device=mel.device)
        self._check(_L().taco_gl_mel_to_mag(self._h, mel.data_ptr(), n_frames.data_ptr() if n_frames is not None else None, B, F,
                                            C.byref(self.norm), mag.data_ptr(), stream), "taco_gl_mel_to_mag")
        return mag

    def griffin_lim(self, mag, n_frames=None, turns=None, seed=0, utterance0=0, iters=None, preemphasis=None, out=None, stream=None):
        import torch
        B, F, _ = mag.shape
        stride = self.hop * (F - 1)
        audio = out if out is not None else torch.empty((B, stride), dtype=torch.float32, device=mag.device)
        k = (self.hp.preemphasis if self.hp.preemphasize else 0.0) if preemphasis is None else preemphasis
        self._check(_L().taco_gl_griffin_lim(self._h, mag.data_ptr(), turns.data_ptr() if turns is not None else None,
                                             n_frames.data_ptr() if n_frames is not None else None, int(seed), utterance0, B, F,
                                             self.hp.griffin_lim_iters if iters is None else iters, float(k), audio.data_ptr(),
                                             audio.shape[1], stream), "taco_gl_griffin_lim")
        return audio

    def spectral_distance(self, audio, mag, n_frames=None, stream=None):
        import torch
        B, F, _ = mag.shape
        out = torch.empty((B,), dtype=torch.float32, device=mag.device)
        self._check(_L().taco_gl_spectral_distance(self._h, audio.data_ptr(), audio.shape[1], mag.data_ptr(),
                                                   n_frames.data_ptr() if n_frames is not None else None, B, F, out.data_ptr(),
                                                   stream), "taco_gl_spectral_distance")
        return out

    # ---- host-buffer form ----
    def mel_to_audio(self, mels, seed=0, utterance0=0, iters=None):
        """mels: list of [len_b, num_mels] float arrays -> list of float32 waveforms [hop * (len_b - 1)]."""
        B = len(mels)
        lens = np.asarray([m.shape[0] for m in mels], np.int32)
        F = int(lens.max())
        if B > 1 and all(n == F for n in lens):
            batch = np.ascontiguousarray(np.stack(mels), np.float32)
        else:
            batch = np.zeros((B, F, self.n_mels), np.float32)
            for b, m in enumerate(mels):
                batch[b, :lens[b]] = m
        stride = self.hop * (F - 1)
        audio = np.empty((B, stride), np.float32)
        k = self.hp.preemphasis if self.hp.preemphasize else 0.0
        self._check(_L().taco_mel_to_audio(self._h, batch.ctypes.data, lens.ctypes.data, B, F, C.byref(self.norm), int(seed),
                                           utterance0, self.hp.griffin_lim_iters if iters is None else iters, float(k),
                                           audio.ctypes.data, stride), "taco_mel_to_audio")
        return [audio[b, :self.hop * (lens[b] - 1)] for b in range(B)]   # views of this call's own buffer


_plans = {}


def _plan(hparams, device=0):
    key = (id(hparams), device)
    if key not in _plans:
        _plans[key] = GriffinLim(hparams, device)
    return _plans[key]


def mel_to_audio_serie(mel, hparams, seed=None, device=0):
    """Reference signature (AudioPreprocess.py:152): ``mel`` is [num_mels, frames] (callers pass ``mels.T``) and the
    result a float32 waveform.  A list of such arrays gives a list of waveforms from one batched launch sequence."""
    seed = hparams.tacotron_random_seed if seed is None else seed
    single = not isinstance(mel, (list, tuple))
    mels = [np.ascontiguousarray(np.asarray(m, np.float32).T) for m in ([mel] if single else mel)]
    out = _plan(hparams, device).mel_to_audio(mels, seed=seed)
    return out[0] if single else out


def inv_preemphasis(wav, k, preemphasize=True):
    """AudioPreprocess.py:45-48.  mel_to_audio_serie already applies it on the device when hparams.preemphasize; this
    host form exists because Tacotron_synthesizer.py and train.py import it."""
    if preemphasize:
        from scipy import signal
        return signal.lfilter([1], [1, -k], wav)
    return wav


def save_wav(wav, path, sr):
    """AudioPreprocess.py:321-325 (scales in place like the reference)."""
    from scipy.io import wavfile
    wav *= 32767 / max(0.01, np.max(np.abs(wav)))
    wavfile.write(path, sr, wav.astype(np.int16))
