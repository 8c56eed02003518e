"""Seeded synthetic inputs for the configs of BASELINE.json.

No dataset is reachable (no network), so every config is driven by synthetic
token ids / mel targets of the reference's shapes:

* token ids uniform in [2, n_symbols) + EOS id 1, pad id 0
  (``Utils/TextProcessing/HangulUtils.py:13,54``; ``Utils/Tacotron_feeder.py:14``)
* lengths cycle the G2P token lengths of the reference's ``sentences.txt``
* mel targets N(0,1)*1.5 clipped to [-4,4], 13 frames per token rounded up to r,
  padded with -4.1 (``Utils/Tacotron_feeder.py:17-18,102-103``)
"""
from __future__ import annotations

import numpy as np

from .hparams import num_symbols

# G2P token lengths (incl. EOS) of the 16 lines of the reference's sentences.txt
SENTENCE_LENGTHS = [42, 58, 60, 65, 31, 60, 65, 64, 41, 36, 58, 58, 42, 63, 18, 111]

PAD_ID = 0
EOS_ID = 1
TARGET_PAD = -4.1  # -(max_abs_value + .1), Tacotron_feeder.py:17-18
TOKEN_PAD = 1.0    # Tacotron_feeder.py:22


def _rng(seed):
    return np.random.Generator(np.random.PCG64(seed))


def token_batch(hp, lengths, seed=None, pad_to=None):
    """ids int32 [B, T] (pad 0), lengths int32 [B]; T = max(lengths) unless ``pad_to``."""
    seed = hp.tacotron_data_random_state if seed is None else seed
    rng = _rng(seed)
    nsym = num_symbols(hp)
    lengths = np.asarray(lengths, np.int32)
    T = int(lengths.max()) if pad_to is None else int(pad_to)
    ids = np.full((len(lengths), T), PAD_ID, np.int32)
    for b, n in enumerate(lengths):
        ids[b, : n - 1] = rng.integers(2, nsym, size=n - 1)
        ids[b, n - 1] = EOS_ID
    return ids, lengths


def cycle_lengths(n, lengths=SENTENCE_LENGTHS):
    return [lengths[i % len(lengths)] for i in range(n)]


def round_up(x, m):
    return x if x % m == 0 else x + m - x % m


def mel_targets(hp, token_lengths, seed=None, frames_per_token=13, max_frames=None):
    """mel fp32 [B, F, 80] padded with -4.1, stop targets fp32 [B, F], frame lengths int32 [B]."""
    seed = hp.tacotron_data_random_state + 1 if seed is None else seed
    rng = _rng(seed)
    r = hp.outputs_per_step
    flen = [int(n) * frames_per_token for n in token_lengths]
    if max_frames is not None:
        flen = [min(f, max_frames) for f in flen]
    F = round_up(max(flen), r)
    # token targets are padded to max_len + 1 (Tacotron_feeder.py:85-89); keep one width
    B = len(flen)
    mel = np.full((B, F, hp.num_mels), TARGET_PAD, np.float32)
    stop = np.full((B, F), TOKEN_PAD, np.float32)
    for b, f in enumerate(flen):
        mel[b, :f] = np.clip(rng.standard_normal((f, hp.num_mels)) * 1.5, -hp.max_abs_value, hp.max_abs_value)
        stop[b, : max(f - 1, 0)] = 0.0
    return mel, stop, np.asarray(flen, np.int32)


def reference_mel(hp, batch, frames, seed=None):
    seed = hp.tacotron_data_random_state + 2 if seed is None else seed
    rng = _rng(seed)
    return np.clip(rng.standard_normal((batch, frames, hp.num_mels)) * 1.5,
                   -hp.max_abs_value, hp.max_abs_value).astype(np.float32)


def gst_random_weights(hp, seed=None):
    """softmax(U[0,1)^{heads x tokens}) (Tacotron.py:217-218), seeded and passed explicitly."""
    seed = hp.tacotron_random_seed + 7 if seed is None else seed
    rng = _rng(seed)
    u = rng.random((hp.num_heads, hp.num_gst)).astype(np.float32)
    e = np.exp(u - u.max(axis=1, keepdims=True))
    return (e / e.sum(axis=1, keepdims=True)).astype(np.float32)


def utterance_set(hp, n_utt, seed=None):
    """C3: ``n_utt`` token-length draws = repeats of the sentence lengths, shuffled (seed 1324)."""
    seed = hp.tacotron_data_random_state if seed is None else seed
    rng = _rng(seed)
    lens = np.asarray(cycle_lengths(n_utt), np.int32)
    rng.shuffle(lens)
    return lens
