"""Batch packing either side of the path -- the wire format ``Tacotron.initialize`` accepts.

Host-only NumPy restatement of the helper contract in the reference's Utils/Tacotron_feeder.py
(:11-22 pad values, :74-112 helpers, :25-60 per-device packing): token ids are padded with 0, mel
targets with -(max_abs_value + .1) up to a multiple of ``outputs_per_step``, stop-token targets
with 1.; per-device batches are laid side by side along the TIME axis and described by
``split_infos`` rows [T_in, T_mel, T_stop, T_linear].  The reference's threaded dataset ``Feeder``
is out of scope.
"""
import numpy as np

from .Hyperparams import hparams

_pad = 0
_token_pad = 1.
_target_pad = -(hparams.max_abs_value + .1) if hparams.symmetric_mels else -0.1


def _round_up(x, multiple):
    return -(-x // multiple) * multiple


def _right_pad(a, length, value):
    a = np.asarray(a)
    out = np.full((length,) + a.shape[1:], value, dtype=a.dtype)
    out[:a.shape[0]] = a
    return out


def _pad_input(x, length):
    return _right_pad(x, length, _pad)


def _pad_target(t, length):
    return _right_pad(t, length, _target_pad)


def _pad_token_target(t, length):
    return _right_pad(t, length, _token_pad)


def _prepare_inputs(inputs):
    """-> ([N, T_max] ids, T_max)"""
    longest = max(len(x) for x in inputs)
    return np.stack([_pad_input(x, longest) for x in inputs]), longest


def _prepare_targets(targets, alignment):
    """-> ([N, F, M] padded to a multiple of ``alignment`` frames, F)"""
    frames = _round_up(max(len(t) for t in targets), alignment)
    return np.stack([_pad_target(t, frames) for t in targets]), frames


def _prepare_token_targets(targets, alignment):
    """One extra position so that every utterance sees at least one <stop>=1 (feeder :85-89)."""
    frames = _round_up(max(len(t) for t in targets) + 1, alignment)
    return np.stack([_pad_token_target(t, frames) for t in targets]), frames


def _get_output_lengths(stop_tokens):
    """Frames kept per utterance: index of the first stop token that rounds (half-to-even) to 1,
    else the whole row (feeder :92-95)."""
    lengths = []
    for row in stop_tokens:
        hit = np.flatnonzero(np.round(np.asarray(row)) == 1)
        lengths.append(int(hit[0]) if hit.size else len(row))
    return lengths


def _prepare_batch(batches, outputs_per_step, n_devices=None, shuffle=True):
    """Examples ``(ids, mel [F,M], stop [F], linear|None, n_frames)`` -> the 7-tuple the train feed
    uses: (inputs, input_lengths, mel_targets, token_targets, linear_targets, targets_lengths,
    split_infos), each device's padded block concatenated along axis 1."""
    G = hparams.tacotron_num_gpus if n_devices is None else n_devices
    if len(batches) % G != 0:
        raise AssertionError("batch of %d examples is not divisible by %d devices" % (len(batches), G))
    if shuffle:
        np.random.shuffle(batches)
    per_dev = len(batches) // G
    blocks = {"in": [], "mel": [], "tok": [], "lin": []}
    split_infos = np.zeros((G, 4), np.int32)
    for g in range(G):
        shard = batches[g * per_dev:(g + 1) * per_dev]
        x, split_infos[g, 0] = _prepare_inputs([e[0] for e in shard])
        m, split_infos[g, 1] = _prepare_targets([e[1] for e in shard], outputs_per_step)
        s, split_infos[g, 2] = _prepare_token_targets([e[2] for e in shard], outputs_per_step)
        blocks["in"].append(x)
        blocks["mel"].append(m)
        blocks["tok"].append(s)
        if shard[0][3] is not None:
            lin, split_infos[g, 3] = _prepare_targets([e[3] for e in shard], outputs_per_step)
            blocks["lin"].append(lin)
    cat = lambda k: np.concatenate(blocks[k], axis=1) if blocks[k] else None
    input_lengths = np.asarray([len(e[0]) for e in batches], np.int32)
    targets_lengths = np.asarray([e[-1] for e in batches], np.int32)
    return cat("in"), input_lengths, cat("mel"), cat("tok"), cat("lin"), targets_lengths, split_infos
