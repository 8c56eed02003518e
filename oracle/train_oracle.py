"""TEST INFRASTRUCTURE (oracle) -- not part of the product path.

Training-mode restatement of the reference: ``Tacotron.initialize(is_training=True)`` +
``add_loss`` + ``add_optimizer`` (TacotronModel/modules/Tacotron.py:46-349, 351-446, 448-540) on top of
``TacotronOracle`` (torch-CPU).  Gradients come from torch autograd through the restated forward
pass; the optimizer is restated in NumPy following SURVEY Appendix A.13-14.  Parity is UNPINNED for
the same reason as tacotron_oracle.py (TensorFlow 1.9 cannot run here, the reference ships no
fixtures).

Randomness: the reference draws dropout / zoneout bits from TF's stateful RNG; both sides of the
parity tests use the keyed Philox streams of oracle/philox.py instead (stream numbers below match
tacotron_prosody_multi_gpu_b200/train/step.py).
"""
from __future__ import annotations

from collections import OrderedDict

import numpy as np
import torch

from . import philox
from .tacotron_oracle import TacotronOracle

STREAM_ENC_CONV, STREAM_POSTNET, STREAM_ENC_LSTM, STREAM_DEC_LSTM = 100, 200, 300, 400
NOT_TRAINABLE = ("moving_mean", "moving_variance")
REG_EXCLUDE = ("bias", "Bias", "_projection", "inputs_embedding", "RNN", "LSTM")     # Tacotron.py:419-422


def _mask(seed, stream, shape, p, offset=0):
    n = int(np.prod(shape))
    return torch.from_numpy(philox.keep_mask(seed, stream, n, p, offset).reshape(shape).astype(np.float32))


def train_masks(hp, seed, B, T, F):
    """All Bernoulli keep masks of one training step, from the keyed Philox streams."""
    keep = 1.0 - hp.dropout_rate
    zk = 1.0 - hp.zoneout_rate
    H, D = hp.enc_lstm_hidden_size, hp.decoder_lstm_units
    S = F // hp.outputs_per_step
    PC = hp.postnet_channels + hp.style_embed_depth
    m = dict()
    m["enc_conv"] = [_mask(seed, STREAM_ENC_CONV + i, (B, T, hp.conv_channels), keep) for i in range(hp.conv_num_layers)]
    m["postnet"] = [_mask(seed, STREAM_POSTNET + i, (B, F, PC), keep) for i in range(hp.postnet_num_layers)]
    m["enc_lstm"] = tuple(([_mask(seed, STREAM_ENC_LSTM + 2 * d, (B, H), zk, s * B * H) for s in range(T)],
                           [_mask(seed, STREAM_ENC_LSTM + 2 * d + 1, (B, H), zk, s * B * H) for s in range(T)]) for d in range(2))
    m["dec_lstm"] = [tuple((_mask(seed, STREAM_DEC_LSTM + 2 * l, (B, D), zk, s * B * D),
                            _mask(seed, STREAM_DEC_LSTM + 2 * l + 1, (B, D), zk, s * B * D)) for l in range(2)) for s in range(S)]
    return m


def reg_weight(hp):
    w = hp.tacotron_reg_weight
    if hp.tacotron_scale_regularization:
        w *= 1.0 / (2 * hp.max_abs_value) if hp.symmetric_mels else 1.0 / hp.max_abs_value
    return w


def train_forward(hp, weights, ids, lens, mel_targets, stop_targets, seed, dtype=torch.float32):
    """One tower of the training graph.  Returns (dict of outputs + losses, oracle) with autograd enabled;
    ``orc.w[name].grad`` holds d(loss)/d(variable) after ``out['loss'].backward()``."""
    orc = TacotronOracle(hp, weights, dtype=dtype)
    for name, v in orc.w.items():
        if not name.endswith(NOT_TRAINABLE):
            v.requires_grad_(True)
    B, T = ids.shape
    F = mel_targets.shape[1]
    masks = train_masks(hp, seed, B, T, F)
    masks = {k: (torch.utils._pytree.tree_map(lambda x: x.to(dtype), v)) for k, v in masks.items()}
    tgt = torch.as_tensor(mel_targets, dtype=dtype)
    stop_t = torch.as_tensor(stop_targets, dtype=dtype)
    # Tacotron.py:191-193: in training the reference encoder consumes the mel targets
    memory = orc.encode(ids, lens, reference_mel=tgt, is_training=True, train_masks=masks)
    d = orc.decode(memory, lens, mel_targets=tgt, philox_seed=seed, is_training=True, zo_masks=masks["dec_lstm"])
    mel = orc.postnet(d["decoder_output"], is_training=True, drop_masks=masks["postnet"])
    # add_loss, mask_decoder=False branch (Tacotron.py:388-396)
    before = torch.mean((tgt - d["decoder_output"]) ** 2)
    after = torch.mean((tgt - mel) ** 2)
    lg = d["stop"]
    stop_loss = torch.mean(torch.clamp(lg, min=0) - lg * stop_t + torch.log1p(torch.exp(-lg.abs())))
    reg = sum((orc.w[n] ** 2).sum() / 2 for n in orc.w
              if not n.endswith(NOT_TRAINABLE) and not any(s in n for s in REG_EXCLUDE)) * reg_weight(hp)
    loss = before + after + stop_loss + reg
    out = dict(decoder_output=d["decoder_output"], mel_outputs=mel, stop_token_prediction=lg, alignments=d["alignments"],
               encoder_outputs=memory, before_loss=before, after_loss=after, stop_token_loss=stop_loss,
               regularization_loss=reg, loss=loss)
    return out, orc


def gradients(orc):
    return OrderedDict((n, (v.grad if v.grad is not None else torch.zeros_like(v)).detach().numpy().copy())
                       for n, v in orc.w.items() if not n.endswith(NOT_TRAINABLE))


def learning_rate(hp, global_step):
    """_learning_rate_decay (Tacotron.py:518-540)."""
    if not hp.tacotron_decay_learning_rate:
        return hp.tacotron_initial_learning_rate
    lr = hp.tacotron_initial_learning_rate * hp.tacotron_decay_rate ** (
        (global_step - hp.tacotron_start_decay) / float(hp.tacotron_decay_steps))
    return min(max(lr, hp.tacotron_final_learning_rate), hp.tacotron_initial_learning_rate)


def adam_update(hp, weights, tower_grads, m, v, global_step):
    """add_optimizer (Tacotron.py:448-515): mean over towers, clip_by_global_norm(1.0), TF Adam (App. A.14).
    ``weights``/``m``/``v``: name -> float64 ndarray, updated in place; returns the pre-clip global norm."""
    names = list(tower_grads[0].keys())
    avg = {n: np.mean([np.asarray(g[n], np.float64) for g in tower_grads], axis=0) for n in names}
    norm = float(np.sqrt(sum((a ** 2).sum() for a in avg.values())))
    scale = 1.0 / max(norm, 1.0) if hp.tacotron_clip_gradients else 1.0
    t = global_step + 1
    b1, b2, eps = hp.tacotron_adam_beta1, hp.tacotron_adam_beta2, hp.tacotron_adam_epsilon
    lr_t = learning_rate(hp, global_step) * np.sqrt(1 - b2 ** t) / (1 - b1 ** t)
    for n in names:
        g = avg[n] * scale
        m[n] = b1 * m[n] + (1 - b1) * g
        v[n] = b2 * v[n] + (1 - b2) * g * g
        weights[n] = weights[n] - lr_t * m[n] / (np.sqrt(v[n]) + eps)
    return norm
