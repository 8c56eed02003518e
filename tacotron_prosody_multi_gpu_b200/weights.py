"""Weight container keyed by the reference's variable-scope names, TF layouts.

No checkpoint of the reference is published (README.md:21) so all parity work
runs on identical random-init weights (BASELINE.json north_star).  Shapes,
layouts and default initialisers follow the TF-1.9 calls made at the cited
lines; names follow the variable scopes listed in SURVEY.md Appendix B
(prefix ``model/inference/`` dropped):

* Dense kernels ``[in, out]``; conv1d ``[k, in, out]``; conv2d ``[kh, kw, in, out]``
* LSTM kernel ``[in + h, 4h]`` gate order i, j, f, o (tf.nn.rnn_cell.LSTMCell)
* GRU ``gates`` ``[in + h, 2h]`` (r, u; bias init 1.0), ``candidate`` ``[in + h, h]``

The container is a plain ``dict[str, np.ndarray(float32)]``; ``save``/``load``
use a flat ``.npz`` (the repo's own checkpoint format).
"""
from __future__ import annotations

import math
from collections import OrderedDict

import numpy as np

from .hparams import num_symbols

ENC_CONV = "encoder_convolutions/conv_layer_%d_encoder_convolutions"
POST_CONV = "postnet_convolutions/conv_layer_%d_postnet_convolutions"
ENC_LSTM_FW = "encoder_LSTM/bidirectional_rnn/fw/encoder_fw_lstm"
ENC_LSTM_BW = "encoder_LSTM/bidirectional_rnn/bw/encoder_bw_lstm"
REF_CONV = "ref_encoder/conv2d_%d"
REF_GRU = "ref_encoder/rnn/gru_cell"
REF_DENSE = "ref_encoder/dense"
MHA = "Multihead-attention"
PRENET = "decoder/decoder_prenet/dense_%d"
DEC_LSTM = "decoder/decoder_LSTM/multi_rnn_cell/cell_%d/decoder_LSTM_%d"
LSA = "decoder/Location_Sensitive_Attention"
MEMORY_LAYER = "decoder/memory_layer"
FRAME_PROJ = "decoder/linear_transform_projection/projection_linear_transform_projection"
STOP_PROJ = "decoder/stop_token_projection/projection_stop_token_projection"
POST_PROJ = "postnet_projection/projection_postnet_projection"


def _glorot(rng, shape, fan_in, fan_out):
    lim = math.sqrt(6.0 / (fan_in + fan_out))
    return rng.uniform(-lim, lim, size=shape).astype(np.float32)


def _dense(rng, w, name, n_in, n_out, bias=True):
    w[name + "/kernel"] = _glorot(rng, (n_in, n_out), n_in, n_out)
    if bias:
        w[name + "/bias"] = np.zeros((n_out,), np.float32)


def _bn(rng, w, name, c, nontrivial):
    bn = name + "/batch_normalization"
    w[bn + "/gamma"] = np.ones((c,), np.float32)
    w[bn + "/beta"] = np.zeros((c,), np.float32)
    if nontrivial:
        # non-trivial moving statistics so BN folding is exercised
        w[bn + "/moving_mean"] = (rng.standard_normal(c) * 0.1).astype(np.float32)
        w[bn + "/moving_variance"] = rng.uniform(0.5, 1.5, size=c).astype(np.float32)
    else:
        w[bn + "/moving_mean"] = np.zeros((c,), np.float32)
        w[bn + "/moving_variance"] = np.ones((c,), np.float32)


def _conv1d(rng, w, name, k, c_in, c_out, nontrivial_bn):
    w[name + "/conv1d/kernel"] = _glorot(rng, (k, c_in, c_out), k * c_in, k * c_out)
    w[name + "/conv1d/bias"] = np.zeros((c_out,), np.float32)
    _bn(rng, w, name, c_out, nontrivial_bn)


def _lstm(rng, w, name, n_in, h):
    w[name + "/kernel"] = _glorot(rng, (n_in + h, 4 * h), n_in + h, 4 * h)
    w[name + "/bias"] = np.zeros((4 * h,), np.float32)


def init_weights(hp, seed=None, nontrivial_bn=True, perturb=0.0):
    """Random-init every variable of ``Tacotron.initialize`` (Tacotron.py:171-309).

    ``perturb`` > 0 additionally replaces every all-zero bias / unit gamma by a
    small random value (test-only: zero biases would hide indexing bugs).
    """
    seed = hp.tacotron_random_seed if seed is None else seed
    rng = np.random.Generator(np.random.PCG64(seed))
    w = OrderedDict()
    E = hp.embedding_dim
    C = hp.conv_channels
    H = hp.enc_lstm_hidden_size
    k = int(hp.conv_kernel_size[0])
    style = hp.style_embed_depth
    tok_d = style // hp.num_heads
    mem = 2 * H + style
    A = hp.attention_dim
    D = hp.decoder_lstm_units
    M = hp.num_mels
    r = hp.outputs_per_step

    # Tacotron.py:173-181
    w["style_tokens"] = np.clip(rng.standard_normal((hp.num_gst, tok_d)), -2.0, 2.0).astype(np.float32) * 0.5
    nsym = num_symbols(hp)
    w["inputs_embedding"] = _glorot(rng, (nsym, E), nsym, E)
    # Encoder.py:40-46 / :145-168
    c_in = E
    for i in range(hp.conv_num_layers):
        _conv1d(rng, w, ENC_CONV % (i + 1), k, c_in, C, nontrivial_bn)
        c_in = C
    # Encoder.py:64-65 (input = conv channels)
    _lstm(rng, w, ENC_LSTM_FW, C, H)
    _lstm(rng, w, ENC_LSTM_BW, C, H)
    # Tacotron.py:12-29 reference encoder
    c_in = 1
    freq = M
    for i, c_out in enumerate(hp.reference_filters):
        name = REF_CONV % i
        w[name + "/conv2d/kernel"] = _glorot(rng, (3, 3, c_in, c_out), 9 * c_in, 9 * c_out)
        w[name + "/conv2d/bias"] = np.zeros((c_out,), np.float32)
        _bn(rng, w, name, c_out, nontrivial_bn)
        c_in = c_out
        freq = (freq + 1) // 2
    gru_in = freq * c_in
    R = hp.reference_depth
    w[REF_GRU + "/gates/kernel"] = _glorot(rng, (gru_in + R, 2 * R), gru_in + R, 2 * R)
    w[REF_GRU + "/gates/bias"] = np.ones((2 * R,), np.float32)
    w[REF_GRU + "/candidate/kernel"] = _glorot(rng, (gru_in + R, R), gru_in + R, R)
    w[REF_GRU + "/candidate/bias"] = np.zeros((R,), np.float32)
    _dense(rng, w, REF_DENSE, R, 128)
    # Attention.py:243-341 GST multi-head attention (conv1d k=1 for q and k)
    U = hp.style_att_dim
    w[MHA + "/conv1d/kernel"] = _glorot(rng, (1, 128, U), 128, U)
    w[MHA + "/conv1d/bias"] = np.zeros((U,), np.float32)
    w[MHA + "/conv1d_1/kernel"] = _glorot(rng, (1, tok_d, U), tok_d, U)
    w[MHA + "/conv1d_1/bias"] = np.zeros((U,), np.float32)
    hu = U // hp.num_heads
    w[MHA + "/attention_v"] = _glorot(rng, (hu,), hu, hu)
    w[MHA + "/attention_g"] = np.asarray(math.sqrt(1.0 / hu), np.float32).reshape(())
    w[MHA + "/attention_b"] = np.zeros((hu,), np.float32)
    # Decoder.py:342-352 prenet
    n_in = M
    for i, size in enumerate(hp.prenet_layers):
        _dense(rng, w, PRENET % (i + 1), n_in, size)
        n_in = size
    # Decoder.py:312-317 decoder LSTMs: layer 1 input = prenet + memory width
    _lstm(rng, w, DEC_LSTM % (0, 1), n_in + mem, D)
    _lstm(rng, w, DEC_LSTM % (1, 2), D, D)
    # Attention.py:153-166 (+ Bahdanau base memory/query layers), :394-399
    _dense(rng, w, MEMORY_LAYER, mem, A, bias=False)
    _dense(rng, w, LSA + "/query_layer", D, A, bias=False)
    kl = int(hp.attention_kernel[0])
    F = hp.attention_filters
    w[LSA + "/location_features_convolution/kernel"] = _glorot(rng, (kl, 1, F), kl, kl * F)
    w[LSA + "/location_features_convolution/bias"] = np.zeros((F,), np.float32)
    _dense(rng, w, LSA + "/location_features_layer", F, A, bias=False)
    w[LSA + "/attention_variable"] = _glorot(rng, (A,), A, A)
    w[LSA + "/attention_bias"] = np.zeros((A,), np.float32)
    # Decoder.py:371, :404 projections on [LSTM_output, context]
    _dense(rng, w, FRAME_PROJ, D + mem, M * r)
    _dense(rng, w, STOP_PROJ, D + mem, r)
    # Postnet.py:22,28-35 (width = postnet_channels + style_embed_depth)
    PC = hp.postnet_channels + hp.style_embed_depth
    kp = int(hp.postnet_kernel_size[0])
    c_in = M
    for i in range(hp.postnet_num_layers):
        _conv1d(rng, w, POST_CONV % (i + 1), kp, c_in, PC, nontrivial_bn)
        c_in = PC
    _dense(rng, w, POST_PROJ, PC, M)

    if perturb > 0.0:
        for name in list(w.keys()):
            v = w[name]
            if name.endswith("/bias") or name.endswith("/beta") or name.endswith("attention_bias") \
                    or name.endswith("attention_b"):
                w[name] = (v + rng.standard_normal(v.shape) * perturb).astype(np.float32)
            elif name.endswith("/gamma"):
                w[name] = (v * rng.uniform(1.0 - perturb, 1.0 + perturb, size=v.shape)).astype(np.float32)
    return w


TRAINABLE_EXCLUDE = ("moving_mean", "moving_variance")


def count_trainable(w):
    return int(sum(v.size for n, v in w.items() if not n.endswith(TRAINABLE_EXCLUDE)))


def set_stop_bias(w, value):
    """Throughput configs: push the stop logit far from 0 so decoding runs to max_iters."""
    w = OrderedDict(w)
    w[STOP_PROJ + "/bias"] = np.full_like(w[STOP_PROJ + "/bias"], value)
    return w


def scale_stop_projection(w, scale, bias):
    """Stop-parity configs: scale the stop kernel and shift the bias."""
    w = OrderedDict(w)
    w[STOP_PROJ + "/kernel"] = (w[STOP_PROJ + "/kernel"] * scale).astype(np.float32)
    w[STOP_PROJ + "/bias"] = np.full_like(w[STOP_PROJ + "/bias"], bias)
    return w


def save(path, w):
    np.savez(path, **{k.replace("/", "|"): v for k, v in w.items()})


def load(path):
    z = np.load(path)
    return OrderedDict((k.replace("|", "/"), z[k]) for k in z.files)
