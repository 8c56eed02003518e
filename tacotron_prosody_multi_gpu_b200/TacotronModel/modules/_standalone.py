"""Standalone calls of the reference's sub-modules (SURVEY 8b "Modules" row) on the stateless device primitives of
include/taco_b200_train.h (``taco_op_*``, bound in train/ops.py).

The product path never comes through here: synthesis runs the fused stages (``taco_encode`` / ``taco_decode`` /
``taco_postnet``), training runs train/step.py.  These functions exist so that a caller who drives the reference's classes one
by one -- ``LocationSensitiveAttention(...)(query, state)``, ``Prenet(...)(x)``, ``DecoderRNN(...)(x, states)``,
``FrameProjection(...)(x)``, ``StopProjection(...)(x)``, ``ZoneoutLSTMCell(...)(x, state)``,
``MultiheadAttention(...).multi_head_attention()``, ``EncoderConvolution(...)(x)``, ``EncoderLSTM(...)(x, lengths)`` -- gets the
same numbers as the reference module, computed on the GPU in fp32 (exact-parity kernels).  No CPU path: ``CudaOps`` raises
without a CUDA device / the library.
"""
from __future__ import annotations

import numpy as np

from ... import weights as W

_cache = {"key": None, "ops": None, "dev": {}}


def _state(hp):
    """(ops, name -> fp32 device tensor) for the installed weight container of this process."""
    from ... import runtime
    from ...train.ops import CudaOps
    w = runtime.weights(hp)
    key = (id(w), runtime.device_index())
    if _cache["key"] != key:
        _cache.update(key=key, ops=CudaOps(runtime.device_index()), dev={})
    return _cache["ops"], w


def _w(hp, name):
    o, w = _state(hp)
    d = _cache["dev"]
    if name not in d:
        d[name] = o.to_device(np.ascontiguousarray(np.asarray(w[name], np.float32)))
    return d[name]


def _dev(o, x, dtype=None):
    import torch
    if isinstance(x, torch.Tensor):
        x = x.to(o.device) if hasattr(o, "device") else x.cuda()
        return x.contiguous() if dtype is None else x.to(dtype).contiguous()
    return o.to_device(np.ascontiguousarray(np.asarray(x)), dtype) if dtype is not None else o.to_device(np.ascontiguousarray(np.asarray(x)))


def dense(hp, x, scope, bias=True):
    """tf.layers.dense: y = x @ kernel[in, out] (+ bias).  x [..., in] -> [..., out]."""
    o, _ = _state(hp)
    x = _dev(o, x)
    k = _w(hp, scope + "/kernel")
    lead = x.shape[:-1]
    x2 = x.reshape(-1, x.shape[-1])
    y = o.empty(x2.shape[0], k.shape[1])
    o.gemm(x2, k, y, bias=_w(hp, scope + "/bias") if bias else None)
    return y.view(*lead, k.shape[1])


def prenet(hp, x, philox_seed=0, step=0):
    """Prenet.__call__ (Decoder.py:342-352): 2 x (Dense 256 + ReLU + dropout 0.5 that is ALWAYS on, kept values x 2).  The
    dropout bits are the engine's Philox stream (seed, step, row, layer, unit), so a standalone call reproduces the fused
    decoder's prenet of that step."""
    o, _ = _state(hp)
    x = _dev(o, x)
    B = x.shape[0]
    keep = 1.0 - hp.dropout_rate
    for layer in (0, 1):
        z = dense(hp, x, W.PRENET % (layer + 1))
        out = o.empty(1, B, z.shape[1])
        o.prenet_act_fwd(z.view(1, B, -1), layer, int(philox_seed), int(step), keep, out)
        x = out.view(B, -1)
    return x


def zoneout_lstm(hp, x, state, scope, zoneout):
    """ZoneoutLSTMCell.__call__ in inference mode (Encoder.py:118-142): gates = [x, h] @ kernel + bias, order i, j, f, o, forget
    bias 1.0; returns (raw h, (zoned c, zoned h)) with zoned = (1 - z) new + z old."""
    o, _ = _state(hp)
    x = _dev(o, x)
    c_prev, h_prev = (_dev(o, s) for s in state)
    K = _w(hp, scope + "/kernel")
    In, H = x.shape[1], h_prev.shape[1]
    if K.shape[0] != In + H:
        raise ValueError("%s: kernel rows %d != inputs %d + units %d" % (scope, K.shape[0], In, H))
    B = x.shape[0]
    gates = o.empty(B, 4 * H)
    o.gemm(x, K[:In], gates)
    o.gemm(h_prev, K[In:], gates, beta=1.0)
    cnew, cz, hraw, hz = (o.empty(B, H) for _ in range(4))
    o.lstm_fwd(gates, None, _w(hp, scope + "/bias"), c_prev, h_prev, None, 0, float(zoneout), False, 0, 0, 0, 0, cnew, cz, hraw, hz)
    return hraw, (cz, hz)


def decoder_rnn(hp, x, states, zoneout):
    """DecoderRNN.__call__ (Decoder.py:295-321): MultiRNNCell of two ZoneoutLSTMCells; layer 2 reads layer 1's RAW output."""
    h1, s1 = zoneout_lstm(hp, x, states[0], W.DEC_LSTM % (0, 1), zoneout)
    h2, s2 = zoneout_lstm(hp, h1, states[1], W.DEC_LSTM % (1, 2), zoneout)
    return h2, (s1, s2)


def memory_keys(hp, memory, lengths, mask_encoder=True):
    """BahdanauAttention base (Attention.py:153-158): values = memory zeroed beyond the length, keys = values @ W_mem."""
    import torch
    o, _ = _state(hp)
    values = _dev(o, memory).clone()
    B, T, MD = values.shape
    if mask_encoder and lengths is not None:
        lens = _dev(o, lengths, torch.int32)
        o.mask_rows(values, lens)
    else:
        lens = o.to_device(np.full((B,), T, np.int32))
    k = _w(hp, W.MEMORY_LAYER + "/kernel")
    keys = o.empty(B, T, k.shape[1])
    o.gemm(values.view(B * T, MD), k, keys.view(B * T, -1))
    return values, keys, lens


def location_sensitive_attention(hp, query, cum_prev, values, keys, lens):
    """LocationSensitiveAttention.__call__ (Attention.py:169-210, score :369-401): returns (alignments, next_state) where
    next_state = cumulative alignments (cumulate_weights) -- plus the context of _compute_attention (:345-366)."""
    o, _ = _state(hp)
    q = dense(hp, query, W.LSA + "/query_layer", bias=False)
    cum_prev = _dev(o, cum_prev)
    B, T = cum_prev.shape
    align, cum, ctx = o.empty(B, T), o.empty(B, T), o.empty(B, values.shape[2])
    o.attn_fwd(q, cum_prev, keys, values, lens, _w(hp, W.LSA + "/location_features_convolution/kernel"),
               _w(hp, W.LSA + "/location_features_convolution/bias"), _w(hp, W.LSA + "/location_features_layer/kernel"),
               _w(hp, W.LSA + "/attention_variable"), _w(hp, W.LSA + "/attention_bias"), align, cum, ctx)
    if not hp.cumulative_weights:
        cum = align
    return align, cum, ctx


def gst_multi_head_attention(hp, query, num_heads):
    """MultiheadAttention.multi_head_attention (Attention.py:243-341) with value = tanh(style tokens): query [B, 128] (the
    reference-encoder output) -> style embedding [B, 1, 256]."""
    o, _ = _state(hp)
    ref = _dev(o, query).reshape(-1, query.shape[-1])
    tok = _w(hp, "style_tokens")
    tt = o.empty(*tok.shape)
    o.ew("tanh", tt, tok)
    U = hp.style_att_dim
    gq, gk = o.empty(ref.shape[0], U), o.empty(tok.shape[0], U)
    o.gemm(ref, _w(hp, W.MHA + "/conv1d/kernel")[0], gq, bias=_w(hp, W.MHA + "/conv1d/bias"))
    o.gemm(tt, _w(hp, W.MHA + "/conv1d_1/kernel")[0], gk, bias=_w(hp, W.MHA + "/conv1d_1/bias"))
    wts, style = o.empty(ref.shape[0], num_heads, tok.shape[0]), o.empty(ref.shape[0], num_heads * tok.shape[1])
    o.gst_fwd(gq, gk, _w(hp, W.MHA + "/attention_v"), _w(hp, W.MHA + "/attention_g"), _w(hp, W.MHA + "/attention_b"), tt, num_heads, wts, style)
    return style.view(ref.shape[0], 1, -1), wts


def _folded_bn(hp, scope, conv="conv1d"):
    """Inference batch norm (moving statistics, eps 1e-3, Encoder.py:165) folded into the convolution that precedes it:
    w' = w * g / sqrt(var + eps) per output channel, b' = (b - mean) * g / sqrt(var + eps) + beta (host, float64, cached)."""
    o, w = _state(hp)
    key = "__folded__/" + scope
    d = _cache["dev"]
    if key not in d:
        k = np.asarray(w[scope + "/%s/kernel" % conv], np.float64)
        b = np.asarray(w[scope + "/%s/bias" % conv], np.float64)
        bn = scope + "/batch_normalization/"
        s = np.asarray(w[bn + "gamma"], np.float64) / np.sqrt(np.asarray(w[bn + "moving_variance"], np.float64) + 1e-3)
        d[key] = (o.to_device((k * s).astype(np.float32)), o.to_device(((b - np.asarray(w[bn + "moving_mean"], np.float64)) * s
                                                                       + np.asarray(w[bn + "beta"], np.float64)).astype(np.float32)))
    return d[key]


def _relu_(o, x):
    o.clip_fwd(x, 0.0, 3.0e38, x)
    return x


def encoder_convolution(hp, x):
    """EncoderConvolution.__call__ in inference mode (Encoder.py:40-46,145-168): conv_num_layers x (conv1d k5 'same' + BN + ReLU)
    on embedded inputs [B, T, 512]."""
    o, _ = _state(hp)
    x = _dev(o, x)
    for i in range(hp.conv_num_layers):
        k, b = _folded_bn(hp, W.ENC_CONV % (i + 1))
        y = o.empty(x.shape[0], x.shape[1], k.shape[2])
        o.conv1d_fwd(x, k, b, y)
        x = _relu_(o, y)
    return x


def encoder_lstm(hp, x, lengths, zoneout):
    """EncoderLSTM.__call__ in inference mode (Encoder.py:68-77): bidirectional ZoneoutLSTM over the valid length of every row,
    outputs fw || bw raw h, zero beyond the length; the backward direction runs on the length-reversed sequence."""
    import torch
    o, _ = _state(hp)
    x = _dev(o, x)
    B, T, C = x.shape
    lens = _dev(o, lengths, torch.int32)
    H = hp.enc_lstm_hidden_size
    out = o.zeros(B, T, 2 * H)
    for d, (scope, rev) in enumerate(((W.ENC_LSTM_FW, False), (W.ENC_LSTM_BW, True))):
        K = _w(hp, scope + "/kernel")
        xw = o.empty(B, T, 4 * H)
        o.gemm(x.view(B * T, C), K[:C], xw.view(B * T, 4 * H), bias=_w(hp, scope + "/bias"))
        cz, hz = o.zeros(2, B, H), o.zeros(2, B, H)
        pre, acts, cnew, hraw = o.empty(B, 4 * H), o.empty(B, 4 * H), o.empty(B, H), o.empty(B, H)
        for s in range(T):
            a, b = s & 1, (s + 1) & 1
            o.gather_rows(xw, lens, s, rev, pre)
            o.gemm(hz[a], K[C:], acts)
            o.lstm_fwd(acts, pre, None, cz[a], hz[a], lens, s, float(zoneout), False, 0, 0, 0, 0, cnew, cz[b], hraw, hz[b])
            o.scatter_rows(hraw, lens, s, rev, out[:, :, d * H:(d + 1) * H])
    return out


def reference_encoder(hp, mel):
    """reference_encoder (Tacotron.py:12-29) in inference mode: 6 x (conv2d 3x3 stride 2 'same' + BN + ReLU) -> GRU(128) over
    the remaining time steps (no length mask) -> last output -> Dense(128, tanh).  mel [B, Tm, 80] -> [B, 128]."""
    o, _ = _state(hp)
    x = _dev(o, mel)
    B, Tm, M = x.shape
    x = x.reshape(B, Tm, M, 1).contiguous()
    for i in range(len(hp.reference_filters)):
        k, b = _folded_bn(hp, W.REF_CONV % i, conv="conv2d")
        Ho, Wo = (x.shape[1] + 1) // 2, (x.shape[2] + 1) // 2
        y = o.empty(B, Ho, Wo, k.shape[3])
        o.conv2d_fwd(x, k, b, y)
        x = _relu_(o, y)
    _, Tp, Wp, Cp = x.shape
    In, R = Wp * Cp, hp.reference_depth
    Xt = o.empty(Tp, B, In)
    Xt.copy_(x.view(B, Tp, In).transpose(0, 1))
    Wg, Wc = _w(hp, W.REF_GRU + "/gates/kernel"), _w(hp, W.REF_GRU + "/candidate/kernel")
    GX, CX = o.empty(Tp, B, 2 * R), o.empty(Tp, B, R)
    o.gemm(Xt.view(Tp * B, In), Wg[:In], GX.view(Tp * B, 2 * R), bias=_w(hp, W.REF_GRU + "/gates/bias"))
    o.gemm(Xt.view(Tp * B, In), Wc[:In], CX.view(Tp * B, R), bias=_w(hp, W.REF_GRU + "/candidate/bias"))
    h, rh = o.zeros(2, B, R), o.empty(B, R)
    for s in range(Tp):                       # GRUCell: reset applied before the candidate matmul
        a, b = s & 1, (s + 1) & 1
        o.gemm(h[a], Wg[In:], GX[s], beta=1.0)
        o.gru_pre(GX[s], h[a], rh)
        o.gemm(rh, Wc[In:], CX[s], beta=1.0)
        o.gru_out(CX[s], GX[s], h[a], h[b])
    refz, ref = o.empty(B, 128), o.empty(B, 128)
    o.gemm(h[Tp & 1], _w(hp, W.REF_DENSE + "/kernel"), refz, bias=_w(hp, W.REF_DENSE + "/bias"))
    o.ew("tanh", ref, refz)
    return ref
