"""TEST INFRASTRUCTURE (oracle) -- not part of the product path.

Independent NumPy (fp64, explicit index loops where layout matters) re-derivation
of one decoder step and of the TF-semantics corner cases, used ONLY to pin
oracle/tacotron_oracle.py: two restatements written in
different styles must agree to 1e-6.  Follows Decoder.py:246-292,
Attention.py:169-210,345-401, Encoder.py:118-142 and Appendix A of SURVEY.md.
"""
import numpy as np

from .tacotron_oracle import (W_PRENET, W_DEC_LSTM, W_LSA, W_FRAME, W_STOP)


def sigmoid(x):
    return 1.0 / (1.0 + np.exp(-x))


def lstm_cell_np(x, c, h, kernel, bias):
    """App. A.1, written per gate with explicit slicing."""
    H = h.shape[1]
    z = np.concatenate([x, h], axis=1).dot(kernel) + bias
    i, j, f, o = z[:, 0:H], z[:, H:2 * H], z[:, 2 * H:3 * H], z[:, 3 * H:4 * H]
    new_c = sigmoid(f + 1.0) * c + sigmoid(i) * np.tanh(j)
    new_h = sigmoid(o) * np.tanh(new_c)
    return new_c, new_h


def location_conv_np(cum, kernel, bias):
    """conv1d k=31 'same' over [B,T,1] with explicit loops (App. A.4: pad 15/15, cross-correlation)."""
    B, T = cum.shape
    K, _, Fn = kernel.shape
    half = (K - 1) // 2
    out = np.zeros((B, T, Fn))
    for b in range(B):
        for t in range(T):
            acc = bias.astype(np.float64).copy()
            for k in range(K):
                src = t + k - half
                if 0 <= src < T:
                    acc += cum[b, src] * kernel[k, 0, :]
            out[b, t] = acc
    return out


def decoder_step_np(w, hp, x, st, values, keys, lengths, keep_masks):
    """One step; ``st`` = dict(c1,h1,c2,h2,attention,alignments) of fp64 arrays."""
    W = {k: np.asarray(v, np.float64) for k, v in w.items()}
    z = hp.zoneout_rate
    p = x
    for i in range(len(hp.prenet_layers)):
        p = np.maximum(p.dot(W[(W_PRENET % (i + 1)) + "/kernel"]) + W[(W_PRENET % (i + 1)) + "/bias"], 0.0)
        p = p * keep_masks[:, i].astype(np.float64) * (1.0 / (1.0 - hp.dropout_rate))
    lstm_in = np.concatenate([p, st["attention"]], axis=1)
    nc1, nh1 = lstm_cell_np(lstm_in, st["c1"], st["h1"], W[(W_DEC_LSTM % (0, 1)) + "/kernel"], W[(W_DEC_LSTM % (0, 1)) + "/bias"])
    c1 = (1 - z) * nc1 + z * st["c1"]
    h1 = (1 - z) * nh1 + z * st["h1"]
    nc2, nh2 = lstm_cell_np(nh1, st["c2"], st["h2"], W[(W_DEC_LSTM % (1, 2)) + "/kernel"], W[(W_DEC_LSTM % (1, 2)) + "/bias"])
    c2 = (1 - z) * nc2 + z * st["c2"]
    h2 = (1 - z) * nh2 + z * st["h2"]
    q = nh2.dot(W[W_LSA + "/query_layer/kernel"])
    f = location_conv_np(st["alignments"], W[W_LSA + "/location_features_convolution/kernel"],
                         W[W_LSA + "/location_features_convolution/bias"])
    loc = f.dot(W[W_LSA + "/location_features_layer/kernel"])
    B, T = st["alignments"].shape
    energy = np.empty((B, T))
    va, ba = W[W_LSA + "/attention_variable"], W[W_LSA + "/attention_bias"]
    for b in range(B):
        for t in range(T):
            energy[b, t] = np.sum(va * np.tanh(keys[b, t] + q[b] + loc[b, t] + ba))
    align = np.zeros((B, T))
    for b in range(B):
        n = int(lengths[b])
        e = energy[b, :n]
        e = np.exp(e - e.max())
        align[b, :n] = e / e.sum()
    cum = st["alignments"] + align
    ctx = np.einsum("bt,btd->bd", align, values)
    pin = np.concatenate([nh2, ctx], axis=1)
    frames = pin.dot(W[W_FRAME + "/kernel"]) + W[W_FRAME + "/bias"]
    stop_logits = pin.dot(W[W_STOP + "/kernel"]) + W[W_STOP + "/bias"]
    nst = dict(c1=c1, h1=h1, c2=c2, h2=h2, attention=ctx, alignments=cum)
    return frames, stop_logits, align, nst


def same_pad_stride2(n, k=3):
    """App. A.5: (out, pad_before, pad_after) for tf 'same' stride-2."""
    out = -(-n // 2)
    total = max((out - 1) * 2 + k - n, 0)
    return out, total // 2, total - total // 2


def conv2d_same_s2_np(x, kernel, bias):
    """Explicit-loop NHWC 3x3 stride-2 SAME conv (small inputs only)."""
    B, H, Wd, C = x.shape
    kh, kw, _, Co = kernel.shape
    Ho, pt, _ = same_pad_stride2(H, kh)
    Wo, pl, _ = same_pad_stride2(Wd, kw)
    out = np.zeros((B, Ho, Wo, Co))
    for b in range(B):
        for i in range(Ho):
            for j in range(Wo):
                acc = bias.astype(np.float64).copy()
                for di in range(kh):
                    for dj in range(kw):
                        y, xx = i * 2 + di - pt, j * 2 + dj - pl
                        if 0 <= y < H and 0 <= xx < Wd:
                            acc += x[b, y, xx, :].dot(kernel[di, dj])
                out[b, i, j] = acc
    return out
