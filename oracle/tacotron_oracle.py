"""TEST INFRASTRUCTURE (oracle) -- not part of the product path.

CPU restatement (PyTorch-CPU, fp32 by default, fp64 on request) of the
reference's Tacotron-2 mel-generation path.  The reference itself cannot run
here: it needs tensorflow-gpu==1.9.0 + tf.contrib (requirements.txt:5), which is
neither installed nor installable, and it ships no tests, golden vectors or
fixtures.  **Parity is therefore UNPINNED**: this file follows
the reference line by line plus the TF-1.9 op semantics of SURVEY.md Appendix A,
and is cross-checked only against an independent NumPy re-derivation
(oracle/numpy_step.py), known-answer micro cases (tests/test_oracle_*.py) and
fp64 self-consistency.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl
reference legs may import this module.

Each function cites the reference file:line it restates (paths relative to the
reference root).
"""
from __future__ import annotations

import math
from collections import namedtuple

import numpy as np
import torch
import torch.nn.functional as F

from . import philox

W_ENC_CONV = "encoder_convolutions/conv_layer_%d_encoder_convolutions"
W_POST_CONV = "postnet_convolutions/conv_layer_%d_postnet_convolutions"
W_ENC_FW = "encoder_LSTM/bidirectional_rnn/fw/encoder_fw_lstm"
W_ENC_BW = "encoder_LSTM/bidirectional_rnn/bw/encoder_bw_lstm"
W_REF_CONV = "ref_encoder/conv2d_%d"
W_REF_GRU = "ref_encoder/rnn/gru_cell"
W_REF_DENSE = "ref_encoder/dense"
W_MHA = "Multihead-attention"
W_PRENET = "decoder/decoder_prenet/dense_%d"
W_DEC_LSTM = "decoder/decoder_LSTM/multi_rnn_cell/cell_%d/decoder_LSTM_%d"
W_LSA = "decoder/Location_Sensitive_Attention"
W_MEMORY = "decoder/memory_layer"
W_FRAME = "decoder/linear_transform_projection/projection_linear_transform_projection"
W_STOP = "decoder/stop_token_projection/projection_stop_token_projection"
W_POST_PROJ = "postnet_projection/projection_postnet_projection"

BN_EPS = 1e-3  # tf.layers.batch_normalization default epsilon (Appendix A.6)

DecoderState = namedtuple("DecoderState", "c1 h1 c2 h2 attention alignments time")


def _round_bf16(x):
    return x.to(torch.bfloat16).to(x.dtype)


class TacotronOracle(object):
    """Module-for-module restatement of ``Tacotron.initialize`` (Tacotron.py:46-349)."""

    def __init__(self, hp, weights, dtype=torch.float32, conv_bf16=False, dec_bf16=False):
        """``conv_bf16``: round conv/dense GEMM operands of the encoder-conv / postnet stacks to
        bf16 (fp32 accumulate), mirroring the tensor-core kernels' operand precision (this includes the
        hoisted input projection of the encoder LSTMs; their recurrent contraction stays fp32).
        ``dec_bf16``: same for the decoder step's contractions (prenet, both LSTM layers, query
        layer, frame/stop projection), the memory-layer keys and the attention values of the context reduction; recurrent
        state, gate non-linearities, energies, softmax and cumulative alignments stay fp32.  In this mirror mode the
        free-running feedback (Helpers.py:56 -> Decoder.py:346) is contracted the way the cluster decoder does it: the last
        frame's projection and prenet dense_1 precomposed into one [1024 x 256] kernel (see ``_folded_dense1``) -- in exact
        arithmetic the same function; the plain fp32 oracle keeps the reference's two separate contractions."""
        self.hp = hp
        self.dtype = dtype
        self.conv_bf16 = conv_bf16
        self.dec_bf16 = dec_bf16
        self.w = {k: torch.from_numpy(np.asarray(v, np.float32)).to(dtype) for k, v in weights.items()}

    # ------------------------------------------------------------------ primitives
    def _conv1d_same(self, x, kernel, bias, round_ops=False):
        """tf.layers.conv1d(padding='same') NWC, kernel [k,in,out] (Encoder.py:159-164; App. A.4)."""
        k = kernel.shape[0]
        left = (k - 1) // 2
        right = k - 1 - left
        if round_ops:
            x = _round_bf16(x)
            kernel = _round_bf16(kernel)
        xp = F.pad(x.transpose(1, 2), (left, right))           # [B, Cin, T+k-1]
        wt = kernel.permute(2, 1, 0).contiguous()                # [Cout, Cin, k] cross-correlation
        y = F.conv1d(xp, wt, bias)
        return y.transpose(1, 2)

    def _bn(self, x, scope, is_training):
        """tf.layers.batch_normalization axis -1, eps 1e-3 (Encoder.py:165; App. A.6)."""
        g = self.w[scope + "/batch_normalization/gamma"]
        b = self.w[scope + "/batch_normalization/beta"]
        if is_training:
            red = tuple(range(x.dim() - 1))
            mean = x.mean(dim=red)
            var = x.var(dim=red, unbiased=False)
        else:
            mean = self.w[scope + "/batch_normalization/moving_mean"]
            var = self.w[scope + "/batch_normalization/moving_variance"]
        return g * (x - mean) / torch.sqrt(var + BN_EPS) + b

    def _conv_block(self, x, scope, activation, is_training, drop_mask=None):
        """Encoder.conv1d block: conv -> BN -> activation -> dropout (Encoder.py:145-168)."""
        y = self._conv1d_same(x, self.w[scope + "/conv1d/kernel"], self.w[scope + "/conv1d/bias"],
                              round_ops=self.conv_bf16)
        y = self._bn(y, scope, is_training)
        y = activation(y)
        if is_training:
            if drop_mask is None:
                raise ValueError("training-mode conv block needs an explicit dropout keep mask")
            keep = 1.0 - self.hp.dropout_rate
            y = y * drop_mask.to(y.dtype) / keep
        return y

    def _dmm(self, x, w):
        """Decoder contraction; operands rounded to bf16 in the ``dec_bf16`` mirror mode."""
        if self.dec_bf16:
            return _round_bf16(x) @ _round_bf16(w)
        return x @ w

    @staticmethod
    def lstm_cell(x, c, h, kernel, bias, round_ops=False, round_input_only=False):
        """tf.nn.rnn_cell.LSTMCell (Encoder.py:104; App. A.1): gates i,j,f,o; forget_bias 1.0.
        ``round_input_only``: bf16 operands for the INPUT projection only (the encoder's hoisted x . W[0:in]
        runs on the tensor cores in the bf16 modes; the recurrent h . W[in:] stays fp32)."""
        if round_input_only:
            n = x.shape[1]
            z = _round_bf16(x) @ _round_bf16(kernel[:n]) + h @ kernel[n:] + bias
        else:
            xh = torch.cat([x, h], dim=1)
            if round_ops:
                xh, kernel = _round_bf16(xh), _round_bf16(kernel)
            z = xh @ kernel + bias
        i, j, f, o = torch.chunk(z, 4, dim=1)
        new_c = torch.sigmoid(f + 1.0) * c + torch.sigmoid(i) * torch.tanh(j)
        new_h = torch.sigmoid(o) * torch.tanh(new_c)
        return new_c, new_h

    def zoneout_lstm(self, x, c, h, kernel, bias, is_training=False, masks=None, round_ops=False, round_input_only=False):
        """ZoneoutLSTMCell.__call__ (Encoder.py:118-142; App. A.2): output = RAW new_h, state zoned."""
        z = self.hp.zoneout_rate
        new_c, new_h = self.lstm_cell(x, c, h, kernel, bias, round_ops, round_input_only)
        if is_training:
            mc, mh = masks  # Bernoulli(1-z) keep masks; (1-z)*dropout(d, 1-z) == mask*d (Encoder.py:135-136)
            zc = c + mc.to(c.dtype) * (new_c - c)
            zh = h + mh.to(h.dtype) * (new_h - h)
        else:
            zc = (1 - z) * new_c + z * c
            zh = (1 - z) * new_h + z * h
        return new_h, zc, zh

    # ------------------------------------------------------------------ encoder
    def embed(self, ids):
        """tf.nn.embedding_lookup (Tacotron.py:179-181)."""
        return self.w["inputs_embedding"][torch.as_tensor(ids, dtype=torch.long)]

    def encoder_convs(self, x, is_training=False, drop_masks=None):
        """EncoderConvolution.__call__ (Encoder.py:40-46): 3x conv block with ReLU."""
        for i in range(self.hp.conv_num_layers):
            m = None if drop_masks is None else drop_masks[i]
            x = self._conv_block(x, W_ENC_CONV % (i + 1), torch.relu, is_training, m)
        return x

    def _lstm_direction(self, x, lengths, scope, reverse, is_training=False, masks=None):
        """dynamic_rnn with sequence_length (App. A.3): zero output / state copy-through beyond length."""
        B, T, _ = x.shape
        H = self.hp.enc_lstm_hidden_size
        kernel, bias = self.w[scope + "/kernel"], self.w[scope + "/bias"]
        c = torch.zeros(B, H, dtype=x.dtype)
        h = torch.zeros(B, H, dtype=x.dtype)
        out = torch.zeros(B, T, H, dtype=x.dtype)
        lengths = torch.as_tensor(lengths, dtype=torch.long)
        rows = torch.arange(B)
        for s in range(T):
            valid = s < lengths                                            # [B]
            # reverse_sequence by length: step s reads position len-1-s (valid rows only)
            pos = torch.where(valid, (lengths - 1 - s) if reverse else torch.full_like(lengths, s),
                              torch.zeros_like(lengths))
            xt = x[rows, pos]
            m = None if masks is None else (masks[0][s], masks[1][s])
            raw_h, zc, zh = self.zoneout_lstm(xt, c, h, kernel, bias, is_training, m, round_input_only=self.conv_bf16)
            v = valid[:, None]
            c = torch.where(v, zc, c)
            h = torch.where(v, zh, h)
            out[rows[valid], pos[valid]] = raw_h[valid]
        return out

    def encoder_lstm(self, x, lengths, is_training=False, masks=None):
        """EncoderLSTM.__call__ (Encoder.py:68-77): bidirectional_dynamic_rnn, concat fw|bw."""
        fw = self._lstm_direction(x, lengths, W_ENC_FW, False, is_training, None if masks is None else masks[0])
        bw = self._lstm_direction(x, lengths, W_ENC_BW, True, is_training, None if masks is None else masks[1])
        return torch.cat([fw, bw], dim=2)

    # ------------------------------------------------------------------ reference encoder + GST
    def _conv2d_same_s2(self, x, kernel, bias):
        """tf.layers.conv2d(3x3, strides 2, 'same') NHWC (Encoder.py:173-178; App. A.5).
        x: [B, H, W, C]."""
        kh, kw = kernel.shape[0], kernel.shape[1]
        H, W = x.shape[1], x.shape[2]

        def pads(n, k):
            out = -(-n // 2)
            total = max((out - 1) * 2 + k - n, 0)
            return total // 2, total - total // 2

        pt, pb = pads(H, kh)
        pl, pr = pads(W, kw)
        xn = F.pad(x.permute(0, 3, 1, 2), (pl, pr, pt, pb))
        y = F.conv2d(xn, kernel.permute(3, 2, 0, 1).contiguous(), bias, stride=2)
        return y.permute(0, 2, 3, 1)

    def gru_cell(self, x, h):
        """tf.contrib.rnn.GRUCell (Tacotron.py:201; App. A.9): reset applied BEFORE the candidate matmul."""
        R = h.shape[1]
        g = torch.sigmoid(torch.cat([x, h], 1) @ self.w[W_REF_GRU + "/gates/kernel"] + self.w[W_REF_GRU + "/gates/bias"])
        r, u = g[:, :R], g[:, R:]
        c = torch.tanh(torch.cat([x, r * h], 1) @ self.w[W_REF_GRU + "/candidate/kernel"]
                       + self.w[W_REF_GRU + "/candidate/bias"])
        return u * h + (1 - u) * c

    def reference_encoder(self, mel, is_training=False):
        """reference_encoder (Tacotron.py:12-29): 6x(conv2d s2 + BN + ReLU) -> GRU -> last -> dense tanh."""
        x = torch.as_tensor(mel, dtype=self.dtype).unsqueeze(-1)            # [B, Tm, 80, 1]
        for i in range(len(self.hp.reference_filters)):
            scope = W_REF_CONV % i
            x = self._conv2d_same_s2(x, self.w[scope + "/conv2d/kernel"], self.w[scope + "/conv2d/bias"])
            x = torch.relu(self._bn(x, scope, is_training))
        B, Tp, Wp, Cp = x.shape
        x = x.reshape(B, Tp, Wp * Cp)
        h = torch.zeros(B, self.hp.reference_depth, dtype=self.dtype)
        for t in range(Tp):                                                  # no length mask (Tacotron.py:24-27)
            h = self.gru_cell(x[:, t], h)
        return torch.tanh(h @ self.w[W_REF_DENSE + "/kernel"] + self.w[W_REF_DENSE + "/bias"])

    def gst_attention(self, ref):
        """MultiheadAttention.multi_head_attention, mlp_attention, normalize=True (Attention.py:243-341)."""
        hp = self.hp
        B = ref.shape[0]
        nh = hp.num_heads
        tokens = torch.tanh(self.w["style_tokens"])                          # value, [10, 64] tiled over batch
        q = ref @ self.w[W_MHA + "/conv1d/kernel"][0] + self.w[W_MHA + "/conv1d/bias"]            # [B, U]
        k = tokens @ self.w[W_MHA + "/conv1d_1/kernel"][0] + self.w[W_MHA + "/conv1d_1/bias"]     # [10, U]
        U = q.shape[1]
        hu = U // nh
        qs = q.reshape(B, nh, 1, hu)                                         # [B, nh, 1, hu]
        ks = k.reshape(-1, nh, hu).permute(1, 0, 2).unsqueeze(0)             # [1, nh, 10, hu]
        v = self.w[W_MHA + "/attention_v"]
        g = self.w[W_MHA + "/attention_g"]
        b = self.w[W_MHA + "/attention_b"]
        normed_v = g * v * torch.rsqrt(torch.sum(v * v))
        add = torch.sum(normed_v * torch.tanh(ks + qs + b), dim=-1)          # [B, nh, 10]
        wts = torch.softmax(add, dim=-1)
        ctx = wts @ tokens                                                   # [B, nh, 64]
        return ctx.reshape(B, 1, nh * tokens.shape[1])

    def style_random(self, gst_weights):
        """No-reference branch (Tacotron.py:216-222): softmaxed random weights [heads, tokens] @ tanh(tokens)."""
        rw = torch.as_tensor(gst_weights, dtype=self.dtype)
        return (rw @ torch.tanh(self.w["style_tokens"])).reshape(1, 1, -1)

    def encode(self, ids, lengths, reference_mel=None, gst_weights=None, is_training=False, train_masks=None):
        """Embedding -> EncoderCell -> style concat (Tacotron.py:179-223). Returns memory [B, T, 768]."""
        x = self.embed(ids)
        conv_masks = lstm_masks = None
        if is_training:
            conv_masks, lstm_masks = train_masks["enc_conv"], train_masks["enc_lstm"]
        x = self.encoder_convs(x, is_training, conv_masks)
        enc = self.encoder_lstm(x, lengths, is_training, lstm_masks)
        B, T, _ = enc.shape
        if reference_mel is not None:
            ref = self.reference_encoder(reference_mel, is_training)
            self.refnet_outputs = ref
            style = self.gst_attention(ref) if self.hp.use_gst else ref.unsqueeze(1)
            style = style.expand(B, T, style.shape[-1])
        else:
            if gst_weights is None:
                raise ValueError("no reference mel: explicit GST weights [heads, tokens] must be injected")
            style = self.style_random(gst_weights).expand(B, T, -1)
        return torch.cat([enc, style], dim=-1)

    # ------------------------------------------------------------------ attention
    def attention_init(self, memory, lengths):
        """BahdanauAttention base (Attention.py:153-158; App. A.10): masked values, keys computed once."""
        B, T, _ = memory.shape
        lengths = torch.as_tensor(lengths, dtype=torch.long)
        mask = (torch.arange(T)[None, :] < lengths[:, None])
        if self.hp.mask_encoder:
            values = memory * mask[:, :, None].to(memory.dtype)
        else:
            values = memory
            mask = torch.ones_like(mask)
        keys = self._dmm(values, self.w[W_MEMORY + "/kernel"])   # bf16 operands in the dec_bf16 mirror mode
        return values, keys, mask

    def attention_step(self, query, cum, values, keys, mask):
        """LocationSensitiveAttention.__call__ + _compute_attention (Attention.py:169-210,345-366,369-401)."""
        q = self._dmm(query, self.w[W_LSA + "/query_layer/kernel"])                           # [B, A]
        f = self._conv1d_same(cum.unsqueeze(2), self.w[W_LSA + "/location_features_convolution/kernel"],
                              self.w[W_LSA + "/location_features_convolution/bias"])          # [B, T, 32]
        loc = f @ self.w[W_LSA + "/location_features_layer/kernel"]                           # [B, T, A]
        v_a = self.w[W_LSA + "/attention_variable"]
        b_a = self.w[W_LSA + "/attention_bias"]
        energy = torch.sum(v_a * torch.tanh(keys + q.unsqueeze(1) + loc + b_a), dim=2)        # [B, T]
        # TF 1.9 _BaseAttentionMechanism wraps EVERY probability_fn as probability_fn(_maybe_mask_score(score, lengths, -inf)):
        # the -inf mask is applied before the smoothing normalisation too (sigmoid(-inf) = 0)
        energy = torch.where(mask, energy, torch.full_like(energy, -math.inf))
        if self.hp.smoothing:
            sg = torch.sigmoid(energy)                      # Attention.py:405-425 (_smoothing_normalization)
            align = sg / sg.sum(dim=-1, keepdim=True)
        else:
            align = torch.softmax(energy, dim=-1)
        next_cum = align + cum if self.hp.cumulative_weights else align
        vals = _round_bf16(values) if self.dec_bf16 else values
        context = torch.bmm(align.unsqueeze(1), vals).squeeze(1)                              # [B, 768]
        return context, align, next_cum

    # ------------------------------------------------------------------ decoder
    def _folded_dense1(self):
        """bf16 mirror only: (Wo[:, -M:] . Wp1, bo[-M:] . Wp1 + bp1) composed in float64, as csrc/api.cu pack_decoder3 does."""
        if getattr(self, "_fold", None) is None:
            M = self.hp.num_mels
            wo = self.w[W_FRAME + "/kernel"].double()[:, -M:]
            bo = self.w[W_FRAME + "/bias"].double()[-M:]
            wp = self.w[(W_PRENET % 1) + "/kernel"].double()
            bp = self.w[(W_PRENET % 1) + "/bias"].double()
            self._fold = ((wo @ wp).float().to(self.dtype), (bo @ wp + bp).float().to(self.dtype))
        return self._fold

    def prenet(self, x, keep_masks, fold_in=None):
        """Prenet.__call__ (Decoder.py:342-352): dense+ReLU+dropout(0.5, ALWAYS on). keep_masks: [B, L, units].
        ``fold_in`` (bf16 mirror, free-running steps > 0): the previous step's projection input [h2raw ; context]; dense_1 is
        then evaluated through the precomposed kernel instead of from ``x``."""
        keep = 1.0 - self.hp.dropout_rate
        for i in range(len(self.hp.prenet_layers)):
            if i == 0 and fold_in is not None:
                a, b = self._folded_dense1()
                x = torch.relu(_round_bf16(fold_in) @ _round_bf16(a) + b)
            else:
                x = torch.relu(self._dmm(x, self.w[(W_PRENET % (i + 1)) + "/kernel"]) + self.w[(W_PRENET % (i + 1)) + "/bias"])
            x = x * keep_masks[:, i].to(x.dtype) / keep
        return x

    def zero_state(self, B, T):
        """TacotronDecoderCell.zero_state (Decoder.py:212-244)."""
        D = self.hp.decoder_lstm_units
        mem = 2 * self.hp.enc_lstm_hidden_size + self.hp.style_embed_depth
        z = lambda n: torch.zeros(B, n, dtype=self.dtype)
        return DecoderState(z(D), z(D), z(D), z(D), z(mem), z(T), 0)

    def decoder_step(self, x, state, values, keys, mask, keep_masks, is_training=False, zo_masks=None, fold_in=None):
        """TacotronDecoderCell.__call__ (Decoder.py:246-292). Returns (frames [B,160], stop [B,2] (logits), state)."""
        p = self.prenet(x, keep_masks, fold_in)
        lstm_in = torch.cat([p, state.attention], dim=-1)
        k1, b1 = self.w[(W_DEC_LSTM % (0, 1)) + "/kernel"], self.w[(W_DEC_LSTM % (0, 1)) + "/bias"]
        k2, b2 = self.w[(W_DEC_LSTM % (1, 2)) + "/kernel"], self.w[(W_DEC_LSTM % (1, 2)) + "/bias"]
        m1 = m2 = None
        if is_training:
            m1, m2 = zo_masks
        h1raw, c1, h1 = self.zoneout_lstm(lstm_in, state.c1, state.h1, k1, b1, is_training, m1, self.dec_bf16)
        h2raw, c2, h2 = self.zoneout_lstm(h1raw, state.c2, state.h2, k2, b2, is_training, m2, self.dec_bf16)
        context, align, cum = self.attention_step(h2raw, state.alignments, values, keys, mask)
        proj_in = torch.cat([h2raw, context], dim=-1)
        self._last_proj_in = proj_in
        frames = self._dmm(proj_in, self.w[W_FRAME + "/kernel"]) + self.w[W_FRAME + "/bias"]
        stop_logits = self._dmm(proj_in, self.w[W_STOP + "/kernel"]) + self.w[W_STOP + "/bias"]
        nxt = DecoderState(c1, h1, c2, h2, context, cum, state.time + 1)
        return frames, stop_logits, align, nxt

    def decode(self, memory, lengths, mel_targets=None, max_iters=None, prenet_masks=None, philox_seed=None,
               is_training=False, zo_masks=None):
        """dynamic_decode(CustomDecoder(cell, helper)) (Tacotron.py:262-284,312; Helpers.py; App. A.11-12).

        Teacher-forced when ``mel_targets`` is given (TacoTrainingHelper, ratio 1.0), else free-running
        (TacoTestHelper, batch-global stop).  Returns dict with decoder_output [B,2S,80] (clipped),
        stop [B,2S] (sigmoid unless training), alignments [B,T,S], steps S, raw frames.
        """
        hp = self.hp
        B, T, _ = memory.shape
        r, M = hp.outputs_per_step, hp.num_mels
        values, keys, mask = self.attention_init(memory, lengths)
        state = self.zero_state(B, T)
        x = torch.zeros(B, M, dtype=self.dtype)                               # <GO> frame (Helpers.py:131-133)
        teacher = mel_targets is not None
        if teacher:
            tgt = torch.as_tensor(mel_targets, dtype=self.dtype)[:, r - 1::r, :]   # Helpers.py:77
            S_max = tgt.shape[1]
            if max_iters is not None:
                S_max = min(S_max, max_iters)
        else:
            S_max = hp.max_iters if max_iters is None else max_iters
        units = hp.prenet_layers[0]
        frames_l, stop_l, align_l = [], [], []
        for t in range(S_max):
            if prenet_masks is not None:
                km = torch.from_numpy(np.ascontiguousarray(prenet_masks[t]))
            elif philox_seed is not None:
                km = torch.from_numpy(np.stack(
                    [philox.prenet_keep_mask(philox_seed, t, B, units, l) for l in range(len(hp.prenet_layers))], 1))
            else:
                raise ValueError("prenet dropout is always on: pass explicit masks or a philox seed")
            zm = None if zo_masks is None else zo_masks[t]
            fold_in = self._last_proj_in if (self.dec_bf16 and not teacher and t > 0) else None
            frames, stop_logits, align, state = self.decoder_step(x, state, values, keys, mask, km, is_training, zm, fold_in)
            stop = stop_logits if is_training else torch.sigmoid(stop_logits)   # Decoder.py:408-410
            frames_l.append(frames)
            stop_l.append(stop)
            align_l.append(align)
            if teacher:
                x = tgt[:, t]                                                  # Helpers.py:121-124 (ratio 1.0)
            else:
                x = frames[:, -M:]                                             # Helpers.py:56
                fin = torch.round(stop) > 0                                    # tf.round half-to-even, Helpers.py:40
                done = bool(fin.any()) if hp.stop_at_any else bool(fin.all())
                if done:
                    break
        S = len(frames_l)
        frames = torch.stack(frames_l, 1).reshape(B, S * r, M)                 # Tacotron.py:280
        stop = torch.stack(stop_l, 1).reshape(B, S * r)                        # :281
        dec_out = frames
        if hp.clip_outputs:
            lo = -hp.max_abs_value - hp.lower_bound_decay if hp.symmetric_mels else -hp.lower_bound_decay
            dec_out = torch.clamp(frames, lo, hp.max_abs_value)                # :282-284
        align = torch.stack(align_l, 0).permute(1, 2, 0).contiguous()         # :312 [S,B,T] -> [B,T,S]
        return dict(decoder_output=dec_out, frames_raw=frames, stop=stop, alignments=align, steps=S,
                    final_state=state)

    # ------------------------------------------------------------------ postnet
    def postnet(self, dec_out, is_training=False, drop_masks=None):
        """Postnet + projection + residual + clip (Postnet.py:28-36; Tacotron.py:287-309)."""
        hp = self.hp
        x = dec_out
        n = hp.postnet_num_layers
        for i in range(n):
            act = torch.tanh if i < n - 1 else (lambda v: v)
            m = None if drop_masks is None else drop_masks[i]
            x = self._conv_block(x, W_POST_CONV % (i + 1), act, is_training, m)
        kern, bias = self.w[W_POST_PROJ + "/kernel"], self.w[W_POST_PROJ + "/bias"]
        if self.conv_bf16:
            x, kern = _round_bf16(x), _round_bf16(kern)
        mel = dec_out + (x @ kern + bias)
        if hp.clip_outputs:
            lo = -hp.max_abs_value - hp.lower_bound_decay if hp.symmetric_mels else -hp.lower_bound_decay
            mel = torch.clamp(mel, lo, hp.max_abs_value)
        return mel

    # ------------------------------------------------------------------ whole model
    def forward(self, ids, lengths, mel_targets=None, reference_mel=None, gst_weights=None, max_iters=None,
                prenet_masks=None, philox_seed=None):
        """Inference / GTA forward of one tower (Appendix D modes 'GTA synth' and 'free-run inference')."""
        with torch.no_grad():
            memory = self.encode(ids, lengths, reference_mel, gst_weights)
            if max_iters is None:
                max_iters = self.hp.max_iters
            d = self.decode(memory, lengths, mel_targets, max_iters, prenet_masks, philox_seed)
            mel = self.postnet(d["decoder_output"])
        d["mel_outputs"] = mel
        d["encoder_outputs"] = memory
        return d


def output_lengths(stop_tokens):
    """_get_output_lengths (Utils/Tacotron_feeder.py:92-95): first index where round(stop) == 1 else len."""
    out = []
    for row in np.round(np.asarray(stop_tokens)).tolist():
        out.append(row.index(1) if 1 in row else len(row))
    return out
