"""Teacher-forced training step of the Tacotron-2 path (config C4 of BASELINE.json).

Restates ``Tacotron.initialize(is_training=True)`` + ``add_loss`` + ``add_optimizer``
(TacotronModel/modules/Tacotron.py:46-349, 351-446, 448-540) as a host-side composition of stateless
device primitives (include/taco_b200_train.h, bound by ``train/ops.py``); torch tensors only carry
device memory.  The backward pass is hand-derived (no autograd anywhere on this path):

* everything that does not depend on the recurrence is hoisted out of the time loops as one large
  contraction over all S*B rows: the whole prenet, the prenet term of the LSTM-1 gates, the frame /
  stop projections, every weight gradient (dW = X^T dG over all steps) and d(values);
* the loops themselves carry only [B, 256..1024] vectors: 2 LSTM cells + one fused
  location-sensitive-attention primitive per decoder step, forward and backward.

Mode semantics (column "train"): batch-statistics BN + moving-stat update,
encoder / postnet dropout 0.5, prenet dropout 0.5, Bernoulli(0.9) zoneout masks on all four LSTMs,
stop projection returns logits, style embedding from the reference encoder run on ``mel_targets``.
All random bits come from counter-based Philox streams keyed by ``seed`` (oracle/philox.py restates
them) -- the injected stand-in for TF's stateful RNG.
"""
from __future__ import annotations

import contextlib

import numpy as np

from .. import weights as W
from .params import ParamStore

ACT_NONE, ACT_RELU, ACT_TANH = 0, 1, 2
BN_MOMENTUM = 0.99      # tf.layers.batch_normalization default (Encoder.py:165)

# Philox mask streams (oracle/train_oracle.py uses the same numbers)
STREAM_ENC_CONV = 100   # + layer
STREAM_POSTNET = 200    # + layer
STREAM_ENC_LSTM = 300   # + 2 * direction (+1 for h)
STREAM_DEC_LSTM = 400   # + 2 * layer (+1 for h)


class TrainStep(object):
    def __init__(self, hp, ops, weights):
        self.hp = hp
        self.ops = ops
        reg = hp.tacotron_reg_weight
        if hp.tacotron_scale_regularization:
            reg *= 1.0 / (2 * hp.max_abs_value) if hp.symmetric_mels else 1.0 / hp.max_abs_value
        # hparams that change the reference's TRAINING graph and are not built here fail loudly (the inference path raises
        # FusedStageError for the same settings) instead of silently training a different model
        if str(getattr(hp, "tacotron_teacher_forcing_mode", "constant")) != "constant":
            raise NotImplementedError("tacotron_teacher_forcing_mode=%r (scheduled teacher forcing, Helpers.py:100-124) is not built: "
                                      "only 'constant' with ratio 1.0" % hp.tacotron_teacher_forcing_mode)
        if float(getattr(hp, "tacotron_teacher_forcing_ratio", 1.0)) != 1.0:
            raise NotImplementedError("tacotron_teacher_forcing_ratio=%r: only 1.0 (pure teacher forcing, Helpers.py:118-124) is built"
                                      % hp.tacotron_teacher_forcing_ratio)
        if bool(getattr(hp, "predict_linear", False)):
            raise NotImplementedError("predict_linear=True (CBHG post-processing net, Tacotron.py:319-337) is out of scope")
        self.store = ParamStore(ops, weights, reg)
        # tacotron_fine_tuning (Tacotron.py:478-480): variables whose name contains 'inputs_embedding' or 'encoder_' are
        # excluded from the update; here a 0/1 mask over the flat gradient, applied before the global norm
        self.update_mask = None
        self.regw_grad = self.store.regw
        if bool(getattr(hp, "tacotron_fine_tuning", False)):
            import numpy as _np
            from .params import FINE_TUNE_FROZEN
            mask = _np.ones(self.store.n, _np.float32)
            for name, (o_, shp) in self.store.layout.items():
                if any(f in name for f in FINE_TUNE_FROZEN):
                    mask[o_:o_ + (int(_np.prod(shp)) if len(shp) else 1)] = 0.0
            self.update_mask = ops.to_device(mask)
            # compute_gradients(var_list=update_vars): the frozen variables get no L2 gradient either (the LOSS keeps their term)
            self.regw_grad = ops.to_device(mask)
            ops.ew("mul", self.regw_grad, self.regw_grad, self.store.regw)
        self.on_tail_ready = None     # set by the data-parallel trainer: called in backward() when store.g[tail_start:] is final
        self.losses = None
        self.update_moving_stats = True
        lo = (-hp.max_abs_value if hp.symmetric_mels else 0.0) - hp.lower_bound_decay
        self.clip = (lo, hp.max_abs_value)

    # ------------------------------------------------------------------ helpers
    def _bn(self, scope):
        P, G, st = self.store.P, self.store.G, self.store.stats
        b = scope + "/batch_normalization"
        return P[b + "/gamma"], P[b + "/beta"], G[b + "/gamma"], G[b + "/beta"], st[b + "/moving_mean"], st[b + "/moving_variance"]

    def _conv_block_fwd(self, x, scope, act, keep, seed, stream, tape, conv2d=False):
        """conv (+bias) -> batch-stat BN -> activation -> dropout (Encoder.py:145-168 / :171-182)."""
        o, P = self.ops, self.store.P
        kind = "/conv2d" if conv2d else "/conv1d"
        w, b = P[scope + kind + "/kernel"], P[scope + kind + "/bias"]
        if conv2d:
            z = o.empty(x.shape[0], (x.shape[1] + 1) // 2, (x.shape[2] + 1) // 2, w.shape[3])
            o.conv2d_fwd(x, w, b, z)
        else:
            z = o.empty(x.shape[0], x.shape[1], w.shape[2])
            o.conv1d_fwd(x, w, b, z)
        C = z.shape[-1]
        y, mean, rstd = o.empty(*z.shape), o.empty(C), o.empty(C)
        gamma, beta, _, _, mm, mv = self._bn(scope)
        o.bn_fwd(z, gamma, beta, act, keep, seed, stream, y, mean, rstd)
        if self.update_moving_stats:
            o.bn_update_moving(mean, rstd, mm, mv, BN_MOMENTUM)
        tape.append(dict(x=x, z=z, y=y, mean=mean, rstd=rstd, scope=scope, act=act, keep=keep, stream=stream, conv2d=conv2d))
        return y

    def _conv_block_bwd(self, dy, rec, seed, need_dx=True):
        o, P, G = self.ops, self.store.P, self.store.G
        scope = rec["scope"]
        kind = "/conv2d" if rec["conv2d"] else "/conv1d"
        gamma, _, dgamma, dbeta, _, _ = self._bn(scope)
        dz = o.empty(*rec["z"].shape)
        o.bn_bwd(dy, rec["z"], rec["y"], rec["mean"], rec["rstd"], gamma, rec["act"], rec["keep"], seed, rec["stream"],
                 dz, dgamma, dbeta)
        o.colsum(dz, G[scope + kind + "/bias"])
        w = P[scope + kind + "/kernel"]
        dx = None
        if rec["conv2d"]:
            o.conv2d_wgrad(rec["x"], dz, G[scope + kind + "/kernel"])
            if need_dx:
                dx = o.empty(*rec["x"].shape)
                o.conv2d_dgrad(dz, w, dx)
        else:
            o.conv1d_wgrad(rec["x"], dz, G[scope + kind + "/kernel"])
            if need_dx:
                dx = o.empty(*rec["x"].shape)
                o.conv1d_dgrad(dz, w, dx)
        return dx

    def _mark(self, name):
        """Time stamp on the current stream (only when ``self.marks`` is a list: scripts/bench_train.py --phases)."""
        if getattr(self, "marks", None) is not None:
            import torch
            ev = torch.cuda.Event(enable_timing=True)
            ev.record()
            self.marks.append((name, ev))

    def _side(self, i):
        """Launches of an independent branch of the step go to side stream i (CudaOps.side); a no-op for the CPU twin."""
        return self.ops.side(i) if hasattr(self.ops, "side") else contextlib.nullcontext()

    def _join(self):
        if hasattr(self.ops, "join"):
            self.ops.join()

    def _dense_bwd(self, x2, dy2, kernel_name, bias_name, dx2=None, beta=0.0):
        """dW = x^T dy, db = colsum(dy), optionally dx (+)= dy W^T."""
        o, P, G = self.ops, self.store.P, self.store.G
        o.gemm(x2, dy2, G[kernel_name], ta=True)
        if bias_name is not None:
            o.colsum(dy2, G[bias_name])
        if dx2 is not None:
            o.gemm(dy2, P[kernel_name], dx2, tb=True, beta=beta)

    # ------------------------------------------------------------------ forward
    def _check_shapes(self, ids, lens, mel_targets, stop_targets):
        B = ids.shape[0]
        if mel_targets.dim() != 3 or mel_targets.shape[0] != B or mel_targets.shape[2] != self.hp.num_mels:
            raise ValueError("mel_targets must be [B=%d, F, %d], got %s" % (B, self.hp.num_mels, tuple(mel_targets.shape)))
        if mel_targets.shape[1] % self.hp.outputs_per_step != 0:
            raise ValueError("mel_targets frames (%d) must be a multiple of outputs_per_step" % mel_targets.shape[1])
        if tuple(stop_targets.shape) != (B, mel_targets.shape[1]):
            raise ValueError("stop_token_targets must be [B, F] = %s (one per mel frame), got %s"
                             % ((B, mel_targets.shape[1]), tuple(stop_targets.shape)))
        if ids.dim() != 2 or tuple(lens.shape) != (B,):
            raise ValueError("inputs must be [B, T] with input_lengths [B], got %s / %s" % (tuple(ids.shape), tuple(lens.shape)))

    def forward(self, ids, lens, mel_targets, stop_targets, seed):
        """ids int32 [B,T], lens int32 [B], mel_targets fp32 [B,F,80], stop_targets fp32 [B,F] (device
        tensors).  Returns the tower outputs + losses of ``initialize(is_training=True)`` / ``add_loss``."""
        hp, o, P = self.hp, self.ops, self.store.P
        self._check_shapes(ids, lens, mel_targets, stop_targets)
        B, T = ids.shape
        F = mel_targets.shape[1]
        r, M = hp.outputs_per_step, hp.num_mels
        assert F % r == 0, "mel targets are padded to a multiple of outputs_per_step (Tacotron_feeder.py:102-103)"
        S = F // r
        H = hp.enc_lstm_hidden_size
        D = hp.decoder_lstm_units
        A = hp.attention_dim
        zo = hp.zoneout_rate
        keep = 1.0 - hp.dropout_rate
        t = dict(B=B, T=T, F=F, S=S, seed=seed, ids=ids, lens=lens, mel=mel_targets, stop_tgt=stop_targets)

        self._mark("fwd start")
        # ---- embedding + encoder convolutions (Tacotron.py:179-181; Encoder.py:40-46)
        emb = o.empty(B, T, hp.embedding_dim)
        o.embedding_fwd(ids, P["inputs_embedding"], emb)
        t["enc_conv"] = []
        x = emb
        for i in range(hp.conv_num_layers):
            x = self._conv_block_fwd(x, W.ENC_CONV % (i + 1), ACT_RELU, keep, seed, STREAM_ENC_CONV + i, t["enc_conv"])
        C = x.shape[-1]
        x3 = x.view(B * T, C)

        self._mark("fwd encoder convs done")
        # ---- bidirectional zoneout LSTM over the valid length (Encoder.py:68-77,118-142)
        enc_out = o.zeros(B, T, 2 * H)
        t["bilstm"] = []
        for d, (scope, rev) in enumerate(((W.ENC_LSTM_FW, False), (W.ENC_LSTM_BW, True))):
            K = P[scope + "/kernel"]
            xw = o.empty(B, T, 4 * H)
            o.gemm(x3, K[:C], xw.view(B * T, 4 * H), bias=P[scope + "/bias"])
            acts, cnew, hraw = o.empty(T, B, 4 * H), o.empty(T, B, H), o.empty(T, B, H)
            cz, hz = o.zeros(T + 1, B, H), o.zeros(T + 1, B, H)
            pre = o.empty(B, 4 * H)
            sc, sh = STREAM_ENC_LSTM + 2 * d, STREAM_ENC_LSTM + 2 * d + 1
            fast = hasattr(o, "lstm_loop")         # the same loop sequenced in C (train_loops.cu)
            with self._side(d):                    # the two directions are independent chains: one side stream each
                if fast:
                    o.lstm_loop(False, dict(B=B, T=T, H=H, C=C, reverse=int(rev)), zo, seed, sc, sh, enc_out[:, :, d * H:(d + 1) * H],
                                K=K, lens=lens, xw=xw, acts=acts, cnew=cnew, cz=cz, hz=hz, work=o.empty(B * 9 * H))
                for s in range(0 if not fast else T, T):
                    o.gather_rows(xw, lens, s, rev, pre)
                    o.gemm(hz[s], K[C:], acts[s])
                    o.lstm_fwd(acts[s], pre, None, cz[s], hz[s], lens, s, zo, True, seed, sc, sh, s * B,
                               cnew[s], cz[s + 1], hraw[s], hz[s + 1])
                    o.scatter_rows(hraw[s], lens, s, rev, enc_out[:, :, d * H:(d + 1) * H])
            if getattr(self, "marks", None) is not None:
                with self._side(d):
                    self._mark("  [side] BiLSTM direction %d done" % d)
            t["bilstm"].append(dict(scope=scope, rev=rev, acts=acts, cnew=cnew, cz=cz, hz=hz, sc=sc, sh=sh))
        t["x3"] = x3

        # ---- reference encoder on the mel targets (Tacotron.py:12-29,191-193) -> GST attention (Attention.py:243-341)
        ref_side = self._side(2)                   # independent of the text encoder: its own side stream
        ref_side.__enter__()
        t["ref_conv"] = []
        xr = mel_targets.view(B, F, M, 1)
        for i in range(len(hp.reference_filters)):
            xr = self._conv_block_fwd(xr, W.REF_CONV % i, ACT_RELU, 1.0, seed, 0, t["ref_conv"], conv2d=True)
            self._mark("  [side] ref conv layer done")
        _, Tp, Wp, Cp = xr.shape
        In, R = Wp * Cp, hp.reference_depth
        Xt = o.empty(Tp, B, In)
        Xt.copy_(xr.view(B, Tp, In).transpose(0, 1))
        Wg, Wc = P[W.REF_GRU + "/gates/kernel"], P[W.REF_GRU + "/candidate/kernel"]
        GX, CX = o.empty(Tp, B, 2 * R), o.empty(Tp, B, R)
        o.gemm(Xt.view(Tp * B, In), Wg[:In], GX.view(Tp * B, 2 * R), bias=P[W.REF_GRU + "/gates/bias"])
        o.gemm(Xt.view(Tp * B, In), Wc[:In], CX.view(Tp * B, R), bias=P[W.REF_GRU + "/candidate/bias"])
        hg, rh = o.zeros(Tp + 1, B, R), o.empty(Tp, B, R)
        gru_fast = hasattr(o, "gru_loop")         # the same loop sequenced in C (train_loops.cu)
        if gru_fast:
            o.gru_loop(False, dict(B=B, Tp=Tp, In=In, R=R), Wg=Wg, Wc=Wc, GX=GX, CX=CX, hg=hg, rh=rh)
        for s in range(0 if not gru_fast else Tp, Tp):   # GRUCell: reset applied before the candidate matmul (App. A.9)
            o.gemm(hg[s], Wg[In:], GX[s], beta=1.0)
            o.gru_pre(GX[s], hg[s], rh[s])        # GX[s] <- (r | u)
            o.gemm(rh[s], Wc[In:], CX[s], beta=1.0)
            o.gru_out(CX[s], GX[s], hg[s], hg[s + 1])   # CX[s] <- candidate c
        self._mark("  [side] ref GRU done")
        refz, ref = o.empty(B, 128), o.empty(B, 128)
        o.gemm(hg[Tp], P[W.REF_DENSE + "/kernel"], refz, bias=P[W.REF_DENSE + "/bias"])
        o.ew("tanh", ref, refz)
        t.update(Xt=Xt, GX=GX, CX=CX, hg=hg, rh=rh, ref=ref, Tp=Tp, In=In)
        nh = hp.num_heads
        tok = P["style_tokens"]
        tt = o.empty(*tok.shape)
        o.ew("tanh", tt, tok)
        U = hp.style_att_dim
        gq, gk = o.empty(B, U), o.empty(tok.shape[0], U)
        o.gemm(ref, P[W.MHA + "/conv1d/kernel"][0], gq, bias=P[W.MHA + "/conv1d/bias"])
        o.gemm(tt, P[W.MHA + "/conv1d_1/kernel"][0], gk, bias=P[W.MHA + "/conv1d_1/bias"])
        wts, style = o.empty(B, nh, tok.shape[0]), o.empty(B, nh * tok.shape[1])
        o.gst_fwd(gq, gk, P[W.MHA + "/attention_v"], P[W.MHA + "/attention_g"], P[W.MHA + "/attention_b"], tt, nh, wts, style)
        t.update(tt=tt, gq=gq, gk=gk, wts=wts)
        self._mark("  [side] GST done")
        ref_side.__exit__(None, None, None)
        self._join()                               # both LSTM directions and the style embedding are ready
        SD = style.shape[1]
        MD = 2 * H + SD
        memory = o.empty(B, T, MD)
        memory[:, :, :2 * H].copy_(enc_out)
        o.tile_style(style, memory, 2 * H)
        t["memory"] = memory

        self._mark("fwd BiLSTM + reference encoder joined")
        # ---- attention memory (Attention.py:153-158): masked values, keys computed once
        values = o.empty(B, T, MD)
        values.copy_(memory)
        lens_eff = lens
        if hp.mask_encoder:
            o.mask_rows(values, lens)
        else:
            lens_eff = o.to_device(np.full((B,), T, np.int32))
        keys = o.empty(B, T, A)
        o.gemm(values.view(B * T, MD), P[W.MEMORY_LAYER + "/kernel"], keys.view(B * T, A))
        t.update(values=values, keys=keys, lens_eff=lens_eff, MD=MD)

        self._mark("fwd memory / keys done")
        # ---- decoder, teacher-forced (Decoder.py:246-292; Helpers.py:62-128; ratio 1.0)
        Xin = o.zeros(S, B, M)                                       # <GO> frame, then targets[:, r*t - 1]
        if S > 1:
            Xin[1:].copy_(mel_targets[:, r - 1::r][:, :S - 1].transpose(0, 1))
        U1, U2 = hp.prenet_layers
        Z1, P1, Z2, P2 = o.empty(S, B, U1), o.empty(S, B, U1), o.empty(S, B, U2), o.empty(S, B, U2)
        o.gemm(Xin.view(S * B, M), P[W.PRENET % 1 + "/kernel"], Z1.view(S * B, U1), bias=P[W.PRENET % 1 + "/bias"])
        o.prenet_act_fwd(Z1, 0, seed, 0, keep, P1)
        o.gemm(P1.view(S * B, U1), P[W.PRENET % 2 + "/kernel"], Z2.view(S * B, U2), bias=P[W.PRENET % 2 + "/bias"])
        o.prenet_act_fwd(Z2, 1, seed, 0, keep, P2)
        W1, b1 = P[W.DEC_LSTM % (0, 1) + "/kernel"], P[W.DEC_LSTM % (0, 1) + "/bias"]
        W2, b2 = P[W.DEC_LSTM % (1, 2) + "/kernel"], P[W.DEC_LSTM % (1, 2) + "/bias"]
        Wq = P[W.LSA + "/query_layer/kernel"]
        G1x = o.empty(S, B, 4 * D)
        o.gemm(P2.view(S * B, U2), W1[:U2], G1x.view(S * B, 4 * D), bias=b1)     # prenet term of the LSTM-1 gates, all steps
        att_w = (P[W.LSA + "/location_features_convolution/kernel"], P[W.LSA + "/location_features_convolution/bias"],
                 P[W.LSA + "/location_features_layer/kernel"], P[W.LSA + "/attention_variable"], P[W.LSA + "/attention_bias"])
        acts1, acts2 = o.empty(S, B, 4 * D), o.empty(S, B, 4 * D)
        c1new, c2new, h1raw, h2raw = (o.empty(S, B, D) for _ in range(4))
        c1z, h1z, c2z, h2z = (o.zeros(S + 1, B, D) for _ in range(4))
        ctx, cum = o.zeros(S + 1, B, MD), o.zeros(S + 1, B, T)
        align, q = o.empty(S, B, T), o.empty(S, B, A)
        s1, s2 = STREAM_DEC_LSTM, STREAM_DEC_LSTM + 2
        dims = dict(B=B, T=T, S=S, D=D, MD=MD, A=A, U2=U2, NF=att_w[0].shape[2], KW=att_w[0].shape[0])
        fast = hasattr(o, "decoder_loop")           # the same loop sequenced in C (train_loops.cu)
        th = attn_work = None
        if fast and o.attn_fused_supported(T, D, A, MD, dims["NF"], dims["KW"]):
            # fused attention of the loops (csrc/train_attn_fused.cu): tanh(pre) of every step is kept for the backward pass
            th, attn_work = o.empty(S, B, T, A), o.empty((dims["KW"] + 1) * A + B * (dims["KW"] + 2) * A)
        if fast:
            o.decoder_loop(False, dims, zo, seed, s1, s2, th=th, attn_work=attn_work, W1=W1, W2=W2, b2=b2, Wq=Wq, conv_w=att_w[0], conv_b=att_w[1], Wl=att_w[2],
                           va=att_w[3], ba=att_w[4], keys=keys, values=values, lens=lens_eff, G1x=G1x, acts1=acts1, acts2=acts2,
                           c1new=c1new, c2new=c2new, h1raw=h1raw, h2raw=h2raw, c1z=c1z, h1z=h1z, c2z=c2z, h2z=h2z, ctx=ctx, cum=cum,
                           align=align, q=q)
        for s in range(0 if not fast else S, S):
            o.gemm(ctx[s], W1[U2:U2 + MD], acts1[s])
            o.gemm(h1z[s], W1[U2 + MD:], acts1[s], beta=1.0)
            o.lstm_fwd(acts1[s], G1x[s], None, c1z[s], h1z[s], None, 0, zo, True, seed, s1, s1 + 1, s * B,
                       c1new[s], c1z[s + 1], h1raw[s], h1z[s + 1])
            o.gemm(h1raw[s], W2[:D], acts2[s])
            o.gemm(h2z[s], W2[D:], acts2[s], beta=1.0)
            o.lstm_fwd(acts2[s], None, b2, c2z[s], h2z[s], None, 0, zo, True, seed, s2, s2 + 1, s * B,
                       c2new[s], c2z[s + 1], h2raw[s], h2z[s + 1])
            o.gemm(h2raw[s], Wq, q[s])
            o.attn_fwd(q[s], cum[s], keys, values, lens_eff, *att_w, align[s], cum[s + 1], ctx[s + 1])
        t.update(th=th, attn_work=attn_work, dec_dims=dims, G1x=G1x, Xin=Xin, P1=P1, P2=P2, acts1=acts1, acts2=acts2, c1new=c1new, c2new=c2new, h1raw=h1raw, h2raw=h2raw,
                 c1z=c1z, h1z=h1z, c2z=c2z, h2z=h2z, ctx=ctx, cum=cum, align=align, q=q)
        # frame / stop projections on [LSTM_output, context] for all steps at once (Decoder.py:355-410)
        Wf, Ws = P[W.FRAME_PROJ + "/kernel"], P[W.STOP_PROJ + "/kernel"]
        Y, L = o.empty(S, B, M * r), o.empty(S, B, r)
        h2all, ctxall = h2raw.view(S * B, D), ctx[1:].view(S * B, MD)
        o.gemm(h2all, Wf[:D], Y.view(S * B, M * r), bias=P[W.FRAME_PROJ + "/bias"])
        o.gemm(ctxall, Wf[D:], Y.view(S * B, M * r), beta=1.0)
        o.gemm(h2all, Ws[:D], L.view(S * B, r), bias=P[W.STOP_PROJ + "/bias"])
        o.gemm(ctxall, Ws[D:], L.view(S * B, r), beta=1.0)
        frames = o.empty(B, S, M * r)
        frames.copy_(Y.transpose(0, 1))
        frames = frames.view(B, F, M)                                # Tacotron.py:280
        stop_logits = o.empty(B, S, r)
        stop_logits.copy_(L.transpose(0, 1))
        stop_logits = stop_logits.view(B, F)                         # :281 (logits: StopProjection skips the sigmoid in training)
        dec_out = o.empty(B, F, M)
        o.clip_fwd(frames, self.clip[0], self.clip[1], dec_out)      # :282-284
        t.update(frames=frames, stop_logits=stop_logits, dec_out=dec_out)

        self._mark("fwd decoder (prenet, loop, projections) done")
        # ---- postnet + projection + residual + clip (Postnet.py:28-36; Tacotron.py:287-309)
        t["postnet"] = []
        x = dec_out
        n = hp.postnet_num_layers
        for i in range(n):
            x = self._conv_block_fwd(x, W.POST_CONV % (i + 1), ACT_TANH if i < n - 1 else ACT_NONE, keep, seed,
                                     STREAM_POSTNET + i, t["postnet"])
        PC = x.shape[-1]
        pre_mel = o.empty(B, F, M)
        o.gemm(x.view(B * F, PC), P[W.POST_PROJ + "/kernel"], pre_mel.view(B * F, M), bias=P[W.POST_PROJ + "/bias"])
        o.ew("add", pre_mel, pre_mel, dec_out)
        mel_out = o.empty(B, F, M)
        o.clip_fwd(pre_mel, self.clip[0], self.clip[1], mel_out)
        t.update(post_x=x, pre_mel=pre_mel, mel_out=mel_out)

        self._mark("fwd postnet done")
        # ---- losses (Tacotron.py:388-396,419-446; mask_decoder=False branch)
        if hp.mask_decoder:
            raise NotImplementedError("mask_decoder=True (MaskedMSE, Tacotron.py:373-386) is not the default branch")
        ls = o.zeros(5)                                              # before, after, stop, regularization, total
        t["d_dec_b"], t["d_mel"], t["d_stop"] = o.empty(B, F, M), o.empty(B, F, M), o.empty(B, F)
        o.mse(dec_out, mel_targets, ls[0:1], t["d_dec_b"])
        o.mse(mel_out, mel_targets, ls[1:2], t["d_mel"])
        o.sigmoid_ce(stop_logits, stop_targets, ls[2:3], t["d_stop"])
        o.reg_loss(self.store.p, self.store.regw, ls[3:4])
        self.tape = t
        self.losses = ls
        return dict(decoder_output=dec_out, mel_outputs=mel_out, stop_token_prediction=stop_logits, alignments=align,
                    encoder_outputs=memory, losses=ls)

    # ------------------------------------------------------------------ backward
    def backward(self, tape=None):
        """Hand-derived gradient of loss = before + after + stop (+ regularization, added in apply()) with
        respect to every trainable variable -> the flat gradient buffer ``store.g`` (overwritten).
        ``tape``: the record of the forward pass to differentiate (default: the most recent one)."""
        hp, o, P, G = self.hp, self.ops, self.store.P, self.store.G
        t = self.tape if tape is None else tape
        B, T, F, S, seed, MD = t["B"], t["T"], t["F"], t["S"], t["seed"], t["MD"]
        r, M = hp.outputs_per_step, hp.num_mels
        H, D, A = hp.enc_lstm_hidden_size, hp.decoder_lstm_units, hp.attention_dim
        zo = hp.zoneout_rate
        keep = 1.0 - hp.dropout_rate
        self.store.g.zero_()

        self._mark("bwd start")
        # ---- postnet
        d_pre = o.empty(B, F, M)
        o.clip_bwd(t["d_mel"], t["pre_mel"], self.clip[0], self.clip[1], d_pre)
        d_dec = o.empty(B, F, M)
        o.ew("add", d_dec, t["d_dec_b"], d_pre)                       # MSE(before) + residual branch
        px = t["post_x"]
        PC = px.shape[-1]
        dy = o.empty(B, F, PC)
        self._dense_bwd(px.view(B * F, PC), d_pre.view(B * F, M), W.POST_PROJ + "/kernel", W.POST_PROJ + "/bias", dy.view(B * F, PC))
        for rec in reversed(t["postnet"]):
            dy = self._conv_block_bwd(dy, rec, seed)
        o.ew("add", d_dec, d_dec, dy)
        d_frames = o.empty(B, F, M)
        o.clip_bwd(d_dec, t["frames"], self.clip[0], self.clip[1], d_frames)

        self._mark("bwd postnet done")
        # ---- frame / stop projections (all steps at once)
        dY, dL = o.empty(S, B, M * r), o.empty(S, B, r)
        dY.copy_(d_frames.view(B, S, M * r).transpose(0, 1))
        dL.copy_(t["d_stop"].view(B, S, r).transpose(0, 1))
        dY2, dL2 = dY.view(S * B, M * r), dL.view(S * B, r)
        h2all, ctxall = t["h2raw"].view(S * B, D), t["ctx"][1:].view(S * B, MD)
        Wf, Ws = P[W.FRAME_PROJ + "/kernel"], P[W.STOP_PROJ + "/kernel"]
        GWf, GWs = G[W.FRAME_PROJ + "/kernel"], G[W.STOP_PROJ + "/kernel"]
        o.gemm(h2all, dY2, GWf[:D], ta=True)
        o.gemm(ctxall, dY2, GWf[D:], ta=True)
        o.gemm(h2all, dL2, GWs[:D], ta=True)
        o.gemm(ctxall, dL2, GWs[D:], ta=True)
        o.colsum(dY2, G[W.FRAME_PROJ + "/bias"])
        o.colsum(dL2, G[W.STOP_PROJ + "/bias"])
        dH2, dCtx = o.empty(S, B, D), o.empty(S, B, MD)               # d(h2raw_t), d(ctx_t) from the projections
        o.gemm(dY2, Wf[:D], dH2.view(S * B, D), tb=True)
        o.gemm(dL2, Ws[:D], dH2.view(S * B, D), tb=True, beta=1.0)
        o.gemm(dY2, Wf[D:], dCtx.view(S * B, MD), tb=True)
        o.gemm(dL2, Ws[D:], dCtx.view(S * B, MD), tb=True, beta=1.0)

        self._mark("bwd projections done")
        # ---- decoder BPTT
        U1, U2 = hp.prenet_layers
        W1, W2 = P[W.DEC_LSTM % (0, 1) + "/kernel"], P[W.DEC_LSTM % (1, 2) + "/kernel"]
        Wq = P[W.LSA + "/query_layer/kernel"]
        att_names = (W.LSA + "/location_features_convolution/kernel", W.LSA + "/location_features_convolution/bias",
                     W.LSA + "/location_features_layer/kernel", W.LSA + "/attention_variable", W.LSA + "/attention_bias")
        att_w = tuple(P[n] for n in att_names)
        parts = tuple(o.zeros(B, *P[n].shape) for n in att_names)     # per-row partial sums of the attention weight grads
        keys, values, lens_eff = t["keys"], t["values"], t["lens_eff"]
        dkeys = o.zeros(B, T, A)
        dG1, dG2, dQ = o.empty(S, B, 4 * D), o.empty(S, B, 4 * D), o.empty(S, B, A)
        dc1, dh1z, dc2, dh2z = (o.zeros(B, D) for _ in range(4))
        dc1n, dh1zn, dc2n, dh2zn = (o.empty(B, D) for _ in range(4))
        dh1raw = o.empty(B, D)
        dctx_carry = o.zeros(B, MD)
        dcum, dcum_n = o.zeros(B, T), o.empty(B, T)
        s1, s2 = STREAM_DEC_LSTM, STREAM_DEC_LSTM + 2
        fast = hasattr(o, "decoder_loop")
        if fast:
            o.decoder_loop(True, t["dec_dims"], zo, seed, s1, s2, th=t["th"], attn_work=t["attn_work"],
                           de=o.empty(S, B, T) if t["th"] is not None else None, W1=W1, W2=W2, Wq=Wq, conv_w=att_w[0], conv_b=att_w[1], Wl=att_w[2],
                           va=att_w[3], ba=att_w[4], keys=keys, values=values, lens=lens_eff, acts1=t["acts1"], acts2=t["acts2"],
                           c1new=t["c1new"], c2new=t["c2new"], c1z=t["c1z"], h1z=t["h1z"], c2z=t["c2z"], h2z=t["h2z"], cum=t["cum"],
                           align=t["align"], q=t["q"], dH2=dH2, dCtx=dCtx, dG1=dG1, dG2=dG2, dQ=dQ, dkeys=dkeys, p_conv_w=parts[0],
                           p_conv_b=parts[1], p_Wl=parts[2], p_va=parts[3], p_ba=parts[4], work=o.empty(B * (8 * D + 2 * T + T * t["dec_dims"]["NF"]) + S * B * D + (S + 1) * B * MD))
        for s in range(S - 1 if not fast else -1, -1, -1):
            o.ew("add", dCtx[s], dCtx[s], dctx_carry)                 # total d(ctx_t); kept for d(values)
            o.attn_bwd(dCtx[s], dcum, t["q"][s], t["cum"][s], t["align"][s], keys, values, lens_eff, *att_w,
                       dQ[s], dcum_n, dkeys, *parts)
            dcum, dcum_n = dcum_n, dcum
            o.gemm(dQ[s], Wq, dH2[s], tb=True, beta=1.0)              # d(h2raw_t) += dq Wq^T
            o.lstm_bwd(t["acts2"][s], t["c2z"][s], t["h2z"][s], t["c2new"][s], None, 0, zo, True, seed, s2, s2 + 1, s * B,
                       dH2[s], dc2, dh2z, dG2[s], dc2n, dh2zn)
            o.gemm(dG2[s], W2[:D], dh1raw, tb=True)
            o.gemm(dG2[s], W2[D:], dh2zn, tb=True, beta=1.0)
            o.lstm_bwd(t["acts1"][s], t["c1z"][s], t["h1z"][s], t["c1new"][s], None, 0, zo, True, seed, s1, s1 + 1, s * B,
                       dh1raw, dc1, dh1z, dG1[s], dc1n, dh1zn)
            o.gemm(dG1[s], W1[U2:U2 + MD], dctx_carry, tb=True)
            o.gemm(dG1[s], W1[U2 + MD:], dh1zn, tb=True, beta=1.0)
            dc1, dc1n, dh1z, dh1zn = dc1n, dc1, dh1zn, dh1z
            dc2, dc2n, dh2z, dh2zn = dc2n, dc2, dh2zn, dh2z
        # d(values)[b] = align[:, b, :]^T . d(ctx)[:, b, :]; keys = values . Wmem
        dvalues = o.empty(B, T, MD)
        if hasattr(o, "attn_dvalues") and MD % 64 == 0:
            o.attn_dvalues(t["align"], dCtx, dvalues)              # all rows in one launch
        else:
            for b in range(B):
                o.gemm(t["align"][:, b, :], dCtx[:, b, :], dvalues[b], ta=True)
        self._dense_bwd(values.view(B * T, MD), dkeys.view(B * T, A), W.MEMORY_LAYER + "/kernel", None,
                        dvalues.view(B * T, MD), beta=1.0)
        if hp.mask_encoder:
            o.mask_rows(dvalues, t["lens"])
        dmemory = dvalues

        self._mark("bwd decoder BPTT + d(values) done")
        # ---- style embedding: GST attention -> reference encoder (Attention.py:243-341; Tacotron.py:12-29)
        nh = hp.num_heads
        tt, ref, Tp, In = t["tt"], t["ref"], t["Tp"], t["In"]
        R = hp.reference_depth
        ref_side = self._side(2)                   # reference-encoder gradient chain beside the BiLSTM BPTT
        ref_side.__enter__()
        dstyle = o.empty(B, MD - 2 * H)
        o.sum_style(dmemory, 2 * H, dstyle)
        dq, dk, dtt = o.empty(*t["gq"].shape), o.empty(*t["gk"].shape), o.empty(*tt.shape)
        o.gst_bwd(dstyle, t["gq"], t["gk"], P[W.MHA + "/attention_v"], P[W.MHA + "/attention_g"], P[W.MHA + "/attention_b"],
                  tt, t["wts"], nh, dq, dk, G[W.MHA + "/attention_v"], G[W.MHA + "/attention_g"], G[W.MHA + "/attention_b"], dtt)
        dref = o.empty(B, 128)
        o.gemm(ref, dq, G[W.MHA + "/conv1d/kernel"][0], ta=True)
        o.colsum(dq, G[W.MHA + "/conv1d/bias"])
        o.gemm(dq, P[W.MHA + "/conv1d/kernel"][0], dref, tb=True)
        o.gemm(tt, dk, G[W.MHA + "/conv1d_1/kernel"][0], ta=True)
        o.colsum(dk, G[W.MHA + "/conv1d_1/bias"])
        o.gemm(dk, P[W.MHA + "/conv1d_1/kernel"][0], dtt, tb=True, beta=1.0)
        o.ew("tanh_bwd", G["style_tokens"], dtt, tt)
        drefz = o.empty(B, 128)
        o.ew("tanh_bwd", drefz, dref, ref)
        hg, GX, CX, rh, Xt = t["hg"], t["GX"], t["CX"], t["rh"], t["Xt"]
        dh, dhd, drh = o.empty(B, R), o.empty(B, R), o.empty(B, R)
        self._dense_bwd(hg[Tp], drefz, W.REF_DENSE + "/kernel", W.REF_DENSE + "/bias", dh)
        Wg, Wc = P[W.REF_GRU + "/gates/kernel"], P[W.REF_GRU + "/candidate/kernel"]
        dGX, dCX = o.empty(Tp, B, 2 * R), o.empty(Tp, B, R)
        gru_fast = hasattr(o, "gru_loop")
        if gru_fast:
            o.gru_loop(True, dict(B=B, Tp=Tp, In=In, R=R), Wg=Wg, Wc=Wc, GX=GX, CX=CX, hg=hg, rh=rh, dGX=dGX, dCX=dCX, dh=dh,
                       work=o.empty(2 * B * R))
        for s in range(Tp - 1 if not gru_fast else -1, -1, -1):
            o.gru_out_bwd(dh, GX[s], hg[s], CX[s], dCX[s], dGX[s], dhd)
            o.gemm(dCX[s], Wc[In:], drh, tb=True)
            o.gru_pre_bwd(drh, GX[s], hg[s], dGX[s], dhd)
            o.gemm(dGX[s], Wg[In:], dhd, tb=True, beta=1.0)
            dh, dhd = dhd, dh
        dGX2, dCX2, Xt2 = dGX.view(Tp * B, 2 * R), dCX.view(Tp * B, R), Xt.view(Tp * B, In)
        GWg, GWc = G[W.REF_GRU + "/gates/kernel"], G[W.REF_GRU + "/candidate/kernel"]
        o.gemm(Xt2, dGX2, GWg[:In], ta=True)
        o.gemm(hg[:Tp].view(Tp * B, R), dGX2, GWg[In:], ta=True)
        o.colsum(dGX2, G[W.REF_GRU + "/gates/bias"])
        o.gemm(Xt2, dCX2, GWc[:In], ta=True)
        o.gemm(rh.view(Tp * B, R), dCX2, GWc[In:], ta=True)
        o.colsum(dCX2, G[W.REF_GRU + "/candidate/bias"])
        dXt = o.empty(Tp, B, In)
        o.gemm(dGX2, Wg[:In], dXt.view(Tp * B, In), tb=True)
        o.gemm(dCX2, Wc[:In], dXt.view(Tp * B, In), tb=True, beta=1.0)
        last = t["ref_conv"][-1]["y"]
        dy = o.empty(*last.shape)
        dy.view(B, Tp, In).copy_(dXt.transpose(0, 1))
        for i in range(len(t["ref_conv"]) - 1, -1, -1):
            dy = self._conv_block_bwd(dy, t["ref_conv"][i], seed, need_dx=i > 0)
        ref_side.__exit__(None, None, None)

        # ---- encoder BiLSTM BPTT over the valid lengths
        lens = t["lens"]
        x3 = t["x3"]
        C = x3.shape[1]
        dx3 = o.empty(B * T, C)
        dxws = []
        for d, rec in enumerate(t["bilstm"]):
            K = P[rec["scope"] + "/kernel"]
            GK = G[rec["scope"] + "/kernel"]
            src = dmemory[:, :, d * H:(d + 1) * H]
            with self._side(d):                    # the two directions are independent chains: one side stream each
                dxw = o.zeros(B, T, 4 * H)
                dg = o.empty(T, B, 4 * H)
                dc, dhz, dcn, dhn = o.zeros(B, H), o.zeros(B, H), o.empty(B, H), o.empty(B, H)
                d_hraw = o.empty(B, H)
                fast = hasattr(o, "lstm_loop")
                if fast:
                    o.lstm_loop(True, dict(B=B, T=T, H=H, C=C, reverse=int(rec["rev"])), zo, seed, rec["sc"], rec["sh"], src, K=K, lens=lens,
                                acts=rec["acts"], cnew=rec["cnew"], cz=rec["cz"], hz=rec["hz"], dxw=dxw, dg=dg, work=o.empty(B * 9 * H))
                for s in range(T - 1 if not fast else -1, -1, -1):
                    o.gather_rows(src, lens, s, rec["rev"], d_hraw)
                    o.lstm_bwd(rec["acts"][s], rec["cz"][s], rec["hz"][s], rec["cnew"][s], lens, s, zo, True, seed, rec["sc"], rec["sh"],
                               s * B, d_hraw, dc, dhz, dg[s], dcn, dhn)
                    o.gemm(dg[s], K[C:], dhn, tb=True, beta=1.0)
                    o.scatter_rows(dg[s], lens, s, rec["rev"], dxw)
                    dc, dcn, dhz, dhn = dcn, dc, dhn, dhz
                o.gemm(rec["hz"][:T].view(T * B, H), dg.view(T * B, 4 * H), GK[C:], ta=True)
                dxw2 = dxw.view(B * T, 4 * H)
                o.gemm(x3, dxw2, GK[:C], ta=True)
                o.colsum(dxw2, G[rec["scope"] + "/bias"])
            dxws.append((dxw2, K))
        self._mark("bwd side chains forked")
        # ---- decoder weight gradients and the prenet, all steps at once: on the main stream while the three encoder-side
        # chains above run on their side streams
        dG1_2, dG2_2, dQ2 = dG1.view(S * B, 4 * D), dG2.view(S * B, 4 * D), dQ.view(S * B, A)
        GW1, GW2 = G[W.DEC_LSTM % (0, 1) + "/kernel"], G[W.DEC_LSTM % (1, 2) + "/kernel"]
        o.gemm(t["P2"].view(S * B, U2), dG1_2, GW1[:U2], ta=True)
        o.gemm(t["ctx"][:S].view(S * B, MD), dG1_2, GW1[U2:U2 + MD], ta=True)
        o.gemm(t["h1z"][:S].view(S * B, D), dG1_2, GW1[U2 + MD:], ta=True)
        o.colsum(dG1_2, G[W.DEC_LSTM % (0, 1) + "/bias"])
        o.gemm(t["h1raw"].view(S * B, D), dG2_2, GW2[:D], ta=True)
        o.gemm(t["h2z"][:S].view(S * B, D), dG2_2, GW2[D:], ta=True)
        o.colsum(dG2_2, G[W.DEC_LSTM % (1, 2) + "/bias"])
        o.gemm(h2all, dQ2, G[W.LSA + "/query_layer/kernel"], ta=True)
        for n, part in zip(att_names, parts):
            o.colsum(part.view(B, -1), G[n].view(-1))
        # prenet (hoisted: all steps at once)
        dP2, dZ2, dP1, dZ1 = o.empty(S, B, U2), o.empty(S, B, U2), o.empty(S, B, U1), o.empty(S, B, U1)
        o.gemm(dG1_2, W1[:U2], dP2.view(S * B, U2), tb=True)
        o.prenet_act_bwd(dP2, t["P2"], 1, seed, 0, keep, dZ2)
        self._dense_bwd(t["P1"].view(S * B, U1), dZ2.view(S * B, U2), W.PRENET % 2 + "/kernel", W.PRENET % 2 + "/bias",
                        dP1.view(S * B, U1))
        o.prenet_act_bwd(dP1, t["P1"], 0, seed, 0, keep, dZ1)
        self._dense_bwd(t["Xin"].view(S * B, M), dZ1.view(S * B, U1), W.PRENET % 1 + "/kernel", W.PRENET % 1 + "/bias")
        # every decoder / postnet gradient (store.g[tail_start:]) is final from here on: the data-parallel trainer starts
        # the exchange of that half now, under the rest of the encoder backward
        if self.on_tail_ready is not None:
            self.on_tail_ready()
        self._join()                               # both directions and the reference-encoder chain
        for d, (dxw2, K) in enumerate(dxws):
            o.gemm(dxw2, K[:C], dx3, tb=True, beta=0.0 if d == 0 else 1.0)

        self._mark("bwd weight gradients done, side chains joined")
        # ---- encoder convolutions + embedding
        dy = dx3.view(B, T, C)
        for rec in reversed(t["enc_conv"]):
            dy = self._conv_block_bwd(dy, rec, seed)
        o.embedding_bwd(t["ids"], dy, G["inputs_embedding"])
        return self.store.g

    # ------------------------------------------------------------------ optimizer (Tacotron.py:448-540)
    def learning_rate(self, global_step):
        """_learning_rate_decay (Tacotron.py:518-540): exponential decay after start_decay, clipped to [final, init]."""
        hp = self.hp
        if not hp.tacotron_decay_learning_rate:
            return hp.tacotron_initial_learning_rate
        lr = hp.tacotron_initial_learning_rate * hp.tacotron_decay_rate ** (
            (global_step - hp.tacotron_start_decay) / float(hp.tacotron_decay_steps))
        return min(max(lr, hp.tacotron_final_learning_rate), hp.tacotron_initial_learning_rate)

    def apply(self, global_step, world=1, all_reduce=None, ranks=None):
        """Average the tower gradients (one all-reduce of the flat buffer replaces Tacotron.py:484-501), add the
        L2 term, clip by global norm 1.0 (:507), Adam update (:470-471; TF formulation, eps outside the bias
        correction).  ``global_step`` counts completed updates (0 for the first)."""
        hp, o, st = self.hp, self.ops, self.store
        if all_reduce is not None and world > 1:
            # SUM over ranks of the ONE flat bucket [BN moving statistics | gradients]; a trainer that already started the
            # early half (decoder + postnet gradients) passes a reducer that finishes the rest and joins both
            all_reduce(st.bucket)
            if st.ns:
                # mean over the RANKS that were summed (``world`` also counts the in-process towers of one rank)
                o.ew("scale", st.s, st.s, None, 1.0 / (world if ranks is None else ranks))
        if self.update_mask is not None:
            o.ew("mul", st.g, st.g, self.update_mask)                # tacotron_fine_tuning (Tacotron.py:478-480)
        sumsq = o.zeros(1)
        o.grad_finalize(st.g, st.p, self.regw_grad, 1.0 / world, sumsq)
        tstep = global_step + 1
        b1, b2 = hp.tacotron_adam_beta1, hp.tacotron_adam_beta2
        lr_t = self.learning_rate(global_step) * np.sqrt(1.0 - b2 ** tstep) / (1.0 - b1 ** tstep)
        clip = 1.0 if hp.tacotron_clip_gradients else 0.0
        o.adam_step(st.p, st.g, st.m, st.v, sumsq, clip, float(lr_t), b1, b2, hp.tacotron_adam_epsilon)
        return sumsq
