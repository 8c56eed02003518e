"""TEST INFRASTRUCTURE: torch-CPU twin of the training primitives of libtaco_b200.so (include/taco_b200_train.h).

The training step (tacotron_prosody_multi_gpu_b200/train/step.py) is a host-side composition of
stateless device primitives.  This twin implements every primitive with plain torch-CPU arithmetic
so that (i) the composition -- the hand-derived backward pass -- can be checked against autograd
through the oracle on a machine without a GPU, and (ii) each CUDA primitive can be checked against
its twin on the GPU box.  It is never imported by the product path.
"""
from __future__ import annotations

import numpy as np
import torch
import torch.nn.functional as F

from oracle import philox

BN_EPS = 1e-3
ACT_NONE, ACT_RELU, ACT_TANH = 0, 1, 2


def _keep(seed, stream, n, p, offset=0):
    return torch.from_numpy(philox.keep_mask(seed, stream, n, p, offset).astype(np.float32))


class TorchOps(object):
    device = torch.device("cpu")

    def __init__(self):
        self.launches = 0

    # ---------------------------------------------------------------- allocation (plumbing)
    def empty(self, *shape, dtype=torch.float32):
        return torch.empty(*shape, dtype=dtype)

    def zeros(self, *shape, dtype=torch.float32):
        return torch.zeros(*shape, dtype=dtype)

    def to_device(self, a, dtype=None):
        t = torch.as_tensor(np.asarray(a))
        return t if dtype is None else t.to(dtype)

    # ---------------------------------------------------------------- dense / conv contractions
    def gemm(self, A, B, C, ta=False, tb=False, beta=0.0, bias=None):
        a = A.t() if ta else A
        b = B.t() if tb else B
        r = a @ b
        if bias is not None:
            r = r + bias
        if beta != 0.0:
            r = r + beta * C
        C.copy_(r)

    def conv1d_fwd(self, x, w, bias, y):
        k = w.shape[0]
        left = (k - 1) // 2
        xp = F.pad(x.transpose(1, 2), (left, k - 1 - left))
        y.copy_(F.conv1d(xp, w.permute(2, 1, 0).contiguous(), bias).transpose(1, 2))

    def conv1d_dgrad(self, dy, w, dx):
        k = w.shape[0]
        left = (k - 1) // 2
        T = dy.shape[1]
        # dx[b,t,ci] = sum_{j,co} dy[b, t - j + left, co] * w[j, ci, co]
        dyp = F.pad(dy.transpose(1, 2), (k - 1 - left, left))
        wf = torch.flip(w, dims=[0]).permute(1, 2, 0).contiguous()        # [ci, co, k]
        dx.copy_(F.conv1d(dyp, wf).transpose(1, 2)[:, :T])

    def conv1d_wgrad(self, x, dy, dw):
        k = dw.shape[0]
        left = (k - 1) // 2
        T = x.shape[1]
        xp = F.pad(x, (0, 0, left, k - 1 - left))
        for j in range(k):
            dw[j].copy_(torch.einsum("btc,bto->co", xp[:, j:j + T], dy))

    def colsum(self, x, out, beta=0.0):
        r = x.reshape(-1, x.shape[-1]).sum(0)
        out.copy_(r + beta * out if beta != 0.0 else r)

    # ---------------------------------------------------------------- batch norm (+ activation + dropout)
    def _drop(self, n, keep, seed, stream):
        if keep >= 1.0:
            return None
        return _keep(seed, stream, n, keep) / keep

    def bn_fwd(self, x, gamma, beta, act, keep, seed, stream, y, mean, rstd):
        x2 = x.reshape(-1, x.shape[-1])
        mu = x2.mean(0)
        var = x2.var(0, unbiased=False)
        rs = torch.rsqrt(var + BN_EPS)
        a = (x2 - mu) * rs * gamma + beta
        if act == ACT_RELU:
            a = torch.relu(a)
        elif act == ACT_TANH:
            a = torch.tanh(a)
        d = self._drop(a.numel(), keep, seed, stream)
        if d is not None:
            a = a * d.reshape(a.shape)
        y.copy_(a.reshape(y.shape))
        mean.copy_(mu)
        rstd.copy_(rs)

    def bn_bwd(self, dy, x, y, mean, rstd, gamma, act, keep, seed, stream, dx, dgamma, dbeta):
        C = x.shape[-1]
        x2, y2, dy2 = x.reshape(-1, C), y.reshape(-1, C), dy.reshape(-1, C)
        d = self._drop(x2.numel(), keep, seed, stream)
        da = dy2 if d is None else dy2 * d.reshape(dy2.shape)
        a = y2 if d is None else y2 * keep              # where the mask kept the value; elsewhere da == 0
        if act == ACT_RELU:
            dz = da * (a > 0).to(da.dtype)
        elif act == ACT_TANH:
            dz = da * (1 - a * a)
        else:
            dz = da
        xhat = (x2 - mean) * rstd
        dbeta.copy_(dz.sum(0))
        dgamma.copy_((dz * xhat).sum(0))
        M = x2.shape[0]
        dx.copy_((gamma * rstd * (dz - dbeta / M - xhat * dgamma / M)).reshape(dx.shape))

    def bn_update_moving(self, mean, rstd, mov_mean, mov_var, momentum):
        var = 1.0 / (rstd * rstd) - BN_EPS
        mov_mean.mul_(momentum).add_((1 - momentum) * mean)
        mov_var.mul_(momentum).add_((1 - momentum) * var)

    # ---------------------------------------------------------------- conv2d 3x3 stride 2 'same' (reference encoder)
    @staticmethod
    def _pads(n):
        out = -(-n // 2)
        total = max((out - 1) * 2 + 3 - n, 0)
        return total // 2, total - total // 2

    def _c2d(self, x, w, bias):
        pt, pb = self._pads(x.shape[1])
        pl, pr = self._pads(x.shape[2])
        xn = F.pad(x.permute(0, 3, 1, 2), (pl, pr, pt, pb))
        return F.conv2d(xn, w.permute(3, 2, 0, 1).contiguous(), bias, stride=2).permute(0, 2, 3, 1)

    def conv2d_fwd(self, x, w, bias, y):
        y.copy_(self._c2d(x, w, bias))

    def conv2d_dgrad(self, dy, w, dx):
        xx = torch.zeros_like(dx).requires_grad_(True)
        (gx,) = torch.autograd.grad(self._c2d(xx, w, None), xx, dy)
        dx.copy_(gx)

    def conv2d_wgrad(self, x, dy, dw):
        ww = torch.zeros_like(dw).requires_grad_(True)
        (gw,) = torch.autograd.grad(self._c2d(x, ww, None), ww, dy)
        dw.copy_(gw)

    # ---------------------------------------------------------------- LSTM / zoneout pointwise
    def _zo_masks(self, train, zoneout, seed, stream_c, stream_h, row0, B, H):
        if not train:
            k = torch.full((B, H), 1.0 - zoneout)
            return k, k
        mc = _keep(seed, stream_c, B * H, 1.0 - zoneout, row0 * H).reshape(B, H)
        mh = _keep(seed, stream_h, B * H, 1.0 - zoneout, row0 * H).reshape(B, H)
        return mc, mh

    def lstm_fwd(self, gates, pre, bias, c_prev, h_prev, lens, s, zoneout, train, seed, stream_c, stream_h, row0,
                 cnew, cz, hraw, hz):
        B, H = c_prev.shape
        z = gates.clone()
        if pre is not None:
            z = z + pre
        if bias is not None:
            z = z + bias
        i, j, f, o = torch.chunk(z, 4, dim=1)
        si, tj, sf, so = torch.sigmoid(i), torch.tanh(j), torch.sigmoid(f + 1.0), torch.sigmoid(o)
        gates.copy_(torch.cat([si, tj, sf, so], 1))
        cn = sf * c_prev + si * tj
        hr = so * torch.tanh(cn)
        mc, mh = self._zo_masks(train, zoneout, seed, stream_c, stream_h, row0, B, H)
        czv = c_prev + mc * (cn - c_prev)
        hzv = h_prev + mh * (hr - h_prev)
        if lens is not None:
            v = (s < lens.to(torch.long))[:, None]
            czv = torch.where(v, czv, c_prev)
            hzv = torch.where(v, hzv, h_prev)
            hr = torch.where(v, hr, torch.zeros_like(hr))
        cnew.copy_(cn)
        cz.copy_(czv)
        hraw.copy_(hr)
        hz.copy_(hzv)

    def lstm_bwd(self, acts, c_prev, h_prev, cnew, lens, s, zoneout, train, seed, stream_c, stream_h, row0,
                 d_hraw, d_cz, d_hz, dgates, d_cprev, d_hprev):
        B, H = c_prev.shape
        si, tj, sf, so = torch.chunk(acts, 4, dim=1)
        mc, mh = self._zo_masks(train, zoneout, seed, stream_c, stream_h, row0, B, H)
        tc = torch.tanh(cnew)
        dh_new = d_hraw + mh * d_hz
        dhp = (1 - mh) * d_hz
        dc_new = mc * d_cz + dh_new * so * (1 - tc * tc)
        dcp = (1 - mc) * d_cz + dc_new * sf
        dg = torch.cat([dc_new * tj * si * (1 - si), dc_new * si * (1 - tj * tj), dc_new * c_prev * sf * (1 - sf),
                        dh_new * tc * so * (1 - so)], 1)
        if lens is not None:
            v = (s < lens.to(torch.long))[:, None]
            dg = torch.where(v, dg, torch.zeros_like(dg))
            dcp = torch.where(v, dcp, d_cz)
            dhp = torch.where(v, dhp, d_hz)
        dgates.copy_(dg)
        d_cprev.copy_(dcp)
        d_hprev.copy_(dhp)

    def _pos(self, lens, s, reverse):
        lens = lens.to(torch.long)
        valid = s < lens
        pos = torch.where(valid, (lens - 1 - s) if reverse else torch.full_like(lens, s), torch.zeros_like(lens))
        return valid, pos

    def gather_rows(self, src, lens, s, reverse, out):
        """out[b] = src[b, pos_b] for valid rows (s < len_b; pos = s, or len-1-s when reverse), else 0."""
        valid, pos = self._pos(lens, s, reverse)
        r = src[torch.arange(src.shape[0]), pos]
        out.copy_(torch.where(valid[:, None], r, torch.zeros_like(r)))

    def scatter_rows(self, src, lens, s, reverse, dst):
        """dst[b, pos_b] = src[b] for valid rows (dst may be a column slice of a wider tensor)."""
        valid, pos = self._pos(lens, s, reverse)
        rows = torch.arange(src.shape[0])[valid]
        dst[rows, pos[valid]] = src[valid]

    # ---------------------------------------------------------------- location-sensitive attention step
    @staticmethod
    def _attn(q, cum_prev, keys, values, lens, conv_w, conv_b, Wl, va, ba):
        B, T = cum_prev.shape
        k = conv_w.shape[0]
        left = (k - 1) // 2
        cp = F.pad(cum_prev.unsqueeze(1), (left, k - 1 - left))                       # [B,1,T+k-1]
        f = F.conv1d(cp, conv_w.permute(2, 1, 0).contiguous(), conv_b).transpose(1, 2)  # [B,T,32]
        pre = keys + q.unsqueeze(1) + f @ Wl + ba
        e = (va * torch.tanh(pre)).sum(-1)
        mask = torch.arange(T)[None, :] < lens.to(torch.long)[:, None]
        e = torch.where(mask, e, torch.full_like(e, -float("inf")))
        a = torch.softmax(e, -1)
        return a, cum_prev + a, torch.bmm(a.unsqueeze(1), values).squeeze(1)

    def attn_fwd(self, q, cum_prev, keys, values, lens, conv_w, conv_b, Wl, va, ba, align, cum, ctx):
        a, c, x = self._attn(q, cum_prev, keys, values, lens, conv_w, conv_b, Wl, va, ba)
        align.copy_(a)
        cum.copy_(c)
        ctx.copy_(x)

    def attn_bwd(self, dctx, dcum, q, cum_prev, align, keys, values, lens, conv_w, conv_b, Wl, va, ba,
                 dq, dcum_prev, dkeys, p_conv_w, p_conv_b, p_Wl, p_va, p_ba):
        """Gradients of one attention step given d(ctx_t) and d(cum_t).  dkeys and the per-row weight
        partials p_* [B, ...] are ACCUMULATED; dq and dcum_prev are overwritten.  (dvalues is a
        post-loop batched contraction of the saved alignments with the saved d(ctx).)"""
        ins = [t.detach().clone().requires_grad_(True) for t in (q, cum_prev, keys, conv_w, conv_b, Wl, va, ba)]
        a, c, x = self._attn(ins[0], ins[1], ins[2], values, lens, *ins[3:])
        g = torch.autograd.grad([x, c], ins, [dctx, dcum])
        dq.copy_(g[0])
        dcum_prev.copy_(g[1])
        dkeys.add_(g[2])
        for part, gg in zip((p_conv_w, p_conv_b, p_Wl, p_va, p_ba), g[3:]):
            part[0].add_(gg.reshape(part[0].shape))

    # ---------------------------------------------------------------- GRU pointwise (reference encoder)
    def gru_pre(self, g, h, rh):
        """g [B,2R] pre-activations in -> sigmoid(g) out (r | u); rh = r * h."""
        R = h.shape[1]
        g.copy_(torch.sigmoid(g))
        rh.copy_(g[:, :R] * h)

    def gru_out(self, cpre, ru, h, hnew):
        """cpre [B,R] in -> tanh out (candidate c); hnew = u*h + (1-u)*c."""
        R = h.shape[1]
        cpre.copy_(torch.tanh(cpre))
        u = ru[:, R:]
        hnew.copy_(u * h + (1 - u) * cpre)

    def gru_out_bwd(self, dhnew, ru, h, c, dcpre, dg, dh):
        R = h.shape[1]
        u = ru[:, R:]
        dcpre.copy_(dhnew * (1 - u) * (1 - c * c))
        dg[:, R:].copy_(dhnew * (h - c) * u * (1 - u))
        dh.copy_(dhnew * u)

    def gru_pre_bwd(self, drh, ru, h, dg, dh):
        R = h.shape[1]
        r = ru[:, :R]
        dg[:, :R].copy_(drh * h * r * (1 - r))
        dh.add_(drh * r)

    # ---------------------------------------------------------------- GST multi-head (mlp, weight-normalised v)
    @staticmethod
    def _gst(q, k, v, g, b, tt, nh):
        B, U = q.shape
        N = k.shape[0]
        hu = U // nh
        qs = q.reshape(B, nh, 1, hu)
        ks = k.reshape(N, nh, hu).permute(1, 0, 2).unsqueeze(0)
        nv = g * v * torch.rsqrt(torch.sum(v * v))
        sc = torch.sum(nv * torch.tanh(ks + qs + b), dim=-1)            # [B, nh, N]
        w = torch.softmax(sc, -1)
        return w, (w @ tt).reshape(B, -1)

    def gst_fwd(self, q, k, v, g, b, tt, nh, wts, style):
        w, s = self._gst(q, k, v, g, b, tt, nh)
        wts.copy_(w)
        style.copy_(s)

    def gst_bwd(self, dstyle, q, k, v, g, b, tt, wts, nh, dq, dk, dv, dg, db, dtt):
        ins = [t.detach().clone().requires_grad_(True) for t in (q, k, v, g, b, tt)]
        _, s = self._gst(*ins, nh)
        gr = torch.autograd.grad(s, ins, dstyle)
        for o, x in zip((dq, dk, dv, dg, db, dtt), gr):
            o.copy_(x.reshape(o.shape))

    # ---------------------------------------------------------------- elementwise
    def ew(self, op, out, a, b=None, alpha=1.0):
        if op == "copy":
            r = a
        elif op == "add":           # out = a + alpha * b
            r = a + alpha * b
        elif op == "mul":
            r = a * b
        elif op == "scale":
            r = a * alpha
        elif op == "tanh":
            r = torch.tanh(a)
        elif op == "tanh_bwd":      # a = dy, b = y
            r = a * (1 - b * b)
        else:
            raise ValueError(op)
        out.copy_(r)

    def clip_fwd(self, x, lo, hi, y):
        y.copy_(torch.clamp(x, lo, hi))

    def clip_bwd(self, dy, x, lo, hi, dx):
        dx.copy_(dy * ((x >= lo) & (x <= hi)).to(dy.dtype))

    def mse(self, pred, tgt, loss, dpred):
        """loss[0] = mean((pred - tgt)^2); dpred = 2 (pred - tgt) / n."""
        d = pred - tgt
        loss[0] = (d * d).mean()
        dpred.copy_(2.0 * d / d.numel())

    def sigmoid_ce(self, logits, labels, loss, dlogits):
        """tf.nn.sigmoid_cross_entropy_with_logits, mean over all elements."""
        l, y = logits, labels
        loss[0] = (torch.clamp(l, min=0) - l * y + torch.log1p(torch.exp(-l.abs()))).mean()
        dlogits.copy_((torch.sigmoid(l) - y) / l.numel())

    def embedding_fwd(self, ids, table, out):
        out.copy_(table[ids.to(torch.long)])

    def embedding_bwd(self, ids, dout, dtable):
        dtable.zero_()
        dtable.index_add_(0, ids.reshape(-1).to(torch.long), dout.reshape(-1, dout.shape[-1]))

    def _prenet_mask(self, S, B, U, layer, seed, step0):
        m = np.stack([philox.prenet_keep_mask(seed, step0 + s, B, U, layer) for s in range(S)], 0)
        return torch.from_numpy(m.astype(np.float32))

    def prenet_act_fwd(self, z, layer, seed, step0, keep, out):
        """out = relu(z) * keepbit / keep with the decoder's prenet Philox stream (seed, step, row, layer, unit)."""
        S, B, U = z.shape
        out.copy_(torch.relu(z) * self._prenet_mask(S, B, U, layer, seed, step0) / keep)

    def prenet_act_bwd(self, dout, out, layer, seed, step0, keep, dz):
        S, B, U = out.shape
        dz.copy_(dout * self._prenet_mask(S, B, U, layer, seed, step0) / keep * (out > 0).to(dout.dtype))

    def mask_rows(self, x, lens):
        """x [B,T,C]: zero the rows t >= len_b in place."""
        T = x.shape[1]
        m = (torch.arange(T)[None, :] < lens.to(torch.long)[:, None]).to(x.dtype)
        x.mul_(m[:, :, None])

    def tile_style(self, style, memory, col0):
        memory[:, :, col0:col0 + style.shape[1]] = style[:, None, :]

    def sum_style(self, dmemory, col0, dstyle):
        dstyle.copy_(dmemory[:, :, col0:col0 + dstyle.shape[1]].sum(1))

    # ---------------------------------------------------------------- optimizer
    def grad_finalize(self, g, p, regw, inv_world, sumsq):
        """g = g * inv_world + regw * p (L2 term of the loss, Tacotron.py:419-422); sumsq[0] += sum(g^2)."""
        g.mul_(inv_world).add_(regw * p)
        sumsq[0] += (g.double() ** 2).sum().float()

    def adam_step(self, p, g, m, v, sumsq, clip_norm, lr_t, b1, b2, eps):
        """tf.clip_by_global_norm + tf.train.AdamOptimizer update (lr_t already bias-corrected)."""
        norm = torch.sqrt(sumsq[0])
        scale = clip_norm / torch.clamp(norm, min=clip_norm) if clip_norm > 0 else torch.tensor(1.0)
        gg = g * scale
        m.mul_(b1).add_((1 - b1) * gg)
        v.mul_(b2).add_((1 - b2) * gg * gg)
        p.sub_(lr_t * m / (torch.sqrt(v) + eps))

    def reg_loss(self, p, regw, out):
        """out[0] = sum(regw * p^2) / 2  (tf.nn.l2_loss over the name-filtered set times reg_weight)."""
        out[0] = 0.5 * (regw.double() * p.double() ** 2).sum().float()
