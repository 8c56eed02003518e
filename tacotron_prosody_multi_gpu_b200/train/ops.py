"""ctypes binding of the training primitives (include/taco_b200_train.h) -- the device side of train/step.py.

torch tensors are only the memory carrier: every method passes raw device pointers (and strides where
views are allowed) to a ``taco_op_*`` entry point of libtaco_b200.so on torch's current stream.  There is
no CPU / PyTorch fallback: construction fails without the library or a CUDA device.
"""
from __future__ import annotations

import ctypes as C

import numpy as np
import torch

from .. import _lib

EW = {"copy": 0, "add": 1, "mul": 2, "scale": 3, "tanh": 4, "tanh_bwd": 5}

_P, _I, _F, _LL, _U32, _U64 = C.c_void_p, C.c_int, C.c_float, C.c_longlong, C.c_uint32, C.c_uint64
_SIGS = {
    "taco_op_gemm": [_P, _LL, _P, _LL, _P, _LL, _I, _I, _I, _I, _I, _F, _P, _P],
    "taco_op_conv1d_fwd": [_P, _P, _P, _P, _I, _I, _I, _I, _I, _P],
    "taco_op_conv1d_dgrad": [_P, _P, _P, _I, _I, _I, _I, _I, _P],
    "taco_op_conv1d_wgrad": [_P, _P, _P, _I, _I, _I, _I, _I, _P],
    "taco_op_conv2d_fwd": [_P, _P, _P, _P, _I, _I, _I, _I, _I, _P],
    "taco_op_conv2d_dgrad": [_P, _P, _P, _I, _I, _I, _I, _I, _P],
    "taco_op_conv2d_wgrad": [_P, _P, _P, _I, _I, _I, _I, _I, _P],
    "taco_op_colsum": [_P, _LL, _I, _I, _P, _F, _P],
    "taco_op_bn_fwd": [_P, _P, _P, _I, _U32, _F, _U64, _U32, _P, _P, _P, _I, _I, _P],
    "taco_op_bn_bwd": [_P, _P, _P, _P, _P, _P, _I, _U32, _F, _U64, _U32, _P, _P, _P, _I, _I, _P],
    "taco_op_bn_update_moving": [_P, _P, _P, _P, _F, _I, _P],
    "taco_op_lstm_fwd": [_P, _P, _P, _P, _P, _P, _I, _F, _I, _U32, _U64, _U32, _U32, _I, _P, _P, _P, _P, _I, _I, _P],
    "taco_op_lstm_bwd": [_P, _P, _P, _P, _P, _I, _F, _I, _U32, _U64, _U32, _U32, _I, _P, _P, _P, _P, _P, _P, _I, _I, _P],
    "taco_op_gather_rows": [_P, _LL, _LL, _P, _I, _I, _P, _I, _I, _P],
    "taco_op_scatter_rows": [_P, _P, _I, _I, _P, _LL, _LL, _I, _I, _P],
    "taco_op_attn_fwd": [_P] * 13 + [_I] * 6 + [_P],
    "taco_op_attn_bwd": [_P] * 22 + [_I] * 6 + [_P],
    "taco_op_attn_fold": [_P] * 5 + [_P],
    "taco_op_attn_fwd_fused": [_P] * 15 + [_I] * 4 + [_P],
    "taco_op_attn_bwd_fused": [_P] * 15 + [_I] * 4 + [_P],
    "taco_op_attn_bwd_hoisted": [_P] * 15 + [_I] * 3 + [_P],
    "taco_op_attn_dvalues": [_P] * 3 + [_I] * 4 + [_P],
    "taco_op_gru_pre": [_P, _P, _P, _I, _I, _P],
    "taco_op_gru_out": [_P, _P, _P, _P, _I, _I, _P],
    "taco_op_gru_out_bwd": [_P, _P, _P, _P, _P, _P, _P, _I, _I, _P],
    "taco_op_gru_pre_bwd": [_P, _P, _P, _P, _P, _I, _I, _P],
    "taco_op_gst_fwd": [_P] * 8 + [_I] * 5 + [_P],
    "taco_op_gst_bwd": [_P] * 15 + [_I] * 5 + [_P],
    "taco_op_ew": [_I, _P, _P, _P, _F, _LL, _P],
    "taco_op_clip_fwd": [_P, _F, _F, _P, _LL, _P],
    "taco_op_clip_bwd": [_P, _P, _F, _F, _P, _LL, _P],
    "taco_op_mse": [_P, _P, _P, _P, _LL, _P],
    "taco_op_sigmoid_ce": [_P, _P, _P, _P, _LL, _P],
    "taco_op_embedding_fwd": [_P, _P, _P, _I, _I, _I, _P],
    "taco_op_embedding_bwd": [_P, _P, _P, _I, _I, _I, _P],
    "taco_op_prenet_act_fwd": [_P, _I, _U64, _I, _F, _P, _I, _I, _I, _P],
    "taco_op_prenet_act_bwd": [_P, _P, _I, _U64, _I, _F, _P, _I, _I, _I, _P],
    "taco_op_mask_rows": [_P, _P, _I, _I, _I, _P],
    "taco_op_tile_style": [_P, _P, _I, _I, _I, _I, _I, _P],
    "taco_op_sum_style": [_P, _P, _I, _I, _I, _I, _I, _P],
    "taco_op_reg_loss": [_P, _P, _P, _LL, _P],
    "taco_op_grad_finalize": [_P, _P, _P, _F, _P, _LL, _P],
    "taco_op_adam_step": [_P, _P, _P, _P, _P, _F, _F, _F, _F, _F, _LL, _P],
}
TRAIN_SYMBOLS = tuple(_SIGS.keys())


def keep_threshold(p):
    return min(int(p * 4294967296.0), 0xFFFFFFFF)


def _p(t):
    return 0 if t is None else t.data_ptr()


def _rows(t):
    """2-D view with unit inner stride -> (ptr, leading stride)."""
    assert t.dim() == 2 and (t.stride(1) == 1 or t.shape[1] == 1), (t.shape, t.stride())
    return t.data_ptr(), (t.stride(0) if t.shape[0] > 1 else max(t.shape[1], 1))


def _c(t):
    assert t.is_contiguous(), (t.shape, t.stride())
    return t.data_ptr()


class _Side(object):
    def __init__(self, ops, i):
        self.ops, self.i = ops, i

    def __enter__(self):
        o = self.ops
        if self.i not in o._sides:
            o._sides[self.i] = torch.cuda.Stream(device=o.device)
        s = o._sides[self.i]
        s.wait_stream(torch.cuda.current_stream(o.device))
        self.cm = torch.cuda.stream(s)
        self.cm.__enter__()
        if s not in o._pending:
            o._pending.append(s)
        return s

    def __exit__(self, *exc):
        return self.cm.__exit__(*exc)


class CudaOps(object):
    def __init__(self, device=0):
        if not torch.cuda.is_available():
            raise _lib.TacoError("CudaOps needs a CUDA device (sm_100a); the training path has no CPU fallback")
        self.device = torch.device("cuda", device)
        self.L = _lib.lib()
        for name, sig in _SIGS.items():
            fn = getattr(self.L, name)      # AttributeError if the library was built without the training kernels
            fn.argtypes = sig
            fn.restype = C.c_int
        self.launches = 0
        self._sides, self._pending = {}, []

    # ------------------------------------------------------------------ plumbing
    def empty(self, *shape, dtype=torch.float32):
        return torch.empty(*shape, dtype=dtype, device=self.device)

    def zeros(self, *shape, dtype=torch.float32):
        return torch.zeros(*shape, dtype=dtype, device=self.device)

    def to_device(self, a, dtype=None):
        t = torch.as_tensor(np.asarray(a))
        if dtype is not None:
            t = t.to(dtype)
        return t.to(self.device)

    def _st(self):
        return torch.cuda.current_stream(self.device).cuda_stream

    def _call(self, name, *args):
        rc = getattr(self.L, name)(*args, self._st())
        self.launches += 1
        if rc != 0:
            raise _lib.TacoError("%s failed with code %d (%s)" % (name, rc, "bad argument" if rc == -1 else "CUDA error"))

    # ------------------------------------------------------------------ independent branches of the step on side streams
    def side(self, i):
        """Context manager: the enclosed launches go to side stream ``i``, ordered after everything already queued on the
        current stream.  The text encoder's two LSTM directions and the reference encoder are independent chains of small,
        latency-bound launches; ``join()`` makes the current stream wait for all of them."""
        return _Side(self, i)

    def join(self):
        cur = torch.cuda.current_stream(self.device)
        for s in self._pending:
            cur.wait_stream(s)
        self._pending = []

    # ------------------------------------------------------------------ contractions
    def gemm(self, A, B, Cm, ta=False, tb=False, beta=0.0, bias=None):
        pa, lda = _rows(A)
        pb, ldb = _rows(B)
        pc, ldc = _rows(Cm)
        M, N = Cm.shape
        K = A.shape[0] if ta else A.shape[1]
        assert (A.shape[1] if ta else A.shape[0]) == M and (B.shape[1] if tb else B.shape[0]) == K and \
            (B.shape[0] if tb else B.shape[1]) == N, (A.shape, B.shape, Cm.shape, ta, tb)
        self._call("taco_op_gemm", pa, lda, pb, ldb, pc, ldc, M, N, K, int(ta), int(tb), float(beta), _p(bias))

    def conv1d_fwd(self, x, w, bias, y):
        B, T, Ci = x.shape
        k, _, Co = w.shape
        self._call("taco_op_conv1d_fwd", _c(x), _c(w), _p(bias), _c(y), B, T, Ci, Co, k)

    def conv1d_dgrad(self, dy, w, dx):
        B, T, Co = dy.shape
        k, Ci, _ = w.shape
        self._call("taco_op_conv1d_dgrad", _c(dy), _c(w), _c(dx), B, T, Ci, Co, k)

    def conv1d_wgrad(self, x, dy, dw):
        B, T, Ci = x.shape
        k, _, Co = dw.shape
        self._call("taco_op_conv1d_wgrad", _c(x), _c(dy), _c(dw), B, T, Ci, Co, k)

    def conv2d_fwd(self, x, w, bias, y):
        B, H, Wd, Ci = x.shape
        self._call("taco_op_conv2d_fwd", _c(x), _c(w), _p(bias), _c(y), B, H, Wd, Ci, w.shape[3])

    def conv2d_dgrad(self, dy, w, dx):
        B, H, Wd, Ci = dx.shape
        self._call("taco_op_conv2d_dgrad", _c(dy), _c(w), _c(dx), B, H, Wd, Ci, w.shape[3])

    def conv2d_wgrad(self, x, dy, dw):
        B, H, Wd, Ci = x.shape
        self._call("taco_op_conv2d_wgrad", _c(x), _c(dy), _c(dw), B, H, Wd, Ci, dw.shape[3])

    def colsum(self, x, out, beta=0.0):
        x2 = x.reshape(-1, x.shape[-1]) if x.dim() != 2 else x
        px, ld = _rows(x2)
        self._call("taco_op_colsum", px, ld, x2.shape[0], x2.shape[1], _c(out), float(beta))

    # ------------------------------------------------------------------ batch norm
    def bn_fwd(self, x, gamma, beta, act, keep, seed, stream, y, mean, rstd):
        Cc = x.shape[-1]
        self._call("taco_op_bn_fwd", _c(x), _c(gamma), _c(beta), act, keep_threshold(keep), float(keep), seed, stream,
                   _c(y), _c(mean), _c(rstd), x.numel() // Cc, Cc)

    def bn_bwd(self, dy, x, y, mean, rstd, gamma, act, keep, seed, stream, dx, dgamma, dbeta):
        Cc = x.shape[-1]
        self._call("taco_op_bn_bwd", _c(dy), _c(x), _c(y), _c(mean), _c(rstd), _c(gamma), act, keep_threshold(keep), float(keep),
                   seed, stream, _c(dx), _c(dgamma), _c(dbeta), x.numel() // Cc, Cc)

    def bn_update_moving(self, mean, rstd, mov_mean, mov_var, momentum):
        self._call("taco_op_bn_update_moving", _c(mean), _c(rstd), _c(mov_mean), _c(mov_var), float(momentum), mean.numel())

    # ------------------------------------------------------------------ LSTM
    def lstm_fwd(self, gates, pre, bias, c_prev, h_prev, lens, s, zoneout, train, seed, stream_c, stream_h, row0,
                 cnew, cz, hraw, hz):
        B, H = c_prev.shape
        self._call("taco_op_lstm_fwd", _c(gates), _p(pre), _p(bias), _c(c_prev), _c(h_prev), _p(lens), s, float(zoneout), int(train),
                   keep_threshold(1.0 - zoneout), seed, stream_c, stream_h, row0, _c(cnew), _c(cz), _c(hraw), _c(hz), B, H)

    def lstm_bwd(self, acts, c_prev, h_prev, cnew, lens, s, zoneout, train, seed, stream_c, stream_h, row0,
                 d_hraw, d_cz, d_hz, dgates, d_cprev, d_hprev):
        B, H = c_prev.shape
        self._call("taco_op_lstm_bwd", _c(acts), _c(c_prev), _c(h_prev), _c(cnew), _p(lens), s, float(zoneout), int(train),
                   keep_threshold(1.0 - zoneout), seed, stream_c, stream_h, row0, _c(d_hraw), _c(d_cz), _c(d_hz), _c(dgates),
                   _c(d_cprev), _c(d_hprev), B, H)

    def gather_rows(self, src, lens, s, reverse, out):
        B, _, N = src.shape
        assert src.stride(2) == 1
        self._call("taco_op_gather_rows", src.data_ptr(), src.stride(0), src.stride(1), _c(lens), s, int(reverse), _c(out), B, N)

    def scatter_rows(self, src, lens, s, reverse, dst):
        B, N = src.shape
        assert dst.stride(2) == 1 and dst.shape[2] == N
        self._call("taco_op_scatter_rows", _c(src), _c(lens), s, int(reverse), dst.data_ptr(), dst.stride(0), dst.stride(1), B, N)

    # ------------------------------------------------------------------ attention
    def attn_fwd(self, q, cum_prev, keys, values, lens, conv_w, conv_b, Wl, va, ba, align, cum, ctx):
        B, T, A = keys.shape
        self._call("taco_op_attn_fwd", _c(q), _c(cum_prev), _c(keys), _c(values), _c(lens), _c(conv_w), _c(conv_b), _c(Wl), _c(va),
                   _c(ba), _c(align), _c(cum), _c(ctx), B, T, A, values.shape[2], conv_w.shape[2], conv_w.shape[0])

    def attn_bwd(self, dctx, dcum, q, cum_prev, align, keys, values, lens, conv_w, conv_b, Wl, va, ba,
                 dq, dcum_prev, dkeys, p_conv_w, p_conv_b, p_Wl, p_va, p_ba):
        B, T, A = keys.shape
        self._call("taco_op_attn_bwd", _c(dctx), _c(dcum), _c(q), _c(cum_prev), _c(align), _c(keys), _c(values), _c(lens), _c(conv_w),
                   _c(conv_b), _c(Wl), _c(va), _c(ba), _c(dq), _c(dcum_prev), _c(dkeys), _c(p_conv_w), _c(p_conv_b), _c(p_Wl),
                   _c(p_va), _c(p_ba), _c(self.empty(B * T * conv_w.shape[2])), B, T, A, values.shape[2], conv_w.shape[2], conv_w.shape[0])

    # fused attention of the time loops (train_attn_fused.cu)
    def attn_fused_supported(self, T, D, A, V, NF, KW):
        fn = self.L.taco_op_attn_fused_supported
        fn.argtypes, fn.restype = [C.c_int] * 6, C.c_int
        return bool(fn(T, D, A, V, NF, KW))

    def attn_fold(self, conv_w, conv_b, Wl, U, c0):
        self._call("taco_op_attn_fold", _c(conv_w), _c(conv_b), _c(Wl), _c(U), _c(c0))

    def attn_fwd_fused(self, h2raw, Wq, U, c0, ba, va, cum_prev, keys, values, lens, q, th, align, cum, ctx):
        B, T, _ = keys.shape
        self._call("taco_op_attn_fwd_fused", _c(h2raw), _c(Wq), _c(U), _c(c0), _c(ba), _c(va), _c(cum_prev), _c(keys), _c(values), _c(lens),
                   _c(q), _c(th), _c(align), _c(cum), _c(ctx), B, T, h2raw.shape[1], values.shape[2])

    def attn_bwd_fused(self, dctx, carry, dcum, align, th, values, lens, U, va, Wq, de, dq, dh2, dcum_prev):
        B, T, V = values.shape
        self._call("taco_op_attn_bwd_fused", _c(dctx), _c(carry), _c(dcum), _c(align), _c(th), _c(values), _c(lens), _c(U), _c(va), _c(Wq),
                   _c(de), _c(dq), _c(dh2), _c(dcum_prev), _c(self.empty(B * T)), B, T, dh2.shape[1], V)

    def attn_dvalues(self, align, dctx, dvalues):
        S, B, T = align.shape
        self._call("taco_op_attn_dvalues", _c(align), _c(dctx), _c(dvalues), S, B, T, dctx.shape[2])

    def attn_bwd_hoisted(self, de, th, cum, lens, conv_w, conv_b, Wl, va, dkeys, g_conv_w, g_conv_b, g_Wl, g_va, g_ba):
        S, B, T = de.shape
        self._call("taco_op_attn_bwd_hoisted", _c(de), _c(th), _c(cum), _c(lens), _c(conv_w), _c(conv_b), _c(Wl), _c(va), _c(dkeys),
                   _c(g_conv_w), _c(g_conv_b), _c(g_Wl), _c(g_va), _c(g_ba), _c(self.empty(B * (conv_w.shape[0] + 2) * Wl.shape[1])), S, B, T)

    # ------------------------------------------------------------------ GRU / GST
    def gru_pre(self, g, h, rh):
        self._call("taco_op_gru_pre", _c(g), _c(h), _c(rh), h.shape[0], h.shape[1])

    def gru_out(self, cpre, ru, h, hnew):
        self._call("taco_op_gru_out", _c(cpre), _c(ru), _c(h), _c(hnew), h.shape[0], h.shape[1])

    def gru_out_bwd(self, dhnew, ru, h, c, dcpre, dg, dh):
        self._call("taco_op_gru_out_bwd", _c(dhnew), _c(ru), _c(h), _c(c), _c(dcpre), _c(dg), _c(dh), h.shape[0], h.shape[1])

    def gru_pre_bwd(self, drh, ru, h, dg, dh):
        self._call("taco_op_gru_pre_bwd", _c(drh), _c(ru), _c(h), _c(dg), _c(dh), h.shape[0], h.shape[1])

    def gst_fwd(self, q, k, v, g, b, tt, nh, wts, style):
        self._call("taco_op_gst_fwd", _c(q), _c(k), _c(v), _c(g), _c(b), _c(tt), _c(wts), _c(style), q.shape[0], k.shape[0], q.shape[1],
                   tt.shape[1], nh)

    def gst_bwd(self, dstyle, q, k, v, g, b, tt, wts, nh, dq, dk, dv, dg, db, dtt):
        scratch = self.empty(64)
        self._call("taco_op_gst_bwd", _c(dstyle), _c(q), _c(k), _c(v), _c(g), _c(b), _c(tt), _c(wts), _c(dq), _c(dk), _c(dv), _c(dg),
                   _c(db), _c(dtt), _c(scratch), q.shape[0], k.shape[0], q.shape[1], tt.shape[1], nh)

    # ------------------------------------------------------------------ elementwise / losses / misc
    def ew(self, op, out, a, b=None, alpha=1.0):
        assert out.is_contiguous() and a.is_contiguous() and (b is None or b.is_contiguous())
        self._call("taco_op_ew", EW[op], out.data_ptr(), a.data_ptr(), _p(b), float(alpha), out.numel())

    def clip_fwd(self, x, lo, hi, y):
        self._call("taco_op_clip_fwd", _c(x), float(lo), float(hi), _c(y), x.numel())

    def clip_bwd(self, dy, x, lo, hi, dx):
        self._call("taco_op_clip_bwd", _c(dy), _c(x), float(lo), float(hi), _c(dx), x.numel())

    def mse(self, pred, tgt, loss, dpred):
        self._call("taco_op_mse", _c(pred), _c(tgt), loss.data_ptr(), _c(dpred), pred.numel())

    def sigmoid_ce(self, logits, labels, loss, dlogits):
        self._call("taco_op_sigmoid_ce", _c(logits), _c(labels), loss.data_ptr(), _c(dlogits), logits.numel())

    def embedding_fwd(self, ids, table, out):
        self._call("taco_op_embedding_fwd", _c(ids), _c(table), _c(out), ids.numel(), table.shape[1], table.shape[0])

    def embedding_bwd(self, ids, dout, dtable):
        self._call("taco_op_embedding_bwd", _c(ids), _c(dout), _c(dtable), ids.numel(), dtable.shape[1], dtable.shape[0])

    def prenet_act_fwd(self, z, layer, seed, step0, keep, out):
        S, B, U = z.shape
        self._call("taco_op_prenet_act_fwd", _c(z), layer, seed, step0, float(keep), _c(out), S, B, U)

    def prenet_act_bwd(self, dout, out, layer, seed, step0, keep, dz):
        S, B, U = out.shape
        self._call("taco_op_prenet_act_bwd", _c(dout), _c(out), layer, seed, step0, float(keep), _c(dz), S, B, U)

    def mask_rows(self, x, lens):
        B, T, Cc = x.shape
        self._call("taco_op_mask_rows", _c(x), _c(lens), B, T, Cc)

    def tile_style(self, style, memory, col0):
        B, T, MD = memory.shape
        self._call("taco_op_tile_style", _c(style), _c(memory), B, T, style.shape[1], MD, col0)

    def sum_style(self, dmemory, col0, dstyle):
        B, T, MD = dmemory.shape
        self._call("taco_op_sum_style", _c(dmemory), _c(dstyle), B, T, dstyle.shape[1], MD, col0)

    # ------------------------------------------------------------------ optimizer
    def reg_loss(self, p, regw, out):
        self._call("taco_op_reg_loss", _c(p), _c(regw), out.data_ptr(), p.numel())

    def grad_finalize(self, g, p, regw, inv_world, sumsq):
        self._call("taco_op_grad_finalize", _c(g), _c(p), _c(regw), float(inv_world), _c(sumsq), g.numel())

    def adam_step(self, p, g, m, v, sumsq, clip_norm, lr_t, b1, b2, eps):
        self._call("taco_op_adam_step", _c(p), _c(g), _c(m), _c(v), _c(sumsq), float(clip_norm), float(lr_t), float(b1), float(b2),
                   float(eps), p.numel())


# ---------------------------------------------------------------------- whole time loops sequenced in C (train_loops.cu)
class TacoDecLoop(C.Structure):
    _fields_ = ([(n, C.c_int32) for n in ("B", "T", "S", "D", "MD", "A", "U2", "NF", "KW")] +
                [("zoneout", C.c_float), ("zo_keep_thresh", C.c_uint32), ("seed", C.c_uint64), ("stream1", C.c_uint32),
                 ("stream2", C.c_uint32)] +
                [(n, C.c_void_p) for n in ("W1", "W2", "b2", "Wq", "conv_w", "conv_b", "Wl", "va", "ba", "keys", "values", "lens", "G1x",
                                           "acts1", "acts2", "c1new", "c2new", "h1raw", "h2raw", "c1z", "h1z", "c2z", "h2z", "ctx", "cum",
                                           "align", "q", "dH2", "dCtx", "dG1", "dG2", "dQ", "dkeys", "p_conv_w", "p_conv_b", "p_Wl",
                                           "p_va", "p_ba", "work", "th", "de", "attn_work")])


class TacoLstmLoop(C.Structure):
    _fields_ = ([(n, C.c_int32) for n in ("B", "T", "H", "C", "reverse")] +
                [("zoneout", C.c_float), ("zo_keep_thresh", C.c_uint32), ("seed", C.c_uint64), ("stream_c", C.c_uint32),
                 ("stream_h", C.c_uint32)] +
                [(n, C.c_void_p) for n in ("K", "lens", "xw", "acts", "cnew", "cz", "hz", "out")] +
                [("out_sb", C.c_longlong), ("out_st", C.c_longlong)] +
                [(n, C.c_void_p) for n in ("dxw", "dg", "work")])


class TacoGruLoop(C.Structure):
    _fields_ = ([(n, C.c_int32) for n in ("B", "Tp", "In", "R")] +
                [(n, C.c_void_p) for n in ("Wg", "Wc", "GX", "CX", "hg", "rh", "dGX", "dCX", "dh", "work")])


LOOP_SYMBOLS = ("taco_op_decoder_loop_fwd", "taco_op_decoder_loop_bwd", "taco_op_lstm_loop_fwd", "taco_op_lstm_loop_bwd",
                "taco_op_gru_loop_fwd", "taco_op_gru_loop_bwd")


def _fill(struct, ints, ptrs):
    for k, v in ints.items():
        setattr(struct, k, v)
    for k, t in ptrs.items():
        if t is not None:
            assert t.is_contiguous() or k == "out", k
            setattr(struct, k, t.data_ptr())
    return struct


def _loop_call(self, name, struct, n_launch):
    fn = getattr(self.L, name)
    fn.argtypes = [C.c_void_p, C.c_void_p]
    fn.restype = C.c_int
    rc = fn(C.byref(struct), self._st())
    self.launches += n_launch
    if rc != 0:
        raise _lib.TacoError("%s failed with code %d" % (name, rc))


def _decoder_loop(self, bwd, dims, zoneout, seed, stream1, stream2, **t):
    s = _fill(TacoDecLoop(), dict(dims, zoneout=float(zoneout), zo_keep_thresh=keep_threshold(1.0 - zoneout), seed=seed,
                                   stream1=stream1, stream2=stream2), t)
    fused = t.get("th") is not None and t.get("attn_work") is not None
    per_step = ((5 if fused else 12) if bwd else (5 if fused else 8))
    _loop_call(self, "taco_op_decoder_loop_bwd" if bwd else "taco_op_decoder_loop_fwd", s, dims["S"] * per_step)


def _lstm_loop(self, bwd, dims, zoneout, seed, stream_c, stream_h, out, **t):
    s = _fill(TacoLstmLoop(), dict(dims, zoneout=float(zoneout), zo_keep_thresh=keep_threshold(1.0 - zoneout), seed=seed,
                                    stream_c=stream_c, stream_h=stream_h, out_sb=out.stride(0), out_st=out.stride(1)),
              dict(t, out=out))
    assert out.stride(2) == 1
    _loop_call(self, "taco_op_lstm_loop_bwd" if bwd else "taco_op_lstm_loop_fwd", s, dims["T"] * 2)


def _gru_loop(self, bwd, dims, **t):
    s = _fill(TacoGruLoop(), dims, t)
    _loop_call(self, "taco_op_gru_loop_bwd" if bwd else "taco_op_gru_loop_fwd", s, dims["Tp"] * 4)


CudaOps.gru_loop = _gru_loop
CudaOps.decoder_loop = _decoder_loop
CudaOps.lstm_loop = _lstm_loop
TRAIN_SYMBOLS = tuple(_SIGS.keys()) + LOOP_SYMBOLS + ("taco_op_set_precision", "taco_op_gemm_pair_acc", "taco_op_gemm_nt_split_acc", "taco_op_gemm_acc", "taco_op_attn_fused_supported", "taco_op_attn_bwd_noalloc",
                                                         "taco_op_lstm_fwd_seq", "taco_op_lstm_bwd_seq")


def _set_precision(self, precision):
    """"fp32": SIMT FMA everywhere (parity mode); "bf16": the throughput-bound contractions (conv fwd / dgrad / wgrad and the
    GEMMs over all S*B rows) on mma.sync tensor cores with bf16 operands and fp32 accumulation (process-wide)."""
    code = {"fp32": 0, "bf16": 1}[precision]
    self.L.taco_op_set_precision.argtypes = [C.c_int]
    self.L.taco_op_set_precision.restype = C.c_int
    if self.L.taco_op_set_precision(code) != 0:
        raise _lib.TacoError("taco_op_set_precision failed")
    self.precision = precision


CudaOps.set_precision = _set_precision
