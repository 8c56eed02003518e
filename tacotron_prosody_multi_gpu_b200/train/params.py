"""Flat parameter / gradient / Adam-moment buffers of the training step.

The reference keeps one ``tf.Variable`` per tensor on the CPU parameter server and averages the
per-tower gradients there (Tacotron.py:484-501).  Here every trainable variable is a view into ONE
flat fp32 device buffer (20 425 491 elements at default hparams), so the data-parallel exchange
is a single ``all_reduce`` of the flat gradient buffer and the optimizer is one fused launch.
"""
from __future__ import annotations

from collections import OrderedDict

import numpy as np

NOT_TRAINABLE = ("moving_mean", "moving_variance")
# Tacotron.py:419-422: variables whose NAME contains one of these are not L2-regularised
REG_EXCLUDE = ("bias", "Bias", "_projection", "inputs_embedding", "RNN", "LSTM")


def is_trainable(name):
    return not name.endswith(NOT_TRAINABLE)


def is_regularized(name):
    return is_trainable(name) and not any(s in name for s in REG_EXCLUDE)


class ParamStore(object):
    def __init__(self, ops, weights, reg_weight):
        self.ops = ops
        self.layout = OrderedDict()
        off = 0
        for name, v in weights.items():
            if is_trainable(name):
                v = np.asarray(v, np.float32)
                self.layout[name] = (off, tuple(v.shape))
                off += v.size
        self.n = off
        flat = np.empty(off, np.float32)
        regw = np.zeros(off, np.float32)
        for name, (o, shp) in self.layout.items():
            sz = int(np.prod(shp)) if len(shp) else 1
            flat[o:o + sz] = np.asarray(weights[name], np.float32).reshape(-1)
            if is_regularized(name):
                regw[o:o + sz] = reg_weight
        self.p = ops.to_device(flat)
        self.regw = ops.to_device(regw)
        self.g = ops.zeros(off)
        self.m = ops.zeros(off)
        self.v = ops.zeros(off)
        self.P = self._views(self.p)
        self.G = self._views(self.g)
        # BN moving statistics: one flat buffer too (averaged across ranks by the data-parallel trainer)
        snames = [name for name in weights if not is_trainable(name)]
        sflat = np.concatenate([np.asarray(weights[n], np.float32).reshape(-1) for n in snames]) if snames else np.zeros(0, np.float32)
        self.s = ops.to_device(sflat)
        self.stats = OrderedDict()
        off = 0
        for name in snames:
            sz = int(np.asarray(weights[name]).size)
            self.stats[name] = self.s[off:off + sz]
            off += sz

    def _views(self, flat):
        out = OrderedDict()
        for name, (o, shp) in self.layout.items():
            sz = int(np.prod(shp)) if len(shp) else 1
            out[name] = flat[o:o + sz].view(*shp) if len(shp) else flat[o:o + 1]
        return out

    def export(self):
        """name -> float32 ndarray (the weight-container format of weights.py), moving statistics included."""
        out = OrderedDict()
        for name, (o, shp) in self.layout.items():
            out[name] = self.P[name].detach().cpu().numpy().reshape(shp).copy()
        for name, t in self.stats.items():
            out[name] = t.detach().cpu().numpy().copy()
        return out
