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
# variable scopes whose gradients are complete when the decoder BPTT ends (train/step.py backward): the early exchange half
ALIGN = 4          # floats: variables start on 16-byte boundaries of the flat parameter / gradient / Adam buffers
TAIL_SCOPES = ("decoder/", "postnet_convolutions/", "postnet_projection/")
# Tacotron.py:478-480: with tacotron_fine_tuning only variables NOT matching these are updated
FINE_TUNE_FROZEN = ("inputs_embedding", "encoder_")


def is_trainable(name):
    return not name.endswith(NOT_TRAINABLE)


def is_regularized(name):
    return is_trainable(name) and not any(s in name for s in REG_EXCLUDE)


class ParamStore(object):
    def __init__(self, ops, weights, reg_weight):
        self.ops = ops
        self.layout = OrderedDict()
        # every variable starts on a 16-byte boundary of the flat buffers (the kernels use 16-byte loads of weight rows where the
        # address allows it; the scalar attention_g would otherwise misalign the whole decoder and postnet).  Padding elements are
        # zero in p, g, m, v and regw: they add nothing to the global norm and Adam leaves them at zero.
        off = 0
        n_params = 0
        for name, v in weights.items():
            if is_trainable(name):
                v = np.asarray(v, np.float32)
                off = (off + ALIGN - 1) // ALIGN * ALIGN
                self.layout[name] = (off, tuple(v.shape))
                off += v.size
                n_params += v.size
        off = (off + ALIGN - 1) // ALIGN * ALIGN
        self.n = off                      # length of the flat buffers (padding included)
        self.n_params = n_params          # trainable scalars (20 425 491 with the default hyper-parameters)
        flat = np.zeros(off, np.float32)
        regw = np.zeros(off, np.float32)
        for name, (o, shp) in self.layout.items():
            sz = int(np.prod(shp)) if len(shp) else 1
            flat[o:o + sz] = np.asarray(weights[name], np.float32).reshape(-1)
            if is_regularized(name):
                regw[o:o + sz] = reg_weight
        self.p = ops.to_device(flat)
        self.regw = ops.to_device(regw)
        self.m = ops.zeros(off)
        self.v = ops.zeros(off)
        self.P = self._views(self.p)
        # ONE exchange bucket per step: [BN moving statistics | gradients].  The statistics sit in front so that
        # the bucket splits into two contiguous halves by the time their contents are final in backward():
        #   late  = bucket[: ns + tail_start]  statistics + encoder / style gradients (final at the end of backward)
        #   early = bucket[ns + tail_start :]  decoder + postnet gradients (final once the decoder BPTT is done)
        snames = [name for name in weights if not is_trainable(name)]
        sflat = np.concatenate([np.asarray(weights[n], np.float32).reshape(-1) for n in snames]) if snames else np.zeros(0, np.float32)
        self.ns = int(sflat.size)
        self.bucket = ops.zeros(self.ns + off)
        self.s = self.bucket[:self.ns]
        if self.ns:
            self.s.copy_(ops.to_device(sflat))
        self.g = self.bucket[self.ns:]
        self.G = self._views(self.g)
        names = list(self.layout.keys())
        k = len(names)
        while k > 0 and names[k - 1].startswith(TAIL_SCOPES):
            k -= 1
        self.tail_start = self.layout[names[k]][0] if k < len(names) else off   # first gradient of the early half
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

    # ---- resumable state (tf.train.Saver keeps the Adam slots and global_step in the checkpoint, train.py:60) ----
    def save_state(self, path, global_step):
        """Variables, BN moving statistics, Adam moments and the update count in one .npz; ``load_state`` resumes the
        schedule (lr decay, Adam bias correction) exactly where it stopped."""
        out = {k.replace("/", "|"): v for k, v in self.export().items()}
        out["__adam_m__"] = self.m.detach().cpu().numpy()
        out["__adam_v__"] = self.v.detach().cpu().numpy()
        out["__global_step__"] = np.asarray(int(global_step), np.int64)
        np.savez(path, **out)

    def load_state(self, path):
        """-> global_step.  Raises if the container does not hold exactly this store's variables."""
        z = np.load(path)
        names = {k.replace("|", "/"): k for k in z.files if not k.startswith("__")}
        missing = [n for n in list(self.layout) + list(self.stats) if n not in names]
        if missing:
            raise KeyError("training state %r lacks %d variables, e.g. %s" % (path, len(missing), missing[:3]))
        flat = np.zeros(self.n, np.float32)
        for name, (o, shp) in self.layout.items():
            a = np.asarray(z[names[name]], np.float32)
            if tuple(a.shape) != tuple(shp):
                raise ValueError("training state %r: %s has shape %s, expected %s" % (path, name, a.shape, shp))
            flat[o:o + a.size] = a.reshape(-1)
        self.p.copy_(self.ops.to_device(flat))
        for name, t in self.stats.items():
            t.copy_(self.ops.to_device(np.asarray(z[names[name]], np.float32).reshape(-1)))
        if "__adam_m__" in z.files:
            if int(np.asarray(z["__adam_m__"]).size) != self.n:
                raise ValueError("training state %r holds Adam slots of %d elements, this store's flat layout has %d (written by a "
                                 "version with a different variable alignment): resume from the variables only"
                                 % (path, int(np.asarray(z["__adam_m__"]).size), self.n))
            self.m.copy_(self.ops.to_device(np.asarray(z["__adam_m__"], np.float32)))
            self.v.copy_(self.ops.to_device(np.asarray(z["__adam_v__"], np.float32)))
            return int(z["__global_step__"])
        self.m.zero_()
        self.v.zero_()
        return 0

    def export(self):
        """name -> float32 ndarray (the weight-container format of weights.py), moving statistics included."""
        out = OrderedDict()
        for name, (o, shp) in self.layout.items():
            out[name] = self.P[name].detach().cpu().numpy().reshape(shp).copy()
        for name, t in self.stats.items():
            out[name] = t.detach().cpu().numpy().copy()
        return out
