"""TF-free stand-in for the reference's global ``tf.contrib.training.HParams`` object.

Mirrors the fields the mel-generation hot path reads from
``Utils/Hyperparams.py:4-239`` (network dims :43-63, GST :7-14, run params
:157-172) with the same names and defaults, plus ``.parse("a=b,c=[1,2]")``
(``TacotronModel/synthesize.py:146``) and ``.values()``
(``Utils/Utils.py: hparams_to_string``).  Fields belonging to the vocoders /
audio preprocessing are carried only where a hot-path file touches them.
"""
from __future__ import annotations

import ast
import copy
import re


class HParams(object):
    """Minimal HParams: attribute bag + typed ``parse`` of ``name=value`` lists."""

    def __init__(self, **kwargs):
        object.__setattr__(self, "_fields", {})
        for k, v in kwargs.items():
            self.add_hparam(k, v)

    # -- tf.contrib.training.HParams surface ---------------------------------
    def add_hparam(self, name, value):
        if name in self._fields:
            raise ValueError("Hyperparameter name is reserved: %s" % name)
        self._fields[name] = value

    def set_hparam(self, name, value):
        if name not in self._fields:
            raise KeyError("Unknown hyperparameter: %s" % name)
        self._fields[name] = _cast_like(self._fields[name], value, name)

    def values(self):
        return dict(self._fields)

    def __getattr__(self, name):
        try:
            return object.__getattribute__(self, "_fields")[name]
        except KeyError:
            raise AttributeError(name)

    def __setattr__(self, name, value):
        if name in self._fields:
            self._fields[name] = value
        else:
            self.add_hparam(name, value)

    def __contains__(self, name):
        return name in self._fields

    def copy(self):
        return HParams(**copy.deepcopy(self._fields))

    def parse(self, values):
        """``"name=value,name2=[a,b]"`` typed by the existing field."""
        if not values:
            return self
        for name, raw in _split_assignments(values):
            if name not in self._fields:
                raise ValueError("Unknown hyperparameter type for %s" % name)
            self.set_hparam(name, raw)
        return self

    def __repr__(self):
        return "HParams(%s)" % ", ".join("%s=%r" % kv for kv in sorted(self._fields.items()))


_ASSIGN = re.compile(r"\s*([A-Za-z_][A-Za-z0-9_]*)\s*=\s*(\[[^\]]*\]|[^,]*)\s*(?:,|$)")


def _split_assignments(s):
    pos = 0
    out = []
    while pos < len(s):
        m = _ASSIGN.match(s, pos)
        if not m:
            raise ValueError("Malformed hyperparameter value: %s" % s[pos:])
        out.append((m.group(1), m.group(2).strip()))
        pos = m.end()
    return out


def _parse_scalar(like, raw, name):
    if isinstance(raw, str):
        txt = raw.strip()
        if isinstance(like, bool):
            if txt.lower() in ("true", "1"):
                return True
            if txt.lower() in ("false", "0"):
                return False
            raise ValueError("Could not parse bool hparam %s=%s" % (name, raw))
        if isinstance(like, int) and not isinstance(like, bool):
            return int(txt)
        if isinstance(like, float):
            return float(txt)
        if like is None:
            try:
                return ast.literal_eval(txt)
            except Exception:
                return txt
        return txt.strip("'\"")
    if isinstance(like, bool):
        return bool(raw)
    if isinstance(like, int) and not isinstance(like, bool):
        return int(raw)
    if isinstance(like, float):
        return float(raw)
    return raw


def _cast_like(like, raw, name):
    if isinstance(like, (list, tuple)):
        if isinstance(raw, str):
            txt = raw.strip()
            if not (txt.startswith("[") and txt.endswith("]")):
                raise ValueError("Must pass a list for multi-valued hparam %s" % name)
            items = [t for t in txt[1:-1].split(",") if t.strip() != ""]
        else:
            items = list(raw)
        proto = like[0] if len(like) else None
        vals = [_parse_scalar(proto, it, name) for it in items]
        return type(like)(vals) if isinstance(like, tuple) else vals
    return _parse_scalar(like, raw, name)


def default_hparams():
    """Defaults of ``Utils/Hyperparams.py`` for every field the hot path reads."""
    return HParams(
        # Global style tokens (Hyperparams.py:7-14)
        use_gst=True,
        num_gst=10,
        num_heads=4,
        style_embed_depth=256,
        reference_filters=[32, 32, 64, 64, 128, 128],
        reference_depth=128,
        style_att_type="mlp_attention",
        style_att_dim=128,
        # optimiser / loss (:17-40)
        tacotron_adam_beta1=0.9,
        tacotron_adam_beta2=0.999,
        tacotron_adam_epsilon=1e-6,
        mask_decoder=False,
        mask_encoder=True,
        tacotron_decay_learning_rate=True,
        tacotron_start_decay=40000,
        tacotron_decay_steps=18000,
        tacotron_decay_rate=0.5,
        tacotron_initial_learning_rate=1e-3,
        tacotron_final_learning_rate=1e-4,
        tacotron_teacher_forcing_mode="constant",
        tacotron_teacher_forcing_ratio=1.0,
        tacotron_teacher_forcing_init_ratio=1.0,
        tacotron_teacher_forcing_final_ratio=0.0,
        tacotron_teacher_forcing_start_decay=10000,
        tacotron_teacher_forcing_decay_steps=40000,
        tacotron_teacher_forcing_decay_alpha=None,
        tacotron_reg_weight=1e-6,
        tacotron_scale_regularization=False,
        # network (:43-63)
        conv_num_layers=3,
        conv_kernel_size=(5,),
        conv_channels=512,
        enc_lstm_hidden_size=256,
        zoneout_rate=0.1,
        dropout_rate=0.5,
        smoothing=False,
        attention_dim=128,
        attention_filters=32,
        attention_kernel=(31,),
        cumulative_weights=True,
        embedding_dim=512,
        postnet_kernel_size=(5,),
        postnet_channels=512,
        postnet_num_layers=5,
        prenet_layers=[256, 256],
        decoder_lstm_units=256,
        hangul_type=1,
        # mel / audio fields the path touches (:73-95)
        sample_rate=44100,
        predict_linear=False,
        symmetric_mels=True,
        max_abs_value=4.0,
        num_mels=80,
        num_freq=1025,
        clip_mels_length=True,
        max_mel_frames=1000,
        hop_size=275,
        n_fft=2048,
        win_size=1024,
        frame_shift_ms=None,
        fmin=25,
        fmax=7600,
        min_level_db=-100,
        ref_level_db=20,
        power=1.2,
        griffin_lim_iters=60,
        signal_normalization=True,
        allow_clipping_in_normalization=True,
        use_lws=False,
        preemphasize=False,
        preemphasis=0.97,
        gin_channels=-1,
        # run params (:157-172)
        tacotron_num_gpus=1,
        split_on_cpu=True,
        tacotron_synthesis_batch_size=1,
        tacotron_data_random_state=1324,
        outputs_per_step=2,
        tacotron_random_seed=45454,
        tacotron_swap_with_cpu=False,
        tacotron_batch_size=48,
        max_iters=3000,
        stop_at_any=True,
        clip_outputs=True,
        lower_bound_decay=0.1,
        tacotron_clip_gradients=True,
        base_dir="",
        join_output_audio=True,
        sentences=[],
    )


# the reference exposes one module-level object (Hyperparams.py:4)
hparams = default_hparams()

# type-1 symbol table size (Utils/TextProcessing/HangulUtils.py:13-25): 66 symbols,
# id 0 = pad, id 1 = EOS.  Types 2..5 have 72/56/56/42 symbols.
_SYMBOL_COUNTS = {1: 66, 2: 72, 3: 56, 4: 56, 5: 42}


def num_symbols(hp):
    return _SYMBOL_COUNTS.get(int(hp.hangul_type), _SYMBOL_COUNTS[5])
