"""Attention modules (TacotronModel/modules/Attention.py) as descriptors of fused stages.

``LocationSensitiveAttention`` (Attention.py:96-210, scores :369-401, context :345-366): keys =
memory . W_mem are computed once per utterance batch; every decoder step the persistent kernel
evaluates  e[t] = v_a . tanh(keys[t] + q + loc(cum)[t] + b_a), a masked softmax over t < length,
cum += a, context = a . values -- with the location conv (k31, 32 filters) and its 32 -> 128
projection folded into one Toeplitz contraction.  ``MultiheadAttention`` (Attention.py:214-341) is the
GST style-token attention, executed inside ``taco_encode``.
"""
from .Encoder import FusedStageError


class LocationSensitiveAttention(object):
    def __init__(self, num_units, memory, hparams, mask_encoder=True, memory_sequence_length=None,
                 smoothing=False, cumulate_weights=True, name='LocationSensitiveAttention'):
        if num_units != hparams.attention_dim:
            raise ValueError("num_units must equal hparams.attention_dim (%d)" % hparams.attention_dim)
        self.num_units = num_units
        self.values = memory                     # [B, T, 768] device tensor (zeroed beyond length by the encoder)
        self.hparams = hparams
        self.mask_encoder = mask_encoder
        self.memory_sequence_length = memory_sequence_length if mask_encoder else None
        self.smoothing = smoothing
        self._cumulate = cumulate_weights
        self.synthesis_constraint = False
        self.name = name
        self.batch_size = memory.shape[0]
        self.alignments_size = memory.shape[1]

    state_size = property(lambda self: self.alignments_size)

    def initial_alignments(self, batch_size, dtype=None):
        import torch
        return torch.zeros(batch_size, self.alignments_size, device=self.values.device)

    initial_state = initial_alignments

    def __call__(self, query, state):
        """Attention.py:169-210: (alignments [B, T], next_state) for query = LSTM output [B, 256] and state = previous
        (cumulative) alignments.  Standalone form on the device primitives (``taco_op_attn_fwd``); inside decoding the same step
        is part of the persistent kernel.  The context of ``_compute_attention`` (:345-366) is kept in ``self.last_context``."""
        from . import _standalone as S
        if getattr(self, "_keys", None) is None:
            self._values, self._keys, self._lens = S.memory_keys(self.hparams, self.values, self.memory_sequence_length, self.mask_encoder)
        hp = self.hparams
        if bool(hp.smoothing) != bool(self.smoothing) or bool(hp.cumulative_weights) != bool(self._cumulate):
            raise FusedStageError("smoothing / cumulate_weights must equal the hparams the device kernels were configured with")
        if bool(self.smoothing):
            raise FusedStageError("smoothing=True is implemented in the fused decoder kernels only (taco_decode)")
        align, nxt, ctx = S.location_sensitive_attention(hp, query, state, self._values, self._keys, self._lens)
        self.last_context = ctx
        return align, nxt

    keys = property(lambda self: self._ensure()[1])

    def _ensure(self):
        from . import _standalone as S
        if getattr(self, "_keys", None) is None:
            self._values, self._keys, self._lens = S.memory_keys(self.hparams, self.values, self.memory_sequence_length, self.mask_encoder)
        return self._values, self._keys, self._lens


class MultiheadAttention(object):
    def __init__(self, query, value, num_heads=4, attention_type='mlp_attention', num_units=None, normalize=True):
        self.query = query
        self.value = value
        self.num_heads = num_heads
        self.attention_type = attention_type
        self.num_units = num_units or query.shape[-1]
        self.normalize = normalize
        if attention_type != 'mlp_attention':
            raise FusedStageError("only mlp_attention (the reference default, Hyperparams.py:12) is on the B200 path")

    def multi_head_attention(self, hparams=None):
        """Attention.py:243-260: style embedding [B, 1, num_units * ...] = multi-head MLP attention of the query (reference
        encoder output [B, 1, 128] or [B, 128]) over value = tanh(style tokens).  Standalone form on ``taco_op_gst_fwd`` with
        the installed GST weights; inside synthesis the same computation is part of ``taco_encode``."""
        from ...Utils.Hyperparams import hparams as default_hp
        from . import _standalone as S
        hp = default_hp if hparams is None else hparams
        style, self.attention_weights = S.gst_multi_head_attention(hp, self.query, self.num_heads)
        return style
