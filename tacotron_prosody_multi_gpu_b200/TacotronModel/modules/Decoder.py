"""Decoder modules of the reference (TacotronModel/modules/Decoder.py) over the persistent kernel.

The reference evaluates ``TacotronDecoderCell.__call__`` (Decoder.py:246-292: prenet -> concat
context -> 2 x ZoneoutLSTM -> location-sensitive attention -> frame / stop projection) once per
``dynamic_decode`` iteration as ~40 TF kernel launches.  Here one persistent sm_100a kernel
(``taco_decode``) runs ALL steps with the weights resident on chip; the classes below keep the
reference's names and signatures:

* ``Prenet`` / ``DecoderRNN`` / ``FrameProjection`` / ``StopProjection``: stage descriptors (their
  parameters are read by ``TacotronDecoderCell``); standalone ``__call__`` raises ``FusedStageError``.
* ``TacotronDecoderCell``: ``zero_state`` + single-step ``__call__(inputs, state)`` (one
  ``taco_decode`` launch with S_max = 1 continuing from the state kept in the handle) and
  ``decode_all`` (the whole ``dynamic_decode`` loop in one launch).
* ``CustomDecoder``: ``initialize()`` / ``step(time, inputs, state)`` as in Decoder.py:108-146.
"""
import collections

import numpy as np

from ...Utils.Hyperparams import hparams
from .Encoder import FusedStageError


class CustomDecoderOutput(collections.namedtuple("CustomDecoderOutput", ("rnn_output", "token_output", "sample_id"))):
    pass


class TacotronDecoderCellState(collections.namedtuple(
        "TacotronDecoderCellState", ("cell_state", "attention", "time", "alignments", "alignment_history"))):
    """Decoder.py:419-438.  cell_state = ((c1, h1), (c2, h2)) zoned LSTM states [B,256]; attention =
    previous context [B,768]; alignments = CUMULATIVE alignments [B,T]; alignment_history = list of
    per-step alignments [B,T] (the TensorArray)."""

    def replace(self, **kwargs):
        return super(TacotronDecoderCellState, self)._replace(**kwargs)


class Prenet(object):
    def __init__(self, is_training, layers_sizes=[256, 256], drop_rate=0.5, activation='relu', scope=None):
        self.drop_rate = drop_rate
        self.layers_sizes = layers_sizes
        self.activation = activation
        self.is_training = is_training
        self.scope = 'prenet' if scope is None else scope

    def __call__(self, inputs, philox_seed=None, step=0):
        """Decoder.py:342-352 standalone (device primitives): the dropout bits are the decoder's Philox stream for
        (seed, step), so the result equals the fused kernel's prenet output of that step."""
        from . import _standalone as S
        if list(self.layers_sizes) != list(hparams.prenet_layers) or self.drop_rate != hparams.dropout_rate:
            raise FusedStageError("Prenet layers / drop_rate must equal the hparams the weights were created with")
        return S.prenet(hparams, inputs, hparams.tacotron_random_seed if philox_seed is None else philox_seed, step)


class DecoderRNN(object):
    def __init__(self, is_training, layers=2, size=256, zoneout=0.1, scope='decoder_rnn'):
        if layers != 2:
            raise FusedStageError("the persistent decoder kernel implements the reference's 2-layer DecoderRNN")
        self.is_training = is_training
        self.layers = layers
        self.size = size
        self.zoneout = zoneout
        self.scope = scope

    def __call__(self, input, states):
        """Decoder.py:319-321 standalone, inference mode: (raw LSTM-2 output, ((c1, h1), (c2, h2)) zoned)."""
        from . import _standalone as S
        if self.is_training:
            raise FusedStageError("training-mode DecoderRNN (sampled zoneout masks) runs inside train/step.py")
        return S.decoder_rnn(hparams, input, states, self.zoneout)


class FrameProjection(object):
    def __init__(self, shape=hparams.num_mels, activation=None, scope=None):
        self.shape = shape
        self.activation = activation
        self.scope = 'Linear_projection' if scope is None else scope

    def __call__(self, inputs):
        """Decoder.py:371-380 standalone: Dense(num_mels * r) of [LSTM output ; context] [B, 1024]."""
        from ... import weights as W
        from . import _standalone as S
        return S.dense(hparams, inputs, W.FRAME_PROJ)


class StopProjection(object):
    def __init__(self, is_training, shape=1, activation='sigmoid', scope=None):
        self.is_training = is_training
        self.shape = shape
        self.activation = activation
        self.scope = 'stop_token_projection' if scope is None else scope

    def __call__(self, inputs):
        """Decoder.py:404-410 standalone: Dense(r) then sigmoid unless is_training (logits feed the loss)."""
        import torch
        from ... import weights as W
        from . import _standalone as S
        y = S.dense(hparams, inputs, W.STOP_PROJ)
        return y if self.is_training else torch.sigmoid(y)


class TacotronDecoderCell(object):
    def __init__(self, prenet, attention_mechanism, rnn_cell, frame_projection, stop_projection):
        self._prenet = prenet
        self._attention_mechanism = attention_mechanism
        self._cell = rnn_cell
        self._frame_projection = frame_projection
        self._stop_projection = stop_projection
        self._attention_layer_size = attention_mechanism.values.shape[-1]
        self._hp = attention_mechanism.hparams
        self._live_state = None  # identity of the state the handle currently holds

    output_size = property(lambda self: self._frame_projection.shape)

    @property
    def state_size(self):
        return TacotronDecoderCellState(cell_state=((self._cell.size,) * 2,) * self._cell.layers, time=(),
                                        attention=self._attention_layer_size,
                                        alignments=self._attention_mechanism.alignments_size, alignment_history=())

    def _engine(self):
        from ... import runtime
        return runtime.engine(self._hp)

    def _lens(self):
        import torch
        am = self._attention_mechanism
        if am.memory_sequence_length is None:
            return torch.full((am.batch_size,), am.alignments_size, dtype=torch.int32)
        return torch.as_tensor(np.asarray(am.memory_sequence_length.cpu()) if hasattr(am.memory_sequence_length, "cpu")
                               else np.asarray(am.memory_sequence_length)).to(torch.int32)

    def zero_state(self, batch_size, dtype=None):
        """Decoder.py:212-244."""
        import torch
        am = self._attention_mechanism
        if batch_size != am.batch_size:
            raise ValueError("Non-matching batch sizes between the memory (encoder output) and the requested batch size.")
        dev = am.values.device
        z = lambda n: torch.zeros(batch_size, n, device=dev)
        return TacotronDecoderCellState(cell_state=((z(self._cell.size), z(self._cell.size)),) * 2, time=0,
                                        attention=z(self._attention_layer_size), alignments=z(am.alignments_size),
                                        alignment_history=[])

    def __call__(self, inputs, state, prenet_mask=None, philox_seed=0):
        """One decoder step.  ``state`` must be ``zero_state(...)`` or the state returned by the previous
        call (the recurrent state lives in the C-ABI handle; TacotronDecoderCellState carries copies).
        Returns ((frames [B, r*80], stop [B, r]), next_state)."""
        import torch
        eng = self._engine()
        am = self._attention_mechanism
        B, t = am.batch_size, int(state.time)
        if t > 0 and state is not self._live_state:
            raise FusedStageError("cell(inputs, state): only the state returned by the previous step can be continued "
                                  "(the handle owns the live recurrent state)")
        r, M = self._hp.outputs_per_step, self._hp.num_mels
        # feed `inputs` through the teacher-forcing slot of step t: x(t) = targets[:, r*t - 1] (Helpers.py:121-124)
        tgt = torch.zeros(B, r * (t + 1), M, device=am.values.device)
        if t > 0:
            tgt[:, r * t - 1] = torch.as_tensor(inputs, dtype=torch.float32).to(am.values.device)
        masks = None if prenet_mask is None else np.asarray(prenet_mask, np.uint8).reshape(1, B, 2, self._prenet.layers_sizes[0])
        d = eng.decode(am.values, self._lens(), mel_targets=tgt, max_iters=None, prenet_masks=masks, philox_seed=philox_seed,
                       reset_state=(t == 0), logits_only=bool(self._stop_projection.is_training), single_step_at=t)
        st = eng.decoder_state(B, am.alignments_size)
        nxt = TacotronDecoderCellState(cell_state=((st["c1"], st["h1"]), (st["c2"], st["h2"])), time=t + 1,
                                       attention=st["attention"], alignments=st["alignments"],
                                       alignment_history=list(state.alignment_history) + [d["align_hist"][0]])
        self._live_state = nxt
        return (d["frames"].reshape(B, r * M), d["stop"]), nxt

    def decode_all(self, helper, max_iters=None, prenet_masks=None, philox_seed=0):
        """dynamic_decode(CustomDecoder(cell, helper)) in ONE persistent-kernel launch (Tacotron.py:262-284)."""
        eng = self._engine()
        am = self._attention_mechanism
        tgt = helper.targets_full if helper.mode == "teacher" else None
        if helper.mode == "teacher":
            helper.initialize()
            if helper.ratio != 1.0:
                raise FusedStageError("teacher forcing ratio %.2f: the kernel implements ratio 1.0 (GTA / constant mode) "
                                      "and free running" % helper.ratio)
        return eng.decode(am.values, self._lens(), mel_targets=tgt, max_iters=max_iters, prenet_masks=prenet_masks,
                          philox_seed=philox_seed, logits_only=bool(self._stop_projection.is_training))


class CustomDecoder(object):
    """Decoder.py:32-146: (cell, helper, initial_state) -> ``initialize`` / ``step``."""

    def __init__(self, cell, helper, initial_state, output_layer=None):
        if output_layer is not None:
            raise TypeError("output_layer is not used on the Tacotron path (Tacotron.py:271-273)")
        self._cell = cell
        self._helper = helper
        self._initial_state = initial_state
        self._output_layer = output_layer

    batch_size = property(lambda self: self._helper.batch_size)

    @property
    def output_size(self):
        return CustomDecoderOutput(rnn_output=self._cell.output_size, token_output=self._helper.token_output_size, sample_id=())

    def initialize(self, name=None):
        return self._helper.initialize() + (self._initial_state,)

    def step(self, time, inputs, state, name=None):
        (cell_outputs, stop_token), cell_state = self._cell(inputs, state)
        sample_ids = self._helper.sample(time=time, outputs=cell_outputs, state=cell_state)
        finished, next_inputs, next_state = self._helper.next_inputs(
            time=time, outputs=cell_outputs.cpu().numpy(), state=cell_state, sample_ids=sample_ids,
            stop_token_prediction=stop_token.cpu().numpy())
        return CustomDecoderOutput(cell_outputs, stop_token, sample_ids), next_state, next_inputs, finished

    def decode(self, max_iters=None, **kw):
        """The whole loop (``dynamic_decode(self, maximum_iterations=max_iters)``) in one launch."""
        return self._cell.decode_all(self._helper, max_iters=max_iters, **kw)
