"""Decoder helpers of the reference (Utils/Helpers.py:6-128) as configuration objects.

In the reference a helper is evaluated by ``dynamic_decode`` once per step inside the TF while-loop.
Here the stepping rule is part of the persistent decoder kernel (taco_decode ``mode``):
``TacoTestHelper`` == TACO_MODE_FREE (next input = last predicted frame, batch-global stop when any
rounded stop token is 1, Helpers.py:36-59), ``TacoTrainingHelper`` == TACO_MODE_TEACHER with the
teacher-forcing ratio 1.0 (GTA / constant mode, Helpers.py:99-101,121-124).  The objects keep the
reference's constructor signatures and also offer the host-side ``initialize`` / ``sample`` /
``next_inputs`` for single-step use on NumPy arrays.
"""
import numpy as np


def _go_frames(batch_size, output_dim):
    """All-zero <GO> frames (Helpers.py:131-133)."""
    return np.zeros((batch_size, output_dim), np.float32)


class TacoTestHelper(object):
    mode = "free"

    def __init__(self, batch_size, hparams):
        self._batch_size = batch_size
        self._output_dim = hparams.num_mels
        self._reduction_factor = hparams.outputs_per_step
        self.stop_at_any = hparams.stop_at_any

    batch_size = property(lambda self: self._batch_size)
    token_output_size = property(lambda self: self._reduction_factor)
    sample_ids_dtype = np.int32
    sample_ids_shape = ()

    def initialize(self, name=None):
        return np.zeros(self._batch_size, bool), _go_frames(self._batch_size, self._output_dim)

    def sample(self, time, outputs, state, name=None):
        return np.zeros(self._batch_size, np.int32)

    def next_inputs(self, time, outputs, state, sample_ids, stop_token_prediction, name=None):
        """Helpers.py:36-59; note the reduction is over the whole [B, r] tensor (a scalar)."""
        fin = np.round(np.asarray(stop_token_prediction)).astype(bool)
        finished = np.any(fin) if self.stop_at_any else np.all(fin)
        return finished, np.asarray(outputs)[:, -self._output_dim:], state


class TacoTrainingHelper(object):
    mode = "teacher"

    def __init__(self, batch_size, targets, hparams, gta, evaluating, global_step):
        self._batch_size = batch_size
        self._output_dim = hparams.num_mels
        self._reduction_factor = hparams.outputs_per_step
        self._hparams = hparams
        self.gta = gta
        self.eval = evaluating
        self.global_step = global_step
        self.targets_full = targets
        r = self._reduction_factor
        self._targets = targets[:, r - 1::r, :]                 # every r-th frame is fed back
        self._lengths = np.full(batch_size, self._targets.shape[1], np.int32)
        self._ratio = float(hparams.tacotron_teacher_forcing_ratio)

    batch_size = property(lambda self: self._batch_size)
    token_output_size = property(lambda self: self._reduction_factor)
    sample_ids_dtype = np.int32
    sample_ids_shape = ()

    def initialize(self, name=None):
        if self.gta:
            self._ratio = 1.0
        elif self.eval and getattr(self._hparams, 'natural_eval', False):  # field absent from Hyperparams.py (quirk)
            self._ratio = 0.0
        elif self._hparams.tacotron_teacher_forcing_mode == 'scheduled':
            raise NotImplementedError("scheduled teacher forcing belongs to the training loop (SURVEY 8f row f1)")
        return np.zeros(self._batch_size, bool), _go_frames(self._batch_size, self._output_dim)

    @property
    def ratio(self):
        return self._ratio

    def sample(self, time, outputs, state, name=None):
        return np.zeros(self._batch_size, np.int32)

    def next_inputs(self, time, outputs, state, sample_ids, stop_token_prediction, name=None):
        finished = (time + 1 >= self._lengths)
        if self._ratio >= 1.0:
            nxt = np.asarray(self._targets[:, time, :])
        elif self._ratio <= 0.0:
            nxt = np.asarray(outputs)[:, -self._output_dim:]
        else:
            raise NotImplementedError("stochastic teacher forcing (0 < ratio < 1) is not on the B200 path")
        return finished, nxt, state
