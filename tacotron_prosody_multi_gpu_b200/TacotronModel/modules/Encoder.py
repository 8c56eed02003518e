"""Encoder modules of the reference (TacotronModel/modules/Encoder.py) as stage descriptors.

``EncoderConvolution`` (3 x conv1d k5 + BN + ReLU, Encoder.py:22-46,145-168), ``EncoderLSTM`` /
``ZoneoutLSTMCell`` (BiLSTM 256/dir, Encoder.py:48-142) keep the reference's constructor
signatures.  On the B200 path they are fused into ONE C-ABI call (``taco_encode``: embedding ->
conv stack on tcgen05 / SIMT kernels -> persistent BiLSTM kernel), so the objects describe the
stage and ``EncoderCell.__call__`` runs it; calling a sub-stage on its own raises
``FusedStageError`` naming the composite call that contains it.
"""
import numpy as np


class FusedStageError(NotImplementedError):
    """The stage exists only inside a fused sm_100a kernel; call the composite named in the message."""


class ZoneoutLSTMCell(object):
    """Encoder.py:79-142: LSTM cell with zoneout on c and h; the *output* is the raw h, the carried
    state is zoned (inference: (1-z)*new + z*old).  Gate order i, j, f, o; forget bias 1.0."""

    def __init__(self, num_units, is_training, zoneout_factor_cell=0., zoneout_factor_output=0.,
                 state_is_tuple=True, name=None):
        zm = min(zoneout_factor_output, zoneout_factor_cell)
        zs = max(zoneout_factor_output, zoneout_factor_cell)
        if zm < 0. or zs > 1.:
            raise ValueError('One/both provided Zoneout factors are not in [0, 1]')
        self._num_units = num_units
        self.is_training = is_training
        self._zoneout_cell = zoneout_factor_cell
        self._zoneout_outputs = zoneout_factor_output
        self.state_is_tuple = state_is_tuple
        self.name = name

    state_size = property(lambda self: (self._num_units, self._num_units))
    output_size = property(lambda self: self._num_units)

    _SCOPES = {"encoder_fw_LSTM": "ENC_LSTM_FW", "encoder_bw_LSTM": "ENC_LSTM_BW",
               "decoder_LSTM_1": ("DEC_LSTM", (0, 1)), "decoder_LSTM_2": ("DEC_LSTM", (1, 2))}

    def __call__(self, inputs, state, scope=None):
        """Encoder.py:118-142 standalone, inference mode: (raw h, (zoned c, zoned h)).  The cell's weights are found by its
        reference name (``encoder_fw_LSTM`` / ``encoder_bw_LSTM`` / ``decoder_LSTM_1`` / ``decoder_LSTM_2``)."""
        from ... import weights as W
        from ...Utils.Hyperparams import hparams
        from . import _standalone as S
        if self.is_training:
            raise FusedStageError("training-mode zoneout (sampled masks) runs inside train/step.py")
        ref = self._SCOPES.get(self.name)
        if ref is None:
            raise FusedStageError("ZoneoutLSTMCell(name=%r): no weights of that name (one of %s)" % (self.name, sorted(self._SCOPES)))
        wscope = getattr(W, ref) if isinstance(ref, str) else getattr(W, ref[0]) % ref[1]
        if self._zoneout_cell != self._zoneout_outputs:
            raise FusedStageError("the device cell applies one zoneout factor to c and h (the reference always passes equal factors)")
        return S.zoneout_lstm(hparams, inputs, state, wscope, self._zoneout_cell)


class EncoderConvolution(object):
    def __init__(self, is_training, hparams, activation='relu', scope='encoder_convolutions'):
        if is_training:
            raise FusedStageError("training-mode encoder convolutions (dropout + batch statistics) run inside train/step.py "
                                  "(Tacotron.initialize(is_training=True)); the standalone stage is inference-mode")
        self.is_training = is_training
        self.activation = activation
        self.scope = scope
        self.kernel_size = hparams.conv_kernel_size
        self.channels = hparams.conv_channels
        self.drop_rate = hparams.dropout_rate
        self.num_layers = hparams.conv_num_layers
        self.hparams = hparams

    def __call__(self, inputs):
        """Encoder.py:40-46 standalone on embedded inputs [B, T, 512] (inference: moving statistics, no dropout)."""
        from . import _standalone as S
        return S.encoder_convolution(self.hparams, inputs)


class EncoderLSTM(object):
    def __init__(self, is_training, size=256, zoneout=0.1, scope='encoder_lstm'):
        self.is_training = is_training
        self.size = size
        self.zoneout = zoneout
        self.scope = scope
        self.fw_cell = ZoneoutLSTMCell(size, is_training, zoneout, zoneout, name='encoder_fw_LSTM')
        self.bw_cell = ZoneoutLSTMCell(size, is_training, zoneout, zoneout, name='encoder_bw_LSTM')

    def __call__(self, inputs, input_lengths):
        """Encoder.py:68-77 standalone on conv outputs [B, T, 512] (inference-mode zoneout): [B, T, 2 * size]."""
        from ...Utils.Hyperparams import hparams
        from . import _standalone as S
        if self.is_training:
            raise FusedStageError("training-mode EncoderLSTM (sampled zoneout masks) runs inside train/step.py")
        return S.encoder_lstm(hparams, inputs, input_lengths, self.zoneout)


class EncoderCell(object):
    """Encoder.py:11-20.  ``inputs`` are the int32 token ids [B, T] (the embedding lookup of
    Tacotron.py:179-181 is the first kernel of the fused stage); returns the BiLSTM outputs
    [B, T, 2*size] as a device tensor, zero beyond each length."""

    def __init__(self, EncoderConvolution, EncoderLSTM):
        self.convolution_network = EncoderConvolution
        self.lstm_network = EncoderLSTM
        self.conv_output_shape = None

    def __call__(self, inputs, input_lengths=None, reference_mel=None, gst_weights=None):
        import torch
        from ... import runtime
        from ...synthetic import gst_random_weights
        hp = self.convolution_network.hparams
        ids = torch.as_tensor(np.asarray(inputs) if not isinstance(inputs, torch.Tensor) else inputs)
        if ids.dtype.is_floating_point:
            # the reference's call shape (Encoder.py:11-20): embedded inputs [B, T, 512] -> conv stack -> BiLSTM, composed from
            # the standalone stages (device primitives); token ids take the fused taco_encode path below
            conv_output = self.convolution_network(ids)
            self.conv_output_shape = tuple(conv_output.shape)
            if input_lengths is None:
                input_lengths = np.full(ids.shape[0], ids.shape[1], np.int32)
            return self.lstm_network(conv_output, input_lengths)
        if input_lengths is None:
            input_lengths = np.full(ids.shape[0], ids.shape[1], np.int32)
        lens = torch.as_tensor(np.asarray(input_lengths) if not isinstance(input_lengths, torch.Tensor) else input_lengths)
        eng = runtime.engine(hp)
        if reference_mel is None and gst_weights is None:
            gst_weights = gst_random_weights(hp)
        memory = eng.encode(ids, lens, reference_mel=reference_mel, gst_weights=gst_weights)
        self.conv_output_shape = (ids.shape[0], ids.shape[1], self.convolution_network.channels)
        self.memory = memory  # [B, T, 2*size + style_embed_depth] == `encoder_outputs` (Tacotron.py:224)
        return memory[..., : 2 * self.lstm_network.size]
