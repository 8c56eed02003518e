"""``Tacotron(hparams).initialize(...)`` of the reference (TacotronModel/modules/Tacotron.py:37-349)
over the B200 C-ABI extension.

The reference builds one TF graph "tower" per GPU inside a single process and evaluates it later with
``session.run``; this class executes eagerly: ``initialize`` splits the packed inputs per tower with
``split_infos`` exactly as Tacotron.py:92-130 does, runs each tower through
encoder -> persistent decoder -> postnet on the process's GPU (one process per GPU is the scaling
model, see sharding.py), and leaves the results in the reference's attributes, as lists with one entry
per tower: ``tower_decoder_output`` / ``tower_mel_outputs`` [B,F,80], ``tower_alignments`` [B,T,S],
``tower_stop_token_prediction`` [B,F] (Tacotron.py:314-317) and ``encoder_outputs`` [B,T,768] (:224).

Modes: inference (free running, TacoTestHelper) and GTA / evaluation
(teacher-forced forward, TacoTrainingHelper with ratio 1.0), and training: ``is_training=True`` runs the
train-mode forward of every tower (dropout, batch statistics, Bernoulli zoneout, stop logits, reference
encoder on the mel targets) through the training primitives (train/step.py); ``add_loss`` exposes the
reference's loss attributes and ``add_optimizer`` an ``optimize()`` callable that runs the hand-derived
backward pass of every tower, averages the tower gradients, clips and applies Adam (Tacotron.py:351-540).
"""
import numpy as np

from ...Utils.Helpers import TacoTestHelper, TacoTrainingHelper
from ...Utils.Utils import split_func
from . import Attention, Decoder, Encoder, Postnet
from .Encoder import FusedStageError


def reference_encoder(inputs, hparams, is_training=True, activation='relu'):
    """Tacotron.py:12-29 (6 x conv2d 3x3 s2 + BN + ReLU -> GRU(128) -> Dense tanh): mel [B, Tm, 80] -> [B, 128].  Standalone
    form on the device primitives, inference mode (moving statistics); inside synthesis it is part of ``taco_encode``, in training
    (batch statistics) part of train/step.py."""
    if is_training:
        raise FusedStageError("training-mode reference_encoder (batch statistics) runs inside train/step.py")
    from . import _standalone as S
    return S.reference_encoder(hparams, inputs)


class Tacotron(object):
    def __init__(self, hparams):
        self.hparams = hparams
        self.tower_decoder_output = []
        self.tower_alignments = []
        self.tower_stop_token_prediction = []
        self.tower_mel_outputs = []
        self.tower_linear_outputs = []
        self.tower_ref_output = []

    def initialize(self, inputs, input_lengths, mel_targets=None, stop_token_targets=None, linear_targets=None,
                   targets_lengths=None, GTA=False, global_step=None, is_training=False, is_evaluating=False,
                   split_infos=None, reference_mel=None, gst_weights=None, max_iters=None, prenet_masks=None,
                   philox_seed=None):
        """Arguments up to ``reference_mel`` are the reference's (Tacotron.py:46-48), NumPy arrays or
        tensors in place of placeholders.  Extra, optional: ``gst_weights`` [heads, tokens] (the
        injected stand-in for the ``tf.random_uniform`` style weights of Tacotron.py:216-222 when no
        reference mel is given), ``max_iters`` (default hparams.max_iters), ``prenet_masks`` /
        ``philox_seed`` (the always-on prenet dropout, Decoder.py:350-351)."""
        hp = self.hparams
        # ---- argument checks, Tacotron.py:65-77 ----
        if mel_targets is None and stop_token_targets is not None:
            raise ValueError('no mel targets were provided but token_targets were given')
        if mel_targets is not None and stop_token_targets is None and not GTA:
            raise ValueError('Mel targets are provided without corresponding token_targets')
        if not GTA and hp.predict_linear == True and linear_targets is None and is_training:  # noqa: E712
            raise ValueError('Model is set to use post processing to predict linear spectrograms in training but no linear targets given!')
        if GTA and linear_targets is not None:
            raise ValueError('Linear spectrogram prediction is not supported in GTA mode!')
        if is_training and hp.mask_decoder and targets_lengths is None:
            raise RuntimeError('Model set to mask paddings but no targets lengths provided for the mask!')
        if is_training and is_evaluating:
            raise RuntimeError('Model can not be in training and evaluation modes at the same time!')
        import torch
        from ... import runtime
        from ...synthetic import gst_random_weights
        eng = None if is_training else runtime.engine(hp)
        G = hp.tacotron_num_gpus
        inputs = np.asarray(inputs.cpu() if hasattr(inputs, "cpu") else inputs)
        input_lengths = np.asarray(input_lengths.cpu() if hasattr(input_lengths, "cpu") else input_lengths).astype(np.int32).reshape(-1)
        if split_infos is None:
            if G != 1:
                raise ValueError("split_infos is required when tacotron_num_gpus > 1")
            split_infos = np.asarray([[inputs.shape[1], 0 if mel_targets is None else np.shape(mel_targets)[1], 0, 0]], np.int32)
        split_infos = np.asarray(split_infos, np.int32)
        if split_infos.shape[0] != G:
            raise ValueError("split_infos has %d rows for %d towers" % (split_infos.shape[0], G))
        # ---- tower split, Tacotron.py:92-130 (split_func on the host; per-device batch = inputs.shape[0]) ----
        batch_size = inputs.shape[0]
        if input_lengths.shape[0] != batch_size * G:
            raise ValueError("input_lengths has %d entries, expected %d towers x %d" % (input_lengths.shape[0], G, batch_size))
        tower_input_lengths = np.split(input_lengths, G)
        p_inputs = split_func(inputs, split_infos[:, 0])
        p_mel = None
        if mel_targets is not None:
            mel_np = np.asarray(mel_targets.cpu() if hasattr(mel_targets, "cpu") else mel_targets, np.float32)
            p_mel = split_func(mel_np, split_infos[:, 1])
        self.tower_inputs = [np.ascontiguousarray(p_inputs[i].reshape(batch_size, -1)) for i in range(G)]
        self.tower_input_lengths = tower_input_lengths
        self.tower_mel_targets = [] if p_mel is None else [np.ascontiguousarray(p_mel[i].reshape(batch_size, -1, hp.num_mels)) for i in range(G)]
        self.tower_targets_lengths = None if targets_lengths is None else np.split(np.asarray(targets_lengths), G)
        self.tower_stop_token_targets = []
        self.tower_linear_targets = []
        if is_training:
            return self._initialize_training(stop_token_targets, split_infos, philox_seed, global_step)
        if reference_mel is not None:
            reference_mel = torch.as_tensor(np.asarray(reference_mel, np.float32))
        elif gst_weights is None:
            gst_weights = gst_random_weights(hp)
        seed = hp.tacotron_random_seed if philox_seed is None else philox_seed

        self.tower_decoder_output, self.tower_alignments = [], []
        self.tower_stop_token_prediction, self.tower_mel_outputs = [], []
        tower_encoder_outputs = []
        self.tower_steps = []
        for i in range(G):
            ids = torch.from_numpy(self.tower_inputs[i].astype(np.int32))
            lens = torch.from_numpy(tower_input_lengths[i].astype(np.int32))
            # -- encoder (Tacotron.py:179-224) --
            enc = Encoder.EncoderCell(Encoder.EncoderConvolution(is_training, hparams=hp, scope='encoder_convolutions'),
                                      Encoder.EncoderLSTM(is_training, size=hp.enc_lstm_hidden_size, zoneout=hp.zoneout_rate,
                                                          scope='encoder_LSTM'))
            enc(ids, lens, reference_mel=reference_mel, gst_weights=gst_weights)
            encoder_outputs = enc.memory
            # -- decoder (Tacotron.py:229-284) --
            prenet = Decoder.Prenet(is_training, layers_sizes=hp.prenet_layers, drop_rate=hp.dropout_rate, scope='decoder_prenet')
            attention_mechanism = Attention.LocationSensitiveAttention(
                num_units=hp.attention_dim, memory=encoder_outputs, hparams=hp, mask_encoder=hp.mask_encoder,
                memory_sequence_length=lens, smoothing=hp.smoothing, cumulate_weights=hp.cumulative_weights)
            decoder_lstm = Decoder.DecoderRNN(is_training=is_training, layers=2, size=hp.decoder_lstm_units,
                                              zoneout=hp.zoneout_rate, scope='decoder_LSTM')
            frame_projection = Decoder.FrameProjection(shape=hp.num_mels * hp.outputs_per_step, scope='linear_transform_projection')
            stop_projection = Decoder.StopProjection(is_training, shape=hp.outputs_per_step, scope='stop_token_projection')
            decoder_cell = Decoder.TacotronDecoderCell(prenet, attention_mechanism, decoder_lstm, frame_projection, stop_projection)
            if is_evaluating or GTA:
                tgt = torch.from_numpy(self.tower_mel_targets[i])
                helper = TacoTrainingHelper(batch_size, tgt, hp, GTA, is_evaluating, global_step)
            else:
                helper = TacoTestHelper(batch_size, hp)
            decoder_init_state = decoder_cell.zero_state(batch_size=batch_size)
            iters = hp.max_iters if max_iters is None else max_iters
            d = Decoder.CustomDecoder(decoder_cell, helper, decoder_init_state).decode(
                max_iters=None if (is_evaluating or GTA) else iters, prenet_masks=prenet_masks, philox_seed=seed)
            S = d["steps"]
            F = S * hp.outputs_per_step
            # -- postnet + projection + residual + clip (Tacotron.py:280-309) --
            decoder_output, mel_outputs = Postnet.Postnet(is_training, hparams=hp, scope='postnet_convolutions')(d["frames"][:, :F])
            tower_encoder_outputs.append(encoder_outputs)
            self.tower_decoder_output.append(decoder_output)
            self.tower_alignments.append(eng.alignments(d["align_hist"], S))    # [B, T, S] (Tacotron.py:312)
            self.tower_stop_token_prediction.append(d["stop"][:, :F])
            self.tower_mel_outputs.append(mel_outputs)
            self.tower_steps.append(S)
        self.encoder_outputs = tower_encoder_outputs[-1]   # the reference keeps the last tower's (Tacotron.py:224)
        self.tower_encoder_outputs = tower_encoder_outputs
        self.all_vars = sorted(runtime.weights(hp).keys())
        return self

    # ------------------------------------------------------------------ training (Tacotron.py:46-349 train mode, 351-540)
    def _initialize_training(self, stop_token_targets, split_infos, philox_seed, global_step):
        from ... import runtime
        hp = self.hparams
        G = hp.tacotron_num_gpus
        tr = runtime.trainer(hp)
        o = tr.ops
        batch_size = self.tower_inputs[0].shape[0]
        stop_np = np.asarray(stop_token_targets.cpu() if hasattr(stop_token_targets, "cpu") else stop_token_targets, np.float32)
        if split_infos.shape[1] > 2 and split_infos[:, 2].any() and not np.array_equal(split_infos[:, 2], split_infos[:, 1]):
            # the stop loss pairs one target with every mel frame (Tacotron.py:394-396): the reference's feeder pads len(mel) - 1
            # stop tokens to round_up(max + 1, r), which equals the mel width; anything else would mis-index the labels
            raise ValueError("split_infos: stop-token width %s != mel width %s" % (split_infos[:, 2].tolist(), split_infos[:, 1].tolist()))
        p_stop = split_func(stop_np, split_infos[:, 2] if split_infos.shape[1] > 2 and split_infos[:, 2].any() else split_infos[:, 1])
        self.tower_stop_token_targets = [np.ascontiguousarray(p_stop[i].reshape(batch_size, -1)) for i in range(G)]
        seed = (hp.tacotron_random_seed + 7919 * int(global_step or 0)) if philox_seed is None else philox_seed
        seed += 104729 * G * tr.rank          # every tower of every rank draws its own dropout / zoneout masks
        self._trainer, self._tapes, self._tower_losses = tr, [], []
        self.tower_decoder_output, self.tower_alignments = [], []
        self.tower_stop_token_prediction, self.tower_mel_outputs = [], []
        for i in range(G):
            out = tr.step_fn.forward(o.to_device(self.tower_inputs[i].astype(np.int32)), o.to_device(self.tower_input_lengths[i].astype(np.int32)),
                                     o.to_device(self.tower_mel_targets[i]), o.to_device(self.tower_stop_token_targets[i]), seed + 104729 * i)
            self._tapes.append(tr.step_fn.tape)
            self._tower_losses.append(out["losses"])
            self.tower_decoder_output.append(out["decoder_output"])
            self.tower_mel_outputs.append(out["mel_outputs"])
            self.tower_stop_token_prediction.append(out["stop_token_prediction"])      # logits in training (Decoder.py:408-409)
            self.tower_alignments.append(out["alignments"].permute(1, 2, 0).contiguous())   # [B, T, S] (Tacotron.py:312)
            self.encoder_outputs = out["encoder_outputs"]
        self.all_vars = sorted(tr.step_fn.store.layout.keys())
        return self

    def add_loss(self):
        """Tacotron.py:351-446: per-tower before / after / stop-token / regularization losses and their tower means."""
        if not getattr(self, "_tower_losses", None):
            raise RuntimeError("add_loss: initialize(is_training=True) must have been called")
        ls = [t.detach().cpu().numpy().astype(np.float64) for t in self._tower_losses]
        self.tower_before_loss = [float(l[0]) for l in ls]
        self.tower_after_loss = [float(l[1]) for l in ls]
        self.tower_stop_token_loss = [float(l[2]) for l in ls]
        self.tower_regularization_loss = [float(l[3]) for l in ls]
        self.tower_linear_loss = [0.0 for _ in ls]
        self.tower_loss = [float(l[0] + l[1] + l[2] + l[3]) for l in ls]
        n = float(len(ls))
        self.before_loss = sum(self.tower_before_loss) / n
        self.after_loss = sum(self.tower_after_loss) / n
        self.stop_token_loss = sum(self.tower_stop_token_loss) / n
        self.regularization_loss = sum(self.tower_regularization_loss) / n
        self.linear_loss = 0.0
        self.loss = sum(self.tower_loss) / n
        return self

    def add_optimizer(self, global_step):
        """Tacotron.py:448-515.  Sets ``learning_rate``, ``gradients`` (name -> view of the flat gradient buffer, valid
        after ``optimize()``) and ``optimize``: a callable that back-propagates every tower, averages the tower
        gradients (and, under torch.distributed, all-reduces them across ranks), clips to global norm 1 and applies
        Adam -- the eager equivalent of ``sess.run(model.optimize)``."""
        if not getattr(self, "_tapes", None):
            raise RuntimeError("add_optimizer: initialize(is_training=True) must have been called")
        tr = self._trainer
        st = tr.step_fn
        step_no = int(global_step or 0)
        self.learning_rate = st.learning_rate(step_no)
        self.gradients = st.store.G

        def optimize():
            acc = None
            for k, tape in enumerate(self._tapes):
                # the early exchange half may only start when the LAST tower's gradients are in the buffer
                hook, st.on_tail_ready = st.on_tail_ready, (st.on_tail_ready if len(self._tapes) == 1 else None)
                try:
                    st.backward(tape)
                finally:
                    st.on_tail_ready = hook
                if len(self._tapes) > 1:
                    if acc is None:
                        acc = st.store.g.clone()
                    else:
                        tr.ops.ew("add", acc, acc, st.store.g)
            if acc is not None:
                st.store.g.copy_(acc)
            towers = len(self._tapes) * tr.world
            st.apply(step_no, world=towers, all_reduce=tr._all_reduce if tr.world > 1 else None, ranks=tr.world)
            self._tapes = []
            return step_no + 1

        self.optimize = optimize
        return self
