"""``Synthesizer`` of the reference (Utils/Tacotron_synthesizer.py:13-182) over the B200 engine.

``load`` replaces graph construction + ``saver.restore`` by installing a weight container
(``.npz`` written by weights.save, names = reference variable names) or, when no checkpoint exists,
the TF-default random init with the reference's seed.  ``synthesize`` keeps the reference's contract:
pads the text list to a multiple of ``tacotron_num_gpus`` by repeating the last entry (:65-69), packs
per-device batches along the time axis with ``split_infos`` (:77-84,99-105), runs the model, trims each
mel at the first rounded stop token (:112, feeder :92-95), clips to the T2 output range (:131-135) and
writes ``speech-mel-{basename}.npy`` float32 [len, 80] (:166-167).  With ``log_dir`` the Griffin-Lim wav of every
mel (:170-175) comes from the device path in Utils/AudioProcessing; the matplotlib plots
(:176-181) are emitted only if the reference's plot helpers are importable.
"""
import os

import numpy as np

from ..TacotronModel.modules.Tacotron import Tacotron
from .Infolog import log
from .AudioProcessing.AudioPreprocess import inv_preemphasis, mel_to_audio_serie, save_wav
from .Tacotron_feeder import _get_output_lengths, _prepare_inputs, _prepare_targets


class TextFrontEndUnavailable(RuntimeError):
    """The reference's Korean G2P (Utils/TextProcessing, pure Python) is not importable."""


def _import_reference_g2p():
    from Utils.TextProcessing.HangulUtils import hangul_to_sequence  # the reference's own module, when its root is on sys.path
    return hangul_to_sequence


def _default_text_to_sequence(text, hparams, _import=_import_reference_g2p):
    """Korean G2P front end of the reference (HangulUtils.hangul_to_sequence:49-65 -> KoG2P; pure Python, used UNCHANGED when the
    reference's root is on sys.path).  There is no substitute tokeniser on this path: without the front end a text input raises
    ``TextFrontEndUnavailable`` (pass token-id arrays instead, or opt in to ``synthetic_text_to_sequence`` explicitly with
    ``Synthesizer(synthetic_tokens=True)`` / ``TACO_SYNTHETIC_TOKENS=1``)."""
    try:
        hangul_to_sequence = _import()
    except ImportError as e:
        raise TextFrontEndUnavailable(
            "text input needs the reference's Utils/TextProcessing package on sys.path (%s); pass token-id arrays, or request "
            "synthetic tokens explicitly (Synthesizer(synthetic_tokens=True) or TACO_SYNTHETIC_TOKENS=1)" % e) from e
    return np.asarray(hangul_to_sequence(dir=hparams.base_dir, hangul_text=text, hangul_type=hparams.hangul_type))


def synthetic_text_to_sequence(text, hparams):
    """Explicit opt-in stand-in for benchmarking without the text assets: one deterministic id in [2, n_symbols) per non-blank
    byte, EOS = 1 appended.  NOT a tokeniser of the reference; never selected implicitly."""
    from ..hparams import num_symbols
    n = num_symbols(hparams)
    ids = [2 + (b * 31 + i) % (n - 2) for i, b in enumerate(text.encode("utf-8")) if not chr(b).isspace()]
    return np.asarray(ids + [1], np.int32)


class Synthesizer(object):
    def __init__(self, synthetic_tokens=None):
        if synthetic_tokens is None:
            synthetic_tokens = os.environ.get("TACO_SYNTHETIC_TOKENS", "0") == "1"
        self.text_to_sequence = synthetic_text_to_sequence if synthetic_tokens else _default_text_to_sequence
        if synthetic_tokens:
            log('synthetic token ids requested explicitly: text is NOT passed through the G2P front end')

    def load(self, checkpoint_path, hparams, GTA=False, reference_mel=None, model_name='Tacotron'):
        from .. import runtime, weights
        log('Constructing model: %s' % model_name)
        self.model = Tacotron(hparams)
        self.GTA = GTA
        self.hparams = hparams
        self.reference_mel = None if reference_mel is None else np.asarray(reference_mel, np.float32).reshape(1, -1, hparams.num_mels)
        if checkpoint_path is not None:
            # saver.restore of a missing checkpoint fails in the reference (:49-50); so does this
            if not os.path.isfile(str(checkpoint_path)):
                raise FileNotFoundError('weight container %r does not exist (pass checkpoint_path=None for random-init weights)' % (checkpoint_path,))
            log('Loading checkpoint: %s' % checkpoint_path)
            runtime.set_weights(weights.load(checkpoint_path))
        else:
            log('checkpoint_path=None: random-init weights, seed %d' % hparams.tacotron_random_seed)
        runtime.engine(hparams)   # builds the handle, uploads + packs the weights (fails loudly without a GPU / .so)
        return self

    def synthesize(self, texts, basenames, out_dir, log_dir, mel_filenames, gst_weights=None, max_iters=None):
        hparams = self.hparams
        G = hparams.tacotron_num_gpus
        texts = list(texts)
        basenames = None if basenames is None else list(basenames)
        mel_filenames = None if mel_filenames is None else list(mel_filenames)
        n_real = len(texts)
        while len(texts) % G != 0:                       # last-run scenario (:65-69)
            texts.append(texts[-1])
            if basenames is not None:
                basenames.append(basenames[-1])
            if mel_filenames is not None:
                mel_filenames.append(mel_filenames[-1])
        seqs = [np.asarray(t) if not isinstance(t, str) else np.asarray(self.text_to_sequence(t, hparams)) for t in texts]
        input_lengths = [len(s) for s in seqs]
        per_dev = len(seqs) // G
        blocks, split_infos = [], []
        for g in range(G):
            padded, t_in = _prepare_inputs(seqs[per_dev * g: per_dev * (g + 1)])
            blocks.append(padded)
            split_infos.append([t_in, 0, 0, 0])
        feed = dict(inputs=np.concatenate(blocks, axis=1), input_lengths=np.asarray(input_lengths, np.int32))
        if self.GTA:
            np_targets = [np.load(f) for f in mel_filenames]
            assert len(np_targets) == len(texts)
            tblocks = []
            for g in range(G):
                padded, t_mel = _prepare_targets(np_targets[per_dev * g: per_dev * (g + 1)], hparams.outputs_per_step)
                tblocks.append(padded)
                split_infos[g][1] = t_mel
            feed["mel_targets"] = np.concatenate(tblocks, axis=1)
        feed["split_infos"] = np.asarray(split_infos, np.int32)

        m = self.model.initialize(GTA=self.GTA, reference_mel=self.reference_mel, gst_weights=gst_weights, max_iters=max_iters, **feed)
        # linearise towers (:112-117) -- ONE device->host read per fetched tensor per tower
        mels = [mel for tower in m.tower_mel_outputs for mel in tower.cpu().numpy()]
        alignments = [a for tower in m.tower_alignments for a in tower.cpu().numpy()]
        stop_tokens = [s for tower in m.tower_stop_token_prediction for s in tower.cpu().numpy()]
        target_lengths = _get_output_lengths(stop_tokens)
        mels = [mel[:n, :] for mel, n in zip(mels, target_lengths)]
        lo, hi = (-hparams.max_abs_value, hparams.max_abs_value) if hparams.symmetric_mels else (0, hparams.max_abs_value)
        mels = [np.clip(mel, lo, hi) for mel in mels]
        self.last_alignments, self.last_stop_tokens, self.last_mels = alignments, stop_tokens, mels
        if basenames is None:
            # Tacotron_synthesizer.py:137-149 writes temp.wav and shells out to a player; the waveform is produced and
            # saved, playback is left to the caller
            wav = mel_to_audio_serie(mels[0].T, hparams)
            save_wav(wav, 'temp.wav', sr=hparams.sample_rate)
            return
        saved_mels_paths, speaker_ids = [], []
        for i, mel in enumerate(mels):
            speaker_ids.append('<no_g>')
            mel_filename = os.path.join(out_dir, 'speech-mel-{}.npy'.format(basenames[i]))
            np.save(mel_filename, mel.astype(np.float32), allow_pickle=False)
            saved_mels_paths.append(mel_filename)
            if log_dir is not None:
                self._emit_logs(mel, alignments[i], texts[i], basenames[i], log_dir)
        return saved_mels_paths, speaker_ids

    _warned = False

    def _emit_logs(self, mel, alignment, text, basename, log_dir):
        """:170-181: Griffin-Lim wav (device path, include/taco_b200_audio.h) and, when the reference's matplotlib
        helpers are importable, the alignment / mel plots."""
        hp = self.hparams
        os.makedirs(os.path.join(log_dir, 'wavs'), exist_ok=True)
        if mel.shape[0] >= 5:   # librosa's reflect padding needs hop * (frames - 1) > n_fft / 2
            wav = inv_preemphasis(mel_to_audio_serie(mel.T, hp), hp.preemphasis, hp.preemphasize)
            save_wav(wav, os.path.join(log_dir, 'wavs/speech-wav-{}-mel.wav'.format(basename)), sr=hp.sample_rate)
        try:
            from Utils.Plot import plot_alignment, plot_spectrogram  # the reference's own module, if on sys.path
        except Exception:
            if not Synthesizer._warned:
                log('plot emission skipped (the reference\'s matplotlib helpers are not importable here)')
                Synthesizer._warned = True
            return
        os.makedirs(os.path.join(log_dir, 'plots'), exist_ok=True)
        plot_alignment(alignment, os.path.join(log_dir, 'plots/speech-alignment-{}.png'.format(basename)), info=str(text), split_title=True)
        plot_spectrogram(mel, os.path.join(log_dir, 'plots/speech-mel-{}.png'.format(basename)), info=str(text), split_title=True)
