#!/usr/bin/env python
"""Golden fixtures produced by the REFERENCE'S OWN code (not by the oracle).

The reference's TensorFlow graph cannot run here (TF 1.9), but its pure-Python pieces can.  This script
imports them from /root/reference (build container only) under a tiny stub for the modules they import at module scope
but do not need for the functions called here (``tensorflow`` -- only ``tf.contrib.training.HParams`` is real, an
attribute bag --, ``librosa``, ``matplotlib``), runs them on seeded ragged inputs and writes

    tests/golden/ref_feeder.npz       Utils/Tacotron_feeder.py:25-112  (_prepare_inputs/_targets/_token_targets, _round_up,
                                      _get_output_lengths incl. the half-to-even cases, _prepare_batch with split_infos for 1 and 2
                                      devices under a seeded np.random.shuffle) and Utils/Utils.py:28-35 (split_func)
    tests/golden/ref_hparams.json     every default of Utils/Hyperparams.py:4-239
    tests/golden/ref_text.json        HangulUtils.hangul_to_sequence ids (type 1, KoG2P) of the 16 lines of sentences.txt
                                      (Utils/TextProcessing/HangulUtils.py:49-65, KoG2P.py:153-202)
    tests/golden/ref_audio.npz        inv_preemphasis / preemphasis, _denormalize, _normalize, _db_to_amp, _amp_to_db
                                      (Utils/AudioProcessing/AudioPreprocess.py:39-48,440-484)

tests/test_ref_fixtures_cpu.py compares the repo's host side and the oracle's audio helpers with these files.  Nothing at test
or bench time reads /root/reference; the fixtures are small and committed.
"""
import json
import os
import sys
import types
from unittest import mock

import numpy as np

REF = os.environ.get("TACO_REFERENCE_ROOT", "/root/reference")
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
OUT = os.path.join(ROOT, "tests", "golden")


class _HParams(object):   # the whole of tf.contrib.training.HParams that the code below touches
    def __init__(self, **kw):
        self.__dict__.update(kw)

    def values(self):
        return dict(self.__dict__)


def _install_stubs():
    tf = mock.MagicMock(name="tensorflow")
    tf.contrib.training.HParams = _HParams
    sys.modules["tensorflow"] = tf
    for name in ("librosa", "librosa.filters", "librosa.display", "librosa.output", "matplotlib", "matplotlib.pyplot", "matplotlib.font_manager", "lws"):
        if name not in sys.modules:
            sys.modules[name] = mock.MagicMock(name=name)
    sys.path.insert(0, REF)


def _jsonable(v):
    if isinstance(v, (list, tuple)):
        return [_jsonable(x) for x in v]
    if isinstance(v, (np.integer,)):
        return int(v)
    if isinstance(v, (np.floating,)):
        return float(v)
    return v


def feeder_fixture():
    from Utils import Tacotron_feeder as RF
    from Utils.Hyperparams import hparams
    from Utils.Utils import split_func

    out = {}
    rng = np.random.default_rng(1324)
    tok_lens = [5, 9, 7, 4, 12, 3]
    mel_lens = [13, 26, 17, 8, 31, 6]          # odd and even maxima: both branches of _round_up
    seqs = [rng.integers(2, 66, n).astype(np.int32) for n in tok_lens]
    mels = [rng.standard_normal((n, hparams.num_mels)).astype(np.float32) for n in mel_lens]
    stops = [np.zeros(n - 1, np.float32) for n in mel_lens]     # the reference's convention: len(mel) - 1 zeros (Tacotron_preprocessor)
    for i, s in enumerate(seqs):
        out["seq_%d" % i] = s
    for i, m in enumerate(mels):
        out["mel_%d" % i] = m
    for r in (1, 2, 3):
        x, t = RF._prepare_inputs(seqs)
        out["inputs"] = x
        out["inputs_T"] = np.int64(t)
        m, f = RF._prepare_targets(mels, r)
        out["mel_targets_r%d" % r] = m
        out["mel_targets_F_r%d" % r] = np.int64(f)
        s, fs = RF._prepare_token_targets(stops, r)
        out["token_targets_r%d" % r] = s
        out["token_targets_F_r%d" % r] = np.int64(fs)
    # subsets whose maximum is even / odd
    m, f = RF._prepare_targets(mels[:2], 2)
    out["mel_targets_even_max"] = m
    s, fs = RF._prepare_token_targets(stops[:2], 2)
    out["token_targets_even_max"] = s
    out["round_up_x"] = np.arange(0, 40)
    out["round_up_r2"] = np.asarray([RF._round_up(int(x), 2) for x in range(40)])
    out["round_up_r5"] = np.asarray([RF._round_up(int(x), 5) for x in range(40)])
    out["pad_values"] = np.asarray([RF._pad, RF._target_pad, RF._token_pad], np.float64)
    # stop tokens: exact halves (np.round is half-to-even: 0.5 -> 0, 1.5 -> 2), first crossing mid-row, no crossing, crossing at 0
    st = np.asarray([[0.1, 0.5, 0.49999997, 0.50000006, 0.9, 0.2],
                     [0.2, 0.3, 0.4, 0.49, 0.5, 0.5],
                     [0.7, 0.1, 0.1, 0.1, 0.1, 0.1],
                     [0.0, 0.0, 0.0, 0.0, 0.0, 1.0],
                     [0.5, 0.5, 0.5, 0.5, 0.5, 0.50001]], np.float32)
    out["stop_tokens"] = st
    out["output_lengths"] = np.asarray(RF._get_output_lengths(st), np.int64)
    rs = rng.random((7, 33)).astype(np.float32)
    out["stop_tokens_rand"] = rs
    out["output_lengths_rand"] = np.asarray(RF._get_output_lengths(rs), np.int64)

    # _prepare_batch (Tacotron_feeder.py:25-60): examples (ids, mel, stop, linear, n_frames), seeded shuffle, 1 and 2 devices
    for G in (1, 2):
        hparams.tacotron_num_gpus = G
        ex = [(seqs[i], mels[i], stops[i], mels[i][:, :7].copy(), len(mels[i])) for i in range(6)]
        np.random.seed(1324 + G)
        res = RF._prepare_batch(ex, 2)
        names = ("inputs", "input_lengths", "mel_targets", "token_targets", "linear_targets", "targets_lengths", "split_infos")
        for n, v in zip(names, res):
            out["batch_G%d_%s" % (G, n)] = np.asarray(v)
        parts = split_func(np.asarray(res[0]), np.asarray(res[6])[:, 0])
        for i, p in enumerate(parts):
            out["batch_G%d_split_inputs_%d" % (G, i)] = p
        parts = split_func(np.asarray(res[2]), np.asarray(res[6])[:, 1])
        for i, p in enumerate(parts):
            out["batch_G%d_split_mel_%d" % (G, i)] = p
    hparams.tacotron_num_gpus = 1
    np.savez_compressed(os.path.join(OUT, "ref_feeder.npz"), **out)
    return len(out)


def hparams_fixture():
    from Utils.Hyperparams import hparams
    vals = {k: _jsonable(v) for k, v in sorted(hparams.values().items())}
    with open(os.path.join(OUT, "ref_hparams.json"), "w") as f:
        json.dump(vals, f, indent=0, sort_keys=True, ensure_ascii=False)
    return len(vals)


def text_fixture():
    from Utils.TextProcessing.HangulUtils import hangul_symbol_1, hangul_to_sequence
    with open(os.path.join(REF, "sentences.txt"), encoding="utf-8") as f:
        lines = [ln.strip() for ln in f if ln.strip()]
    cwd = os.getcwd()
    os.chdir(REF)   # KoG2P opens the rulebook relative to `dir`
    try:
        ids = [[int(i) for i in hangul_to_sequence(dir=REF, hangul_text=ln, hangul_type=1)] for ln in lines]
    finally:
        os.chdir(cwd)
    doc = {"source": "sentences.txt (reference root), hangul_type=1", "n_symbols": len(hangul_symbol_1), "pad_id": 0, "eos_id": 1,
           "ids": ids, "lengths": [len(x) for x in ids]}
    with open(os.path.join(OUT, "ref_text.json"), "w", encoding="utf-8") as f:
        json.dump(doc, f, ensure_ascii=False, indent=0)
    return doc["lengths"]


def audio_fixture():
    from Utils.AudioProcessing import AudioPreprocess as RA
    from Utils.Hyperparams import hparams
    rng = np.random.default_rng(45454)
    out = {}
    wav = (rng.standard_normal(4001) * 0.3).astype(np.float64)
    out["wav"] = wav
    out["preemphasis_k097"] = RA.preemphasis(wav, 0.97, True)
    out["inv_preemphasis_k097"] = RA.inv_preemphasis(wav, 0.97, True)
    out["inv_preemphasis_off"] = RA.inv_preemphasis(wav, 0.97, False)
    mel = np.clip(rng.standard_normal((80, 37)) * 2.5, -5.0, 5.0).astype(np.float32)   # beyond +-max_abs_value: exercises the clip
    out["mel"] = mel
    cfgs = {"sym_clip": (True, True), "sym_noclip": (True, False), "asym_clip": (False, True), "asym_noclip": (False, False)}
    for name, (sym, clip) in cfgs.items():
        hparams.symmetric_mels, hparams.allow_clipping_in_normalization = sym, clip
        d = RA._denormalize(mel, hparams)
        out["denormalize_" + name] = d
        # without clipping the reference asserts its input range (AudioPreprocess.py:461): feed it in-range decibels there
        out["normalize_" + name] = RA._normalize(d if clip else np.clip(d, hparams.min_level_db, 0.0), hparams)
    hparams.symmetric_mels, hparams.allow_clipping_in_normalization = True, True
    db = np.linspace(-120.0, 30.0, 61)
    out["db"] = db
    out["db_to_amp"] = RA._db_to_amp(db)
    amp = np.geomspace(1e-7, 10.0, 50)
    out["amp"] = amp
    out["amp_to_db"] = RA._amp_to_db(amp, hparams)
    out["audio_hparams"] = np.asarray([hparams.max_abs_value, hparams.min_level_db, hparams.ref_level_db, hparams.power, hparams.preemphasis], np.float64)
    np.savez_compressed(os.path.join(OUT, "ref_audio.npz"), **out)
    return len(out)


def main():
    if not os.path.isdir(REF):
        raise SystemExit("reference tree %s not present: fixtures can only be regenerated in the build container" % REF)
    _install_stubs()
    os.makedirs(OUT, exist_ok=True)
    print("feeder arrays:", feeder_fixture())
    print("hparams:", hparams_fixture())
    print("token lengths of sentences.txt:", text_fixture())
    print("audio arrays:", audio_fixture())


if __name__ == "__main__":
    main()
