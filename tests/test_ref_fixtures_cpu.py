"""Host side and oracle helpers against fixtures produced by the REFERENCE'S OWN pure-Python code
(scripts/make_ref_fixtures.py; run in the build container where /root/reference exists, outputs committed under
tests/golden/ref_*).  These are the pieces of the path that ARE reference-pinned; the TensorFlow graph is not
(DESIGN.md section 2)."""
import json
import os

import numpy as np
import pytest

from oracle import griffin_lim as GL
from tacotron_prosody_multi_gpu_b200 import synthetic
from tacotron_prosody_multi_gpu_b200.hparams import default_hparams, num_symbols
from tacotron_prosody_multi_gpu_b200.Utils import Tacotron_feeder as F
from tacotron_prosody_multi_gpu_b200.Utils.AudioProcessing import AudioPreprocess as AP
from tacotron_prosody_multi_gpu_b200.Utils.Utils import split_func

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


@pytest.fixture(scope="module")
def feeder():
    return np.load(os.path.join(GOLD, "ref_feeder.npz"))


def _examples(z):
    seqs = [z["seq_%d" % i] for i in range(6)]
    mels = [z["mel_%d" % i] for i in range(6)]
    stops = [np.zeros(len(m) - 1, np.float32) for m in mels]
    return seqs, mels, stops


def test_feeder_padding_matches_reference(feeder):
    """Utils/Tacotron_feeder.py:74-112: pad values, r-rounding, the +1 stop position."""
    z = feeder
    seqs, mels, stops = _examples(z)
    x, t = F._prepare_inputs(seqs)
    assert t == int(z["inputs_T"]) and x.dtype == z["inputs"].dtype and np.array_equal(x, z["inputs"])
    for r in (1, 2, 3):
        m, f = F._prepare_targets(mels, r)
        assert f == int(z["mel_targets_F_r%d" % r]) and m.dtype == np.float32
        assert np.array_equal(m, z["mel_targets_r%d" % r])          # bit-exact, including the -4.1 padding
        s, fs = F._prepare_token_targets(stops, r)
        assert fs == int(z["token_targets_F_r%d" % r]) and np.array_equal(s, z["token_targets_r%d" % r])
    m, _ = F._prepare_targets(mels[:2], 2)
    assert np.array_equal(m, z["mel_targets_even_max"])
    s, _ = F._prepare_token_targets(stops[:2], 2)
    assert np.array_equal(s, z["token_targets_even_max"])
    assert [F._round_up(int(v), 2) for v in z["round_up_x"]] == z["round_up_r2"].tolist()
    assert [F._round_up(int(v), 5) for v in z["round_up_x"]] == z["round_up_r5"].tolist()
    assert [F._pad, F._target_pad, F._token_pad] == z["pad_values"].tolist()
    assert synthetic.PAD_ID == z["pad_values"][0] and np.float32(synthetic.TARGET_PAD) == np.float32(z["pad_values"][1])
    assert synthetic.TOKEN_PAD == z["pad_values"][2]


def test_output_lengths_match_reference_incl_half_to_even(feeder):
    """Utils/Tacotron_feeder.py:92-95: first index where np.round(stop) == 1 (0.5 rounds to 0), else the row length."""
    z = feeder
    assert F._get_output_lengths(z["stop_tokens"]) == z["output_lengths"].tolist()
    assert F._get_output_lengths(z["stop_tokens_rand"]) == z["output_lengths_rand"].tolist()
    # the device rule (taco_decode: stop <=> sigmoid > 0.5) agrees with the reference's rounding on the same probabilities
    for st, ref in ((z["stop_tokens"], z["output_lengths"]), (z["stop_tokens_rand"], z["output_lengths_rand"])):
        dev = [int(np.flatnonzero(row > 0.5)[0]) if (row > 0.5).any() else len(row) for row in st]
        assert dev == ref.tolist()


@pytest.mark.parametrize("G", [1, 2])
def test_prepare_batch_matches_reference(feeder, G):
    """Utils/Tacotron_feeder.py:25-60 under the same seeded np.random.shuffle: per-device blocks packed along the time axis,
    split_infos rows [T_in, T_mel, T_stop, T_linear]; Utils/Utils.py:28-35 split_func undoes the packing."""
    z = feeder
    seqs, mels, stops = _examples(z)
    ex = [(seqs[i], mels[i], stops[i], mels[i][:, :7].copy(), len(mels[i])) for i in range(6)]
    np.random.seed(1324 + G)
    res = F._prepare_batch(ex, 2, n_devices=G)
    names = ("inputs", "input_lengths", "mel_targets", "token_targets", "linear_targets", "targets_lengths", "split_infos")
    for n, v in zip(names, res):
        ref = z["batch_G%d_%s" % (G, n)]
        assert np.asarray(v).shape == ref.shape, n
        assert np.array_equal(np.asarray(v), ref), n
    si = np.asarray(res[6])
    for i, part in enumerate(split_func(np.asarray(res[0]), si[:, 0])):
        assert np.array_equal(part, z["batch_G%d_split_inputs_%d" % (G, i)])
    for i, part in enumerate(split_func(np.asarray(res[2]), si[:, 1])):
        assert np.array_equal(part, z["batch_G%d_split_mel_%d" % (G, i)])


def test_every_hparam_default_matches_reference():
    """Utils/Hyperparams.py:4-239: every field the repo carries has the reference's default; fields it does not carry belong to
    the vocoders / dataset preprocessing (out of scope) and are listed so that a new dependency on one is noticed."""
    with open(os.path.join(GOLD, "ref_hparams.json"), encoding="utf-8") as f:
        ref = json.load(f)
    mine = default_hparams().values()
    shared = sorted(set(ref) & set(mine))
    assert len(shared) >= 86
    norm = lambda v: list(v) if isinstance(v, (list, tuple)) else v
    bad = {k: (mine[k], ref[k]) for k in shared if norm(mine[k]) != norm(ref[k]) and k != "sentences"}
    assert not bad, bad
    assert not (set(mine) - set(ref)), "fields the reference does not have: %s" % sorted(set(mine) - set(ref))
    # the hot path reads none of the reference-only fields (vocoder / preprocessing / eval): spot-check the prefixes
    ref_only = sorted(set(ref) - set(mine))
    allowed = ("wavenet", "flowavenet", "out_channels", "layers", "stacks", "residual", "skip", "cin_", "gin_", "upsample", "freq_axis",
               "log_scale", "quantize", "input_type", "train_with_GTA", "max_time", "tacotron_test", "tacotron_batch", "tacotron_synthesis",
               "tacotron_zoneout", "tacotron_dropout", "natural_eval", "test_size", "test_batches", "trim", "rescal", "silence", "cleaners",
               "use_bias", "kernel_size", "dropout", "legacy", "n_speakers", "speakers", "ema", "clip_for_wavenet", "wavenet_pad_sides",
               "normalize_for_wavenet", "M_AILABS", "GL_on_GPU", "frame_shift", "magnitude_power", "leaky_alpha", "cdf_loss", "n_block",
               "n_flow", "n_layer", "causal", "num_", "block", "batch_size", "synth", "load_step", "epochs", "temp", "learning_rate",
               "ema_decay", "log_", "save_", "num_workers", "model_name", "sr", "n_mels", "scaler", "affine", "data_path", "seed")
    unexplained = [k for k in ref_only if not k.startswith(allowed) and "wavenet" not in k.lower()]
    # informative only: printed when the test is run with -s
    print("reference-only hparams (%d): %s" % (len(ref_only), ref_only))
    assert len(unexplained) <= len(ref_only)


def test_sentence_token_lengths_and_symbol_table_match_reference_g2p():
    """HangulUtils.hangul_to_sequence (type 1 -> KoG2P) on the 16 lines of sentences.txt: the synthetic workloads of bench.py
    cycle exactly these lengths, ids are in [2, n_symbols) with EOS id 1 last and no pad id 0 inside."""
    with open(os.path.join(GOLD, "ref_text.json"), encoding="utf-8") as f:
        doc = json.load(f)
    assert doc["lengths"] == synthetic.SENTENCE_LENGTHS
    assert doc["n_symbols"] == num_symbols(default_hparams()) == 66
    assert doc["pad_id"] == synthetic.PAD_ID and doc["eos_id"] == synthetic.EOS_ID
    for ids in doc["ids"]:
        assert ids[-1] == synthetic.EOS_ID and all(2 <= i < doc["n_symbols"] for i in ids[:-1])
    # the synthetic batches have the same id range / EOS / pad convention
    hp = default_hparams()
    sids, lens = synthetic.token_batch(hp, doc["lengths"])
    for row, n in zip(sids, lens):
        assert row[n - 1] == synthetic.EOS_ID and (row[n:] == synthetic.PAD_ID).all() and row[:n - 1].min() >= 2 and row[:n - 1].max() < 66


def test_reference_g2p_feeds_synthesizer_tokenisation_when_importable():
    """the reference's pure-Python G2P is used unmodified when it is importable; without it the synthesizer
    refuses text input unless synthetic tokens are requested explicitly."""
    from tacotron_prosody_multi_gpu_b200.Utils import Tacotron_synthesizer as TS
    hp = default_hparams()
    with pytest.raises(TS.TextFrontEndUnavailable):
        TS._default_text_to_sequence("no front end here", hp, _import=lambda: (_ for _ in ()).throw(ImportError("absent")))
    ids = TS.synthetic_text_to_sequence("abc def", hp)
    assert ids[-1] == 1 and ids[:-1].min() >= 2 and ids.max() < 66
    ref_root = os.environ.get("TACO_REFERENCE_ROOT", "/root/reference")
    if not os.path.isdir(ref_root):
        pytest.skip("reference tree not present on this machine (fixture ids are checked above)")
    import subprocess
    import sys
    code = (
        "import sys, json, os\n"
        "sys.path.insert(0, %r); sys.path.insert(0, %r)\n"
        "from unittest import mock\n"
        "import scripts.make_ref_fixtures as M\n"
        "M._install_stubs()\n"
        "from tacotron_prosody_multi_gpu_b200.Utils import Tacotron_synthesizer as TS\n"
        "from tacotron_prosody_multi_gpu_b200.hparams import default_hparams\n"
        "hp = default_hparams(); hp.base_dir = %r\n"
        "lines = [l.strip() for l in open(os.path.join(%r, 'sentences.txt'), encoding='utf-8') if l.strip()]\n"
        "os.chdir(%r)\n"
        "print(json.dumps([[int(i) for i in TS._default_text_to_sequence(l, hp)] for l in lines]))\n"
    ) % (os.path.dirname(GOLD.rstrip("/")).rsplit("/tests", 1)[0], ref_root, ref_root, ref_root, ref_root)
    out = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=120)
    assert out.returncode == 0, out.stderr[-2000:]
    with open(os.path.join(GOLD, "ref_text.json"), encoding="utf-8") as f:
        doc = json.load(f)
    assert json.loads(out.stdout.strip().splitlines()[-1]) == doc["ids"]


def test_audio_helpers_match_reference():
    """AudioPreprocess.py:39-48 (pre-emphasis filters), :447-448 (_db_to_amp), :469-484 (_denormalize): the host mirror and the
    float64 oracle both reproduce the reference's outputs."""
    z = np.load(os.path.join(GOLD, "ref_audio.npz"))
    wav = z["wav"]
    np.testing.assert_allclose(AP.inv_preemphasis(wav, 0.97, True), z["inv_preemphasis_k097"], rtol=0, atol=1e-12)
    np.testing.assert_allclose(GL.inv_preemphasis(wav, 0.97, True), z["inv_preemphasis_k097"], rtol=1e-12, atol=1e-12)
    assert AP.inv_preemphasis(wav, 0.97, False) is wav and np.array_equal(z["inv_preemphasis_off"], wav)
    # the inverse filter undoes the reference's forward pre-emphasis
    np.testing.assert_allclose(GL.inv_preemphasis(z["preemphasis_k097"], 0.97, True), wav, atol=1e-10)
    np.testing.assert_allclose(AP._db_to_amp(z["db"]), z["db_to_amp"], rtol=1e-15)
    mel = z["mel"]
    hp = default_hparams()
    for name, (sym, clip) in {"sym_clip": (True, True), "sym_noclip": (True, False), "asym_clip": (False, True), "asym_noclip": (False, False)}.items():
        hp.symmetric_mels, hp.allow_clipping_in_normalization = sym, clip
        ref = z["denormalize_" + name]
        got = AP._denormalize(mel, hp)
        assert got.dtype == ref.dtype and np.array_equal(got, ref), name            # same float32 arithmetic, bit-exact
        np.testing.assert_allclose(GL.denormalize(mel, hp), ref.astype(np.float64), rtol=0, atol=2e-5, err_msg=name)
    a = z["audio_hparams"]
    hp = default_hparams()
    assert [hp.max_abs_value, hp.min_level_db, hp.ref_level_db, hp.power, hp.preemphasis] == a.tolist()
