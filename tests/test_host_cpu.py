"""Host-side mirror of the reference interface: packing, helpers, argument errors, sharding plan (no GPU)."""
import numpy as np
import pytest

from tacotron_prosody_multi_gpu_b200 import runtime, sharding, synthetic
from tacotron_prosody_multi_gpu_b200.hparams import default_hparams
from tacotron_prosody_multi_gpu_b200.Utils import Helpers, Tacotron_feeder as F
from tacotron_prosody_multi_gpu_b200.Utils.Utils import split_func


def test_feeder_padding_and_round_up():
    ids, t = F._prepare_inputs([np.arange(1, 4), np.arange(1, 6)])
    assert t == 5 and ids.tolist() == [[1, 2, 3, 0, 0], [1, 2, 3, 4, 5]]
    mel, f = F._prepare_targets([np.ones((3, 80), np.float32), np.ones((5, 80), np.float32)], 2)
    assert f == 6 and mel.shape == (2, 6, 80)
    assert np.all(mel[0, 3:] == np.float32(-4.1)) and np.all(mel[1, 5:] == np.float32(-4.1))  # -(max_abs_value + .1)
    tok, ft = F._prepare_token_targets([np.zeros(3), np.zeros(6)], 2)
    assert ft == 8 and tok[0].tolist() == [0, 0, 0, 1, 1, 1, 1, 1] and tok[1, 6:].tolist() == [1, 1]
    assert [F._round_up(x, 2) for x in (4, 5)] == [4, 6]


def test_output_lengths_round_half_to_even():
    # np.round(0.5) == 0 (tf.round semantics): 0.5 does not stop, 0.5000001 does
    st = np.array([[0.1, 0.5, 0.50001, 0.9], [0.2, 0.3, 0.4, 0.49]], np.float32)
    assert F._get_output_lengths(st) == [2, 4]


def test_prepare_batch_packs_along_time_and_split_func_inverts():
    rng = np.random.default_rng(0)
    ex = []
    for n in (5, 9, 7, 4):
        ex.append((rng.integers(2, 60, n).astype(np.int32), rng.standard_normal((2 * n + 1, 80)).astype(np.float32), np.zeros(2 * n + 1, np.float32), None,
                   2 * n + 1))
    inputs, in_len, mel, tok, lin, tgt_len, si = F._prepare_batch(list(ex), 2, n_devices=2, shuffle=False)
    assert lin is None and si.shape == (2, 4) and si[:, 0].tolist() == [9, 7] and si[:, 1].tolist() == [20, 16]
    assert inputs.shape == (2, 16) and mel.shape == (2, 36, 80) and tok.shape == (2, 20 + 16)
    parts = split_func(inputs, si[:, 0])
    assert parts[0].shape == (2, 9) and parts[1].shape == (2, 7)
    assert parts[1][0, :7].tolist() == ex[2][0].tolist()
    assert in_len.tolist() == [5, 9, 7, 4] and tgt_len.tolist() == [11, 19, 15, 9]
    with pytest.raises(AssertionError):
        F._prepare_batch(list(ex[:3]), 2, n_devices=2)


def test_helpers_host_semantics():
    hp = default_hparams()
    h = Helpers.TacoTestHelper(3, hp)
    fin, go = h.initialize()
    assert not fin.any() and go.shape == (3, 80) and not go.any()
    out = np.arange(3 * 160, dtype=np.float32).reshape(3, 160)
    f, nxt, _ = h.next_inputs(0, out, None, None, np.array([[0.1, 0.2], [0.4, 0.5], [0.3, 0.2]]))
    assert not f and np.array_equal(nxt, out[:, -80:])
    f, _, _ = h.next_inputs(0, out, None, None, np.array([[0.1, 0.2], [0.4, 0.51], [0.3, 0.2]]))
    assert f  # batch-global: ANY rounded stop token of ANY row (Helpers.py:51-52)
    tg = np.random.default_rng(1).standard_normal((3, 8, 80)).astype(np.float32)
    th = Helpers.TacoTrainingHelper(3, tg, hp, True, False, None)
    th.initialize()
    assert th.ratio == 1.0
    f, nxt, _ = th.next_inputs(1, out, None, None, None)
    assert np.array_equal(nxt, tg[:, 3]) and not f.any()      # targets[:, r-1::r][:, 1] == targets[:, 3]
    assert th.next_inputs(3, out, None, None, None)[0].all()  # time + 1 >= S = 4


def test_tacotron_argument_errors_match_reference():
    from tacotron_prosody_multi_gpu_b200.TacotronModel.modules.Tacotron import Tacotron
    from tacotron_prosody_multi_gpu_b200.TacotronModel.modules.Encoder import FusedStageError
    hp = default_hparams()
    ids, lens = np.zeros((2, 3), np.int32), np.array([3, 3])
    m = Tacotron(hp)
    with pytest.raises(ValueError, match="no mel targets"):
        m.initialize(ids, lens, stop_token_targets=np.zeros((2, 4)))
    with pytest.raises(ValueError, match="without corresponding token_targets"):
        m.initialize(ids, lens, mel_targets=np.zeros((2, 4, 80)))
    with pytest.raises(ValueError, match="GTA mode"):
        m.initialize(ids, lens, mel_targets=np.zeros((2, 4, 80)), GTA=True, linear_targets=np.zeros((2, 4, 1025)))
    with pytest.raises(RuntimeError, match="training and evaluation"):
        m.initialize(ids, lens, mel_targets=np.zeros((2, 4, 80)), stop_token_targets=np.zeros((2, 4)), is_training=True, is_evaluating=True)
    hp2 = default_hparams()
    hp2.mask_decoder = True
    with pytest.raises(RuntimeError, match="targets lengths"):
        Tacotron(hp2).initialize(ids, lens, mel_targets=np.zeros((2, 4, 80)), stop_token_targets=np.zeros((2, 4)), is_training=True)
    import torch
    if not torch.cuda.is_available():
        from tacotron_prosody_multi_gpu_b200._lib import TacoError
        with pytest.raises(TacoError):     # training runs on the device primitives only: no CPU fallback
            m.initialize(ids, lens, mel_targets=np.zeros((2, 4, 80)), stop_token_targets=np.zeros((2, 4)), is_training=True)
    with pytest.raises(RuntimeError, match="initialize"):
        Tacotron(hp).add_loss()


def test_install_aliases_exposes_reference_import_paths():
    runtime.install_aliases()
    from TacotronModel.modules.Tacotron import Tacotron            # noqa: F401
    from TacotronModel.modules import Attention, Decoder, Encoder, Postnet  # noqa: F401
    from Utils.Tacotron_synthesizer import Synthesizer              # noqa: F401
    from Utils.Helpers import TacoTestHelper, TacoTrainingHelper    # noqa: F401
    from Utils.Hyperparams import hparams
    assert hparams.outputs_per_step == 2 and hparams.tacotron_random_seed == 45454
    assert Decoder.Prenet(False, [256, 256]).scope == 'prenet'
    with pytest.raises(ValueError):
        Encoder.ZoneoutLSTMCell(256, False, 1.5, 0.1)


def test_no_cpu_fallback():
    import torch
    from tacotron_prosody_multi_gpu_b200._lib import TacoError
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(TacoError):
        runtime.engine(default_hparams())


def test_shard_plan_is_rank_invariant_and_balanced():
    hp = default_hparams()
    lengths = synthetic.utterance_set(hp, 4096).tolist()
    ref_batches = None
    for world in (1, 2, 4, 8):
        batches, owner = sharding.shard(lengths, world)
        assert len(batches) == 64 and all(len(b) == 64 for b in batches)
        if ref_batches is None:
            ref_batches = batches
        assert all(np.array_equal(a, b) for a, b in zip(batches, ref_batches))   # composition independent of G
        assert sorted(np.concatenate(batches).tolist()) == list(range(4096))
        costs = np.array([sharding.batch_cost(lengths, b) for b in batches])
        load = np.array([costs[owner == r].sum() for r in range(world)])
        assert load.max() <= load.mean() + costs.max()                            # LPT bound
    # within a batch the lengths are contiguous in sorted order
    mx = [max(lengths[i] for i in b) for b in ref_batches]
    assert mx == sorted(mx)


def test_shard_empty_and_ragged():
    batches, owner = sharding.shard([], 4)
    assert batches == [] and len(owner) == 0
    batches, owner = sharding.shard([5, 7, 3], 8, batch_size=2)
    assert [len(b) for b in batches] == [2, 1] and set(owner.tolist()) <= set(range(8))
