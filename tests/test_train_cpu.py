"""Host logic of the training step on CPU: the hand-derived backward pass of train/step.py, composed from
the torch-CPU twin of the device primitives (tests/cpu_train_ops.py), against autograd through the oracle."""
import numpy as np
import pytest
import torch

from cpu_train_ops import TorchOps
from oracle import train_oracle
from tacotron_prosody_multi_gpu_b200 import synthetic, weights
from tacotron_prosody_multi_gpu_b200.hparams import default_hparams
from tacotron_prosody_multi_gpu_b200.train.params import is_regularized
from tacotron_prosody_multi_gpu_b200.train.step import TrainStep


def small_problem(hp, lengths=(7, 5, 6), frames_per_token=2, seed=3):
    ids, lens = synthetic.token_batch(hp, list(lengths))
    mel, stop, _ = synthetic.mel_targets(hp, lens, frames_per_token=frames_per_token)
    return ids, lens, mel, stop, seed


@pytest.fixture(scope="module")
def problem():
    hp = default_hparams()
    w = weights.init_weights(hp, perturb=0.05)
    ids, lens, mel, stop, seed = small_problem(hp)
    out, orc = train_oracle.train_forward(hp, w, ids, lens, mel, stop, seed)
    out["loss"].backward()
    return hp, w, (ids, lens, mel, stop, seed), out, train_oracle.gradients(orc)


def _run(hp, w, data):
    ids, lens, mel, stop, seed = data
    ops = TorchOps()
    step = TrainStep(hp, ops, w)
    out = step.forward(torch.from_numpy(ids), torch.from_numpy(lens), torch.from_numpy(mel), torch.from_numpy(stop), seed)
    step.backward()
    return step, out


def test_regularization_filter_matches_reference_name_rule():
    assert is_regularized("decoder/Location_Sensitive_Attention/attention_variable")       # included despite the comment
    assert is_regularized("encoder_convolutions/conv_layer_1_encoder_convolutions/batch_normalization/beta")
    assert is_regularized("ref_encoder/rnn/gru_cell/gates/kernel") and is_regularized("Multihead-attention/attention_b")
    assert not is_regularized("decoder/decoder_LSTM/multi_rnn_cell/cell_0/decoder_LSTM_1/kernel")
    assert not is_regularized("postnet_projection/projection_postnet_projection/kernel")
    assert not is_regularized("decoder/Location_Sensitive_Attention/attention_bias")
    assert not is_regularized("inputs_embedding") and not is_regularized("ref_encoder/rnn/gru_cell/gates/bias")


def test_training_forward_and_losses_match_oracle(problem):
    hp, w, data, ref, _ = problem
    step, out = _run(hp, w, data)
    np.testing.assert_allclose(out["encoder_outputs"].numpy(), ref["encoder_outputs"].detach().numpy(), atol=2e-5)
    np.testing.assert_allclose(out["decoder_output"].numpy(), ref["decoder_output"].detach().numpy(), atol=5e-5)
    np.testing.assert_allclose(out["mel_outputs"].numpy(), ref["mel_outputs"].detach().numpy(), atol=1e-4)
    np.testing.assert_allclose(out["stop_token_prediction"].numpy(), ref["stop_token_prediction"].detach().numpy(), atol=5e-5)
    np.testing.assert_allclose(out["alignments"].permute(1, 2, 0).numpy(), ref["alignments"].detach().numpy(), atol=1e-5)
    ls = out["losses"].numpy()
    for i, k in enumerate(("before_loss", "after_loss", "stop_token_loss", "regularization_loss")):
        np.testing.assert_allclose(ls[i], float(ref[k].detach()), rtol=2e-5)


def test_hand_derived_backward_matches_autograd(problem):
    hp, w, data, ref, ref_grads = problem
    step, out = _run(hp, w, data)
    st = step.store
    # the L2 term enters the product path in apply(); add it here to compare with d(loss)/d(var)
    worst = ("", 0.0)
    for name, g_ref in ref_grads.items():
        g = st.G[name].numpy().reshape(g_ref.shape).astype(np.float64)
        if is_regularized(name):
            g = g + train_oracle.reg_weight(hp) * np.asarray(w[name], np.float64)
        scale = max(np.abs(g_ref).max(), 1e-3)   # conv biases in front of BN have exactly-zero gradients (rounding noise)
        err = np.abs(g - g_ref).max() / scale
        if err > worst[1]:
            worst = (name, err)
        assert err < 2e-3, "%s: rel err %.3e (|g|max %.3e)" % (name, err, scale)
    print("worst gradient:", worst)


def test_apply_matches_reference_optimizer(problem):
    hp, w, data, ref, ref_grads = problem
    step, out = _run(hp, w, data)
    step.apply(global_step=0)
    wn = {n: np.asarray(v, np.float64).copy() for n, v in w.items() if n in ref_grads}
    m = {n: np.zeros_like(v) for n, v in wn.items()}
    v = {n: np.zeros_like(x) for n, x in wn.items()}
    norm = train_oracle.adam_update(hp, wn, [ref_grads], m, v, 0)
    assert norm > 0
    new = step.store.export()
    for n in wn:
        np.testing.assert_allclose(new[n], wn[n], atol=2e-5, err_msg=n)
    # moving statistics moved towards the batch statistics (UPDATE_OPS, Tacotron.py:513)
    k = "postnet_convolutions/conv_layer_1_postnet_convolutions/batch_normalization/moving_mean"
    assert np.abs(new[k] - np.asarray(w[k])).max() > 0


def test_unsupported_training_hparams_raise():
    """Settings that change the reference's training graph and are not built must fail loudly (Helpers.py:100-124)."""
    w = weights.init_weights(default_hparams(), perturb=0.05)
    for field, value in (("tacotron_teacher_forcing_mode", "scheduled"), ("tacotron_teacher_forcing_ratio", 0.5), ("predict_linear", True)):
        hp = default_hparams()
        setattr(hp, field, value)
        with pytest.raises(NotImplementedError):
            TrainStep(hp, TorchOps(), w)


def test_forward_rejects_mismatched_stop_targets(problem):
    hp, w, data, _, _ = problem
    ids, lens, mel, stop, seed = data
    step = TrainStep(hp, TorchOps(), w)
    wide = np.concatenate([stop, np.ones((stop.shape[0], 2), np.float32)], axis=1)      # F + 2 columns
    with pytest.raises(ValueError):
        step.forward(torch.from_numpy(ids), torch.from_numpy(lens), torch.from_numpy(mel), torch.from_numpy(wide), seed)


def test_fine_tuning_freezes_embedding_and_encoder(problem):
    """tacotron_fine_tuning (Tacotron.py:478-480): variables named *inputs_embedding* / *encoder_* keep their values."""
    hp0, w, data, _, _ = problem
    hp = default_hparams()
    hp.tacotron_fine_tuning = True
    step, _ = _run(hp, w, data)
    step.apply(global_step=0)
    new = step.store.export()
    frozen = [n for n in step.store.layout if "inputs_embedding" in n or "encoder_" in n]
    assert "inputs_embedding" in frozen and any(n.startswith("encoder_LSTM/") for n in frozen) and len(frozen) == 17
    assert not any(n.startswith("ref_encoder/") for n in frozen)     # 'ref_encoder/...' does not contain 'encoder_'
    for n in frozen:
        assert np.array_equal(new[n], np.asarray(w[n], np.float32)), n
    moved = "decoder/decoder_prenet/dense_2/kernel"
    assert np.abs(new[moved] - np.asarray(w[moved])).max() > 0


def test_training_state_round_trip_resumes_adam_and_schedule(problem, tmp_path):
    """Checkpoint = variables + BN statistics + Adam moments + global_step (tf.train.Saver, train.py:60): two updates in one
    run equal one update, save, load into a fresh store, second update."""
    hp, w, data, _, _ = problem
    ids, lens, mel, stop, seed = data
    args = tuple(torch.from_numpy(a) for a in (ids, lens, mel, stop))

    def update(step, k):
        step.forward(*args, seed + k)
        step.backward()
        step.apply(global_step=k)

    a = TrainStep(hp, TorchOps(), w)
    update(a, 0)
    path = str(tmp_path / "state.npz")
    a.store.save_state(path, 1)
    update(a, 1)
    b = TrainStep(hp, TorchOps(), weights.init_weights(hp, perturb=0.0))       # different initial values: everything must come from the file
    assert b.store.load_state(path) == 1
    update(b, 1)
    ea, eb = a.store.export(), b.store.export()
    for n in ea:
        assert np.array_equal(ea[n], eb[n]), n
    assert torch.equal(a.store.m, b.store.m) and torch.equal(a.store.v, b.store.v)
    # a plain weight container (no Adam slots) loads with zero moments and step 0
    plain = str(tmp_path / "weights.npz")
    weights.save(plain, ea)
    c = TrainStep(hp, TorchOps(), w)
    assert c.store.load_state(plain) == 0 and float(c.store.m.abs().max()) == 0.0
