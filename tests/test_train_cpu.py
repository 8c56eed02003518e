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
