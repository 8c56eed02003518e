"""Pins for the CPU oracle: NumPy re-derivation agreement, TF-semantics
known-answer cases, structural invariants.  The reference has no tests of its own."""
import numpy as np
import pytest
import torch

from oracle import numpy_step, philox
from oracle.tacotron_oracle import TacotronOracle, output_lengths
from tacotron_prosody_multi_gpu_b200 import synthetic, weights
from tacotron_prosody_multi_gpu_b200.hparams import default_hparams


@pytest.fixture(scope="module")
def setup():
    hp = default_hparams()
    w = weights.init_weights(hp, perturb=0.05)
    return hp, w


def test_parameter_count_matches_survey(setup):
    hp, w = setup
    assert weights.count_trainable(w) == 20425491
    dec = sum(w[k].size for k in w if k.startswith("decoder/") and "memory_layer" not in k)
    assert dec == 2127778                                   # per-step resident set


def test_philox_known_answers():
    # Random123 kat_vectors, philox4x32 10 rounds
    assert [int(x) for x in philox.philox4x32_10(0, 0, 0, 0, 0, 0)] == [0x6627e8d5, 0xe169c58d, 0xbc57ac4c, 0x9b00dbd8]
    assert [int(x) for x in philox.philox4x32_10(0xffffffff, 0xffffffff, 0xffffffff, 0xffffffff, 0xffffffff, 0xffffffff)] == \
        [0x408f276d, 0x41c83b0e, 0xa20bc7c6, 0x6d5451fd]
    assert [int(x) for x in philox.philox4x32_10(0x243f6a88, 0x85a308d3, 0x13198a2e, 0x03707344, 0xa4093822, 0x299f31d0)] == \
        [0xd16cfe09, 0x94fdcceb, 0x5001e420, 0x24126ea1]
    m = philox.prenet_masks(45454, 50, 8)
    assert m.shape == (50, 8, 2, 256) and abs(m.mean() - 0.5) < 0.01


def test_decoder_step_matches_numpy_rederivation(setup):
    hp, w = setup
    o = TacotronOracle(hp, w, dtype=torch.float64)
    rng = np.random.default_rng(0)
    B, T = 3, 23
    lengths = np.array([23, 9, 17])
    memory = torch.from_numpy(rng.standard_normal((B, T, 768)))
    values, keys, mask = o.attention_init(memory, lengths)
    st = o.zero_state(B, T)
    stn = dict(c1=np.zeros((B, 256)), h1=np.zeros((B, 256)), c2=np.zeros((B, 256)), h2=np.zeros((B, 256)),
               attention=np.zeros((B, 768)), alignments=np.zeros((B, T)))
    x = torch.from_numpy(rng.standard_normal((B, 80)))
    for step in range(3):
        km = philox.prenet_masks(7, 1, B, step0=step)[0]
        f, s, a, st = o.decoder_step(x, st, values, keys, mask, torch.from_numpy(km))
        fn, sn, an, stn = numpy_step.decoder_step_np(w, hp, x.numpy(), stn, values.numpy(), keys.numpy(), lengths, km)
        np.testing.assert_allclose(f.numpy(), fn, atol=1e-9)
        np.testing.assert_allclose(s.numpy(), sn, atol=1e-9)
        np.testing.assert_allclose(a.numpy(), an, atol=1e-12)
        np.testing.assert_allclose(st.alignments.numpy(), stn["alignments"], atol=1e-12)
        x = f[:, -80:]
    # fp32 oracle vs fp64: bounds the fp32 error used for the kernel tolerances
    o32 = TacotronOracle(hp, w)
    st32 = o32.zero_state(B, T)
    v32, k32, m32 = o32.attention_init(memory.float(), lengths)
    f32, _, a32, _ = o32.decoder_step(torch.zeros(B, 80), st32, v32, k32, m32, torch.from_numpy(philox.prenet_masks(7, 1, B)[0]))
    f64, _, a64, _ = o.decoder_step(torch.zeros(B, 80, dtype=torch.float64), o.zero_state(B, T), values, keys, mask,
                                    torch.from_numpy(philox.prenet_masks(7, 1, B)[0]))
    assert (f32.double() - f64).abs().max() < 1e-5 and (a32.double() - a64).abs().max() < 1e-6


def test_same_padding_tables():
    # Appendix A.5: freq axis 80->40->20->10->5->3->2; even sizes pad (0,1), odd (1,1)
    sizes = [80]
    for _ in range(6):
        out, pb, pa = numpy_step.same_pad_stride2(sizes[-1])
        assert (pb, pa) == ((0, 1) if sizes[-1] % 2 == 0 else (1, 1))
        sizes.append(out)
    assert sizes == [80, 40, 20, 10, 5, 3, 2]


def test_conv2d_same_stride2_matches_loops(setup):
    hp, w = setup
    o = TacotronOracle(hp, w, dtype=torch.float64)
    rng = np.random.default_rng(1)
    for H, Wd in [(7, 5), (8, 6), (5, 3), (3, 2)]:
        x = rng.standard_normal((2, H, Wd, 3))
        k = rng.standard_normal((3, 3, 3, 4))
        b = rng.standard_normal(4)
        y = o._conv2d_same_s2(torch.from_numpy(x), torch.from_numpy(k), torch.from_numpy(b)).numpy()
        np.testing.assert_allclose(y, numpy_step.conv2d_same_s2_np(x, k, b), atol=1e-10)


def test_conv1d_same_k5_k31(setup):
    hp, w = setup
    o = TacotronOracle(hp, w, dtype=torch.float64)
    rng = np.random.default_rng(2)
    cum = rng.random((2, 40))
    k = rng.standard_normal((31, 1, 32))
    b = rng.standard_normal(32)
    y = o._conv1d_same(torch.from_numpy(cum).unsqueeze(2), torch.from_numpy(k), torch.from_numpy(b)).numpy()
    np.testing.assert_allclose(y, numpy_step.location_conv_np(cum, k, b), atol=1e-10)
    # k=5: pad 2/2, tap 2 is the centre
    x = torch.zeros(1, 9, 1, dtype=torch.float64); x[0, 4, 0] = 1.0
    k5 = torch.arange(5, dtype=torch.float64).reshape(5, 1, 1)
    y5 = o._conv1d_same(x, k5, None)[0, :, 0].tolist()
    assert y5 == [0, 0, 4, 3, 2, 1, 0, 0, 0]


def test_round_half_even_and_stop_lengths():
    assert float(torch.round(torch.tensor(0.5))) == 0.0          # p == 0.5 does NOT stop (App. A.12)
    assert float(torch.round(torch.tensor(0.5000001))) == 1.0
    assert output_lengths(np.array([[0.1, 0.5, 0.51, 0.9], [0.2, 0.3, 0.4, 0.5]])) == [2, 4]


def test_zoneout_and_raw_output(setup):
    hp, w = setup
    o = TacotronOracle(hp, w, dtype=torch.float64)
    rng = np.random.default_rng(3)
    x = torch.from_numpy(rng.standard_normal((2, 8)))
    c = torch.from_numpy(rng.standard_normal((2, 4)))
    h = torch.from_numpy(rng.standard_normal((2, 4)))
    k = torch.from_numpy(rng.standard_normal((12, 16)))
    b = torch.from_numpy(rng.standard_normal(16))
    raw, zc, zh = o.zoneout_lstm(x, c, h, k, b)
    nc, nh = numpy_step.lstm_cell_np(x.numpy(), c.numpy(), h.numpy(), k.numpy(), b.numpy())
    np.testing.assert_allclose(raw.numpy(), nh, atol=1e-12)                       # output is RAW h
    np.testing.assert_allclose(zc.numpy(), 0.9 * nc + 0.1 * c.numpy(), atol=1e-12)
    np.testing.assert_allclose(zh.numpy(), 0.9 * nh + 0.1 * h.numpy(), atol=1e-12)


def test_bilstm_masking(setup):
    hp, w = setup
    o = TacotronOracle(hp, w)
    rng = np.random.default_rng(4)
    x = torch.from_numpy(rng.standard_normal((2, 7, 512)).astype(np.float32))
    out = o.encoder_lstm(x, [7, 4])
    assert out.shape == (2, 7, 512)
    assert float(out[1, 4:].abs().max()) == 0.0                                    # zero beyond length
    # row 1 must equal running it alone at its own length (padding-independence of the LSTM)
    alone = o.encoder_lstm(x[1:2, :4], [4])
    np.testing.assert_allclose(out[1, :4].numpy(), alone[0].numpy(), atol=1e-6)


def test_forward_invariants(setup):
    hp, w = setup
    w2 = weights.set_stop_bias(w, -10.0)
    o = TacotronOracle(hp, w2)
    ids, lens = synthetic.token_batch(hp, [20, 11, 16])
    d = o.forward(ids, lens, gst_weights=synthetic.gst_random_weights(hp), max_iters=6, philox_seed=5)
    S = d["steps"]
    assert S == 6
    assert d["alignments"].shape == (3, 20, S) and d["stop"].shape == (3, 2 * S)
    assert d["mel_outputs"].shape == (3, 2 * S, 80) and d["encoder_outputs"].shape == (3, 20, 768)
    a = d["alignments"].numpy()
    np.testing.assert_allclose(a.sum(1), 1.0, atol=1e-5)
    assert a[1, 11:].max() == 0.0 and a[2, 16:].max() == 0.0                       # exact zeros beyond length
    assert float(d["mel_outputs"].max()) <= 4.0 and float(d["mel_outputs"].min()) >= -4.1
    # batch-global stop: S = first any-stop step + 1
    w3 = weights.scale_stop_projection(w, 0.0, -10.0)
    o3 = TacotronOracle(hp, w3)
    d3 = o3.forward(ids, lens, gst_weights=synthetic.gst_random_weights(hp), max_iters=4, philox_seed=5)
    assert d3["steps"] == 4
    w4 = weights.scale_stop_projection(w, 0.0, 10.0)
    d4 = TacotronOracle(hp, w4).forward(ids, lens, gst_weights=synthetic.gst_random_weights(hp), max_iters=4, philox_seed=5)
    assert d4["steps"] == 1


def test_teacher_forced_steps(setup):
    hp, w = setup
    o = TacotronOracle(hp, w)
    ids, lens = synthetic.token_batch(hp, [6, 4])
    mel, stop, flen = synthetic.mel_targets(hp, lens, frames_per_token=3)
    d = o.forward(ids, lens, mel_targets=mel, gst_weights=synthetic.gst_random_weights(hp), philox_seed=1)
    assert d["steps"] == mel.shape[1] // hp.outputs_per_step
    # teacher forcing: step t>0 input is target frame r*t-1, so outputs differ from free-running
    d2 = o.forward(ids, lens, gst_weights=synthetic.gst_random_weights(hp), philox_seed=1, max_iters=d["steps"])
    assert not np.allclose(d["frames_raw"][:, 2:4].numpy(), d2["frames_raw"][:, 2:4].numpy())
    np.testing.assert_allclose(d["frames_raw"][:, :2].numpy(), d2["frames_raw"][:, :2].numpy(), atol=1e-6)


def test_reference_encoder_and_gst_shapes(setup):
    hp, w = setup
    o = TacotronOracle(hp, w)
    ref = synthetic.reference_mel(hp, 2, 130)
    r = o.reference_encoder(ref)
    assert r.shape == (2, 128)
    s = o.gst_attention(r)
    assert s.shape == (2, 1, 256)
    # each head's context is a convex combination of tanh(tokens)
    tok = torch.tanh(o.w["style_tokens"])
    assert float(s.max()) <= float(tok.max()) + 1e-6 and float(s.min()) >= float(tok.min()) - 1e-6
    rw = synthetic.gst_random_weights(hp)
    np.testing.assert_allclose(rw.sum(1), 1.0, atol=1e-6)
    assert o.style_random(rw).shape == (1, 1, 256)
