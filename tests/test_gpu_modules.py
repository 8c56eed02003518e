"""Standalone calls of the reference's sub-modules (SURVEY 8b "Modules" row) against the matching oracle method, same random-init
weights and seeded inputs.  fp32 device primitives vs the fp32 oracle: absolute tolerances stated per check."""
import numpy as np
import pytest
import torch

from oracle import philox
from oracle.tacotron_oracle import TacotronOracle
from tacotron_prosody_multi_gpu_b200 import runtime, synthetic, weights
from tacotron_prosody_multi_gpu_b200 import weights as W
from tacotron_prosody_multi_gpu_b200.Utils.Hyperparams import hparams as HP
from tacotron_prosody_multi_gpu_b200.TacotronModel.modules import Attention, Decoder, Encoder
from tacotron_prosody_multi_gpu_b200.TacotronModel.modules.Tacotron import reference_encoder

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def env():
    w = weights.set_stop_bias(weights.init_weights(HP, perturb=0.05), -10.0)
    runtime.set_weights(w)
    yield HP, w, TacotronOracle(HP, w)
    runtime.close()


def _t(a):
    return a.detach().cpu().numpy()


def test_prenet_standalone_matches_oracle_and_fused_philox(env):
    hp, w, orc = env
    x = (np.random.default_rng(0).standard_normal((5, 80)) * 1.5).astype(np.float32)
    masks = philox.prenet_masks(77, 4, 5)               # [S, B, 2, 256]: the decoder's stream for seed 77
    with torch.no_grad():
        ref = orc.prenet(torch.from_numpy(x), torch.from_numpy(masks[3].astype(np.float32)))
    got = Decoder.Prenet(False, layers_sizes=hp.prenet_layers, drop_rate=hp.dropout_rate)(x, philox_seed=77, step=3)
    np.testing.assert_allclose(_t(got), ref.numpy(), atol=1e-5)


def test_decoder_rnn_and_zoneout_cell_standalone(env):
    hp, w, orc = env
    rng = np.random.default_rng(1)
    B = 6
    x = rng.standard_normal((B, 1024)).astype(np.float32)
    st = [tuple(rng.standard_normal((B, 256)).astype(np.float32) * 0.3 for _ in range(2)) for _ in range(2)]
    h2, ((c1, h1), (c2, h2z)) = Decoder.DecoderRNN(False, layers=2, size=256, zoneout=hp.zoneout_rate)(x, st)
    with torch.no_grad():
        r1, rc1, rh1 = orc.zoneout_lstm(torch.from_numpy(x), torch.from_numpy(st[0][0]), torch.from_numpy(st[0][1]),
                                        orc.w[W.DEC_LSTM % (0, 1) + "/kernel"], orc.w[W.DEC_LSTM % (0, 1) + "/bias"])
        r2, rc2, rh2 = orc.zoneout_lstm(r1, torch.from_numpy(st[1][0]), torch.from_numpy(st[1][1]),
                                        orc.w[W.DEC_LSTM % (1, 2) + "/kernel"], orc.w[W.DEC_LSTM % (1, 2) + "/bias"])
    for a, b in ((h2, r2), (c1, rc1), (h1, rh1), (c2, rc2), (h2z, rh2)):
        np.testing.assert_allclose(_t(a), b.numpy(), atol=2e-5)
    cell = Encoder.ZoneoutLSTMCell(256, False, hp.zoneout_rate, hp.zoneout_rate, name="decoder_LSTM_1")
    out, (cz, hz) = cell(x, st[0])
    np.testing.assert_allclose(_t(out), r1.numpy(), atol=2e-5)
    np.testing.assert_allclose(_t(cz), rc1.numpy(), atol=2e-5)


def test_projections_standalone(env):
    hp, w, orc = env
    x = np.random.default_rng(2).standard_normal((4, 1024)).astype(np.float32)
    fr = Decoder.FrameProjection(hp.num_mels * hp.outputs_per_step)(x)
    np.testing.assert_allclose(_t(fr), x @ w[W.FRAME_PROJ + "/kernel"] + w[W.FRAME_PROJ + "/bias"], atol=2e-5)
    logits = x @ w[W.STOP_PROJ + "/kernel"] + w[W.STOP_PROJ + "/bias"]
    np.testing.assert_allclose(_t(Decoder.StopProjection(True, shape=hp.outputs_per_step)(x)), logits, atol=2e-5)
    np.testing.assert_allclose(_t(Decoder.StopProjection(False, shape=hp.outputs_per_step)(x)), 1 / (1 + np.exp(-logits)), atol=2e-6)


def test_location_sensitive_attention_call(env):
    hp, w, orc = env
    rng = np.random.default_rng(3)
    lengths = np.asarray([17, 9, 13, 5], np.int32)
    B, T = 4, 17
    memory = (rng.standard_normal((B, T, 768)) * 0.5).astype(np.float32)
    am = Attention.LocationSensitiveAttention(hp.attention_dim, torch.from_numpy(memory).cuda(), hparams=hp, mask_encoder=True,
                                              memory_sequence_length=lengths, smoothing=hp.smoothing, cumulate_weights=hp.cumulative_weights)
    with torch.no_grad():
        values, keys, mask = orc.attention_init(torch.from_numpy(memory), lengths)
    cum_o = torch.zeros(B, T)
    state = am.initial_alignments(B)
    for step in range(3):
        q = rng.standard_normal((B, 256)).astype(np.float32)
        with torch.no_grad():
            ctx_o, al_o, cum_o = orc.attention_step(torch.from_numpy(q), cum_o, values, keys, mask)
        al, state = am(q, state)
        np.testing.assert_allclose(_t(al), al_o.numpy(), atol=2e-6)
        np.testing.assert_allclose(_t(state), cum_o.numpy(), atol=5e-6)
        np.testing.assert_allclose(_t(am.last_context), ctx_o.numpy(), atol=2e-5)
        for b, n in enumerate(lengths):
            assert float(al[b, n:].abs().max()) == 0.0 if n < T else True


def test_gst_multi_head_attention_and_reference_encoder(env):
    hp, w, orc = env
    mel = synthetic.reference_mel(hp, 3, 150)
    with torch.no_grad():
        ref_o = orc.reference_encoder(torch.from_numpy(mel))
        style_o = orc.gst_attention(ref_o)
    ref = reference_encoder(mel, hp, is_training=False)
    np.testing.assert_allclose(_t(ref), ref_o.numpy(), atol=5e-5)
    style = Attention.MultiheadAttention(ref.view(3, 1, -1), None, num_heads=hp.num_heads, num_units=hp.style_att_dim).multi_head_attention(hp)
    np.testing.assert_allclose(_t(style).reshape(3, -1), style_o.numpy().reshape(3, -1), atol=5e-5)


def test_encoder_cell_on_embedded_inputs(env):
    """The reference's call shape (Encoder.py:11-20): EncoderCell(conv, lstm)(embedded_inputs, input_lengths)."""
    hp, w, orc = env
    ids, lens = synthetic.token_batch(hp, [14, 9, 11])
    with torch.no_grad():
        emb = orc.embed(ids)
        conv_o = orc.encoder_convs(emb)
        out_o = orc.encoder_lstm(conv_o, lens)
    cell = Encoder.EncoderCell(Encoder.EncoderConvolution(False, hparams=hp), Encoder.EncoderLSTM(False, size=hp.enc_lstm_hidden_size, zoneout=hp.zoneout_rate))
    out = cell(emb.numpy(), lens)
    assert cell.conv_output_shape == (3, 14, 512)
    np.testing.assert_allclose(_t(out), out_o.numpy(), atol=1e-4)
    assert float(out[1, 9:].abs().max()) == 0.0
