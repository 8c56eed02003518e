"""Parity of the CUDA path (through the C ABI) against the CPU oracle on identical random-init
weights and seeded synthetic inputs.  fp32 kernels: tolerances are absolute, stated per test."""
import numpy as np
import pytest
import torch

from oracle import philox
from oracle.tacotron_oracle import TacotronOracle, output_lengths
from tacotron_prosody_multi_gpu_b200 import synthetic, weights
from tacotron_prosody_multi_gpu_b200.hparams import default_hparams

pytestmark = pytest.mark.gpu

ATOL_F32 = 2e-4      # teacher-forced fp32 kernels vs fp32 oracle (rtol 1e-4 / atol 1e-5 per op;
                     # accumulated over <= 40 recurrent steps)
ATOL_ALIGN = 2e-5


def pick_stop_bias(hp, base, ids, lens, gw, max_iters, philox_seed, scale=8.0):
    """Stop-parity config: the stop bias only shifts the logits (free-running feedback does
    not read them), so run the oracle once without stopping, then place the bias in the widest gap
    between the running max of the pre-stop logits and the logit that crosses.  Returns
    (weights, S, |logit| margin on every decision up to and including the stop step)."""
    w0 = weights.scale_stop_projection(base, scale, -30.0)
    ref = TacotronOracle(hp, w0).forward(ids, lens, gst_weights=gw, max_iters=max_iters, philox_seed=philox_seed)
    assert ref["steps"] == max_iters
    logit = torch.logit(ref["stop"].double().clamp(1e-300, 1.0)) + 30.0        # [B, 2S]
    m = logit.reshape(logit.shape[0], max_iters, -1).amax(dim=(0, 2)).numpy()    # per-step max over batch, r
    best = None
    run = m[0]
    for s in range(1, max_iters - 1):
        if s >= 3 and m[s] > run and (best is None or m[s] - run > best[0]):
            best = (m[s] - run, s, -(m[s] + run) / 2.0)
        run = max(run, m[s])
    assert best is not None
    gap, s, bias = best
    return weights.scale_stop_projection(base, scale, float(bias)), s + 1, gap / 2.0


@pytest.fixture(scope="module")
def ctx():
    from tacotron_prosody_multi_gpu_b200.engine import TacoEngine
    hp = default_hparams()
    w = weights.set_stop_bias(weights.init_weights(hp, perturb=0.05), -10.0)
    eng = TacoEngine(hp, w)
    return hp, w, eng, TacotronOracle(hp, w)


def _t(a):
    return a.detach().cpu().numpy()


def test_encoder_parity(ctx):
    hp, w, eng, orc = ctx
    ids, lens = synthetic.token_batch(hp, [20, 11, 16, 7, 20])
    gw = synthetic.gst_random_weights(hp)
    ref = orc.encode(ids, lens, gst_weights=gw).numpy()
    mem = eng.encode(torch.from_numpy(ids), torch.from_numpy(lens), gst_weights=gw)
    np.testing.assert_allclose(_t(mem), ref, atol=1e-4)
    assert float(mem[1, 11:, :512].abs().max()) == 0.0       # zero LSTM output beyond length


def test_reference_encoder_gst_parity(ctx):
    hp, w, eng, orc = ctx
    ids, lens = synthetic.token_batch(hp, [9, 6])
    rm = synthetic.reference_mel(hp, 2, 150)
    ref = orc.encode(ids, lens, reference_mel=rm).numpy()
    mem = eng.encode(torch.from_numpy(ids), torch.from_numpy(lens), reference_mel=torch.from_numpy(rm))
    np.testing.assert_allclose(_t(mem)[:, :, 512:], ref[:, :, 512:], atol=5e-5)
    np.testing.assert_allclose(_t(mem), ref, atol=1e-4)


@pytest.mark.parametrize("lengths", [[13], [20, 11, 16], [9] * 6 + [14] * 5, [10, 25] * 10, [12, 30, 7, 22] * 16])
def test_decoder_teacher_forced_parity(ctx, lengths):
    """Covers every batch-tile configuration: RBP 4 / 4 / 16 / 32 (U=2) and B=64 (GB=2, U=4)."""
    hp, w, eng, orc = ctx
    B = len(lengths)
    ids, lens = synthetic.token_batch(hp, lengths)
    T = ids.shape[1]
    rng = np.random.default_rng(B)
    memory = (rng.standard_normal((B, T, 768)) * 0.5).astype(np.float32)
    mel, _, _ = synthetic.mel_targets(hp, lens, frames_per_token=1, max_frames=24)
    S = mel.shape[1] // 2
    masks = philox.prenet_masks(11, S, B)
    with torch.no_grad():
        ref = orc.decode(torch.from_numpy(memory), lens, mel_targets=mel, prenet_masks=masks)
    d = eng.decode(torch.from_numpy(memory).cuda(), torch.from_numpy(lens), mel_targets=torch.from_numpy(mel),
                   prenet_masks=masks)
    assert d["steps"] == ref["steps"] == S
    np.testing.assert_allclose(_t(d["frames"]), ref["frames_raw"].numpy(), atol=ATOL_F32)
    np.testing.assert_allclose(_t(d["stop"]), ref["stop"].numpy(), atol=ATOL_F32)
    al = _t(eng.alignments(d["align_hist"], S))
    np.testing.assert_allclose(al, ref["alignments"].numpy(), atol=ATOL_ALIGN)
    for b, n in enumerate(lens):
        assert al[b, n:].max(initial=0.0) == 0.0              # exact zeros beyond length (-inf mask)
    np.testing.assert_allclose(al.sum(1), 1.0, atol=1e-5)
    st = eng.decoder_state(B, T)
    fs = ref["final_state"]
    np.testing.assert_allclose(_t(st["c1"]), fs.c1.numpy(), atol=ATOL_F32)
    np.testing.assert_allclose(_t(st["h2"]), fs.h2.numpy(), atol=ATOL_F32)
    np.testing.assert_allclose(_t(st["attention"]), fs.attention.numpy(), atol=ATOL_F32)
    np.testing.assert_allclose(_t(st["alignments"]), fs.alignments.numpy(), atol=1e-4)


def test_philox_in_kernel_equals_explicit_masks(ctx):
    hp, w, eng, orc = ctx
    ids, lens = synthetic.token_batch(hp, [10, 8, 12])
    rng = np.random.default_rng(5)
    memory = torch.from_numpy((rng.standard_normal((3, 12, 768)) * 0.5).astype(np.float32)).cuda()
    S = 7
    a = eng.decode(memory, torch.from_numpy(lens), max_iters=S, philox_seed=0x1234567890)
    b = eng.decode(memory, torch.from_numpy(lens), max_iters=S, prenet_masks=philox.prenet_masks(0x1234567890, S, 3))
    assert a["steps"] == b["steps"] == S
    assert torch.equal(a["frames"], b["frames"]) and torch.equal(a["align_hist"], b["align_hist"])


def test_free_running_stop_step_bit_exact(ctx):
    """Batch-global stop (Helpers.py:51-52): S == first step where any stop prob > 0.5, +1."""
    hp, _, _, _ = ctx
    from tacotron_prosody_multi_gpu_b200.engine import TacoEngine
    base = weights.init_weights(hp, perturb=0.05)
    ids, lens = synthetic.token_batch(hp, [14, 9, 12, 11])
    gw = synthetic.gst_random_weights(hp)
    w, S, margin = pick_stop_bias(hp, base, ids, lens, gw, max_iters=40, philox_seed=99)
    assert 3 <= S < 40 and margin >= 0.01, (S, margin)
    ref = TacotronOracle(hp, w).forward(ids, lens, gst_weights=gw, max_iters=40, philox_seed=99)
    assert ref["steps"] == S
    eng = TacoEngine(hp, w)
    out = eng.forward(torch.from_numpy(ids), torch.from_numpy(lens), gst_weights=gw, max_iters=40, philox_seed=99)
    assert out["steps"] == S                                               # bit-exact stop step
    assert out["mel_outputs"].shape == (4, 2 * S, 80) and out["alignments"].shape == (4, 14, S)
    np.testing.assert_allclose(_t(out["mel_outputs"]), ref["mel_outputs"].numpy(), atol=5e-4)
    assert output_lengths(_t(out["stop"])) == output_lengths(ref["stop"].numpy())   # bit-exact output lengths
    eng.close()


def test_postnet_parity(ctx):
    hp, w, eng, orc = ctx
    rng = np.random.default_rng(3)
    frames = (rng.standard_normal((3, 38, 80)) * 3.0).astype(np.float32)      # exercises the clip
    with torch.no_grad():
        dec_ref = torch.clamp(torch.from_numpy(frames), -4.1, 4.0)
        mel_ref = orc.postnet(dec_ref).numpy()
    dec, mel = eng.postnet(torch.from_numpy(frames).cuda(), 38)
    assert torch.equal(dec.cpu(), dec_ref)
    np.testing.assert_allclose(_t(mel), mel_ref, atol=2e-4)
    # F < F_stride: only the first F frames are consumed
    dec2, mel2 = eng.postnet(torch.from_numpy(frames).cuda(), 20)
    with torch.no_grad():
        mel_ref2 = orc.postnet(dec_ref[:, :20]).numpy()
    np.testing.assert_allclose(_t(mel2), mel_ref2, atol=2e-4)


def test_full_forward_c1_like(ctx):
    """C1-shaped (B=8 teacher-forced / GTA) at a size the oracle finishes in seconds."""
    hp, w, eng, orc = ctx
    ids, lens = synthetic.token_batch(hp, synthetic.SENTENCE_LENGTHS[:8])
    mel, _, _ = synthetic.mel_targets(hp, lens, frames_per_token=1)
    gw = synthetic.gst_random_weights(hp)
    ref = orc.forward(ids, lens, mel_targets=mel, gst_weights=gw, philox_seed=hp.tacotron_random_seed)
    out = eng.forward(torch.from_numpy(ids), torch.from_numpy(lens), mel_targets=torch.from_numpy(mel), gst_weights=gw,
                      philox_seed=hp.tacotron_random_seed)
    assert out["steps"] == ref["steps"] == mel.shape[1] // 2
    np.testing.assert_allclose(_t(out["encoder_outputs"]), ref["encoder_outputs"].numpy(), atol=2e-4)
    np.testing.assert_allclose(_t(out["decoder_output"]), ref["decoder_output"].numpy(), atol=5e-4)
    np.testing.assert_allclose(_t(out["mel_outputs"]), ref["mel_outputs"].numpy(), atol=1e-3)
    np.testing.assert_allclose(_t(out["alignments"]), ref["alignments"].numpy(), atol=5e-5)
    np.testing.assert_allclose(_t(out["stop"]), ref["stop"].numpy(), atol=5e-4)


def test_cell_stepping_matches_one_launch(ctx):
    """taco_decode(reset_state=0, S_max=1) == `cell(inputs, state)` stepping (Decoder.py:246-292)."""
    hp, w, eng, orc = ctx
    ids, lens = synthetic.token_batch(hp, [10, 8])
    rng = np.random.default_rng(8)
    memory = torch.from_numpy((rng.standard_normal((2, 10, 768)) * 0.5).astype(np.float32)).cuda()
    S = 5
    whole = eng.decode(memory, torch.from_numpy(lens), max_iters=S, philox_seed=3)
    frames = []
    for t in range(S):
        d = eng.decode(memory, torch.from_numpy(lens), max_iters=1, philox_seed=3, reset_state=(t == 0))
        frames.append(d["frames"].clone())
    assert torch.equal(torch.cat(frames, 1), whole["frames"])


# ------------------------------------------------------------------ tcgen05 (bf16 operand) conv path
ATOL_TC_MIRROR = 1.2e-2  # tcgen05 kernels vs the oracle with the SAME bf16 operand rounding (conv_bf16=True):
                        # differences are fp32 accumulation order + bf16 re-rounding of near-tie activations
ATOL_TC_FP32 = 6e-2     # vs the fp32 oracle: bf16 operand rounding through 6 stacked contractions (K = 3840)


@pytest.mark.parametrize("B,F", [(3, 38), (2, 300), (5, 128)])
def test_postnet_tcgen05_parity(ctx, B, F):
    hp, w, _, orc = ctx
    from tacotron_prosody_multi_gpu_b200.engine import TacoEngine
    eng = TacoEngine(hp, w, precision="bf16")
    rng = np.random.default_rng(3 + F)
    frames = (rng.standard_normal((B, F + 6, 80)) * 3.0).astype(np.float32)
    with torch.no_grad():
        dec_ref = torch.clamp(torch.from_numpy(frames[:, :F]), -4.1, 4.0)
        mel_mirror = TacotronOracle(hp, w, conv_bf16=True).postnet(dec_ref).numpy()
        mel_fp32 = orc.postnet(dec_ref).numpy()
    dec, mel = eng.postnet(torch.from_numpy(frames).cuda(), F)
    assert torch.equal(dec.cpu(), dec_ref)
    e_mirror = float(np.abs(_t(mel) - mel_mirror).max())
    e_fp32 = float(np.abs(_t(mel) - mel_fp32).max())
    e_round = float(np.abs(mel_mirror - mel_fp32).max())
    print("tcgen05 postnet B=%d F=%d: |gpu-mirror|=%.2e |gpu-fp32|=%.2e |mirror-fp32|=%.2e rms(gpu-mirror)=%.2e" % (
        B, F, e_mirror, e_fp32, e_round, float(np.sqrt(np.mean((_t(mel) - mel_mirror) ** 2)))))
    np.testing.assert_allclose(_t(mel), mel_mirror, atol=ATOL_TC_MIRROR)
    np.testing.assert_allclose(_t(mel), mel_fp32, atol=ATOL_TC_FP32)
    # the kernel is closer to the bf16-rounding oracle than that oracle is to the fp32 one
    assert e_mirror < e_round
    eng.close()


# ------------------------------------------------------------------ cluster decoder, decoder3.cu (bf16 tensor-core contractions)
# Tolerances of TACO_PRECISION_BF16 against the oracle that rounds the SAME operands to bf16 (dec_bf16=True):
# remaining differences are fp32 accumulation order, tanh.approx in the energies (|err| <= 2^-10.99 per PTX ISA),
# the hi/lo-split fused location contraction (~2^-16 relative) and bf16 re-rounding of near-tie activations.
ATOL_D2_FRAMES = 3e-2
ATOL_D2_ALIGN = 5e-3
# ... and against the plain fp32 oracle (bf16 operand rounding through <= 24 recurrent steps)
ATOL_D2_FRAMES_FP32 = 1e-1
ATOL_D2_ALIGN_FP32 = 2e-2


@pytest.fixture(scope="module")
def ctx2():
    from tacotron_prosody_multi_gpu_b200.engine import TacoEngine
    hp = default_hparams()
    w = weights.set_stop_bias(weights.init_weights(hp, perturb=0.05), -10.0)
    eng = TacoEngine(hp, w, precision="bf16")
    return hp, w, eng, TacotronOracle(hp, w, conv_bf16=True, dec_bf16=True), TacotronOracle(hp, w)


@pytest.mark.parametrize("lengths", [[13], [20, 11, 16], [9] * 6 + [14] * 5, [10, 25] * 10, [12, 30, 7, 22] * 16])
def test_cluster_decoder_teacher_forced_parity(ctx2, lengths):
    """Covers MT = 1 / 1 / 1 / 2 / 4 M-tiles and NS = 96 / 32 / 8 / 6 / 2 context slices."""
    hp, w, eng, mir, orc = ctx2
    B = len(lengths)
    ids, lens = synthetic.token_batch(hp, lengths)
    T = ids.shape[1]
    rng = np.random.default_rng(B)
    memory = (rng.standard_normal((B, T, 768)) * 0.5).astype(np.float32)
    mel, _, _ = synthetic.mel_targets(hp, lens, frames_per_token=1, max_frames=24)
    S = mel.shape[1] // 2
    masks = philox.prenet_masks(11, S, B)
    with torch.no_grad():
        ref = mir.decode(torch.from_numpy(memory), lens, mel_targets=mel, prenet_masks=masks)
        ref32 = orc.decode(torch.from_numpy(memory), lens, mel_targets=mel, prenet_masks=masks)
    d = eng.decode(torch.from_numpy(memory).cuda(), torch.from_numpy(lens), mel_targets=torch.from_numpy(mel),
                   prenet_masks=masks)
    assert d["steps"] == ref["steps"] == S
    al = _t(eng.alignments(d["align_hist"], S))
    print("cluster decoder B=%d: |frames-mirror|=%.2e |frames-fp32|=%.2e |mirror-fp32|=%.2e |align-mirror|=%.2e" % (
        B, np.abs(_t(d["frames"]) - ref["frames_raw"].numpy()).max(), np.abs(_t(d["frames"]) - ref32["frames_raw"].numpy()).max(),
        np.abs(ref["frames_raw"].numpy() - ref32["frames_raw"].numpy()).max(), np.abs(al - ref["alignments"].numpy()).max()))
    np.testing.assert_allclose(_t(d["frames"]), ref["frames_raw"].numpy(), atol=ATOL_D2_FRAMES)
    np.testing.assert_allclose(_t(d["stop"]), ref["stop"].numpy(), atol=ATOL_D2_FRAMES)
    np.testing.assert_allclose(al, ref["alignments"].numpy(), atol=ATOL_D2_ALIGN)
    np.testing.assert_allclose(_t(d["frames"]), ref32["frames_raw"].numpy(), atol=ATOL_D2_FRAMES_FP32)
    np.testing.assert_allclose(al, ref32["alignments"].numpy(), atol=ATOL_D2_ALIGN_FP32)
    for b, n in enumerate(lens):
        assert al[b, n:].max(initial=0.0) == 0.0              # exact zeros beyond length (-inf mask)
    np.testing.assert_allclose(al.sum(1), 1.0, atol=1e-5)
    st = eng.decoder_state(B, T)
    fs = ref["final_state"]
    np.testing.assert_allclose(_t(st["c1"]), fs.c1.numpy(), atol=ATOL_D2_FRAMES)
    np.testing.assert_allclose(_t(st["h2"]), fs.h2.numpy(), atol=ATOL_D2_FRAMES)
    np.testing.assert_allclose(_t(st["attention"]), fs.attention.numpy(), atol=ATOL_D2_FRAMES)
    np.testing.assert_allclose(_t(st["alignments"]), fs.alignments.numpy(), atol=5e-2)


def test_cluster_decoder_philox_equals_explicit_masks(ctx2):
    hp, w, eng, _, _ = ctx2
    ids, lens = synthetic.token_batch(hp, [10, 8, 12])
    rng = np.random.default_rng(5)
    memory = torch.from_numpy((rng.standard_normal((3, 12, 768)) * 0.5).astype(np.float32)).cuda()
    S = 7
    a = eng.decode(memory, torch.from_numpy(lens), max_iters=S, philox_seed=0x1234567890)
    b = eng.decode(memory, torch.from_numpy(lens), max_iters=S, prenet_masks=philox.prenet_masks(0x1234567890, S, 3))
    assert a["steps"] == b["steps"] == S
    assert torch.equal(a["frames"], b["frames"]) and torch.equal(a["align_hist"], b["align_hist"])


def test_cluster_decoder_cell_stepping_matches_one_launch(ctx2):
    hp, w, eng, _, _ = ctx2
    ids, lens = synthetic.token_batch(hp, [10, 8])
    rng = np.random.default_rng(8)
    memory = torch.from_numpy((rng.standard_normal((2, 10, 768)) * 0.5).astype(np.float32)).cuda()
    S = 5
    whole = eng.decode(memory, torch.from_numpy(lens), max_iters=S, philox_seed=3)
    frames = []
    for t in range(S):
        d = eng.decode(memory, torch.from_numpy(lens), max_iters=1, philox_seed=3, reset_state=(t == 0))
        frames.append(d["frames"].clone())
    assert torch.equal(torch.cat(frames, 1), whole["frames"])


def test_cluster_decoder_free_running_stop_step_bit_exact(ctx2):
    """Batch-global stop: S and output lengths equal the bf16-mirroring oracle's exactly."""
    hp, _, _, _, _ = ctx2
    from tacotron_prosody_multi_gpu_b200.engine import TacoEngine
    base = weights.init_weights(hp, perturb=0.05)
    ids, lens = synthetic.token_batch(hp, [14, 9, 12, 11])
    gw = synthetic.gst_random_weights(hp)
    w, S, margin = pick_stop_bias(hp, base, ids, lens, gw, max_iters=40, philox_seed=99)
    mir = TacotronOracle(hp, w, conv_bf16=True, dec_bf16=True)
    ref = mir.forward(ids, lens, gst_weights=gw, max_iters=40, philox_seed=99)
    eng = TacoEngine(hp, w, precision="bf16")
    out = eng.forward(torch.from_numpy(ids), torch.from_numpy(lens), gst_weights=gw, max_iters=40, philox_seed=99)
    print("cluster decoder free-run: S gpu=%d mirror=%d fp32-oracle=%d margin=%.3f" % (out["steps"], ref["steps"], S, margin))
    assert out["steps"] == ref["steps"]
    assert out["mel_outputs"].shape == (4, 2 * ref["steps"], 80)
    np.testing.assert_allclose(_t(out["mel_outputs"]), ref["mel_outputs"].numpy(), atol=5e-2)
    assert output_lengths(_t(out["stop"])) == output_lengths(ref["stop"].numpy())
    eng.close()


# ------------------------------------------------------------------ full-size configurations of BASELINE.json
def test_c2_batch64_T111_teacher_forced_parity(ctx2):
    """C2 shape (B=64, the sentences.txt lengths x4, T=111) on the cluster decoder, teacher-forced so the
    oracle finishes in seconds; covers the 7/8-cluster row split and the in-register `pre` path (T <= 128)."""
    hp, w, eng, mir, orc = ctx2
    lengths = synthetic.cycle_lengths(64)
    ids, lens = synthetic.token_batch(hp, lengths)
    B, T = ids.shape
    assert (B, T) == (64, 111)
    rng = np.random.default_rng(64)
    memory = (rng.standard_normal((B, T, 768)) * 0.5).astype(np.float32)
    mel, _, _ = synthetic.mel_targets(hp, lens, frames_per_token=1, max_frames=32)
    S = mel.shape[1] // 2
    with torch.no_grad():
        ref = mir.decode(torch.from_numpy(memory), lens, mel_targets=mel, philox_seed=21)
    d = eng.decode(torch.from_numpy(memory).cuda(), torch.from_numpy(lens), mel_targets=torch.from_numpy(mel), philox_seed=21)
    assert d["steps"] == ref["steps"] == S
    al = _t(eng.alignments(d["align_hist"], S))
    np.testing.assert_allclose(_t(d["frames"]), ref["frames_raw"].numpy(), atol=ATOL_D2_FRAMES)
    np.testing.assert_allclose(al, ref["alignments"].numpy(), atol=ATOL_D2_ALIGN)
    for b, n in enumerate(lens):
        assert al[b, n:].max(initial=0.0) == 0.0
    np.testing.assert_allclose(al.sum(1), 1.0, atol=1e-5)


def test_c5_long_form_reference_encoder(ctx2):
    """C5: B=16, T=300 tokens, prosody reference encoder on a [16, 2000, 80] reference mel, S=1000 decoder
    steps.  (i) the first 48 steps teacher-forced against the bf16-mirroring oracle (T > 128: the `pre`
    scratch path of the cluster decoder); (ii) the full 1000 free-running steps through size-independent
    properties: S == max_iters, alignment rows are distributions over the valid positions and exactly 0
    beyond the length, outputs finite and inside the clip range, frames(t) of a second identical launch
    are bit-identical (Philox dropout is keyed by (seed, step, row), not by launch)."""
    hp, w, eng, mir, orc = ctx2
    lengths = [300, 287, 300, 251, 199, 300, 123, 64, 300, 18, 277, 300, 150, 290, 300, 42]
    ids, lens = synthetic.token_batch(hp, lengths)
    B, T = ids.shape
    rm = synthetic.reference_mel(hp, B, 2000)
    with torch.no_grad():
        enc_ref = mir.encode(ids, lens, reference_mel=rm)
    mem = eng.encode(torch.from_numpy(ids), torch.from_numpy(lens), reference_mel=torch.from_numpy(rm))
    np.testing.assert_allclose(_t(mem)[:, :, 512:], enc_ref.numpy()[:, :, 512:], atol=2e-3)   # style embedding
    np.testing.assert_allclose(_t(mem), enc_ref.numpy(), atol=ATOL_TC_FP32)
    # (i) teacher-forced prefix from the ORACLE's memory so the decoder is compared on identical inputs
    mel, _, _ = synthetic.mel_targets(hp, lens, frames_per_token=1, max_frames=96)
    S = mel.shape[1] // 2
    with torch.no_grad():
        ref = mir.decode(enc_ref, lens, mel_targets=mel, philox_seed=5)
    d = eng.decode(enc_ref.cuda(), torch.from_numpy(lens), mel_targets=torch.from_numpy(mel), philox_seed=5)
    assert d["steps"] == ref["steps"] == S
    al = _t(eng.alignments(d["align_hist"], S))
    np.testing.assert_allclose(_t(d["frames"]), ref["frames_raw"].numpy(), atol=ATOL_D2_FRAMES)
    np.testing.assert_allclose(al, ref["alignments"].numpy(), atol=ATOL_D2_ALIGN)
    # (ii) the whole long-form pass
    out = eng.forward(torch.from_numpy(ids), torch.from_numpy(lens), reference_mel=torch.from_numpy(rm), max_iters=1000, philox_seed=5)
    assert out["steps"] == 1000
    assert out["mel_outputs"].shape == (B, 2000, 80) and out["alignments"].shape == (B, T, 1000)
    al = _t(out["alignments"])
    assert np.isfinite(al).all() and np.isfinite(_t(out["mel_outputs"])).all()
    np.testing.assert_allclose(al.sum(1), 1.0, atol=2e-5)
    for b, n in enumerate(lens):
        assert al[b, n:].max(initial=0.0) == 0.0
    mo = _t(out["mel_outputs"])
    assert mo.min() >= -4.1 - 1e-6 and mo.max() <= 4.0 + 1e-6
    out2 = eng.forward(torch.from_numpy(ids), torch.from_numpy(lens), reference_mel=torch.from_numpy(rm), max_iters=1000, philox_seed=5)
    assert torch.equal(out["frames_raw"], out2["frames_raw"])


@pytest.mark.parametrize("precision", ["fp32", "bf16"])
def test_smoothing_normalisation_ragged_batch(precision):
    """hparams.smoothing=True (Attention.py:405-425): sigmoid(e) / sum sigmoid(e) with the TF 1.9 length mask applied BEFORE the
    normalisation (_maybe_mask_score wraps every probability_fn): padded positions get exactly 0 on a ragged batch."""
    from tacotron_prosody_multi_gpu_b200.engine import TacoEngine
    hp = default_hparams()
    hp.smoothing = True
    w = weights.set_stop_bias(weights.init_weights(hp, perturb=0.05), -10.0)
    lengths = [21, 9, 14, 21, 5, 17, 12, 3, 19, 11, 8]
    ids, lens = synthetic.token_batch(hp, lengths)
    B, T = ids.shape
    memory = (np.random.default_rng(7).standard_normal((B, T, 768)) * 0.5).astype(np.float32)
    mel, _, _ = synthetic.mel_targets(hp, np.full(B, T, np.int32), frames_per_token=2, max_frames=40)
    S = mel.shape[1] // 2
    bf = precision == "bf16"
    with torch.no_grad():
        ref = TacotronOracle(hp, w, conv_bf16=bf, dec_bf16=bf).decode(torch.from_numpy(memory), lens, mel_targets=mel, philox_seed=4)
    eng = TacoEngine(hp, w, precision=precision)
    d = eng.decode(torch.from_numpy(memory).cuda(), torch.from_numpy(lens), mel_targets=torch.from_numpy(mel), philox_seed=4)
    assert d["steps"] == ref["steps"] == S
    al = _t(eng.alignments(d["align_hist"], S))
    np.testing.assert_allclose(al, ref["alignments"].numpy(), atol=ATOL_D2_ALIGN if bf else ATOL_ALIGN)
    np.testing.assert_allclose(_t(d["frames"]), ref["frames_raw"].numpy(), atol=ATOL_D2_FRAMES if bf else ATOL_F32)
    for b, n in enumerate(lens):
        assert al[b, n:].max(initial=0.0) == 0.0
    np.testing.assert_allclose(al.sum(1), 1.0, atol=1e-5)
    eng.close()


def test_bf16_mode_beyond_cluster_decoder_limits_is_exact_and_announced(capfd):
    """T > 512 does not fit the bf16 cluster decoder: the batch is decoded by the fp32 persistent kernel (fp32 tolerances hold)
    and the library says so on stderr -- never a silent change of kernel."""
    from tacotron_prosody_multi_gpu_b200.engine import TacoEngine
    hp = default_hparams()
    w = weights.set_stop_bias(weights.init_weights(hp, perturb=0.05), -10.0)
    ids, lens = synthetic.token_batch(hp, [530, 64])
    B, T = ids.shape
    memory = (np.random.default_rng(9).standard_normal((B, T, 768)) * 0.5).astype(np.float32)
    S = 6
    with torch.no_grad():
        ref = TacotronOracle(hp, w).decode(torch.from_numpy(memory), lens, max_iters=S, philox_seed=4)
    eng = TacoEngine(hp, w, precision="bf16")
    d = eng.decode(torch.from_numpy(memory).cuda(), torch.from_numpy(lens), max_iters=S, philox_seed=4)
    assert d["steps"] == S
    np.testing.assert_allclose(_t(d["frames"]), ref["frames_raw"].numpy(), atol=ATOL_F32)
    err = capfd.readouterr().err
    assert "does not fit the bf16 cluster decoder" in err
    eng.close()


def test_encoder_bilstm_32_cta_grid_equals_default(ctx):
    """taco_set_encoder_ctas(32): the BiLSTM with 16 units per CTA (the grid that fits beside a running decoder) against the oracle
    and against the default 128-CTA grid at B = 64."""
    hp, w, eng, orc = ctx
    ids, lens = synthetic.token_batch(hp, synthetic.cycle_lengths(64)[:40] + [7, 3] * 12)
    gw = synthetic.gst_random_weights(hp)
    a = _t(eng.encode(torch.from_numpy(ids), torch.from_numpy(lens), gst_weights=gw))
    eng.set_encoder_ctas(32)
    try:
        b = _t(eng.encode(torch.from_numpy(ids), torch.from_numpy(lens), gst_weights=gw))
    finally:
        eng.set_encoder_ctas(0)
    ref = orc.encode(ids, lens, gst_weights=gw).numpy()
    np.testing.assert_allclose(b, ref, atol=1e-4)
    np.testing.assert_allclose(b, a, atol=2e-5)
    for i, n in enumerate(lens):
        assert float(np.abs(b[i, n:, :512]).max(initial=0.0)) == 0.0
