"""Parity of the CUDA decoder against the CPU oracle AT THE SIZES THE BENCH IS MEASURED ON (BASELINE.json configs C2 / C5):
the full 720 (1000) decoder steps, teacher-forced and free-running, bf16 product path and fp32 path, plus a batch-global stop
that fires mid-sequence at B=64.  The tolerances are 4-5x the maxima of the error-vs-step curves committed under
profiles/r02_parity_curves.json (scripts/parity_curves.py): the error saturates after ~100 steps, it does not grow with S.

  bf16 kernels vs the oracle that rounds the SAME operands to bf16:  frames 3e-3, alignments 2e-4   (measured 6.5e-4 / 3.0e-5)
  bf16 kernels vs the plain fp32 oracle:                             frames 1.2e-2, alignments 2e-3 (measured 2.6e-3 / 3.8e-4)
  fp32 kernels vs the fp32 oracle:                                   frames 5e-6, alignments 2e-6   (measured 6.3e-7 / 1.9e-7)
"""
import numpy as np
import pytest
import torch

from oracle.tacotron_oracle import TacotronOracle, output_lengths
from tacotron_prosody_multi_gpu_b200 import synthetic, weights
from tacotron_prosody_multi_gpu_b200.hparams import default_hparams

pytestmark = pytest.mark.gpu

TOL = {"mirror": (3e-3, 2e-4), "fp32": (1.2e-2, 2e-3), "exact": (5e-6, 2e-6)}
MARKS = (16, 100, 360, 720, 1000)
C5_LENGTHS = [300, 287, 300, 251, 199, 300, 123, 64, 300, 18, 277, 300, 150, 290, 300, 42]


@pytest.fixture(scope="module")
def base():
    hp = default_hparams()
    return hp, weights.set_stop_bias(weights.init_weights(hp, perturb=0.05), -10.0)


def _engine(hp, w, precision):
    from tacotron_prosody_multi_gpu_b200.engine import TacoEngine
    return TacoEngine(hp, w, precision=precision)


def _check_prefixes(name, gpu, ref, step_axis, per_step, tol):
    """max |gpu - ref| over the first m steps for every mark m <= S must stay within tol (the curve is flat, not growing)."""
    d = np.abs(np.asarray(gpu, np.float64) - np.asarray(ref, np.float64))
    d = np.moveaxis(d, step_axis, 0)
    d = d.reshape(d.shape[0] // per_step, -1).max(axis=1)
    S = d.shape[0]
    got = {m: float(d[:m].max()) for m in MARKS if m <= S}
    print("%s: max err by step %s (tol %.1e)" % (name, {k: "%.2e" % v for k, v in got.items()}, tol))
    assert float(d.max()) <= tol, (name, got)
    return got


def _decode_case(hp, w, lengths, S, teacher, precision, seed=21):
    ids, lens = synthetic.token_batch(hp, lengths)
    B, T = ids.shape
    memory = (np.random.default_rng(64).standard_normal((B, T, 768)) * 0.5).astype(np.float32)
    mel = None
    if teacher:
        mel, _, _ = synthetic.mel_targets(hp, np.full(B, T, np.int32), frames_per_token=13, max_frames=2 * S)
        mel = mel[:, :2 * S]
    eng = _engine(hp, w, precision)
    d = eng.decode(torch.from_numpy(memory).cuda(), torch.from_numpy(lens), mel_targets=None if mel is None else torch.from_numpy(mel),
                   max_iters=S, philox_seed=seed)
    assert d["steps"] == S
    out = dict(frames=d["frames"].cpu().numpy(), stop=d["stop"].cpu().numpy(), align=eng.alignments(d["align_hist"], S).cpu().numpy())
    eng.close()
    refs = {}
    with torch.no_grad():
        if precision == "bf16":
            refs["mirror"] = TacotronOracle(hp, w, conv_bf16=True, dec_bf16=True).decode(torch.from_numpy(memory), lens, mel_targets=mel, max_iters=S, philox_seed=seed)
            refs["fp32"] = TacotronOracle(hp, w).decode(torch.from_numpy(memory), lens, mel_targets=mel, max_iters=S, philox_seed=seed)
        else:
            refs["exact"] = TacotronOracle(hp, w).decode(torch.from_numpy(memory), lens, mel_targets=mel, max_iters=S, philox_seed=seed)
    for k, ref in refs.items():
        assert ref["steps"] == S
        tf, ta = TOL[k]
        _check_prefixes("frames vs %s" % k, out["frames"], ref["frames_raw"].numpy(), 1, 2, tf)
        _check_prefixes("stop vs %s" % k, out["stop"], ref["stop"].numpy(), 1, 2, tf)
        _check_prefixes("alignments vs %s" % k, out["align"], ref["alignments"].numpy(), 2, 1, ta)
    for b, n in enumerate(lens):
        assert out["align"][b, n:].max(initial=0.0) == 0.0          # exact zeros beyond the length (-inf mask)
    np.testing.assert_allclose(out["align"].sum(1), 1.0, atol=2e-5)


@pytest.mark.parametrize("teacher", [True, False], ids=["teacher_forced", "free_running"])
def test_c2_b64_all_720_steps_bf16(base, teacher):
    """The headline configuration itself: B=64, T=111, S=720, bf16 product path (decoder3 + bf16 values)."""
    hp, w = base
    _decode_case(hp, w, synthetic.cycle_lengths(64), 720, teacher, "bf16")


def test_c2_b64_all_720_steps_fp32(base):
    hp, w = base
    _decode_case(hp, w, synthetic.cycle_lengths(64), 720, True, "fp32")


@pytest.mark.parametrize("precision", ["bf16", "fp32"])
@pytest.mark.parametrize("T,S", [(42, 277), (111, 720)])
def test_c2_batch1_free_running(base, precision, T, S):
    """synthesize.py --mode=inference runs one sentence at a time (TacotronModel/synthesize.py:68-88): B=1, both precisions."""
    hp, w = base
    _decode_case(hp, w, [T], S, False, precision)


def test_c5_all_1000_steps_teacher_forced_bf16(base):
    """C5 decoder: B=16, T=300 (the T > 128 scratch path of the cluster decoder), 1000 steps."""
    hp, w = base
    _decode_case(hp, w, C5_LENGTHS, 1000, True, "bf16")


def test_c2_b64_stop_fires_mid_sequence_bit_exact(base):
    """Batch-global stop at B=64 (Utils/Helpers.py:36-59): the stop kernel is scaled x8 and the bias placed in the widest gap of
    the oracle's per-step maximum logit, so that the first crossing lands mid-sequence with a margin; the GPU must stop
    at exactly the same step and trim every utterance to exactly the same length (Utils/Tacotron_feeder.py:92-95)."""
    hp, w0 = base
    S_cap = 720
    ids, lens = synthetic.token_batch(hp, synthetic.cycle_lengths(64))
    memory = (np.random.default_rng(64).standard_normal((64, ids.shape[1], 768)) * 0.5).astype(np.float32)
    mem_t = torch.from_numpy(memory)
    with torch.no_grad():
        ref0 = TacotronOracle(hp, weights.scale_stop_projection(w0, 8.0, -30.0), conv_bf16=True, dec_bf16=True).decode(
            mem_t, lens, max_iters=S_cap, philox_seed=99)
    assert ref0["steps"] == S_cap
    logit = torch.logit(ref0["stop"].double().clamp(1e-300, 1.0)) + 30.0
    m = logit.reshape(64, S_cap, -1).amax(dim=(0, 2)).numpy()
    best, run = None, m[:100].max()
    for s in range(100, S_cap - 1):
        if m[s] > run and (best is None or m[s] - run > best[0]):
            best = (m[s] - run, s, -(m[s] + run) / 2.0)
        run = max(run, m[s])
    assert best is not None
    gap, s_stop, bias = best
    assert gap / 2.0 > 1e-2, "stop margin %.2e too small for a bit-exact comparison" % (gap / 2.0)
    w = weights.scale_stop_projection(w0, 8.0, float(bias))
    with torch.no_grad():
        ref = TacotronOracle(hp, w, conv_bf16=True, dec_bf16=True).decode(mem_t, lens, max_iters=S_cap, philox_seed=99)
    S = ref["steps"]
    assert S == s_stop + 1 and 100 < S < S_cap
    eng = _engine(hp, w, "bf16")
    d = eng.decode(mem_t.cuda(), torch.from_numpy(lens), max_iters=S_cap, philox_seed=99)
    print("B=64 stop: oracle S=%d gpu S=%d logit margin %.3e" % (S, d["steps"], gap / 2.0))
    assert d["steps"] == S                                                                # bit-exact stop step
    stop_gpu = d["stop"].cpu().numpy()[:, :2 * S]
    assert output_lengths(stop_gpu) == output_lengths(ref["stop"].numpy())                # bit-exact output lengths
    np.testing.assert_allclose(d["frames"].cpu().numpy()[:, :2 * S], ref["frames_raw"].numpy(), atol=TOL["mirror"][0])
    eng.close()
