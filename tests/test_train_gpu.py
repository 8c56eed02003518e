"""Training primitives and the whole training step on the GPU (through the C ABI of include/taco_b200_train.h)
against (i) the torch-CPU twin of each primitive and (ii) autograd through the oracle."""
import numpy as np
import pytest
import torch

from cpu_train_ops import TorchOps
from oracle import train_oracle
from tacotron_prosody_multi_gpu_b200 import synthetic, weights
from tacotron_prosody_multi_gpu_b200.hparams import default_hparams
from tacotron_prosody_multi_gpu_b200.train.params import is_regularized
from tacotron_prosody_multi_gpu_b200.train.step import TrainStep

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ops():
    from tacotron_prosody_multi_gpu_b200.train.ops import CudaOps
    return CudaOps(0), TorchOps()


def both(ops, name, args, atol=1e-5, rtol=1e-4):
    """Run primitive ``name`` on the GPU and on the twin with identical arguments (tensors are cloned to each
    side) and compare EVERY tensor argument afterwards (outputs are written in place)."""
    cu, tw = ops
    a_gpu = [a.clone().cuda() if isinstance(a, torch.Tensor) else a for a in args]
    a_cpu = [a.clone() if isinstance(a, torch.Tensor) else a for a in args]
    getattr(cu, name)(*a_gpu)
    getattr(tw, name)(*a_cpu)
    torch.cuda.synchronize()
    for i, (g, c) in enumerate(zip(a_gpu, a_cpu)):
        if isinstance(c, torch.Tensor) and c.dtype == torch.float32:
            np.testing.assert_allclose(g.cpu().numpy(), c.numpy(), atol=atol, rtol=rtol, err_msg="%s arg %d" % (name, i))
    return a_gpu, a_cpu


def R(*shape, seed=0, scale=1.0):
    g = torch.Generator().manual_seed(seed + sum(shape))
    return torch.randn(*shape, generator=g) * scale


@pytest.mark.parametrize("M,N,K,ta,tb,beta,bias", [(5, 7, 9, 0, 0, 0.0, 0), (33, 65, 130, 1, 0, 0.0, 1), (64, 1024, 768, 0, 1, 1.0, 0),
                                                    (300, 257, 45, 1, 1, 0.5, 1), (1280, 1024, 333, 1, 0, 0.0, 0), (2000, 1100, 64, 0, 0, 0.0, 1)])
def test_gemm(ops, M, N, K, ta, tb, beta, bias):
    A = R(K, M) if ta else R(M, K)
    B = R(N, K) if tb else R(K, N)
    both(ops, "gemm", [A, B, R(M, N, seed=3), bool(ta), bool(tb), beta, R(N) if bias else None], atol=2e-4 * np.sqrt(K))


def test_gemm_strided_views(ops):
    cu, tw = ops
    big = R(40, 96)
    Cg, Cc = torch.zeros(20, 64).cuda(), torch.zeros(20, 64)
    Bm = R(96, 64)
    cu.gemm(big.cuda()[10:30], Bm.cuda(), Cg)
    tw.gemm(big[10:30], Bm, Cc)
    np.testing.assert_allclose(Cg.cpu().numpy(), Cc.numpy(), atol=1e-4)
    # column-block destination + strided rows ([S,B,T] -> [:, b, :])
    X = R(6, 3, 10)
    D = R(6, 3, 8)
    og, oc = torch.zeros(10, 8).cuda(), torch.zeros(10, 8)
    cu.gemm(X.cuda()[:, 1, :], D.cuda()[:, 1, :], og, ta=True)
    tw.gemm(X[:, 1, :], D[:, 1, :], oc, ta=True)
    np.testing.assert_allclose(og.cpu().numpy(), oc.numpy(), atol=1e-4)


@pytest.mark.parametrize("B,T,Ci,Co,k", [(2, 7, 8, 12, 5), (3, 130, 80, 768, 5), (2, 50, 768, 80, 5), (4, 33, 512, 512, 5), (2, 9, 1, 32, 31)])
def test_conv1d_fwd_dgrad_wgrad(ops, B, T, Ci, Co, k):
    x, w, dy = R(B, T, Ci), R(k, Ci, Co, scale=0.1), R(B, T, Co)
    both(ops, "conv1d_fwd", [x, w, R(Co), torch.zeros(B, T, Co)], atol=1e-3)
    both(ops, "conv1d_dgrad", [dy, w, torch.zeros(B, T, Ci)], atol=1e-3)
    both(ops, "conv1d_wgrad", [x, dy, torch.zeros(k, Ci, Co)], atol=2e-3)


@pytest.mark.parametrize("B,H,W,Ci,Co", [(2, 12, 80, 1, 32), (2, 7, 5, 32, 64), (3, 3, 3, 64, 128), (2, 16, 10, 32, 32)])
def test_conv2d_fwd_dgrad_wgrad(ops, B, H, W, Ci, Co):
    Ho, Wo = (H + 1) // 2, (W + 1) // 2
    x, w, dy = R(B, H, W, Ci), R(3, 3, Ci, Co, scale=0.2), R(B, Ho, Wo, Co)
    both(ops, "conv2d_fwd", [x, w, R(Co), torch.zeros(B, Ho, Wo, Co)], atol=1e-4)
    both(ops, "conv2d_dgrad", [dy, w, torch.zeros(B, H, W, Ci)], atol=1e-4)
    both(ops, "conv2d_wgrad", [x, dy, torch.zeros(3, 3, Ci, Co)], atol=5e-4)


@pytest.mark.parametrize("C", [96, 30], ids=["c96_16byte_kernels", "c30_scalar_kernels"])
@pytest.mark.parametrize("act,keep", [(0, 1.0), (1, 0.5), (2, 0.5), (1, 1.0)])
def test_bn_fwd_bwd_colsum(ops, act, keep, C):
    M = 1234
    x = R(M, C) * 2 + 0.5
    gamma, beta = R(C) * 0.1 + 1, R(C, seed=9) * 0.1
    ag, ac = both(ops, "bn_fwd", [x, gamma, beta, act, keep, 77, 105, torch.zeros(M, C), torch.zeros(C), torch.zeros(C)], atol=2e-5)
    y, mean, rstd = ac[7], ac[8], ac[9]
    both(ops, "bn_bwd", [R(M, C, seed=4), x, y, mean, rstd, gamma, act, keep, 77, 105, torch.zeros(M, C), torch.zeros(C), torch.zeros(C)],
         atol=2e-4)
    both(ops, "bn_update_moving", [mean, rstd, R(C), R(C).abs(), 0.99])
    both(ops, "colsum", [x, torch.ones(C), 0.0], atol=1e-3)
    both(ops, "colsum", [x, torch.ones(C), 1.0], atol=1e-3)


@pytest.mark.parametrize("train,use_lens", [(True, False), (True, True), (False, True)])
def test_lstm_fwd_bwd_gather_scatter(ops, train, use_lens):
    B, H, T = 5, 256, 6
    lens = torch.tensor([6, 2, 4, 1, 3], dtype=torch.int32) if use_lens else None
    s = 2
    gates, pre, bias = R(B, 4 * H), R(B, 4 * H, seed=1), R(4 * H, seed=2)
    c, h = R(B, H, seed=3), R(B, H, seed=4)
    outs = [torch.zeros(B, H) for _ in range(4)]
    ag, ac = both(ops, "lstm_fwd", [gates, pre, bias, c, h, lens, s, 0.1, train, 5, 300, 301, 7 * B] + outs)
    acts, cnew = ac[0], ac[13]
    both(ops, "lstm_bwd", [acts, c, h, cnew, lens, s, 0.1, train, 5, 300, 301, 7 * B, R(B, H, seed=5), R(B, H, seed=6), R(B, H, seed=7),
                           torch.zeros(B, 4 * H), torch.zeros(B, H), torch.zeros(B, H)])
    if use_lens:
        src = R(B, T, 64)
        for rev in (False, True):
            both(ops, "gather_rows", [src, lens, s, rev, torch.ones(B, 64)])
            both(ops, "scatter_rows", [R(B, 64), lens, s, rev, torch.zeros(B, T, 64)])
        cu, tw = ops
        wide_g, wide_c = torch.zeros(B, T, 128).cuda(), torch.zeros(B, T, 128)
        v = R(B, 64)
        cu.scatter_rows(v.cuda(), lens.cuda(), s, True, wide_g[:, :, 64:])
        tw.scatter_rows(v, lens, s, True, wide_c[:, :, 64:])
        assert torch.equal(wide_g.cpu(), wide_c)
        og, oc = torch.zeros(B, 64).cuda(), torch.zeros(B, 64)
        cu.gather_rows(wide_g[:, :, 64:], lens.cuda(), s, True, og)
        tw.gather_rows(wide_c[:, :, 64:], lens, s, True, oc)
        assert torch.equal(og.cpu(), oc)


@pytest.mark.parametrize("T", [9, 111, 300])
def test_attention_step_fwd_bwd(ops, T):
    B, A, V = 4, 128, 768
    lens = torch.tensor([T, max(1, T // 2), max(1, T - 3), 1], dtype=torch.int32)
    q, keys, values = R(B, A), R(B, T, A), R(B, T, V)
    cum = torch.rand(B, T, generator=torch.Generator().manual_seed(1)) * (torch.arange(T)[None, :] < lens[:, None])
    cw, cb, Wl, va, ba = R(31, 1, 32, scale=0.3), R(32, scale=0.1), R(32, A, scale=0.2), R(A, scale=0.3), R(A, scale=0.1)
    ag, ac = both(ops, "attn_fwd", [q, cum, keys, values, lens, cw, cb, Wl, va, ba, torch.zeros(B, T), torch.zeros(B, T), torch.zeros(B, V)],
                  atol=2e-5)
    align = ac[10]
    parts = [torch.zeros(B, 31, 1, 32), torch.zeros(B, 32), torch.zeros(B, 32, A), torch.zeros(B, A), torch.zeros(B, A)]
    args = [R(B, V, seed=2), R(B, T, seed=3), q, cum, align, keys, values, lens, cw, cb, Wl, va, ba,
            torch.zeros(B, A), torch.zeros(B, T), R(B, T, A, seed=5) * 0.01] + parts
    cu, tw = ops
    a_gpu = [a.clone().cuda() if isinstance(a, torch.Tensor) else a for a in args]
    a_cpu = [a.clone() if isinstance(a, torch.Tensor) else a for a in args]
    for _ in range(2):                     # twice: dkeys / partials accumulate
        cu.attn_bwd(*a_gpu)
        tw.attn_bwd(*a_cpu)
    for i in (13, 14, 15):
        np.testing.assert_allclose(a_gpu[i].cpu().numpy(), a_cpu[i].numpy(), atol=3e-5, rtol=1e-4, err_msg="attn_bwd arg %d" % i)
    for i in range(16, 21):                # the twin puts the batch total in row 0: compare sums over rows
        np.testing.assert_allclose(a_gpu[i].sum(0).cpu().numpy(), a_cpu[i].sum(0).numpy(), atol=2e-4, rtol=1e-4, err_msg="attn_bwd partial %d" % i)


@pytest.mark.parametrize("T,D", [(9, 256), (111, 256), (230, 64)])
def test_fused_attention_of_the_time_loops_matches_the_op_by_op_kernels(ops, T, D):
    """train_attn_fused.cu (one launch forward, two backward per decoder step, weight / key gradients hoisted out of the BPTT loop)
    against the op-by-op kernels checked above, over a 3-step chain: the carried cumulative alignments forward, the carried
    d(cum) backward; the hoisted gradients against the per-step accumulation.  Same fp32 arithmetic, different summation
    order (folded 31-tap filter bank): 2e-5 absolute on O(1) values, 1e-4 relative on the accumulated gradients."""
    cu, _ = ops
    S, B, A, V = 3, 4, 128, 768
    lens = torch.tensor([T, max(1, T // 2), max(1, T - 3), 1], dtype=torch.int32).cuda()
    keys, values = R(B, T, A).cuda(), R(B, T, V).cuda()
    cw, cb, Wl, va, ba = (x.cuda() for x in (R(31, 1, 32, scale=0.3), R(32, scale=0.1), R(32, A, scale=0.2), R(A, scale=0.3), R(A, scale=0.1)))
    Wq, h2 = R(D, A, scale=0.1).cuda(), R(S, B, D, seed=4).cuda()
    z = lambda *sh: torch.zeros(*sh, device="cuda")
    # forward, op by op
    q0, al0, cum0, ctx0 = z(S, B, A), z(S, B, T), z(S + 1, B, T), z(S, B, V)
    for s in range(S):
        q0[s] = h2[s] @ Wq
        cu.attn_fwd(q0[s], cum0[s], keys, values, lens, cw, cb, Wl, va, ba, al0[s], cum0[s + 1], ctx0[s])
    # forward, fused
    U, c0 = z(31, A), z(A)
    cu.attn_fold(cw, cb, Wl, U, c0)
    q1, al1, cum1, ctx1, th = z(S, B, A), z(S, B, T), z(S + 1, B, T), z(S, B, V), torch.full((S, B, T, A), float("nan"), device="cuda")
    for s in range(S):
        cu.attn_fwd_fused(h2[s], Wq, U, c0, ba, va, cum1[s], keys, values, lens, q1[s], th[s], al1[s], cum1[s + 1], ctx1[s])
    for name, a, b in (("q", q1, q0), ("align", al1, al0), ("cum", cum1, cum0), ("ctx", ctx1, ctx0)):
        np.testing.assert_allclose(a.cpu().numpy(), b.cpu().numpy(), atol=2e-5, rtol=1e-5, err_msg=name)
    assert float(al1[:, 1, max(1, T // 2):].abs().sum()) == 0.0          # exact zeros beyond the length
    # backward chain, op by op
    dctx_in, carry = R(S, B, V, seed=2).cuda(), R(S, B, V, seed=6).cuda() * 0.5
    parts = [z(B, 31, 1, 32), z(B, 32), z(B, 32, A), z(B, A), z(B, A)]
    dkeys0, dq0, dh20 = z(B, T, A), z(S, B, A), R(S, B, D, seed=8).cuda()
    dh2_init = dh20.clone()
    tot0 = dctx_in + carry
    dcum, dcum_n = R(B, T, seed=3).cuda(), z(B, T)
    dcum_init = dcum.clone()
    for s in range(S - 1, -1, -1):
        cu.attn_bwd(tot0[s], dcum, q0[s], cum0[s], al0[s], keys, values, lens, cw, cb, Wl, va, ba, dq0[s], dcum_n, dkeys0, *parts)
        dh20[s] += dq0[s] @ Wq.t()
        dcum, dcum_n = dcum_n, dcum
    dcum_end0 = dcum.clone()
    # backward chain, fused + hoisted
    tot1, de, dq1, dh21 = dctx_in.clone(), z(S, B, T), z(S, B, A), dh2_init.clone()
    dcum, dcum_n = dcum_init.clone(), z(B, T)
    for s in range(S - 1, -1, -1):
        cu.attn_bwd_fused(tot1[s], carry[s], dcum, al1[s], th[s], values, lens, U, va, Wq, de[s], dq1[s], dh21[s], dcum_n)
        dcum, dcum_n = dcum_n, dcum
    g = [z(31, 1, 32), z(32), z(32, A), z(A), z(A)]
    dkeys1 = z(B, T, A)
    cu.attn_bwd_hoisted(de, th, cum1[:S].contiguous(), lens, cw, cb, Wl, va, dkeys1, *g)
    torch.cuda.synchronize()
    for name, a, b in (("total dctx", tot1, tot0), ("dq", dq1, dq0), ("dh2", dh21, dh20), ("dcum", dcum, dcum_end0), ("dkeys", dkeys1, dkeys0)):
        np.testing.assert_allclose(a.cpu().numpy(), b.cpu().numpy(), atol=3e-5, rtol=1e-4, err_msg=name)
    for name, a, b in zip(("d conv_w", "d conv_b", "d Wl", "d va", "d ba"), g, parts):
        ref = b.sum(0).cpu().numpy()
        np.testing.assert_allclose(a.cpu().numpy(), ref, atol=1e-4 * max(1.0, float(np.abs(ref).max())), rtol=1e-4, err_msg=name)


@pytest.mark.parametrize("S,B,T,V", [(7, 3, 9, 64), (50, 4, 111, 768), (33, 2, 70, 128)])
def test_attn_dvalues_matches_einsum(ops, S, B, T, V):
    """d(values)[b,t,c] = sum_s align[s,b,t] d(ctx)[s,b,c] in one launch vs the fp64 contraction (fp32 accumulation over S)."""
    cu, _ = ops
    al, dc = torch.rand(S, B, T, generator=torch.Generator().manual_seed(3)), R(S, B, V, seed=5)
    out = torch.full((B, T, V), float("nan")).cuda()
    cu.attn_dvalues(al.cuda(), dc.cuda(), out)
    ref = torch.einsum("sbt,sbc->btc", al.double(), dc.double()).numpy()
    np.testing.assert_allclose(out.cpu().numpy(), ref, atol=2e-5 * max(1.0, float(np.abs(ref).max())), rtol=1e-5)


def test_gru_gst_pointwise(ops):
    B, Rr = 6, 128
    g, h = R(B, 2 * Rr), R(B, Rr, seed=1)
    ag, ac = both(ops, "gru_pre", [g, h, torch.zeros(B, Rr)])
    ru = ac[0]
    ag, ac = both(ops, "gru_out", [R(B, Rr, seed=2), ru, h, torch.zeros(B, Rr)])
    c = ac[0]
    ag, ac = both(ops, "gru_out_bwd", [R(B, Rr, seed=3), ru, h, c, torch.zeros(B, Rr), torch.zeros(B, 2 * Rr), torch.zeros(B, Rr)])
    both(ops, "gru_pre_bwd", [R(B, Rr, seed=4), ru, h, ac[5], ac[6]])
    N, U, D, nh = 10, 128, 64, 4
    q, k, v, gg, b, tt = R(B, U), R(N, U), R(32), torch.tensor([0.3]), R(32, scale=0.1), torch.tanh(R(N, D))
    ag, ac = both(ops, "gst_fwd", [q, k, v, gg, b, tt, nh, torch.zeros(B, nh, N), torch.zeros(B, nh * D)], atol=1e-5)
    both(ops, "gst_bwd", [R(B, nh * D, seed=8), q, k, v, gg, b, tt, ac[7], nh, torch.zeros(B, U), torch.zeros(N, U), torch.zeros(32),
                          torch.zeros(1), torch.zeros(32), torch.zeros(N, D)], atol=2e-5)


def test_elementwise_losses_misc(ops):
    a, b = R(1000), R(1000, seed=1)
    for op in ("copy", "add", "mul", "scale", "tanh", "tanh_bwd"):
        both(ops, "ew", [op, torch.zeros(1000), a, torch.tanh(b), 0.7])
    both(ops, "clip_fwd", [a * 3, -4.1, 4.0, torch.zeros(1000)])
    both(ops, "clip_bwd", [b, a * 3, -4.1, 4.0, torch.zeros(1000)])
    both(ops, "mse", [R(7, 50, 80), R(7, 50, 80, seed=2), torch.zeros(1), torch.zeros(7, 50, 80)], atol=1e-6, rtol=1e-4)
    both(ops, "sigmoid_ce", [R(7, 50) * 4, (R(7, 50, seed=3) > 0).float(), torch.zeros(1), torch.zeros(7, 50)], atol=1e-6, rtol=1e-4)
    ids = torch.randint(0, 66, (3, 17), dtype=torch.int32)
    table = R(66, 512)
    both(ops, "embedding_fwd", [ids, table, torch.zeros(3, 17, 512)])
    both(ops, "embedding_bwd", [ids, R(3, 17, 512), torch.ones(66, 512)], atol=1e-4)
    z = R(5, 3, 256)
    ag, ac = both(ops, "prenet_act_fwd", [z, 1, 45454, 0, 0.5, torch.zeros(5, 3, 256)])
    both(ops, "prenet_act_bwd", [R(5, 3, 256, seed=2), ac[5], 1, 45454, 0, 0.5, torch.zeros(5, 3, 256)])
    lens = torch.tensor([4, 9, 1], dtype=torch.int32)
    both(ops, "mask_rows", [R(3, 9, 40), lens])
    both(ops, "tile_style", [R(3, 256), torch.zeros(3, 9, 768), 512])
    both(ops, "sum_style", [R(3, 9, 768), 512, torch.zeros(3, 256)], atol=1e-4)
    n = 100003
    p, g, regw = R(n), R(n, seed=1) * 0.01, (R(n, seed=2) > 0).float() * 1e-3
    both(ops, "reg_loss", [p, regw, torch.zeros(1)], rtol=1e-3)
    ag, ac = both(ops, "grad_finalize", [g, p, regw, 0.5, torch.zeros(1)], rtol=1e-3)
    both(ops, "adam_step", [p, ac[0], R(n, seed=3) * 0.01, R(n, seed=4).abs() * 1e-4, ac[4], 1.0, 1e-3, 0.9, 0.999, 1e-6], atol=1e-5)


# ------------------------------------------------------------------ the whole step
def _problem(hp, lengths, fpt, seed):
    ids, lens = synthetic.token_batch(hp, list(lengths))
    mel, stop, _ = synthetic.mel_targets(hp, lens, frames_per_token=fpt)
    return ids, lens, mel, stop, seed


@pytest.mark.parametrize("lengths,fpt", [((7, 5, 6), 2), ((18, 31, 36, 41), 4)])
def test_training_step_matches_oracle_autograd(ops, lengths, fpt):
    cu, _ = ops
    hp = default_hparams()
    w = weights.init_weights(hp, perturb=0.05)
    ids, lens, mel, stop, seed = _problem(hp, lengths, fpt, 3)
    ref, orc = train_oracle.train_forward(hp, w, ids, lens, mel, stop, seed)
    ref["loss"].backward()
    ref_grads = train_oracle.gradients(orc)
    step = TrainStep(hp, cu, w)
    out = step.forward(cu.to_device(ids), cu.to_device(lens), cu.to_device(mel), cu.to_device(stop), seed)
    step.backward()
    t = lambda x: x.detach().cpu().numpy()
    np.testing.assert_allclose(t(out["encoder_outputs"]), t(ref["encoder_outputs"]), atol=5e-5)
    np.testing.assert_allclose(t(out["decoder_output"]), t(ref["decoder_output"]), atol=2e-4)
    np.testing.assert_allclose(t(out["mel_outputs"]), t(ref["mel_outputs"]), atol=5e-4)
    np.testing.assert_allclose(t(out["stop_token_prediction"]), t(ref["stop_token_prediction"]), atol=2e-4)
    np.testing.assert_allclose(t(out["alignments"].permute(1, 2, 0)), t(ref["alignments"]), atol=2e-5)
    ls = t(out["losses"])
    for i, k in enumerate(("before_loss", "after_loss", "stop_token_loss", "regularization_loss")):
        np.testing.assert_allclose(ls[i], float(ref[k].detach()), rtol=1e-4)
    worst = ("", 0.0)
    for name, g_ref in ref_grads.items():
        g = t(step.store.G[name]).reshape(g_ref.shape).astype(np.float64)
        if is_regularized(name):
            g = g + train_oracle.reg_weight(hp) * np.asarray(w[name], np.float64)
        scale = max(np.abs(g_ref).max(), 1e-3)
        err = np.abs(g - g_ref).max() / scale
        worst = max(worst, (name, err), key=lambda z: z[1])
        assert err < 5e-3, "%s: rel err %.3e (|g|max %.3e)" % (name, err, scale)
    print("worst gradient:", worst, "launches:", cu.launches)
    # optimizer
    step.apply(global_step=0)
    wn = {n: np.asarray(v, np.float64).copy() for n, v in w.items() if n in ref_grads}
    m = {n: np.zeros_like(v) for n, v in wn.items()}
    v = {n: np.zeros_like(x) for n, x in wn.items()}
    train_oracle.adam_update(hp, wn, [ref_grads], m, v, 0)
    new = step.store.export()
    for n in wn:
        np.testing.assert_allclose(new[n], wn[n], atol=5e-5, err_msg=n)


# ------------------------------------------------------------------ bf16 tensor-core mode of the large contractions
def _bf(x):
    return x.to(torch.bfloat16).to(torch.float32)


def test_bf16_mma_contractions_match_bf16_rounded_operands(ops):
    """TACO_TRAIN_BF16: the result equals an fp32 contraction of the bf16-ROUNDED operands up to accumulation order."""
    cu, tw = ops
    cu.set_precision("bf16")
    try:
        for (M, N, K, ta, tb, beta, bias) in [(600, 520, 600, 0, 0, 0.0, 1), (1280, 1024, 4099, 1, 0, 0.0, 0), (2000, 1100, 512, 0, 1, 1.0, 1),
                                              (256, 1024, 16000, 1, 0, 0.0, 0)]:
            A = R(K, M) if ta else R(M, K)
            Bm = R(N, K) if tb else R(K, N)
            C0, b = R(M, N, seed=3), (R(N) if bias else None)
            Cg = C0.clone().cuda()
            cu.gemm(A.cuda(), Bm.cuda(), Cg, bool(ta), bool(tb), beta, None if b is None else b.cuda())
            Cc = C0.clone()
            tw.gemm(_bf(A), _bf(Bm), Cc, bool(ta), bool(tb), beta, b)
            np.testing.assert_allclose(Cg.cpu().numpy(), Cc.numpy(), atol=2e-4 * np.sqrt(K), rtol=1e-4)
        for (B, T, Ci, Co, k) in [(3, 130, 80, 768, 5), (4, 200, 768, 768, 5), (20, 200, 768, 80, 5), (4, 111, 512, 512, 5)]:
            x, w, dy = R(B, T, Ci), R(k, Ci, Co, scale=0.1), R(B, T, Co)
            for name, args, mnk in (("conv1d_fwd", [x, w, R(Co), torch.zeros(B, T, Co)], (B * T, Co, k * Ci)),
                                    ("conv1d_dgrad", [dy, w, torch.zeros(B, T, Ci)], (B * T, Ci, k * Co)),
                                    ("conv1d_wgrad", [x, dy, torch.zeros(k, Ci, Co)], (k * Ci, Co, B * T))):
                # the library routes a contraction to the tensor cores only when it is throughput-bound (train_gemm.cu: use_mma)
                routed = mnk[0] * mnk[1] >= 64 ** 3 and mnk[0] * mnk[1] * mnk[2] >= 2 ** 24
                ag = [a.clone().cuda() for a in args]
                ac = [(_bf(a) if (routed and i in (0, 1)) else a.clone()) for i, a in enumerate(args)]
                getattr(cu, name)(*ag)
                getattr(tw, name)(*ac)
                np.testing.assert_allclose(ag[-1].cpu().numpy(), ac[-1].numpy(), atol=3e-3, rtol=1e-3, err_msg=name)
    finally:
        cu.set_precision("fp32")


def test_training_step_bf16_mode_close_to_fp32_oracle(ops):
    """bf16 operands in the large contractions: losses within 2 %, every large gradient tensor within 6 % (relative L2; worst are the reference-encoder kernels behind the GRU + GST chain) of
    autograd through the fp32 oracle; stated tolerance of the mixed-precision mode."""
    cu, _ = ops
    hp = default_hparams()
    w = weights.init_weights(hp, perturb=0.05)
    ids, lens, mel, stop, seed = _problem(hp, (18, 31, 36, 41), 4, 3)
    ref, orc = train_oracle.train_forward(hp, w, ids, lens, mel, stop, seed)
    ref["loss"].backward()
    ref_grads = train_oracle.gradients(orc)
    cu.set_precision("bf16")
    try:
        step = TrainStep(hp, cu, w)
        out = step.forward(cu.to_device(ids), cu.to_device(lens), cu.to_device(mel), cu.to_device(stop), seed)
        step.backward()
        torch.cuda.synchronize()
    finally:
        cu.set_precision("fp32")
    ls = out["losses"].cpu().numpy()
    for i, k in enumerate(("before_loss", "after_loss", "stop_token_loss")):
        np.testing.assert_allclose(ls[i], float(ref[k].detach()), rtol=2e-2)
    worst = ("", 0.0)
    for name, g_ref in ref_grads.items():
        if g_ref.size < 1024 or np.linalg.norm(g_ref) < 1e-4:
            continue
        g = step.store.G[name].cpu().numpy().reshape(g_ref.shape).astype(np.float64)
        if is_regularized(name):
            g = g + train_oracle.reg_weight(hp) * np.asarray(w[name], np.float64)
        err = np.linalg.norm(g - g_ref) / np.linalg.norm(g_ref)
        worst = max(worst, (name, err), key=lambda z: z[1])
        assert err < 6e-2, "%s: relative L2 error %.3e" % (name, err)
    print("bf16 mode worst gradient (relative L2):", worst)


@pytest.mark.parametrize("Ci,Co,T", [(768, 768, 200), (80, 768, 199), (512, 512, 130)])
def test_conv1d_tcgen05_training_paths_match_bf16_reference(Ci, Co, T):
    """bf16 training mode: conv1d forward, data gradient and weight gradient run on the tcgen05 kernel of the synthesis path
    (conv_tc.cu; operands packed / transposed to bf16 per call) when B*T >= 4096 rows.  Reference: float64 convolutions of the
    same bf16-rounded operands; tolerance = fp32 accumulation order over K <= 6400 (relative 2e-4 of the output scale)."""
    from tacotron_prosody_multi_gpu_b200.train.ops import CudaOps
    o = CudaOps(0)
    o.set_precision("bf16")
    try:
        B, k = 32, 5
        g = torch.Generator(device="cuda").manual_seed(Ci + T)
        x = torch.randn(B, T, Ci, device="cuda", generator=g)
        w = torch.randn(k, Ci, Co, device="cuda", generator=g) * 0.05
        b = torch.randn(Co, device="cuda", generator=g)
        dy = torch.randn(B, T, Co, device="cuda", generator=g) * 0.1
        # views at an odd float offset, like the parameter / gradient views of the flat training buffer
        flat = torch.zeros(k * Ci * Co + Co + 3, device="cuda")
        wv = flat[1:1 + k * Ci * Co].view(k, Ci, Co)
        wv.copy_(w)
        bv = flat[1 + k * Ci * Co:1 + k * Ci * Co + Co]
        bv.copy_(b)
        gflat = torch.zeros(k * Ci * Co + 3, device="cuda")
        dw = gflat[3:3 + k * Ci * Co].view(k, Ci, Co)
        y, dx = o.empty(B, T, Co), o.empty(B, T, Ci)
        o.conv1d_fwd(x, wv, bv, y)
        o.conv1d_dgrad(dy, wv, dx)
        o.conv1d_wgrad(x, dy, dw)
        xr, wr, dyr = x.bfloat16().double(), w.bfloat16().double(), dy.bfloat16().double()
        ref_y = torch.nn.functional.conv1d(xr.transpose(1, 2), wr.permute(2, 1, 0), b.double(), padding=2).transpose(1, 2)
        ref_dx = torch.nn.functional.conv_transpose1d(dyr.transpose(1, 2), wr.permute(2, 1, 0), padding=2).transpose(1, 2)
        xp = torch.nn.functional.pad(xr, (0, 0, 2, 2))
        ref_dw = torch.stack([torch.einsum("btc,bto->co", xp[:, tap:tap + T], dyr) for tap in range(k)])
        for name, got, ref in (("fwd", y, ref_y), ("dgrad", dx, ref_dx), ("wgrad", dw, ref_dw)):
            err = float((got.double() - ref).abs().max())
            scale = float(ref.abs().max())
            print("%s Ci=%d Co=%d: max err %.2e of %.2e" % (name, Ci, Co, err, scale))
            assert err <= 2e-4 * scale, (name, err, scale)
    finally:
        o.set_precision("fp32")
