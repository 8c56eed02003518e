"""taco_op_attn_bwd_noalloc (include/taco_b200_train.h): the allocation-free variant of the attention-step gradient, scratch in a
static device array.  Called through the C ABI (ctypes on libtaco_b200.so) with caller-owned inputs and outputs and compared
with the existing entry point taco_op_attn_bwd (caller-provided scratch) on the same inputs, two calls in a row with different
shapes (the second call sees the first call's leftovers in the shared array).  Tolerances are those of
tests/test_train_gpu.py::test_attention_step_fwd_bwd: 3e-5 / 1e-4 on dq, d(cum), d(keys); 2e-4 / 1e-4 on the weight-gradient
partials (atomic accumulation order).  Collected last on purpose (file name)."""
import ctypes as C
import re
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

A, V, NF, KW = 128, 768, 32, 31
HEADER = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "include", "taco_b200_train.h")


def _limits():
    src = open(HEADER).read()
    m = re.search(r"#define\s+TACO_ATTN_BWD_NOALLOC_MAX_FLOATS\s+\(1\s*<<\s*(\d+)\)", src)
    e = re.search(r"#define\s+TACO_OP_ERR_SCRATCH_LIMIT\s+\((-?\d+)\)", src)
    return 1 << int(m.group(1)), int(e.group(1))


def _lib():
    from tacotron_prosody_multi_gpu_b200 import _lib
    L = _lib.lib()
    P, I = C.c_void_p, C.c_int
    L.taco_op_attn_bwd.argtypes = [P] * 22 + [I] * 6 + [P]
    L.taco_op_attn_bwd.restype = I
    L.taco_op_attn_bwd_noalloc.argtypes = [P] * 21 + [I] * 6 + [P]
    L.taco_op_attn_bwd_noalloc.restype = I
    L.taco_op_attn_fwd.argtypes = [P] * 13 + [I] * 6 + [P]
    L.taco_op_attn_fwd.restype = I
    return L


def _problem(B, T, seed):
    g = torch.Generator().manual_seed(seed)
    R = lambda *s, scale=1.0: (torch.randn(*s, generator=g) * scale).cuda()
    lens = torch.tensor([max(1, T - (7 * b) % T) for b in range(B)], dtype=torch.int32)
    lens[0] = T
    mask = (torch.arange(T)[None, :] < lens[:, None]).float()
    cum = (torch.rand(B, T, generator=g) * mask).cuda()
    ins = dict(dctx=R(B, V), dcum=R(B, T), q=R(B, A), cum=cum, keys=R(B, T, A), values=R(B, T, V), lens=lens.cuda(),
               cw=R(KW, 1, NF, scale=0.3), cb=R(NF, scale=0.1), Wl=R(NF, A, scale=0.2), va=R(A, scale=0.3), ba=R(A, scale=0.1))
    ins["dkeys0"] = R(B, T, A, scale=0.01)          # d(keys) accumulates: both entry points start from the same values
    return ins


def _outputs(B, T, ins):
    z = lambda *s: torch.zeros(*s, device="cuda")
    return dict(dq=z(B, A), dcum_prev=z(B, T), dkeys=ins["dkeys0"].clone(), p_cw=z(B, KW, 1, NF), p_cb=z(B, NF), p_Wl=z(B, NF, A),
                p_va=z(B, A), p_ba=z(B, A))


def _run(L, B, T, seed):
    ins = _problem(B, T, seed)
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    p = lambda t: C.c_void_p(t.data_ptr())
    align, cum_n, ctx = torch.zeros(B, T, device="cuda"), torch.zeros(B, T, device="cuda"), torch.zeros(B, V, device="cuda")
    rc = L.taco_op_attn_fwd(p(ins["q"]), p(ins["cum"]), p(ins["keys"]), p(ins["values"]), p(ins["lens"]), p(ins["cw"]), p(ins["cb"]),
                            p(ins["Wl"]), p(ins["va"]), p(ins["ba"]), p(align), p(cum_n), p(ctx), B, T, A, V, NF, KW, st)
    assert rc == 0
    common = lambda o: [p(ins["dctx"]), p(ins["dcum"]), p(ins["q"]), p(ins["cum"]), p(align), p(ins["keys"]), p(ins["values"]),
                        p(ins["lens"]), p(ins["cw"]), p(ins["cb"]), p(ins["Wl"]), p(ins["va"]), p(ins["ba"]), p(o["dq"]),
                        p(o["dcum_prev"]), p(o["dkeys"]), p(o["p_cw"]), p(o["p_cb"]), p(o["p_Wl"]), p(o["p_va"]), p(o["p_ba"])]
    ref, new = _outputs(B, T, ins), _outputs(B, T, ins)
    scratch = torch.full((B * T * NF,), float("nan"), device="cuda")      # allocated BEFORE the allocation-free call
    torch.cuda.synchronize()
    assert L.taco_op_attn_bwd(*common(ref), p(scratch), B, T, A, V, NF, KW, st) == 0
    mem0 = torch.cuda.memory_allocated()
    assert L.taco_op_attn_bwd_noalloc(*common(new), B, T, A, V, NF, KW, st) == 0
    torch.cuda.synchronize()
    assert torch.cuda.memory_allocated() == mem0
    for k in ("dq", "dcum_prev", "dkeys"):
        np.testing.assert_allclose(new[k].cpu().numpy(), ref[k].cpu().numpy(), atol=3e-5, rtol=1e-4, err_msg=k)
    for k in ("p_cw", "p_cb", "p_Wl", "p_va", "p_ba"):
        np.testing.assert_allclose(new[k].cpu().numpy(), ref[k].cpu().numpy(), atol=2e-4, rtol=1e-4, err_msg=k)
    assert float(ref["dq"].abs().max()) > 0 and float(ref["p_Wl"].abs().max()) > 0       # the comparison is not 0 == 0


def test_attn_bwd_noalloc_matches_attn_bwd_two_calls_in_a_row():
    L = _lib()
    _run(L, 8, 230, seed=1)         # larger problem first: its leftovers stay in the shared array ...
    _run(L, 5, 57, seed=2)          # ... under a smaller one with a different layout
    _run(L, 8, 230, seed=3)


def test_attn_bwd_noalloc_rejects_sizes_beyond_the_static_array():
    L = _lib()
    max_floats, err = _limits()
    B, T = 64, max_floats // (64 * NF) + 1          # B*T*NF just above the limit
    assert B * T * NF > max_floats
    one = torch.zeros(1, device="cuda")              # never dereferenced: the size check comes first
    p = C.c_void_p(one.data_ptr())
    rc = L.taco_op_attn_bwd_noalloc(*([p] * 21), B, T, A, V, NF, KW, C.c_void_p(torch.cuda.current_stream().cuda_stream))
    assert rc == err
    torch.cuda.synchronize()                         # nothing was launched, the context is healthy
    _run(L, 4, 40, seed=4)
