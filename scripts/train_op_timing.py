#!/usr/bin/env python
"""Warm-cache timing of the per-step primitives of the training recurrence at the C4 shape (B=32, T=111, D=256, MD=768):
each primitive is captured N times back to back into a CUDA graph and the replay is timed with CUDA events (no profiler: ncu launch lists are
cold-cache and serialised).  Prints microseconds per launch, i.e. what a decoder step pays for each of them."""
import ctypes as C
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from tacotron_prosody_multi_gpu_b200 import synthetic  # noqa: E402
from tacotron_prosody_multi_gpu_b200.train.ops import CudaOps  # noqa: E402

N = int(sys.argv[1]) if len(sys.argv) > 1 else 300
B, T, D, MD, A, U2 = 32, 111, 256, 768, 128, 256
G = 4 * D
cu = CudaOps(0)
g = torch.Generator(device="cuda").manual_seed(1)
R = lambda *s: torch.randn(*s, device="cuda", generator=g) * 0.1
Z = lambda *s: torch.zeros(*s, device="cuda")
lens = torch.tensor([min(T, n) for n in synthetic.cycle_lengths(B)], dtype=torch.int32).cuda()
keys, values = R(B, T, A), R(B, T, MD)
cw, cb, Wl, va, ba, Wq = R(31, 1, 32), R(32), R(32, A), R(A), R(A), R(D, A)
W1, W2 = R(U2 + MD + D, G), R(2 * D, G)
ctx, h1, h2 = R(B, MD), R(B, D), R(B, D)
acts, pre, c_prev = Z(B, G), R(B, G), R(B, D)
cum = torch.rand(B, T, device="cuda") * (torch.arange(T, device="cuda")[None] < lens[:, None])
U, c0 = Z(31, A), Z(A)
cu.attn_fold(cw, cb, Wl, U, c0)
q, al, cum2, cx, th = Z(B, A), Z(B, T), Z(B, T), Z(B, MD), Z(B, T, A)
de, dq, dh2, dcn, dkeys = Z(B, T), Z(B, A), Z(B, D), Z(B, T), Z(B, T, A)
parts = [Z(B, 31, 1, 32), Z(B, 32), Z(B, 32, A), Z(B, A), Z(B, A)]
dctx, carry, dcum = R(B, MD), R(B, MD), R(B, T)
outs = [Z(B, D) for _ in range(8)]
dG = R(B, G)
P = lambda t: C.c_void_p(t.data_ptr())
L = cu.L
for n_ in ("taco_op_gemm_pair_acc", "taco_op_gemm_nt_split_acc"):
    getattr(L, n_).restype = C.c_int


def pair():
    L.taco_op_gemm_pair_acc(P(ctx), C.c_longlong(MD), MD, P(h1), C.c_longlong(D), D, P(W1[U2:]), C.c_longlong(G), P(acts), C.c_longlong(G),
                            B, G, C.c_void_p(cu._st()))


def nt_split():
    L.taco_op_gemm_nt_split_acc(P(dG), C.c_longlong(G), P(W1[U2:]), C.c_longlong(G), P(dctx), C.c_longlong(MD), MD, P(outs[0]),
                                C.c_longlong(D), D, B, G, C.c_void_p(cu._st()))


cases = [
    ("gemm_pair_acc  [ctx|h1] . W1 (K=1024, N=1024)", pair),
    ("gemm_nt_split_acc dG . W1^T (K=1024 -> 768|256)", nt_split),
    ("gemm q = h2 . Wq", lambda: cu.gemm(h2, Wq, q)),
    ("lstm_fwd", lambda: cu.lstm_fwd(acts, pre, None, c_prev, h1, None, 0, 0.1, True, 7, 1, 2, 0, outs[0], outs[1], outs[2], outs[3])),
    ("lstm_bwd", lambda: cu.lstm_bwd(acts, c_prev, h1, outs[0], None, 0, 0.1, True, 7, 1, 2, 0, outs[2], outs[1], outs[3], dG, outs[4], outs[5])),
    ("attn_fwd (3 kernels)", lambda: cu.attn_fwd(q, cum, keys, values, lens, cw, cb, Wl, va, ba, al, cum2, cx)),
    ("attn_fwd_fused (1 kernel, with q)", lambda: cu.attn_fwd_fused(h2, Wq, U, c0, ba, va, cum, keys, values, lens, q, th, al, cum2, cx)),
    ("attn_bwd (4 kernels)", lambda: cu.attn_bwd(dctx, dcum, q, cum, al, keys, values, lens, cw, cb, Wl, va, ba, dq, dcn, dkeys, *parts)),
    ("attn_bwd_fused (2 kernels, with dh2)", lambda: cu.attn_bwd_fused(dctx, carry, dcum, al, th, values, lens, U, va, Wq, de, dq, dh2, dcn)),
    ("ew add", lambda: cu.ew("add", dctx, dctx, carry)),
]
ONLY = os.environ.get("TRAIN_OP_ONLY")           # substring of a case name: run just that case, plain launches (for ncu)
for name, fn in cases:
    if ONLY and ONLY not in name:
        continue
    for _ in range(5):
        fn()
    torch.cuda.synchronize()
    if ONLY:
        for _ in range(N):
            fn()
        torch.cuda.synchronize()
        continue
    graph = torch.cuda.CUDAGraph()          # N dependent launches replayed from a graph: device time, no host launch cost
    with torch.cuda.graph(graph):
        for _ in range(N):
            fn()
    graph.replay()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    graph.replay()
    e1.record()
    torch.cuda.synchronize()
    print("%-52s %7.2f us" % (name, 1e3 * e0.elapsed_time(e1) / N))
