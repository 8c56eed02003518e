"""world_size-2 CPU (gloo) test of the N>1 host path: every rank derives the same shard plan, the shards
partition the utterances, the max-over-ranks timing reduction of bench.py works, no data-path collective."""
import os
import socket
import subprocess
import sys
import textwrap

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

WORKER = textwrap.dedent("""
    import os, sys, json
    sys.path.insert(0, %r)
    import numpy as np, torch, torch.distributed as dist
    from tacotron_prosody_multi_gpu_b200 import sharding, synthetic
    from tacotron_prosody_multi_gpu_b200.hparams import default_hparams
    dist.init_process_group("gloo")
    rank, world = dist.get_rank(), dist.get_world_size()
    hp = default_hparams()
    lengths = synthetic.utterance_set(hp, 512).tolist()
    batches, owner = sharding.shard(lengths, world)
    mine = sorted(int(i) for b, idx in enumerate(batches) if owner[b] == rank for i in idx)
    # gather only the BOOKKEEPING (index lists) to check the partition; mel data never crosses ranks
    got = [None] * world
    dist.all_gather_object(got, mine)
    t = torch.tensor([float(rank + 1)], dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)          # bench.py's max-over-ranks time
    plan = torch.tensor(owner.tolist())
    ref = plan.clone()
    dist.broadcast(ref, 0)
    if rank == 0:
        print(json.dumps({"world": world, "covered": sorted(sum(got, [])) == list(range(512)),
                          "disjoint": len(set(sum(got, []))) == 512, "sizes": [len(g) for g in got], "max": t.item()}))
    assert torch.equal(plan, ref)
    dist.destroy_process_group()
""") % ROOT


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def test_two_rank_shard_plan_gloo(tmp_path):
    script = tmp_path / "worker.py"
    script.write_text(WORKER)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1",
           "--master-port", str(_free_port()), str(script)]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=240)
    assert r.returncode == 0, r.stderr[-2000:]
    import json
    line = [l for l in r.stdout.splitlines() if l.startswith("{")][-1]
    d = json.loads(line)
    assert d["world"] == 2 and d["covered"] and d["disjoint"] and d["max"] == 2.0
    assert sum(d["sizes"]) == 512 and min(d["sizes"]) >= 64   # cost-balanced (LPT), not count-balanced


TRAIN_WORKER = textwrap.dedent("""
    import os, sys, json
    sys.path.insert(0, %r); sys.path.insert(0, os.path.join(%r, "tests"))
    import numpy as np, torch, torch.distributed as dist
    from cpu_train_ops import TorchOps
    from oracle import train_oracle
    from tacotron_prosody_multi_gpu_b200 import synthetic, weights
    from tacotron_prosody_multi_gpu_b200.hparams import default_hparams
    from tacotron_prosody_multi_gpu_b200.train.trainer import DataParallelTrainer
    dist.init_process_group("gloo")
    rank, world = dist.get_rank(), dist.get_world_size()
    hp = default_hparams()
    w = weights.init_weights(hp, perturb=0.05)
    shards = [((6, 4), 11), ((5, 7), 12)]                     # (token lengths, dropout seed) per rank
    def data(r):
        ids, lens = synthetic.token_batch(hp, list(shards[r][0]), seed=100 + r)
        mel, stop, _ = synthetic.mel_targets(hp, lens, frames_per_token=2, seed=200 + r)
        return ids, lens, mel, stop, shards[r][1]
    ids, lens, mel, stop, seed = data(rank)
    tr = DataParallelTrainer(hp, w, ops=TorchOps())
    ls = tr.step(torch.from_numpy(ids), torch.from_numpy(lens), torch.from_numpy(mel), torch.from_numpy(stop), seed)
    new = tr.export()
    # every rank must hold identical weights after the update
    flat = torch.from_numpy(np.concatenate([new[k].reshape(-1) for k in sorted(new)]))
    ref = flat.clone(); dist.broadcast(ref, 0)
    same = bool(torch.equal(flat, ref))
    ok = True
    if rank == 0:   # the reference's semantics: mean of the tower gradients -> clip -> Adam (Tacotron.py:484-515)
        grads = []
        for r in range(world):
            out, orc = train_oracle.train_forward(hp, w, *data(r))
            out["loss"].backward()
            grads.append(train_oracle.gradients(orc))
        wn = {n: np.asarray(w[n], np.float64).copy() for n in grads[0]}
        m = {n: np.zeros_like(v) for n, v in wn.items()}; v = {n: np.zeros_like(x) for n, x in wn.items()}
        train_oracle.adam_update(hp, wn, grads, m, v, 0)
        err = max(float(np.abs(new[n] - wn[n]).max()) for n in wn)
        ok = err < 5e-5
        print(json.dumps({"world": world, "same": same, "max_err": err, "loss": float(ls[4])}))
    assert same and ok
    dist.destroy_process_group()
""") % (ROOT, ROOT)


def test_two_rank_data_parallel_training_step_gloo(tmp_path):
    """The N>1 training path on CPU: 2 ranks x 1 tower with the torch-CPU twin of the device primitives, one
    all-reduce of the flat gradient buffer; the result equals the reference's tower-averaged update."""
    script = tmp_path / "train_worker.py"
    script.write_text(TRAIN_WORKER)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1",
           "--master-port", str(_free_port()), str(script)]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, (r.stdout[-1500:], r.stderr[-2500:])
    import json
    d = json.loads([l for l in r.stdout.splitlines() if l.startswith("{")][-1])
    assert d["world"] == 2 and d["same"] and d["max_err"] < 5e-5


CLAIM_WORKER = textwrap.dedent("""
    import os, sys, json
    sys.path.insert(0, %r)
    import numpy as np, torch.distributed as dist
    from tacotron_prosody_multi_gpu_b200 import sharding
    from tacotron_prosody_multi_gpu_b200.hparams import default_hparams
    dist.init_process_group("gloo")
    rank, world = dist.get_rank(), dist.get_world_size()
    hp = default_hparams()
    rng = np.random.default_rng(5)
    seqs = [np.arange(2, 2 + n).astype(np.int32) for n in rng.integers(3, 40, size=70)]
    class FakeEngine:                       # stands in for TacoEngine.forward_host: one frame per token, no GPU
        def forward_host(self, ids, lens, **kw):
            B, T = ids.shape
            mel = np.repeat(ids[:, :, None].astype(np.float32), 80, axis=2) * 0.01
            stop = np.zeros((B, T), np.float32)
            for b, n in enumerate(lens):
                stop[b, n - 1:] = 1.0
            return dict(mel_outputs=mel, stop=stop, steps=T)
    store = dist.distributed_c10d._get_default_store()
    mine = sharding.synthesize_batches(FakeEngine(), hp, seqs, rank, world, batch_size=8, gst_weights=np.zeros((4, 10), np.float32),
                                       max_iters=50, philox_seed=1, store=store, store_key="t/claim")
    got = [None] * world
    dist.all_gather_object(got, {k: (v["mel"].shape[0], float(v["mel"].sum())) for k, v in mine.items()})
    if rank == 0:
        merged = {}
        for g in got:
            assert not (set(g) & set(merged)), "an utterance was synthesised twice"
            merged.update(g)
        ref = sharding.synthesize_batches(FakeEngine(), hp, seqs, 0, 1, batch_size=8, gst_weights=np.zeros((4, 10), np.float32), max_iters=50, philox_seed=1)
        ok = set(merged) == set(ref) and all(merged[k] == (ref[k]["mel"].shape[0], float(ref[k]["mel"].sum())) for k in ref)
        print(json.dumps({"ok": bool(ok), "n": len(merged), "per_rank": [len(g) for g in got]}))
        assert ok
    dist.destroy_process_group()
""") % (ROOT,)


def test_dynamic_batch_claiming_covers_every_utterance_once_gloo(tmp_path):
    """N > 1 synthesis with batches claimed from a shared store counter: every utterance exactly once, outputs equal to the
    single-rank run (same batch composition), no data-path collective."""
    script = tmp_path / "claim_worker.py"
    script.write_text(CLAIM_WORKER)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1",
           "--master-port", str(_free_port()), str(script)]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, (r.stdout[-1500:], r.stderr[-2500:])
    import json
    d = json.loads([l for l in r.stdout.splitlines() if l.startswith("{")][-1])
    assert d["ok"] and d["n"] == 70 and sum(d["per_rank"]) == 70
