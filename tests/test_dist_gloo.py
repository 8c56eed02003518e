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
