"""The N>1 path on CPU: two processes (gloo) each take their share of the sessions exactly as
bench.py's ranks do, decode it, and rank 0 checks that the shards put together equal the
single-process result.  The decode itself is done by the CPU oracle here (no GPU in this
suite); what is under test is the partitioning, the per-rank xid tables and the merge."""
import importlib
import os
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _worker(rank, world, port, q):
    for p in (ROOT, os.path.join(ROOT, "oracle"), os.path.join(ROOT, "tests")):
        if p not in sys.path:
            sys.path.insert(0, p)
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    import zkoracle
    synth = importlib.import_module("node-zkstream_b200.synth")
    shard = importlib.import_module("node-zkstream_b200.shard")
    c = synth.cfg2_mixed(6, 120, seed=31)           # every rank derives the same corpus from the seed
    d, off, fl, hs = c.arrays()
    xt = c.xidtable()
    s0, s1 = shard.partition_sessions(off, world)[rank]
    sd, so, sf, sh, sx = shard.shard_inputs(d, off, fl, hs, xt, s0, s1)
    cols = zkoracle.decode(sd, so, sf, sh, sx)
    # what bench.py reduces: frames per rank (SUM) and time (MAX)
    t = torch.tensor([cols.n_frames, int(off[s1]) - int(off[s0])], dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.SUM)
    gathered = [None] * world
    dist.all_gather_object(gathered, (s0, s1, cols.n_frames, cols.xid.tolist(), cols.meta.tolist()))
    if rank == 0:
        full = zkoracle.decode(d, off, fl, hs, xt)
        ok = int(t[0].item()) == full.n_frames and int(t[1].item()) == len(c.data)
        xids = sum((g[3] for g in gathered), [])
        metas = sum((g[4] for g in gathered), [])
        ok = ok and xids == full.xid.tolist() and metas == full.meta.tolist()
        ok = ok and gathered[0][0] == 0 and gathered[-1][1] == len(off) - 1
        q.put(ok)
    dist.destroy_process_group()


def test_two_rank_sharded_decode():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29600 + os.getpid() % 300
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(120)
        assert p.exitcode == 0
    assert q.get(timeout=10) is True
