"""CPU, world_size 2 over gloo: the multi-GPU path shards pairs with no data-path
collective; only the timing reduction (max) and the counts (sum) cross ranks."""
import os
import socket
import sys

import numpy as np
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _worker(rank, world, port, q):
    sys.path.insert(0, ROOT)
    import torch
    from mgl_b200 import shard, synth
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    n = 200
    first, cnt = shard.weak_shard(n, rank)
    d = synth.generate(2, cnt, first=first)
    t = torch.tensor([float(rank + 1)], dtype=torch.float64)
    c = torch.tensor([float(d["cells"]), float(cnt)], dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    dist.all_reduce(c, op=dist.ReduceOp.SUM)
    q.put((rank, first, cnt, synth.pair_bytes(d, 0), t.item(), c.tolist()))
    dist.destroy_process_group()


def test_weak_shards_are_disjoint_and_reduce():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    got = sorted(q.get(timeout=120) for _ in range(2))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    sys.path.insert(0, ROOT)
    from mgl_b200 import synth
    whole = synth.generate(2, 400)
    assert got[0][1:3] == (0, 200) and got[1][1:3] == (200, 200)
    assert got[0][3] == synth.pair_bytes(whole, 0) and got[1][3] == synth.pair_bytes(whole, 200)
    assert got[0][4] == got[1][4] == 2.0                       # max over ranks
    assert got[0][5] == got[1][5] == [float(whole["cells"]), 400.0]


def test_strong_shard_balances_cells():
    from mgl_b200 import shard
    rng = np.random.default_rng(1)
    rl, al = rng.integers(300, 10000, 5000), rng.integers(150, 10000, 5000)
    cuts = shard.strong_shard_by_cells(rl, al, 8)
    assert cuts[0] == 0 and cuts[-1] == 5000 and (np.diff(cuts) > 0).all()
    cells = rl.astype(np.uint64) * al.astype(np.uint64)
    per = np.array([cells[cuts[k]:cuts[k + 1]].sum() for k in range(8)], dtype=np.float64)
    assert per.max() / per.mean() < 1.05


def _ph_worker(rank, world, port, q):
    sys.path.insert(0, ROOT)
    import random
    import torch
    from mgl_b200 import shard
    from oracle import pairhmm_oracle as P
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    rng = random.Random(11)                                          # every rank builds the same job
    haps = [bytes(rng.choice(b"ACGT") for _ in range(rng.randint(20, 90))) for _ in range(3)]
    reads = []
    for _ in range(21):
        n = rng.randint(1, 60)
        reads.append((bytes(rng.choice(b"ACGT") for _ in range(n)), bytes([30] * n), bytes([45] * n), bytes([45] * n), bytes([10] * n)))
    lengths, rb, hb = P.pack(reads, haps)
    first, sub_lengths, sub_reads = shard.pairhmm_shard_reads(lengths, rb, world, rank)
    n_mine = int(sub_lengths[0])
    mine, _ = P.likelihoods(reads[first:first + n_mine], haps)       # the rank's own job (the oracle stands in for the GPU here)
    assert bytes(sub_reads) == b"".join(b"".join(r) for r in reads[first:first + n_mine])
    whole, _ = P.likelihoods(reads, haps)
    cnt = torch.tensor([float(n_mine)], dtype=torch.float64)
    dist.all_reduce(cnt, op=dist.ReduceOp.SUM)                        # only counts and times cross ranks
    q.put((rank, first, n_mine, bool(np.array_equal(mine, whole[first:first + n_mine])), cnt.item()))
    dist.destroy_process_group()


def test_pairhmm_reads_shard_without_exchange():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_ph_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    got = sorted(q.get(timeout=120) for _ in range(2))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert got[0][1] == 0 and got[1][1] == got[0][2] and got[0][2] + got[1][2] == 21
    assert got[0][3] and got[1][3]
    assert got[0][4] == got[1][4] == 21.0
    assert 0.3 < got[0][2] / 21 < 0.7                                # cut by read bases, not by count
