"""Multi-GPU forest merge (seqrush_b200/distributed.py) exercised with world_size=2 over gloo on
CPU: each rank unites its shard of the regions into a private forest, the compressed root arrays
are all-gathered and merged, and every rank must end with the single-process partition."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp


class _HostUF:
    """Stand-in for the device union-find with the same link-to-smaller rule (host logic test only)."""

    def __init__(self, n):
        self.p = np.arange(n, dtype=np.int64)

    def find(self, x):
        while self.p[x] != x:
            self.p[x] = self.p[self.p[x]]
            x = self.p[x]
        return x

    def unite(self, a, b):
        ra, rb = self.find(a), self.find(b)
        if ra != rb:
            self.p[max(ra, rb)] = min(ra, rb)

    def compress(self):
        for i in range(self.p.size):
            self.p[i] = self.find(i)


def _edges(n, m, seed):
    rng = np.random.default_rng(seed)
    return rng.integers(0, n, size=(m, 2))


def _worker(rank, world, port, n, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from seqrush_b200.distributed import merge_forests, shard_pairs
    edges = _edges(n, 400, 3)
    uf = _HostUF(n)
    for a, b in shard_pairs(edges, rank, world):
        uf.unite(int(a), int(b))
    uf.compress()
    t = torch.from_numpy(uf.p.copy())

    def unite_with_peer(peer):
        pr = peer.numpy()
        for i in np.flatnonzero(pr != np.arange(n)):
            uf.unite(int(i), int(pr[i]))

    merge_forests(t, world, dist, unite_with_peer)
    uf.compress()
    q.put((rank, uf.p.tolist()))
    dist.barrier()
    dist.destroy_process_group()


def test_merge_forests_world2():
    n, world = 300, 2
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, world, port, n, q)) for r in range(world)]
    for p in procs:
        p.start()
    got = dict(q.get(timeout=120) for _ in range(world))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    ref = _HostUF(n)
    for a, b in _edges(n, 400, 3):
        ref.unite(int(a), int(b))
    ref.compress()
    assert got[0] == got[1] == ref.p.tolist()
