"""Multi-GPU sharding of the align-and-unite phase: one process per GPU (torch.distributed over
NCCL), the pair list dealt round-robin by unordered pair, a private union-find per GPU over the full Pos
space, and ONE exchange at the end -- all_gather of the compressed parent arrays followed by a
device-side union pass. The host logic is backend-agnostic so that the gloo
CPU tests can cover it with a stand-in union-find."""
import numpy as np


def shard_pairs(pairs: np.ndarray, rank: int, world: int) -> np.ndarray:
    """Deal of the ordered pair list over the ranks (pairs are independent: allwave/src/iterator.rs:185-203).
    The two orders (i, j) and (j, i) of an unordered pair go to the SAME rank, so that rank's launch can take the
    strand choice of one from the other (csrc/kernels.cuh, mirror_*_kernel); the unordered pairs are dealt round-robin
    in the order of their (min, max) key, and a rank keeps the list's own order among its pairs."""
    p = np.asarray(pairs)
    if world <= 1 or p.shape[0] == 0:
        return np.ascontiguousarray(p)
    a = p[:, 0].astype(np.uint64)
    b = p[:, 1].astype(np.uint64)
    key = (np.minimum(a, b) << np.uint64(32)) | np.maximum(a, b)
    order = np.argsort(key, kind="stable")
    ks = key[order]
    gid = np.cumsum(np.concatenate([[True], ks[1:] != ks[:-1]])) - 1
    # unordered pairs present in both orders (or repeated) are dealt round-robin among themselves, the single ones
    # after them, so that the ranks' PAIR counts stay within one unordered pair of each other
    size = np.bincount(gid)
    multi = size > 1
    owner = np.empty(size.shape[0], dtype=np.int64)
    n_multi = int(multi.sum())
    owner[multi] = np.arange(n_multi) % world
    owner[~multi] = (n_multi + np.arange(size.shape[0] - n_multi)) % world
    sel = np.sort(order[owner[gid] == rank])
    return np.ascontiguousarray(p[sel])


def merge_forests(local_roots, world: int, dist, unite_with_peer):
    """Backend-agnostic form of the exchange (the gloo CPU tests drive it with a host union-find).
    local_roots: tensor of this rank's compressed parent array (parent[p] == find(p)).
    Gathers every rank's array and calls unite_with_peer(peer_roots_tensor) for each other rank."""
    import torch

    gathered = [torch.empty_like(local_roots) for _ in range(world)]
    dist.all_gather(gathered, local_roots)
    me = dist.get_rank()
    for r in range(world):
        if r != me:
            unite_with_peer(gathered[r])
    return gathered


_gather_buf = {}


def merge_device_union_finds(uf, dist, timings=None):
    """NCCL path: uf is a seqrush_b200.BidirectedUnionFind. Everything is ordered on the context's
    stream -- compress, ONE all_gather_into_tensor of the compressed parent arrays into a [world, n]
    buffer, ONE merge pass over all peers, compress -- with no host synchronisation: the collective
    is issued with the context's stream current, so NCCL waits for the compress kernel and the merge
    kernel waits for NCCL through stream events. `timings` (optional dict) receives CUDA events."""
    import torch

    world = dist.get_world_size()
    if world == 1:
        uf.compress_async()
        return
    dev = torch.device("cuda", uf.ctx.device)
    stream = torch.cuda.ExternalStream(uf.ctx.stream_ptr(), device=dev)
    n = int(uf.capacity)
    key = (uf.ctx.device, world, n)
    buf = _gather_buf.get(key)
    if buf is None:
        _gather_buf.clear()
        buf = _gather_buf[key] = torch.empty((world, n), dtype=torch.int32, device=dev)
    with torch.cuda.stream(stream):
        if timings is not None:
            ev = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
            ev[0].record(stream)
        uf.compress_async()
        t = uf.as_tensor()
        dist.all_gather_into_tensor(buf.view(-1), t)
        if timings is not None:
            ev[1].record(stream)
        uf.merge_gathered(buf.data_ptr(), world, dist.get_rank())
        uf.compress_async()
        if timings is not None:
            ev[2].record(stream)
            timings["events"] = ev
            timings["gathered_bytes"] = int(buf.numel() * 4)
