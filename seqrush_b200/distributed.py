"""Multi-GPU sharding of the align-and-unite phase: one process per GPU (torch.distributed over
NCCL), the ordered pair list dealt round-robin, a private union-find per GPU over the full Pos
space, and ONE exchange at the end -- all_gather of the compressed parent arrays followed by a
device-side union pass. The host logic is backend-agnostic so that the gloo
CPU tests can cover it with a stand-in union-find."""
import numpy as np


def shard_pairs(pairs: np.ndarray, rank: int, world: int) -> np.ndarray:
    """Round-robin deal of the ordered pair list (pairs are independent: allwave/src/iterator.rs:185-203)."""
    return np.ascontiguousarray(pairs[rank::world])


def merge_forests(local_roots, world: int, dist, unite_with_peer):
    """local_roots: tensor of this rank's compressed parent array (parent[p] == find(p)).
    Gathers every rank's array and calls unite_with_peer(peer_roots_tensor) for each other rank."""
    import torch

    gathered = [torch.empty_like(local_roots) for _ in range(world)]
    dist.all_gather(gathered, local_roots)
    me = dist.get_rank()
    for r in range(world):
        if r != me:
            unite_with_peer(gathered[r])
    return gathered


def merge_device_union_finds(uf, dist):
    """NCCL path: uf is a seqrush_b200.BidirectedUnionFind."""
    import torch

    uf.compress()
    t = uf.as_tensor()
    world = dist.get_world_size()
    if world == 1:
        return
    # the library launches on the context's own stream, NCCL on torch's: order them explicitly
    # (compress -> all_gather -> merge), once per step
    uf.ctx.synchronize()

    def unite_with_peer(peer):
        torch.cuda.current_stream(peer.device).synchronize()
        uf.merge_roots(peer.data_ptr(), peer.numel())

    merge_forests(t, world, dist, unite_with_peer)
    uf.compress()
