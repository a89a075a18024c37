"""Pair-list sparsification -- host-side mirror of allwave::SparsificationStrategy and what hangs off it
(allwave/src/types.rs:64-80, allwave/src/iterator.rs:49-77,256-334, allwave/src/mash.rs:30-165,
allwave/src/knn_graph.rs:10-172) plus seqrush's flag parser (src/seqrush.rs:356-431).

Hashing (Rust's DefaultHasher = SipHash-1-3, zero keys), sketch selection and the n^2 Jaccard counts run in
the CUDA library; this module turns the integer counts into Mash distances and neighbour lists with plain
f64 arithmetic (math.log = the platform libm, like Rust's f64::ln) and assembles the pair lists in the
reference's order. Parity status: the SipHash core is pinned by published vectors; the Rust framing of the
hashed bytes and therefore every concrete pair list is PARITY UNPINNED (no Rust toolchain, no reference test
holds a hash or a pair list) -- see DESIGN.md."""
import ctypes as C
import math
from dataclasses import dataclass
from typing import Optional

import numpy as np

from . import _lib
from .context import default_context

DEFAULT_KMER_SIZE = 15      # allwave/src/mash.rs:12
DEFAULT_SKETCH_SIZE = 1000  # allwave/src/mash.rs:15


@dataclass
class SparsificationStrategy:  # allwave/src/types.rs:64-80
    kind: str = "none"  # none | random | auto | connectivity | tree
    value: float = 1.0  # Random(keep_fraction) / Connectivity(prob)
    k_nearest: int = 0
    k_farthest: int = 0
    random_fraction: float = 0.0
    kmer_size: Optional[int] = None

    @classmethod
    def none(cls):
        return cls("none")


def parse_sparsification(s: str) -> SparsificationStrategy:
    """src/seqrush.rs:356-431 (the reference exits(1) on an error; this raises ValueError)."""
    if s in ("none", "1.0"):
        return SparsificationStrategy("none")
    if s == "auto":
        return SparsificationStrategy("auto")
    if s.startswith("random:"):
        try:
            f = float(s[7:])
        except ValueError:
            raise ValueError("Invalid random factor: %s" % s)
        if not (0.0 < f <= 1.0):
            raise ValueError("Random factor must be in (0.0, 1.0], got %s" % f)
        return SparsificationStrategy("random", f)
    if s.startswith("connectivity:"):
        try:
            p = float(s[13:])
        except ValueError:
            raise ValueError("Invalid connectivity probability: %s" % s)
        if not (0.0 < p <= 1.0):
            raise ValueError("Connectivity probability must be in (0.0, 1.0], got %s" % p)
        return SparsificationStrategy("connectivity", p)
    if s.startswith("tree:"):
        parts = s[5:].split(",")
        if len(parts) > 4:
            raise ValueError("Tree sampling requires 1-4 values: tree:neighbor[,stranger[,random[,k-mer]]], got %s" % s)
        try:
            kn = int(parts[0])
        except ValueError:
            raise ValueError("Invalid neighbor count: %s" % parts[0])
        kf = 0
        if len(parts) >= 2:
            try:
                kf = int(parts[1])
            except ValueError:
                raise ValueError("Invalid stranger count: %s" % parts[1])
        rf = 0.0
        if len(parts) >= 3:
            try:
                rf = float(parts[2])
            except ValueError:
                raise ValueError("Invalid random fraction: %s" % parts[2])
            if rf < 0.0 or rf > 1.0:
                raise ValueError("Random fraction must be in [0.0, 1.0], got %s" % rf)
        ks = 16  # seqrush's default (:409-418), not allwave's 15
        if len(parts) >= 4:
            try:
                ks = int(parts[3])
            except ValueError:
                raise ValueError("Invalid k-mer size: %s" % parts[3])
            if ks == 0:
                raise ValueError("K-mer size must be > 0")
        if kn < 0 or kf < 0 or ks < 0:
            raise ValueError("Invalid sparsification: '%s'" % s)
        return SparsificationStrategy("tree", 0.0, kn, kf, rf, ks)
    try:  # backward compatibility: a plain float is a random factor
        f = float(s)
    except ValueError:
        f = None
    if f is not None and 0.0 < f <= 1.0:
        return SparsificationStrategy("random", f)
    raise ValueError("Invalid sparsification: '%s'. Use 'none', 'auto', 'random:F', 'connectivity:F', or "
                     "'tree:neighbor[,stranger[,random[,k-mer]]]'" % s)


def rust_hash_bytes(b: bytes) -> int:
    """DefaultHasher over a &[u8] (length prefix, then the bytes): allwave/src/mash.rs:110-114."""
    a = np.frombuffer(bytes(b), dtype=np.uint8)
    return int(_lib.lib().srgpu_rust_hash_bytes(a.ctypes.data if a.size else None, a.size))


def rust_hash_str(s: str) -> int:
    """DefaultHasher over a String (bytes, then 0xff): allwave/src/iterator.rs:262-280."""
    return int(_lib.lib().srgpu_rust_hash_str(s.encode()))


def compute_connectivity_probability(n: int, connectivity_prob: float) -> float:  # allwave/src/iterator.rs:300-334
    if n <= 1:
        return 1.0
    x = min(max(connectivity_prob, 0.001), 0.999)
    if n <= 10:
        return {2: 1.0, 3: 0.8, 4: 0.7, 5: 0.6}.get(n, 0.5)
    c = -math.log(-math.log(x))
    p = (math.log(float(n)) + c) / float(n)
    return min(max(p, 0.001), 1.0)


def _ctx_with(sequences, ctx):
    ctx = ctx or default_context()
    ctx.set_sequences([s.seq if hasattr(s, "seq") else s.data for s in sequences])
    return ctx


def sketches_and_jaccard(sequences, kmer_size: int, sketch_size: int = DEFAULT_SKETCH_SIZE, ctx=None, uploaded=False):
    """-> (sketches[n][sketch_size] u64, counts[n], inter[n][n], uni[n][n]) from the device."""
    ctx = ctx or default_context()
    if not uploaded:
        _ctx_with(sequences, ctx)
    n = len(sequences)
    sk = np.zeros((n, sketch_size), dtype=np.uint64)
    cnt = np.zeros(n, dtype=np.uint32)
    inter = np.zeros((n, n), dtype=np.uint32)
    uni = np.zeros((n, n), dtype=np.uint32)
    _lib.check(_lib.lib().srgpu_mash_sketch_jaccard(ctx._h, kmer_size, sketch_size, sk.ctypes.data, cnt.ctypes.data,
                                                    inter.ctypes.data, uni.ctypes.data))
    return sk, cnt, inter, uni


def mash_distance_from_counts(inter: int, uni: int, k: int) -> float:  # KmerSketch::mash_distance, mash.rs:57-73
    jac = 0.0 if uni == 0 else inter / uni
    if jac <= 0.0:
        return 1.0
    ratio = (2.0 * jac) / (1.0 + jac)
    if ratio <= 0.0:
        return 1.0
    return (-1.0 / float(k)) * math.log(ratio)


def compute_distance_matrix_with_params(sequences, k: int, sketch_size: int = DEFAULT_SKETCH_SIZE, ctx=None, uploaded=False):
    """allwave/src/mash.rs:144-165."""
    _, _, inter, uni = sketches_and_jaccard(sequences, k, sketch_size, ctx, uploaded)
    n = len(sequences)
    cache = {}
    d = np.zeros((n, n), dtype=np.float64)
    for i in range(n):
        for j in range(i + 1, n):
            key = (int(inter[i, j]), int(uni[i, j]))
            v = cache.get(key)
            if v is None:
                v = cache[key] = mash_distance_from_counts(key[0], key[1], k)
            d[i, j] = d[j, i] = v
    return d


def build_knn_graph(distance_matrix: np.ndarray, k_neighbors: int, farthest: bool):  # knn_graph.rs:110-142
    n = distance_matrix.shape[0]
    pairs = []
    for i in range(n):
        others = np.array([j for j in range(n) if j != i], dtype=np.int64)
        row = distance_matrix[i, others]
        order = np.argsort(-row if farthest else row, kind="stable")  # sort_by is stable: ties keep index order
        for j in others[order[:min(k_neighbors, others.size)]]:
            pairs.append((i, int(j)))
    return pairs


def pair_filter(ids, fraction: float, exclude_self: bool, ctx=None):
    """Hashed pair filter over the i-major n^2 list: should_include_pair / apply_random_sparsification
    (knn_graph.rs:159-172, iterator.rs:256-284). Returns the kept (i, j) pairs in i-major order."""
    ctx = ctx or default_context()
    n = len(ids)
    names = (C.c_char_p * n)(*[i.encode() for i in ids])
    keep = np.zeros((n, n), dtype=np.uint8)
    _lib.check(_lib.lib().srgpu_pair_hash_filter(ctx._h, names, n, float(fraction), int(bool(exclude_self)), keep.ctypes.data, None))
    ii, jj = np.nonzero(keep)
    return list(zip(ii.tolist(), jj.tolist()))


def extract_tree_pairs_separated(sequences, k_nearest, k_farthest, random_fraction, kmer_size, ctx=None, uploaded=False):
    """knn_graph.rs:54-97: (tree pairs sorted + dedup'd, random pairs not already among them)."""
    if len(sequences) < 2:
        return [], []
    d = compute_distance_matrix_with_params(sequences, kmer_size, 1000, ctx, uploaded)
    tree = []
    if k_nearest > 0:
        tree += build_knn_graph(d, k_nearest, False)
    if k_farthest > 0:
        tree += build_knn_graph(d, k_farthest, True)
    tree = sorted(set(tree))
    rnd = pair_filter([s.id for s in sequences], random_fraction, True, ctx) if random_fraction > 0.0 else []
    ts = set(tree)
    return tree, [p for p in rnd if p not in ts]


def extract_tree_pairs(sequences, k_nearest, k_farthest, random_fraction, kmer_size, ctx=None, uploaded=False):
    """knn_graph.rs:10-50: nearest + farthest + hashed random pairs, sorted, dedup'd (no self pairs)."""
    tree, rnd = extract_tree_pairs_separated(sequences, k_nearest, k_farthest, random_fraction, kmer_size, ctx, uploaded)
    return sorted(set(tree) | set(rnd))


def sparsified_pairs(sequences, strategy: SparsificationStrategy, exclude_self: bool, ctx=None, uploaded=False) -> np.ndarray:
    """The pair list AllPairIterator::with_options ends up with (allwave/src/iterator.rs:37-77)."""
    n = len(sequences)
    ids = [s.id for s in sequences]
    if strategy is None or strategy.kind == "none":
        ii, jj = np.meshgrid(np.arange(n, dtype=np.uint32), np.arange(n, dtype=np.uint32), indexing="ij")
        pr = np.stack([ii.ravel(), jj.ravel()], axis=1)
        if exclude_self:
            pr = pr[pr[:, 0] != pr[:, 1]]
        return np.ascontiguousarray(pr, dtype=np.uint32)
    if strategy.kind in ("random", "auto", "connectivity"):
        keep = (strategy.value if strategy.kind == "random" else
                compute_connectivity_probability(n, 0.95 if strategy.kind == "auto" else strategy.value))
        pr = pair_filter(ids, keep, exclude_self, ctx)
    elif strategy.kind == "tree":
        pr = extract_tree_pairs(sequences, strategy.k_nearest, strategy.k_farthest, strategy.random_fraction,
                                strategy.kmer_size if strategy.kmer_size is not None else DEFAULT_KMER_SIZE, ctx, uploaded)
    else:
        raise ValueError("unknown sparsification strategy %r" % strategy.kind)
    return np.ascontiguousarray(np.array(pr, dtype=np.uint32).reshape(-1, 2))
