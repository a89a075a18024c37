"""Deterministic synthetic inputs of the shapes BASELINE.json names: a random
ACGT ancestor, each sequence mutated independently at per-base rate d split 80 % substitution /
10 % 1-base deletion / 10 % 1-base insertion; optionally some sequences reverse-complemented and
carrying one inverted segment (config 3)."""
import numpy as np

_ACGT = np.frombuffer(b"ACGT", dtype=np.uint8)
_COMP = np.zeros(256, dtype=np.uint8)
for _a, _b in zip(b"ACGT", b"TGCA"):
    _COMP[_a] = _b


def mutate(rng, anc: np.ndarray, d: float) -> np.ndarray:
    n = anc.size
    r = rng.random(n)
    sub = r < 0.8 * d
    dele = (r >= 0.8 * d) & (r < 0.9 * d)
    ins = (r >= 0.9 * d) & (r < d)
    out = anc.copy()
    # substitution to one of the three other bases
    idx = np.searchsorted(_ACGT, out[sub])
    out[sub] = _ACGT[(idx + rng.integers(1, 4, size=int(sub.sum()))) % 4]
    keep = ~dele
    reps = keep.astype(np.int64) + ins.astype(np.int64) * keep.astype(np.int64)
    res = np.repeat(out, reps)
    # inserted copies (second of each repeated pair) get a random base
    pos = np.cumsum(reps) - 1
    ins_pos = pos[ins & keep]
    res[ins_pos] = _ACGT[rng.integers(0, 4, size=ins_pos.size)]
    return res


def revcomp(a: np.ndarray) -> np.ndarray:
    return _COMP[a[::-1]]


def make_sequences(n: int, length: int, divergence: float, seed: int, rc_fraction: float = 0.0,
                   inversion_fraction: float = 0.0):
    """Returns (ids, list of uint8 arrays). IDs are seq%05d (they feed the reference's pair hash)."""
    rng = np.random.default_rng(seed)
    anc = _ACGT[rng.integers(0, 4, size=length)]
    ids, seqs = [], []
    for i in range(n):
        s = mutate(rng, anc, divergence)
        if inversion_fraction > 0 and rng.random() < inversion_fraction:
            seg = int(rng.integers(1000, 5001))
            if seg < s.size:
                a = int(rng.integers(0, s.size - seg))
                s = s.copy()
                s[a:a + seg] = revcomp(s[a:a + seg])
        if rc_fraction > 0 and rng.random() < rc_fraction:
            s = revcomp(s)
        ids.append("seq%05d" % i)
        seqs.append(np.ascontiguousarray(s))
    return ids, seqs


CONFIGS = {
    # name: (n, length, per-sequence divergence, seed, rc_fraction, inversion_fraction)
    "c2": (100, 50_000, 0.01, 2, 0.0, 0.0),
    "c3": (200, 50_000, 0.01, 3, 0.2, 0.1),
    "c4": (1000, 100_000, 0.01, 4, 0.0, 0.0),
    "c5": (2000, 200_000, 0.05, 5, 0.0, 0.0),
}


def make_config(name: str, n=None, length=None):
    nn, L, d, seed, rcf, invf = CONFIGS[name]
    return make_sequences(n or nn, length or L, d, seed, rcf, invf)


def write_fasta(path, ids, seqs):
    with open(path, "wb") as f:
        for i, s in zip(ids, seqs):
            f.write(b">" + i.encode() + b"\n" + s.tobytes() + b"\n")


def all_pairs(n: int) -> np.ndarray:
    """n^2 ordered pairs, i-major, self pairs included (allwave/src/iterator.rs:37-46)."""
    ii, jj = np.meshgrid(np.arange(n, dtype=np.uint32), np.arange(n, dtype=np.uint32), indexing="ij")
    return np.ascontiguousarray(np.stack([ii.ravel(), jj.ravel()], axis=1), dtype=np.uint32)


# Parity sets at the benchmarked configurations: tests/golden/bench_<name>.npz holds the reference's
# outputs for exactly these inputs (scripts/make_golden_bench.py).
BENCH_SETS = {
    # name: (config, n taken from the front of the config's sequence list, description)
    "c2": ("c2", 100, "BASELINE.json configs[1] in full: 100 x 50 kb, 1 %, all 10^4 ordered pairs"),
    "c3s": ("c3", 40, "configs[2] slice: first 40 of 200 x 50 kb (20 % reverse-complemented, inversions), 1600 ordered pairs"),
    "c4s": ("c4", 24, "configs[3] slice: first 24 of 1000 x 100 kb, 576 ordered pairs (256-thread launch shape)"),
    "gmem": ("c2", 6, "6 x 52 kb with IUPAC bytes: 8-bit symbols, staged sequences exceed the shared-memory budget"),
    "c5s": ("c5", 6, "configs[4] slice: first 6 of 2000 x 200 kb at 5 %, the 30 non-self ordered pairs"),
}


def bench_set(name: str):
    """-> (ids, seqs, pairs, description) of a parity set. Sequences are a prefix of the named
    config's list (make_sequences draws them one after another from one generator, so a prefix of
    the full config is reproduced by asking for fewer)."""
    cfg, n, desc = BENCH_SETS[name]
    if name == "gmem":
        ids, seqs = make_sequences(n, 52_000, 0.01, 77)
        seqs = [s.copy() for s in seqs]
        for k, s in enumerate(seqs):  # IUPAC codes and soft-masked stretches: byte equality, allwave RC -> 'N'
            s[1000 + 10 * k:1004 + 10 * k] = np.frombuffer(b"RYKM", dtype=np.uint8)
            s[20_000:20_050] = np.frombuffer(s[20_000:20_050].tobytes().lower(), dtype=np.uint8)
        seqs[3] = np.ascontiguousarray(revcomp_iupac_safe(seqs[3]))
        return ids, seqs, all_pairs(n), desc
    ids, seqs = make_config(cfg, n=n)
    pairs = all_pairs(n)
    if name == "c5s":
        pairs = np.ascontiguousarray(pairs[pairs[:, 0] != pairs[:, 1]])
    return ids, seqs, pairs, desc


def revcomp_iupac_safe(a: np.ndarray) -> np.ndarray:
    """Reverse complement that keeps non-ACGT bytes as they are (so the set has a reverse-strand
    sequence whose IUPAC bytes survive)."""
    tbl = np.arange(256, dtype=np.uint8)
    for x, y in zip(b"ACGTacgt", b"TGCAtgca"):
        tbl[x] = y
    return tbl[a[::-1]]
