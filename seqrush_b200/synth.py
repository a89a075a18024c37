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
