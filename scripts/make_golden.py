#!/usr/bin/env python
"""Generate tests/golden/* -- run in the development container (needs /root/reference and
oracle/_ref built by `make -C oracle`). The outputs are committed; the GPU box only reads them.

 1. wfa2_biwfa_affine2p.json.gz : the reference's own 305 BiWFA gap-affine-2p golden vectors
      (inputs WFA2-lib/tests/wfa.utest.seq, expected `score<TAB>CIGAR` lines
      WFA2-lib/tests/wfa.utest.check/test.biwfa.affine2p.alg; penalties 0,4,6,2,24,1 per
      WFA2-lib/tools/align_benchmark/align_benchmark_params.c:64-71), re-packed as JSON.
 2. seqrush_pairs.json.gz : pairs under seqrush's penalties 0,5,8,2,24,1 with the outputs of the
      REFERENCE ITSELF (oracle/_ref = WFA2-lib + edlib compiled from /root/reference): orientation,
      both edit distances, raw score, run-length CIGAR. Covers 1 % / 5 % / 15 % divergence, long
      indels, N / lower-case bytes, whole-sequence RC, ragged lengths, 1-base and identical inputs.
 4. c1_gfa.npz : configs[0] as GFA from the restated builder / compaction / writer (see config1_gfa).
 3. DRB1-3123.fa.gz + c1_expected.json.gz : BASELINE.json configs[0] (HLA-zoo DRB1-3123, 12 seqs,
      -k 0): per ordered pair orientation / score / CIGAR digest from oracle/_ref, the unite-call
      count and a digest of the canonical partition from the restated glue.
"""
import gzip
import hashlib
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import _oracle  # noqa: E402
from seqrush_b200 import synth  # noqa: E402

REF = "/root/reference"
OUT = os.path.join(ROOT, "tests", "golden")


def rle(cig: bytes) -> str:
    if not cig:
        return ""
    a = np.frombuffer(cig, dtype=np.uint8)
    cut = np.flatnonzero(a[1:] != a[:-1]) + 1
    st = np.concatenate(([0], cut))
    en = np.concatenate((cut, [a.size]))
    return "".join("%d%s" % (e - s, chr(a[s])) for s, e in zip(st, en))


def dump(name, obj):
    with gzip.GzipFile(os.path.join(OUT, name), "wb", mtime=0) as f:
        f.write(json.dumps(obj, separators=(",", ":")).encode())
    print("wrote", name, os.path.getsize(os.path.join(OUT, name)), "bytes")


def wfa2_goldens():
    seqs = open(os.path.join(REF, "WFA2-lib/tests/wfa.utest.seq")).read().split("\n")
    alg = open(os.path.join(REF, "WFA2-lib/tests/wfa.utest.check/test.biwfa.affine2p.alg")).read().strip().split("\n")
    pats = [l[1:] for l in seqs if l.startswith(">")]
    txts = [l[1:] for l in seqs if l.startswith("<")]
    assert len(pats) == len(txts) == len(alg) == 305
    vec = []
    for p, t, a in zip(pats, txts, alg):
        sc, cg = a.split("\t")
        vec.append({"p": p, "t": t, "score": int(sc), "cigar": cg})
    dump("wfa2_biwfa_affine2p.json.gz", {"penalties": [0, 4, 6, 2, 24, 1], "source": "WFA2-lib/tests/wfa.utest.seq + "
          "wfa.utest.check/test.biwfa.affine2p.alg (SAM-style run-length CIGAR with M/X/I/D)", "vectors": vec})


def seqrush_pairs():
    orc = _oracle.Oracle()
    rng = np.random.default_rng(20261104)
    cases = []

    def add(tag, q, t):
        q = np.ascontiguousarray(q, dtype=np.uint8)
        t = np.ascontiguousarray(t, dtype=np.uint8)
        r = orc.align_pair(q, t)
        cases.append({"tag": tag, "q": q.tobytes().decode("latin1"), "t": t.tobytes().decode("latin1"),
                      "is_reverse": int(r["is_reverse"]), "status": r["status"], "score": r["score"],
                      "fwd_ed": r["fwd_ed"], "rev_ed": r["rev_ed"], "cigar": rle(r["cigar"])})

    acgt = np.frombuffer(b"ACGT", dtype=np.uint8)
    for L, d, n in ((300, 0.02, 6), (2000, 0.01, 6), (2000, 0.05, 6), (8000, 0.01, 4), (8000, 0.05, 3), (3000, 0.15, 3),
                    (20000, 0.02, 2)):
        for i in range(n):
            anc = acgt[rng.integers(0, 4, size=L)]
            a, b = synth.mutate(rng, anc, d), synth.mutate(rng, anc, d)
            add("div%g_L%d" % (d, L), a, b)
    # whole-sequence reverse complement and inversions
    for L, d in ((1500, 0.01), (6000, 0.02), (6000, 0.05)):
        anc = acgt[rng.integers(0, 4, size=L)]
        a, b = synth.mutate(rng, anc, d), synth.mutate(rng, anc, d)
        add("rc_L%d" % L, synth.revcomp(a), b)
        c = b.copy()
        c[L // 3:L // 3 + L // 5] = synth.revcomp(c[L // 3:L // 3 + L // 5])
        add("inv_L%d" % L, a, c)
    # long indels (exercise the second affine piece) and ragged lengths
    for L, gap in ((4000, 40), (4000, 400), (10000, 1500)):
        anc = acgt[rng.integers(0, 4, size=L)]
        a = synth.mutate(rng, anc, 0.01)
        b = synth.mutate(rng, anc, 0.01)
        cut = int(rng.integers(L // 4, L // 2))
        add("del%d_L%d" % (gap, L), a, np.concatenate([b[:cut], b[cut + gap:]]))
        add("ins%d_L%d" % (gap, L), np.concatenate([a[:cut], a[cut + gap:]]), b)
    # non-ACGT bytes: N runs, lower case (byte equality is case-sensitive; allwave RC maps other -> N)
    for L in (1200, 5000):
        anc = acgt[rng.integers(0, 4, size=L)]
        a, b = synth.mutate(rng, anc, 0.02), synth.mutate(rng, anc, 0.02)
        a = a.copy()
        b = b.copy()
        a[100:160] = ord("N")
        b[100:150] = ord("N")
        lo = slice(L // 2, L // 2 + 80)
        a[lo] = np.frombuffer(a[lo].tobytes().lower(), dtype=np.uint8)
        add("N_lower_L%d" % L, a, b)
        add("N_lower_rc_L%d" % L, synth.revcomp(b), a)  # RC query with N/lower-case target
        q = a.copy()
        q[10:14] = np.frombuffer(b"RYKM", dtype=np.uint8)
        add("iupac_rc_L%d" % L, q[::-1].copy(), b)
    # degenerate shapes
    add("one_base_eq", np.frombuffer(b"A", dtype=np.uint8), np.frombuffer(b"A", dtype=np.uint8))
    add("one_base_ne", np.frombuffer(b"A", dtype=np.uint8), np.frombuffer(b"C", dtype=np.uint8))
    add("kat_test_wfa2_cigar_debug", np.frombuffer(b"ATCGATCG", dtype=np.uint8), np.frombuffer(b"ATCGATCGATCG", dtype=np.uint8))
    anc = acgt[rng.integers(0, 4, size=700)]
    add("identical_700", anc, anc)
    add("short_vs_long", anc[:90], acgt[rng.integers(0, 4, size=900)])
    add("long_vs_short", acgt[rng.integers(0, 4, size=900)], anc[:90])
    add("unrelated_400", acgt[rng.integers(0, 4, size=400)], acgt[rng.integers(0, 4, size=450)])
    add("homopolymer", np.full(500, ord("A"), dtype=np.uint8), np.full(530, ord("A"), dtype=np.uint8))
    add("palindromic_tie", np.frombuffer(b"ACGT" * 60, dtype=np.uint8), np.frombuffer(b"ACGT" * 60, dtype=np.uint8))
    dump("seqrush_pairs.json.gz", {"penalties": [0, 5, 8, 2, 24, 1], "source": "oracle/_ref (reference WFA2-lib + edlib "
          "compiled from /root/reference), generated by scripts/make_golden.py; CIGAR in WFA convention, run-length",
          "cases": cases})


def load_fasta(path):
    ids, seqs, cur = [], [], None
    for line in open(path, "rb"):
        line = line.rstrip(b"\r\n")
        if line.startswith(b">"):
            ids.append(line[1:].split()[0].decode())
            cur = []
            seqs.append(cur)
        elif cur is not None:
            cur.append(line.strip())
    return ids, [np.frombuffer(b"".join(s), dtype=np.uint8) for s in seqs]


def config1():
    src = os.path.join(REF, "HLA-zoo/seqs/DRB1-3123.fa")
    raw = open(src, "rb").read()
    with gzip.GzipFile(os.path.join(OUT, "DRB1-3123.fa.gz"), "wb", mtime=0) as f:
        f.write(raw)
    ids, seqs = load_fasta(src)
    n = len(seqs)
    pairs = np.array([(i, j) for i in range(n) for j in range(n)], dtype=np.uint32)
    orc = _oracle.Oracle()
    tot = int(sum(s.size for s in seqs))
    per_pair = []
    for k in (0, 11):
        uf = _oracle.OracleUF(tot)
        uf.unite_fr_all(tot)
        res, st = orc.run_pairs(seqs, pairs, min_match_len=k, threads=8, uf=uf)
        canon = uf.canonical(2 * tot)
        entry = {"min_match_len": k, "unite_calls": int(st["unite_calls"]),
                 "components": int(uf.count_components(tot)),
                 "partition_sha1": hashlib.sha1(canon.astype("<u8").tobytes()).hexdigest()}
        per_pair.append(entry)
    pp = []
    for (i, j) in pairs:
        r = orc.align_pair(seqs[i], seqs[j])
        pp.append({"q": int(i), "t": int(j), "is_reverse": int(r["is_reverse"]), "score": r["score"],
                   "fwd_ed": r["fwd_ed"], "rev_ed": r["rev_ed"], "cigar_len": len(r["cigar"]),
                   "cigar_sha1": hashlib.sha1(r["cigar"]).hexdigest()})
    dump("c1_expected.json.gz", {"fasta": "DRB1-3123.fa.gz (HLA-zoo/seqs/DRB1-3123.fa)", "penalties": [0, 5, 8, 2, 24, 1],
                                 "total_length": tot, "pairs": pp, "partitions": per_pair,
                                 "source": "oracle/_ref + oracle/oracle.cpp, scripts/make_golden.py"})


def config1_gfa():
    """BASELINE.json configs[0] as a GFA: the restated write_bidirected_gfa (oracle/oracle.cpp, literal
    builder + compact + renumber + validate + write_gfa, --skip-sgd --skip-groom --skip-topo) over the
    union-find the restated glue reaches when the 144 pairs are processed SEQUENTIALLY in the reference's
    order (uf_rush roots depend on union order; this is the reference at -t 1). Stored: the compacted GFA text
    (L lines sorted), the sha256 of the --no-compact text and the union-find root of every position, which is what
    fixes the node bases (src/bidirected_builder.rs:176-182)."""
    ids, seqs = load_fasta(os.path.join(REF, "HLA-zoo/seqs/DRB1-3123.fa"))
    n = len(seqs)
    pairs = np.array([(i, j) for i in range(n) for j in range(n)], dtype=np.uint32)
    orc = _oracle.Oracle()
    tot = int(sum(s.size for s in seqs))
    uf = _oracle.OracleUF(tot)
    uf.unite_fr_all(tot)
    orc.run_pairs(seqs, pairs, min_match_len=0, threads=1, uf=uf, want_results=False)
    raw, rep, nn = _oracle.build_gfa(uf, seqs, ids, no_compact=True)
    gfa, _, _ = _oracle.build_gfa(uf, seqs, ids, no_compact=False)
    np.savez_compressed(os.path.join(OUT, "c1_gfa.npz"), representatives=rep.astype(np.uint32),
                        gfa=np.frombuffer(gfa.encode(), np.uint8),
                        gfa_no_compact_sha256=np.frombuffer(hashlib.sha256(raw.encode()).hexdigest().encode(), np.uint8),
                        uncompacted_nodes=np.array([nn], np.uint64))
    print("wrote c1_gfa.npz", os.path.getsize(os.path.join(OUT, "c1_gfa.npz")), "bytes;", nn, "nodes before compaction,",
          sum(1 for l in gfa.splitlines() if l.startswith("S")), "after; sha256", hashlib.sha256(gfa.encode()).hexdigest())


if __name__ == "__main__":
    os.makedirs(OUT, exist_ok=True)
    if len(sys.argv) > 1 and sys.argv[1] == "gfa":
        config1_gfa()
        sys.exit(0)
    wfa2_goldens()
    seqrush_pairs()
    config1()
    config1_gfa()
