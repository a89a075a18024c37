#!/usr/bin/env python
"""Golden vectors AT THE BENCHMARKED CONFIGURATIONS -- run in the development container (needs
oracle/_ref built from /root/reference by `make -C oracle`). Outputs tests/golden/bench_<set>.npz,
committed; the GPU box only reads them (tests/test_gpu_bench_parity.py, bench.py's `parity` key).

Every record is an output of the REFERENCE's own arithmetic (oracle/_ref = WFA2-lib + edlib compiled
from /root/reference) driven through the restated Rust glue (oracle/oracle.cpp: align_pair's strand
rule, cigar_bytes_to_string, process_alignment, uf_rush), on the same deterministic synthetic
inputs bench.py uses (seqrush_b200/synth.py, fixed seeds):

  c2    BASELINE.json configs[1] IN FULL: 100 x 50 kb, all 10^4 ordered pairs (the bench workload)
  c3s   configs[2] slice: first 40 of the 200 x 50 kb sequences (whole-sequence RC + inversions), 1600 pairs
  c4s   configs[3] slice: first 24 of the 1000 x 100 kb sequences, 576 pairs (the 256-thread launch shape)
  gmem  6 x 52 kb carrying IUPAC bytes, so the symbols pack at 8 bits and the two staged sequences
        exceed the shared-memory budget: the global-memory sequence path of align_kernel
  c5s   configs[4] slice: 6 x 200 kb at 5 % divergence, the 30 non-self ordered pairs

Per ordered pair: strand, raw cigar->score (INT32_MIN quirk included), both edit distances, CIGAR
length and sha1 of the CIGAR bytes (WFA convention M/X/I/D). Per set: unite-call count, component
count, sha1 of the canonical partition (min Pos per component, u64 LE) and the value
srgpu_uf_checksum must return for that partition (sum over p < 2*total+2 of mix64((p<<32) ^ root(p))).
"""
import argparse
import hashlib
import json
import os
import sys
import time
from concurrent.futures import ThreadPoolExecutor

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import _oracle  # noqa: E402
from seqrush_b200 import synth  # noqa: E402

OUT = os.path.join(ROOT, "tests", "golden")
PEN = (5, 8, 2, 24, 1)


def mix64(x):
    x = x.astype(np.uint64)
    x ^= x >> np.uint64(33)
    x *= np.uint64(0xff51afd7ed558ccd)
    x ^= x >> np.uint64(33)
    x *= np.uint64(0xc4ceb9fe1a85ec53)
    x ^= x >> np.uint64(33)
    return x


def uf_checksum_of(canon, capacity):
    """What srgpu_uf_checksum returns for a partition whose canonical labels (min Pos) are `canon`
    over [0, len(canon)) and identity above (seqrush_b200/csrc/kernels.cuh uf_checksum_kernel)."""
    root = np.arange(capacity, dtype=np.uint64)
    root[:canon.size] = canon
    with np.errstate(over="ignore"):
        v = mix64((np.arange(capacity, dtype=np.uint64) << np.uint64(32)) ^ root)
        return int(v.sum(dtype=np.uint64))


def seq_digest(seqs):
    h = hashlib.sha1()
    for s in seqs:
        h.update(np.uint64(s.size).tobytes())
        h.update(s.tobytes())
    return h.hexdigest()


def make_set(name):
    """-> (ids, seqs, pairs, description): seqrush_b200/synth.py bench_set, shared with the tests and bench.py."""
    return synth.bench_set(name)


def run_set(name, threads):
    ids, seqs, pairs, desc = make_set(name)
    orc = _oracle.Oracle()
    tot = int(sum(s.size for s in seqs))
    off = np.concatenate(([0], np.cumsum([s.size for s in seqs])))[:-1]
    uf = _oracle.OracleUF(tot)
    uf.unite_fr_all(tot)
    n = pairs.shape[0]
    is_rev = np.zeros(n, np.uint8)
    score = np.zeros(n, np.int32)
    fwd = np.zeros(n, np.int32)
    rev = np.zeros(n, np.int32)
    clen = np.zeros(n, np.int32)
    sha = np.zeros((n, 20), np.uint8)
    t0 = time.time()

    def work(i):
        q, t = int(pairs[i, 0]), int(pairs[i, 1])
        r = orc.align_pair(seqs[q], seqs[t], PEN)
        assert r["status"] == 0
        is_rev[i] = r["is_reverse"]
        score[i] = r["score"]
        fwd[i] = r["fwd_ed"]
        rev[i] = r["rev_ed"]
        clen[i] = len(r["cigar"])
        sha[i] = np.frombuffer(hashlib.sha1(r["cigar"]).digest(), np.uint8)
        cs = _oracle.cigar_bytes_to_string(r["cigar"])
        rc = uf.process_alignment(cs, seqs[q], int(off[q]), seqs[t], int(off[t]), 0, r["is_reverse"])
        assert rc == 0, "validate_match failed (the reference would panic)"
        return i

    with ThreadPoolExecutor(threads) as ex:
        for k, _ in enumerate(ex.map(work, range(n))):
            if k % 500 == 0:
                print("  %s: %d/%d pairs, %.0f s" % (name, k, n, time.time() - t0), flush=True)
    canon = uf.canonical(2 * tot)
    meta = {
        "set": name, "description": desc, "n_seqs": len(seqs), "total_length": tot, "seq_sha1": seq_digest(seqs),
        "penalties": [0] + list(PEN), "min_match_len": 0, "unite_calls": int(uf.unite_calls()),  # unite_fr_all is not counted (oracle.cpp)
        "components": int(uf.count_components(tot)),
        "partition_sha1": hashlib.sha1(canon.astype("<u8").tobytes()).hexdigest(),
        "uf_checksum": "%016x" % uf_checksum_of(canon, 2 * tot + 2),
        "n_reverse": int(is_rev.sum()), "seconds": time.time() - t0, "threads": threads,
        "source": "oracle/_ref (reference WFA2-lib + edlib compiled from /root/reference) + oracle/oracle.cpp, "
                  "scripts/make_golden_bench.py",
    }
    np.savez_compressed(os.path.join(OUT, "bench_%s.npz" % name), meta=np.frombuffer(json.dumps(meta).encode(), np.uint8),
                        pairs=pairs, is_reverse=is_rev, score=score, fwd_ed=fwd, rev_ed=rev, cigar_len=clen,
                        cigar_sha1=sha)
    print(json.dumps(meta), flush=True)


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("sets", nargs="*", default=["gmem", "c3s", "c4s", "c2"])
    ap.add_argument("--threads", type=int, default=os.cpu_count() or 8)
    a = ap.parse_args()
    os.makedirs(OUT, exist_ok=True)
    for s in a.sets:
        run_set(s, a.threads)
