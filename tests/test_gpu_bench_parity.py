"""Parity AT THE BENCHMARKED CONFIGURATIONS: the CUDA path through the C ABI against the reference's
outputs committed in tests/golden/bench_<set>.npz (scripts/make_golden_bench.py: oracle/_ref =
WFA2-lib + edlib compiled from the reference, plus the restated Rust glue), on the very inputs
bench.py times. Bit-exact: strand, raw score, edit distance, CIGAR length and sha1 for EVERY ordered
pair; unite-call count, component count, canonical partition sha1 and srgpu_uf_checksum per set.

  c2    BASELINE.json configs[1] in full (10^4 pairs at 50 kb: the bench workload, 128x4 launch shape)
  c3s   configs[2] slice, 40 x 50 kb with reverse-complemented sequences and inversions
  c4s   configs[3] slice, 24 x 100 kb (the 256x2 launch shape used above 70 kb)
  gmem  6 x 52 kb with IUPAC bytes: 8-bit symbols, sequences read from global memory (no staging)
"""
import hashlib
import json
import os

import numpy as np
import pytest

from conftest import GOLDEN

pytestmark = pytest.mark.gpu


def load_bench_golden(name):
    z = np.load(os.path.join(GOLDEN, "bench_%s.npz" % name))
    meta = json.loads(bytes(z["meta"]).decode())
    return meta, z


@pytest.fixture(scope="module")
def s(built):
    import torch
    assert torch.cuda.is_available(), "gpu tests need a CUDA device"
    import seqrush_b200 as s
    return s


def seq_digest(seqs):
    h = hashlib.sha1()
    for a in seqs:
        h.update(np.uint64(a.size).tobytes())
        h.update(a.tobytes())
    return h.hexdigest()


@pytest.mark.parametrize("name", ["gmem", "c3s", "c4s", "c2"])
def test_bench_set_matches_reference(s, name):
    from seqrush_b200 import synth
    meta, z = load_bench_golden(name)
    ids, seqs, pairs, _ = synth.bench_set(name)
    assert seq_digest(seqs) == meta["seq_sha1"], "synthetic inputs drifted from the ones the golden was made on"
    assert np.array_equal(pairs, z["pairs"])
    ctx = s.Context(0)
    try:
        # (1) every ordered pair through srgpu_align_pairs (the AllPairIterator -> AlignmentResult stream)
        it = s.AllPairIterator([s.AllwaveSequence(i, a.tobytes()) for i, a in zip(ids, seqs)], s.AlignmentParams(), False,
                               False, pairs=pairs, ctx=ctx)
        k = 0
        bad = []
        for rs in it._batches(True, 1024):
            for r in rs:
                ok = (int(r.is_reverse) == int(z["is_reverse"][k]) and r.score == int(z["score"][k])
                      and r.edit_distance == min(int(z["fwd_ed"][k]), int(z["rev_ed"][k]))
                      and len(r.cigar_bytes) == int(z["cigar_len"][k])
                      and hashlib.sha1(r.cigar_bytes).digest() == z["cigar_sha1"][k].tobytes())
                if not ok:
                    bad.append((k, int(pairs[k, 0]), int(pairs[k, 1])))
                k += 1
        assert k == pairs.shape[0]
        assert not bad, "%d of %d pairs differ from the reference, first: %s" % (len(bad), k, bad[:5])
        assert ctx.stats()["failed_pairs"] == 0
        # (2) the fused path: align + CIGAR walk + unite, then the partition
        sr = s.SeqRush([s.Sequence(i, a.tobytes()) for i, a in zip(ids, seqs)], ctx=ctx)
        st = sr.align_and_unite(s.Args(min_match_length=meta["min_match_len"]), pairs=pairs)
        assert st["failed_pairs"] == 0
        assert st["unite_ops"] == meta["unite_calls"]
        assert sr.count_components() == meta["components"]
        assert "%016x" % sr.union_find.checksum() == meta["uf_checksum"]
        canon = sr.canonical_partition()
        assert hashlib.sha1(canon.astype("<u8").tobytes()).hexdigest() == meta["partition_sha1"]
    finally:
        ctx.close()


@pytest.mark.parametrize("split", [0, 250, 400, 700, 3000, 100000])
def test_split_threshold_does_not_change_results(s, split):
    """Every split threshold -- none (one CTA per pair from start to end, also the retry path), the smallest
    allowed, and one above every score (the whole pair becomes one one-warp task) -- gives the reference's
    records and partition on the c3 slice (reverse-complemented sequences, inversions)."""
    from seqrush_b200 import synth
    meta, z = load_bench_golden("c3s")
    ids, seqs, pairs, _ = synth.bench_set("c3s")
    sel = np.arange(0, pairs.shape[0], 5 if split in (0, 700) else 23)  # a few hundred / a few dozen pairs
    sub = np.ascontiguousarray(pairs[sel])
    ctx = s.Context(0)
    try:
        ctx.set_split(split)
        it = s.AllPairIterator([s.AllwaveSequence(i, a.tobytes()) for i, a in zip(ids, seqs)], s.AlignmentParams(), False,
                               False, pairs=sub, ctx=ctx)
        k = 0
        for rs in it._batches(True, 256):
            for r in rs:
                g = int(sel[k])
                assert int(r.is_reverse) == int(z["is_reverse"][g]) and r.score == int(z["score"][g]), (split, g)
                assert len(r.cigar_bytes) == int(z["cigar_len"][g]), (split, g)
                assert hashlib.sha1(r.cigar_bytes).digest() == z["cigar_sha1"][g].tobytes(), (split, g)
                k += 1
        assert k == sel.size
        assert ctx.stats()["failed_pairs"] == 0
    finally:
        ctx.close()


def test_split_threshold_argument_checked(s):
    ctx = s.Context(0)
    try:
        with pytest.raises(Exception):
            ctx.set_split(100)
        ctx.set_split(0)
        ctx.set_split(250)
    finally:
        ctx.close()


@pytest.mark.parametrize("env", [{"SRGPU_TEST_TASK_CAP": "40"}, {"SRGPU_TEST_LEAF_CAP": "200"}, {"SRGPU_TEST_SUB_W": "320"}, {"SRGPU_TEST_ASM_MAX": "16"}, {"SRGPU_TWO_PASS": "0"},
                                 {"SRGPU_MIRROR_ORIENT": "0"}])
def test_split_buffers_overflow_and_variants(s, env, monkeypatch):
    """Pairs whose sub-problems do not fit the task / leaf arrays are re-run on the monolithic path (tiny arrays force
    it for most pairs of the launch); the one-pass form of the one-warp kernel and both-orders strand choice give the
    same records. All against the c3 slice golden."""
    from seqrush_b200 import synth
    for k, v in env.items():
        monkeypatch.setenv(k, v)
    meta, z = load_bench_golden("c3s")
    ids, seqs, pairs, _ = synth.bench_set("c3s")
    sel = np.arange(0, pairs.shape[0], 17)
    sub = np.ascontiguousarray(pairs[sel])
    ctx = s.Context(0)  # reads the environment
    try:
        it = s.AllPairIterator([s.AllwaveSequence(i, a.tobytes()) for i, a in zip(ids, seqs)], s.AlignmentParams(), False,
                               False, pairs=sub, ctx=ctx)
        k = 0
        for rs in it._batches(True, 256):
            for r in rs:
                g = int(sel[k])
                assert int(r.is_reverse) == int(z["is_reverse"][g]) and r.score == int(z["score"][g]), (env, g)
                assert hashlib.sha1(r.cigar_bytes).digest() == z["cigar_sha1"][g].tobytes(), (env, g)
                k += 1
        assert k == sel.size
        st = ctx.stats()
        assert st["failed_pairs"] == 0
        if any(k.startswith("SRGPU_TEST_") for k in env):
            assert st["retries"] > 0, "the tiny arrays were meant to overflow"
    finally:
        ctx.close()
