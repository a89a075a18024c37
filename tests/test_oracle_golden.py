"""The oracle pinned against the reference's own golden vectors and known answers (CPU only)."""
import numpy as np
import pytest

import _oracle
from conftest import expand_rle, rle


def test_ref_reproduces_wfa2_biwfa_affine2p_goldens(oracle, wfa2_golden):
    """WFA2-lib/tests/wfa.utest.check/test.biwfa.affine2p.alg: all 305 `score<TAB>CIGAR` lines."""
    pen = tuple(wfa2_golden["penalties"][1:])
    assert len(wfa2_golden["vectors"]) == 305
    for i, v in enumerate(wfa2_golden["vectors"]):
        st, score, cig = oracle.wfa_only(v["p"].encode(), v["t"].encode(), pen=pen)
        assert st == 0
        assert rle(cig) == v["cigar"], i
        # the golden score column is the penalty recomputed from the CIGAR (align_benchmark prints it)
        assert -_cigar_penalty(cig, pen) == v["score"], i


def _cigar_penalty(cig, pen):
    x, o1, e1, o2, e2 = pen
    tot, i, n = 0, 0, len(cig)
    while i < n:
        op = cig[i]
        j = i
        while j < n and cig[j] == op:
            j += 1
        ln = j - i
        if op == ord("X"):
            tot += x * ln
        elif op in (ord("I"), ord("D")):
            tot += min(o1 + e1 * ln, o2 + e2 * ln)
        i = j
    return tot


def test_ref_reproduces_committed_seqrush_pairs(oracle, seqrush_golden):
    """The fixtures made by scripts/make_golden.py replay identically with whatever ISA variant of
    oracle/_ref this host selects (avx512 / avx2 / generic kernels must agree)."""
    for c in seqrush_golden["cases"]:
        r = oracle.align_pair(c["q"].encode("latin1"), c["t"].encode("latin1"))
        assert (int(r["is_reverse"]), r["status"], r["score"], r["fwd_ed"], r["rev_ed"]) == \
               (c["is_reverse"], c["status"], c["score"], c["fwd_ed"], c["rev_ed"]), c["tag"]
        assert rle(r["cigar"]) == c["cigar"], c["tag"]


def test_isa_variants_agree(built, seqrush_golden):
    import os
    outs = []
    for isa in ("generic", "avx2", "avx512"):
        if not os.path.exists(_oracle.ref_path(isa)):
            continue
        if isa != "generic" and isa == "avx512" and _oracle.ref_isa() != "avx512":
            continue
        if isa == "avx2" and _oracle.ref_isa() == "generic":
            continue
        o = _oracle.Oracle(isa)
        outs.append([o.align_pair(c["q"].encode("latin1"), c["t"].encode("latin1"))["cigar"]
                     for c in seqrush_golden["cases"][:20]])
    for other in outs[1:]:
        assert other == outs[0]


def test_kat_wfa2_cigar_debug(oracle):
    """tests/test_wfa2_cigar_debug.rs:8-29: ATCGATCG vs ATCGATCGATCG, affine 0,5,8,2 -> MMMMMMMMIIII."""
    st, score, cig = oracle.wfa_only(b"ATCGATCG", b"ATCGATCGATCG")
    assert st == 0 and cig == b"MMMMMMMMIIII"


def test_score_is_optimal_vs_quadratic_dp(oracle):
    rng = np.random.default_rng(5)
    acgt = np.frombuffer(b"ACGT", dtype=np.uint8)
    from seqrush_b200 import synth
    for L, d in ((40, 0.1), (150, 0.05), (300, 0.2), (500, 0.03)):
        for _ in range(4):
            anc = acgt[rng.integers(0, 4, size=L)]
            a, b = synth.mutate(rng, anc, d), synth.mutate(rng, anc, d)
            st, _, cig = oracle.wfa_only(a, b)
            assert st == 0
            assert _cigar_penalty(cig, (5, 8, 2, 24, 1)) == _oracle.swg2p_penalty(a, b)


def test_score_quirk_int32_min_on_base_case(oracle):
    """WFA2-lib leaves cigar->score at INT32_MIN when level 0 goes through the base case
    (wavefront_bialign.c:1214-1219 + alignment/cigar.c:79-89): e.g. every self pair."""
    s = b"ACGTTGCA" * 40
    st, score, cig = oracle.wfa_only(s, s)
    assert st == 0 and score == _oracle.INT32_MIN and cig == b"M" * len(s)


# ---- restated Rust glue: the reference's unit-test known answers ----
def test_uf_kat_forward_region(built):
    """src/bidirected_union_find.rs:191-215."""
    uf = _oracle.OracleUF(1000)
    uf.unite_matching_region(100, 200, 10, 15, 5, False, 100)
    assert uf.same(110 * 2, 215 * 2) and uf.same(114 * 2, 219 * 2)
    assert not uf.same(115 * 2, 220 * 2)


def test_uf_kat_reverse_region(built):
    """src/bidirected_union_find.rs:217-257: make_pos(139,true) ~ make_pos(215,false)."""
    uf = _oracle.OracleUF(1000)
    uf.unite_matching_region(100, 200, 10, 15, 3, True, 50)
    assert uf.same(139 * 2 + 1, 215 * 2) and uf.same(137 * 2 + 1, 217 * 2)
    assert not uf.same(139 * 2, 215 * 2)  # a bare union-find keeps the strands apart


def test_uf_capacity_and_fr(built):
    """src/bidirected_union_find.rs:16-24 (capacity 2*max_offset+2); src/seqrush.rs:324-328."""
    uf = _oracle.OracleUF(10)
    assert uf.capacity == 22
    assert not uf.same(4, 5)
    uf.unite_fr_all(10)
    assert uf.same(4, 5) and not uf.same(4, 6)
    assert uf.count_components(10) == 10
    # uf_rush: equal rank -> the smaller index becomes the child, so the odd Pos is the root
    assert uf.find(4) == 5


def test_process_alignment_semantics(built):
    """src/seqrush.rs:1134-1481: runs span consecutive M/= ops, are flushed by X/I/D and mismatching
    bases inside M, and are united only when >= min_match_len."""
    q = b"ACGTACGTAC"
    t = b"ACGTTCGTAC"
    for k, expect in ((0, [0, 1, 2, 3, 5, 6, 7, 8, 9]), (5, [5, 6, 7, 8, 9]), (6, [])):
        uf = _oracle.OracleUF(20)
        assert uf.process_alignment("10M", q, 0, t, 10, k, False) == 0
        got = [i for i in range(10) if uf.same(2 * i, 2 * (10 + i))]
        assert got == expect, k
    uf = _oracle.OracleUF(20)
    uf.process_alignment("4=1X5=", q, 0, t, 10, 0, False)
    assert [i for i in range(10) if uf.same(2 * i, 2 * (10 + i))] == [0, 1, 2, 3, 5, 6, 7, 8, 9]
    # 'I' consumes the query, 'D' the target (SAM convention after cigar_bytes_to_string's swap)
    uf = _oracle.OracleUF(20)
    uf.process_alignment("2=1I2=1D2=", b"ACGTACGT", 0, b"ACTAGCGT"[:8], 8, 0, False)
    assert uf.same(0, 16) and uf.same(2, 18) and uf.same(2 * 3, 2 * (8 + 2)) and uf.same(2 * 4, 2 * (8 + 3))


def test_cigar_bytes_to_string(built):
    """allwave/src/alignment.rs:327-356: RLE, M->'=', I<->D."""
    assert _oracle.cigar_bytes_to_string(b"MMMMXIIDMM") == "4=1X2D1I2="
    assert _oracle.cigar_bytes_to_string(b"") == ""


def test_c1_golden_shape(c1_golden):
    """BASELINE.json configs[0]: 12 sequences, 144 ordered pairs, 22 reverse-complement, 30 INT32_MIN scores
    (measured with the compiled reference)."""
    exp, ids, seqs = c1_golden
    assert len(seqs) == 12 and sum(len(s) for s in seqs) == 163416 == exp["total_length"]
    assert len(exp["pairs"]) == 144
    assert sum(p["is_reverse"] for p in exp["pairs"]) == 22
    assert sum(1 for p in exp["pairs"] if p["score"] == _oracle.INT32_MIN) == 30
    assert exp["partitions"][0]["unite_calls"] == 1621982
