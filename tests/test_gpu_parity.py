"""Parity tests proper: the CUDA path, called through the C ABI (seqrush_b200 mirrors -> ctypes ->
libseqrush_b200.so), against the oracle and the committed goldens. Bit-exact everywhere: strand,
raw score, CIGAR bytes, unite-call counts, canonical partition."""
import hashlib

import numpy as np
import pytest

from conftest import expand_rle, rle

pytestmark = pytest.mark.gpu

INT32_MIN, INT32_MAX = -2**31, 2**31 - 1


@pytest.fixture(scope="module")
def s(built):
    import torch
    assert torch.cuda.is_available(), "gpu tests need a CUDA device"
    import seqrush_b200 as s
    return s


def _align(s, seq_bytes, pairs, params=None):
    it = s.AllPairIterator([s.AllwaveSequence("s%d" % i, b) for i, b in enumerate(seq_bytes)],
                           params or s.AlignmentParams(), False, False, pairs=pairs, ctx=s.Context(0))
    res = it.align_all()
    it.ctx.close()
    return res


def test_wfa2_goldens_through_abi(s, wfa2_golden, oracle):
    """The reference's 305 BiWFA gap-affine-2p vectors (penalties 0,4,6,2,24,1). The ABI always runs the
    strand step first, so vectors whose reverse complement is closer are checked against the oracle's
    align_pair instead of the forward golden line."""
    vec = wfa2_golden["vectors"]
    seqs, pairs = [], []
    for v in vec:
        seqs += [v["p"].encode(), v["t"].encode()]
        pairs.append((len(seqs) - 2, len(seqs) - 1))
    pen = wfa2_golden["penalties"]
    res = _align(s, seqs, pairs, s.AlignmentParams(*pen))
    n_fwd = 0
    for v, r in zip(vec, res):
        o = oracle.align_pair(v["p"].encode(), v["t"].encode(), pen=tuple(pen[1:]))
        assert r.is_reverse == o["is_reverse"]
        assert r.cigar_bytes == o["cigar"] and r.score == o["score"]
        if not r.is_reverse:
            assert rle(r.cigar_bytes) == v["cigar"] and -r.penalty == v["score"]
            n_fwd += 1
    assert n_fwd >= 290


def test_seqrush_pairs_goldens(s, seqrush_golden):
    cases = seqrush_golden["cases"]
    seqs, pairs = [], []
    for c in cases:
        seqs += [c["q"].encode("latin1"), c["t"].encode("latin1")]
        pairs.append((len(seqs) - 2, len(seqs) - 1))
    res = _align(s, seqs, pairs)
    for c, r in zip(cases, res):
        assert int(r.is_reverse) == c["is_reverse"], c["tag"]
        assert r.edit_distance == min(c["fwd_ed"], c["rev_ed"]), c["tag"]
        assert r.score == c["score"], c["tag"]
        assert rle(r.cigar_bytes) == c["cigar"], c["tag"]
        cig = r.cigar_bytes
        assert r.num_matches == cig.count(b"M") and r.alignment_length == cig.count(b"M") + cig.count(b"X")
        assert r.query_end == len(c["q"]) and r.target_end == len(c["t"])  # end-to-end: consumes both fully


def test_config1_hla_drb1_full(s, c1_golden, oracle):
    """BASELINE.json configs[0]: 144 ordered pairs of HLA-zoo DRB1-3123 (944 N bytes, 22 RC pairs, 30
    INT32_MIN scores), -k 0 and -k 11: strand, score, CIGAR digest, unite calls, canonical partition."""
    exp, ids, seqs = c1_golden
    n = len(seqs)
    pairs = [(i, j) for i in range(n) for j in range(n)]
    res = _align(s, seqs, pairs)
    for e, r in zip(exp["pairs"], res):
        assert (r.query_idx, r.target_idx) == (e["q"], e["t"])
        assert int(r.is_reverse) == e["is_reverse"] and r.score == e["score"]
        assert r.edit_distance == min(e["fwd_ed"], e["rev_ed"])
        assert len(r.cigar_bytes) == e["cigar_len"]
        assert hashlib.sha1(r.cigar_bytes).hexdigest() == e["cigar_sha1"], (e["q"], e["t"])
    for part in exp["partitions"]:
        ctx = s.Context(0)
        sr = s.SeqRush([s.Sequence(i, d) for i, d in zip(ids, seqs)], ctx=ctx)
        st = sr.align_and_unite(s.Args(min_match_length=part["min_match_len"]))
        assert st["unite_ops"] == part["unite_calls"]
        canon = sr.canonical_partition()
        assert hashlib.sha1(canon.astype("<u8").tobytes()).hexdigest() == part["partition_sha1"]
        assert sr.count_components() == part["components"]
        ctx.close()


@pytest.mark.parametrize("cfg", [(8, 6000, 0.01, 11, 0.0, 0.0), (10, 4000, 0.03, 12, 0.3, 0.3), (5, 15000, 0.02, 13, 0.2, 0.0)])
def test_seeded_sets_vs_oracle(s, oracle, cfg):
    """Seeded synthetic sets (incl. RC sequences + inversions): every ordered pair and the partition."""
    import _oracle
    from seqrush_b200 import synth
    ids, seqs = synth.make_sequences(*cfg)
    ctx = s.Context(0)
    sr = s.SeqRush([s.Sequence(i, a.tobytes()) for i, a in zip(ids, seqs)], ctx=ctx)
    pairs = sr.all_pairs()
    res = _align(s, [a.tobytes() for a in seqs], pairs)
    ores, _ = oracle.run_pairs(seqs, pairs, threads=4)
    for k, r in enumerate(res):
        o = oracle.align_pair(seqs[r.query_idx], seqs[r.target_idx]) if k % 7 == 0 else None
        assert int(r.is_reverse) == ores[k][0] and r.score == ores[k][2] and len(r.cigar_bytes) == ores[k][3]
        if o is not None:
            assert r.cigar_bytes == o["cigar"]
    for k in (0, 16):
        sr.union_find.close()
        sr.union_find = s.BidirectedUnionFind(sr.total_length, ctx)
        sr.union_find.unite_fr_all(sr.total_length)
        st = sr.align_and_unite(s.Args(min_match_length=k))
        ouf = _oracle.OracleUF(sr.total_length)
        ouf.unite_fr_all(sr.total_length)
        _, ost = oracle.run_pairs(seqs, pairs, min_match_len=k, threads=4, uf=ouf, want_results=False)
        assert st["unite_ops"] == ost["unite_calls"]
        assert np.array_equal(sr.canonical_partition(), ouf.canonical(2 * sr.total_length))
        assert sr.count_components() == ouf.count_components(sr.total_length)
    ctx.close()


def test_ragged_and_degenerate_inputs(s, oracle):
    seqs = [b"A", b"C", b"ACGT" * 30, b"ACGT" * 30 + b"TT", b"G" * 150, b"ACGTNNNNacgtRYKM" * 9, b"T" * 7]
    pairs = [(i, j) for i in range(len(seqs)) for j in range(len(seqs))]
    res = _align(s, seqs, pairs)
    for r in res:
        o = oracle.align_pair(seqs[r.query_idx], seqs[r.target_idx])
        assert (r.is_reverse, r.score, r.cigar_bytes) == (o["is_reverse"], o["score"], o["cigar"]), (r.query_idx, r.target_idx)
    assert _align(s, seqs, np.zeros((0, 2), dtype=np.uint32)) == []


def test_error_behaviour(s):
    ctx = s.Context(0)
    with pytest.raises(s.SrgpuError) as e:
        ctx.set_sequences([b"ACGT", b""])  # src/seqrush.rs:311-316 panics on empty sequences
    assert e.value.code == -2
    with pytest.raises(ValueError):
        s.SeqRush([s.Sequence("a", b"")], ctx=ctx)
    it = s.AllPairIterator([s.AllwaveSequence("a", b"ACGT")], s.AlignmentParams(), False, False, pairs=[(0, 3)], ctx=ctx)
    with pytest.raises(s.SrgpuError) as e:
        it.align_all()
    assert e.value.code == -2
    uf = s.BidirectedUnionFind(10, ctx)
    with pytest.raises(s.SrgpuError) as e:
        uf.unite(3, 22)  # uf_rush asserts x < size (lib.rs:113)
    assert e.value.code == -3
    with pytest.raises(s.SrgpuError):
        uf.find(22)
    with pytest.raises(s.SrgpuError):
        s.AllPairIterator([s.AllwaveSequence("a", b"ACGT")], s.AlignmentParams(1, 5, 8, 2, 24, 1), False, False,
                          pairs=[(0, 0)], ctx=ctx).align_all()
    ctx.close()


def test_union_find_kats(s):
    """src/bidirected_union_find.rs:137-257, tests/test_graph_construction.rs:98-139."""
    ctx = s.Context(0)
    uf = s.BidirectedUnionFind.new(1000, ctx)
    assert uf.capacity == 2002
    assert not uf.same(220, 430) and uf.same(7, 7)
    uf.unite_matching_region(100, 200, 10, 15, 5, False, 100)
    assert uf.same(110 * 2, 215 * 2) and uf.same(114 * 2, 219 * 2) and not uf.same(115 * 2, 220 * 2)
    uf2 = s.BidirectedUnionFind(1000, ctx)
    uf2.unite_matching_region(100, 200, 10, 15, 3, True, 50)
    assert uf2.same(139 * 2 + 1, 215 * 2) and uf2.same(137 * 2 + 1, 217 * 2)
    assert not uf2.same(139 * 2, 215 * 2)      # no implicit forward/reverse unite on a bare union-find
    uf3 = s.BidirectedUnionFind(1000, ctx)
    uf3.unite_matching_region_seq2_rc(100, 200, 10, 15, 3, True, 50)
    assert uf3.same(110 * 2, (200 + 50 - 1 - 15) * 2 + 1) and uf3.same(112 * 2, (200 + 50 - 1 - 17) * 2 + 1)
    # transitivity + canonical representative = min Pos
    uf.unite(2, 40)
    uf.unite(40, 600)
    assert uf.same(2, 600) and uf.find(600) == 2 and uf.find(40) == 2
    uf.unite(9, 9)
    assert uf.find(9) == 9
    uf.unite_fr_all(500)
    assert uf.same(4, 5) and uf.count_components(500) <= 500
    # UFRush::clear (uf_rush lib.rs:211-216): every Pos is its own set again
    uf.clear()
    assert not uf.same(2, 600) and not uf.same(4, 5) and uf.find(600) == 600
    assert uf.count_components(500) == 500
    ctx.close()


def test_uf_concurrent_stress(s):
    """uf_rush lib.rs:330-374 in spirit: many concurrent unites -> exactly one component, and the batched
    region form equals the sequential oracle."""
    import _oracle
    ctx = s.Context(0)
    n = 200_000
    uf = s.BidirectedUnionFind(n, ctx)
    rng = np.random.default_rng(3)
    regs, ouf = [], _oracle.OracleUF(n)
    for _ in range(400):
        ln = int(rng.integers(1, 3000))
        a, b = int(rng.integers(0, n - ln)), int(rng.integers(0, n - ln))
        rc = bool(rng.integers(0, 2))
        regs.append((a, 0, 0, b, ln, True, ln) if rc else (0, 0, a, b, ln, False, ln))
        o1, o2, s1, s2, l, r, l1 = regs[-1]
        ouf.unite_matching_region(o1, o2, s1, s2, l, r, l1)
    uf.unite_regions(regs)
    assert np.array_equal(uf.roots(0, uf.capacity), ouf.canonical(uf.capacity))
    chain = s.BidirectedUnionFind(n, ctx)
    chain.unite_regions([(0, 0, 0, 1, n - 1, False, 0)])  # unite(2i, 2i+2) for all i at once
    roots = chain.roots(0, 2 * n)
    assert (roots[0::2] == 0).all() and (roots[1::2] == np.arange(1, 2 * n, 2)).all()
    ctx.close()


def test_process_alignment_paf_path(s, oracle):
    """src/seqrush.rs:510-609 shares process_alignment: 'M' ops hiding mismatches, RC query, starts."""
    import _oracle
    q = b"ACGTACGTACGGTTACAGGATTACA"
    t = b"ACGTTCGTACGGTTACAGCATTACA"
    for cigar, k, rc, qs, ts in (("25M", 0, False, 0, 0), ("25M", 6, False, 0, 0), ("4=1X20=", 0, False, 0, 0),
                                 ("10M2I3M2D8M", 0, False, 0, 0), ("20M", 3, True, 2, 3), ("5=1X5=", 0, True, 0, 0)):
        ctx = s.Context(0)
        sr = s.SeqRush([s.Sequence("q", q), s.Sequence("t", t)], ctx=ctx)
        sr.union_find.close()
        sr.union_find = s.BidirectedUnionFind(sr.total_length, ctx)
        sr.process_alignment(cigar, 0, 1, k, rc, qs, ts)
        ouf = _oracle.OracleUF(len(q) + len(t))
        ouf.process_alignment(cigar, q, 0, t, len(q), k, rc, qs, ts)
        assert np.array_equal(sr.union_find.roots(0, 2 * sr.total_length), ouf.canonical(2 * sr.total_length)), cigar
        ctx.close()


def test_process_alignments_batched(s, oracle):
    """srgpu_process_alignments: many records in one launch against the restated process_alignment
    (src/seqrush.rs:1134-1481). 'M' runs that hide mismatches (stretches shorter / longer than
    min_match_len, across 32-base word edges), adjacent '='/'M' ops continuing one run, ops that only
    flush ('N', 'S'), a missing count, an 'M' that runs past the end of a sequence, RC queries with
    IUPAC bytes, and PAF start coordinates."""
    import _oracle
    from seqrush_b200 import synth
    rng = np.random.default_rng(99)
    acgt = np.frombuffer(b"ACGT", dtype=np.uint8)
    anc = acgt[rng.integers(0, 4, size=6000)]
    a = synth.mutate(rng, anc, 0.03)
    b = synth.mutate(rng, anc, 0.03)
    c = synth.revcomp(synth.mutate(rng, anc, 0.02)).copy()
    c[100:104] = np.frombuffer(b"RYKM", dtype=np.uint8)
    seqs = [a.tobytes(), b.tobytes(), c.tobytes(), b"ACGTACGTAC", b"ACGTTCGTAC"]
    n1 = min(len(seqs[0]), len(seqs[1]))
    recs = [
        ("%dM" % n1, 0, 1, False, 0, 0),                       # one long 'M' hiding every mismatch and indel shift
        ("%dM" % (n1 - 50), 1, 0, False, 7, 3),                # starts from PAF coordinates
        ("2000=1000M5X3I2D%dM" % 900, 0, 1, False, 0, 0),      # '=' then 'M' continue one run; X/I/D flush
        ("100M1N100M2S50=H10M", 0, 1, False, 10, 10),          # letters that only flush; 'H' with a missing count
        ("%dM" % (len(seqs[0]) + 500), 0, 1, False, 0, 0),     # runs past both ends: skipped, not an error
        ("%dM" % min(len(seqs[2]), len(seqs[1])), 2, 1, True, 0, 0),  # RC query with IUPAC bytes (kept by seqrush's RC)
        ("4=1X5=", 3, 4, False, 0, 0),
        ("M", 3, 4, False, 2, 2),                              # count 0 -> 1
        ("", 3, 4, False, 0, 0),                               # no cg:Z: tag
        ("31M1X1M1X33M1X64M", 0, 0, False, 0, 0),              # self record: word-edge stretches
    ]
    res = _align(s, [seqs[0], seqs[1]], [(0, 1)])
    recs.append((s.cigar_bytes_to_string(res[0].cigar_bytes), 0, 1, False, 0, 0))  # a real alignment
    lens = [len(x) for x in seqs]
    offs = np.concatenate(([0], np.cumsum(lens)))[:-1]
    tot = int(sum(lens))
    for k in (0, 1, 2, 5, 33, 100, 5000):
        ctx = s.Context(0)
        sr = s.SeqRush([s.Sequence("s%d" % i, d) for i, d in enumerate(seqs)], ctx=ctx)
        sr.union_find.close()
        sr.union_find = s.BidirectedUnionFind(sr.total_length, ctx)
        sr.process_alignments(recs, k)
        ouf = _oracle.OracleUF(tot)
        for cigar, qi, ti, rc, qs, ts in recs:
            ouf.process_alignment(cigar, seqs[qi], int(offs[qi]), seqs[ti], int(offs[ti]), k, rc, qs, ts)
        assert ctx.stats()["unite_ops"] == ouf.unite_calls(), k
        assert np.array_equal(sr.union_find.roots(0, 2 * tot), ouf.canonical(2 * tot)), k
        ctx.close()


def test_properties_at_bench_scale(s):
    """Size-independent properties on a slice of BASELINE.json configs[1] (100 x 50 kb, 1 %): the
    partition does not depend on pair order or launch configuration, re-running is idempotent, CIGARs
    consume both sequences, and the raw score equals minus the penalty recomputed from the CIGAR."""
    from seqrush_b200 import synth
    ids, seqs = synth.make_config("c2", n=12)
    ctx = s.Context(0)
    sr = s.SeqRush([s.Sequence(i, a.tobytes()) for i, a in zip(ids, seqs)], ctx=ctx)
    pairs = sr.all_pairs()
    st = sr.align_and_unite(s.Args(min_match_length=0), want_results=True)
    res = sr.last_results
    for i in range(len(pairs)):
        r = res[i]
        assert r.status == 0 and r.query_end == seqs[r.query_idx].size and r.target_end == seqs[r.target_idx].size
        assert r.score in (-r.penalty, INT32_MIN)
        if r.query_idx != r.target_idx:
            assert r.score == -r.penalty and 1000 < r.penalty < 12000
    ck = sr.union_find.checksum()
    comps = sr.count_components()
    sr.align_and_unite(s.Args(min_match_length=0))        # idempotent
    assert sr.union_find.checksum() == ck
    ctx2 = s.Context(0)
    ctx2.configure(threads=64, ctas_per_sm=8, ring_diagonals=4096)   # narrow ring: exercises the retry path
    sr2 = s.SeqRush([s.Sequence(i, a.tobytes()) for i, a in zip(ids, seqs)], ctx=ctx2)
    perm = np.random.default_rng(0).permutation(len(pairs))
    st2 = sr2.align_and_unite(s.Args(min_match_length=0), pairs=pairs[perm])
    assert sr2.union_find.checksum() == ck and sr2.count_components() == comps
    assert st2["unite_ops"] == st["unite_ops"] and st2["wf_cells"] == st["wf_cells"]
    assert st2["retries"] >= 1
    ctx.close()
    ctx2.close()


def test_forest_merge_on_device(s):
    """Multi-GPU merge path on one device: two private forests over disjoint pair shards, merged through
    compress + merge_roots, equal the single forest (the N>1 collective only moves the arrays)."""
    from seqrush_b200 import synth
    from seqrush_b200.distributed import shard_pairs
    ids, seqs = synth.make_sequences(6, 5000, 0.02, 21, 0.3, 0.0)
    ctx = s.Context(0)
    mk = lambda: s.SeqRush([s.Sequence(i, a.tobytes()) for i, a in zip(ids, seqs)], ctx=ctx)
    whole = mk()
    pairs = whole.all_pairs()
    whole.align_and_unite(pairs=pairs)
    a, b = mk(), mk()
    a.align_and_unite(pairs=shard_pairs(pairs, 0, 2))
    b.align_and_unite(pairs=shard_pairs(pairs, 1, 2))
    a.union_find.compress()
    b.union_find.compress()
    a.union_find.merge_roots(b.union_find.device_ptr(), b.union_find.capacity)
    a.union_find.compress()
    assert np.array_equal(a.canonical_partition(), whole.canonical_partition())
    t = a.union_find.as_tensor()
    assert t.is_cuda and t.numel() == a.union_find.capacity
    ctx.close()


def test_paf_out_and_back_in(s, tmp_path):
    """--output-alignments then --paf (src/seqrush.rs:676-716, :510-609): the partition from the fused
    path, from the pass that also writes the PAF, and from reading that PAF back are the same."""
    from seqrush_b200 import synth
    ids, seqs = synth.make_sequences(6, 3000, 0.02, 21, 0.4, 0.0)
    mk = lambda ctx: s.SeqRush([s.Sequence(i, a.tobytes()) for i, a in zip(ids, seqs)], ctx=ctx)
    paf = str(tmp_path / "out.paf")
    for k in (0, 8):
        ctx = s.Context(0)
        a, b, c = mk(ctx), mk(ctx), mk(ctx)  # the context holds one sequence set; these three are the same set
        a.align_and_unite(s.Args(min_match_length=k))
        b.align_and_unite(s.Args(min_match_length=k, output_alignments=paf))
        lines = open(paf).read().splitlines()
        assert len(lines) == len(ids) ** 2
        f = lines[1].split("\t")
        assert len(f) == 14 and f[0] == ids[0] and f[5] == ids[1] and f[4] in "+-"
        assert f[12].startswith("gi:f:") and f[13].startswith("cg:Z:")  # allwave/src/lib.rs:95-111
        assert int(f[10]) == max(int(f[3]) - int(f[2]), int(f[8]) - int(f[7]))
        assert any(l.split("\t")[4] == "-" for l in lines)
        with open(paf, "a") as fh:  # lines the reference skips: empty, short, unknown name
            fh.write("\n" + "x\t1\t0\n" + lines[1].replace(ids[0], "nobody") + "\n")
        assert c.align_and_unite(s.Args(min_match_length=k, paf=paf)) == len(ids) ** 2
        pa = a.canonical_partition()
        assert np.array_equal(pa, b.canonical_partition()) and np.array_equal(pa, c.canonical_partition())
        ctx.close()


def test_iterative_early_stop(s, oracle):
    """--iterative (src/seqrush.rs:868-1132) with explicit pair lists: the grouped device driver stops at
    the same random pair as the reference's pair-by-pair loop (restated over the oracle) and ends with
    the same partition."""
    import _oracle
    from seqrush_b200 import synth
    ids, seqs = synth.make_sequences(8, 2000, 0.02, 31, 0.25, 0.0)
    n = len(ids)
    tree = [(i, i + 1) for i in range(n - 1)]
    rng = np.random.default_rng(5)
    rand = [(int(a), int(b)) for a, b in rng.integers(0, n, size=(400, 2)) if a != b]
    ctx = s.Context(0)
    sr = s.SeqRush([s.Sequence(i, a.tobytes()) for i, a in zip(ids, seqs)], ctx=ctx)
    out = sr.align_and_unite_iterative(s.Args(), tree, rand)
    # the reference's loop: one pair at a time, check every 10 pairs, stop after 10 unchanged counts
    ouf = _oracle.OracleUF(sr.total_length)
    ouf.unite_fr_all(sr.total_length)
    four = lambda i, j: np.array([(i, i), (i, j), (j, i), (j, j)], dtype=np.uint32)
    for i, j in tree:
        oracle.run_pairs(seqs, four(i, j), threads=4, uf=ouf, want_results=False)
    post_tree = ouf.count_components(sr.total_length)
    prev, stable, done = post_tree, 0, 0
    for k, (i, j) in enumerate(rand):
        oracle.run_pairs(seqs, four(i, j), threads=4, uf=ouf, want_results=False)
        done = k + 1
        if (k + 1) % 10 == 0:
            cnt = ouf.count_components(sr.total_length)
            if cnt == prev:
                stable += 1
                if stable >= 10:
                    break
            else:
                stable = 0
            prev = cnt
    assert out["post_tree_components"] == post_tree
    assert out["random_pairs_processed"] == done and out["stabilized"] == (done < len(rand))
    assert 100 <= done < len(rand)  # the case does stop early
    assert out["final_components"] == ouf.count_components(sr.total_length)
    assert np.array_equal(sr.canonical_partition(), ouf.canonical(2 * sr.total_length))
    ctx.close()


def _gfa_flip_normalise(txt_ours, txt_ref):
    """--no-compact GFAs of the same partition differ only by the orientation of nodes whose base the two
    sides read from different members of a mixed-orientation component: flip those nodes in `txt_ref`
    (complement the base, toggle the sign of every step and edge end) and return the re-sorted text."""
    comp = {"A": "T", "T": "A", "C": "G", "G": "C"}
    base_ours = {f[1]: f[2] for f in (l.split("\t") for l in txt_ours.splitlines()) if f[0] == "S"}
    flip = set()
    out, links = [], []
    for l in txt_ref.splitlines():
        f = l.split("\t")
        if f[0] == "S":
            if base_ours[f[1]] != f[2]:
                assert comp.get(f[2].upper()) == base_ours[f[1]].upper(), (f, base_ours[f[1]])
                flip.add(f[1])
                f[2] = base_ours[f[1]]
            out.append("\t".join(f))
        elif f[0] == "L":
            a, b = (f[1], f[2]), (f[3], f[4])
            a = (a[0], "-" if (a[1] == "+") == (a[0] in flip) else "+") if a[0] in flip else a
            b = (b[0], "-" if (b[1] == "+") == (b[0] in flip) else "+") if b[0] in flip else b
            links.append((a, b))
        elif f[0] == "P":
            steps = [(x[:-1], x[-1]) for x in f[2].split(",")]
            steps = [(n, ("-" if o == "+" else "+") if n in flip else o) for n, o in steps]
            out.append("P\t%s\t%s\t*" % (f[1], ",".join(n + o for n, o in steps)))
        else:
            out.append(l)
    return out, links, flip


def test_graph_build_and_gfa(s, oracle, c1_golden):
    """Row f1: srgpu_graph_build (device: roots, first-seen numbering, steps, edge set) + compaction + GFA
    writer against the restated reference (oracle/oracle.cpp) on the GPU's own partition.
      * forward-only set: byte-identical GFA, --no-compact and compacted, with the library's own representatives;
      * set with reverse-complemented sequences: byte-identical once the builder is given the sequential uf_rush
        roots as representatives; with its own representatives the --no-compact graph is the same up to the
        orientation of mixed-component nodes;
      * BASELINE.json configs[0] (HLA DRB1-3123, 22 RC pairs): byte-identical to the committed reference GFA."""
    import _oracle
    from seqrush_b200 import synth
    from test_graph_host import oracle_partition
    for cfg, mixed in (((6, 3000, 0.02, 41, 0.0, 0.0), False), ((7, 2500, 0.03, 42, 0.45, 0.3), True)):
        ids, seqs = synth.make_sequences(*cfg)
        ctx = s.Context(0)
        sr = s.SeqRush([s.Sequence(i, a.tobytes()) for i, a in zip(ids, seqs)], ctx=ctx)
        sr.align_and_unite(s.Args(min_match_length=0))
        uf = oracle_partition(oracle, seqs)
        assert np.array_equal(sr.canonical_partition(), uf.canonical(2 * sr.total_length))
        raw, rep, nn = _oracle.build_gfa(uf, seqs, ids, no_compact=True)
        want, _, _ = _oracle.build_gfa(uf, seqs, ids, no_compact=False)
        g = sr.build_bidirected_graph(representatives=rep)
        assert g.stats()["uncompacted_nodes"] == nn and g.node_count == nn
        assert g.gfa_text() == raw
        assert g.compact().gfa_text() == want
        g.close()
        g = sr.build_bidirected_graph()  # the library's canonical representatives
        mine = g.gfa_text()
        if not mixed:
            assert mine == raw and g.compact().gfa_text() == want
        else:
            out, links, flip = _gfa_flip_normalise(mine, raw)
            assert flip, "the set has mixed-orientation components"
            mine_l = [l for l in mine.splitlines() if not l.startswith("L")]
            assert mine_l == out
            canon = lambda a, b: min((a, b), ((b[0], "-" if b[1] == "+" else "+"), (a[0], "-" if a[1] == "+" else "+")))
            mine_links = sorted(canon((f[1], f[2]), (f[3], f[4])) for f in (l.split("\t") for l in mine.splitlines()) if f[0] == "L")
            assert mine_links == sorted(canon(a, b) for a, b in links)
            first = {}
            for p in range(len(ids)):  # first step through every node is '+'
                for h in g.path(p):
                    first.setdefault(h.node_id, h.is_reverse)
            assert not any(first.values())
        g.close()
        ctx.close()
    # configs[0]
    import hashlib
    import os
    from conftest import GOLDEN
    exp, ids, seqs = c1_golden
    z = np.load(os.path.join(GOLDEN, "c1_gfa.npz"))
    ctx = s.Context(0)
    sr = s.SeqRush([s.Sequence(i, d) for i, d in zip(ids, seqs)], ctx=ctx)
    sr.align_and_unite(s.Args(min_match_length=0))
    g = sr.build_bidirected_graph(representatives=z["representatives"].astype(np.uint64))
    assert g.node_count == int(z["uncompacted_nodes"][0])
    assert hashlib.sha256(g.gfa_text().encode()).hexdigest() == bytes(z["gfa_no_compact_sha256"]).decode()
    g.compact()
    assert g.gfa_text() == bytes(z["gfa"]).decode()
    out = sr.write_bidirected_gfa(os.path.join(os.environ.get("TMPDIR", "/tmp"), "c1_own.gfa"))  # own representatives
    assert out.node_count > 0 and out.stats()["uncompacted_nodes"] == int(z["uncompacted_nodes"][0])
    out.close()
    g.close()
    ctx.close()


def test_alignment_modes_and_aligner_trait(s, oracle):
    """allwave's three modes (allwave/src/alignment.rs:203-231) and negative match scores through the C ABI against
    WFA2-lib itself (oracle/_ref, gap-affine / gap-affine-2p BiWFA with the same penalties): raw score and CIGAR;
    then the Aligner trait surface (src/aligner.rs:12-33,64-96; src/aligner/allwave_impl.rs:95-148)."""
    from seqrush_b200 import synth
    rng = np.random.default_rng(77)
    acgt = np.frombuffer(b"ACGT", dtype=np.uint8)
    seqs = []
    for L, d in ((90, 0.05), (700, 0.03), (3000, 0.02), (9000, 0.04), (2500, 0.10)):
        anc = acgt[rng.integers(0, 4, size=L)]
        seqs += [synth.mutate(rng, anc, d).tobytes(), synth.mutate(rng, anc, d).tobytes()]
    anc = acgt[rng.integers(0, 4, size=4000)]
    a = synth.mutate(rng, anc, 0.01)
    seqs += [a.tobytes(), np.concatenate([a[:1500], a[1900:]]).tobytes()]  # a 400-base gap
    pairs = [(i, i + 1) for i in range(0, len(seqs), 2)] + [(i + 1, i) for i in range(0, len(seqs), 2)]
    cases = [
        (s.AlignmentParams(0, 5, 8, 2, None, None), (0, 5, 8, 2, 0, 0), False),     # single-piece affine
        (s.AlignmentParams(0, 4, 6, 2, None, None), (0, 4, 6, 2, 0, 0), False),
        (s.AlignmentParams.edit_distance(), (0, 1, 1, 1, 0, 0), False),              # "edit distance" = affine 1,1,1
        (s.AlignmentParams(0, 3, 3, 3, None, None), (0, 3, 3, 3, 0, 0), False),
        (s.AlignmentParams(-1, 4, 6, 2, 24, 1), (-1, 4, 6, 2, 24, 1), True),         # match < 0: Eizenga's adjustment
        (s.AlignmentParams(-2, 4, 4, 2, 24, 1), (-2, 4, 4, 2, 24, 1), True),         # the trait's defaults with the sign WFA2-lib takes
        (s.AlignmentParams(-1, 3, 5, 1, None, None), (-1, 3, 5, 1, 0, 0), False),
    ]
    for params, pen6, two in cases:
        res = _align(s, seqs, pairs, params)
        for r in res:
            q = seqs[r.query_idx]
            if r.is_reverse:
                q = s.reverse_complement(q)
            st, score, cig = oracle.wfa_mode(q, seqs[r.target_idx], pen6, two)
            assert st == 0 and r.cigar_bytes == cig and r.score == score, (pen6, r.query_idx, r.target_idx, r.score, score)
    with pytest.raises(s.SrgpuError, match="Match score must be negative or zero"):
        _align(s, seqs[:2], [(0, 1)], s.AlignmentParams(2, 4, 4, 2, 24, 1))
    # the trait
    ctx = s.Context(0)
    al = s.AllwaveAligner.with_params(1, False, s.AlignmentParams(), ctx=ctx)
    recs = al.align_sequences([s.AlignmentSequence("s%d" % i, b) for i, b in enumerate(seqs[:6])])
    assert len(recs) == 36 and isinstance(s.create_aligner("AllWave", 1, False, None, ctx=ctx), s.AllwaveAligner)
    k = 0
    for i in range(6):
        for j in range(6):
            o = oracle.align_pair(seqs[i], seqs[j])
            r = recs[k]
            k += 1
            assert (r.query_name, r.target_name, r.query_len, r.target_len) == ("s%d" % i, "s%d" % j, len(seqs[i]), len(seqs[j]))
            assert r.strand == ("-" if o["is_reverse"] else "+") and r.cigar == s.cigar_bytes_to_string(o["cigar"])
            assert (r.query_start, r.query_end, r.target_start, r.target_end) == (0, len(seqs[i]), 0, len(seqs[j]))
    with pytest.raises(s.SrgpuError, match="Match score"):
        s.create_aligner("allwave", ctx=ctx).align_sequences([s.AlignmentSequence("a", seqs[0]), s.AlignmentSequence("b", seqs[1])])
    with pytest.raises(RuntimeError):
        s.create_aligner("sweepga")
    with pytest.raises(ValueError):
        s.create_aligner("minimap2")
    assert s.AllwaveAligner.get_prioritized_pairs([s.AlignmentSequence("a", seqs[0]), s.AlignmentSequence("b", seqs[1])], "none", ctx=ctx) == [(0, 0), (0, 1), (1, 0), (1, 1)]
    ctx.close()


def test_union_find_rejects_a_second_sequence_set(s):
    """A context holds one sequence table: a forest filled against one set must not be united against another
    that was uploaded into the same context in between (its Pos space differs); srgpu_uf_clear resets that."""
    rng = np.random.default_rng(5)
    def mk(n, L):
        return [s.Sequence("s%d" % i, bytes(rng.choice(np.frombuffer(b"ACGT", dtype=np.uint8), size=L))) for i in range(n)]
    ctx = s.Context(0)
    try:
        a = s.SeqRush(mk(3, 300), ctx=ctx)
        a.align_and_unite(s.Args(min_match_length=0))
        b = s.SeqRush(mk(3, 300), ctx=ctx)  # replaces the context's sequence table
        b.align_and_unite(s.Args(min_match_length=0))
        with pytest.raises(Exception, match="different sequence set"):
            a.align_and_unite(s.Args(min_match_length=0))
        a.union_find.clear()
        ctx.set_sequences([q.data for q in a.sequences])
        a.align_and_unite(s.Args(min_match_length=0))
    finally:
        ctx.close()


def test_strand_choice_once_per_unordered_pair(s, oracle):
    """The strand choice of (t, q) is copied from (q, t) only where that is exact -- both sequences over {A,C,G,T,N} --
    and computed otherwise: a list with both orders, repeats and self pairs over sequences in upper case, lower case,
    with N and with IUPAC bytes, and palindromic / reverse-complement-equal content where fwd == rev (forward wins).
    (The set also holds UNRELATED sequences of 1500 and 1200 bases: their wavefronts reach the corner diagonals, which
    once overflowed a full-width ring row of 2 * (1500 + 4) = 47 * 64 diagonals and failed the call.)"""
    rng = np.random.default_rng(23)
    acgt = np.frombuffer(b"ACGT", dtype=np.uint8)
    comp = {65: 84, 67: 71, 71: 67, 84: 65, 78: 78}
    def rc(a):
        return np.array([comp[int(x)] for x in a[::-1]], dtype=np.uint8)
    def mutate(a, rate):
        a = a.copy()
        idx = rng.choice(a.size, size=max(1, int(a.size * rate)), replace=False)
        a[idx] = rng.choice(acgt, size=idx.size)
        return a
    base = rng.choice(acgt, size=1500)
    half = rng.choice(acgt, size=600)
    pal = np.concatenate([half, rc(half)])              # equal to its own reverse complement: fwd == rev against itself
    seqs = [base, mutate(base, 0.02), rc(mutate(base, 0.03)), pal, mutate(pal, 0.01),
            mutate(base, 0.02).tobytes().lower(),   # lower case: RC is not an involution
            None, None]
    withn = mutate(base, 0.02); withn[100:110] = ord("N"); seqs[6] = withn
    iupac = mutate(base, 0.02); iupac[200] = ord("R"); iupac[900] = ord("Y"); seqs[7] = iupac
    seqs = [np.frombuffer(x, dtype=np.uint8) if isinstance(x, bytes) else x for x in seqs]
    n = len(seqs)
    pairs = np.array([(i, j) for i in range(n) for j in range(n)] + [(1, 0), (0, 1), (3, 4), (4, 3), (2, 2)], dtype=np.uint32)
    pairs = pairs[rng.permutation(pairs.shape[0])]
    res = _align(s, [a.tobytes() for a in seqs], pairs)
    ores, _ = oracle.run_pairs(seqs, pairs, threads=4)
    n_rev = 0
    for k, r in enumerate(res):
        assert (r.query_idx, r.target_idx) == (int(pairs[k, 0]), int(pairs[k, 1]))
        assert int(r.is_reverse) == ores[k][0] and r.score == ores[k][2] and len(r.cigar_bytes) == ores[k][3], (k, pairs[k])
        n_rev += int(r.is_reverse)
    assert n_rev > 0

