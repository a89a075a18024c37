"""Host-side mirror of the reference interfaces (no device needed)."""
import numpy as np
import pytest

import seqrush_b200 as s
from seqrush_b200 import pos
from seqrush_b200.distributed import shard_pairs


def test_pos_roundtrip():
    """src/pos.rs:89-160."""
    p = pos.make_pos(42, False)
    assert pos.offset(p) == 42 and not pos.is_rev(p) and pos.orientation_char(p) == "+"
    r = pos.make_pos(42, True)
    assert pos.offset(r) == 42 and pos.is_rev(r) and pos.orientation_char(r) == "-"
    assert pos.flip_orientation(p) == r and pos.flip_orientation(r) == p
    assert pos.incr_pos(p) == pos.make_pos(43, False) and pos.decr_pos(p) == pos.make_pos(41, False)
    assert pos.incr_pos(r) == pos.make_pos(41, True) and pos.decr_pos(r) == pos.make_pos(43, True)
    assert pos.incr_pos(pos.make_pos(0, True)) == pos.make_pos(0, True)
    assert pos.decr_pos(pos.make_pos(0, False)) == pos.make_pos(0, False)
    assert [pos.rc_base(b) for b in b"ACGTNacgtX"] == list(b"TGCANTGCAX")


def test_allwave_helpers():
    assert s.reverse_complement(b"ACGTNacgtRY") == b"NNACGTNACGT"  # allwave/src/alignment.rs:174-186
    assert s.cigar_bytes_to_string(b"MMMMXIIDMM") == "4=1X2D1I2="
    assert s.cigar_bytes_to_string(b"") == ""
    p = s.parse_scores("0,5,8,2,24,1")
    assert (p.match_score, p.mismatch_penalty, p.gap_open, p.gap_extend, p.gap2_open, p.gap2_extend) == (0, 5, 8, 2, 24, 1)
    assert s.parse_scores("0,5,8,2").gap2_open is None
    with pytest.raises(ValueError):
        s.parse_scores("1,2,3")
    # AlignmentMode::from_params (allwave/src/types.rs:101-111) and what reaches the C ABI for each mode
    assert s.parse_scores("0,5,8,2,24,1").mode() == "TwoPieceAffine" and s.parse_scores("0,5,8,2").mode() == "SinglePieceAffine"
    assert s.AlignmentParams.edit_distance().mode() == "EditDistance"
    c = s.parse_scores("0,5,8,2").to_c()
    assert (c.mismatch, c.gap_open1, c.gap_ext1, c.gap_open2, c.gap_ext2) == (5, 8, 2, -1, -1)
    c = s.AlignmentParams(0, 3, 3, 3, None, None).to_c()  # "edit distance" = affine x, x, x (allwave/src/alignment.rs:205-213)
    assert (c.mismatch, c.gap_open1, c.gap_ext1, c.gap_ext2) == (3, 3, 3, -1)


def test_paf_line():
    """allwave/src/lib.rs:71-112. Expected lines are written out from the reference's format string
    "{}\t{}\t{}\t{}\t{}\t{}\t{}\t{}\t{}\t{}\t{}\t{}\tgi:f:{:.6}\tcg:Z:{}" with
    block_len = max(query_end - query_start, target_end - target_start) and identity =
    num_matches / alignment_length (0.0 when alignment_length == 0), not from the repo's own reader."""
    seqs = [s.AllwaveSequence("q", b"ACGTACGT"), s.AllwaveSequence("t", b"ACGTTCGT"), s.AllwaveSequence("long", b"ACGTACGTACGT")]
    # forward, one mismatch: 7 matches over 8 aligned columns
    r = s.AlignmentResult(0, 1, 0, 8, 0, 8, False, b"MMMMXMMM", -5, 7, 8)
    assert s.alignment_to_paf(r, seqs) == "q\t8\t0\t8\t+\tt\t8\t0\t8\t7\t8\t60\tgi:f:0.875000\tcg:Z:4=1X3="
    # reverse strand with gaps (test_wfa2_cigar_debug's MMMMMMMMIIII: WFA 'I' consumes the target and is
    # written as 'D'): block_len = max(8, 12) = 12 although only 8 columns are M/X
    r = s.AlignmentResult(0, 2, 0, 8, 0, 12, True, b"MMMMMMMMIIII", -16, 8, 8)
    assert s.alignment_to_paf(r, seqs) == "q\t8\t0\t8\t-\tlong\t12\t0\t12\t8\t12\t60\tgi:f:1.000000\tcg:Z:8=4D"
    # block_len follows the longer side, here the query (WFA 'D' consumes the query -> 'I')
    r = s.AlignmentResult(2, 0, 0, 12, 0, 8, False, b"MMMMDDDDMMMM", -16, 8, 8)
    assert s.alignment_to_paf(r, seqs) == "long\t12\t0\t12\t+\tq\t8\t0\t8\t8\t12\t60\tgi:f:1.000000\tcg:Z:4=4I4="
    # failed alignment (allwave/src/alignment.rs:40-55): empty CIGAR, alignment_length 0 -> identity 0.0
    r = s.AlignmentResult(0, 1, 0, 0, 0, 0, False, b"", 2**31 - 1, 0, 0)
    assert s.alignment_to_paf(r, seqs) == "q\t8\t0\t0\t+\tt\t8\t0\t0\t0\t0\t60\tgi:f:0.000000\tcg:Z:"
    # identity is rounded to six decimals like Rust's {:.6}
    r = s.AlignmentResult(0, 1, 0, 8, 0, 8, False, b"MMXMMXMM", -10, 2, 3)
    assert "\tgi:f:0.666667\t" in s.alignment_to_paf(r, seqs)
    # what the writer emits is what the --paf reader takes back (src/seqrush.rs:529-560 looks for cg:Z: in any tag)
    from seqrush_b200.seqrush import parse_paf_line
    r = s.AlignmentResult(0, 1, 0, 8, 0, 8, True, b"MMMMXMMM", -5, 7, 8)
    assert parse_paf_line(s.alignment_to_paf(r, seqs) + "\n") == ("q", 0, 8, "-", "t", 0, 8, "4=1X3=")


def test_load_sequences(tmp_path):
    """src/seqrush.rs:1801-1837: id = first word, offsets = running sum, case preserved."""
    fa = tmp_path / "x.fa"
    fa.write_text(">s1 desc here\nACGT\nacgt\n>s2\nNNNN\n")
    seqs = s.load_sequences(str(fa))
    assert [(q.id, q.data, q.offset) for q in seqs] == [("s1", b"ACGTacgt", 0), ("s2", b"NNNN", 8)]


def test_pair_order_and_sharding():
    """allwave/src/iterator.rs:37-46: i-major ordered pairs, self pairs kept when exclude_self=false."""
    from bench import all_pairs
    pr = all_pairs(3)
    assert pr.tolist() == [[0, 0], [0, 1], [0, 2], [1, 0], [1, 1], [1, 2], [2, 0], [2, 1], [2, 2]]
    parts = [shard_pairs(pr, r, 2) for r in range(2)]
    assert sorted(map(tuple, np.concatenate(parts).tolist())) == sorted(map(tuple, pr.tolist()))
    assert abs(parts[0].shape[0] - parts[1].shape[0]) <= 1


def test_synth_is_deterministic():
    from seqrush_b200 import synth
    a = synth.make_config("c2", n=3, length=2000)[1]
    b = synth.make_config("c2", n=3, length=2000)[1]
    assert all(np.array_equal(x, y) for x, y in zip(a, b))
    assert all(set(np.unique(x).tolist()) <= set(b"ACGT") for x in a)


def test_bench_reference_arm_prints_one_json_line():
    """bench.py contract: exactly one JSON line on stdout (library banners must not land there)."""
    import json, os, subprocess, sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    if not os.path.exists(os.path.join(root, "oracle", "_ref", "libsrref_generic.so")):
        pytest.skip("oracle/_ref not built")
    r = subprocess.run([sys.executable, os.path.join(root, "bench.py"), "--impl", "reference", "--steps", "1", "--warmup", "0",
                        "--cpu-sample", "2", "--workload", "tiny"], stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True,
                       timeout=300)
    assert r.returncode == 0, r.stderr[-2000:]
    lines = [ln for ln in r.stdout.splitlines() if ln.strip()]
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["cpu_baseline"]["kind"] == "reference" and d["value"] > 0
    assert d["e2e"]["h2d_bytes_per_step"] == 0 and d["unit"] == "Mbp/s"


def test_parse_paf_line():
    """src/seqrush.rs:529-560: short lines skipped, cg:Z: looked up among the optional fields only."""
    from seqrush_b200.seqrush import parse_paf_line
    line = "q\t25\t2\t22\t-\tt\t30\t3\t23\t18\t20\t60\ttp:A:P\tcg:Z:10=1X9=\n"
    assert parse_paf_line(line) == ("q", 2, 22, "-", "t", 3, 23, "10=1X9=")
    assert parse_paf_line("\n") is None
    assert parse_paf_line("q\t25\t0\t25\t+\tt\t30\t0\t25\t25\t25\n") is None  # 11 fields
    assert parse_paf_line("q\t25\t0\t25\t+\tt\t30\t0\t25\t25\t25\t60\n")[-1] == ""  # no CIGAR tag
    with pytest.raises(ValueError):
        parse_paf_line("q\tx\t0\t25\t+\tt\t30\t0\t25\t25\t25\t60\n")


def test_shard_pairs_keeps_both_orders_on_one_rank():
    """seqrush_b200.distributed.shard_pairs: a partition of the list, both orders of an unordered pair on the same rank,
    the list's own order kept inside a rank, sizes within one unordered pair of each other."""
    import numpy as np
    from seqrush_b200.distributed import shard_pairs
    from seqrush_b200 import synth
    pairs = synth.all_pairs(37)
    rng = np.random.default_rng(3)
    pairs = np.ascontiguousarray(pairs[np.sort(rng.choice(pairs.shape[0], size=900, replace=False))])  # a sparsified list
    for world in (1, 2, 3, 8):
        seen, sizes = set(), []
        everything = set(map(tuple, pairs.tolist()))
        for rank in range(world):
            mine = shard_pairs(pairs, rank, world)
            got = list(map(tuple, mine.tolist()))
            assert got == [p for p in map(tuple, pairs.tolist()) if p in set(got)]  # order of the list
            st = set(got)
            assert not (st & seen)
            for (a, b) in st:
                if (b, a) in everything:
                    assert (b, a) in st
            seen |= st
            sizes.append(len(got))
        assert seen == everything
        assert max(sizes) - min(sizes) <= 3
    assert shard_pairs(pairs[:0], 0, 4).shape[0] == 0
