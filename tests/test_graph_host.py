"""Graph build, compaction and GFA text (row f1) on the CPU: the library's compaction / renumbering /
validation / writer (native host code, seqrush_b200/csrc/graph_host.hpp) against the literal restatement of
the reference's BidirectedGraph::compact / merge_component_v2 / write_gfa (oracle/oracle.cpp, quadratic like the
reference) on the same uncompacted graphs, plus the reference's known answers for the builder."""
import numpy as np
import pytest

import _oracle
import seqrush_b200 as s
from seqrush_b200 import synth


def parse_gfa(txt):
    nodes, edges, paths = {}, [], []
    for line in txt.splitlines():
        f = line.split("\t")
        if f[0] == "S":
            nodes[int(f[1])] = f[2]
        elif f[0] == "L":
            edges.append(((int(f[1]) << 1) | (f[2] == "-"), (int(f[3]) << 1) | (f[4] == "-")))
        elif f[0] == "P":
            paths.append((f[1], [(int(x[:-1]) << 1) | (x[-1] == "-") for x in f[2].split(",")]))
    return nodes, edges, paths


def oracle_partition(oracle, seqs, pairs=None, k=0):
    tot = int(sum(a.size for a in seqs))
    uf = _oracle.OracleUF(tot)
    uf.unite_fr_all(tot)
    oracle.run_pairs(seqs, synth.all_pairs(len(seqs)) if pairs is None else pairs, min_match_len=k, threads=1, uf=uf,
                     want_results=False)
    return uf


def test_builder_known_answers(oracle):
    """src/lib.rs:571-604 (five identical 150 bp sequences -> exactly 150 nodes, identical paths) and
    :339-354 (tandem duplications: path lengths 100 / 110 / 120 / 115 with single-base nodes), --no-compact."""
    rng = np.random.default_rng(777)
    acgt = np.frombuffer(b"ACGT", dtype=np.uint8)
    base = acgt[rng.integers(0, 4, size=150)]
    seqs = [base.copy() for _ in range(5)]
    ids = ["seq%d" % (i + 1) for i in range(5)]
    uf = oracle_partition(oracle, seqs)
    txt, _, nn = _oracle.build_gfa(uf, seqs, ids, no_compact=True)
    nodes, edges, paths = parse_gfa(txt)
    assert nn == 150 and len(nodes) == 150 and len(paths) == 5
    assert all(p[1] == paths[0][1] and len(p[1]) == 150 for p in paths)
    assert paths[0][1] == [(i + 1) << 1 for i in range(150)]  # 1+,2+,...: first-seen numbering, tests/bidirected_tests.rs:38-89

    def dup(a, pos, ln, times):  # add_tandem_duplication: the segment repeated `times` times in total
        return np.concatenate([a[:pos]] + [a[pos:pos + ln]] * times + [a[pos + ln:]])
    b = acgt[np.random.default_rng(999).integers(0, 4, size=100)]
    seqs = [b, dup(b, 40, 10, 2), dup(b, 40, 10, 3), dup(b, 60, 5, 4)]
    uf = oracle_partition(oracle, seqs)
    txt, _, _ = _oracle.build_gfa(uf, seqs, ["seq1", "seq2", "seq3", "seq4"], no_compact=True)
    assert [len(p[1]) for p in parse_gfa(txt)[2]] == [100, 110, 120, 115]


@pytest.mark.parametrize("cfg", [(5, 300, 0.03, 1, 0.0, 0.0), (6, 400, 0.05, 2, 0.4, 0.4), (4, 1200, 0.02, 3, 0.5, 0.0),
                                 (8, 150, 0.10, 4, 0.3, 0.0), (3, 2000, 0.01, 5, 0.0, 0.0), (7, 250, 0.06, 6, 0.5, 0.5)])
@pytest.mark.parametrize("k", [0, 8])
def test_compaction_matches_restated_reference(built, oracle, cfg, k):
    """The uncompacted graph of the restated builder goes through the library's compact + renumber + validate +
    write; the text must equal the restated reference's compacted GFA byte for byte (L lines sorted on both sides).
    Covers forward-only sets and sets with reverse-complemented sequences / inversions (mixed-orientation nodes)."""
    ids, seqs = synth.make_sequences(*cfg)
    uf = oracle_partition(oracle, seqs, k=k)
    raw, _, nn = _oracle.build_gfa(uf, seqs, ids, no_compact=True)
    want, _, _ = _oracle.build_gfa(uf, seqs, ids, no_compact=False)
    nodes, edges, paths = parse_gfa(raw)
    assert sorted(nodes) == list(range(1, nn + 1))
    g = s.BidirectedGraph.from_parts(bytes(ord(nodes[i + 1]) for i in range(nn)), [p[1] for p in paths], edges,
                                     [a.tobytes() for a in seqs], ids)
    assert g.gfa_text() == raw          # writer + validation on the uncompacted graph
    g.compact()
    assert g.gfa_text() == want
    st = g.stats()
    assert st["uncompacted_nodes"] == nn and g.node_count == sum(1 for l in want.splitlines() if l.startswith("S"))
    g.close()


def test_validation_rejects_a_wrong_path(built):
    g = s.BidirectedGraph.from_parts(b"ACG", [[2, 4, 6]], [(2, 4), (4, 6)], [b"ACT"], ["x"])
    with pytest.raises(s.SrgpuError, match="does not match original sequence"):
        g.gfa_text()
    g.close()
    g = s.BidirectedGraph.from_parts(b"ACG", [[2, 4, 7]], [(2, 4), (4, 7)], [b"ACC"], ["x"])  # 3- spells C
    assert g.gfa_text().splitlines()[-1] == "P\tx\t1+,2+,3-\t*"
    g.close()
