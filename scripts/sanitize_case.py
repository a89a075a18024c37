"""A small workload that touches every kernel of the library, for compute-sanitizer (GPU box only):
  compute-sanitizer --tool memcheck|racecheck|initcheck python scripts/sanitize_case.py [big]
smoke set (5 x 4 kb with reverse-complemented sequences and inversions): orientation + BiWFA + backtrace, fused
union, CIGAR expansion, batched CIGAR walk, union-find entry points, graph build, sketches + pair filter;
`big` adds two 50 kb pairs (dual forward/reverse step with the parked reverse step, cp.async row staging,
bulk-copy sequence staging, deep recursion)."""
import sys
sys.path.insert(0, ".")
import numpy as np
import seqrush_b200 as s
from seqrush_b200 import synth, sparsify

big = len(sys.argv) > 1 and sys.argv[1] == "big"
ids, seqs = synth.make_sequences(5, 4000, 0.02, 7, 0.4, 0.4)
ctx = s.Context(0)
sr = s.SeqRush([s.Sequence(i, a.tobytes()) for i, a in zip(ids, seqs)], ctx=ctx)
st = sr.align_and_unite(s.Args(min_match_length=0))
print("fused:", st["pairs"], "pairs,", st["unite_ops"], "unite ops, components", sr.count_components(), flush=True)
aw = [s.AllwaveSequence(i, a.tobytes()) for i, a in zip(ids, seqs)]
it = s.AllPairIterator(aw, s.AlignmentParams(), False, False, pairs=sr.all_pairs()[:8], ctx=ctx)
it._uploaded = True
res = it.align_all()
sr2 = s.SeqRush([s.Sequence(i, a.tobytes()) for i, a in zip(ids, seqs)], ctx=ctx)
sr2.process_alignments([(s.cigar_bytes_to_string(r.cigar_bytes), r.query_idx, r.target_idx, r.is_reverse, 0, 0) for r in res], 5)
sr2.union_find.unite_matching_region(0, 4000, 10, 20, 50, False, 0)
sr2.union_find.unite_regions([(0, 8000, 5, 5, 100, True, 4000)])
print("paf walk + regions: checksum %016x" % sr2.union_find.checksum(), sr2.union_find.find(3), sr2.union_find.same(0, 1), flush=True)
g = sr.build_bidirected_graph()
g.compact()
print("graph:", g.node_count, "nodes", g.edge_count, "edges", len(g.gfa_text()), "bytes of GFA", flush=True)
print("pairs from tree:2,1,0.3,11:", len(sparsify.extract_tree_pairs(aw, 2, 1, 0.3, 11, ctx, uploaded=True)), flush=True)
sr.union_find.compress()
print("roots:", int(sr.canonical_partition().sum() % 1000003), flush=True)
if big:
    ids, seqs = synth.make_config("c2", n=3)
    srb = s.SeqRush([s.Sequence(i, a.tobytes()) for i, a in zip(ids, seqs)], ctx=ctx)
    stb = srb.align_and_unite(s.Args(), pairs=[(0, 1), (2, 1)])
    print("big:", stb["wf_cells"], "cells, checksum %016x" % srb.union_find.checksum(), flush=True)
ctx.close()
print("done", flush=True)
