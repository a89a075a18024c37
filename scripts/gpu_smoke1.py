import sys, time, numpy as np
sys.path.insert(0, "."); sys.path.insert(0, "tests")
import seqrush_b200 as s
from seqrush_b200 import synth
import _oracle
t0=time.time()
ids, seqs = synth.make_sequences(6, 3000, 0.02, 1, 0.3, 0.3)
orc = _oracle.Oracle()
sr = s.SeqRush([s.Sequence(i, a.tobytes()) for i, a in zip(ids, seqs)])
pairs = sr.all_pairs()
it = s.AllPairIterator([s.AllwaveSequence(i, a.tobytes()) for i, a in zip(ids, seqs)], s.AlignmentParams(), False, False, ctx=sr.ctx)
res = it.align_all()
print("aligned", len(res), "in", time.time()-t0)
bad = 0
for r in res:
    o = orc.align_pair(seqs[r.query_idx], seqs[r.target_idx])
    okc = (o["cigar"] == r.cigar_bytes); oks = (o["score"] == r.score); okr = (o["is_reverse"] == r.is_reverse)
    if not (okc and oks and okr):
        bad += 1
        print("MISMATCH", r.query_idx, r.target_idx, okc, oks, okr, o["score"], r.score, o["is_reverse"], r.is_reverse, len(o["cigar"]), len(r.cigar_bytes), min(o["fwd_ed"], o["rev_ed"]), r.edit_distance)
print("pairs", len(res), "bad", bad)
st = sr.align_and_unite(s.Args(min_match_length=0))
print(st)
ouf = _oracle.OracleUF(sr.total_length); ouf.unite_fr_all(sr.total_length)
_, ost = orc.run_pairs(seqs, pairs, uf=ouf, threads=4)
print(ost)
mine = sr.canonical_partition(); ref = ouf.canonical(2*sr.total_length)
print("partition equal:", bool((mine == ref).all()), "unite ops", st["unite_ops"], ost["unite_calls"], "components", sr.count_components(), ouf.count_components(sr.total_length))
