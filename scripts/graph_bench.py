"""Time the graph build from the partition (row f1) on a bench workload (GPU box only)."""
import sys, time, hashlib
sys.path.insert(0, ".")
import numpy as np
import seqrush_b200 as s
from seqrush_b200 import synth
name = sys.argv[1] if len(sys.argv) > 1 else "c2"
ids, seqs, pairs, desc = synth.bench_set(name)
ctx = s.Context(0)
sr = s.SeqRush([s.Sequence(i, a.tobytes()) for i, a in zip(ids, seqs)], ctx=ctx)
t0 = time.time(); st = sr.align_and_unite(s.Args(min_match_length=0), pairs=pairs); t1 = time.time()
print("%s: align+unite %.2f s (%d pairs), %d components" % (name, t1 - t0, pairs.shape[0], sr.count_components()), flush=True)
t0 = time.time(); g = sr.build_bidirected_graph(); t1 = time.time()
print("graph build (device passes + download): %.3f s, %d nodes, %d edges, %d steps" % (t1 - t0, g.node_count, g.edge_count, sr.total_length), flush=True)
t0 = time.time(); g.compact(); t1 = time.time()
print("compact + renumber (host): %.3f s -> %d nodes, %d edges, %s" % (t1 - t0, g.node_count, g.edge_count, g.stats()), flush=True)
t0 = time.time(); txt = g.gfa_text(); t1 = time.time()
print("validate + GFA text: %.3f s, %d bytes, sha256 %s" % (t1 - t0, len(txt), hashlib.sha256(txt.encode()).hexdigest()), flush=True)
