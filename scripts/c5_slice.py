"""BASELINE.json configs[4] shape on a slice (GPU box only): n x 200 kb at 5 % divergence, pair list from
-x tree:3,3,0.1 (MinHash sketches + k-nearest/k-farthest + hashed random pairs on the device), align + unite."""
import sys, time, json
sys.path.insert(0, ".")
import numpy as np
import seqrush_b200 as s
from seqrush_b200 import synth
n = int(sys.argv[1]) if len(sys.argv) > 1 else 16
ids, seqs = synth.make_config("c5", n=n)
ctx = s.Context(0)
sr = s.SeqRush([s.Sequence(i, a.tobytes()) for i, a in zip(ids, seqs)], ctx=ctx)
args = s.Args(sparsification="tree:3,3,0.1", min_match_length=0)
t0 = time.time(); pairs = sr.pair_list(args); t1 = time.time()
print("pair list from tree:3,3,0.1 over %d x 200 kb: %d pairs in %.3f s" % (n, pairs.shape[0], t1 - t0), flush=True)
t0 = time.time(); st = sr.align_and_unite(args, pairs=pairs); t1 = time.time()
out = {"workload": "c5 slice: %d x 200 kb, 5 %%, tree:3,3,0.1" % n, "pairs": int(pairs.shape[0]), "seconds": t1 - t0,
       "aligned_Mbp_per_s": st["aligned_bp"] / 1e6 / (t1 - t0), "wf_cells": int(st["wf_cells"]),
       "cells_per_s": st["wf_cells"] / (st["align_ms"] / 1e3), "align_ms": st["align_ms"], "orient_ms": st["orient_ms"],
       "union_ms": st["union_ms"], "retries": st["retries"], "failed_pairs": st["failed_pairs"],
       "components": sr.count_components(), "partition_checksum": "%016x" % sr.union_find.checksum()}
print(json.dumps(out), flush=True)
