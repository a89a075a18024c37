"""Sweep launch configurations of the align kernel on a slice of a workload (GPU box only)."""
import sys, time, json
sys.path.insert(0, ".")
import numpy as np, torch
import seqrush_b200 as s
from bench import WORKLOADS
from seqrush_b200 import synth

wl = sys.argv[1] if len(sys.argv) > 1 else "c2"
npairs = int(sys.argv[2]) if len(sys.argv) > 2 else 5000
configs = [tuple(int(x) for x in c.split("x")) for c in (sys.argv[3].split(",") if len(sys.argv) > 3 else ["256x2"])]
cfg_name, n_seq, L = WORKLOADS[wl][:3]
ids, seqs = synth.make_config(cfg_name, n=n_seq, length=L)
pairs = synth.all_pairs(len(seqs))
rng = np.random.default_rng(0)
pairs = pairs[np.sort(rng.choice(pairs.shape[0], size=min(npairs, pairs.shape[0]), replace=False))]
ctx = s.Context(0)
sr = s.SeqRush([s.Sequence(i, a.tobytes()) for i, a in zip(ids, seqs)], ctx=ctx)
d_pairs = torch.from_numpy(pairs.view(np.int32).copy()).cuda()
ref_ck = None
for cfg in configs:
    th, ct = cfg[0], cfg[1]
    ring = cfg[2] if len(cfg) > 2 else 0
    ctx.configure(threads=th, ctas_per_sm=ct, ring_diagonals=ring)
    for rep in range(2):
        sr.union_find.close()
        sr.union_find = s.BidirectedUnionFind(sr.total_length, ctx)
        sr.union_find.unite_fr_all(sr.total_length)
        t0 = time.time()
        st = sr.align_and_unite_device(d_pairs.data_ptr(), pairs.shape[0])
        dt = time.time() - t0
    ck = sr.union_find.checksum()
    ref_ck = ref_ck or ck
    print("cfg %s: %.3f s  align_ms %.1f orient_ms %.1f union_ms %.2f  launches %d retries %d  %.1f Mbp/s  cells/s %.3g  checksum %s"
          % (cfg, dt, st["align_ms"], st["orient_ms"], st["union_ms"], st["kernel_launches"], st["retries"],
             st["aligned_bp"] / 1e6 / dt, st["wf_cells"] / (st["align_ms"] / 1e3), "same" if ck == ref_ck else "DIFF"), flush=True)
