"""Run one rank's shard of a multi-GPU bench problem on a single GPU (debugging aid)."""
import sys
sys.path.insert(0, ".")
import numpy as np, torch
import seqrush_b200 as s
from bench import make_workload
from seqrush_b200.distributed import shard_pairs

wl, world, rank = sys.argv[1], int(sys.argv[2]), int(sys.argv[3])
scaling = sys.argv[4] if len(sys.argv) > 4 else "weak"
ids, seqs, pairs, cfg = make_workload(wl, world, scaling)
mine = shard_pairs(pairs, rank, world)
print("shard", mine.shape, mine[:6].tolist(), flush=True)
ctx = s.Context(0)
sr = s.SeqRush([s.Sequence(i, a.tobytes()) for i, a in zip(ids, seqs)], ctx=ctx)
d_pairs = torch.from_numpy(mine.view(np.int32).copy()).cuda()
for rep in range(2):
    st = sr.align_and_unite_device(d_pairs.data_ptr(), mine.shape[0])
    print("rep", rep, {k: st[k] for k in ("align_ms", "orient_ms", "retries", "failed_pairs", "kernel_launches")}, "%016x" % sr.union_find.checksum(), flush=True)
