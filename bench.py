#!/usr/bin/env python
"""bench.py -- align+unite phase throughput on synthetic sequences (BASELINE.json metric).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--workload c2|c3|c4s] [--impl ours|reference]

A step = one pass of the hot path (strand choice + BiWFA + CIGAR walk + unite) over the whole
ordered pair list of the workload. N>1 is launched by torch.distributed.run, one rank per GPU:
pairs are dealt round-robin, each GPU unites into a private forest, and the step ends with the
NCCL all_gather + device merge of the forests. Rank 0 prints ONE JSON line.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

# workloads: name -> (synth config, n, length, description)
WORKLOADS = {
    "c2": ("c2", 100, 50_000, "synthetic 100 x 50 kb, 1% SNP/indel divergence per sequence, all 10^4 ordered pairs (BASELINE.json configs[1])"),
    "c3": ("c3", 200, 50_000, "synthetic 200 x 50 kb, 20% reverse-complemented + inversions, all ordered pairs (configs[2])"),
    "c4s": ("c4", 100, 100_000, "synthetic 100 x 100 kb slice of configs[3] (1000 x 100 kb), all ordered pairs"),
    "tiny": ("c2", 16, 20_000, "synthetic 16 x 20 kb (debug)"),
}
PEN = (5, 8, 2, 24, 1)  # seqrush -S 0,5,8,2,24,1 (src/seqrush.rs:45)


def load_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            d = json.load(open(p))
            return float(d.get("hbm_gbs", 6650.0)), float(d.get("sm_max_mhz", 1965.0)), "measured"
        except Exception:
            pass
    return 6650.0, 1965.0, "fallback"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md recipe)."""

    def __init__(self, index=0):
        self.index = index
        self.proc = None
        self.lines = []

    def start(self):
        q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
             "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
             "clocks_event_reasons.sw_power_cap")
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + q,
                                          "--format=csv,noheader,nounits", "-lms", "1000"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=3)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0]))
                mx.append(float(f[1]))
            except ValueError:
                continue
            for nm, v in zip(names, f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def make_workload(name):
    from seqrush_b200 import synth
    cfg, n, L, desc = WORKLOADS[name]
    ids, seqs = synth.make_config(cfg, n=n, length=L)
    return ids, seqs, desc


def all_pairs(n):
    import numpy as np
    ii, jj = np.meshgrid(np.arange(n, dtype=np.uint32), np.arange(n, dtype=np.uint32), indexing="ij")
    return np.ascontiguousarray(np.stack([ii.ravel(), jj.ravel()], axis=1), dtype=np.uint32)


def cpu_baseline(seqs, pairs, sample_pairs, threads, seed=0):
    """Reference CPU path (oracle/_ref = the reference's WFA2-lib + edlib compiled from its own
    sources, plus the restated Rust glue) on a bounded, fixed-seed sample of the same pair list."""
    import numpy as np
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import _oracle
    rng = np.random.default_rng(seed)
    # sample among non-self pairs in proportion (self pairs are trivial); keep the n:n^2 ratio
    idx = rng.choice(pairs.shape[0], size=min(sample_pairs, pairs.shape[0]), replace=False)
    sub = pairs[np.sort(idx)]
    orc = _oracle.Oracle()
    tot = int(sum(int(s.size) for s in seqs))
    uf = _oracle.OracleUF(tot)
    uf.unite_fr_all(tot)
    _, st = orc.run_pairs(seqs, sub, pen=PEN, min_match_len=0, threads=threads, uf=uf, want_results=False)
    return st, sub.shape[0]


def run_reference(args):
    """--impl reference: the reference's own CPU implementation of the path on the host cores."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    ids, seqs, desc = make_workload(args.workload)
    pairs = all_pairs(len(seqs))
    cores = os.cpu_count() or 1
    sample = args.cpu_sample or max(cores, 64)
    for _ in range(args.warmup and 1):
        cpu_baseline(seqs, pairs, max(cores // 2, 4), cores, seed=99)
    t0 = time.time()
    bp = ops = 0
    secs = 0.0
    for k in range(args.steps):
        st, ns = cpu_baseline(seqs, pairs, sample, cores, seed=k)
        bp += st["aligned_bp"]
        ops += st["unite_calls"]
        secs += st["seconds"]
    val = bp / 1e6 / secs
    out = {
        "impl": "reference", "metric": "aligned_Mbp_per_s", "value": val, "unit": "Mbp/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * secs / args.steps, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "int32", "data": "synthetic",
        "unite_ops_per_s": ops / secs,
        "config": {"workload": desc, "penalties": "0,5,8,2,24,1", "min_match_len": 0},
        "cpu_baseline": {"value": val, "unit": "Mbp/s", "cores": cores, "kind": "reference",
                         "sample": "%d random ordered pairs of the workload per step (fixed seeds), oracle/_ref = "
                                   "reference WFA2-lib + edlib compiled from its sources + restated Rust glue" % sample},
        "e2e": {"value": val, "unit": "Mbp/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    emit(out)


_JSON_OUT = None


def emit(obj):
    """The one JSON line of the contract, on the process's original stdout."""
    f = _JSON_OUT or sys.stdout
    f.write(json.dumps(obj) + "\n")
    f.flush()


def main():
    # stdout carries exactly one JSON line: anything else a library prints there (the NCCL version
    # banner, device printf) is sent to stderr by pointing fd 1 at fd 2 and keeping the original
    global _JSON_OUT
    sys.stdout.flush()
    _JSON_OUT = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--workload", default="c2", choices=sorted(WORKLOADS))
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--cpu-sample", type=int, default=0)
    ap.add_argument("--threads", type=int, default=0)
    ap.add_argument("--ctas-per-sm", type=int, default=0)
    ap.add_argument("--ring", type=int, default=-1)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference(args)

    import numpy as np
    import torch
    import seqrush_b200 as s

    def trace(msg):
        if os.environ.get("BENCH_TRACE"):
            print("[bench rank %s] %s" % (os.environ.get("RANK", "0"), msg), file=sys.stderr, flush=True)

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (no CPU fallback)")
    torch.cuda.set_device(local)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
        trace("process group up")

    ids, seqs, desc = make_workload(args.workload)
    nseq = len(seqs)
    pairs_all = all_pairs(nseq)
    from seqrush_b200.distributed import shard_pairs, merge_device_union_finds
    # weak scaling: every rank runs the full pair list of its own replica of the workload?  No --
    # the path shards: the ordered pair list is dealt round-robin, per-GPU work = pairs/N. To keep
    # per-GPU work fixed as N grows ("weak"), the pair list is replicated N times (N passes of the
    # workload), so each rank always aligns one full workload's worth of pairs.
    job_pairs = np.concatenate([pairs_all] * world, axis=0) if world > 1 else pairs_all
    my_pairs = shard_pairs(job_pairs, rank, world)

    stream = torch.cuda.Stream()
    ctx = s.Context(local, stream.cuda_stream)
    if args.threads or args.ctas_per_sm or args.ring >= 0:
        ctx.configure(threads=args.threads, ctas_per_sm=args.ctas_per_sm, ring_diagonals=args.ring)
    sr = s.SeqRush([s.Sequence(i, a.tobytes()) for i, a in zip(ids, seqs)], ctx=ctx)
    total_len = sr.total_length
    d_pairs = torch.from_numpy(my_pairs.view(np.int32).copy()).to("cuda:%d" % local)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda:%d" % local)
    sargs = s.Args(min_match_length=0)

    def reset_uf():
        sr.union_find.clear()  # UFRush::clear, then SeqRush::new's unite(2i, 2i+1) loop (src/seqrush.rs:324-328)
        sr.union_find.unite_fr_all(total_len)

    def step_device():
        reset_uf()
        trace("union-find reset")
        st = sr.align_and_unite_device(d_pairs.data_ptr(), my_pairs.shape[0], sargs)
        trace("align+unite done: %.0f ms align, %d launches, %d retries" % (st["align_ms"], st["kernel_launches"], st["retries"]))
        if dist is not None:
            merge_device_union_finds(sr.union_find, dist)
            trace("forests merged")
        return st

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- device-resident timing (value) ----
    trace("sequences resident, %d pairs on this rank" % my_pairs.shape[0])
    for _ in range(args.warmup):
        step_device()
        trace("warm-up step done")
        flush.fill_(1)
    barrier()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    agg = {"align_ms": 0.0, "union_ms": 0.0, "orient_ms": 0.0, "wf_cells": 0, "orient_cells": 0, "unite_ops": 0, "unite_links": 0,
           "launches": 0, "aligned_bp": 0, "align_launches": 0, "union_launches": 0}
    step_ms = []
    for k in range(args.steps):
        flush.fill_(k & 255)  # L2 flush between timed iterations (256 MB > 126 MB L2); outside the timed events
        barrier()
        e0.record(stream)
        st = step_device()
        e1.record(stream)
        barrier()
        step_ms.append(e0.elapsed_time(e1))
        trace("timed step %d: %.1f ms" % (k, step_ms[-1]))
        for kk in ("align_ms", "union_ms", "orient_ms", "wf_cells", "orient_cells", "unite_ops", "unite_links", "aligned_bp",
                   "align_launches", "union_launches"):
            agg[kk] += st[kk]
        agg["launches"] += st["kernel_launches"] + 3  # + uf init, unite_fr_all, (compress/merge counted below)
    clocks = sampler.stop() if rank == 0 else None
    t = torch.tensor([sum(step_ms)], dtype=torch.float64, device="cuda:%d" % local)
    if dist is not None:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    total_ms = float(t.item())
    checksum = sr.union_find.checksum()
    comps = sr.count_components()

    # ---- end to end through the public API with host buffers (e2e) ----
    host_seqs = [s.Sequence(i, a.tobytes()) for i, a in zip(ids, seqs)]
    h2d = int(sum(a.size for a in seqs)) + int(my_pairs.nbytes)
    d2h = 2 * total_len * 8
    e2e_ms = []
    for k in range(max(1, min(args.steps, 2))):
        flush.fill_(k & 255)
        barrier()
        e0.record(stream)
        sr2 = s.SeqRush(host_seqs, ctx=ctx)           # H2D of the sequences + union-find init
        sr2.align_and_unite(sargs, pairs=my_pairs)      # H2D of the pair list, kernels
        if dist is not None:
            merge_device_union_finds(sr2.union_find, dist)
        part = sr2.canonical_partition()                # D2H of the canonical partition
        e1.record(stream)
        barrier()
        e2e_ms.append(e0.elapsed_time(e1))
        trace("e2e step %d: %.1f ms" % (k, e2e_ms[-1]))
        sr2.union_find.close()
    t = torch.tensor([sum(e2e_ms) / len(e2e_ms)], dtype=torch.float64, device="cuda:%d" % local)
    if dist is not None:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    e2e_step_ms = float(t.item())

    # totals over ranks
    tot = torch.tensor([agg["aligned_bp"], agg["unite_ops"], agg["wf_cells"], agg["orient_cells"], agg["unite_links"]],
                       dtype=torch.float64, device="cuda:%d" % local)
    if dist is not None:
        dist.all_reduce(tot)
    bp, ops, cells, ocells, links = [float(x) for x in tot.tolist()]
    if rank != 0:
        if dist is not None:
            dist.destroy_process_group()
        return
    hbm_peak, sm_max_mhz, peak_kind = load_peaks()
    secs = total_ms / 1e3
    value = bp / 1e6 / secs
    job_bp_step = bp / args.steps
    # roofline of the dominant kernel (align_kernel), rank 0's launches, device time from CUDA events
    # recorded by the library on its launch stream (srgpu_stats.align_ms)
    al_s = agg["align_ms"] / 1e3
    r0_cells, r0_ocells = agg["wf_cells"], agg["orient_cells"]
    n_launch = max(1, agg["align_launches"])
    BYTES_PER_CELL = 48  # 7 source offsets read + 5 offsets written, int32 (DESIGN.md "roofline")
    # measured DRAM traffic of the align kernel per wavefront cell: dram__bytes_read.sum + dram__bytes_write.sum
    # = 855.5 GB over the 2.052e10 cells of a 600-pair c2 launch (ncu --set full, profiles/r01j_ncu_align_kernel_c2.txt);
    # scaled by this launch's cell count because the full launch is too long to replay under ncu
    DRAM_BYTES_PER_CELL_NCU = 855.5e9 / 2.052e10
    alg_bytes = BYTES_PER_CELL * r0_cells  # the strand choice runs in its own (small) launch: orient_kernel
    INT_OPS_PER_CELL = 17 + 6  # 17 int ops per cell + 6 per extend word (>= 1 word per cell)
    int_peak = 148 * 4 * 32 * sm_max_mhz * 1e6
    un_s = max(agg["union_ms"] / 1e3, 1e-9)
    r0_ops, r0_links = agg["unite_ops"], agg["unite_links"]
    out = {
        "metric": "aligned_Mbp_per_s", "value": value, "unit": "Mbp/s", "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": total_ms / args.steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "int32", "data": "synthetic",
        "unite_ops_per_s": ops / secs,
        "config": {"workload": desc, "penalties": "0,5,8,2,24,1", "min_match_len": 0,
                   "pairs_per_step": int(job_pairs.shape[0]), "aligned_Mbp_per_step": job_bp_step / 1e6,
                   "sharding": "pairs round-robin over ranks, forests merged by all_gather + device union" if world > 1 else "single GPU",
                   "l2": "256 MB flush write between timed steps; per-CTA wavefront workspaces >> 126 MB L2"},
        "e2e": {"value": job_bp_step / 1e6 / (e2e_step_ms / 1e3), "unit": "Mbp/s", "h2d_bytes_per_step": h2d,
                "d2h_bytes_per_step": d2h, "ms_per_step": e2e_step_ms},
        "gpu_launches": int(agg["launches"]),
        "clocks": clocks,
        "roofline": {"kernel": "align_kernel (orientation + BiWFA + backtrace)", "bound": "hbm",
                     "achieved": alg_bytes / al_s / 1e9, "peak": hbm_peak, "unit": "GB/s",
                     "frac": alg_bytes / al_s / 1e9 / hbm_peak,
                     "traffic": DRAM_BYTES_PER_CELL_NCU * r0_cells / n_launch if args.workload == "c2" else None,
                     "traffic_unit": "bytes per launch (ncu-measured bytes per cell x cells of this launch)",
                     "peak_kind": peak_kind,
                     "algorithmic_bytes_per_launch": alg_bytes / n_launch, "avg_launch_ms": agg["align_ms"] / n_launch,
                     "share_of_step": agg["align_ms"] / sum(step_ms),
                     "orient_kernel_share_of_step": agg["orient_ms"] / sum(step_ms)},
        "roofline_int_issue": {"kernel": "align_kernel", "bound": "int_issue",
                               "achieved": INT_OPS_PER_CELL * r0_cells / al_s / 1e9,
                               "peak": int_peak / 1e9, "unit": "Gop/s",
                               "frac": INT_OPS_PER_CELL * r0_cells / al_s / int_peak,
                               "cells_per_s": r0_cells / al_s},
        "roofline_union": {"kernel": "union_kernel", "bound": "hbm",
                           "achieved": (r0_ops * 8 + r0_links * 4) / un_s / 1e9, "peak": hbm_peak, "unit": "GB/s",
                           "frac": (r0_ops * 8 + r0_links * 4) / un_s / 1e9 / hbm_peak,
                           "unite_ops_per_s_kernel": r0_ops / un_s, "avg_launch_ms": agg["union_ms"] / max(1, agg["union_launches"])},
        "parity": {"partition_checksum": "%016x" % checksum, "components": int(comps)},
    }
    if not args.no_cpu_baseline and world == 1:
        cores = os.cpu_count() or 1
        sample = args.cpu_sample or max(cores, 64)
        st, ns = cpu_baseline(seqs, pairs_all, sample, cores)
        out["cpu_baseline"] = {"value": st["aligned_bp"] / 1e6 / st["seconds"], "unit": "Mbp/s", "cores": cores,
                               "kind": "reference",
                               "sample": "%d random ordered pairs of the workload, oracle/_ref (reference WFA2-lib + "
                                         "edlib compiled from its sources) + restated Rust glue, %d threads" % (ns, cores),
                               "unite_ops_per_s": st["unite_calls"] / st["seconds"]}
    emit(out)
    if dist is not None:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
