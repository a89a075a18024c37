"""SeqRush pipeline driver for the `--aligner allwave` align-and-unite phase -- host-side mirror of
src/seqrush.rs (Sequence :272-277, load_sequences :1801-1837, SeqRush::new :308-336,
align_and_unite_with_allwave :611-757, count_components :341-353). Graph construction, layout and
GFA output are outside this hot path."""
import ctypes as C
from dataclasses import dataclass
from typing import List, Optional

import numpy as np

from . import _lib
from .allwave import AlignmentParams, parse_scores
from .context import Context, default_context
from .union_find import BidirectedUnionFind


@dataclass
class Sequence:  # src/seqrush.rs:272-277
    id: str
    data: bytes
    offset: int = 0


@dataclass
class Args:  # the subset of src/seqrush.rs:17-152 that reaches the hot path
    scores: str = "0,5,8,2,24,1"        # -S, :45
    min_match_length: int = 0           # -k
    aligner: str = "allwave"
    sparsification: str = "none"
    threads: int = 1                    # kept for interface parity; the device schedules pairs itself
    verbose: bool = False
    output_alignments: Optional[str] = None  # --output-alignments, :693-716
    paf: Optional[str] = None                # --paf, :463-466


def load_sequences(path: str) -> List[Sequence]:
    """FASTA reader with the reference's conventions (src/seqrush.rs:1801-1837): id = first word of
    the header, offset = running sum of lengths, sequence bytes kept as they are (case preserved)."""
    seqs, cur_id, cur, off = [], None, [], 0
    with open(path, "rb") as f:
        for line in f:
            line = line.rstrip(b"\r\n")
            if line.startswith(b">"):
                if cur_id is not None:
                    data = b"".join(cur)
                    seqs.append(Sequence(cur_id, data, off))
                    off += len(data)
                hdr = line[1:].decode("utf-8", "replace")
                cur_id = hdr.split()[0] if hdr.split() else ""
                cur = []
            elif cur_id is not None:
                cur.append(line.strip())
    if cur_id is not None:
        data = b"".join(cur)
        seqs.append(Sequence(cur_id, data, off))
    return seqs


def parse_paf_line(line: str):
    """One PAF record as src/seqrush.rs:529-560 reads it: returns (query_name, query_start, query_end,
    strand, target_name, target_start, target_end, cigar) or None for an empty line or one with fewer
    than 12 fields (the reference warns and skips). A record without a cg:Z: tag yields cigar ""
    (:552-555). Non-numeric coordinates raise ValueError where the reference's unwrap() panics."""
    line = line.rstrip("\r\n")
    if not line:
        return None
    f = line.split("\t")
    if len(f) < 12:
        return None
    int(f[1]), int(f[6])  # parsed and dropped by the reference too (_query_len, _target_len)
    cigar = ""
    for tag in f[12:]:
        if tag.startswith("cg:Z:"):
            cigar = tag[5:]
            break
    return f[0], int(f[2]), int(f[3]), f[4], f[5], int(f[7]), int(f[8]), cigar


class SeqRush:
    def __init__(self, sequences: List[Sequence], ctx: Optional[Context] = None):  # ::new, :308-336
        for s in sequences:
            if len(s.data) == 0:
                raise ValueError("Empty sequences are not allowed: sequence '%s' has length 0" % s.id)
        off = 0
        for s in sequences:  # offsets are the running sum of lengths (:1816-1818)
            s.offset = off
            off += len(s.data)
        self.sequences = sequences
        self.total_length = off
        self.ctx = ctx or default_context()
        self.ctx.set_sequences([s.data for s in sequences])
        self.union_find = BidirectedUnionFind(self.total_length, self.ctx)
        self.union_find.unite_fr_all(self.total_length)  # :324-328
        self.last_results = None

    @classmethod
    def new(cls, sequences, ctx=None):
        return cls(sequences, ctx)

    def all_pairs(self) -> np.ndarray:
        """n^2 ordered pairs including self pairs, i-major (allwave/src/iterator.rs:37-46 with
        exclude_self=false, src/seqrush.rs:728-735)."""
        n = len(self.sequences)
        ii, jj = np.meshgrid(np.arange(n, dtype=np.uint32), np.arange(n, dtype=np.uint32), indexing="ij")
        return np.ascontiguousarray(np.stack([ii.ravel(), jj.ravel()], axis=1), dtype=np.uint32)

    def pair_list(self, args: Args = None) -> np.ndarray:
        """The pairs align_and_unite_with_allwave hands to AllPairIterator::with_options (:664-675, :728-735):
        the n^2 i-major list including self pairs, thinned by -x/--sparsify (parse_sparsification, :356-431;
        a string the reference cannot parse falls back to 'none' with a message, :668-674)."""
        import sys
        from .allwave import Sequence as AwSequence
        from .sparsify import parse_sparsification, sparsified_pairs, SparsificationStrategy
        args = args or Args()
        spec = args.sparsification if args.sparsification not in (None, "None") else "none"
        try:
            strat = parse_sparsification(spec)
        except ValueError as e:
            print("Error parsing sparsification: %s. Using 'none'" % e, file=sys.stderr)
            strat = SparsificationStrategy.none()
        if strat.kind == "none":
            return self.all_pairs()
        aw = [AwSequence(s.id, s.data) for s in self.sequences]
        return sparsified_pairs(aw, strat, False, self.ctx, uploaded=True)  # the context holds these sequences

    def tree_pairs_separated(self, args: Args = None):
        """extract_tree_pairs_separated for --iterative (:900-935): (tree pairs, random pairs)."""
        from .allwave import Sequence as AwSequence
        from .sparsify import parse_sparsification, extract_tree_pairs_separated
        import sys
        try:
            strat = parse_sparsification((args or Args()).sparsification)
        except ValueError as e:
            print("Error parsing sparsification: %s. Using default tree:3,3,0.1,16" % e, file=sys.stderr)
            strat = None
        if strat is None or strat.kind != "tree":  # :925-935
            if strat is not None:
                print("Note: Iterative mode works best with tree sampling. Using default tree:3,3,0.1,16", file=sys.stderr)
            kn, kf, rf, ks = 3, 3, 0.1, 16
        else:
            kn, kf, rf, ks = strat.k_nearest, strat.k_farthest, strat.random_fraction, strat.kmer_size or 16
        aw = [AwSequence(s.id, s.data) for s in self.sequences]
        return extract_tree_pairs_separated(aw, kn, kf, rf, ks, self.ctx, uploaded=True)

    def align_and_unite(self, args: Args = None, pairs=None, want_results=False):  # :460-508 -> :611-757
        args = args or Args()
        if args.paf:  # :463-466: a PAF file replaces the alignment phase
            return self.align_and_unite_from_paf(args.paf, args)
        if args.aligner != "allwave":
            raise SystemExit("only --aligner allwave is implemented on the device")  # :504 exits(1)
        if args.output_alignments:
            return self._align_write_paf_and_unite(args, pairs)
        if pairs is None:
            pairs = self.pair_list(args)
        pairs = np.ascontiguousarray(np.asarray(pairs, dtype=np.uint32).reshape(-1, 2))
        pen = parse_scores(args.scores).to_c()
        n = pairs.shape[0]
        res = (_lib.Result * n)() if want_results else None
        _lib.check(_lib.lib().srgpu_align_and_unite(self.ctx._h, self.union_find._h, pairs.ctypes.data, n,
                                                    C.byref(pen), args.min_match_length, res))
        self.last_results = res
        return self.ctx.stats()

    def align_and_unite_device(self, d_pairs_ptr: int, n_pairs: int, args: Args = None):
        """Same with the pair list already in HBM (used by bench.py's device-resident timing)."""
        args = args or Args()
        pen = parse_scores(args.scores).to_c()
        _lib.check(_lib.lib().srgpu_align_and_unite_device(self.ctx._h, self.union_find._h, d_pairs_ptr, n_pairs,
                                                           C.byref(pen), args.min_match_length))
        return self.ctx.stats()

    def process_alignment(self, cigar: str, seq1_idx: int, seq2_idx: int, min_match_len: int, query_is_rc: bool,
                          query_start: int = 0, target_start: int = 0):  # :1134-1481
        _lib.check(_lib.lib().srgpu_process_alignment(self.ctx._h, self.union_find._h, cigar.encode(), seq1_idx,
                                                      seq2_idx, min_match_len, int(bool(query_is_rc)), query_start,
                                                      target_start))

    def process_alignments(self, records, min_match_len: int):
        """Batched process_alignment: records = iterable of (cigar, seq1_idx, seq2_idx, query_is_rc,
        query_start, target_start); one multi-CTA launch walks them all (srgpu_process_alignments)."""
        records = list(records)
        arr = (_lib.PafRecord * len(records))()
        for i, (cigar, qi, ti, rc, qs, ts) in enumerate(records):
            arr[i] = _lib.PafRecord(qi, ti, int(bool(rc)), 0, qs, ts, cigar.encode())
        _lib.check(_lib.lib().srgpu_process_alignments(self.ctx._h, self.union_find._h, arr, len(records), min_match_len))

    def align_and_unite_from_paf(self, paf_path: str, args: Args = None, batch: int = 65536) -> int:  # :510-609
        """Unite from the alignments of a PAF file (ids looked up by name, unknown names and short lines
        skipped with a warning, strand "-" = query reverse-complemented); the CIGARs are walked on the
        device, `batch` records per launch. Returns the number of alignments processed."""
        import sys
        args = args or Args()
        idx = {s.id: i for i, s in enumerate(self.sequences)}
        n, pending = 0, []
        with open(paf_path, "r") as f:
            for line in f:
                rec = parse_paf_line(line)
                if rec is None:
                    if line.strip():
                        print("Warning: Invalid PAF line (not enough fields): %s" % line.rstrip(), file=sys.stderr)
                    continue
                qn, qs, _qe, strand, tn, ts, _te, cigar = rec
                if qn not in idx or tn not in idx:
                    print("Warning: Unknown sequence name(s) in PAF: %s or %s" % (qn, tn), file=sys.stderr)
                    continue
                pending.append((cigar, idx[qn], idx[tn], strand == "-", qs, ts))
                n += 1
                if len(pending) >= batch:
                    self.process_alignments(pending, args.min_match_length)
                    pending = []
        if pending:
            self.process_alignments(pending, args.min_match_length)
        return n

    def _align_write_paf_and_unite(self, args: Args, pairs=None, batch: int = 1024):
        """--output-alignments (:676-716). The reference aligns every pair twice (once for the PAF, once
        for the graph); here ONE pass over the pair list yields the CIGARs, which go to the file while
        the same pass unites on the device from the run-length CIGARs (srgpu_align_pairs_and_unite).
        Records are written in pair order."""
        from .allwave import AllPairIterator, Sequence as AwSequence, alignment_to_paf
        aw = [AwSequence(s.id, s.data) for s in self.sequences]
        it = AllPairIterator.with_options(aw, parse_scores(args.scores), False, False, None,
                                          pairs=self.all_pairs() if pairs is None else pairs, ctx=self.ctx)
        it._uploaded = True  # the context already holds these sequences (SeqRush.__init__)
        with open(args.output_alignments, "w") as f:
            for rs in it._batches(True, batch, unite_into=self.union_find, min_match_len=args.min_match_length):
                for r in rs:
                    f.write(alignment_to_paf(r, aw) + "\n")
        return self.ctx.stats()

    STABILITY_THRESHOLD = 10  # :1038
    CHECK_INTERVAL = 10       # :1039

    @staticmethod
    def _expand_pair(i: int, j: int):
        """One (i, j) of the iterative mode is a two-sequence AllPairIterator with self pairs
        (:976-991): (i,i), (i,j), (j,i), (j,j) in that order."""
        return [(i, i), (i, j), (j, i), (j, j)]

    def align_and_unite_iterative(self, args: Args = None, tree_pairs=None, random_pairs=None):  # :868-1132
        """--iterative. Without explicit lists the tree pairs and the random pairs come from
        extract_tree_pairs_separated over the -x strategy (default tree:3,3,0.1,16; MinHash sketches and the
        hashed pair filter on the device, seqrush_b200/sparsify.py), like the reference (:900-945). Phase 1
        runs every tree pair in one launch; phase 2 runs the random pairs CHECK_INTERVAL at a time and
        counts components on the device after each group, stopping after STABILITY_THRESHOLD unchanged
        counts in a row — the reference's check points, so the same prefix of random_pairs is processed
        and the partition is the reference's. Returns a dict with the counts and pairs processed."""
        args = args or Args()
        if args.output_alignments:
            raise NotImplementedError("--iterative with --output-alignments: write the PAF with align_and_unite")
        if tree_pairs is None and random_pairs is None:
            tree_pairs, random_pairs = self.tree_pairs_separated(args)
        tree_pairs = [(int(i), int(j)) for i, j in (tree_pairs or ())]
        random_pairs = [(int(i), int(j)) for i, j in (random_pairs or ())]
        sub = Args(scores=args.scores, min_match_length=args.min_match_length)
        if tree_pairs:
            self.align_and_unite(sub, pairs=[q for p in tree_pairs for q in self._expand_pair(*p)])
        post_tree = self.count_components()
        prev, stable, done, stabilized = post_tree, 0, 0, False
        ci = self.CHECK_INTERVAL
        for b0 in range(0, len(random_pairs), ci):
            grp = random_pairs[b0:b0 + ci]
            self.align_and_unite(sub, pairs=[q for p in grp for q in self._expand_pair(*p)])
            done += len(grp)
            if len(grp) < ci:  # the tail shorter than CHECK_INTERVAL is processed without a check (:1089)
                break
            cnt = self.count_components()
            if cnt == prev:
                stable += 1
                if stable >= self.STABILITY_THRESHOLD:
                    stabilized = True
                    break
            else:
                stable = 0
            prev = cnt
        return {"post_tree_components": post_tree, "final_components": self.count_components(),
                "random_pairs_processed": done, "random_pairs_skipped": len(random_pairs) - done,
                "stabilized": stabilized}

    def build_bidirected_graph(self, representatives=None):  # src/bidirected_builder.rs:17-289
        from .graph import build_bidirected_graph
        return build_bidirected_graph(self, representatives)

    def write_bidirected_gfa(self, output_path: str, no_compact: bool = False, representatives=None):
        """src/bidirected_gfa_writer.rs:8-167 with --skip-sgd --skip-groom --skip-topo: build, compact and
        renumber unless no_compact, validate the paths against the sequences, write. Returns the graph."""
        g = self.build_bidirected_graph(representatives)
        if not no_compact:
            g.compact()
        g.write_gfa(output_path)
        return g

    def count_components(self) -> int:  # :341-353
        return self.union_find.count_components(self.total_length)

    def canonical_partition(self) -> np.ndarray:
        """label[p] = minimum Pos of p's component, for p in 0..2*total_length."""
        return self.union_find.roots(0, 2 * self.total_length)
