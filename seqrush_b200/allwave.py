"""Host-side mirror of the allwave surface the seqrush hot path uses (allwave/src/types.rs,
allwave/src/iterator.rs, allwave/src/alignment.rs, allwave/src/lib.rs): same names, argument
meaning and failure behaviour, with the per-pair arithmetic running in the CUDA library."""
import ctypes as C
from dataclasses import dataclass, field
from typing import List, Optional

import numpy as np

from . import _lib
from .context import default_context

INT32_MAX = 2**31 - 1


@dataclass
class Sequence:  # allwave/src/types.rs:6-10
    id: str
    seq: bytes


@dataclass
class AlignmentParams:  # allwave/src/types.rs:37-61 (defaults = seqrush -S 0,5,8,2,24,1)
    match_score: int = 0
    mismatch_penalty: int = 5
    gap_open: int = 8
    gap_extend: int = 2
    gap2_open: Optional[int] = 24
    gap2_extend: Optional[int] = 1
    max_divergence: Optional[float] = None  # never read by the reference either

    def mode(self) -> str:  # AlignmentMode::from_params, allwave/src/types.rs:101-111
        if self.gap2_open is not None and self.gap2_extend is not None:
            return "TwoPieceAffine"
        if self.gap_open == self.gap_extend and self.gap_open == self.mismatch_penalty:
            return "EditDistance"
        return "SinglePieceAffine"

    def to_c(self) -> "_lib.Penalties":
        m = self.mode()
        if m == "TwoPieceAffine":
            return _lib.Penalties(self.match_score, self.mismatch_penalty, self.gap_open, self.gap_extend,
                                  self.gap2_open, self.gap2_extend)
        if m == "EditDistance":  # allwave/src/alignment.rs:205-213: affine with x, x, x
            return _lib.Penalties(self.match_score, self.mismatch_penalty, self.mismatch_penalty, self.mismatch_penalty, -1, -1)
        return _lib.Penalties(self.match_score, self.mismatch_penalty, self.gap_open, self.gap_extend, -1, -1)

    @classmethod
    def edit_distance(cls):  # allwave/src/types.rs:49-61
        return cls(0, 1, 1, 1, None, None)


def parse_scores(s: str) -> AlignmentParams:  # allwave/src/lib.rs:116-153
    parts = [int(x) for x in s.split(",")]
    if len(parts) == 4:
        return AlignmentParams(parts[0], parts[1], parts[2], parts[3], None, None)
    if len(parts) == 6:
        return AlignmentParams(*parts)
    raise ValueError("scores must be match,mismatch,gap_open,gap_extend[,gap2_open,gap2_extend]")


@dataclass
class AlignmentResult:  # allwave/src/types.rs:14-35
    query_idx: int
    target_idx: int
    query_start: int
    query_end: int
    target_start: int
    target_end: int
    is_reverse: bool
    cigar_bytes: bytes
    score: int
    num_matches: int
    alignment_length: int
    penalty: int = 0          # extra: penalty recomputed from the CIGAR
    edit_distance: int = 0    # extra: min(fwd, rc) unit-cost distance from the orientation step


def reverse_complement(seq: bytes) -> bytes:  # allwave/src/alignment.rs:174-186
    tbl = bytearray(b"N" * 256)
    for a, b in (("A", "T"), ("a", "T"), ("T", "A"), ("t", "A"), ("C", "G"), ("c", "G"), ("G", "C"), ("g", "C")):
        tbl[ord(a)] = ord(b)
    return bytes(seq)[::-1].translate(bytes(tbl))


def cigar_bytes_to_string(cigar_bytes: bytes) -> str:  # allwave/src/alignment.rs:327-356
    if not cigar_bytes:
        return ""
    a = np.frombuffer(cigar_bytes, dtype=np.uint8)
    cut = np.flatnonzero(a[1:] != a[:-1]) + 1
    starts = np.concatenate(([0], cut))
    ends = np.concatenate((cut, [a.size]))
    m = {ord("M"): "=", ord("X"): "X", ord("I"): "D", ord("D"): "I"}
    return "".join("%d%s" % (e - s, m.get(int(a[s]), "?")) for s, e in zip(starts, ends))


def alignment_to_paf(r: AlignmentResult, seqs: List[Sequence]) -> str:
    """allwave/src/lib.rs:71-112, field for field: the 12 mandatory PAF columns with
    block_len = max(query_end - query_start, target_end - target_start) (:77-79), mapping quality 60,
    then ``gi:f:{identity:.6}`` (identity = num_matches / alignment_length, 0.0 for an empty
    alignment, :82-86) and ``cg:Z:`` + cigar_bytes_to_string (:89, :95-111)."""
    q, t = seqs[r.query_idx], seqs[r.target_idx]
    block_len = max(r.target_end - r.target_start, r.query_end - r.query_start)
    identity = (r.num_matches / r.alignment_length) if r.alignment_length > 0 else 0.0
    return "%s\t%d\t%d\t%d\t%s\t%s\t%d\t%d\t%d\t%d\t%d\t%d\tgi:f:%.6f\tcg:Z:%s" % (
        q.id, len(q.seq), r.query_start, r.query_end, "-" if r.is_reverse else "+", t.id, len(t.seq),
        r.target_start, r.target_end, r.num_matches, block_len, 60, identity, cigar_bytes_to_string(r.cigar_bytes))


class AllPairIterator:
    """allwave/src/iterator.rs:12-253. ``with_options(sequences, params, exclude_self,
    use_mash_orientation, sparsification)``; iterating yields AlignmentResult in pair order
    (the reference's rayon order is unspecified). ``sparsification`` is a SparsificationStrategy or the
    string form of seqrush's -x flag (none | auto | random:F | connectivity:F | tree:kn[,kf[,rand[,k]]]); an
    explicit ``pairs`` list overrides it. Mash orientation is not supported (the seqrush hot path passes false)."""

    def __init__(self, sequences: List[Sequence], params: AlignmentParams, exclude_self=True,
                 use_mash_orientation=False, sparsification=None, pairs=None, ctx=None):
        if use_mash_orientation:
            raise NotImplementedError("mash orientation is outside the seqrush hot path (seqrush.rs:733 passes false)")
        from .sparsify import SparsificationStrategy, parse_sparsification, sparsified_pairs
        if isinstance(sparsification, str):
            sparsification = parse_sparsification(sparsification)
        self.sequences = sequences
        self.params = params
        self.ctx = ctx or default_context()
        self._uploaded = False
        if pairs is None:
            # allwave/src/iterator.rs:37-77: i-major n^2 list (self pairs unless excluded), then the strategy
            strat = sparsification or SparsificationStrategy.none()
            if strat.kind == "tree":  # the sketches need the sequences on the device
                self.ctx.set_sequences([s.seq for s in sequences])
                self._uploaded = True
            self.pairs = sparsified_pairs(sequences, strat, exclude_self, self.ctx, uploaded=self._uploaded)
        else:
            self.pairs = np.ascontiguousarray(np.asarray(pairs, dtype=np.uint32).reshape(-1, 2))

    @classmethod
    def with_options(cls, sequences, params, exclude_self, use_mash_orientation, sparsification=None, **kw):
        return cls(sequences, params, exclude_self, use_mash_orientation, sparsification, **kw)

    def with_orientation_params(self, _params):  # ignored downstream by the reference too (alignment.rs:22)
        return self

    def __len__(self):
        return int(self.pairs.shape[0])

    def align_all(self, want_cigars=True, batch=4096):
        """Run every pair; returns (list[AlignmentResult])."""
        out = []
        for r in self._batches(want_cigars, batch):
            out.extend(r)
        return out

    def _batches(self, want_cigars, batch, unite_into=None, min_match_len=0):
        """Batches of AlignmentResult in pair order. With ``unite_into`` (a BidirectedUnionFind on the
        same context) every batch is also united on the device from the same alignment pass
        (srgpu_align_pairs_and_unite)."""
        L = _lib.lib()
        if not self._uploaded:
            self.ctx.set_sequences([s.seq for s in self.sequences])
            self._uploaded = True
        pen = self.params.to_c()
        for b0 in range(0, len(self), batch):
            pr = np.ascontiguousarray(self.pairs[b0:b0 + batch])
            n = pr.shape[0]
            res = (_lib.Result * n)()
            arena = C.c_void_p()
            alen = C.c_uint64()
            if unite_into is None:
                _lib.check(L.srgpu_align_pairs(self.ctx._h, pr.ctypes.data, n, C.byref(pen), res,
                                               C.byref(arena) if want_cigars else None, C.byref(alen)))
            else:
                _lib.check(L.srgpu_align_pairs_and_unite(self.ctx._h, unite_into._h, pr.ctypes.data, n, C.byref(pen),
                                                         min_match_len, res, C.byref(arena) if want_cigars else None,
                                                         C.byref(alen)))
            buf = C.string_at(arena, alen.value) if want_cigars and arena.value else b""
            if arena.value:
                L.srgpu_free(arena)
            rs = []
            for i in range(n):
                r = res[i]
                failed = r.status != 0
                cig = b"" if failed else buf[r.cigar_off:r.cigar_off + r.cigar_len]
                rs.append(AlignmentResult(
                    query_idx=int(r.query_idx), target_idx=int(r.target_idx), query_start=0,
                    query_end=0 if failed else int(r.query_end), target_start=0,
                    target_end=0 if failed else int(r.target_end), is_reverse=bool(r.is_reverse), cigar_bytes=cig,
                    score=INT32_MAX if failed else int(r.score), num_matches=int(r.num_matches),
                    alignment_length=int(r.alignment_length), penalty=int(r.penalty),
                    edit_distance=int(r.edit_distance)))
            yield rs

    def __iter__(self):
        for rs in self._batches(True, 1024):
            yield from rs

    into_par_iter = __iter__


def align_pair(query: Sequence, target: Sequence, query_idx: int, target_idx: int, params: AlignmentParams,
               _orientation_params=None, use_mash_orientation=False, ctx=None) -> AlignmentResult:
    """allwave::align_pair (allwave/src/alignment.rs:16-57) for one pair."""
    it = AllPairIterator([query, target], params, False, use_mash_orientation, pairs=[(0, 1)], ctx=ctx)
    r = it.align_all()[0]
    r.query_idx, r.target_idx = query_idx, target_idx
    return r
