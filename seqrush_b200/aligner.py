"""The `Aligner` trait surface of src/aligner.rs (AlignmentSequence :5-9, AlignmentRecord :12-24, trait Aligner
:27-33, AlignerBackend :36-60, create_aligner :64-96) and its AllWave implementation
(src/aligner/allwave_impl.rs:7-148), over the CUDA library: records carry RLE CIGAR strings in the
reference's convention ('=' / 'X' / 'I' = query only / 'D' = target only), strand '+' / '-'.

The production pipeline does not go through this trait (src/seqrush.rs:728-738 drives AllPairIterator
directly); it is what src/seqrush_clean.rs:140-148,258-291 uses, single-threaded."""
from abc import ABC, abstractmethod
from dataclasses import dataclass
from typing import List, Optional

from .allwave import AlignmentParams, AllPairIterator, cigar_bytes_to_string
from .allwave import Sequence as AwSequence


@dataclass
class AlignmentSequence:  # src/aligner.rs:5-9
    id: str
    seq: bytes


@dataclass
class AlignmentRecord:  # src/aligner.rs:12-24
    query_name: str
    query_len: int
    query_start: int
    query_end: int
    strand: str
    target_name: str
    target_len: int
    target_start: int
    target_end: int
    cigar: str


class Aligner(ABC):  # src/aligner.rs:27-33
    @abstractmethod
    def align_sequences(self, sequences: List[AlignmentSequence]) -> List[AlignmentRecord]:
        ...


class AlignerBackend:  # src/aligner.rs:36-60
    AllWave = "allwave"
    SweepGA = "sweepga"

    @staticmethod
    def from_str(s: str) -> str:
        v = s.lower()
        if v in (AlignerBackend.AllWave, AlignerBackend.SweepGA):
            return v
        raise ValueError("Unknown aligner: %s" % s)


class AllwaveAligner(Aligner):  # src/aligner/allwave_impl.rs:7-148
    # allwave_impl.rs:15-23. NOTE: match_score = 2 is positive; WFA2-lib refuses positive match scores
    # (wavefront_penalties.c:141-144 prints "Match score must be negative or zero" and exits), so with these
    # defaults the reference can only work if lib_wfa2@819b82cd (source not in the tree) flips the sign --
    # parity unpinned. The CUDA library answers like WFA2-lib: SRGPU_E_ARG. Use with_params for real runs.
    DEFAULT_PARAMS = (2, 4, 4, 2, 24, 1)

    def __init__(self, threads: int = 1, verbose: bool = False, params: Optional[AlignmentParams] = None, ctx=None):
        self.threads = threads  # kept for interface parity: the device schedules pairs itself
        self.verbose = verbose
        self.params = params or AlignmentParams(*self.DEFAULT_PARAMS)
        self.ctx = ctx

    @classmethod
    def new(cls, threads: int, verbose: bool, ctx=None):
        return cls(threads, verbose, None, ctx)

    @classmethod
    def with_params(cls, threads: int, verbose: bool, params: AlignmentParams, ctx=None):
        return cls(threads, verbose, params, ctx)

    @staticmethod
    def get_prioritized_pairs(sequences: List[AlignmentSequence], sparsification, verbose: bool = False, ctx=None):
        """allwave_impl.rs:44-92: the pair list AllPairIterator::with_options builds for a strategy (self pairs
        included, no alignment performed)."""
        aw = [AwSequence(s.id, s.seq) for s in sequences]
        it = AllPairIterator.with_options(aw, AlignmentParams(*AllwaveAligner.DEFAULT_PARAMS), False, False, sparsification, ctx=ctx)
        return [(int(i), int(j)) for i, j in it.pairs]

    def align_sequences(self, sequences: List[AlignmentSequence]) -> List[AlignmentRecord]:  # allwave_impl.rs:95-148
        aw = [AwSequence(s.id, s.seq) for s in sequences]
        it = AllPairIterator.with_options(aw, self.params, False, False, None, ctx=self.ctx)
        out = []
        for a in it:
            q, t = aw[a.query_idx], aw[a.target_idx]
            out.append(AlignmentRecord(q.id, len(q.seq), a.query_start, a.query_end, "-" if a.is_reverse else "+",
                                       t.id, len(t.seq), a.target_start, a.target_end, cigar_bytes_to_string(a.cigar_bytes)))
        return out


def create_aligner(backend: str, threads: int = 1, verbose: bool = False, frequency: Optional[int] = None, ctx=None) -> Aligner:
    """src/aligner.rs:64-96. Only the AllWave backend exists here (SweepGA/FastGA is a different aligner,
    out of scope: SURVEY.md 2 row 16)."""
    backend = AlignerBackend.from_str(backend)
    if backend == AlignerBackend.AllWave:
        return AllwaveAligner.new(threads, verbose, ctx)
    raise RuntimeError("SweepGA aligner not available. Rebuild with --features use-sweepga")
