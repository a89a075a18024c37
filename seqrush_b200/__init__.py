"""seqrush_b200 -- B200-native (sm_100a) implementation of the all-vs-all align-and-unite hot path
of pangenome/seqrush (`--aligner allwave`): strand choice by exact edit distance, gap-affine-2p
BiWFA with on-device backtrace, CIGAR walk and lock-free union of matched Pos pairs. The compute
lives in seqrush_b200/csrc (CUDA) behind the C ABI of include/seqrush_b200.h; the modules here
mirror the reference's Rust interfaces for that path."""
from . import pos
from ._lib import build, SrgpuError
from .allwave import (AlignmentParams, AlignmentResult, AllPairIterator, align_pair, alignment_to_paf,
                      cigar_bytes_to_string, parse_scores, reverse_complement)
from .allwave import Sequence as AllwaveSequence
from .aligner import (Aligner, AlignerBackend, AlignmentRecord, AlignmentSequence, AllwaveAligner, create_aligner)
from .context import Context, default_context
from .graph import BidirectedGraph, Handle
from .sparsify import SparsificationStrategy, parse_sparsification
from .seqrush import Args, SeqRush, Sequence, load_sequences
from .union_find import BidirectedUnionFind

__all__ = ["pos", "build", "SrgpuError", "AlignmentParams", "AlignmentResult", "AllPairIterator", "align_pair",
           "alignment_to_paf", "cigar_bytes_to_string", "parse_scores", "reverse_complement", "AllwaveSequence",
           "Aligner", "AlignerBackend", "AlignmentRecord", "AlignmentSequence", "AllwaveAligner", "create_aligner",
           "Context", "default_context", "BidirectedGraph", "Handle", "SparsificationStrategy", "parse_sparsification", "Args", "SeqRush", "Sequence", "load_sequences", "BidirectedUnionFind"]
