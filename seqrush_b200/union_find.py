"""BidirectedUnionFind -- host-side mirror of src/bidirected_union_find.rs:7-135 over the
HBM-resident union-find of the CUDA library (same method names, argument meaning and error
behaviour; ``find`` returns the canonical representative = minimum Pos of the component)."""
import ctypes as C

import numpy as np

from . import _lib
from .context import default_context


class BidirectedUnionFind:
    def __init__(self, max_offset: int, ctx=None):  # ::new, :16-24
        self.ctx = ctx or default_context()
        self._h = C.c_void_p()
        _lib.check(_lib.lib().srgpu_uf_new(self.ctx._h, max_offset, C.byref(self._h)))
        self.capacity = _lib.lib().srgpu_uf_capacity(self._h)
        self.ctx._children.add(self)

    @classmethod
    def new(cls, max_offset: int, ctx=None):
        return cls(max_offset, ctx)

    def clear(self):  # UFRush::clear (uf_rush lib.rs:211-216)
        _lib.check(_lib.lib().srgpu_uf_clear(self._h))

    def close(self):
        if getattr(self, "_h", None) and self._h.value:
            _lib.lib().srgpu_uf_free(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def find(self, pos: int) -> int:  # :27-31 (out of range: the reference asserts, uf_rush lib.rs:113)
        out = C.c_uint64()
        _lib.check(_lib.lib().srgpu_uf_find(self._h, pos, C.byref(out)))
        return out.value

    def unite(self, pos1: int, pos2: int) -> None:  # :35-43
        _lib.check(_lib.lib().srgpu_uf_unite(self._h, pos1, pos2))

    def same(self, pos1: int, pos2: int) -> bool:  # :46-54
        out = C.c_int()
        _lib.check(_lib.lib().srgpu_uf_same(self._h, pos1, pos2, C.byref(out)))
        return bool(out.value)

    def unite_matching_region(self, seq1_offset, seq2_offset, seq1_local_start, seq2_local_start, match_length,
                              seq1_is_rc, seq1_len):  # :60-98
        _lib.check(_lib.lib().srgpu_uf_unite_matching_region(self._h, seq1_offset, seq2_offset, seq1_local_start,
                                                             seq2_local_start, match_length, int(bool(seq1_is_rc)),
                                                             seq1_len))

    def unite_matching_region_seq2_rc(self, seq1_offset, seq2_offset, seq1_local_start, seq2_local_start,
                                      match_length, seq2_is_rc, seq2_len):  # :103-129
        _lib.check(_lib.lib().srgpu_uf_unite_matching_region_seq2_rc(self._h, seq1_offset, seq2_offset,
                                                                     seq1_local_start, seq2_local_start, match_length,
                                                                     int(bool(seq2_is_rc)), seq2_len))

    # ---- batched forms (what a Rust host would call instead of per-Pos FFI round trips) ----
    def unite_regions(self, regions):
        """regions: iterable of (seq1_offset, seq2_offset, s1, s2, len, seq1_is_rc, seq1_len)."""
        regions = list(regions)
        arr = (_lib.Region * len(regions))()
        for i, r in enumerate(regions):
            arr[i] = _lib.Region(r[0], r[1], r[2], r[3], r[4], r[6], int(bool(r[5])), 0)
        _lib.check(_lib.lib().srgpu_uf_unite_regions(self._h, arr, len(regions)))

    def unite_fr_all(self, total_length: int):
        """SeqRush::new's loop (src/seqrush.rs:324-328)."""
        _lib.check(_lib.lib().srgpu_uf_unite_fr_all(self._h, total_length))

    def roots(self, first=0, n=None) -> np.ndarray:
        n = self.capacity - first if n is None else n
        out = np.empty(n, dtype=np.uint64)
        _lib.check(_lib.lib().srgpu_uf_roots(self._h, first, n, out.ctypes.data))
        return out

    def count_components(self, total_length: int) -> int:  # src/seqrush.rs:341-353
        out = C.c_uint64()
        _lib.check(_lib.lib().srgpu_uf_count_components(self._h, total_length, C.byref(out)))
        return out.value

    def checksum(self) -> int:
        out = C.c_uint64()
        _lib.check(_lib.lib().srgpu_uf_checksum(self._h, C.byref(out)))
        return out.value

    def compress(self):
        _lib.check(_lib.lib().srgpu_uf_compress(self._h))

    def compress_async(self):
        _lib.check(_lib.lib().srgpu_uf_compress_async(self._h))

    def merge_gathered(self, d_gathered_ptr: int, world: int, my_rank: int):
        """One pass over the [world][capacity] gathered parent arrays (stream-ordered, no host sync)."""
        _lib.check(_lib.lib().srgpu_uf_merge_gathered(self._h, d_gathered_ptr, self.capacity, world, my_rank))

    def device_ptr(self) -> int:
        return _lib.lib().srgpu_uf_device_ptr(self._h)

    def merge_roots(self, d_peer_roots_ptr: int, n: int):
        _lib.check(_lib.lib().srgpu_uf_merge_roots(self._h, d_peer_roots_ptr, n))

    def as_tensor(self):
        """Zero-copy torch view (int32 bit pattern of the uint32 parents) for NCCL collectives."""
        import torch

        class _Arr:
            pass

        a = _Arr()
        a.__cuda_array_interface__ = {"shape": (int(self.capacity),), "typestr": "<i4",
                                      "data": (int(self.device_ptr()), False), "version": 2}
        return torch.as_tensor(a, device="cuda:%d" % self.ctx.device)
