"""Device context: one per GPU (owns the stream, the uploaded sequence set and the workspaces)."""
import ctypes as C
import weakref

import numpy as np

from . import _lib

_default = None


class Context:
    def __init__(self, device: int = 0, stream: int = 0):
        self._h = C.c_void_p()
        self.device = device
        _lib.check(_lib.lib().srgpu_ctx_create(device, C.c_void_p(stream or None), C.byref(self._h)))
        self._seq_keep = None
        self.n_seqs = 0
        self._children = weakref.WeakSet()  # union-finds living on this context: freed before it

    def close(self):
        for ch in list(getattr(self, "_children", ())):
            ch.close()
        if getattr(self, "_h", None) and self._h.value:
            _lib.lib().srgpu_ctx_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def configure(self, threads=0, ctas_per_sm=0, ring_diagonals=-1):
        _lib.check(_lib.lib().srgpu_ctx_configure(self._h, threads, ctas_per_sm, ring_diagonals))

    def set_split(self, split_score: int):
        """0: every pair stays in one CTA; >= 250: sub-problems at or below that score go to one-warp CTAs."""
        _lib.check(_lib.lib().srgpu_ctx_set_split(self._h, int(split_score)))

    def stream_ptr(self) -> int:
        """cudaStream_t of every launch of this context."""
        return int(_lib.lib().srgpu_ctx_stream(self._h) or 0)

    def synchronize(self):
        _lib.check(_lib.lib().srgpu_ctx_synchronize(self._h))

    def set_sequences(self, seqs):
        """seqs: list of bytes / numpy uint8 arrays (array order defines the Pos offsets)."""
        keep = [np.frombuffer(s, dtype=np.uint8) if isinstance(s, (bytes, bytearray)) else np.ascontiguousarray(s, dtype=np.uint8)
                for s in seqs]
        views = (_lib.SeqView * len(keep))()
        for i, a in enumerate(keep):
            views[i].data = a.ctypes.data
            views[i].len = a.size
        _lib.check(_lib.lib().srgpu_set_sequences(self._h, views, len(keep)))
        self._seq_keep = keep
        self.n_seqs = len(keep)
        self.seq_lens = [int(a.size) for a in keep]

    def total_length(self) -> int:
        return _lib.lib().srgpu_total_length(self._h)

    def stats(self) -> dict:
        st = _lib.Stats()
        _lib.check(_lib.lib().srgpu_ctx_stats(self._h, C.byref(st)))
        return {k: getattr(st, k) for k, _ in _lib.Stats._fields_}


def default_context() -> Context:
    global _default
    if _default is None:
        _default = Context(0)
    return _default
