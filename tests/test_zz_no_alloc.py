"""srgpu_uf_count_components_static -- the allocation-free variant of count_components (scratch in a file-scope
__device__ array, no counterpart in the reference) -- through the C ABI
against the existing srgpu_uf_count_components: equal counts (integers: exact) on two calls in a row, on a second,
smaller union-find afterwards (the shared scratch keeps the previous call's bits and must be cleared), and a clear
error above the stated maximum. Collected last on purpose (pytest -x): a failure here hides no other test."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def s(built):
    import torch
    assert torch.cuda.is_available(), "gpu tests need a CUDA device"
    import seqrush_b200 as s
    return s


def _counts(L, uf, total):
    a, b = C.c_uint64(), C.c_uint64()
    from seqrush_b200 import _lib
    _lib.check(L.srgpu_uf_count_components(uf._h, total, C.byref(a)))
    _lib.check(L.srgpu_uf_count_components_static(uf._h, total, C.byref(b)))
    return a.value, b.value


def test_count_components_static_matches_existing(s):
    from seqrush_b200 import _lib, synth
    L = _lib.lib()
    assert L.srgpu_uf_count_components_static_max_capacity() == 32 * (1 << 20)
    ctx = s.Context(0)
    try:
        # a real partition: 6 x 3 kb with reverse-complemented sequences, aligned and united
        ids, seqs = synth.make_sequences(6, 3000, 0.03, 91, 0.4, 0.0)
        sr = s.SeqRush([s.Sequence(i, a.tobytes()) for i, a in zip(ids, seqs)], ctx=ctx)
        total = sr.total_length
        want0, got0 = _counts(L, sr.union_find, total)          # call 1: only forward ~ reverse united
        assert want0 == got0 == total
        sr.align_and_unite(s.Args(min_match_length=0))
        want1, got1 = _counts(L, sr.union_find, total)          # call 2, straight after: far fewer components
        assert want1 == got1 and 0 < got1 < total
        want2, got2 = _counts(L, sr.union_find, total)          # and again: the array held call 2's bits
        assert (want2, got2) == (want1, got1)
        # a smaller, unrelated union-find next: stale bits of the larger one must not be counted as seen
        n = 5000
        uf = s.BidirectedUnionFind(n, ctx)
        rng = np.random.default_rng(3)
        for a, b in rng.integers(0, 2 * n, size=(300, 2)):
            uf.unite(int(a), int(b))
        w, g = _counts(L, uf, n)
        assert w == g and g <= n
        w_part, g_part = _counts(L, uf, n // 3)                 # a prefix of the positions only
        assert w_part == g_part <= g
        z = C.c_uint64(77)
        _lib.check(L.srgpu_uf_count_components_static(uf._h, 0, C.byref(z)))
        assert z.value == 0
        uf.close()
        # beyond the stated maximum: refused with a clear message, nothing launched
        big = s.BidirectedUnionFind(17_000_000, ctx)            # capacity 34 000 002 Pos > 33 554 432
        out = C.c_uint64()
        rc = L.srgpu_uf_count_components_static(big._h, 10, C.byref(out))
        assert rc == -2 and b"exceeds the static scratch" in L.srgpu_last_error()
        _lib.check(L.srgpu_uf_count_components(big._h, 10, C.byref(out)))
        assert out.value == 10
        big.close()
    finally:
        ctx.close()
