"""Control logic of the CUDA aligner (seqrush_b200/csrc/wfa_core.cuh, orient_core.cuh) compiled for
the host with a one-thread "CTA" (tests/host_emul) and checked against the goldens. This is a
CPU-side logic test only; the product library never contains the host build."""
import ctypes as C
import os

import numpy as np
import pytest

from conftest import ROOT, expand_rle, rle


@pytest.fixture(scope="module")
def emul(built):
    L = C.CDLL(os.path.join(ROOT, "tests", "host_emul", "libwfa_emul.so"))
    L.emul_wfa_align.argtypes = [C.c_char_p, C.c_int, C.c_char_p, C.c_int] + [C.c_int] * 8 + [C.c_char_p, C.POINTER(C.c_int), C.POINTER(C.c_int)]
    L.emul_orient.argtypes = [C.c_char_p, C.c_char_p, C.c_int, C.c_char_p, C.c_int, C.c_int, C.POINTER(C.c_int)]
    return L


def _align(L, p, t, pen, W=0, cap=0, lbits=0):
    W = W or 2 * (max(len(p), len(t)) + 8)
    W = (W + 63) & ~63
    cap = cap or max(512, 100 * pen[0] + 8) + 8
    buf = C.create_string_buffer(len(p) + len(t) + 16)
    n, sc = C.c_int(), C.c_int()
    st = L.emul_wfa_align(p, len(p), t, len(t), *pen, W, cap, lbits, buf, C.byref(n), C.byref(sc))
    return st, sc.value, buf.raw[:n.value]


def test_emul_wfa2_goldens(emul, wfa2_golden):
    pen = tuple(wfa2_golden["penalties"][1:])
    for i, v in enumerate(wfa2_golden["vectors"]):
        for lbits in ((1, 2, 3) if i % 10 == 0 else (0,)):  # 2-, 4- and 8-bit packings must agree
            st, score, cig = _align(emul, v["p"].encode(), v["t"].encode(), pen, lbits=lbits)
            assert st == 0 and rle(cig) == v["cigar"], (i, lbits)


def test_emul_seqrush_pairs(emul, seqrush_golden):
    from seqrush_b200.allwave import reverse_complement
    n = 0
    for c in seqrush_golden["cases"]:
        q, t = c["q"].encode("latin1"), c["t"].encode("latin1")
        if max(len(q), len(t)) > 8100:
            continue
        rc = reverse_complement(q)
        d = C.c_int()
        is_rev = emul.emul_orient(q, rc, len(q), t, len(t), 0, C.byref(d))
        assert is_rev == c["is_reverse"], c["tag"]
        assert d.value == min(c["fwd_ed"], c["rev_ed"]), c["tag"]
        st, score, cig = _align(emul, rc if is_rev else q, t, (5, 8, 2, 24, 1))
        assert st == 0 and score == c["score"] and rle(cig) == c["cigar"], c["tag"]
        n += 1
    assert n >= 40


def test_emul_narrow_ring_reports_overflow(emul):
    rng = np.random.default_rng(1)
    a = bytes(np.frombuffer(b"ACGT", dtype=np.uint8)[rng.integers(0, 4, 3000)])
    b = bytes(np.frombuffer(b"ACGT", dtype=np.uint8)[rng.integers(0, 4, 3000)])
    st, _, _ = _align(emul, a, b, (5, 8, 2, 24, 1), W=256)
    assert st == 2  # ST_WS_DIAG_OVERFLOW: the host retries with a wider ring, never a wrong answer


def test_emul_split_path_equals_monolithic(emul):
    """The device's split path -- bisect down to split_T, bisect every task down to the base-case bound, finish every
    leaf with wf_align_leaf, join the pieces in order -- gives the monolithic path's CIGAR and score (same core
    functions, host build; small tables and a narrow ring for the tasks like the one-warp kernel)."""
    import numpy as np
    emul.emul_wfa_align_split.argtypes = [C.c_char_p, C.c_int, C.c_char_p, C.c_int] + [C.c_int] * 10 + \
        [C.c_char_p, C.POINTER(C.c_int), C.POINTER(C.c_int), C.POINTER(C.c_int), C.POINTER(C.c_int)]
    rng = np.random.default_rng(77)
    acgt = np.frombuffer(b"ACGT", dtype=np.uint8)

    def mutate(a, rate):
        out = []
        for ch in a:
            r = rng.random()
            if r < rate / 3:
                out.append(int(rng.choice(acgt)))
            elif r < 2 * rate / 3:
                continue
            elif r < rate:
                out.append(int(ch))
                out.extend(int(x) for x in rng.choice(acgt, size=1 + (int(rng.integers(40)) if rng.random() < 0.1 else 0)))
            else:
                out.append(int(ch))
        return bytes(out)

    pen = (5, 8, 2, 24, 1)
    cases = [(9000, 0.03, 1000), (7000, 0.05, 1000), (6000, 0.04, 400), (5000, 0.02, 250), (800, 0.02, 1000), (3000, 0.10, 600)]
    deep = 0
    for L, rate, split_T in cases:
        a = bytes(rng.choice(acgt, size=L))
        b = mutate(a, rate)
        st0, sc0, cig0 = _align(emul, a, b, pen)
        assert st0 == 0
        W = (2 * (max(len(a), len(b)) + 16) + 63) & ~63
        buf = C.create_string_buffer(len(a) + len(b) + 64)
        n, sc, nt, nl = C.c_int(), C.c_int(), C.c_int(), C.c_int()
        st = emul.emul_wfa_align_split(a, len(a), b, len(b), *pen, W, 520, 0, split_T, 250, buf, C.byref(n), C.byref(sc),
                                       C.byref(nt), C.byref(nl))
        assert st == 0, (L, rate, split_T, st)
        assert buf.raw[:n.value] == cig0, (L, rate, split_T)
        assert sc.value == sc0, (L, rate, split_T, sc.value, sc0)
        assert nl.value >= nt.value >= 1
        deep += int(nt.value > 1 and nl.value > nt.value)
    assert deep >= 3, "the cases were meant to exercise both levels of the split"
