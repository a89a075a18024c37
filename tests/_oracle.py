"""ctypes access to the oracle (oracle/liboracle.so + oracle/_ref/libsrref_<isa>.so).
TEST INFRASTRUCTURE: imported only by tests/, __graft_entry__.smoke() and bench.py's CPU-baseline legs."""
import ctypes as C
import os

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")
INT32_MAX = 2**31 - 1
INT32_MIN = -2**31


def ref_isa():
    flags = ""
    try:
        with open("/proc/cpuinfo") as f:
            for line in f:
                if line.startswith("flags"):
                    flags = line
                    break
    except OSError:
        pass
    fl = set(flags.split())
    if {"avx512f", "avx512cd", "avx512vl", "avx512bw", "avx512dq"} <= fl:
        return "avx512"
    if {"avx2", "bmi1", "bmi2"} <= fl or {"avx2", "bmi2"} <= fl:
        return "avx2"
    return "generic"


def ref_path(isa=None):
    return os.path.join(ORACLE_DIR, "_ref", "libsrref_%s.so" % (isa or ref_isa()))


class RunStats(C.Structure):
    _fields_ = [("seconds", C.c_double), ("edlib_seconds", C.c_double), ("wfa_seconds", C.c_double),
                ("unite_seconds", C.c_double), ("aligned_bp", C.c_uint64), ("unite_calls", C.c_uint64),
                ("pairs", C.c_uint64), ("cigar_columns", C.c_uint64)]


_lib = None


def lib():
    global _lib
    if _lib is not None:
        return _lib
    L = C.CDLL(os.path.join(ORACLE_DIR, "liboracle.so"))
    u64, vp, i32 = C.c_uint64, C.c_void_p, C.c_int
    L.orc_open.restype = vp
    L.orc_open.argtypes = [C.c_char_p]
    L.orc_allwave_rc.argtypes = [vp, u64, vp]
    L.orc_align_pair.argtypes = [vp, vp, i32, vp, i32, C.POINTER(i32), C.c_char_p] + [C.POINTER(i32)] * 5
    L.orc_wfa_only.argtypes = [vp, vp, i32, vp, i32, vp, i32, C.c_char_p, C.POINTER(i32), C.POINTER(i32)]
    L.orc_edlib.argtypes = [vp, vp, i32, vp, i32]
    L.orc_wfa_mode.argtypes = [vp, vp, i32, vp, i32, vp, i32, C.c_char_p, C.POINTER(i32), C.POINTER(i32)]
    L.orc_cigar_bytes_to_string.argtypes = [C.c_char_p, i32, C.c_char_p, i32]
    L.orc_uf_new.restype = vp
    L.orc_uf_new.argtypes = [u64]
    L.orc_uf_free.argtypes = [vp]
    L.orc_uf_capacity.restype = u64
    L.orc_uf_capacity.argtypes = [vp]
    L.orc_uf_find.restype = u64
    L.orc_uf_find.argtypes = [vp, u64]
    L.orc_uf_unite.argtypes = [vp, u64, u64]
    L.orc_uf_same.argtypes = [vp, u64, u64]
    L.orc_uf_unite_matching_region.argtypes = [vp, u64, u64, u64, u64, u64, i32, u64]
    L.orc_uf_unite_matching_region_seq2_rc.argtypes = [vp, u64, u64, u64, u64, u64, i32, u64]
    L.orc_uf_unite_calls.restype = u64
    L.orc_uf_unite_calls.argtypes = [vp]
    L.orc_uf_unite_fr_all.argtypes = [vp, u64]
    L.orc_uf_canonical.argtypes = [vp, u64, vp]
    L.orc_uf_count_components.restype = u64
    L.orc_uf_count_components.argtypes = [vp, u64]
    L.orc_process_alignment.argtypes = [vp, C.c_char_p, vp, u64, u64, vp, u64, u64, u64, i32, u64, u64]
    L.orc_run_pairs.argtypes = [vp, vp, vp, vp, C.c_uint32, vp, u64, C.POINTER(i32), u64, i32, vp, C.POINTER(RunStats)]
    L.orc_swg2p_penalty.argtypes = [vp, i32, vp, i32, C.POINTER(i32)]
    L.orc_siphash.restype = u64
    L.orc_siphash.argtypes = [C.c_char_p, u64, u64, u64, i32, i32]
    L.orc_rust_hash_bytes.restype = u64
    L.orc_rust_hash_bytes.argtypes = [C.c_char_p, u64]
    L.orc_rust_hash_str.restype = u64
    L.orc_rust_hash_str.argtypes = [C.c_char_p]
    L.orc_sketch.restype = u64
    L.orc_sketch.argtypes = [vp, u64, u64, u64, vp]
    L.orc_tree_pairs.restype = u64
    L.orc_tree_pairs.argtypes = [vp, vp, vp, C.c_uint32, u64, u64, C.c_double, u64, i32, vp, u64, C.POINTER(u64), vp]
    L.orc_random_pairs.restype = u64
    L.orc_random_pairs.argtypes = [vp, C.c_uint32, i32, C.c_double, vp, u64]
    L.orc_connectivity_probability.restype = C.c_double
    L.orc_connectivity_probability.argtypes = [u64, C.c_double]
    L.orc_build_gfa.restype = vp
    L.orc_build_gfa.argtypes = [vp, vp, vp, C.c_uint32, vp, i32, vp, C.POINTER(u64)]
    L.orc_free.argtypes = [vp]
    _lib = L
    return L


class Oracle:
    """The reference CPU path: edlib orientation + WFA2-lib BiWFA + restated Rust glue."""

    def __init__(self, isa=None):
        self.h = lib().orc_open(ref_path(isa).encode())
        if not self.h:
            raise RuntimeError("cannot load %s (build it with `make -C oracle`)" % ref_path(isa))

    @staticmethod
    def _arr(b):
        return np.frombuffer(b, dtype=np.uint8) if isinstance(b, (bytes, bytearray)) else np.ascontiguousarray(b, dtype=np.uint8)

    def align_pair(self, q, t, pen=(5, 8, 2, 24, 1)):
        """Returns dict(is_reverse, status, score, cigar(bytes), fwd_ed, rev_ed)."""
        q, t = self._arr(q), self._arr(t)
        buf = C.create_string_buffer(q.size + t.size + 16)
        n, rev, score, fe, re_ = C.c_int(), C.c_int(), C.c_int(), C.c_int(), C.c_int()
        p = (C.c_int * 5)(*pen)
        st = lib().orc_align_pair(self.h, q.ctypes.data, q.size, t.ctypes.data, t.size, p, buf, C.byref(n),
                                  C.byref(rev), C.byref(score), C.byref(fe), C.byref(re_))
        return dict(status=st, is_reverse=bool(rev.value), score=score.value, cigar=buf.raw[:n.value],
                    fwd_ed=fe.value, rev_ed=re_.value)

    def wfa_only(self, p, t, pen=(5, 8, 2, 24, 1), memory_mode=3):
        """Raw WFA2-lib call (pattern, text) without the orientation step: (status, score, cigar bytes)."""
        p, t = self._arr(p), self._arr(t)
        buf = C.create_string_buffer(p.size + t.size + 16)
        n, score = C.c_int(), C.c_int()
        st = lib().orc_wfa_only(self.h, p.ctypes.data, p.size, t.ctypes.data, t.size, (C.c_int * 5)(*pen), memory_mode,
                                buf, C.byref(n), C.byref(score))
        return st, score.value, buf.raw[:n.value]

    def wfa_mode(self, p, t, pen6, two_piece=True):
        """BiWFA under any of allwave's modes: pen6 = (match <= 0, x, o1, e1, o2, e2) -> (status, score, cigar)."""
        p, t = self._arr(p), self._arr(t)
        buf = C.create_string_buffer(p.size + t.size + 16)
        n, score = C.c_int(), C.c_int()
        st = lib().orc_wfa_mode(self.h, p.ctypes.data, p.size, t.ctypes.data, t.size, (C.c_int * 6)(*pen6),
                                int(bool(two_piece)), buf, C.byref(n), C.byref(score))
        if st == -100:
            raise RuntimeError("oracle/_ref predates ref_wfa_new_mode: rebuild it with `make -C oracle all`")
        return st, score.value, buf.raw[:n.value]

    def edlib(self, q, t):
        q, t = self._arr(q), self._arr(t)
        return lib().orc_edlib(self.h, q.ctypes.data, q.size, t.ctypes.data, t.size)

    def run_pairs(self, seqs, pairs, pen=(5, 8, 2, 24, 1), min_match_len=0, threads=1, uf=None, want_results=True):
        """The pair-dispatch loop on CPU threads. Returns (results[n,6] or None, stats dict)."""
        arrs = [self._arr(s) for s in seqs]
        ptrs = (C.c_void_p * len(arrs))(*[a.ctypes.data for a in arrs])
        lens = np.array([a.size for a in arrs], dtype=np.uint64)
        pr = np.ascontiguousarray(np.asarray(pairs, dtype=np.uint32).reshape(-1, 2))
        res = np.zeros((pr.shape[0], 6), dtype=np.int32) if want_results else None
        st = RunStats()
        p = (C.c_int * 5)(*pen)
        rc = lib().orc_run_pairs(self.h, uf.h if uf is not None else None, ptrs, lens.ctypes.data, len(arrs),
                                 pr.ctypes.data, pr.shape[0], p, min_match_len, threads,
                                 res.ctypes.data if res is not None else None, C.byref(st))
        if rc != 0:
            raise RuntimeError("oracle: validate_match failed (the reference would panic)")
        return res, {k: getattr(st, k) for k, _ in RunStats._fields_}


class OracleUF:
    """Restated BidirectedUnionFind over a restated uf_rush."""

    def __init__(self, max_offset):
        self.h = lib().orc_uf_new(max_offset)
        self.capacity = lib().orc_uf_capacity(self.h)

    def __del__(self):
        try:
            lib().orc_uf_free(self.h)
        except Exception:
            pass

    def find(self, p):
        return lib().orc_uf_find(self.h, p)

    def unite(self, a, b):
        lib().orc_uf_unite(self.h, a, b)

    def same(self, a, b):
        return bool(lib().orc_uf_same(self.h, a, b))

    def unite_matching_region(self, o1, o2, s1, s2, ln, rc, len1):
        lib().orc_uf_unite_matching_region(self.h, o1, o2, s1, s2, ln, int(bool(rc)), len1)

    def unite_matching_region_seq2_rc(self, o1, o2, s1, s2, ln, rc, len2):
        lib().orc_uf_unite_matching_region_seq2_rc(self.h, o1, o2, s1, s2, ln, int(bool(rc)), len2)

    def unite_fr_all(self, total_len):
        lib().orc_uf_unite_fr_all(self.h, total_len)

    def canonical(self, n=None):
        n = self.capacity if n is None else n
        out = np.empty(n, dtype=np.uint64)
        lib().orc_uf_canonical(self.h, n, out.ctypes.data)
        return out

    def count_components(self, total_len):
        return lib().orc_uf_count_components(self.h, total_len)

    def unite_calls(self):
        return lib().orc_uf_unite_calls(self.h)

    def process_alignment(self, cigar, d1, o1, d2, o2, min_match_len, query_is_rc, qstart=0, tstart=0):
        a, b = Oracle._arr(d1), Oracle._arr(d2)
        return lib().orc_process_alignment(self.h, cigar.encode(), a.ctypes.data, a.size, o1, b.ctypes.data, b.size,
                                           o2, min_match_len, int(bool(query_is_rc)), qstart, tstart)


def siphash(data: bytes, k0=0, k1=0, c=1, d=3):
    return lib().orc_siphash(data, len(data), k0, k1, c, d)


def sketch(seq, k, sketch_size=1000):
    a = Oracle._arr(seq)
    out = np.zeros(sketch_size, dtype=np.uint64)
    n = lib().orc_sketch(a.ctypes.data, a.size, k, sketch_size, out.ctypes.data)
    return out[:n]


def tree_pairs(seqs, ids, k_nearest, k_farthest, random_fraction, kmer_size, separated=False):
    """extract_tree_pairs[_separated] restated: (pairs, n_tree, distance matrix)."""
    arrs = [Oracle._arr(s) for s in seqs]
    n = len(arrs)
    ptrs = (C.c_void_p * n)(*[a.ctypes.data for a in arrs])
    lens = np.array([a.size for a in arrs], dtype=np.uint64)
    names = (C.c_char_p * n)(*[i.encode() for i in ids])
    out = np.zeros((n * n + 1, 2), dtype=np.uint32)
    nt = C.c_uint64()
    dist = np.zeros((n, n), dtype=np.float64)
    m = lib().orc_tree_pairs(ptrs, lens.ctypes.data, names, n, k_nearest, k_farthest, random_fraction, kmer_size,
                             int(bool(separated)), out.ctypes.data, out.shape[0], C.byref(nt), dist.ctypes.data)
    return out[:m].copy(), nt.value, dist


def random_pairs(ids, exclude_self, keep_fraction):
    n = len(ids)
    names = (C.c_char_p * n)(*[i.encode() for i in ids])
    out = np.zeros((n * n + 1, 2), dtype=np.uint32)
    m = lib().orc_random_pairs(names, n, int(bool(exclude_self)), keep_fraction, out.ctypes.data, out.shape[0])
    return out[:m].copy()


def build_gfa(uf, seqs, ids, no_compact=False):
    """Restated write_bidirected_gfa (--skip-sgd --skip-groom --skip-topo): build from the union-find,
    compact + renumber unless no_compact, validate, write (L lines sorted).
    Returns (gfa text, representative Pos of every position = what fixes the node bases, #nodes before compaction)."""
    arrs = [Oracle._arr(s) for s in seqs]
    ptrs = (C.c_void_p * len(arrs))(*[a.ctypes.data for a in arrs])
    lens = np.array([a.size for a in arrs], dtype=np.uint64)
    names = (C.c_char_p * len(ids))(*[i.encode() for i in ids])
    rep = np.zeros(int(lens.sum()), dtype=np.uint64)
    nn = C.c_uint64()
    p = lib().orc_build_gfa(uf.h, ptrs, lens.ctypes.data, len(arrs), names, int(bool(no_compact)), rep.ctypes.data, C.byref(nn))
    if not p:
        raise RuntimeError("oracle: path validation failed (validate_paths_match_sequences)")
    txt = C.string_at(p).decode()
    lib().orc_free(p)
    return txt, rep, nn.value


def swg2p_penalty(p, t, pen=(5, 8, 2, 24, 1)):
    a, b = Oracle._arr(p), Oracle._arr(t)
    return lib().orc_swg2p_penalty(a.ctypes.data, a.size, b.ctypes.data, b.size, (C.c_int * 5)(*pen))


def cigar_bytes_to_string(b):
    out = C.create_string_buffer(16 * len(b) + 16)
    n = lib().orc_cigar_bytes_to_string(b, len(b), out, len(out))
    return out.raw[:n].decode()
