"""ctypes binding of the C ABI in include/seqrush_b200.h.

The shared library is built in-tree (``seqrush_b200/libseqrush_b200.so``) by ``build()`` with
``nvcc -gencode arch=compute_100a,code=sm_100a``. There is no CPU fallback: importing the binding
without the library, or calling it without a CUDA device, raises.
"""
import ctypes as C
import os
import subprocess

_HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(_HERE)
LIB_PATH = os.environ.get("SRGPU_LIB") or os.path.join(_HERE, "libseqrush_b200.so")
CSRC = os.path.join(_HERE, "csrc")
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
              "-shared", "-Xcompiler", "-fPIC", "-diag-suppress", "550"]

# every symbol include/seqrush_b200.h declares (checked by tests/test_abi.py)
EXPORTS = [
    "srgpu_last_error", "srgpu_ctx_create", "srgpu_ctx_destroy", "srgpu_ctx_synchronize", "srgpu_ctx_stream",
    "srgpu_ctx_configure", "srgpu_ctx_set_split", "srgpu_uf_compress_async", "srgpu_uf_merge_gathered",
    "srgpu_ctx_stats", "srgpu_set_sequences", "srgpu_total_length", "srgpu_align_pairs", "srgpu_free",
    "srgpu_align_and_unite", "srgpu_align_and_unite_device", "srgpu_align_pairs_and_unite",
    "srgpu_process_alignment", "srgpu_process_alignments", "srgpu_uf_new",
    "srgpu_uf_free", "srgpu_uf_capacity", "srgpu_uf_clear", "srgpu_uf_find", "srgpu_uf_same", "srgpu_uf_unite",
    "srgpu_uf_unite_matching_region", "srgpu_uf_unite_matching_region_seq2_rc", "srgpu_uf_unite_regions",
    "srgpu_uf_unite_fr_all", "srgpu_uf_roots", "srgpu_uf_count_components", "srgpu_uf_checksum",
    "srgpu_uf_device_ptr", "srgpu_uf_compress", "srgpu_uf_merge_roots",
    "srgpu_uf_count_components_static", "srgpu_uf_count_components_static_max_capacity",
    "srgpu_rust_hash_bytes", "srgpu_rust_hash_str", "srgpu_mash_sketch_jaccard", "srgpu_pair_hash_filter",
    "srgpu_graph_build", "srgpu_graph_from_parts", "srgpu_graph_compact", "srgpu_graph_node_count", "srgpu_graph_edge_count",
    "srgpu_graph_path_count", "srgpu_graph_path_length", "srgpu_graph_path_steps", "srgpu_graph_node_sequence",
    "srgpu_graph_edges", "srgpu_graph_gfa_text", "srgpu_graph_write_gfa", "srgpu_graph_stats", "srgpu_graph_free",
]


def _sources():
    return [os.path.join(CSRC, f) for f in sorted(os.listdir(CSRC))] + [os.path.join(ROOT, "include", "seqrush_b200.h")]


def build(force=False, verbose=False):
    """Compile the CUDA library for sm_100a (cross-compiles without a GPU)."""
    if not force and os.path.exists(LIB_PATH):
        newest = max(os.path.getmtime(s) for s in _sources())
        if os.path.getmtime(LIB_PATH) >= newest:
            return LIB_PATH
    cmd = ["nvcc"] + NVCC_FLAGS + (["-Xptxas", "-v"] if verbose else []) + ["-o", LIB_PATH, os.path.join(CSRC, "capi.cu")]
    r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if r.returncode != 0:
        raise RuntimeError("nvcc failed:\n" + r.stdout)
    if verbose:
        print(r.stdout)
    return LIB_PATH


class SeqView(C.Structure):
    _fields_ = [("data", C.c_void_p), ("len", C.c_uint64)]


class Penalties(C.Structure):
    _fields_ = [("match_score", C.c_int32), ("mismatch", C.c_int32), ("gap_open1", C.c_int32),
                ("gap_ext1", C.c_int32), ("gap_open2", C.c_int32), ("gap_ext2", C.c_int32)]


class Result(C.Structure):
    _fields_ = [("query_idx", C.c_uint32), ("target_idx", C.c_uint32), ("is_reverse", C.c_int32),
                ("status", C.c_int32), ("score", C.c_int32), ("penalty", C.c_int32),
                ("num_matches", C.c_uint32), ("alignment_length", C.c_uint32), ("query_end", C.c_uint32),
                ("target_end", C.c_uint32), ("edit_distance", C.c_uint32), ("n_runs", C.c_uint32),
                ("cigar_off", C.c_uint64), ("cigar_len", C.c_uint64), ("wf_cells", C.c_uint64)]


class Stats(C.Structure):
    _fields_ = [("pairs", C.c_uint64), ("aligned_bp", C.c_uint64), ("wf_cells", C.c_uint64),
                ("orient_cells", C.c_uint64), ("unite_ops", C.c_uint64), ("unite_links", C.c_uint64),
                ("runs", C.c_uint64), ("kernel_launches", C.c_uint32), ("align_launches", C.c_uint32),
                ("union_launches", C.c_uint32), ("retries", C.c_uint32), ("align_ms", C.c_float),
                ("union_ms", C.c_float), ("total_ms", C.c_float), ("orient_ms", C.c_float),
                ("failed_pairs", C.c_uint32), ("reserved0", C.c_uint32)]


class Region(C.Structure):
    _fields_ = [("seq1_offset", C.c_uint64), ("seq2_offset", C.c_uint64), ("seq1_local_start", C.c_uint64),
                ("seq2_local_start", C.c_uint64), ("match_length", C.c_uint64), ("seq1_len", C.c_uint64),
                ("seq1_is_rc", C.c_int32), ("seq2_rc_form", C.c_int32)]


class PafRecord(C.Structure):
    _fields_ = [("query_idx", C.c_uint32), ("target_idx", C.c_uint32), ("query_is_rc", C.c_int32),
                ("reserved", C.c_int32), ("query_start", C.c_uint64), ("target_start", C.c_uint64),
                ("cigar", C.c_char_p)]


_lib = None


class SrgpuError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__("srgpu error %d: %s" % (code, msg))
        self.code = code


def lib():
    """Load the CUDA library; fails loudly when it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError("seqrush_b200: %s is missing -- run `python -c 'import __graft_entry__ as g; g.build()'` "
                           "(there is no CPU fallback)" % LIB_PATH)
    L = C.CDLL(LIB_PATH)
    u64, u32, i32, vp = C.c_uint64, C.c_uint32, C.c_int, C.c_void_p
    L.srgpu_last_error.restype = C.c_char_p
    L.srgpu_ctx_create.argtypes = [i32, vp, C.POINTER(vp)]
    L.srgpu_ctx_destroy.argtypes = [vp]
    L.srgpu_ctx_destroy.restype = None
    L.srgpu_ctx_synchronize.argtypes = [vp]
    L.srgpu_ctx_configure.argtypes = [vp, i32, i32, i32]
    L.srgpu_ctx_set_split.argtypes = [vp, i32]
    L.srgpu_ctx_stats.argtypes = [vp, C.POINTER(Stats)]
    L.srgpu_set_sequences.argtypes = [vp, C.POINTER(SeqView), u32]
    L.srgpu_total_length.argtypes = [vp]
    L.srgpu_total_length.restype = u64
    L.srgpu_align_pairs.argtypes = [vp, vp, u64, C.POINTER(Penalties), C.POINTER(Result), C.POINTER(vp), C.POINTER(u64)]
    L.srgpu_free.argtypes = [vp]
    L.srgpu_free.restype = None
    L.srgpu_align_and_unite.argtypes = [vp, vp, vp, u64, C.POINTER(Penalties), u64, C.POINTER(Result)]
    L.srgpu_align_and_unite_device.argtypes = [vp, vp, vp, u64, C.POINTER(Penalties), u64]
    L.srgpu_process_alignment.argtypes = [vp, vp, C.c_char_p, u32, u32, u64, i32, u64, u64]
    L.srgpu_process_alignments.argtypes = [vp, vp, C.POINTER(PafRecord), u64, u64]
    L.srgpu_align_pairs_and_unite.argtypes = [vp, vp, vp, u64, C.POINTER(Penalties), u64, C.POINTER(Result),
                                              C.POINTER(vp), C.POINTER(u64)]
    L.srgpu_uf_new.argtypes = [vp, u64, C.POINTER(vp)]
    L.srgpu_uf_free.argtypes = [vp]
    L.srgpu_uf_free.restype = None
    L.srgpu_uf_capacity.argtypes = [vp]
    L.srgpu_uf_capacity.restype = u64
    L.srgpu_uf_clear.argtypes = [vp]
    L.srgpu_uf_find.argtypes = [vp, u64, C.POINTER(u64)]
    L.srgpu_uf_same.argtypes = [vp, u64, u64, C.POINTER(i32)]
    L.srgpu_uf_unite.argtypes = [vp, u64, u64]
    L.srgpu_uf_unite_matching_region.argtypes = [vp, u64, u64, u64, u64, u64, i32, u64]
    L.srgpu_uf_unite_matching_region_seq2_rc.argtypes = [vp, u64, u64, u64, u64, u64, i32, u64]
    L.srgpu_uf_unite_regions.argtypes = [vp, C.POINTER(Region), u64]
    L.srgpu_uf_unite_fr_all.argtypes = [vp, u64]
    L.srgpu_uf_roots.argtypes = [vp, u64, u64, vp]
    L.srgpu_uf_count_components.argtypes = [vp, u64, C.POINTER(u64)]
    L.srgpu_uf_checksum.argtypes = [vp, C.POINTER(u64)]
    L.srgpu_uf_count_components_static.argtypes = [vp, u64, C.POINTER(u64)]
    L.srgpu_uf_count_components_static_max_capacity.argtypes = []
    L.srgpu_uf_count_components_static_max_capacity.restype = u64
    L.srgpu_uf_device_ptr.argtypes = [vp]
    L.srgpu_uf_device_ptr.restype = vp
    L.srgpu_uf_compress.argtypes = [vp]
    L.srgpu_uf_merge_roots.argtypes = [vp, vp, u64]
    L.srgpu_uf_compress_async.argtypes = [vp]
    L.srgpu_uf_merge_gathered.argtypes = [vp, vp, u64, i32, i32]
    L.srgpu_ctx_stream.argtypes = [vp]
    L.srgpu_ctx_stream.restype = vp
    L.srgpu_rust_hash_bytes.argtypes = [vp, u64]
    L.srgpu_rust_hash_bytes.restype = u64
    L.srgpu_rust_hash_str.argtypes = [C.c_char_p]
    L.srgpu_rust_hash_str.restype = u64
    L.srgpu_mash_sketch_jaccard.argtypes = [vp, u32, u32, vp, vp, vp, vp]
    L.srgpu_pair_hash_filter.argtypes = [vp, vp, u32, C.c_double, i32, vp, vp]
    L.srgpu_graph_build.argtypes = [vp, vp, vp, C.POINTER(vp)]
    L.srgpu_graph_from_parts.argtypes = [u64, vp, u32, vp, vp, u64, vp, C.POINTER(SeqView), C.POINTER(vp)]
    L.srgpu_graph_compact.argtypes = [vp]
    for nm in ("srgpu_graph_node_count", "srgpu_graph_edge_count", "srgpu_graph_path_count"):
        getattr(L, nm).argtypes = [vp]
        getattr(L, nm).restype = u64
    L.srgpu_graph_path_length.argtypes = [vp, u32]
    L.srgpu_graph_path_length.restype = u64
    L.srgpu_graph_path_steps.argtypes = [vp, u32, vp, u64]
    L.srgpu_graph_node_sequence.argtypes = [vp, u64, vp, u64]
    L.srgpu_graph_node_sequence.restype = C.c_int64
    L.srgpu_graph_edges.argtypes = [vp, vp, u64]
    L.srgpu_graph_gfa_text.argtypes = [vp, vp, u32, C.POINTER(vp), C.POINTER(u64)]
    L.srgpu_graph_write_gfa.argtypes = [vp, vp, u32, C.c_char_p]
    L.srgpu_graph_stats.argtypes = [vp, C.POINTER(u64), C.POINTER(u32), C.POINTER(u32)]
    L.srgpu_graph_free.argtypes = [vp]
    L.srgpu_graph_free.restype = None
    _lib = L
    return L


def check(rc):
    if rc != 0:
        raise SrgpuError(rc, lib().srgpu_last_error().decode("utf-8", "replace"))
