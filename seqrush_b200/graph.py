"""BidirectedGraph -- host-side mirror of the graph the reference builds from the union-find
(src/bidirected_builder.rs:17-289), compacts (src/bidirected_ops.rs:75-478) and writes as GFA
(src/bidirected_ops.rs:880-925; driver src/bidirected_gfa_writer.rs:8-167 with --skip-sgd
--skip-groom --skip-topo). Node numbering, node bases, path steps and the edge set are computed on
the device (srgpu_graph_build); compaction, validation and the writer are native host code."""
import ctypes as C

import numpy as np

from . import _lib


class Handle(int):
    """(node_id << 1) | is_reverse, src/bidirected_graph.rs:6-48."""

    @property
    def node_id(self):
        return int(self) >> 1

    @property
    def is_reverse(self):
        return bool(int(self) & 1)

    def __str__(self):
        return "%d%s" % (self.node_id, "-" if self.is_reverse else "+")


class BidirectedGraph:
    def __init__(self, handle, names):
        self._h = handle
        self.names = list(names)

    @classmethod
    def from_parts(cls, node_bases: bytes, paths, edges, sequences, names):
        """Uncompacted graph from host data: node i+1 has base node_bases[i]; paths = list of handle lists;
        edges = list of (from, to) handles in their stored form; sequences = what the paths must spell."""
        lens = np.array([len(p) for p in paths], dtype=np.uint64)
        steps = np.array([int(h) for p in paths for h in p], dtype=np.uint64)
        ft = np.array([int(x) for e in edges for x in e], dtype=np.uint64)
        keep = [np.frombuffer(bytes(s), dtype=np.uint8) for s in sequences]
        views = (_lib.SeqView * len(keep))()
        for i, a in enumerate(keep):
            views[i].data = a.ctypes.data
            views[i].len = a.size
        nb = np.frombuffer(bytes(node_bases), dtype=np.uint8)
        h = C.c_void_p()
        _lib.check(_lib.lib().srgpu_graph_from_parts(nb.size, nb.ctypes.data, len(paths), lens.ctypes.data,
                                                     steps.ctypes.data, len(edges), ft.ctypes.data, views, C.byref(h)))
        return cls(h, names)

    def close(self):
        if getattr(self, "_h", None) and self._h.value:
            _lib.lib().srgpu_graph_free(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def compact(self):
        """BidirectedGraph::compact followed by renumber_nodes_sequentially (what write_bidirected_gfa does
        unless --no-compact)."""
        _lib.check(_lib.lib().srgpu_graph_compact(self._h))
        return self

    @property
    def node_count(self) -> int:
        return int(_lib.lib().srgpu_graph_node_count(self._h))

    @property
    def edge_count(self) -> int:
        return int(_lib.lib().srgpu_graph_edge_count(self._h))

    def path(self, i: int):
        n = int(_lib.lib().srgpu_graph_path_length(self._h, i))
        out = np.empty(n, dtype=np.uint64)
        _lib.check(_lib.lib().srgpu_graph_path_steps(self._h, i, out.ctypes.data, n))
        return [Handle(int(x)) for x in out]

    def node_sequence(self, node_id: int) -> bytes:
        n = int(_lib.lib().srgpu_graph_node_sequence(self._h, node_id, None, 0))
        if n < 0:
            raise KeyError(node_id)
        buf = C.create_string_buffer(max(1, n))
        _lib.lib().srgpu_graph_node_sequence(self._h, node_id, buf, n)
        return buf.raw[:n]

    def edges(self):
        n = self.edge_count
        out = np.empty(2 * n, dtype=np.uint64)
        _lib.check(_lib.lib().srgpu_graph_edges(self._h, out.ctypes.data, n))
        return [(Handle(int(out[2 * i])), Handle(int(out[2 * i + 1]))) for i in range(n)]

    def stats(self) -> dict:
        a, b, c = C.c_uint64(), C.c_uint32(), C.c_uint32()
        _lib.check(_lib.lib().srgpu_graph_stats(self._h, C.byref(a), C.byref(b), C.byref(c)))
        return {"uncompacted_nodes": a.value, "compaction_rounds": b.value, "merged_chains": c.value}

    def _names(self):
        return (C.c_char_p * len(self.names))(*[n.encode() for n in self.names])

    def gfa_text(self) -> str:
        """validate_paths_match_sequences + write_gfa (L lines sorted)."""
        t, n = C.c_void_p(), C.c_uint64()
        _lib.check(_lib.lib().srgpu_graph_gfa_text(self._h, self._names(), len(self.names), C.byref(t), C.byref(n)))
        s = C.string_at(t, n.value).decode()
        _lib.lib().srgpu_free(t)
        return s

    def write_gfa(self, path: str):
        _lib.check(_lib.lib().srgpu_graph_write_gfa(self._h, self._names(), len(self.names), path.encode()))


def build_bidirected_graph(seqrush, representatives=None) -> BidirectedGraph:
    """SeqRush::build_bidirected_graph_with_options(verbose, apply_sort=false) on the device.
    representatives: optional uint64 array (one Pos per position) that fixes the node bases -- see
    include/seqrush_b200.h; None = the library's canonical first-seen representative."""
    rep = None
    if representatives is not None:
        rep = np.ascontiguousarray(representatives, dtype=np.uint64)
        if rep.size != seqrush.total_length:
            raise ValueError("representatives: one Pos per position expected")
    h = C.c_void_p()
    _lib.check(_lib.lib().srgpu_graph_build(seqrush.ctx._h, seqrush.union_find._h, rep.ctypes.data if rep is not None else None,
                                            C.byref(h)))
    return BidirectedGraph(h, [s.id for s in seqrush.sequences])
