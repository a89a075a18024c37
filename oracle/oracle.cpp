// oracle/oracle.cpp -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
//
// CPU restatement of the reference's Rust glue on the align-and-unite hot path, driving the
// reference's own native arithmetic (WFA2-lib + edlib compiled from /root/reference into
// oracle/_ref/libsrref_<isa>.so by oracle/Makefile). The Rust crates cannot be built in this
// environment (no cargo/rustc), so every function below restates one Rust function and cites it.
// Parity status: the WFA2-lib arithmetic is pinned by its own 305 BiWFA gap-affine-2p golden
// vectors (tests/golden/); the Rust glue restated here is pinned by the reference's unit-test
// known answers (src/bidirected_union_find.rs:191-257, src/pos.rs:89-160,
// tests/test_wfa2_cigar_debug.rs:29). lib_wfa2@819b82cd (the Rust FFI shim) is not in the tree:
// it is assumed to wrap the same WFA2-lib as the top-level submodule -- "parity unpinned" at that
// boundary, see DESIGN.md.
//
// Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
// load this library.
#include <dlfcn.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#include <math.h>
#include <atomic>
#include <string>
#include <thread>
#include <vector>
#include <chrono>

typedef void* (*fn_wfa_new)(int, int, int, int, int, int);
typedef void (*fn_wfa_free)(void*);
typedef int (*fn_wfa_align)(void*, const char*, int, const char*, int, char*, int*, int*);
typedef int (*fn_edlib)(const char*, int, const char*, int);

struct Oracle {
  void* dl = nullptr;
  fn_wfa_new wfa_new = nullptr;
  fn_wfa_free wfa_free = nullptr;
  fn_wfa_align wfa_align = nullptr;
  fn_edlib edlib = nullptr;
};

extern "C" void* orc_open(const char* ref_so) {
  void* dl = dlopen(ref_so, RTLD_NOW | RTLD_LOCAL);
  if (!dl) return nullptr;
  Oracle* o = new Oracle();
  o->dl = dl;
  o->wfa_new = (fn_wfa_new)dlsym(dl, "ref_wfa_new");
  o->wfa_free = (fn_wfa_free)dlsym(dl, "ref_wfa_free");
  o->wfa_align = (fn_wfa_align)dlsym(dl, "ref_wfa_align");
  o->edlib = (fn_edlib)dlsym(dl, "ref_edlib_nw_distance");
  if (!o->wfa_new || !o->wfa_free || !o->wfa_align || !o->edlib) {
    delete o;
    return nullptr;
  }
  return o;
}
extern "C" void orc_close(void* h) {
  Oracle* o = (Oracle*)h;
  if (!o) return;
  delete o;  // the shared object stays mapped (harmless for a test process)
}

// allwave/src/alignment.rs:174-186
static void allwave_rc(const uint8_t* s, size_t n, std::vector<uint8_t>& out) {
  out.resize(n);
  for (size_t i = 0; i < n; ++i) {
    const uint8_t b = s[n - 1 - i];
    uint8_t c;
    switch (b) {
      case 'A': case 'a': c = 'T'; break;
      case 'T': case 't': c = 'A'; break;
      case 'C': case 'c': c = 'G'; break;
      case 'G': case 'g': c = 'C'; break;
      default: c = 'N'; break;
    }
    out[i] = c;
  }
}
extern "C" void orc_allwave_rc(const uint8_t* s, uint64_t n, uint8_t* out) {
  std::vector<uint8_t> v;
  allwave_rc(s, n, v);
  memcpy(out, v.data(), n);
}

// ---------------------------------------------------------------------------
// allwave::align_pair (allwave/src/alignment.rs:16-57): orientation by two exact edlib NW
// distances (:148-171, forward iff fwd <= rev), then BiWFA through WFA2-lib (:189-276).
// cigar_out needs qlen+tlen bytes. Returns the WFA status (0 = completed).
// ---------------------------------------------------------------------------
struct PairRes {
  int is_reverse, status, score, fwd_ed, rev_ed;
  int cigar_len;
};
static int align_pair_impl(Oracle* o, void* wf, const uint8_t* q, int qlen, const uint8_t* t, int tlen, char* cigar_out,
                           PairRes* r, double* t_edlib, double* t_wfa) {
  auto t0 = std::chrono::steady_clock::now();
  std::vector<uint8_t> rc;
  allwave_rc(q, (size_t)qlen, rc);
  r->fwd_ed = o->edlib((const char*)q, qlen, (const char*)t, tlen);
  r->rev_ed = o->edlib((const char*)rc.data(), qlen, (const char*)t, tlen);
  r->is_reverse = !(r->fwd_ed <= r->rev_ed);
  auto t1 = std::chrono::steady_clock::now();
  const uint8_t* qq = r->is_reverse ? rc.data() : q;
  int n = 0, score = 0;
  r->status = o->wfa_align(wf, (const char*)qq, qlen, (const char*)t, tlen, cigar_out, &n, &score);
  auto t2 = std::chrono::steady_clock::now();
  if (r->status != 0) {  // allwave/src/alignment.rs:40-55
    n = 0;
    score = INT32_MAX;
  }
  r->cigar_len = n;
  r->score = score;
  if (t_edlib) *t_edlib += std::chrono::duration<double>(t1 - t0).count();
  if (t_wfa) *t_wfa += std::chrono::duration<double>(t2 - t1).count();
  return r->status;
}

extern "C" int orc_align_pair(void* h, const uint8_t* q, int qlen, const uint8_t* t, int tlen, const int* pen5,
                              char* cigar_out, int* cigar_len, int* is_reverse, int* score, int* fwd_ed, int* rev_ed) {
  Oracle* o = (Oracle*)h;
  void* wf = o->wfa_new(pen5[0], pen5[1], pen5[2], pen5[3], pen5[4], 3);
  PairRes r;
  align_pair_impl(o, wf, q, qlen, t, tlen, cigar_out, &r, nullptr, nullptr);
  o->wfa_free(wf);
  *cigar_len = r.cigar_len;
  *is_reverse = r.is_reverse;
  *score = r.score;
  *fwd_ed = r.fwd_ed;
  *rev_ed = r.rev_ed;
  return r.status;
}

// raw BiWFA call without the orientation step (used to replay WFA2-lib's own golden vectors,
// WFA2-lib/tests/wfa.utest.check/test.biwfa.affine2p.alg). memory_mode 3 = ultralow (BiWFA), 0 = high.
extern "C" int orc_wfa_only(void* h, const uint8_t* p, int plen, const uint8_t* t, int tlen, const int* pen5,
                            int memory_mode, char* cigar_out, int* cigar_len, int* score) {
  Oracle* o = (Oracle*)h;
  void* wf = o->wfa_new(pen5[0], pen5[1], pen5[2], pen5[3], pen5[4], memory_mode);
  const int st = o->wfa_align(wf, (const char*)p, plen, (const char*)t, tlen, cigar_out, cigar_len, score);
  o->wfa_free(wf);
  return st;
}
// the same for any of allwave's modes (allwave/src/alignment.rs:203-231): pen6 = match (<= 0), x, o1, e1, o2, e2
typedef void* (*fn_wfa_new_mode)(int, int, int, int, int, int, int, int);
extern "C" int orc_wfa_mode(void* h, const uint8_t* p, int plen, const uint8_t* t, int tlen, const int* pen6, int two_piece,
                            char* cigar_out, int* cigar_len, int* score) {
  Oracle* o = (Oracle*)h;
  fn_wfa_new_mode mk = (fn_wfa_new_mode)dlsym(o->dl, "ref_wfa_new_mode");
  if (!mk) return -100;
  void* wf = mk(pen6[0], pen6[1], pen6[2], pen6[3], pen6[4], pen6[5], two_piece, 3);
  const int st = o->wfa_align(wf, (const char*)p, plen, (const char*)t, tlen, cigar_out, cigar_len, score);
  o->wfa_free(wf);
  return st;
}
extern "C" int orc_edlib(void* h, const uint8_t* q, int qlen, const uint8_t* t, int tlen) {
  return ((Oracle*)h)->edlib((const char*)q, qlen, (const char*)t, tlen);
}

// allwave::cigar_bytes_to_string (allwave/src/alignment.rs:327-356): RLE, M->'=', I<->D swapped
static void cigar_bytes_to_string(const char* b, int n, std::string& out) {
  out.clear();
  int i = 0;
  while (i < n) {
    const char op = b[i];
    int j = i + 1;
    while (j < n && b[j] == op) ++j;
    char c = '?';
    if (op == 'M') c = '='; else if (op == 'X') c = 'X'; else if (op == 'I') c = 'D'; else if (op == 'D') c = 'I';
    out += std::to_string(j - i);
    out += c;
    i = j;
  }
}
extern "C" int orc_cigar_bytes_to_string(const char* b, int n, char* out, int out_cap) {
  std::string s;
  cigar_bytes_to_string(b, n, s);
  if ((int)s.size() + 1 > out_cap) return -1;
  memcpy(out, s.c_str(), s.size() + 1);
  return (int)s.size();
}

// ---------------------------------------------------------------------------
// uf_rush 0.2.1 (uf_rush-0.2.1/src/lib.rs): 58-bit parent + 6-bit rank per node (:228-240),
// path-halving find (:112-134), union by rank with index tie-break (:159-209). Atomic so the
// multi-threaded CPU baseline exercises the same CAS protocol as the reference.
// ---------------------------------------------------------------------------
struct UFRush {
  std::vector<std::atomic<uint64_t>> nodes;
  explicit UFRush(size_t n) : nodes(n) {
    for (size_t i = 0; i < n; ++i) nodes[i].store((uint64_t)i, std::memory_order_relaxed);
  }
  static uint64_t parent(uint64_t n) { return n & ((1ull << 58) - 1); }
  static uint64_t rank(uint64_t n) { return n >> 58; }
  static uint64_t encode(uint64_t p, uint64_t r) { return p | (r << 58); }
  uint64_t find(uint64_t x) {
    uint64_t xn = nodes[x].load(std::memory_order_relaxed);
    while (x != parent(xn)) {
      const uint64_t xp = parent(xn);
      const uint64_t xpn = nodes[xp].load(std::memory_order_relaxed);
      const uint64_t xpp = parent(xpn);
      uint64_t expect = xn;
      nodes[x].compare_exchange_weak(expect, encode(xpp, rank(xn)), std::memory_order_release, std::memory_order_relaxed);
      x = xpp;
      xn = nodes[x].load(std::memory_order_relaxed);
    }
    return x;
  }
  bool unite(uint64_t x, uint64_t y) {
    while (true) {
      uint64_t xr = find(x), yr = find(y);
      if (xr == yr) return false;
      const uint64_t xn = nodes[xr].load(std::memory_order_relaxed), yn = nodes[yr].load(std::memory_order_relaxed);
      uint64_t xk = rank(xn), yk = rank(yn);
      if (xk > yk || (xk == yk && xr > yr)) {
        std::swap(xr, yr);
        std::swap(xk, yk);
      }
      uint64_t cur = encode(xr, xk);
      if (nodes[xr].compare_exchange_strong(cur, encode(yr, xk), std::memory_order_release, std::memory_order_acquire)) {
        if (xk == yk) {
          uint64_t c2 = encode(yr, yk);
          nodes[yr].compare_exchange_weak(c2, encode(yr, yk + 1), std::memory_order_release, std::memory_order_relaxed);
        }
        return true;
      }
    }
  }
};

// BidirectedUnionFind (src/bidirected_union_find.rs:7-135)
struct BUF {
  UFRush uf;
  std::atomic<uint64_t> unite_calls{0};
  explicit BUF(uint64_t max_offset) : uf((max_offset << 1) + 2) {}  // :16-24
  void unite(uint64_t a, uint64_t b) {                              // :35-43
    unite_calls.fetch_add(1, std::memory_order_relaxed);
    if (a != b) uf.unite(a, b);
  }
  // :60-98
  void unite_matching_region(uint64_t o1, uint64_t o2, uint64_t s1, uint64_t s2, uint64_t len, bool rc, uint64_t len1) {
    for (uint64_t i = 0; i < len; ++i) {
      if (rc) {
        const uint64_t fwd_local = len1 - 1 - (s1 + i);
        unite(((o1 + fwd_local) << 1) | 1, (o2 + s2 + i) << 1);
      } else {
        unite((o1 + s1 + i) << 1, (o2 + s2 + i) << 1);
      }
    }
  }
  // :103-129
  void unite_matching_region_seq2_rc(uint64_t o1, uint64_t o2, uint64_t s1, uint64_t s2, uint64_t len, bool rc, uint64_t len2) {
    for (uint64_t i = 0; i < len; ++i) {
      if (rc) {
        unite((o1 + s1 + i) << 1, ((o2 + (len2 - 1 - (s2 + i))) << 1) | 1);
      } else {
        unite((o1 + s1 + i) << 1, (o2 + s2 + i) << 1);
      }
    }
  }
};

extern "C" void* orc_uf_new(uint64_t max_offset) { return new BUF(max_offset); }
extern "C" void orc_uf_free(void* u) { delete (BUF*)u; }
extern "C" uint64_t orc_uf_capacity(void* u) { return ((BUF*)u)->uf.nodes.size(); }
extern "C" uint64_t orc_uf_find(void* u, uint64_t p) { return ((BUF*)u)->uf.find(p); }
extern "C" void orc_uf_unite(void* u, uint64_t a, uint64_t b) { ((BUF*)u)->unite(a, b); }
extern "C" int orc_uf_same(void* u, uint64_t a, uint64_t b) {  // :46-54
  if (a == b) return 1;
  BUF* f = (BUF*)u;
  return f->uf.find(a) == f->uf.find(b);
}
extern "C" void orc_uf_unite_matching_region(void* u, uint64_t o1, uint64_t o2, uint64_t s1, uint64_t s2, uint64_t len,
                                             int rc, uint64_t len1) {
  ((BUF*)u)->unite_matching_region(o1, o2, s1, s2, len, rc != 0, len1);
}
extern "C" void orc_uf_unite_matching_region_seq2_rc(void* u, uint64_t o1, uint64_t o2, uint64_t s1, uint64_t s2,
                                                     uint64_t len, int rc, uint64_t len2) {
  ((BUF*)u)->unite_matching_region_seq2_rc(o1, o2, s1, s2, len, rc != 0, len2);
}
extern "C" uint64_t orc_uf_unite_calls(void* u) { return ((BUF*)u)->unite_calls.load(); }
// SeqRush::new's loop (src/seqrush.rs:324-328)
extern "C" void orc_uf_unite_fr_all(void* u, uint64_t total_len) {
  BUF* f = (BUF*)u;
  for (uint64_t i = 0; i < total_len; ++i) if (2 * i != 2 * i + 1) f->uf.unite(2 * i, 2 * i + 1);
}
// canonical labelling: min Pos of each component (the comparison rule of SURVEY.md 8(c).3)
extern "C" void orc_uf_canonical(void* u, uint64_t n, uint64_t* out) {
  BUF* f = (BUF*)u;
  std::vector<uint64_t> minpos(f->uf.nodes.size(), UINT64_MAX);
  for (uint64_t p = 0; p < n; ++p) {
    const uint64_t r = f->uf.find(p);
    if (p < minpos[r]) minpos[r] = p;
  }
  for (uint64_t p = 0; p < n; ++p) out[p] = minpos[f->uf.find(p)];
}
// count_components (src/seqrush.rs:341-353)
extern "C" uint64_t orc_uf_count_components(void* u, uint64_t total_len) {
  BUF* f = (BUF*)u;
  std::vector<uint8_t> seen(f->uf.nodes.size(), 0);
  uint64_t c = 0;
  for (uint64_t i = 0; i < total_len; ++i) {
    const uint64_t r = f->uf.find(2 * i);
    if (!seen[r]) {
      seen[r] = 1;
      ++c;
    }
  }
  return c;
}

// ---------------------------------------------------------------------------
// SeqRush::process_alignment (src/seqrush.rs:1134-1481). Returns -1 where the reference panics
// in validate_match (:1184-1205), else 0.
// ---------------------------------------------------------------------------
struct SeqV {
  const uint8_t* data;
  uint64_t len;
  uint64_t offset;
};
static inline uint8_t get_query_base(const SeqV& s, uint64_t local_pos, bool rc) {  // :1162-1176
  if (!rc) return s.data[local_pos];
  const uint8_t b = s.data[s.len - 1 - local_pos];
  switch (b) {
    case 'A': case 'a': return 'T';
    case 'T': case 't': return 'A';
    case 'C': case 'c': return 'G';
    case 'G': case 'g': return 'C';
    default: return b;
  }
}
static int process_alignment(BUF* uf, const char* cigar, const SeqV& s1, const SeqV& s2, uint64_t min_match_len,
                             bool query_is_rc, uint64_t query_start, uint64_t target_start) {
  uint64_t pos1 = query_start, pos2 = target_start, count = 0;
  bool in_run = false;
  uint64_t rs1 = 0, rs2 = 0, rl = 0;
  int rcode = 0;
  auto flush = [&]() {
    if (in_run && rl >= min_match_len) {
      for (uint64_t i = 0; i < rl; ++i)  // validate_match :1179-1207
        if (get_query_base(s1, rs1 + i, query_is_rc) != s2.data[rs2 + i]) rcode = -1;
      uf->unite_matching_region(s1.offset, s2.offset, rs1, rs2, rl, query_is_rc, s1.len);
    }
  };
  for (const char* p = cigar; *p; ++p) {
    const char ch = *p;
    if (ch >= '0' && ch <= '9') {
      count = count * 10 + (uint64_t)(ch - '0');
      continue;
    }
    if (count == 0) count = 1;
    if (ch == 'M' || ch == '=') {
      for (uint64_t k = 0; k < count; ++k) {
        if (pos1 + k < s1.len && pos2 + k < s2.len) {
          if (get_query_base(s1, pos1 + k, query_is_rc) == s2.data[pos2 + k]) {
            if (!in_run) {
              in_run = true;
              rs1 = pos1 + k;
              rs2 = pos2 + k;
              rl = 1;
            } else {
              ++rl;
            }
          } else {
            flush();
            in_run = false;
            rl = 0;
          }
        }
      }
      pos1 += count;
      pos2 += count;
    } else {
      flush();
      in_run = false;
      rl = 0;
      if (ch == 'X') {
        pos1 += count;
        pos2 += count;
      } else if (ch == 'I') {
        pos1 += count;
      } else if (ch == 'D') {
        pos2 += count;
      }
    }
    count = 0;
  }
  flush();
  return rcode;
}
extern "C" int orc_process_alignment(void* u, const char* cigar, const uint8_t* d1, uint64_t l1, uint64_t o1,
                                     const uint8_t* d2, uint64_t l2, uint64_t o2, uint64_t min_match_len, int query_is_rc,
                                     uint64_t qstart, uint64_t tstart) {
  SeqV a{d1, l1, o1}, b{d2, l2, o2};
  return process_alignment((BUF*)u, cigar, a, b, min_match_len, query_is_rc != 0, qstart, tstart);
}

// ---------------------------------------------------------------------------
// Graph build from the union-find, compaction and GFA text -- literal restatement of
//   SeqRush::build_bidirected_graph_with_options (src/bidirected_builder.rs:17-289, apply_sort=false),
//   BidirectedGraph::{add_node, add_edge, build_path, verify_path_edges, compact, find_simple_components,
//   merge_component_v2, renumber_nodes_sequentially, apply_node_id_mapping, write_gfa}
//   (src/bidirected_ops.rs:22-89, 91-478, 804-825, 871-925, 1049-1068), Handle / BiEdge / reverse_complement
//   (src/bidirected_graph.rs), and the driver write_bidirected_gfa (src/bidirected_gfa_writer.rs:8-167) with
//   --skip-sgd --skip-groom --skip-topo (the Ygs phases are then no-ops).
// Quadratic where the reference is quadratic (key scan :96-119, are_perfect_neighbors scanning every path per
// candidate, merge_component_v2 scanning every path and edge per chain): small inputs and configs[0] only.
// HashMap/HashSet iteration order is unspecified in Rust; it is observable in exactly two places: the key scan of
// the builder (first match wins -- on the production path nothing ever matches, distinct roots are distinct
// components) and the order of the L lines in write_gfa, which are therefore emitted SORTED here and compared
// sorted everywhere (SURVEY.md 8(c) rule 4).
// ---------------------------------------------------------------------------
#include <algorithm>
#include <map>
#include <set>
#include <unordered_map>
#include <unordered_set>
namespace ograph {
typedef uint64_t Handle;  // (node_id << 1) | is_reverse, src/bidirected_graph.rs:6-48
static inline Handle mk(uint64_t id, bool rev) { return (id << 1) | (rev ? 1u : 0u); }
static inline uint64_t nid(Handle h) { return h >> 1; }
static inline bool isrev(Handle h) { return (h & 1) != 0; }
static inline Handle flip(Handle h) { return h ^ 1; }
typedef std::pair<Handle, Handle> Edge;
static std::vector<uint8_t> revcomp(const std::vector<uint8_t>& s) {  // src/bidirected_graph.rs:58-71
  std::vector<uint8_t> o(s.rbegin(), s.rend());
  for (auto& b : o) {
    switch (b) {
      case 'A': case 'a': b = 'T'; break;
      case 'T': case 't': b = 'A'; break;
      case 'C': case 'c': b = 'G'; break;
      case 'G': case 'g': b = 'C'; break;
      case 'N': case 'n': b = 'N'; break;
      default: break;
    }
  }
  return o;
}
struct Graph {
  std::vector<std::vector<uint8_t>> seq;  // index = node id
  std::vector<char> present;
  std::set<Edge> edges;  // a set; iteration order is never relied on
  std::vector<std::pair<std::string, std::vector<Handle>>> paths;
  void add_node(uint64_t id, std::vector<uint8_t> s) {  // :804-810
    if (id >= seq.size()) {
      seq.resize(id + 1);
      present.resize(id + 1, 0);
    }
    seq[id] = std::move(s);
    present[id] = 1;
  }
  void add_edge(Handle from, Handle to) {  // :813-825
    const Edge e(from, to), c(flip(to), flip(from));
    if (!edges.count(e) && !edges.count(c)) edges.insert(e);
  }
  void verify_path_edges() {  // :1049-1068
    for (auto& p : paths)
      for (size_t i = 0; i + 1 < p.second.size(); ++i) {
        const Edge e(p.second[i], p.second[i + 1]), c(flip(p.second[i + 1]), flip(p.second[i]));
        if (!edges.count(e) && !edges.count(c)) edges.insert(e);
      }
  }
  // :113-275
  std::vector<std::vector<Handle>> find_simple_components() const {
    std::vector<std::vector<Handle>> components;
    std::unordered_set<Handle> visited;
    std::unordered_set<uint64_t> merged_nodes;
    std::unordered_map<Handle, std::vector<Handle>> fwd, bwd;
    for (const Edge& e : edges) {
      fwd[e.first].push_back(e.second);
      bwd[e.second].push_back(e.first);
      fwd[flip(e.second)].push_back(flip(e.first));
      bwd[flip(e.first)].push_back(flip(e.second));
    }
    auto deg = [](const std::unordered_map<Handle, std::vector<Handle>>& m, Handle h) -> size_t {
      auto it = m.find(h);
      return it == m.end() ? 0 : it->second.size();
    };
    auto perfect = [&](Handle from, Handle to) -> bool {  // :141-190
      for (auto& p : paths) {
        const std::vector<Handle>& st = p.second;
        size_t visits = 0, trans = 0;
        for (size_t i = 0; i < st.size(); ++i)
          if (st[i] == from) {
            ++visits;
            if (i + 1 < st.size()) {
              if (st[i + 1] == to) ++trans;
              else return false;
            } else {
              return false;
            }
          }
        if (visits > 0 && visits != trans) return false;
        const Handle fr = flip(from), tr = flip(to);
        size_t v2 = 0, t2 = 0;
        for (size_t i = 0; i < st.size(); ++i)
          if (st[i] == tr) {
            ++v2;
            if (i + 1 < st.size()) {
              if (st[i + 1] == fr) ++t2;
              else return false;
            } else {
              return false;
            }
          }
        if (v2 > 0 && v2 != t2) return false;
      }
      return true;
    };
    for (uint64_t id = 0; id < seq.size(); ++id) {
      if (!present[id]) continue;
      for (int o = 0; o < 2; ++o) {
        const Handle handle = mk(id, o == 1);
        if (visited.count(handle)) continue;
        if (deg(fwd, handle) == 1) {
          std::vector<Handle> chain{handle};
          visited.insert(handle);
          Handle cur = handle;
          while (true) {
            auto it = fwd.find(cur);
            if (it == fwd.end()) break;
            if (it->second.size() != 1) break;
            const Handle next = it->second[0];
            if (deg(bwd, next) != 1 || visited.count(next)) break;
            if (!perfect(cur, next)) break;
            chain.push_back(next);
            visited.insert(next);
            cur = next;
            if (deg(fwd, next) != 1) break;
          }
          if (chain.size() >= 2) {
            bool already = false;
            for (Handle h : chain)
              if (merged_nodes.count(nid(h))) {
                already = true;
                break;
              }
            if (!already) {
              for (Handle h : chain) merged_nodes.insert(nid(h));
              components.push_back(chain);
            }
          }
        }
      }
    }
    return components;
  }
  // :277-478
  bool merge_component_v2(const std::vector<Handle>& handles) {
    if (handles.size() < 2) return false;
    const size_t n = handles.size();
    std::unordered_map<Handle, size_t> mapping;  // handle -> chain position (the other tuple fields are unused)
    std::vector<uint8_t> new_seq;
    for (Handle h : handles) {
      if (nid(h) < seq.size() && present[nid(h)]) {
        if (isrev(h)) {
          auto r = revcomp(seq[nid(h)]);
          new_seq.insert(new_seq.end(), r.begin(), r.end());
        } else {
          new_seq.insert(new_seq.end(), seq[nid(h)].begin(), seq[nid(h)].end());
        }
      }
    }
    const uint64_t new_id = seq.size();  // next_node_id(), :692-694
    const Handle nf = mk(new_id, false), nr = mk(new_id, true);
    for (size_t i = 0; i < n; ++i) {
      mapping[handles[i]] = i;
      mapping[flip(handles[i])] = n - 1 - i;
    }
    std::vector<Handle> rev_chain(n);
    for (size_t i = 0; i < n; ++i) rev_chain[i] = flip(handles[n - 1 - i]);
    for (auto& p : paths) {
      const std::vector<Handle>& st = p.second;
      size_t i = 0;
      while (i < st.size()) {
        auto it = mapping.find(st[i]);
        if (it != mapping.end()) {
          if (it->second == 0) {
            if (i + n <= st.size()) {
              bool complete = true;
              for (size_t j = 0; j < n; ++j)
                if (st[i + j] != handles[j]) {
                  complete = false;
                  break;
                }
              if (complete) {
                i += n;
                continue;
              }
            }
          }
          if (st[i] == rev_chain[0]) {
            if (i + n <= st.size()) {
              bool complete = true;
              for (size_t j = 0; j < n; ++j)
                if (st[i + j] != rev_chain[j]) {
                  complete = false;
                  break;
                }
              if (complete) {
                i += n;
                continue;
              }
            }
          }
          return false;
        }
        ++i;
      }
    }
    add_node(new_id, new_seq);
    for (auto& p : paths) {
      std::vector<Handle>& st = p.second;
      std::vector<Handle> ns;
      size_t i = 0;
      while (i < st.size()) {
        if (i + n <= st.size()) {
          bool f = true;
          for (size_t j = 0; j < n; ++j)
            if (st[i + j] != handles[j]) {
              f = false;
              break;
            }
          if (f) {
            ns.push_back(nf);
            i += n;
            continue;
          }
          bool r = true;
          for (size_t j = 0; j < n; ++j)
            if (st[i + j] != rev_chain[j]) {
              r = false;
              break;
            }
          if (r) {
            ns.push_back(nr);
            i += n;
            continue;
          }
        }
        ns.push_back(st[i]);
        ++i;
      }
      st.swap(ns);
    }
    const Handle first = handles[0], last = handles[n - 1];
    std::unordered_set<uint64_t> merged_ids;
    for (Handle h : handles) merged_ids.insert(nid(h));
    std::set<Edge> ne;
    for (const Edge& e : edges) {
      const bool fi = merged_ids.count(nid(e.first)) != 0, ti = merged_ids.count(nid(e.second)) != 0;
      if (fi && ti) continue;
      if (!fi && !ti) {
        ne.insert(e);
      } else if (fi && !ti) {
        if (e.first == last) ne.insert(Edge(nf, e.second));
        if (e.first == flip(first)) ne.insert(Edge(nr, e.second));
      } else {
        if (e.second == first) ne.insert(Edge(e.first, nf));
        if (e.second == flip(last)) ne.insert(Edge(e.first, nr));
      }
    }
    edges.swap(ne);
    for (Handle h : handles)
      if (nid(h) < seq.size()) present[nid(h)] = 0;
    return true;
  }
  void compact() {  // :91-112
    bool compacted = true;
    while (compacted) {
      compacted = false;
      for (auto& comp : find_simple_components())
        if (comp.size() >= 2 && merge_component_v2(comp)) compacted = true;
    }
  }
  void renumber() {  // :75-89 + :22-73
    std::vector<uint64_t> map(seq.size(), 0);
    uint64_t next = 1;
    for (uint64_t id = 0; id < seq.size(); ++id)
      if (present[id]) map[id] = next++;
    std::vector<std::vector<uint8_t>> ns(next);
    std::vector<char> np(next, 0);
    for (uint64_t id = 0; id < seq.size(); ++id)
      if (present[id]) {
        ns[map[id]] = seq[id];
        np[map[id]] = 1;
      }
    if (next == 1) {  // max_new_id = 0 -> vec![None; 1]
      ns.assign(1, {});
      np.assign(1, 0);
    }
    seq.swap(ns);
    present.swap(np);
    auto mh = [&](Handle h) -> Handle { return mk(nid(h) < map.size() && map[nid(h)] ? map[nid(h)] : nid(h), isrev(h)); };
    std::set<Edge> ne;
    for (const Edge& e : edges) ne.insert(Edge(mh(e.first), mh(e.second)));
    edges.swap(ne);
    for (auto& p : paths)
      for (auto& h : p.second) h = mh(h);
  }
  std::string gfa() const {  // :880-925, L lines sorted
    std::string o = "H\tVN:Z:1.0\n";
    for (uint64_t id = 0; id < seq.size(); ++id)
      if (present[id]) {
        o += "S\t" + std::to_string(id) + "\t";
        o.append((const char*)seq[id].data(), seq[id].size());
        o += "\n";
      }
    for (const Edge& e : edges)
      o += "L\t" + std::to_string(nid(e.first)) + "\t" + (isrev(e.first) ? "-" : "+") + "\t" + std::to_string(nid(e.second)) +
           "\t" + (isrev(e.second) ? "-" : "+") + "\t0M\n";
    for (auto& p : paths) {
      o += "P\t" + p.first + "\t";
      for (size_t i = 0; i < p.second.size(); ++i) {
        if (i) o += ",";
        o += std::to_string(nid(p.second[i]));
        o += isrev(p.second[i]) ? "-" : "+";
      }
      o += "\t*\n";
    }
    return o;
  }
};
}  // namespace ograph

// build (+ compact + renumber unless no_compact) + validate_paths_match_sequences + write_gfa.
// ids: n_seqs NUL-terminated names. rep_out (optional, total_len entries): the union-find representative
// (uf.find of the forward Pos) of every position at build time -- what decides the node base (:176-182).
// Returns a malloc'd NUL-terminated GFA text (free with orc_free), nullptr when path validation fails.
extern "C" char* orc_build_gfa(void* u, const uint8_t* const* seq_data, const uint64_t* seq_len, uint32_t n_seqs,
                               const char* const* ids, int no_compact, uint64_t* rep_out, uint64_t* n_nodes_uncompacted) {
  BUF* uf = (BUF*)u;
  ograph::Graph g;
  std::vector<uint64_t> off(n_seqs);
  uint64_t acc = 0;
  for (uint32_t i = 0; i < n_seqs; ++i) {
    off[i] = acc;
    acc += seq_len[i];
  }
  auto same = [&](uint64_t a, uint64_t b) { return a == b || uf->uf.find(a) == uf->uf.find(b); };
  std::unordered_map<uint64_t, uint64_t> union_to_node;
  uint64_t next_node_id = 1;
  for (uint32_t si = 0; si < n_seqs; ++si) {
    std::vector<ograph::Handle> steps;
    for (uint64_t i = 0; i < seq_len[si]; ++i) {
      const uint64_t gp = off[si] + i;
      const uint64_t pos_fwd = gp << 1, pos_rev = (gp << 1) | 1;
      const uint64_t union_fwd = uf->uf.find(pos_fwd), union_rev = uf->uf.find(pos_rev);
      if (rep_out) rep_out[gp] = union_fwd;
      uint64_t union_rep;
      if (union_to_node.count(union_fwd)) {
        union_rep = union_fwd;
      } else if (union_to_node.count(union_rev)) {
        union_rep = union_rev;
      } else {
        const uint64_t fwd_root = uf->uf.find(pos_fwd), rev_root = uf->uf.find(pos_rev);
        if (union_to_node.count(fwd_root)) {
          union_rep = fwd_root;
        } else if (union_to_node.count(rev_root)) {
          union_rep = rev_root;
        } else {
          bool found = false;
          uint64_t existing = 0;
          for (auto& kv : union_to_node) {  // :96-119
            if (same(union_fwd, kv.first) || same(union_rev, kv.first)) {
              found = true;
              existing = kv.first;
              break;
            }
          }
          union_rep = found ? existing : union_fwd;
        }
      }
      uint64_t node_id;
      uint8_t node_base;
      auto it = union_to_node.find(union_rep);
      if (it != union_to_node.end()) {
        node_id = it->second;
        node_base = (node_id < g.seq.size() && g.present[node_id]) ? g.seq[node_id][0] : seq_data[si][i];
      } else {
        node_id = next_node_id++;
        union_to_node[union_rep] = node_id;
        if (same(pos_fwd, union_rep) && !union_to_node.count(union_fwd)) union_to_node[union_fwd] = node_id;
        if (same(pos_rev, union_rep) && !union_to_node.count(union_rev)) union_to_node[union_rev] = node_id;
        uint8_t base = seq_data[si][i];
        const uint64_t ro = union_rep >> 1;
        for (uint32_t k = 0; k < n_seqs; ++k)  // find_sequence_for_position, :327-333
          if (ro >= off[k] && ro < off[k] + seq_len[k]) {
            base = seq_data[k][ro - off[k]];
            break;
          }
        g.add_node(node_id, std::vector<uint8_t>{base});
        node_base = base;
      }
      auto up = [](uint8_t c) -> uint8_t { return (c >= 'a' && c <= 'z') ? (uint8_t)(c - 32) : c; };
      const uint8_t e = up(seq_data[si][i]), nb = up(node_base);
      const bool need_reverse = (nb == 'A' && e == 'T') || (nb == 'T' && e == 'A') || (nb == 'C' && e == 'G') || (nb == 'G' && e == 'C');
      steps.push_back(ograph::mk(node_id, need_reverse));
    }
    g.paths.emplace_back(std::string(ids[si]), std::move(steps));
  }
  for (auto& p : g.paths)
    for (size_t i = 0; i + 1 < p.second.size(); ++i) g.add_edge(p.second[i], p.second[i + 1]);
  g.verify_path_edges();
  g.verify_path_edges();
  if (n_nodes_uncompacted) *n_nodes_uncompacted = next_node_id - 1;
  if (!no_compact) {
    g.compact();
    g.renumber();
  }
  // validate_paths_match_sequences (src/seqrush.rs:1741-1798)
  for (uint32_t si = 0; si < n_seqs; ++si) {
    std::vector<uint8_t> ps;
    for (ograph::Handle h : g.paths[si].second) {
      const auto& ns = g.seq[ograph::nid(h)];
      if (ograph::isrev(h)) {
        auto r = ograph::revcomp(ns);
        ps.insert(ps.end(), r.begin(), r.end());
      } else {
        ps.insert(ps.end(), ns.begin(), ns.end());
      }
    }
    if (ps.size() != seq_len[si] || memcmp(ps.data(), seq_data[si], ps.size()) != 0) return nullptr;
  }
  const std::string t = g.gfa();
  char* out = (char*)malloc(t.size() + 1);
  memcpy(out, t.c_str(), t.size() + 1);
  return out;
}
extern "C" void orc_free(void* p) { free(p); }

// ---------------------------------------------------------------------------
// Pair-list sparsification (allwave/src/iterator.rs:49-77,256-284; allwave/src/mash.rs:30-165;
// allwave/src/knn_graph.rs:10-172). Everything hashes through Rust's
// std::collections::hash_map::DefaultHasher = SipHash-1-3 with zero keys (std source is not in the
// reference tree; restated from the published SipHash definition, Aumasson & Bernstein 2012, with
// c = 1 compression round and d = 3 finalisation rounds). Framing of the hashed bytes as in Rust's
// `Hash` impls: `[u8]` writes the length as a little-endian u64 and then the bytes; `str`/`String`
// writes the bytes and then one 0xff byte. PARITY UNPINNED for the framing: no reference test holds a
// hash value or a pair list (allwave/src/knn_graph.rs:198-384 asserts counts and ranges only); the
// SipHash core is pinned by the published SipHash-2-4 vector and the SipHash-1-3 vectors of Rust's
// own test-suite (tests/test_sparsify.py).
// ---------------------------------------------------------------------------
static inline uint64_t rotl64(uint64_t x, int b) { return (x << b) | (x >> (64 - b)); }
static uint64_t siphash(const uint8_t* in, size_t len, uint64_t k0, uint64_t k1, int c_rounds, int d_rounds) {
  uint64_t v0 = 0x736f6d6570736575ull ^ k0, v1 = 0x646f72616e646f6dull ^ k1, v2 = 0x6c7967656e657261ull ^ k0,
           v3 = 0x7465646279746573ull ^ k1;
  auto round = [&]() {
    v0 += v1; v1 = rotl64(v1, 13); v1 ^= v0; v0 = rotl64(v0, 32);
    v2 += v3; v3 = rotl64(v3, 16); v3 ^= v2;
    v0 += v3; v3 = rotl64(v3, 21); v3 ^= v0;
    v2 += v1; v1 = rotl64(v1, 17); v1 ^= v2; v2 = rotl64(v2, 32);
  };
  const size_t nw = len / 8;
  for (size_t w = 0; w < nw; ++w) {
    uint64_t m = 0;
    for (int i = 0; i < 8; ++i) m |= (uint64_t)in[8 * w + i] << (8 * i);
    v3 ^= m;
    for (int r = 0; r < c_rounds; ++r) round();
    v0 ^= m;
  }
  uint64_t b = (uint64_t)(len & 0xff) << 56;
  for (size_t i = 0; i < (len & 7); ++i) b |= (uint64_t)in[8 * nw + i] << (8 * i);
  v3 ^= b;
  for (int r = 0; r < c_rounds; ++r) round();
  v0 ^= b;
  v2 ^= 0xff;
  for (int r = 0; r < d_rounds; ++r) round();
  return v0 ^ v1 ^ v2 ^ v3;
}
extern "C" uint64_t orc_siphash(const uint8_t* in, uint64_t len, uint64_t k0, uint64_t k1, int c_rounds, int d_rounds) {
  return siphash(in, (size_t)len, k0, k1, c_rounds, d_rounds);
}
// DefaultHasher over a `&[u8]` (Hash for [u8]: length prefix, then the bytes) and over a `String`
static uint64_t rust_hash_bytes(const uint8_t* b, size_t n) {
  std::vector<uint8_t> buf(8 + n);
  for (int i = 0; i < 8; ++i) buf[i] = (uint8_t)((uint64_t)n >> (8 * i));
  memcpy(buf.data() + 8, b, n);
  return siphash(buf.data(), buf.size(), 0, 0, 1, 3);
}
static uint64_t rust_hash_str(const std::string& s) {
  std::vector<uint8_t> buf(s.begin(), s.end());
  buf.push_back(0xff);
  return siphash(buf.data(), buf.size(), 0, 0, 1, 3);
}
extern "C" uint64_t orc_rust_hash_bytes(const uint8_t* b, uint64_t n) { return rust_hash_bytes(b, (size_t)n); }
extern "C" uint64_t orc_rust_hash_str(const char* s) { return rust_hash_str(std::string(s)); }

// sketch_sequence (allwave/src/mash.rs:78-107): every k-mer without a non-ACGT byte, canonical = the smaller
// of the two HASHES, sorted, truncated to sketch_size (duplicates are kept). Returns the count written.
static bool is_dna_base(uint8_t b) {
  if (b >= 'a' && b <= 'z') b -= 32;
  return b == 'A' || b == 'C' || b == 'G' || b == 'T';
}
static std::vector<uint64_t> sketch_sequence(const uint8_t* s, size_t n, size_t k, size_t sketch_size) {
  std::vector<uint64_t> hashes;
  if (n < k) return hashes;
  std::vector<uint8_t> rc(k);
  for (size_t i = 0; i + k <= n; ++i) {
    bool ok = true;
    for (size_t j = 0; j < k; ++j)
      if (!is_dna_base(s[i + j])) {
        ok = false;
        break;
      }
    if (!ok) continue;
    for (size_t j = 0; j < k; ++j) {  // reverse_complement_kmer, :124-136
      uint8_t b = s[i + k - 1 - j], u = (b >= 'a' && b <= 'z') ? (uint8_t)(b - 32) : b;
      rc[j] = u == 'A' ? 'T' : u == 'T' ? 'A' : u == 'C' ? 'G' : u == 'G' ? 'C' : b;
    }
    const uint64_t hf = rust_hash_bytes(s + i, k), hr = rust_hash_bytes(rc.data(), k);
    hashes.push_back(hf < hr ? hf : hr);
  }
  std::sort(hashes.begin(), hashes.end());
  if (hashes.size() > sketch_size) hashes.resize(sketch_size);
  return hashes;
}
extern "C" uint64_t orc_sketch(const uint8_t* s, uint64_t n, uint64_t k, uint64_t sketch_size, uint64_t* out) {
  const std::vector<uint64_t> h = sketch_sequence(s, (size_t)n, (size_t)k, (size_t)sketch_size);
  memcpy(out, h.data(), h.size() * 8);
  return h.size();
}
// KmerSketch::jaccard + mash_distance (:40-73)
static double mash_distance(const std::vector<uint64_t>& a, const std::vector<uint64_t>& b, size_t k, uint64_t* inter_out,
                            uint64_t* uni_out) {
  std::set<uint64_t> s1(a.begin(), a.end()), s2(b.begin(), b.end());
  uint64_t inter = 0;
  for (uint64_t x : s1) inter += s2.count(x);
  const uint64_t uni = s1.size() + s2.size() - inter;
  if (inter_out) *inter_out = inter;
  if (uni_out) *uni_out = uni;
  const double jac = uni == 0 ? 0.0 : (double)inter / (double)uni;
  if (jac <= 0.0) return 1.0;
  const double ratio = (2.0 * jac) / (1.0 + jac);
  if (ratio <= 0.0) return 1.0;
  return (-1.0 / (double)k) * log(ratio);
}
static bool should_include_pair(const std::string& a, const std::string& b, double fraction) {  // knn_graph.rs:159-172
  const uint64_t h = rust_hash_str(a + ":" + b);
  const double normalized = (double)h / 18446744073709551615.0;  // u64::MAX as f64 == 2^64
  return normalized < fraction;
}
// extract_tree_pairs / extract_tree_pairs_separated (knn_graph.rs:10-97). pairs_out gets up to cap (i, j) pairs;
// returns the number of pairs (tree pairs first when `separated`, *n_tree = their count).
extern "C" uint64_t orc_tree_pairs(const uint8_t* const* seq_data, const uint64_t* seq_len, const char* const* ids, uint32_t n,
                                   uint64_t k_nearest, uint64_t k_farthest, double random_fraction, uint64_t kmer_size,
                                   int separated, uint32_t* pairs_out, uint64_t cap, uint64_t* n_tree, double* dist_out) {
  typedef std::pair<uint64_t, uint64_t> P;
  std::vector<P> tree, rnd;
  if (n_tree) *n_tree = 0;
  if (n < 2) return 0;
  std::vector<std::vector<uint64_t>> sk(n);
  for (uint32_t i = 0; i < n; ++i) sk[i] = sketch_sequence(seq_data[i], (size_t)seq_len[i], (size_t)kmer_size, 1000);
  std::vector<std::vector<double>> d(n, std::vector<double>(n, 0.0));
  for (uint32_t i = 0; i < n; ++i)
    for (uint32_t j = i + 1; j < n; ++j) d[i][j] = d[j][i] = mash_distance(sk[i], sk[j], (size_t)kmer_size, nullptr, nullptr);
  if (dist_out)
    for (uint32_t i = 0; i < n; ++i)
      for (uint32_t j = 0; j < n; ++j) dist_out[(size_t)i * n + j] = d[i][j];
  auto knn = [&](uint64_t kk, bool farthest) {  // build_knn_graph, :110-142 (sort_by is stable)
    for (uint32_t i = 0; i < n; ++i) {
      std::vector<std::pair<double, uint64_t>> nb;
      for (uint32_t j = 0; j < n; ++j)
        if (j != i) nb.emplace_back(d[i][j], j);
      if (farthest) std::stable_sort(nb.begin(), nb.end(), [](const auto& a, const auto& b) { return a.first > b.first; });
      else std::stable_sort(nb.begin(), nb.end(), [](const auto& a, const auto& b) { return a.first < b.first; });
      for (uint64_t t = 0; t < kk && t < nb.size(); ++t) tree.emplace_back(i, nb[t].second);
    }
  };
  if (k_nearest > 0) knn(k_nearest, false);
  if (k_farthest > 0) knn(k_farthest, true);
  if (random_fraction > 0.0)
    for (uint32_t i = 0; i < n; ++i)
      for (uint32_t j = 0; j < n; ++j)
        if (i != j && should_include_pair(ids[i], ids[j], random_fraction)) rnd.emplace_back(i, j);
  std::vector<P> out;
  if (!separated) {
    tree.insert(tree.end(), rnd.begin(), rnd.end());
    std::sort(tree.begin(), tree.end());
    tree.erase(std::unique(tree.begin(), tree.end()), tree.end());
    out = tree;
    if (n_tree) *n_tree = out.size();
  } else {
    std::sort(tree.begin(), tree.end());
    tree.erase(std::unique(tree.begin(), tree.end()), tree.end());
    out = tree;
    if (n_tree) *n_tree = tree.size();
    for (const P& p : rnd)
      if (!std::binary_search(tree.begin(), tree.end(), p)) out.push_back(p);
  }
  for (size_t i = 0; i < out.size() && i < cap; ++i) {
    pairs_out[2 * i] = (uint32_t)out[i].first;
    pairs_out[2 * i + 1] = (uint32_t)out[i].second;
  }
  return out.size();
}
// apply_random_sparsification (iterator.rs:256-284) over the n^2 (or n^2 - n) i-major pair list
extern "C" uint64_t orc_random_pairs(const char* const* ids, uint32_t n, int exclude_self, double keep_fraction,
                                     uint32_t* pairs_out, uint64_t cap) {
  uint64_t m = 0;
  for (uint32_t i = 0; i < n; ++i)
    for (uint32_t j = 0; j < n; ++j) {
      if (exclude_self && i == j) continue;
      if (should_include_pair(ids[i], ids[j], keep_fraction)) {
        if (m < cap) {
          pairs_out[2 * m] = i;
          pairs_out[2 * m + 1] = j;
        }
        ++m;
      }
    }
  return m;
}
// compute_connectivity_probability (iterator.rs:300-334)
extern "C" double orc_connectivity_probability(uint64_t n, double connectivity_prob) {
  if (n <= 1) return 1.0;
  double x = connectivity_prob < 0.001 ? 0.001 : connectivity_prob > 0.999 ? 0.999 : connectivity_prob;
  if (n <= 10) return n == 2 ? 1.0 : n == 3 ? 0.8 : n == 4 ? 0.7 : n == 5 ? 0.6 : 0.5;
  const double nf = (double)n, c = -log(-log(x)), p = (log(nf) + c) / nf;
  return p < 0.001 ? 0.001 : p > 1.0 ? 1.0 : p;
}

// ---------------------------------------------------------------------------
// The pair-dispatch loop (src/seqrush.rs:611-757 with allwave/src/iterator.rs:185-203): one OS
// thread per core instead of rayon, a thread-local aligner like allwave's thread_local
// (allwave/src/alignment.rs:10-13). Used as the timed CPU baseline.
// results (optional): per pair {is_reverse, status, score, cigar_len, fwd_ed, rev_ed}
// ---------------------------------------------------------------------------
struct RunStats {
  double seconds, edlib_seconds, wfa_seconds, unite_seconds;
  uint64_t aligned_bp, unite_calls, pairs, cigar_columns;
};
extern "C" int orc_run_pairs(void* h, void* u, const uint8_t* const* seq_data, const uint64_t* seq_len, uint32_t n_seqs,
                             const uint32_t* pairs, uint64_t n_pairs, const int* pen5, uint64_t min_match_len,
                             int n_threads, int32_t* results6, RunStats* stats) {
  Oracle* o = (Oracle*)h;
  BUF* uf = (BUF*)u;
  std::vector<uint64_t> off(n_seqs);
  uint64_t acc = 0, maxlen = 0;
  for (uint32_t i = 0; i < n_seqs; ++i) {  // src/seqrush.rs:1816-1818
    off[i] = acc;
    acc += seq_len[i];
    if (seq_len[i] > maxlen) maxlen = seq_len[i];
  }
  if (n_threads < 1) n_threads = 1;
  std::atomic<uint64_t> next{0};
  std::atomic<int> bad{0};
  std::vector<double> te(n_threads, 0.0), tw(n_threads, 0.0), tu(n_threads, 0.0);
  std::vector<uint64_t> bp(n_threads, 0), cols(n_threads, 0);
  const uint64_t calls0 = uf ? uf->unite_calls.load() : 0;
  auto t0 = std::chrono::steady_clock::now();
  auto worker = [&](int tid) {
    void* wf = o->wfa_new(pen5[0], pen5[1], pen5[2], pen5[3], pen5[4], 3);
    std::vector<char> cigar(2 * maxlen + 16);
    std::string cs;
    while (true) {
      const uint64_t i = next.fetch_add(1);
      if (i >= n_pairs) break;
      const uint32_t qi = pairs[2 * i], ti = pairs[2 * i + 1];
      PairRes r;
      align_pair_impl(o, wf, seq_data[qi], (int)seq_len[qi], seq_data[ti], (int)seq_len[ti], cigar.data(), &r, &te[tid], &tw[tid]);
      bp[tid] += seq_len[qi];
      cols[tid] += (uint64_t)r.cigar_len;
      if (results6) {
        int32_t* rr = results6 + 6 * i;
        rr[0] = r.is_reverse; rr[1] = r.status; rr[2] = r.score; rr[3] = r.cigar_len; rr[4] = r.fwd_ed; rr[5] = r.rev_ed;
      }
      if (uf) {
        auto a0 = std::chrono::steady_clock::now();
        cigar_bytes_to_string(cigar.data(), r.cigar_len, cs);
        SeqV s1{seq_data[qi], seq_len[qi], off[qi]}, s2{seq_data[ti], seq_len[ti], off[ti]};
        if (process_alignment(uf, cs.c_str(), s1, s2, min_match_len, r.is_reverse != 0, 0, 0) != 0) bad.store(1);
        tu[tid] += std::chrono::duration<double>(std::chrono::steady_clock::now() - a0).count();
      }
    }
    o->wfa_free(wf);
  };
  std::vector<std::thread> th;
  for (int t = 1; t < n_threads; ++t) th.emplace_back(worker, t);
  worker(0);
  for (auto& t : th) t.join();
  if (stats) {
    stats->seconds = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    stats->edlib_seconds = stats->wfa_seconds = stats->unite_seconds = 0;
    stats->aligned_bp = stats->cigar_columns = 0;
    for (int t = 0; t < n_threads; ++t) {
      stats->edlib_seconds += te[t];
      stats->wfa_seconds += tw[t];
      stats->unite_seconds += tu[t];
      stats->aligned_bp += bp[t];
      stats->cigar_columns += cols[t];
    }
    stats->unite_calls = uf ? uf->unite_calls.load() - calls0 : 0;
    stats->pairs = n_pairs;
  }
  return bad.load() ? -1 : 0;
}

// ---------------------------------------------------------------------------
// Independent check of the optimal gap-affine-2p penalty: plain O(n*m) dynamic programme
// (Smith-Waterman-Gotoh with two affine pieces), for small cases only. Restates the scoring
// model of WFA2-lib/alignment/affine2p_penalties.h, not its wavefront algorithm.
// ---------------------------------------------------------------------------
extern "C" int orc_swg2p_penalty(const uint8_t* p, int n, const uint8_t* t, int m, const int* pen5) {
  const int x = pen5[0], o1 = pen5[1], e1 = pen5[2], o2 = pen5[3], e2 = pen5[4];
  const int INF = 1 << 29;
  std::vector<int> M(m + 1), I1(m + 1), I2(m + 1), D1(m + 1), D2(m + 1), pM(m + 1), pD1(m + 1), pD2(m + 1);
  for (int i = 0; i <= n; ++i) {
    for (int j = 0; j <= m; ++j) {
      int i1 = INF, i2 = INF, d1 = INF, d2 = INF, mm = INF;
      if (j > 0) {  // gap consuming text
        i1 = std::min(M[j - 1] + o1 + e1, I1[j - 1] + e1);
        i2 = std::min(M[j - 1] + o2 + e2, I2[j - 1] + e2);
      }
      if (i > 0) {  // gap consuming pattern
        d1 = std::min(pM[j] + o1 + e1, pD1[j] + e1);
        d2 = std::min(pM[j] + o2 + e2, pD2[j] + e2);
      }
      if (i == 0 && j == 0) mm = 0;
      if (i > 0 && j > 0) mm = pM[j - 1] + (p[i - 1] == t[j - 1] ? 0 : x);
      mm = std::min(std::min(mm, std::min(i1, i2)), std::min(d1, d2));
      M[j] = mm; I1[j] = i1; I2[j] = i2; D1[j] = d1; D2[j] = d2;
    }
    pM = M; pD1 = D1; pD2 = D2;
  }
  return M[m];
}
