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
