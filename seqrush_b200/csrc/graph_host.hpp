// graph_host.hpp -- host half of the graph build: the BidirectedGraph the device passes produced
// (node bases, oriented steps, de-duplicated edges), the reference's compaction and renumbering
// (BidirectedGraph::compact / find_simple_components / merge_component_v2 / renumber_nodes_sequentially,
// src/bidirected_ops.rs:75-478), path validation (src/seqrush.rs:1741-1798) and the GFA writer
// (src/bidirected_ops.rs:880-925).
//
// Compaction keeps the reference's DECISIONS -- handles visited in node-id order (forward then reverse),
// chains grown forwards only, a chain pushed only when none of its nodes was merged earlier in the round,
// chains merged in discovery order with the new node id = current size of the node table, rounds repeated
// until one merges nothing -- because node ids, node orientations and the stored form of every edge follow
// from them. What it does not keep is the cost: the reference rescans every path for every candidate pair
// (are_perfect_neighbors) and every path and every edge for every chain (merge_component_v2). Here a round
// builds, in one pass over the steps, the successor summary of every handle (none / always the same / not
// unique or at a path end) that answers are_perfect_neighbors in O(1), and an occurrence index per node that
// validates a chain in time proportional to its occurrences; edges are kept with per-node incidence lists so a
// merge touches only the edges of its chain; all path rewrites of a round are applied in one pass at its end
// (the chains of a round are node-disjoint, so this equals the reference's one-chain-at-a-time rewriting).
#pragma once
#include <stdint.h>

#include <algorithm>
#include <string>
#include <unordered_map>
#include <unordered_set>
#include <vector>

namespace srb {

typedef uint32_t GHandle;  // (node_id << 1) | is_reverse, src/bidirected_graph.rs:6-48
static inline GHandle gh_flip(GHandle h) { return h ^ 1u; }
static inline uint32_t gh_id(GHandle h) { return h >> 1; }
static inline uint64_t gedge(GHandle f, GHandle t) { return ((uint64_t)f << 32) | t; }

static inline void g_revcomp_append(std::string& out, const std::string& s) {  // src/bidirected_graph.rs:58-71
  for (size_t i = s.size(); i-- > 0;) {
    char b = s[i];
    switch (b) {
      case 'A': case 'a': b = 'T'; break;
      case 'T': case 't': b = 'A'; break;
      case 'C': case 'c': b = 'G'; break;
      case 'G': case 'g': b = 'C'; break;
      case 'N': case 'n': b = 'N'; break;
      default: break;
    }
    out.push_back(b);
  }
}

struct HostGraph {
  std::vector<std::string> seq;  // index = node id (0 unused)
  std::vector<uint8_t> present;
  std::unordered_set<uint64_t> edges;       // stored form (from << 32 | to)
  std::vector<std::vector<GHandle>> paths;  // one per input sequence, in input order
  uint64_t uncompacted_nodes = 0;
  uint32_t rounds = 0, merged_chains = 0;

  uint64_t node_count() const {
    uint64_t c = 0;
    for (uint8_t p : present) c += p;
    return c;
  }

  // ---- one round of find_simple_components + merge_component_v2 over all its chains ----
  bool compact_round() {
    const size_t nh = seq.size() * 2;
    // adjacency summary per handle (src/bidirected_ops.rs:119-139): every stored edge (f, t) also acts as
    // (t^1, f^1); a handle's unique successor is only ever used when its out-degree is exactly 1
    std::vector<uint32_t> outc(nh, 0), inc(nh, 0);
    std::vector<GHandle> out1(nh, 0);
    for (uint64_t e : edges) {
      const GHandle f = (GHandle)(e >> 32), t = (GHandle)e;
      outc[f]++;
      out1[f] = t;
      inc[t]++;
      outc[gh_flip(t)]++;
      out1[gh_flip(t)] = gh_flip(f);
      inc[gh_flip(f)]++;
    }
    // successor summary per handle over all paths: 0 = never visited, 1 = always followed by succ[h],
    // 2 = followed by different handles or visited at the end of a path
    std::vector<uint8_t> sstate(nh, 0);
    std::vector<GHandle> succ(nh, 0);
    std::vector<uint32_t> occ_cnt(seq.size(), 0);
    for (const auto& st : paths)
      for (size_t i = 0; i < st.size(); ++i) {
        const GHandle h = st[i];
        occ_cnt[gh_id(h)]++;
        if (i + 1 < st.size()) {
          const GHandle n = st[i + 1];
          if (sstate[h] == 0) {
            sstate[h] = 1;
            succ[h] = n;
          } else if (sstate[h] == 1 && succ[h] != n) {
            sstate[h] = 2;
          }
        } else {
          sstate[h] = 2;
        }
      }
    auto follows = [&](GHandle h, GHandle x) { return sstate[h] == 0 || (sstate[h] == 1 && succ[h] == x); };
    auto perfect = [&](GHandle from, GHandle to) {  // are_perfect_neighbors, :141-190
      return follows(from, to) && follows(gh_flip(to), gh_flip(from));
    };
    // chains (:192-275)
    std::vector<std::vector<GHandle>> comps;
    std::vector<uint8_t> visited(nh, 0), merged(seq.size(), 0);
    for (uint32_t id = 0; id < seq.size(); ++id) {
      if (!present[id]) continue;
      for (uint32_t o = 0; o < 2; ++o) {
        const GHandle h = (id << 1) | o;
        if (visited[h]) continue;
        if (outc[h] != 1) continue;
        std::vector<GHandle> chain{h};
        visited[h] = 1;
        GHandle cur = h;
        while (true) {
          if (outc[cur] != 1) break;
          const GHandle next = out1[cur];
          if (inc[next] != 1 || visited[next]) break;
          if (!perfect(cur, next)) break;
          chain.push_back(next);
          visited[next] = 1;
          cur = next;
          if (outc[next] != 1) break;
        }
        if (chain.size() >= 2) {
          bool already = false;
          for (GHandle c : chain)
            if (merged[gh_id(c)]) {
              already = true;
              break;
            }
          if (!already) {
            for (GHandle c : chain) merged[gh_id(c)] = 1;
            comps.push_back(std::move(chain));
          }
        }
      }
    }
    if (comps.empty()) return false;
    // occurrence index of the nodes that sit in some chain: (path, step index), CSR over node ids
    std::vector<uint64_t> occ_off(seq.size() + 1, 0);
    for (uint32_t id = 0; id < seq.size(); ++id) occ_off[id + 1] = occ_off[id] + (merged[id] ? occ_cnt[id] : 0);
    std::vector<std::pair<uint32_t, uint32_t>> occ(occ_off.back());
    {
      std::vector<uint64_t> fill(occ_off.begin(), occ_off.end() - 1);
      for (uint32_t p = 0; p < paths.size(); ++p)
        for (uint32_t i = 0; i < paths[p].size(); ++i) {
          const uint32_t id = gh_id(paths[p][i]);
          if (merged[id]) occ[fill[id]++] = std::make_pair(p, i);
        }
    }
    // per-node incidence lists of the edges (only what chains need is touched later)
    std::unordered_map<uint32_t, std::vector<uint64_t>> incident;
    auto inc_add = [&](uint64_t e) {
      incident[gh_id((GHandle)(e >> 32))].push_back(e);
      if (gh_id((GHandle)(e >> 32)) != gh_id((GHandle)e)) incident[gh_id((GHandle)e)].push_back(e);
    };
    for (uint64_t e : edges) {
      const uint32_t a = gh_id((GHandle)(e >> 32)), b = gh_id((GHandle)e);
      if (merged[a] || merged[b]) inc_add(e);
    }
    // start handle -> (merged chain, direction) for the path rewrite at the end of the round
    std::unordered_map<GHandle, std::pair<uint32_t, uint32_t>> start_of;  // handle -> (index into done, 0 fwd / 1 rev)
    struct Done {
      uint32_t n;
      GHandle nf;
    };
    std::vector<Done> done;
    bool any = false;
    // all path rewrites of the node-distinct chains of a round, in one pass
    auto flush_rewrites = [&]() {
      if (done.empty()) return;
      for (auto& st : paths) {
        size_t w = 0, i = 0;
        while (i < st.size()) {
          auto it = start_of.find(st[i]);
          if (it != start_of.end()) {
            const Done& d = done[it->second.first];
            st[w++] = d.nf | it->second.second;
            i += d.n;
          } else {
            st[w++] = st[i++];
          }
        }
        st.resize(w);
      }
      start_of.clear();
      done.clear();
    };
    bool literal = false;  // set once a chain had to be rewritten in place: the occurrence index is stale then
    for (const auto& chain : comps) {
      const size_t n = chain.size();
      std::vector<GHandle> rev_chain(n);
      for (size_t i = 0; i < n; ++i) rev_chain[i] = gh_flip(chain[n - 1 - i]);
      // --- validation (:306-345): every step that is a chain member (either orientation) must lie in a
      // complete forward or reverse occurrence of the chain ---
      bool distinct = true;
      {
        std::unordered_set<uint32_t> ids;
        for (GHandle c : chain)
          if (!ids.insert(gh_id(c)).second) distinct = false;
      }
      if (!distinct && !literal) {  // a node twice in one chain (hairpin): the reference's literal scan from here on
        flush_rewrites();
        literal = true;
      }
      bool ok = true;
      if (!literal) {
        uint64_t total = 0, blocks = 0;
        for (GHandle c : chain) total += occ_off[gh_id(c) + 1] - occ_off[gh_id(c)];
        auto count_blocks = [&](const std::vector<GHandle>& pat) {
          const uint32_t id0 = gh_id(pat[0]);
          for (uint64_t k = occ_off[id0]; k < occ_off[id0 + 1]; ++k) {
            const auto& st = paths[occ[k].first];
            const size_t i = occ[k].second;
            if (st[i] != pat[0] || i + n > st.size()) continue;
            bool m = true;
            for (size_t j = 1; j < n; ++j)
              if (st[i + j] != pat[j]) {
                m = false;
                break;
              }
            if (m) ++blocks;
          }
        };
        count_blocks(chain);
        count_blocks(rev_chain);
        ok = blocks * n == total;
      } else {
        std::unordered_map<GHandle, size_t> pos;
        for (size_t i = 0; i < n; ++i) {
          pos[chain[i]] = i;
          pos[gh_flip(chain[i])] = n - 1 - i;
        }
        for (const auto& st : paths) {
          size_t i = 0;
          while (ok && i < st.size()) {
            auto it = pos.find(st[i]);
            if (it == pos.end()) {
              ++i;
              continue;
            }
            bool skipped = false;
            if (it->second == 0 && i + n <= st.size() && std::equal(chain.begin(), chain.end(), st.begin() + i)) {
              i += n;
              skipped = true;
            }
            if (!skipped && st[i] == rev_chain[0] && i + n <= st.size() && std::equal(rev_chain.begin(), rev_chain.end(), st.begin() + i)) {
              i += n;
              skipped = true;
            }
            if (!skipped) ok = false;
          }
          if (!ok) break;
        }
      }
      if (!ok) continue;
      any = true;
      ++merged_chains;
      // --- new node (:287-303, :347) ---
      std::string ns;
      for (GHandle c : chain) {
        if (gh_id(c) < seq.size() && present[gh_id(c)]) {
          if (c & 1u) g_revcomp_append(ns, seq[gh_id(c)]);
          else ns += seq[gh_id(c)];
        }
      }
      const uint32_t new_id = (uint32_t)seq.size();
      seq.push_back(std::move(ns));
      present.push_back(1);
      const GHandle nf = new_id << 1, nr = nf | 1u;
      if (!literal) {
        start_of[chain[0]] = std::make_pair((uint32_t)done.size(), 0u);
        start_of[rev_chain[0]] = std::make_pair((uint32_t)done.size(), 1u);
        done.push_back(Done{(uint32_t)n, nf});
      } else {
        // rewrite right away, literally (:349-402)
        for (auto& st : paths) {
          std::vector<GHandle> out;
          size_t i = 0;
          while (i < st.size()) {
            if (i + n <= st.size() && std::equal(chain.begin(), chain.end(), st.begin() + i)) {
              out.push_back(nf);
              i += n;
            } else if (i + n <= st.size() && std::equal(rev_chain.begin(), rev_chain.end(), st.begin() + i)) {
              out.push_back(nr);
              i += n;
            } else {
              out.push_back(st[i++]);
            }
          }
          st.swap(out);
        }
      }
      // --- edges (:404-466) ---
      const GHandle first = chain[0], last = chain[n - 1];
      std::unordered_set<uint32_t> ids;
      for (GHandle c : chain) ids.insert(gh_id(c));
      std::vector<uint64_t> touched;
      for (uint32_t id : ids) {
        auto it = incident.find(id);
        if (it == incident.end()) continue;
        for (uint64_t e : it->second) touched.push_back(e);
      }
      std::vector<uint64_t> add;
      for (uint64_t e : touched) {
        if (!edges.erase(e)) continue;  // listed under both endpoints, or already rewritten away
        const GHandle f = (GHandle)(e >> 32), t = (GHandle)e;
        const bool fi = ids.count(gh_id(f)) != 0, ti = ids.count(gh_id(t)) != 0;
        if (fi && ti) continue;
        if (fi) {
          if (f == last) add.push_back(gedge(nf, t));
          if (f == gh_flip(first)) add.push_back(gedge(nr, t));
        } else {
          if (t == first) add.push_back(gedge(f, nf));
          if (t == gh_flip(last)) add.push_back(gedge(f, nr));
        }
      }
      for (uint64_t e : add)
        if (edges.insert(e).second) inc_add(e);
      for (GHandle c : chain) present[gh_id(c)] = 0;
    }
    flush_rewrites();
    return any;
  }

  void compact() {  // :91-112
    while (true) {
      ++rounds;
      if (!compact_round()) break;
    }
  }

  void renumber() {  // :75-89 with apply_node_id_mapping :22-73
    std::vector<uint32_t> map(seq.size(), 0);
    uint32_t next = 1;
    for (uint32_t id = 0; id < seq.size(); ++id)
      if (present[id]) map[id] = next++;
    std::vector<std::string> ns(next);
    std::vector<uint8_t> np(next, 0);
    for (uint32_t id = 0; id < seq.size(); ++id)
      if (present[id]) {
        ns[map[id]].swap(seq[id]);
        np[map[id]] = 1;
      }
    seq.swap(ns);
    present.swap(np);
    auto mh = [&](GHandle h) -> GHandle { return (map[gh_id(h)] << 1) | (h & 1u); };
    std::unordered_set<uint64_t> ne;
    ne.reserve(edges.size() * 2);
    for (uint64_t e : edges) ne.insert(gedge(mh((GHandle)(e >> 32)), mh((GHandle)e)));
    edges.swap(ne);
    for (auto& st : paths)
      for (auto& h : st) h = mh(h);
  }

  // validate_paths_match_sequences (src/seqrush.rs:1741-1798): -1 when all paths spell their sequences,
  // else the index of the first path that does not
  long validate(const std::vector<const uint8_t*>& data, const std::vector<uint64_t>& len) const {
    for (size_t p = 0; p < paths.size(); ++p) {
      uint64_t at = 0;
      std::string piece;
      for (GHandle h : paths[p]) {
        const std::string& s = seq[gh_id(h)];
        if (at + s.size() > len[p]) return (long)p;
        if (h & 1u) {
          piece.clear();
          g_revcomp_append(piece, s);
          if (memcmp(piece.data(), data[p] + at, s.size()) != 0) return (long)p;
        } else if (memcmp(s.data(), data[p] + at, s.size()) != 0) {
          return (long)p;
        }
        at += s.size();
      }
      if (at != len[p]) return (long)p;
    }
    return -1;
  }

  // write_gfa (src/bidirected_ops.rs:880-925). The reference emits the L lines in HashSet iteration order
  // (unspecified, different in every process); they are written sorted by (from, to) handle here.
  void gfa_text(const std::vector<std::string>& names, std::string& o) const {
    o.clear();
    o += "H\tVN:Z:1.0\n";
    char buf[96];
    for (uint32_t id = 0; id < seq.size(); ++id)
      if (present[id]) {
        o += "S\t";
        o += std::to_string(id);
        o += '\t';
        o += seq[id];
        o += '\n';
      }
    std::vector<uint64_t> es(edges.begin(), edges.end());
    std::sort(es.begin(), es.end());
    for (uint64_t e : es) {
      const GHandle f = (GHandle)(e >> 32), t = (GHandle)e;
      const int n = snprintf(buf, sizeof buf, "L\t%u\t%c\t%u\t%c\t0M\n", gh_id(f), (f & 1u) ? '-' : '+', gh_id(t), (t & 1u) ? '-' : '+');
      o.append(buf, (size_t)n);
    }
    for (size_t p = 0; p < paths.size(); ++p) {
      o += "P\t";
      o += names[p];
      o += '\t';
      bool firsts = true;
      for (GHandle h : paths[p]) {
        if (!firsts) o += ',';
        firsts = false;
        const int n = snprintf(buf, sizeof buf, "%u%c", gh_id(h), (h & 1u) ? '-' : '+');
        o.append(buf, (size_t)n);
      }
      o += "\t*\n";
    }
  }
};

}  // namespace srb
