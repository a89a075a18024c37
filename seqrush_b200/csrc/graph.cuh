// graph.cuh -- device half of the graph build from the union-find partition
// (reference: SeqRush::build_bidirected_graph_with_options, src/bidirected_builder.rs:17-289).
//
// The reference walks every base of every sequence in order, does two finds per base (:48-49), numbers a
// component when it is first seen (:134-137), gives the node the forward base at the offset of the
// component's union-find root (:176-182) and orients the path step by a complement test (:190-199); for
// every component it has not seen it first scans ALL keys of its map with same() (:96-119) -- a scan that
// never matches on the production path (distinct roots are distinct components once the unions are
// quiescent), which makes the build quadratic in the number of nodes. Here the same outputs come from
// bulk passes over HBM:
//   1. canonical root of every position (min Pos of its component: the offset of the FIRST-SEEN member,
//      because positions are walked in offset order),
//   2. an exclusive prefix sum over "is first-seen" = the reference's node numbering,
//   3. node base + oriented step of every position,
//   4. the edge set with the reference's complement-aware de-duplication (add_edge, src/bidirected_ops.rs:
//      813-825: an edge is stored in the form of its FIRST occurrence in path order unless that form or its
//      complement is already there) through a hash table keyed by the canonical form with an atomicMin on the
//      global step index.
#pragma once
#include "kernels.cuh"

namespace srb {

struct GraphArgs {
  SeqTable seqs;
  uint32_t* parent;
  unsigned long long total_len;
  const unsigned long long* rep_in;  // optional: representative Pos per position (decides the node base), else nullptr
  uint32_t* comp;                    // [total_len] offset of the first-seen member of the component
  uint32_t* first_flag;              // [total_len] 1 when the position is the first-seen member
  uint32_t* rank;                    // [total_len] exclusive prefix sum of first_flag
  uint32_t* steps;                   // [total_len] (node_id << 1) | need_reverse
  uint8_t* node_base;                // [n_nodes]
  unsigned long long* err;           // [0] positions whose forward and reverse Pos are not united, [1] bad representatives
};

__device__ __forceinline__ uint32_t seq_of_offset(const SeqTable& t, unsigned long long off) {
  uint32_t lo = 0, hi = t.n;  // last sequence with pos_off <= off
  while (hi - lo > 1) {
    const uint32_t mid = (lo + hi) >> 1;
    if (t.pos_off[mid] <= off) lo = mid;
    else hi = mid;
  }
  return lo;
}
__device__ __forceinline__ uint8_t base_at_offset(const SeqTable& t, unsigned long long off) {
  const uint32_t s = seq_of_offset(t, off);
  return t.fwd[t.byte_off[s] + (off - t.pos_off[s])];
}

__global__ void graph_roots_kernel(GraphArgs a) {
  unsigned long long bad = 0;
  for (unsigned long long i = blockIdx.x * (unsigned long long)blockDim.x + threadIdx.x; i < a.total_len;
       i += (unsigned long long)gridDim.x * blockDim.x) {
    const uint32_t r = uf_find(a.parent, (uint32_t)(2 * i));
    const uint32_t r2 = uf_find(a.parent, (uint32_t)(2 * i + 1));
    if (r != r2) ++bad;  // production invariant of SeqRush::new (src/seqrush.rs:324-328)
    a.comp[i] = r >> 1;
    a.first_flag[i] = ((unsigned long long)(r >> 1) == i) ? 1u : 0u;
  }
  if (bad) atomicAdd(&a.err[0], bad);
}

// ---- exclusive prefix sum of uint32 flags, three passes (blocks of SCAN_BLOCK elements) ----
static constexpr int SCAN_THREADS = 256, SCAN_PER = 8, SCAN_BLOCK = SCAN_THREADS * SCAN_PER;
__global__ void scan_block_sums_kernel(const uint32_t* in, unsigned long long n, uint32_t* block_sums) {
  __shared__ uint32_t s[SCAN_THREADS / 32];
  const unsigned long long base = (unsigned long long)blockIdx.x * SCAN_BLOCK;
  uint32_t v = 0;
  for (int j = 0; j < SCAN_PER; ++j) {
    const unsigned long long i = base + (unsigned long long)j * SCAN_THREADS + threadIdx.x;
    if (i < n) v += in[i];
  }
  v = __reduce_add_sync(0xffffffffu, v);
  if ((threadIdx.x & 31) == 0) s[threadIdx.x >> 5] = v;
  __syncthreads();
  if (threadIdx.x == 0) {
    uint32_t t = 0;
    for (int w = 0; w < SCAN_THREADS / 32; ++w) t += s[w];
    block_sums[blockIdx.x] = t;
  }
}
__global__ void scan_sums_kernel(uint32_t* block_sums, unsigned long long nb, unsigned long long* total) {
  // one CTA: sequential over tiles of blockDim elements with a carried prefix
  __shared__ uint32_t s[1024];
  __shared__ uint32_t carry;
  if (threadIdx.x == 0) carry = 0;
  __syncthreads();
  for (unsigned long long t0 = 0; t0 < nb; t0 += blockDim.x) {
    const unsigned long long i = t0 + threadIdx.x;
    const uint32_t v = i < nb ? block_sums[i] : 0u;
    s[threadIdx.x] = v;
    __syncthreads();
    for (int ofs = 1; ofs < (int)blockDim.x; ofs <<= 1) {
      uint32_t x = 0;
      if ((int)threadIdx.x >= ofs) x = s[threadIdx.x - ofs];
      __syncthreads();
      s[threadIdx.x] += x;
      __syncthreads();
    }
    if (i < nb) block_sums[i] = carry + s[threadIdx.x] - v;
    __syncthreads();
    if (threadIdx.x == blockDim.x - 1) carry += s[threadIdx.x];
    __syncthreads();
  }
  if (threadIdx.x == 0) *total = carry;
}
__global__ void scan_apply_kernel(const uint32_t* in, unsigned long long n, const uint32_t* block_sums, uint32_t* out) {
  __shared__ uint32_t s[SCAN_THREADS];
  const unsigned long long base = (unsigned long long)blockIdx.x * SCAN_BLOCK + (unsigned long long)threadIdx.x * SCAN_PER;
  uint32_t v[SCAN_PER], sum = 0;
  for (int j = 0; j < SCAN_PER; ++j) {
    v[j] = (base + j < n) ? in[base + j] : 0u;
    sum += v[j];
  }
  s[threadIdx.x] = sum;
  __syncthreads();
  for (int ofs = 1; ofs < SCAN_THREADS; ofs <<= 1) {
    uint32_t x = 0;
    if ((int)threadIdx.x >= ofs) x = s[threadIdx.x - ofs];
    __syncthreads();
    s[threadIdx.x] += x;
    __syncthreads();
  }
  uint32_t run = block_sums[blockIdx.x] + s[threadIdx.x] - sum;
  for (int j = 0; j < SCAN_PER; ++j) {
    if (base + j < n) out[base + j] = run;
    run += v[j];
  }
}

__device__ __forceinline__ uint8_t ascii_upper(uint8_t c) { return (c >= 'a' && c <= 'z') ? (uint8_t)(c - 32) : c; }

__global__ void graph_steps_kernel(GraphArgs a) {
  unsigned long long bad = 0;
  for (unsigned long long i = blockIdx.x * (unsigned long long)blockDim.x + threadIdx.x; i < a.total_len;
       i += (unsigned long long)gridDim.x * blockDim.x) {
    const uint32_t first = a.comp[i];
    const uint32_t node = a.rank[first] + 1;  // ids start at 1 (:24)
    unsigned long long base_off = first;
    if (a.rep_in) {
      const unsigned long long rp = a.rep_in[first];
      if ((rp >> 1) >= a.total_len || (uf_find(a.parent, (uint32_t)rp) >> 1) != first) ++bad;
      else base_off = rp >> 1;
    }
    const uint8_t nb = base_at_offset(a.seqs, base_off);  // base at the representative's offset (:176-182)
    const uint8_t mine = base_at_offset(a.seqs, i);
    const uint8_t u = ascii_upper(nb), e = ascii_upper(mine);
    const bool rev = (u == 'A' && e == 'T') || (u == 'T' && e == 'A') || (u == 'C' && e == 'G') || (u == 'G' && e == 'C');  // :190-199
    a.steps[i] = (node << 1) | (rev ? 1u : 0u);
    if ((unsigned long long)first == i) a.node_base[node - 1] = nb;
  }
  if (bad) atomicAdd(&a.err[1], bad);
}

// ---- edge set ----
struct EdgeArgs {
  SeqTable seqs;
  const uint32_t* steps;
  unsigned long long total_len;
  unsigned long long* keys;  // [cap] canonical packed edge, EMPTY when free
  uint32_t* first_at;        // [cap] smallest global step index that produced the edge
  unsigned long long cap;    // power of two
  uint2* out;                // unique edges in their stored form
  unsigned long long* n_out;
};
static constexpr unsigned long long EDGE_EMPTY = ~0ull;
__device__ __forceinline__ unsigned long long edge_pack(uint32_t f, uint32_t t) { return ((unsigned long long)f << 32) | t; }
__device__ __forceinline__ unsigned long long hash64(unsigned long long x) {
  x ^= x >> 33;
  x *= 0xff51afd7ed558ccdULL;
  x ^= x >> 33;
  x *= 0xc4ceb9fe1a85ec53ULL;
  x ^= x >> 33;
  return x;
}
__global__ void graph_edges_insert_kernel(EdgeArgs a) {
  for (unsigned long long i = blockIdx.x * (unsigned long long)blockDim.x + threadIdx.x; i + 1 < a.total_len;
       i += (unsigned long long)gridDim.x * blockDim.x) {
    const uint32_t s = seq_of_offset(a.seqs, i);
    if (i + 1 >= a.seqs.pos_off[s] + a.seqs.len[s]) continue;  // last step of its path
    const uint32_t f = a.steps[i], t = a.steps[i + 1];
    const unsigned long long e = edge_pack(f, t), c = edge_pack(t ^ 1u, f ^ 1u);
    const unsigned long long key = e < c ? e : c;
    unsigned long long slot = hash64(key) & (a.cap - 1);
    while (true) {
      const unsigned long long old = atomicCAS(&a.keys[slot], EDGE_EMPTY, key);
      if (old == EDGE_EMPTY || old == key) {
        atomicMin(&a.first_at[slot], (uint32_t)i);
        break;
      }
      slot = (slot + 1) & (a.cap - 1);
    }
  }
}
__global__ void graph_edges_collect_kernel(EdgeArgs a) {
  for (unsigned long long sl = blockIdx.x * (unsigned long long)blockDim.x + threadIdx.x; sl < a.cap;
       sl += (unsigned long long)gridDim.x * blockDim.x) {
    if (a.keys[sl] == EDGE_EMPTY) continue;
    const uint32_t g = a.first_at[sl];
    const unsigned long long o = atomicAdd(a.n_out, 1ull);
    a.out[o] = make_uint2(a.steps[g], a.steps[g + 1]);
  }
}

}  // namespace srb
