// sparsify.cuh -- device side of pair-list sparsification (row f2 / a5):
//   * MinHash sketches of every sequence (allwave/src/mash.rs:78-107): all k-mers without a non-ACGT
//     byte, hashed in both orientations with Rust's DefaultHasher (SipHash-1-3, zero keys, `[u8]` framing
//     = little-endian u64 length, then the bytes), canonical = the smaller of the two hashes, the
//     sketch_size smallest kept in ascending order (duplicates kept);
//   * Jaccard counts of every sequence pair over the distinct values of the two sketches (:40-55);
//   * the hashed random-pair filter on "{id_i}:{id_j}" (`String` framing = bytes, then 0xff)
//     (allwave/src/iterator.rs:256-284, allwave/src/knn_graph.rs:145-172).
// The f64 part (Mash distance, k-nearest / k-farthest selection) is host code over the integer counts
// these kernels return (seqrush_b200/sparsify.py): n^2 values, not a hot loop.
#pragma once
#include "kernels.cuh"

namespace srb {

struct Sip13 {  // streaming SipHash-1-3 with zero keys
  unsigned long long v0, v1, v2, v3, m;
  unsigned int nbytes;
  __host__ __device__ __forceinline__ void init() {
    v0 = 0x736f6d6570736575ull;
    v1 = 0x646f72616e646f6dull;
    v2 = 0x6c7967656e657261ull;
    v3 = 0x7465646279746573ull;
    m = 0;
    nbytes = 0;
  }
  static __host__ __device__ __forceinline__ unsigned long long rotl(unsigned long long x, int b) { return (x << b) | (x >> (64 - b)); }
  __host__ __device__ __forceinline__ void round() {
    v0 += v1; v1 = rotl(v1, 13); v1 ^= v0; v0 = rotl(v0, 32);
    v2 += v3; v3 = rotl(v3, 16); v3 ^= v2;
    v0 += v3; v3 = rotl(v3, 21); v3 ^= v0;
    v2 += v1; v1 = rotl(v1, 17); v1 ^= v2; v2 = rotl(v2, 32);
  }
  __host__ __device__ __forceinline__ void word(unsigned long long w) {  // only when nbytes is a multiple of 8
    v3 ^= w;
    round();
    v0 ^= w;
    nbytes += 8;
  }
  __host__ __device__ __forceinline__ void byte(unsigned char b) {
    m |= (unsigned long long)b << (8 * (nbytes & 7u));
    ++nbytes;
    if ((nbytes & 7u) == 0) {
      v3 ^= m;
      round();
      v0 ^= m;
      m = 0;
    }
  }
  __host__ __device__ __forceinline__ unsigned long long finish() {
    const unsigned long long b = ((unsigned long long)(nbytes & 0xffu) << 56) | m;
    v3 ^= b;
    round();
    v0 ^= b;
    v2 ^= 0xff;
    round();
    round();
    round();
    return v0 ^ v1 ^ v2 ^ v3;
  }
};

__device__ __forceinline__ bool is_dna_base(uint8_t b) {
  b = (b >= 'a' && b <= 'z') ? (uint8_t)(b - 32) : b;
  return b == 'A' || b == 'C' || b == 'G' || b == 'T';
}
__device__ __forceinline__ uint8_t rc_kmer_base(uint8_t b) {  // reverse_complement_kmer, mash.rs:124-136
  const uint8_t u = (b >= 'a' && b <= 'z') ? (uint8_t)(b - 32) : b;
  return u == 'A' ? 'T' : u == 'T' ? 'A' : u == 'C' ? 'G' : u == 'G' ? 'C' : b;
}

// canonical hashes of every valid k-mer of every sequence; hashes of sequence s go to
// out[hash_off[s] ...) in arbitrary order, their number to counts[s]
__global__ void kmer_hash_kernel(SeqTable seqs, uint32_t k, const unsigned long long* __restrict__ hash_off,
                                 unsigned long long* out, unsigned int* counts) {
  for (uint32_t s = blockIdx.y; s < seqs.n; s += gridDim.y) {
    const uint32_t L = seqs.len[s];
    if (L < k) continue;
    const uint8_t* p = seqs.fwd + seqs.byte_off[s];
    unsigned long long* dst = out + hash_off[s];
    const uint32_t nk = L - k + 1;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < nk; i += gridDim.x * blockDim.x) {
      bool ok = true;
      for (uint32_t j = 0; j < k; ++j) ok = ok && is_dna_base(p[i + j]);
      if (!ok) continue;
      Sip13 hf, hr;
      hf.init();
      hr.init();
      hf.word((unsigned long long)k);  // Hash for [u8]: write_length_prefix
      hr.word((unsigned long long)k);
      for (uint32_t j = 0; j < k; ++j) {
        hf.byte(p[i + j]);
        hr.byte(rc_kmer_base(p[i + k - 1 - j]));
      }
      const unsigned long long a = hf.finish(), b = hr.finish();
      const unsigned int at = atomicAdd(&counts[s], 1u);
      dst[at] = a < b ? a : b;
    }
  }
}

// the `want` smallest of each sequence's hashes, ascending, duplicates kept. One CTA per sequence:
// an 8-pass most-significant-byte-first radix select finds the value of rank `m` and how many copies of
// it belong to the sketch, the selected values are gathered into shared memory and bitonic-sorted.
static constexpr int SKETCH_MAX = 4096;
template <int THREADS>
__global__ void __launch_bounds__(THREADS) sketch_select_kernel(const unsigned long long* __restrict__ hashes,
                                                                const unsigned long long* __restrict__ hash_off,
                                                                const unsigned int* __restrict__ counts, uint32_t n_seqs,
                                                                uint32_t want, unsigned long long* sketches,
                                                                unsigned int* sketch_counts) {
  __shared__ unsigned int hist[256];
  __shared__ unsigned long long sel[SKETCH_MAX];
  __shared__ unsigned long long s_prefix;
  __shared__ unsigned int s_rank, s_fill, s_eq;
  for (uint32_t s = blockIdx.x; s < n_seqs; s += gridDim.x) {
    const unsigned long long* h = hashes + hash_off[s];
    const unsigned int cnt = counts[s];
    const unsigned int m = cnt < want ? cnt : want;
    if (threadIdx.x == 0) sketch_counts[s] = m;
    if (m == 0) continue;
    unsigned long long thr = ~0ull;
    unsigned int need_eq = 0;
    if (cnt > want) {
      __syncthreads();
      if (threadIdx.x == 0) {
        s_prefix = 0;
        s_rank = m - 1;  // zero-based rank of the largest kept value
      }
      for (int pass = 0; pass < 8; ++pass) {
        const int shift = 56 - 8 * pass;
        for (int b = threadIdx.x; b < 256; b += THREADS) hist[b] = 0;
        __syncthreads();
        const unsigned long long prefix = s_prefix;
        const unsigned long long mask = pass == 0 ? 0ull : (~0ull << (shift + 8));
        for (unsigned int i = threadIdx.x; i < cnt; i += THREADS) {
          const unsigned long long v = h[i];
          if ((v & mask) == prefix) atomicAdd(&hist[(unsigned int)(v >> shift) & 255u], 1u);
        }
        __syncthreads();
        if (threadIdx.x == 0) {
          unsigned int r = s_rank, b = 0;
          while (hist[b] <= r) {
            r -= hist[b];
            ++b;
          }
          s_rank = r;
          s_prefix = prefix | ((unsigned long long)b << shift);
        }
        __syncthreads();
      }
      thr = s_prefix;          // value of rank m-1
      need_eq = s_rank + 1;    // copies of thr inside the sketch
    }
    __syncthreads();
    if (threadIdx.x == 0) {
      s_fill = 0;
      s_eq = 0;
    }
    __syncthreads();
    for (unsigned int i = threadIdx.x; i < cnt; i += THREADS) {
      const unsigned long long v = h[i];
      if (cnt <= want || v < thr) {
        sel[atomicAdd(&s_fill, 1u)] = v;
      } else if (v == thr) {
        if (atomicAdd(&s_eq, 1u) < need_eq) sel[atomicAdd(&s_fill, 1u)] = v;
      }
    }
    __syncthreads();
    // bitonic sort of the m selected values (padded with the maximum to a power of two)
    unsigned int p2 = 1;
    while (p2 < m) p2 <<= 1;
    for (unsigned int i = m + threadIdx.x; i < p2; i += THREADS) sel[i] = ~0ull;
    __syncthreads();
    for (unsigned int kk = 2; kk <= p2; kk <<= 1)
      for (unsigned int j = kk >> 1; j > 0; j >>= 1) {
        for (unsigned int i = threadIdx.x; i < p2; i += THREADS) {
          const unsigned int x = i ^ j;
          if (x > i) {
            const unsigned long long a = sel[i], b = sel[x];
            const bool up = (i & kk) == 0;
            if ((a > b) == up) {
              sel[i] = b;
              sel[x] = a;
            }
          }
        }
        __syncthreads();
      }
    for (unsigned int i = threadIdx.x; i < m; i += THREADS) sketches[(unsigned long long)s * want + i] = sel[i];
    __syncthreads();
  }
}

// |A n B| and |A u B| over the distinct values of two ascending sketches; one thread per pair i < j
__global__ void jaccard_kernel(const unsigned long long* __restrict__ sketches, const unsigned int* __restrict__ sketch_counts,
                               uint32_t n, uint32_t want, unsigned int* inter, unsigned int* uni) {
  const unsigned long long total = (unsigned long long)n * n;
  for (unsigned long long t = blockIdx.x * (unsigned long long)blockDim.x + threadIdx.x; t < total;
       t += (unsigned long long)gridDim.x * blockDim.x) {
    const uint32_t i = (uint32_t)(t / n), j = (uint32_t)(t % n);
    if (j <= i) continue;
    const unsigned long long* a = sketches + (unsigned long long)i * want;
    const unsigned long long* b = sketches + (unsigned long long)j * want;
    const unsigned int na = sketch_counts[i], nb = sketch_counts[j];
    unsigned int x = 0, y = 0, both = 0, all = 0;
    while (x < na || y < nb) {
      unsigned long long v;
      bool ina = false, inb = false;
      if (y >= nb || (x < na && a[x] <= b[y])) {
        v = a[x];
        ina = true;
        if (y < nb && b[y] == v) inb = true;
      } else {
        v = b[y];
        inb = true;
      }
      while (x < na && a[x] == v) ++x;
      while (y < nb && b[y] == v) ++y;
      ++all;
      if (ina && inb) ++both;
    }
    inter[t] = both;
    uni[t] = all;
    inter[(unsigned long long)j * n + i] = both;
    uni[(unsigned long long)j * n + i] = all;
  }
}

// should_include_pair: DefaultHasher over the String "{id_i}:{id_j}" -> hash / 2^64 < fraction
__global__ void pair_hash_kernel(const uint8_t* __restrict__ ids, const unsigned long long* __restrict__ id_off, uint32_t n,
                                 double fraction, int exclude_self, uint8_t* keep, unsigned long long* hashes) {
  const unsigned long long total = (unsigned long long)n * n;
  for (unsigned long long t = blockIdx.x * (unsigned long long)blockDim.x + threadIdx.x; t < total;
       t += (unsigned long long)gridDim.x * blockDim.x) {
    const uint32_t i = (uint32_t)(t / n), j = (uint32_t)(t % n);
    Sip13 h;
    h.init();
    for (unsigned long long p = id_off[i]; p < id_off[i + 1]; ++p) h.byte(ids[p]);
    h.byte((unsigned char)':');
    for (unsigned long long p = id_off[j]; p < id_off[j + 1]; ++p) h.byte(ids[p]);
    h.byte(0xff);  // Hash for str: write_str appends 0xff
    const unsigned long long v = h.finish();
    if (hashes) hashes[t] = v;
    const double normalized = __ull2double_rn(v) * 0x1p-64;  // (hash as f64) / (u64::MAX as f64): the divisor is 2^64, exact
    keep[t] = (!(exclude_self && i == j) && normalized < fraction) ? 1 : 0;
  }
}

}  // namespace srb
