// kernels.cuh -- device kernels of the align-and-unite path (sm_100a).
//   align_kernel   : persistent CTAs, one pair at a time per CTA: strand choice (orient_core.cuh)
//                    -> BiWFA gap-affine-2p with on-device backtrace (wfa_core.cuh) -> run-length
//                    CIGAR copied into the chunk arena.
//   union_kernel   : one CTA per pair: walks the run-length CIGAR, unites exact-match runs >= k
//                    in the HBM-resident parent array (atomicCAS link-to-smaller + path halving).
//   expand_kernel  : run-length CIGAR -> one byte per column ('M','X','I','D').
//   uf_* kernels   : BidirectedUnionFind entry points (src/bidirected_union_find.rs).
#pragma once
#include <cuda_runtime.h>
#include <cstdio>
#include "orient_core.cuh"

namespace srb {

// ---------------------------------------------------------------------------
// Union-find on a uint32 parent array. Links always point from the larger root to the smaller
// one, so parent[x] <= x, cycles are impossible and the final representative of a component is
// its minimum Pos (the canonical label north_star asks for). find() halves paths with plain
// stores (an ancestor stays an ancestor forever); roots only change through atomicCAS.
// Reference semantics: uf_rush lib.rs:112-134 (find), 159-209 (unite) -- partition-equivalent.
// ---------------------------------------------------------------------------
__device__ __forceinline__ uint32_t uf_find(uint32_t* __restrict__ parent, uint32_t x) {
  uint32_t p = parent[x];
  while (p != x) {
    SR_ASSERT(p < x, "parent[x] <= x (links point to the smaller index)");
    const uint32_t gp = parent[p];
    if (gp != p) parent[x] = gp;
    x = p;
    p = gp;
  }
  return x;
}
// returns true when two roots were linked
// second half of a unite whose two parent reads (pa = parent[a], pb = parent[b]) are already there
__device__ __forceinline__ bool uf_unite_from(uint32_t* __restrict__ parent, uint32_t a, uint32_t b, uint32_t pa,
                                              uint32_t pb) {
  // most unite calls of an all-vs-all run are redundant (the two Pos already share a set, usually
  // as children of the same root after path halving): the two reads settle those
  if (a == b || pa == pb || pa == b || pb == a) return false;  // a == b: src/bidirected_union_find.rs:40
  uint32_t ra = uf_find(parent, pa), rb = uf_find(parent, pb);
  while (ra != rb) {
    if (ra < rb) {
      const uint32_t t = ra;
      ra = rb;
      rb = t;
    }
    const uint32_t old = atomicCAS(&parent[ra], ra, rb);
    if (old == ra) return true;
    ra = uf_find(parent, old);
    rb = uf_find(parent, rb);
  }
  return false;
}
__device__ __forceinline__ bool uf_unite(uint32_t* __restrict__ parent, uint32_t a, uint32_t b) {
  if (a == b) return false;
  return uf_unite_from(parent, a, b, parent[a], parent[b]);
}

// ---------------------------------------------------------------------------
struct SeqTable {
  const uint8_t* fwd;        // padded concatenation, forward bytes (union kernel / process_alignment)
  const uint64_t* byte_off;  // start of sequence i inside fwd
  const uint32_t* len;
  const uint64_t* pos_off;   // offset in the concatenated Pos space (src/seqrush.rs:1816-1818)
  const uint8_t* rc_safe;    // 1 when every byte is in ACGTacgtN: allwave RC == seqrush per-base RC
  const uint8_t* rc_inv;     // 1 when every byte is in ACGTN: allwave RC is an involution on the sequence
  // packed copies for the aligner: 1 << lbits bits per symbol, forward and allwave reverse
  // complement (allwave/src/alignment.rs:174-186); word 0 of sequence i at pk_off[i] (16-byte
  // aligned, 4 guard words on both sides)
  const uint32_t* pk_fwd;
  const uint32_t* pk_rc;
  const uint64_t* pk_off;
  int lbits;
  uint32_t n;
};

struct PairOut {  // device-side result record
  uint32_t q, t;
  int32_t is_reverse, status, score, penalty;
  uint32_t num_matches, n_mismatch, n_ins, n_del;
  uint32_t edit_distance, n_runs;
  unsigned long long run_off;
  unsigned long long cells, ocells;
};

struct AlignArgs {
  SeqTable seqs;
  const uint2* pairs;
  uint32_t n_pairs;
  Penalties pen;
  // per-CTA workspace (indexed by blockIdx.x)
  int32_t* ring;
  size_t ring_stride;  // ints per CTA = 2*(scope+1)*5*W
  int W;               // widest ring row this launch can address
  int fixed_w;         // 1: every pair uses W (retry launches); 0: rows are sized per pair from its edit distance
  int32_t* full;
  size_t full_stride;  // ints per CTA = 5*cap*cap
  WfMeta* full_meta;   // cap*5 per CTA
  int cap;
  int32_t* orient;
  size_t orient_stride;  // ints per CTA = 4*row
  int orient_row;
  uint32_t* runs_tmp;
  uint32_t* bt_tmp;
  uint32_t runs_cap;  // per CTA
  uint32_t smem_seq_words;  // words per staged sequence (0: sequences are read from global memory)
  // outputs
  PairOut* out;
  uint32_t* arena;
  unsigned long long arena_cap;
  unsigned long long* arena_head;
  unsigned int* next_pair;
  const uint32_t* todo;  // optional indirection (retry lists); nullptr = identity
  const uint32_t* mirror_of;  // orient_kernel: pairs whose strand choice is copied from the mirrored pair (else 0xffffffff); may be null
  // split mode (leaf_T > 0): align_kernel only splits the big levels of the BiWFA recursion and emits the small
  // sub-problems as tasks; subalign_kernel solves them one warp per task, assemble_kernel joins the pieces in order
  int leaf_T;
  struct Frame* tasks;          // [task_cap]
  uint32_t* task_pair;          // [task_cap] slot of the pair (index into out[]) the task belongs to
  struct TaskOut* task_out;     // [task_cap]
  struct SplitRec* split;       // [n_pairs] by pair slot
  unsigned int task_cap;
  unsigned int* task_count;     // tasks emitted so far
  unsigned int* next_task;      // subalign_kernel's work counter
  // two-pass mode of the one-warp kernel: its first launch only bisects the tasks down to the base-case threshold
  // and emits what is left as second-level tasks ("leaves"), its second launch finishes those. A launch of
  // subalign_kernel emits when etasks != nullptr; assemble_kernel walks pair -> tasks -> leaves when it is set.
  int leaf_T2;                  // threshold of the first pass (250: the reference's base-case bound)
  struct Frame* etasks;         // [etask_cap] second-level tasks
  uint32_t* etask_pair;         // [etask_cap]
  struct TaskOut* etask_out;    // [etask_cap] results of the second pass
  struct SplitRec* esplit;      // [task_cap] by first-level task: its leaves
  unsigned int etask_cap;
  unsigned int* etask_count;
  unsigned int asm_max;         // pieces assemble_kernel joins per pair (<= ASM_MAX_TASKS; tests lower it)
};

struct TaskOut {
  unsigned long long run_off;
  unsigned long long cells;
  uint32_t n_runs;
  int32_t status, score;
  int32_t pad;
};
struct SplitRec {
  unsigned long long cells;  // cells of the levels align_kernel solved itself
  uint32_t first_task, n_tasks;
  int32_t status, score;     // score: level-0 score when align_kernel found the level-0 breakpoint, else INT32_MIN
};

static constexpr int ST_ARENA_OVERFLOW = 6;
static constexpr int ST_TASK_OVERFLOW = 8;  // split mode: task / leaf array full -> the pair is re-run on the monolithic path

// ---- bulk asynchronous copy global -> shared (TMA engine, 1-D form) completing on an mbarrier ----
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(dst)),
               "l"(src), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  uint32_t done = 0;
  const long long t_start = clock64();
  while (!done) {
    // a bulk copy that never lands would spin here forever: fail loudly instead (about 5 s)
    if (clock64() - t_start > 10000000000ll) {
      printf("seqrush_b200: bulk copy of a sequence pair did not complete (CTA %d)\n", (int)blockIdx.x);
      __trap();
    }
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
        "selp.u32 %0, 1, 0, p;\n"
        "}\n"
        : "=r"(done)
        : "r"(smem_u32(bar)), "r"(parity)
        : "memory");
  }
}

#ifdef WFA_PROFILE
// cycle counters of one CTA (measurement build only)
__device__ __noinline__ void prof_print(const CtaShared& sh, long long prof_start, const char* tag) {
  printf("[prof %s]\n", tag);
    const unsigned long long* pf = sh.prof;
    const double n = pf[0] ? (double)pf[0] : 1.0;
    printf("[prof cta %d] ring steps %llu  total cyc %lld | per step: top-barrier %.0f  setup %.0f  loop %.0f  finish %.0f  "
           "(barrier wait %.0f) | avg width %.0f\n",
           (int)blockIdx.x, pf[0], clock64() - prof_start, pf[1] / n, pf[2] / n, pf[3] / n, pf[4] / n, pf[5] / n, pf[7] / n);
    printf("[prof cta %d] passes %llu, chunks per pass %.2f, passes with <= 1 chunk per warp %.1f%%\n", (int)blockIdx.x, pf[16],
           pf[16] ? (double)pf[17] / pf[16] : 0.0, pf[16] ? 100.0 * pf[18] / pf[16] : 0.0);
    {
      const unsigned long long* q = sh.prof2;
      printf("[prof cta %d] Mcycles: find_breakpoint L0 %.1f L1 %.1f L2 %.1f L3 %.1f L4+ %.1f | base case %.1f (%llu calls, backtrace %.1f) | "
             "overlap %.1f, close phase %.1f (%llu overlap calls)\n", (int)blockIdx.x, q[0] / 1e6, q[1] / 1e6, q[2] / 1e6, q[3] / 1e6, q[4] / 1e6,
             q[5] / 1e6, q[6], q[7] / 1e6, q[8] / 1e6, q[9] / 1e6, q[10]);
      printf("[prof cta %d] Mcells: L0 %.1f L1 %.1f L2 %.1f L3 %.1f L4+ %.1f\n", (int)blockIdx.x, q[11] / 1e6, q[12] / 1e6, q[13] / 1e6,
             q[14] / 1e6, q[15] / 1e6);
    }
    printf("[prof cta %d] strand choice %llu cyc (sum of edit distances %llu), BiWFA %llu cyc\n", (int)blockIdx.x, pf[13], pf[15], pf[14]);
    printf("[prof cta %d] warp1 interior chunks %llu: compute %.0f cyc, extend %.0f cyc | warp0 fringe chunks %llu: %.0f cyc\n",
           (int)blockIdx.x, pf[8], pf[9] / (pf[8] ? (double)pf[8] : 1.0), pf[10] / (pf[8] ? (double)pf[8] : 1.0), pf[11],
           pf[12] / (pf[11] ? (double)pf[11] : 1.0));
}
#endif

// ---------------------------------------------------------------------------
// Strand choice once per UNORDERED pair. For sequences over {A,C,G,T,N} allwave's reverse complement R is an
// involution and a bijection on the alphabet, so for the two orders of a pair
//   ED(a, b) = ED(b, a)   and   ED(R(b), a) = ED(R(R(b)), R(a)) = ED(b, R(a)) = ED(R(a), b):
// both edit distances of (b, a) equal those of (a, b), hence the same strand and the same minimum
// (allwave/src/alignment.rs:148-171 compares exactly these two numbers). A pair (q, t) with q > t whose mirror
// (t, q) is in the same launch takes the mirror's result; everything else is computed as before.
// ---------------------------------------------------------------------------
struct MirrorArgs {
  const uint2* pairs;
  uint32_t n;
  const uint8_t* rc_inv;
  unsigned long long* keys;  // open-addressing table, ~0 = empty
  uint32_t* vals;
  uint32_t mask;             // slots - 1 (power of two)
  uint32_t* mirror_of;       // [n]
  PairOut* out;
};
__device__ __forceinline__ uint32_t mirror_hash(unsigned long long key, uint32_t mask) {
  return (uint32_t)((key * 0x9E3779B97F4A7C15ull) >> 32) & mask;
}
__global__ void mirror_insert_kernel(MirrorArgs m) {
  for (uint32_t p = blockIdx.x * blockDim.x + threadIdx.x; p < m.n; p += gridDim.x * blockDim.x) {
    const uint2 pr = m.pairs[p];
    m.mirror_of[p] = 0xffffffffu;
    if (pr.x >= pr.y || !m.rc_inv[pr.x] || !m.rc_inv[pr.y]) continue;
    const unsigned long long key = ((unsigned long long)pr.x << 32) | pr.y;
    uint32_t h = mirror_hash(key, m.mask);
    while (true) {
      const unsigned long long prev = atomicCAS(&m.keys[h], ~0ull, key);
      if (prev == ~0ull) {
        m.vals[h] = p;
        break;
      }
      if (prev == key) break;  // the same pair twice in the list: the first one serves
      h = (h + 1) & m.mask;
    }
  }
}
__global__ void mirror_lookup_kernel(MirrorArgs m) {
  for (uint32_t p = blockIdx.x * blockDim.x + threadIdx.x; p < m.n; p += gridDim.x * blockDim.x) {
    const uint2 pr = m.pairs[p];
    if (pr.x <= pr.y || !m.rc_inv[pr.x] || !m.rc_inv[pr.y]) continue;
    const unsigned long long key = ((unsigned long long)pr.y << 32) | pr.x;
    uint32_t h = mirror_hash(key, m.mask);
    while (true) {
      const unsigned long long k = m.keys[h];
      if (k == key) {
        m.mirror_of[p] = m.vals[h];
        break;
      }
      if (k == ~0ull) break;
      h = (h + 1) & m.mask;
    }
  }
}
__global__ void mirror_copy_kernel(MirrorArgs m) {
  for (uint32_t p = blockIdx.x * blockDim.x + threadIdx.x; p < m.n; p += gridDim.x * blockDim.x) {
    const uint32_t src = m.mirror_of[p];
    if (src == 0xffffffffu) continue;
    m.out[p].is_reverse = m.out[src].is_reverse;
    m.out[p].edit_distance = m.out[src].edit_distance;
    m.out[p].ocells = 0;
  }
}

// ---------------------------------------------------------------------------
// Strand choice for every pair of a launch (allwave/src/alignment.rs:24-30,148-171), one pair per
// CTA, persistent CTAs. It runs as its own launch because it needs three staged sequences (query
// forward, query reverse complement, target) while the aligner needs two: keeping it out of
// align_kernel cuts that kernel's shared memory per CTA by a third (more resident CTAs for long
// sequences) and its instruction footprint. Writes is_reverse / edit_distance / ocells of out[pi].
// ---------------------------------------------------------------------------
template <int THREADS>
__global__ void __launch_bounds__(THREADS) orient_kernel(AlignArgs a) {
  extern __shared__ __align__(16) uint32_t s_seq[];  // [Pf | Pr | T], smem_seq_words each
  __shared__ CtaShared sh;
  __shared__ unsigned int s_pair;
  __shared__ __align__(8) uint64_t s_bar;
  Ctx cx;
  cx.sh = &sh;
  cx.tid = threadIdx.x;
  cx.nt = THREADS;
  cx.parity = 0;
  cx.stage = nullptr;
  int32_t* obuf = a.orient + (size_t)blockIdx.x * a.orient_stride;
  const int lbits = a.seqs.lbits;
  const bool staged = a.smem_seq_words != 0;
  uint32_t phase = 0;
  if (staged && threadIdx.x == 0) mbar_init(&s_bar, 1);
  while (true) {
    __syncthreads();
    if (threadIdx.x == 0) s_pair = atomicAdd(a.next_pair, 1u);
    __syncthreads();
    const unsigned int slot = s_pair;
    if (slot >= a.n_pairs) break;
    const unsigned int pi = a.todo ? a.todo[slot] : slot;
    if (a.mirror_of && a.mirror_of[pi] != 0xffffffffu) continue;  // copied from the pair (t, q) by mirror_copy_kernel
    const uint2 pr = a.pairs[pi];
    const uint32_t plen = a.seqs.len[pr.x], tlen = a.seqs.len[pr.y];
    const uint32_t* Pf = a.seqs.pk_fwd + a.seqs.pk_off[pr.x];
    const uint32_t* Pr = a.seqs.pk_rc + a.seqs.pk_off[pr.x];
    const uint32_t* T = a.seqs.pk_fwd + a.seqs.pk_off[pr.y];
    if (staged) {
      const uint32_t wp = 8 + (((((plen << lbits) + 31) >> 5) + 3) & ~3u);
      const uint32_t wt = 8 + (((((tlen << lbits) + 31) >> 5) + 3) & ~3u);
      if (threadIdx.x == 0) {
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        mbar_expect_tx(&s_bar, (2 * wp + wt) * 4);
        bulk_g2s(s_seq, Pf - 4, wp * 4, &s_bar);
        bulk_g2s(s_seq + a.smem_seq_words, Pr - 4, wp * 4, &s_bar);
        bulk_g2s(s_seq + 2 * a.smem_seq_words, T - 4, wt * 4, &s_bar);
      }
      mbar_wait(&s_bar, phase);
      phase ^= 1;
      Pf = s_seq + 4;
      Pr = s_seq + a.smem_seq_words + 4;
      T = s_seq + 2 * a.smem_seq_words + 4;
    }
    const OrientResult orr = wf_orient(cx, Pf, Pr, T, lbits, (int)plen, (int)tlen, obuf, a.orient_row);
    __syncthreads();
    if (threadIdx.x == 0) {
      a.out[pi].is_reverse = orr.is_reverse;
      a.out[pi].edit_distance = (uint32_t)orr.distance;
      a.out[pi].ocells = sh.ocells;
    }
  }
}

__global__ void max_edit_distance_kernel(const PairOut* __restrict__ out, uint32_t n, unsigned int* result) {
  unsigned int m = 0;
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) m = max(m, out[i].edit_distance);
  m = __reduce_max_sync(0xffffffffu, m);
  if ((threadIdx.x & 31) == 0 && m) atomicMax(result, m);
}

template <int THREADS, int MIN_CTAS>
__global__ void __launch_bounds__(THREADS, MIN_CTAS) align_kernel(AlignArgs a) {
  // dynamic shared memory: [query strand chosen by orient_kernel | target] (smem_seq_words each; absent when the
  // sequences are read from global memory through L1) followed by the per-warp row staging buffers of the score step
  extern __shared__ __align__(16) uint32_t s_seq[];
  __shared__ CtaShared sh;
  __shared__ CtaTables<MAX_SCOPE, STACK_MAX> s_tables;
  __shared__ unsigned int s_pair;
  __shared__ unsigned int s_cnt[8];
  __shared__ __align__(8) uint64_t s_bar;
  if (threadIdx.x == 0) cta_bind(&sh, &s_tables);
  Ctx cx;
  cx.sh = &sh;
  cx.tid = threadIdx.x;
  cx.nt = THREADS;
  cx.parity = 0;
  cx.stage = reinterpret_cast<int32_t*>(s_seq + 2 * (size_t)a.smem_seq_words);
  const size_t b = blockIdx.x;
  PairWorkspace ws;
  ws.ring_f = a.ring + b * a.ring_stride;
  ws.full = a.full + b * a.full_stride;
  ws.full_meta = a.full_meta + b * (size_t)a.cap * 5;
  ws.cap = a.cap;
  RunOut ro;
  ro.runs = a.runs_tmp + b * (size_t)a.runs_cap;
  ro.cap = a.runs_cap;
  ro.bt = a.bt_tmp + b * (size_t)a.runs_cap;
  ro.bt_cap = a.runs_cap;
  const int lbits = a.seqs.lbits;
  const bool staged = a.smem_seq_words != 0;
  uint32_t phase = 0;
  if (staged && threadIdx.x == 0) mbar_init(&s_bar, 1);
#ifdef WFA_PROFILE
  if (threadIdx.x == 0)
    for (int i = 0; i < 20; ++i) sh.prof[i] = sh.prof2[i] = 0;
  const long long prof_start = clock64();
#endif
  while (true) {
    __syncthreads();
    if (threadIdx.x == 0) s_pair = atomicAdd(a.next_pair, 1u);
    __syncthreads();
    const unsigned int slot = s_pair;
    if (slot >= a.n_pairs) break;
    const unsigned int pi = a.todo ? a.todo[slot] : slot;
    const uint2 pr = a.pairs[pi];
    const uint32_t plen = a.seqs.len[pr.x], tlen = a.seqs.len[pr.y];
    // 1. strand: chosen by orient_kernel (allwave/src/alignment.rs:24-30)
    OrientResult orr;
    orr.is_reverse = a.out[pi].is_reverse;
    orr.distance = (int)a.out[pi].edit_distance;
    const unsigned long long pair_ocells = a.out[pi].ocells;
    const uint32_t* P = (orr.is_reverse ? a.seqs.pk_rc : a.seqs.pk_fwd) + a.seqs.pk_off[pr.x];
    const uint32_t* T = a.seqs.pk_fwd + a.seqs.pk_off[pr.y];
    if (staged) {
      // stage the two packed sequences (with their guard words) into shared memory
      const uint32_t wp = 8 + (((((plen << lbits) + 31) >> 5) + 3) & ~3u);
      const uint32_t wt = 8 + (((((tlen << lbits) + 31) >> 5) + 3) & ~3u);
      if (threadIdx.x == 0) {
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        mbar_expect_tx(&s_bar, (wp + wt) * 4);
        bulk_g2s(s_seq, P - 4, wp * 4, &s_bar);
        bulk_g2s(s_seq + a.smem_seq_words, T - 4, wt * 4, &s_bar);
      }
      mbar_wait(&s_bar, phase);
      phase ^= 1;
      P = s_seq + 4;
      T = s_seq + a.smem_seq_words + 4;
    }
#ifdef WFA_PROFILE
    const long long t_or0 = clock64(), t_or1 = t_or0;
#endif
    // 2. BiWFA (allwave/src/alignment.rs:189-276). Ring rows are sized from the edit distance just
    //    found (penalty <= 10 per edit, about half of it per direction); a pair that still
    //    overflows is reported and re-run by the host with full-width rows.
    int W = a.W;
    if (!a.fixed_w) {
      const long long est = ((8ll * orr.distance + 512) + 63) & ~63ll;
      if (est < W) W = (int)est;
    }
    ws.W = W;
    ws.ring_r = ws.ring_f + (size_t)(a.pen.scope + 1) * 5 * W;
    SeqView sv;
    sv.p = P;
    sv.t = T;
    sv.lbits = lbits;
    // an alignment with `distance` unit edits costs at most max(x, cheapest one-base gap) per edit
    const long long per_edit = max(a.pen.x, min(a.pen.o1 + a.pen.e1, a.pen.o2 + a.pen.e2));
    if (a.leaf_T > 0) {
      // split mode: the run buffer of this CTA holds the leaves (32-byte frames) instead of runs
      Frame* leaves = reinterpret_cast<Frame*>(ro.runs);
      const uint32_t leaf_cap = (uint32_t)(((size_t)a.runs_cap * sizeof(uint32_t)) / sizeof(Frame));
      const PairResult rs = wf_align_from(cx, sv, wf_root_frame((int)plen, (int)tlen), a.pen, ws, ro,
                                          per_edit * (long long)orr.distance, leaves, leaf_cap, a.leaf_T);
      __syncthreads();
      if (threadIdx.x == 0) {
        int st = rs.status;
        unsigned int first = 0;
        if (st == ST_OK) {
          first = atomicAdd(a.task_count, rs.n_runs);
          if ((unsigned long long)first + rs.n_runs > a.task_cap) st = ST_TASK_OVERFLOW;
        }
        sh.bcast[3] = st;
        sh.bcast[4] = (int)first;
        SplitRec sr;
        sr.cells = sh.cells;
        sr.first_task = first;
        sr.n_tasks = st == ST_OK ? rs.n_runs : 0;
        sr.status = st;
        sr.score = rs.score;
        a.split[pi] = sr;
      }
      __syncthreads();
      if (sh.bcast[3] == ST_OK || sh.bcast[3] == ST_TASK_OVERFLOW) {
        // (on overflow the slots this pair was given below the capacity are filled all the same: the one-warp kernel
        // walks [0, min(count, cap)) and must find valid frames there; the pair itself is re-run monolithically)
        const unsigned int first = (unsigned int)sh.bcast[4];
        const uint32_t n_fit = first < a.task_cap ? min(rs.n_runs, a.task_cap - first) : 0u;
        for (uint32_t i = threadIdx.x; i < n_fit; i += THREADS) {
          a.tasks[first + i] = leaves[i];
          a.task_pair[first + i] = pi;
        }
      }
      continue;
    }
    const PairResult r = wf_align_pair(cx, sv, (int)plen, (int)tlen, a.pen, ws, ro, per_edit * (long long)orr.distance);
#ifdef WFA_PROFILE
    if (threadIdx.x == 0) {
      sh.prof[13] += t_or1 - t_or0;         // strand choice
      sh.prof[14] += clock64() - t_or1;     // BiWFA
      sh.prof[15] += (unsigned long long)orr.distance;
    }
#endif
    // 3. finalize: CIGAR statistics (allwave/src/alignment.rs:279-324), penalty
    //    (WFA2-lib/alignment/cigar.c:303-343), copy the runs into the chunk arena
    __syncthreads();
    if (threadIdx.x < 8) s_cnt[threadIdx.x] = 0;
    if (threadIdx.x == 0) {
      unsigned long long off = 0;
      int st = r.status;
      if (st == ST_OK) {
        off = atomicAdd(a.arena_head, (unsigned long long)r.n_runs);
        if (off + r.n_runs > a.arena_cap) st = ST_ARENA_OVERFLOW;
      }
      sh.bcast[3] = st;
      sh.bcast64[0] = off;
    }
    __syncthreads();
    const int st = sh.bcast[3];
    const unsigned long long off = sh.bcast64[0];
    uint32_t c_m = 0, c_x = 0, c_i = 0, c_d = 0, c_pen = 0;
    if (st == ST_OK) {
      for (uint32_t i = threadIdx.x; i < r.n_runs; i += THREADS) {
        const uint32_t run = ro.runs[i];
        SR_ASSERT(off + i < a.arena_cap, "CIGAR arena");
        a.arena[off + i] = run;
        const uint32_t op = run & 3u, len = run >> 2;
        if (op == OP_M) {
          c_m += len;
        } else if (op == OP_X) {
          c_x += len;
          c_pen += len * (uint32_t)a.pen.x;
        } else {
          if (op == OP_I) c_i += len; else c_d += len;
          const uint32_t s1 = (uint32_t)a.pen.o1 + (uint32_t)a.pen.e1 * len;
          const uint32_t s2 = (uint32_t)a.pen.o2 + (uint32_t)a.pen.e2 * len;
          c_pen += s1 < s2 ? s1 : s2;
        }
      }
      c_m = __reduce_add_sync(0xffffffffu, c_m);
      c_x = __reduce_add_sync(0xffffffffu, c_x);
      c_i = __reduce_add_sync(0xffffffffu, c_i);
      c_d = __reduce_add_sync(0xffffffffu, c_d);
      c_pen = __reduce_add_sync(0xffffffffu, c_pen);
      if ((threadIdx.x & 31) == 0) {
        atomicAdd(&s_cnt[0], c_m);
        atomicAdd(&s_cnt[1], c_x);
        atomicAdd(&s_cnt[2], c_i);
        atomicAdd(&s_cnt[3], c_d);
        atomicAdd(&s_cnt[4], c_pen);
      }
    }
    __syncthreads();
    if (threadIdx.x == 0) {
      PairOut o;
      o.q = pr.x;
      o.t = pr.y;
      o.is_reverse = orr.is_reverse;
      o.status = st;
      o.score = (st == ST_OK) ? r.score : INT32_MAX;
      o.penalty = (int32_t)s_cnt[4];
      o.num_matches = s_cnt[0];
      o.n_mismatch = s_cnt[1];
      o.n_ins = s_cnt[2];
      o.n_del = s_cnt[3];
      o.edit_distance = (uint32_t)orr.distance;
      o.n_runs = (st == ST_OK) ? r.n_runs : 0;
      o.run_off = off;
      o.cells = sh.cells;
      o.ocells = pair_ocells;
      a.out[pi] = o;
    }
  }
#ifdef WFA_PROFILE
  if (threadIdx.x == 0 && blockIdx.x < 4) prof_print(sh, prof_start, "align_kernel");
#endif
}

// ---------------------------------------------------------------------------
// Split mode, second launch: one task = one sub-problem (frame) of a pair, solved completely -- its own breakpoint
// searches, recursion, base cases, backtraces -- by a one-warp CTA (persistent CTAs pulling tasks from a counter).
// Narrow sub-problems leave three of four warps of a 128-thread CTA idle while every score step still pays three
// CTA barriers (measured per pair at 50 kb / 1 %: recursion levels >= 2 and base cases were 51 % of the cycles for
// 13 % of the cells, profiles/r02s_prof_phases.log); with a warp per task the "barriers" are warp-level and four
// times as many sub-problems are in flight. Sequences are read from global memory; the runs of a task go to the arena.
// ---------------------------------------------------------------------------
static constexpr int SUB_SLOTS = 32;  // ring slots of the one-warp kernel (scope + 1 <= 32, else the launch stays monolithic)
static constexpr int SUB_CTAS = 16;   // one-warp CTAs per SM (128 registers; the tables below keep shared memory at 10 KB per CTA)
static constexpr int SUB_STACK = 32;  // a task starts below split_T: a few levels deep at most; overflow -> monolithic retry
static constexpr unsigned int LEAF_RUNS_CAP = 8192;  // run / backtrace buffer of a leaf CTA (uint32 each)
static constexpr int LEAF_CTAS = 32;  // one-warp CTAs per SM of the leaf launch (<= 64 registers, no staging buffer)
template <int THREADS, int MIN_CTAS, bool LEAF>
__global__ void __launch_bounds__(THREADS, MIN_CTAS) subalign_kernel(AlignArgs a) {
  extern __shared__ __align__(16) uint32_t s_dyn[];  // row staging buffers of the score step (none when LEAF)
  __shared__ CtaShared sh;
  __shared__ CtaTables<LEAF ? 1 : SUB_SLOTS, LEAF ? 2 : SUB_STACK> s_tables;
  __shared__ unsigned int s_task;
  if (threadIdx.x == 0) cta_bind(&sh, &s_tables);
  Ctx cx;
  cx.sh = &sh;
  cx.tid = threadIdx.x;
  cx.nt = THREADS;
  cx.parity = 0;
  cx.stage = reinterpret_cast<int32_t*>(s_dyn);
  const size_t b = blockIdx.x;
  PairWorkspace ws;
  ws.ring_f = a.ring + b * a.ring_stride;
  ws.W = a.W;
  ws.ring_r = ws.ring_f + (size_t)(a.pen.scope + 1) * 5 * a.W;
  ws.full = a.full + b * a.full_stride;
  ws.full_meta = a.full_meta + b * (size_t)a.cap * 5;
  ws.cap = a.cap;
  RunOut ro;
  ro.runs = a.runs_tmp + b * (size_t)a.runs_cap;
  ro.cap = a.runs_cap;
  ro.bt = a.bt_tmp + b * (size_t)a.runs_cap;
  ro.bt_cap = a.runs_cap;
#ifdef WFA_PROFILE
  if (threadIdx.x == 0)
    for (int i = 0; i < 20; ++i) sh.prof[i] = sh.prof2[i] = 0;
  const long long prof_start = clock64();
#endif
  const unsigned int n_tasks = min(*a.task_count, a.task_cap);
  while (true) {
    __syncthreads();
    if (threadIdx.x == 0) s_task = atomicAdd(a.next_task, 1u);
    __syncthreads();
    const unsigned int ti = s_task;
    if (ti >= n_tasks) break;
    const Frame f = a.tasks[ti];
    const uint32_t pi = a.task_pair[ti];
    const uint2 pr = a.pairs[pi];
    const int is_rev = a.out[pi].is_reverse;
    SeqView sv;
    sv.p = (is_rev ? a.seqs.pk_rc : a.seqs.pk_fwd) + a.seqs.pk_off[pr.x];
    sv.t = a.seqs.pk_fwd + a.seqs.pk_off[pr.y];
    sv.lbits = a.seqs.lbits;
    const long long per_edit = max(a.pen.x, min(a.pen.o1 + a.pen.e1, a.pen.o2 + a.pen.e2));
    if (!LEAF && a.etasks != nullptr) {
      // first pass of the two-pass mode: bisect only; the run buffer holds the emitted leaves (32-byte frames)
      Frame* leaves = reinterpret_cast<Frame*>(ro.runs);
      const uint32_t leaf_cap = (uint32_t)(((size_t)a.runs_cap * sizeof(uint32_t)) / sizeof(Frame));
      const PairResult rs = wf_align_from(cx, sv, f, a.pen, ws, ro, per_edit * (long long)a.out[pi].edit_distance, leaves,
                                          leaf_cap, a.leaf_T2);
      __syncthreads();
      if (threadIdx.x == 0) {
        int st = rs.status;
        if (st == ST_STACK_OVERFLOW) st = ST_WS_DIAG_OVERFLOW;  // this kernel's small stack: monolithic retry
        unsigned int first = 0;
        if (st == ST_OK) {
          first = atomicAdd(a.etask_count, rs.n_runs);
          if ((unsigned long long)first + rs.n_runs > a.etask_cap) st = ST_TASK_OVERFLOW;
        }
        sh.bcast[3] = st;
        sh.bcast[4] = (int)first;
        SplitRec sr;
        sr.cells = sh.cells;
        sr.first_task = first;
        sr.n_tasks = st == ST_OK ? rs.n_runs : 0;
        sr.status = st;
        sr.score = rs.score;
        a.esplit[ti] = sr;
      }
      __syncthreads();
      if (sh.bcast[3] == ST_OK || sh.bcast[3] == ST_TASK_OVERFLOW) {
        const unsigned int first = (unsigned int)sh.bcast[4];
        const uint32_t n_fit = first < a.etask_cap ? min(rs.n_runs, a.etask_cap - first) : 0u;
        for (uint32_t i = threadIdx.x; i < n_fit; i += THREADS) {
          a.etasks[first + i] = leaves[i];
          a.etask_pair[first + i] = pi;
        }
      }
      continue;
    }
    const PairResult r = LEAF ? wf_align_leaf(cx, sv, f, a.pen, ws, ro)
                              : wf_align_from(cx, sv, f, a.pen, ws, ro, per_edit * (long long)a.out[pi].edit_distance, nullptr, 0, 0);
    __syncthreads();
    if (threadIdx.x == 0) {
      unsigned long long off = 0;
      int st = r.status;
      if (st == ST_STACK_OVERFLOW) st = ST_WS_DIAG_OVERFLOW;  // this kernel's small stack: the pair is retried monolithically
      if (LEAF && st == ST_WS_RUNS_OVERFLOW) st = ST_TASK_OVERFLOW;  // the leaf launch's short run buffers: same
      if (st == ST_OK) {
        off = atomicAdd(a.arena_head, (unsigned long long)r.n_runs);
        if (off + r.n_runs > a.arena_cap) st = ST_ARENA_OVERFLOW;
      }
      TaskOut to;
      to.run_off = off;
      to.cells = sh.cells;
      to.n_runs = st == ST_OK ? r.n_runs : 0;
      to.status = st;
      to.score = r.score;
      to.pad = 0;
      a.task_out[ti] = to;
      sh.bcast[3] = st;
      sh.bcast64[0] = off;
    }
    __syncthreads();
    if (sh.bcast[3] == ST_OK) {
      const unsigned long long off = sh.bcast64[0];
      for (uint32_t i = threadIdx.x; i < r.n_runs; i += THREADS) a.arena[off + i] = ro.runs[i];
    }
  }
#ifdef WFA_PROFILE
  if (threadIdx.x == 0 && blockIdx.x < 4) prof_print(sh, prof_start, "subalign_kernel");
#endif
}

// ---------------------------------------------------------------------------
// Split mode, third launch: one CTA per pair joins the run lists of its tasks in order into one canonical list
// (adjacent runs never share an op: a piece whose first run continues the previous piece's last run is merged into
// it, like cigar_append_forward does byte by byte, WFA2-lib/alignment/cigar.c:125-137), computes the CIGAR
// statistics (allwave/src/alignment.rs:279-324) and the penalty, and writes the pair's result record.
// ---------------------------------------------------------------------------
static constexpr int ASM_MAX_TASKS = 1024;
template <int THREADS>
__global__ void __launch_bounds__(THREADS) assemble_kernel(AlignArgs a) {
  __shared__ unsigned long long s_src[ASM_MAX_TASKS];  // first copied run of the piece
  __shared__ uint32_t s_dst[ASM_MAX_TASKS], s_cnt_t[ASM_MAX_TASKS];
  __shared__ uint32_t s_patch_pos[ASM_MAX_TASKS], s_patch_val[ASM_MAX_TASKS];
  __shared__ unsigned int s_cnt[8];
  __shared__ unsigned long long s_off, s_cells;
  __shared__ int s_st, s_score;
  __shared__ uint32_t s_total, s_npatch, s_npieces;
  for (uint32_t slot = blockIdx.x; slot < a.n_pairs; slot += gridDim.x) {
    const uint32_t pi = a.todo ? a.todo[slot] : slot;
    const SplitRec sr = a.split[pi];
    __syncthreads();
    if (threadIdx.x < 8) s_cnt[threadIdx.x] = 0;
    if (threadIdx.x == 0) {
      int st = sr.status;
      int score = sr.score;
      unsigned long long cells = sr.cells;
      uint32_t out_n = 0, npatch = 0;
      if (st == ST_OK && sr.n_tasks > a.asm_max) st = ST_TASK_OVERFLOW;  // more pieces than this kernel joins: monolithic retry
      s_npieces = 0;
      if (st == ST_OK) {
        bool have_last = false;
        uint32_t last_pos = 0, last_val = 0;
        uint32_t n_pieces = 0;  // pieces = tasks (one-pass) or leaves (two-pass), in alignment order
        const bool two = a.etasks != nullptr;
        for (uint32_t t = 0; t < sr.n_tasks && st == ST_OK; ++t) {
          SplitRec lr;  // the pieces of task t
          lr.first_task = sr.first_task + t;
          lr.n_tasks = 1;
          if (two) {
            lr = a.esplit[sr.first_task + t];
            cells += lr.cells;
            if (lr.status != ST_OK) {
              st = lr.status;
              break;
            }
            if (a.tasks[sr.first_task + t].level == 0 && lr.score != INT32_MIN) score = lr.score;  // whole pair, bisected there
            if (n_pieces + lr.n_tasks > a.asm_max) {
              st = ST_TASK_OVERFLOW;  // (long, divergent pairs: > 1024 leaves) the pair is re-run monolithically
              break;
            }
          }
          for (uint32_t l = 0; l < lr.n_tasks; ++l) {
            const TaskOut to = two ? a.etask_out[lr.first_task + l] : a.task_out[lr.first_task + l];
            const Frame tf = two ? a.etasks[lr.first_task + l] : a.tasks[lr.first_task + l];
            const uint32_t x = n_pieces++;
            cells += to.cells;
            s_cnt_t[x] = 0;
            s_dst[x] = out_n;
            s_src[x] = to.run_off;
            if (to.status != ST_OK) {
              if (st == ST_OK || to.status == ST_WS_DIAG_OVERFLOW) st = to.status;
              continue;
            }
            if (tf.level == 0) score = to.score;  // the whole pair was one piece
            if (to.n_runs == 0) continue;
            const uint32_t r0 = a.arena[to.run_off];
            uint32_t skip = 0;
            if (have_last && ((r0 ^ last_val) & 3u) == 0) {
              last_val += (r0 >> 2) << 2;
              skip = 1;
            }
            if (to.n_runs - skip == 0) continue;  // the whole piece went into the open run
            if (have_last) {
              s_patch_pos[npatch] = last_pos;
              s_patch_val[npatch++] = last_val;
            }
            s_src[x] = to.run_off + skip;
            s_cnt_t[x] = to.n_runs - skip;
            out_n += to.n_runs - skip;
            last_pos = out_n - 1;
            last_val = a.arena[to.run_off + to.n_runs - 1];
            have_last = true;
          }
        }
        s_npieces = n_pieces;
        if (have_last) {
          s_patch_pos[npatch] = last_pos;
          s_patch_val[npatch++] = last_val;
        }
      }
      unsigned long long off = 0;
      if (st == ST_OK) {
        off = atomicAdd(a.arena_head, (unsigned long long)out_n);
        if (off + out_n > a.arena_cap) st = ST_ARENA_OVERFLOW;
      }
      s_st = st;
      s_score = score;
      s_cells = cells;
      s_off = off;
      s_total = out_n;
      s_npatch = npatch;
    }
    __syncthreads();
    const int st = s_st;
    const unsigned long long off = s_off;
    if (st == ST_OK) {
      for (uint32_t t = 0; t < s_npieces; ++t) {
        const uint32_t n = s_cnt_t[t], d = s_dst[t];
        const unsigned long long src = s_src[t];
        for (uint32_t i = threadIdx.x; i < n; i += THREADS) a.arena[off + d + i] = a.arena[src + i];
      }
      __syncthreads();
      for (uint32_t i = threadIdx.x; i < s_npatch; i += THREADS) a.arena[off + s_patch_pos[i]] = s_patch_val[i];
      __syncthreads();
      uint32_t c_m = 0, c_x = 0, c_i = 0, c_d = 0, c_pen = 0;
      for (uint32_t i = threadIdx.x; i < s_total; i += THREADS) {
        const uint32_t run = a.arena[off + i];
        const uint32_t op = run & 3u, len = run >> 2;
        if (op == OP_M) {
          c_m += len;
        } else if (op == OP_X) {
          c_x += len;
          c_pen += len * (uint32_t)a.pen.x;
        } else {
          if (op == OP_I) c_i += len; else c_d += len;
          const uint32_t s1 = (uint32_t)a.pen.o1 + (uint32_t)a.pen.e1 * len;
          const uint32_t s2 = (uint32_t)a.pen.o2 + (uint32_t)a.pen.e2 * len;
          c_pen += s1 < s2 ? s1 : s2;
        }
      }
      c_m = __reduce_add_sync(0xffffffffu, c_m);
      c_x = __reduce_add_sync(0xffffffffu, c_x);
      c_i = __reduce_add_sync(0xffffffffu, c_i);
      c_d = __reduce_add_sync(0xffffffffu, c_d);
      c_pen = __reduce_add_sync(0xffffffffu, c_pen);
      if ((threadIdx.x & 31) == 0) {
        atomicAdd(&s_cnt[0], c_m);
        atomicAdd(&s_cnt[1], c_x);
        atomicAdd(&s_cnt[2], c_i);
        atomicAdd(&s_cnt[3], c_d);
        atomicAdd(&s_cnt[4], c_pen);
      }
    }
    __syncthreads();
    if (threadIdx.x == 0) {
      const uint2 pr = a.pairs[pi];
      PairOut o;
      o.q = pr.x;
      o.t = pr.y;
      o.is_reverse = a.out[pi].is_reverse;
      o.edit_distance = a.out[pi].edit_distance;
      o.ocells = a.out[pi].ocells;
      o.status = st;
      o.score = (st == ST_OK) ? s_score : INT32_MAX;
      o.penalty = (int32_t)s_cnt[4];
      o.num_matches = s_cnt[0];
      o.n_mismatch = s_cnt[1];
      o.n_ins = s_cnt[2];
      o.n_del = s_cnt[3];
      o.n_runs = (st == ST_OK) ? s_total : 0;
      o.run_off = off;
      o.cells = s_cells;
      a.out[pi] = o;
    }
  }
}

// ---------------------------------------------------------------------------
// Pack every sequence, forward and allwave reverse complement (allwave/src/alignment.rs:174-186:
// ACGT/acgt -> upper-case complement, anything else -> 'N'), at 1 << lbits bits per symbol through
// an injective byte -> code table (identity for 8 bits). One thread per output word.
// ---------------------------------------------------------------------------
struct CodeTable {
  uint8_t code[256];
};
__device__ __forceinline__ uint8_t allwave_comp(uint8_t c) {
  switch (c) {
    case 'A': case 'a': return 'T';
    case 'T': case 't': return 'A';
    case 'C': case 'c': return 'G';
    case 'G': case 'g': return 'C';
    default: return 'N';
  }
}
__global__ void pack_kernel(const uint8_t* __restrict__ fwd, const uint64_t* __restrict__ byte_off,
                            const uint32_t* __restrict__ len, const uint64_t* __restrict__ pk_off,
                            uint32_t* __restrict__ pk_fwd, uint32_t* __restrict__ pk_rc, CodeTable tab, int lbits, uint32_t n) {
  __shared__ uint8_t s_code[256];
  for (int i = threadIdx.x; i < 256; i += blockDim.x) s_code[i] = tab.code[i];
  __syncthreads();
  const uint32_t per = 32u >> lbits, bits = 1u << lbits;
  for (uint32_t s = blockIdx.y; s < n; s += gridDim.y) {
    const uint32_t L = len[s];
    const uint32_t nw = (L + per - 1) / per;
    const uint8_t* src = fwd + byte_off[s];
    uint32_t* df = pk_fwd + pk_off[s];
    uint32_t* dr = pk_rc + pk_off[s];
    for (uint32_t w = blockIdx.x * blockDim.x + threadIdx.x; w < nw; w += gridDim.x * blockDim.x) {
      uint32_t vf = 0, vr = 0;
      for (uint32_t j = 0; j < per; ++j) {
        const uint32_t i = w * per + j;
        if (i < L) {
          vf |= (uint32_t)s_code[src[i]] << (j * bits);
          vr |= (uint32_t)s_code[allwave_comp(src[L - 1 - i])] << (j * bits);
        }
      }
      df[w] = vf;
      dr[w] = vr;
    }
  }
}

// seqrush's per-base query RC (src/seqrush.rs:1162-1176): unknown bytes are kept
__device__ __forceinline__ uint8_t seqrush_query_base(const uint8_t* q, uint32_t qlen, uint32_t local_pos, bool rc) {
  if (!rc) return q[local_pos];
  const uint8_t c = q[qlen - 1 - local_pos];
  switch (c) {
    case 'A': case 'a': return 'T';
    case 'T': case 't': return 'A';
    case 'C': case 'c': return 'G';
    case 'G': case 'g': return 'C';
    default: return c;
  }
}

// ---------------------------------------------------------------------------
// Union kernel: CIGAR walk + unite (src/seqrush.rs:1134-1481 + src/bidirected_union_find.rs:60-98).
// The run list is canonical, so one 'M' run is exactly one of the reference's match runs (its
// run accumulator spans consecutive M/= ops, :1214-1218, and is flushed by any other op).
// Forward pairs and RC pairs whose query is rc_safe need no per-base re-check: WFA 'M' columns
// compared the very same bytes. Other RC pairs take the serial per-base path.
// ---------------------------------------------------------------------------
struct UnionArgs {
  SeqTable seqs;
  const PairOut* out;
  const uint32_t* arena;
  uint32_t n_pairs;
  const uint32_t* todo;  // optional indirection (retry lists)
  uint32_t* parent;
  unsigned long long min_match_len;
  unsigned long long* counters;  // [0] unite ops, [1] links
};

static constexpr int UNION_TILE = 1024;

template <int THREADS>
__global__ void __launch_bounds__(THREADS) union_kernel(UnionArgs a) {
  __shared__ uint32_t s_qpos[UNION_TILE];   // query position at run start (alignment coordinates)
  __shared__ uint32_t s_tpos[UNION_TILE];
  __shared__ uint32_t s_cum[UNION_TILE + 1];  // exclusive prefix of united bases inside the tile
  __shared__ uint32_t s_scan[3][THREADS];
  __shared__ uint32_t s_carry[3];
  unsigned long long ops = 0, links = 0;
  for (uint32_t slot = blockIdx.x; slot < a.n_pairs; slot += gridDim.x) {
    const uint32_t pi = a.todo ? a.todo[slot] : slot;
    const PairOut po = a.out[pi];
    if (po.status != ST_OK || po.n_runs == 0) continue;
    if (po.q == po.t && !po.is_reverse) {
      // self pair aligned forward: every unite(p,p) is skipped by the reference
      // (src/bidirected_union_find.rs:40); still counted as issued unite calls
      if (threadIdx.x == 0) {
        unsigned long long m = 0;
        for (uint32_t i = 0; i < po.n_runs; ++i) {
          const uint32_t run = a.arena[po.run_off + i];
          if ((run & 3u) == OP_M && (run >> 2) >= a.min_match_len) m += run >> 2;
        }
        ops += m;
      }
      continue;
    }
    const uint32_t qlen = a.seqs.len[po.q];
    const uint64_t qoff = a.seqs.pos_off[po.q], toff = a.seqs.pos_off[po.t];
    const bool rc = po.is_reverse != 0;
    if (rc && !a.seqs.rc_safe[po.q]) {
      // serial exact path with the per-base equality re-check (src/seqrush.rs:1267-1358)
      if (threadIdx.x == 0) {
        const uint8_t* q = a.seqs.fwd + a.seqs.byte_off[po.q];
        const uint8_t* t = a.seqs.fwd + a.seqs.byte_off[po.t];
        uint32_t p1 = 0, p2 = 0;
        bool in_run = false;
        uint32_t rs1 = 0, rs2 = 0, rl = 0;
        for (uint32_t i = 0; i <= po.n_runs; ++i) {
          const bool last = (i == po.n_runs);
          const uint32_t run = last ? 0 : a.arena[po.run_off + i];
          const uint32_t op = run & 3u, len = run >> 2;
          if (!last && op == OP_M) {
            for (uint32_t k = 0; k < len; ++k) {
              const bool eq = seqrush_query_base(q, qlen, p1 + k, true) == t[p2 + k];
              if (eq) {
                if (!in_run) {
                  in_run = true;
                  rs1 = p1 + k;
                  rs2 = p2 + k;
                  rl = 1;
                } else {
                  ++rl;
                }
              }
              if (!eq) {
                if (in_run && rl >= a.min_match_len) {
                  for (uint32_t j = 0; j < rl; ++j) {
                    const uint32_t x = (uint32_t)(2 * (qoff + qlen - 1 - (rs1 + j)) + 1);
                    const uint32_t y = (uint32_t)(2 * (toff + rs2 + j));
                    ++ops;
                    if (uf_unite(a.parent, x, y)) ++links;
                  }
                }
                in_run = false;
                rl = 0;
              }
            }
            p1 += len;
            p2 += len;
          } else {
            if (in_run && rl >= a.min_match_len) {
              for (uint32_t j = 0; j < rl; ++j) {
                const uint32_t x = (uint32_t)(2 * (qoff + qlen - 1 - (rs1 + j)) + 1);
                const uint32_t y = (uint32_t)(2 * (toff + rs2 + j));
                ++ops;
                if (uf_unite(a.parent, x, y)) ++links;
              }
            }
            in_run = false;
            rl = 0;
            if (op == OP_X) {
              p1 += len;
              p2 += len;
            } else if (op == OP_D) {
              p1 += len;  // WFA 'D' consumes the query (allwave/src/alignment.rs:304-319)
            } else if (op == OP_I) {
              p2 += len;
            }
          }
        }
      }
      continue;
    }
    // parallel path
    __syncthreads();
    if (threadIdx.x < 3) s_carry[threadIdx.x] = 0;
    __syncthreads();
    for (uint32_t tile0 = 0; tile0 < po.n_runs; tile0 += UNION_TILE) {
      const uint32_t nt = min((uint32_t)UNION_TILE, po.n_runs - tile0);
      constexpr int PER = UNION_TILE / THREADS;
      // each thread owns PER consecutive runs
      uint32_t runs[PER];
      uint32_t dq = 0, dt = 0, du = 0;
#pragma unroll
      for (int j = 0; j < PER; ++j) {
        const uint32_t i = threadIdx.x * PER + j;
        const uint32_t run = (i < nt) ? a.arena[po.run_off + tile0 + i] : 0u;
        runs[j] = run;
        const uint32_t op = run & 3u, len = run >> 2;
        if (op == OP_M || op == OP_X) {
          dq += len;
          dt += len;
        } else if (op == OP_D) {
          dq += len;
        } else {
          dt += len;
        }
        if (op == OP_M && len >= a.min_match_len) du += len;
      }
      s_scan[0][threadIdx.x] = dq;
      s_scan[1][threadIdx.x] = dt;
      s_scan[2][threadIdx.x] = du;
      __syncthreads();
      // Hillis-Steele inclusive scan over THREADS partials (3 lanes of data)
      for (int ofs = 1; ofs < THREADS; ofs <<= 1) {
        uint32_t v0 = 0, v1 = 0, v2 = 0;
        if ((int)threadIdx.x >= ofs) {
          v0 = s_scan[0][threadIdx.x - ofs];
          v1 = s_scan[1][threadIdx.x - ofs];
          v2 = s_scan[2][threadIdx.x - ofs];
        }
        __syncthreads();
        s_scan[0][threadIdx.x] += v0;
        s_scan[1][threadIdx.x] += v1;
        s_scan[2][threadIdx.x] += v2;
        __syncthreads();
      }
      uint32_t q0 = s_carry[0] + s_scan[0][threadIdx.x] - dq;
      uint32_t t0 = s_carry[1] + s_scan[1][threadIdx.x] - dt;
      uint32_t u0 = s_scan[2][threadIdx.x] - du;
      const uint32_t tile_total = s_scan[2][THREADS - 1];
      const uint32_t tq = s_scan[0][THREADS - 1], tt = s_scan[1][THREADS - 1];
#pragma unroll
      for (int j = 0; j < PER; ++j) {
        const uint32_t i = threadIdx.x * PER + j;
        const uint32_t run = runs[j];
        const uint32_t op = run & 3u, len = run >> 2;
        s_qpos[i] = q0;
        s_tpos[i] = t0;
        s_cum[i] = u0;
        if (op == OP_M || op == OP_X) {
          q0 += len;
          t0 += len;
        } else if (op == OP_D) {
          q0 += len;
        } else {
          t0 += len;
        }
        if (op == OP_M && len >= a.min_match_len) u0 += len;
      }
      if (threadIdx.x == THREADS - 1) s_cum[UNION_TILE] = tile_total;
      __syncthreads();
      // flattened loop over the united bases of this tile: consecutive threads take consecutive
      // bases, so the parent[] reads of a warp fall into one or two 128-byte lines per side
      // run index of base u = largest i with s_cum[i] <= u; u grows by THREADS per iteration, so the
      // index only moves forward: a short linear advance replaces a binary search per base
      // Four bases per thread and iteration: their eight parent reads are issued together (the
      // kernel is bound by the latency of these scattered 4-byte reads, not by bandwidth).
      uint32_t lo = 0;
      constexpr int UNR = 4;
      for (uint32_t u0 = threadIdx.x; u0 < tile_total; u0 += UNR * THREADS) {
        uint32_t x[UNR], y[UNR], px[UNR], py[UNR];
#pragma unroll
        for (int j = 0; j < UNR; ++j) {
          const uint32_t u = u0 + j * THREADS;
          x[j] = y[j] = 0;
          if (u < tile_total) {
            while (s_cum[lo + 1] <= u) ++lo;
            const uint32_t within = u - s_cum[lo];
            const uint32_t p1 = s_qpos[lo] + within, p2 = s_tpos[lo] + within;
            x[j] = rc ? (uint32_t)(2 * (qoff + qlen - 1 - p1) + 1) : (uint32_t)(2 * (qoff + p1));
            y[j] = (uint32_t)(2 * (toff + p2));
            ++ops;
          }
          px[j] = a.parent[x[j]];
          py[j] = a.parent[y[j]];
        }
#pragma unroll
        for (int j = 0; j < UNR; ++j)
          if (uf_unite_from(a.parent, x[j], y[j], px[j], py[j])) ++links;
      }
      __syncthreads();
      if (threadIdx.x == 0) {
        s_carry[0] += tq;
        s_carry[1] += tt;
      }
      __syncthreads();
    }
  }
  // counters
  ops = __reduce_add_sync(0xffffffffu, (unsigned)(ops & 0xffffffffu)) +
        ((unsigned long long)__reduce_add_sync(0xffffffffu, (unsigned)(ops >> 32)) << 32);
  links = __reduce_add_sync(0xffffffffu, (unsigned)(links & 0xffffffffu)) +
          ((unsigned long long)__reduce_add_sync(0xffffffffu, (unsigned)(links >> 32)) << 32);
  if ((threadIdx.x & 31) == 0) {
    if (ops) atomicAdd(&a.counters[0], ops);
    if (links) atomicAdd(&a.counters[1], links);
  }
}

// ---------------------------------------------------------------------------
// run-length CIGAR -> bytes. One CTA per pair; byte_off[pair] precomputed on the host.
// ---------------------------------------------------------------------------
__global__ void expand_kernel(const PairOut* __restrict__ out, const uint32_t* __restrict__ arena,
                              const unsigned long long* __restrict__ byte_off, uint8_t* __restrict__ bytes,
                              uint32_t n_pairs) {
  __shared__ unsigned long long s_pos;
  for (uint32_t pi = blockIdx.x; pi < n_pairs; pi += gridDim.x) {
    const PairOut po = out[pi];
    if (po.status != ST_OK) continue;
    uint8_t* dst = bytes + byte_off[pi];
    // runs are few (thousands); walk them in order, all threads fill each run
    unsigned long long pos = 0;
    for (uint32_t i = 0; i < po.n_runs; ++i) {
      const uint32_t run = arena[po.run_off + i];
      const uint32_t len = run >> 2;
      const uint8_t ch = "MXID"[run & 3u];
      for (uint32_t j = threadIdx.x; j < len; j += blockDim.x) dst[pos + j] = ch;
      pos += len;
    }
    (void)s_pos;
  }
}

// ---------------------------------------------------------------------------
// BidirectedUnionFind kernels
// ---------------------------------------------------------------------------
__global__ void uf_init_kernel(uint32_t* parent, uint64_t n) {
  for (uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x)
    parent[i] = (uint32_t)i;
}
__global__ void uf_unite_fr_kernel(uint32_t* parent, uint64_t total_len, unsigned long long* counters) {
  unsigned long long links = 0;
  for (uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; i < total_len; i += (uint64_t)gridDim.x * blockDim.x)
    if (uf_unite(parent, (uint32_t)(2 * i), (uint32_t)(2 * i + 1))) ++links;
  if (links) atomicAdd(&counters[1], links);
}
struct RegionDev {
  unsigned long long seq1_offset, seq2_offset, s1, s2, len, seq1_len;
  int seq1_is_rc, seq2_rc_form;
};
__global__ void uf_regions_kernel(uint32_t* parent, const RegionDev* regs, uint64_t n, uint64_t capacity,
                                  unsigned long long* counters) {
  for (uint64_t r = blockIdx.x; r < n; r += gridDim.x) {
    const RegionDev g = regs[r];
    for (unsigned long long i = threadIdx.x; i < g.len; i += blockDim.x) {
      unsigned long long x, y;
      if (!g.seq2_rc_form) {
        // src/bidirected_union_find.rs:73-97
        if (g.seq1_is_rc) {
          x = 2 * (g.seq1_offset + (g.seq1_len - 1 - (g.s1 + i))) + 1;
          y = 2 * (g.seq2_offset + g.s2 + i);
        } else {
          x = 2 * (g.seq1_offset + g.s1 + i);
          y = 2 * (g.seq2_offset + g.s2 + i);
        }
      } else {
        // src/bidirected_union_find.rs:113-128
        x = 2 * (g.seq1_offset + g.s1 + i);
        if (g.seq1_is_rc) y = 2 * (g.seq2_offset + (g.seq1_len - 1 - (g.s2 + i))) + 1;
        else y = 2 * (g.seq2_offset + g.s2 + i);
      }
      if (x >= capacity || y >= capacity) {
        atomicAdd(&counters[3], 1ull);
        continue;
      }
      atomicAdd(&counters[0], 1ull);
      if (uf_unite(parent, (uint32_t)x, (uint32_t)y)) atomicAdd(&counters[1], 1ull);
    }
  }
}
__global__ void uf_roots_kernel(uint32_t* parent, uint64_t first, uint64_t n, uint64_t* out) {
  for (uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x)
    out[i] = uf_find(parent, (uint32_t)(first + i));
}
__global__ void uf_compress_kernel(uint32_t* parent, uint64_t n) {
  for (uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x) {
    const uint32_t r = uf_find(parent, (uint32_t)i);
    if (r != (uint32_t)i) parent[i] = r;
  }
}
__global__ void uf_merge_kernel(uint32_t* parent, const uint32_t* peer, uint64_t n, unsigned long long* counters) {
  unsigned long long links = 0;
  for (uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x) {
    const uint32_t r = peer[i];
    if (r != (uint32_t)i && r < n && uf_unite(parent, (uint32_t)i, r)) ++links;
  }
  if (links) atomicAdd(&counters[1], links);
}
// merge of the gathered forests of every rank in ONE pass (gathered = [world][n] compressed parent arrays)
__global__ void uf_merge_gathered_kernel(uint32_t* parent, const uint32_t* __restrict__ gathered, uint64_t n, int world,
                                         int me, unsigned long long* counters) {
  unsigned long long links = 0;
  for (uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x) {
    for (int r = 0; r < world; ++r) {
      if (r == me) continue;
      const uint32_t root = gathered[(uint64_t)r * n + i];
      if (root != (uint32_t)i && root < n && uf_unite(parent, (uint32_t)i, root)) ++links;
    }
  }
  if (links) atomicAdd(&counters[1], links);
}
__global__ void uf_count_kernel(uint32_t* parent, uint64_t total_len, uint32_t* bitmap, unsigned long long* count) {
  // distinct representatives over forward Pos (src/seqrush.rs:341-353): mark each root once
  unsigned long long c = 0;
  for (uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; i < total_len; i += (uint64_t)gridDim.x * blockDim.x) {
    const uint32_t r = uf_find(parent, (uint32_t)(2 * i));
    const uint32_t bit = 1u << (r & 31);
    if (!(atomicOr(&bitmap[r >> 5], bit) & bit)) ++c;
  }
  if (c) atomicAdd(count, c);
}
__device__ __forceinline__ unsigned long long mix64(unsigned long long x) {
  x ^= x >> 33;
  x *= 0xff51afd7ed558ccdULL;
  x ^= x >> 33;
  x *= 0xc4ceb9fe1a85ec53ULL;
  x ^= x >> 33;
  return x;
}
__global__ void uf_checksum_kernel(uint32_t* parent, uint64_t n, unsigned long long* sum) {
  unsigned long long c = 0;
  for (uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x)
    c += mix64((i << 32) ^ (unsigned long long)uf_find(parent, (uint32_t)i));
  if (c) atomicAdd(sum, c);
}

// ---------------------------------------------------------------------------
// Batched CIGAR walk for alignments that come from elsewhere (--paf, src/seqrush.rs:510-609, and
// any caller of process_alignment, :1134-1481). One CTA per record walks a run list the host parsed
// from the RLE string: op 0 = 'M' or '=' (re-checked base by base like the reference, :1267-1358;
// adjacent M/= ops are merged by the host, they continue one match run, :1214-1218), 1 = 'X',
// 2 = 'I' (query += n), 3 = 'D' (target += n), 4 = any other letter (flushes the run, moves
// nothing, :1436-1437). Tiles of runs are prefix-summed like in union_kernel; the bases of the
// match-type runs of a tile are then flattened in 32-base words over the threads. A thread compares
// its word's bases, and a base is united when the maximal stretch of equal bases around it (inside
// its run: every other op flushes) is at least min_match_len long; only stretches that touch a word
// edge look at the neighbouring bases, at most min_match_len of them.
// ---------------------------------------------------------------------------
struct PafRecordDev {
  uint32_t q, t;
  uint32_t query_is_rc, n_runs;
  unsigned long long qstart, tstart;
  unsigned long long run_off;
};
struct PafArgs {
  SeqTable seqs;
  const PafRecordDev* recs;
  const unsigned long long* runs;  // (count << 3) | op
  uint32_t n_recs;
  uint32_t* parent;
  unsigned long long min_match_len;
  unsigned long long* counters;  // [0] unite ops, [1] links
};

template <int THREADS>
__global__ void __launch_bounds__(THREADS) paf_walk_kernel(PafArgs a) {
  constexpr int TILE = 1024, PER = TILE / THREADS;
  __shared__ unsigned long long s_qpos[TILE], s_tpos[TILE];
  __shared__ uint32_t s_len[TILE];       // bases of a match-type run that lie inside both sequences (0 for other ops)
  __shared__ uint32_t s_wcum[TILE + 1];  // exclusive prefix of 32-base words
  __shared__ unsigned long long s_scan[2][THREADS];
  __shared__ uint32_t s_scanw[THREADS];
  __shared__ unsigned long long s_carry[2];
  unsigned long long ops = 0, links = 0;
  for (uint32_t ri = blockIdx.x; ri < a.n_recs; ri += gridDim.x) {
    const PafRecordDev rec = a.recs[ri];
    const uint8_t* q = a.seqs.fwd + a.seqs.byte_off[rec.q];
    const uint8_t* t = a.seqs.fwd + a.seqs.byte_off[rec.t];
    const unsigned long long qlen = a.seqs.len[rec.q], tlen = a.seqs.len[rec.t];
    const unsigned long long qoff = a.seqs.pos_off[rec.q], toff = a.seqs.pos_off[rec.t];
    const bool rc = rec.query_is_rc != 0;
    const unsigned long long k = a.min_match_len;
    __syncthreads();
    if (threadIdx.x == 0) {
      s_carry[0] = rec.qstart;
      s_carry[1] = rec.tstart;
    }
    __syncthreads();
    for (uint32_t tile0 = 0; tile0 < rec.n_runs; tile0 += TILE) {
      const uint32_t nt = min((uint32_t)TILE, rec.n_runs - tile0);
      unsigned long long runs[PER];
      unsigned long long dq = 0, dt = 0;
#pragma unroll
      for (int j = 0; j < PER; ++j) {
        const uint32_t i = threadIdx.x * PER + j;
        const unsigned long long run = (i < nt) ? a.runs[rec.run_off + tile0 + i] : 4ull;
        runs[j] = run;
        const unsigned op = (unsigned)(run & 7ull);
        const unsigned long long len = run >> 3;
        if (op <= 1u) {
          dq += len;
          dt += len;
        } else if (op == 2u) {
          dq += len;
        } else if (op == 3u) {
          dt += len;
        }
      }
      s_scan[0][threadIdx.x] = dq;
      s_scan[1][threadIdx.x] = dt;
      __syncthreads();
      for (int ofs = 1; ofs < THREADS; ofs <<= 1) {
        unsigned long long v0 = 0, v1 = 0;
        if ((int)threadIdx.x >= ofs) {
          v0 = s_scan[0][threadIdx.x - ofs];
          v1 = s_scan[1][threadIdx.x - ofs];
        }
        __syncthreads();
        s_scan[0][threadIdx.x] += v0;
        s_scan[1][threadIdx.x] += v1;
        __syncthreads();
      }
      unsigned long long q0 = s_carry[0] + s_scan[0][threadIdx.x] - dq;
      unsigned long long t0 = s_carry[1] + s_scan[1][threadIdx.x] - dt;
      const unsigned long long tq = s_scan[0][THREADS - 1], tt = s_scan[1][THREADS - 1];
      uint32_t dw = 0;
#pragma unroll
      for (int j = 0; j < PER; ++j) {
        const uint32_t i = threadIdx.x * PER + j;
        const unsigned op = (unsigned)(runs[j] & 7ull);
        const unsigned long long len = runs[j] >> 3;
        s_qpos[i] = q0;
        s_tpos[i] = t0;
        uint32_t inside = 0;
        if (op == 0u && q0 < qlen && t0 < tlen) {
          // positions beyond either sequence are skipped without ending the run (:1268); they can
          // only be the tail of a run, so clipping the run is the same thing
          unsigned long long m = len;
          if (qlen - q0 < m) m = qlen - q0;
          if (tlen - t0 < m) m = tlen - t0;
          inside = (uint32_t)m;
        }
        s_len[i] = inside;
        dw += (inside + 31u) >> 5;
        if (op <= 1u) {
          q0 += len;
          t0 += len;
        } else if (op == 2u) {
          q0 += len;
        } else if (op == 3u) {
          t0 += len;
        }
      }
      s_scanw[threadIdx.x] = dw;
      __syncthreads();
      for (int ofs = 1; ofs < THREADS; ofs <<= 1) {
        uint32_t v = 0;
        if ((int)threadIdx.x >= ofs) v = s_scanw[threadIdx.x - ofs];
        __syncthreads();
        s_scanw[threadIdx.x] += v;
        __syncthreads();
      }
      {
        uint32_t w0 = s_scanw[threadIdx.x] - dw;
#pragma unroll
        for (int j = 0; j < PER; ++j) {
          const uint32_t i = threadIdx.x * PER + j;
          s_wcum[i] = w0;
          w0 += (s_len[i] + 31u) >> 5;
        }
        if (threadIdx.x == THREADS - 1) s_wcum[TILE] = s_scanw[THREADS - 1];
      }
      __syncthreads();
      const uint32_t words = s_wcum[TILE];
      uint32_t lo = 0;
      for (uint32_t w = threadIdx.x; w < words; w += THREADS) {
        while (s_wcum[lo + 1] <= w) ++lo;  // w only grows: linear advance
        const uint32_t rl = s_len[lo];
        const uint32_t b0 = (w - s_wcum[lo]) << 5;
        const uint32_t nb = min(32u, rl - b0);
        const unsigned long long p1 = s_qpos[lo], p2 = s_tpos[lo];
        auto eq_at = [&](uint32_t i) -> bool {
          return seqrush_query_base(q, (uint32_t)qlen, (uint32_t)(p1 + i), rc) == t[p2 + i];
        };
        uint32_t mask = 0;
        for (uint32_t i = 0; i < nb; ++i) mask |= (eq_at(b0 + i) ? 1u : 0u) << i;
        if (!mask) continue;
        uint32_t take = mask;
        if (k > 1) {
          // stretches of equal bases inside the word; the first / last may continue in the neighbours
          take = 0;
          uint32_t m = mask;
          while (m) {
            const uint32_t s0 = (uint32_t)__ffs((int)m) - 1u;
            uint32_t e0 = s0;  // exclusive end
            while (e0 < nb && ((mask >> e0) & 1u)) ++e0;
            unsigned long long ln = e0 - s0;
            if (ln < k && s0 == 0) {
              uint32_t i = b0;
              while (ln < k && i > 0 && eq_at(i - 1)) {
                --i;
                ++ln;
              }
            }
            if (ln < k && e0 == nb) {
              uint32_t i = b0 + nb;
              while (ln < k && i < rl && eq_at(i)) {
                ++i;
                ++ln;
              }
            }
            const uint32_t bits = (e0 - s0 >= 32u) ? 0xffffffffu : (((1u << (e0 - s0)) - 1u) << s0);
            if (ln >= k) take |= bits;
            m &= ~bits;
          }
        }
        while (take) {
          const uint32_t i = b0 + (uint32_t)__ffs((int)take) - 1u;
          take &= take - 1u;
          const unsigned long long x = rc ? 2ull * (qoff + qlen - 1ull - (p1 + i)) + 1ull : 2ull * (qoff + p1 + i);
          const unsigned long long y = 2ull * (toff + p2 + i);
          ++ops;
          if (uf_unite(a.parent, (uint32_t)x, (uint32_t)y)) ++links;
        }
      }
      __syncthreads();
      if (threadIdx.x == 0) {
        s_carry[0] += tq;
        s_carry[1] += tt;
      }
      __syncthreads();
    }
  }
  ops = __reduce_add_sync(0xffffffffu, (unsigned)(ops & 0xffffffffu)) +
        ((unsigned long long)__reduce_add_sync(0xffffffffu, (unsigned)(ops >> 32)) << 32);
  links = __reduce_add_sync(0xffffffffu, (unsigned)(links & 0xffffffffu)) +
          ((unsigned long long)__reduce_add_sync(0xffffffffu, (unsigned)(links >> 32)) << 32);
  if ((threadIdx.x & 31) == 0) {
    if (ops) atomicAdd(&a.counters[0], ops);
    if (links) atomicAdd(&a.counters[1], links);
  }
}

}  // namespace srb
