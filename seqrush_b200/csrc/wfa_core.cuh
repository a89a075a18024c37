// wfa_core.cuh -- gap-affine-2p bidirectional wavefront alignment (BiWFA), one
// sequence pair per CTA, written for sm_100a.
//
// This is a from-scratch CTA-cooperative formulation of the algorithm the
// reference runs on the CPU through WFA2-lib (cited per function below as
// WFA/<file>:<lines> = /root/reference/WFA2-lib/wavefront/<file>). Bit-exact
// CIGAR parity needs the same *decisions* as the reference (wavefront limits,
// trimming, breakpoint search order, backtrace priority), so those are kept;
// everything about how the work is laid out (diagonals across threads, fused
// compute+extend pass, packed warp reductions for the trims, parallel overlap
// search, run-length CIGAR emitted on device) is B200-first.
//
// The file is also compilable by a plain host C++ compiler with a 1-thread
// "CTA" (WFA_HOST_EMULATION). That build exists ONLY for the CPU-side logic
// tests in tests/ (there is no GPU in the development container); the product
// library never contains it.
#pragma once
#include <stdint.h>

#if defined(__CUDACC__) && !defined(WFA_HOST_EMULATION)
#define WFA_DEVICE 1
#define WFA_FN __device__ __forceinline__
#define WFA_FN_NOINLINE __device__ __noinline__
#else
#define WFA_DEVICE 0
#define WFA_FN inline
#define WFA_FN_NOINLINE inline
#include <string.h>
#endif

// SRGPU_CHECKED build (development only, scripts/build_checked.sh): bounds assertions on every workspace this file
// indexes -- ring rows, staging buffers, base-case store, run buffers -- that trap with a message. It stands in for
// compute-sanitizer's memcheck, which is closed on the GPU pool this was developed on (profiles/r02h_*).
#if defined(SRGPU_CHECKED) && defined(__CUDACC__) && !defined(WFA_HOST_EMULATION)
#include <cstdio>
#define SR_ASSERT(cond, what)                                                                                   \
  do {                                                                                                          \
    if (!(cond)) {                                                                                              \
      printf("SRGPU_CHECKED: %s violated at %s:%d (CTA %d thread %d)\n", what, __FILE__, __LINE__, (int)blockIdx.x, \
             (int)threadIdx.x);                                                                                 \
      __trap();                                                                                                 \
    }                                                                                                           \
  } while (0)
#else
#define SR_ASSERT(cond, what) ((void)0)
#endif

namespace srb {

static constexpr int32_t OFFSET_NULL = INT32_MIN / 2;  // WFA/wavefront_offset.h:44
static constexpr int MAX_SCOPE = 64;                    // ring slots available (>= max_score_scope)
static constexpr int STACK_MAX = 96;                    // explicit recursion stack (frames)

// component ids follow WFA2-lib/alignment/affine2p_penalties.h:51-57
enum : int { C_M = 0, C_I1 = 1, C_I2 = 2, C_D1 = 3, C_D2 = 4 };

// run-length CIGAR element: op in the low 2 bits, length above.
enum : uint32_t { OP_M = 0, OP_X = 1, OP_I = 2, OP_D = 3 };
WFA_FN uint32_t make_run(uint32_t op, uint32_t len) { return (len << 2) | op; }

// status codes surfaced through the C ABI
enum : int {
  ST_OK = 0,
  ST_UNATTAINABLE = 1,      // reference: WF_STATUS_UNATTAINABLE -> allwave returns empty CIGAR, score i32::MAX
  ST_WS_DIAG_OVERFLOW = 2,  // workspace too narrow: host retries with a wider ring
  ST_WS_BASE_OVERFLOW = 3,  // base-case score above full-memory capacity
  ST_WS_RUNS_OVERFLOW = 4,  // run buffer overflow
  ST_STACK_OVERFLOW = 5,
  // 6 = ST_ARENA_OVERFLOW (kernels.cuh)
  ST_RUNAWAY = 7,           // a breakpoint search ran past its score bound: a limit of this library, not a reference outcome
};

struct Penalties {
  int x, o1, e1, o2, e2;  // the penalties WFA2-lib computes with (after Eizenga's adjustment when match < 0)
  int scope;  // max_score_scope, WFA/wavefront_components.c:101-113
  int match;  // the caller's match score (<= 0): only the reported score depends on it (WFA/wavefront_compute.c:108-120)
};

// absent wavefront (NULL pointer in the reference) == (1,-1) like its null wavefront
// (WFA/wavefront.c:110-117); allocated but trimmed to nothing == (2,-2). get() maps both
// to (1,-1), which is what wavefront_compute_get_*wavefront (WFA/wavefront_compute.c:251-295) does.
struct WfMeta {
  int lo, hi;
};

struct Breakpoint {  // WFA/wavefront_bialign.h wf_bialign_breakpoint_t
  int score, score_forward, score_reverse;
  int k_forward, k_reverse;
  int offset_forward, offset_reverse;
  int component;
};

struct Frame {
  int pb, pe, tb, te;  // pattern [pb,pe), text [tb,te)
  int cbegin, cend;
  int score_remaining;
  int level;
};

// Sequence view: packed symbols, 1 << lbits bits each (lbits 1/2/3 = 2/4/8 bits per base), base i
// at bit (i << lbits) of a little-endian uint32 word array. The packing is an injective recoding
// of the bytes, so symbol equality == byte equality (the reference compares raw bytes,
// WFA/wavefront_extend_kernels.c:68-92). Arrays carry >= 1 guard word on both sides; they live
// in shared memory (staged per pair by a bulk copy) or, for oversized pairs, in global memory.
struct SeqView {
  const uint32_t* p;  // pattern (query, possibly reverse-complemented)
  const uint32_t* t;  // text (target)
  int lbits;
};

// ---------------------------------------------------------------------------
// Wavefront storage
// ---------------------------------------------------------------------------
// Modular ring (breakpoint search, score-only; WFA/wavefront_bialigner.c:54-55):
//   row(score,c) = base + ((score % scope)*5 + c) * W, indexed by k + W/2
// Full store (base case, WFA/wavefront_bialigner.c:67-68):
//   row(score,c) = base + c*cap*cap + score*score, indexed by k + score
//   (|k| <= score because every source wavefront is at least one score lower)
struct WfStore {
  int32_t* base;
  WfMeta* meta;  // [slots*5]
  int* ub;       // [slots*5] upper bound of the max antidiagonal (2h-k) over the row (modular only)
  int modular;
  int scope;
  int W;    // modular: diagonals per row
  int cap;  // full: max score + 1
  int cur_s, cur_slot;  // modular: slot of score cur_s (avoids an integer modulo per access)
  WFA_FN void set_cur(int s) {
    if (!modular) return;
    if (s == cur_s + 1) {
      cur_slot = (cur_slot + 1 == scope) ? 0 : cur_slot + 1;
    } else if (s != cur_s) {
      cur_slot = s % scope;
    }
    cur_s = s;
  }
  // valid for cur_s - scope < s <= cur_s
  WFA_FN int slot(int s) const {
    if (!modular) return s;
    const int t = cur_slot - (cur_s - s);
    return t < 0 ? t + scope : t;
  }
#ifdef SRGPU_EXP_SHALLOW  // measurement build only: gap components in a 3-slot ring (valid in the far phase only)
  WFA_FN int off_of(int s, int c) const { return ((c == C_M ? slot(s) : (s % 3)) * 5 + c) * W + (W >> 1); }
#else
  WFA_FN int off_of(int s, int c) const { return (slot(s) * 5 + c) * W + (W >> 1); }  // modular rows only
#endif
  WFA_FN int32_t* row_of_slot(int sl, int c) const { return base + (uint32_t)((sl * 5 + c) * W + (W >> 1)); }
  WFA_FN int32_t* row(int s, int c) const {
    if (modular) {
      SR_ASSERT(s > cur_s - scope && s <= cur_s && slot(s) >= 0 && slot(s) < scope, "ring slot of a live score");
      return row_of_slot(slot(s), c);
    }
    SR_ASSERT(s >= 0 && s < cap && c >= 0 && c < 5, "base-case store row");
    return base + (size_t)c * (size_t)cap * (size_t)cap + (size_t)s * (size_t)s + s;
  }
  WFA_FN WfMeta get(int s, int c) const {
    WfMeta m = meta[slot(s < 0 ? cur_s : s) * 5 + c];
    if (s < 0 || m.lo > m.hi) {
      m.lo = 1;
      m.hi = -1;
    }
    return m;
  }
};

// ---------------------------------------------------------------------------
// CTA context. All control state lives in shared memory (device) / a plain
// struct (host emulation).
// ---------------------------------------------------------------------------
struct RedSlots {
  int mn[5];
  int mx[5];
  int max_ak;
  int pad;
};

struct CtaShared {
  // per-score-slot tables and the recursion stack live in a CtaTables<SLOTS, STACK> of the kernel's choosing
  // (the one-warp kernel of split mode keeps them small to fit more CTAs per SM); bound by cta_bind()
  WfMeta* meta_f;  // [slots * 5]
  WfMeta* meta_r;
  int* ub_f;       // [slots * 5]
  int* ub_r;
  int* ovl;        // [slots * 5] overlap search results (min k_0 per task)
  Frame* stack;    // [stack_max]
  int stack_max;
  int slots_max;
  RedSlots red[2];
  alignas(16) int ds[2][48];  // per direction: limits, interior bounds, row offsets, flags of the running score step
  int chunk_ctr;           // next unclaimed chunk of the running score step (dynamic balance)
  int ovl_any;             // overlap search: some task found a diagonal (else the sequential replay is skipped)
  int end_val[2];          // per direction: offset of the end cell (diagonal tlen-plen, end component)
  int sp;
  Breakpoint bp;
  int status;
  int bcast[8];
  unsigned long long bcast64[2];
  // run-length CIGAR assembly
  uint32_t n_runs;
  uint32_t bt_n;
  // work counters (roofline accounting)
  unsigned long long cells;   // BiWFA cells: one diagonal at one score, all five components
  unsigned long long ocells;  // orientation (unit-cost) cells
#ifdef WFA_PROFILE
  unsigned long long prof[20];  // cycle counters of thread 0 (development only)
  unsigned long long prof2[20]; // phase breakdown: [0..4] find_breakpoint cycles by level (4 = deeper), [5] base case, [6] base calls,
                                // [7] backtrace, [8] overlap, [9] close phase, [10] close steps, [11..15] cells by level
#endif
};

template <int SLOTS, int STACK>
struct CtaTables {
  WfMeta meta_f[SLOTS * 5];
  WfMeta meta_r[SLOTS * 5];
  int ub_f[SLOTS * 5];
  int ub_r[SLOTS * 5];
  int ovl[SLOTS * 5];
  Frame stack[STACK];
};

// one thread; the caller synchronises before the tables are used
template <int SLOTS, int STACK>
WFA_FN void cta_bind(CtaShared* sh, CtaTables<SLOTS, STACK>* t) {
  sh->meta_f = t->meta_f;
  sh->meta_r = t->meta_r;
  sh->ub_f = t->ub_f;
  sh->ub_r = t->ub_r;
  sh->ovl = t->ovl;
  sh->stack = t->stack;
  sh->stack_max = STACK;
  sh->slots_max = SLOTS;
}

struct Ctx {
  CtaShared* sh;
  int tid, nt;
  int parity;
  int32_t* stage;  // device: shared-memory staging of the source-row slices of a chunk (STAGE_INTS per warp)
};

// Staged source rows of one chunk: seven rows (o1, o2, i1, i2, d1, d2, mis in the order of DS_O1..DS_MIS),
// each the 128 diagonals of the chunk plus four on either side (the k-1 / k+1 neighbours of the
// chunk's edge cells), as 16-byte segments.
static constexpr int STAGE_ROW = 32 * 4 + 8;
static constexpr int STAGE_INTS = 7 * STAGE_ROW;

#if WFA_DEVICE
WFA_FN void cta_sync() { __syncthreads(); }
WFA_FN int warp_min(int v) { return __reduce_min_sync(0xffffffffu, v); }
WFA_FN int warp_max(int v) { return __reduce_max_sync(0xffffffffu, v); }
WFA_FN void sh_atomic_min(int* a, int v) { atomicMin(a, v); }
WFA_FN void sh_atomic_max(int* a, int v) { atomicMax(a, v); }
WFA_FN int lane_id() { return threadIdx.x & 31; }
#else
WFA_FN void cta_sync() {}
WFA_FN int warp_min(int v) { return v; }
WFA_FN int warp_max(int v) { return v; }
WFA_FN void sh_atomic_min(int* a, int v) {
  if (v < *a) *a = v;
}
WFA_FN void sh_atomic_max(int* a, int v) {
  if (v > *a) *a = v;
}
WFA_FN int lane_id() { return 0; }
#endif

WFA_FN int imin(int a, int b) { return a < b ? a : b; }
WFA_FN int imax(int a, int b) { return a > b ? a : b; }

// ---------------------------------------------------------------------------
// Longest common prefix (the WFA "extend" inner kernel,
// WFA/wavefront_extend_kernels.c:68-92). Explicit length bound replaces the
// reference's '!' / '?' sentinels (WFA/wavefront_sequences.c:37-39,274-310).
// Forward: P[v+i] vs T[h+i]. Reverse aligner works on reversed (not
// complemented) copies (WFA/wavefront_sequences.c:83-101): P[-1-v-i] vs T[-1-h-i]
// relative to the end pointers, i.e. a common *suffix* walk on the originals.
// ---------------------------------------------------------------------------
WFA_FN uint32_t funnel_r(uint32_t lo, uint32_t hi, uint32_t sh) {
#if WFA_DEVICE
  return __funnelshift_r(lo, hi, sh);
#else
  sh &= 31;
  return sh ? ((lo >> sh) | (hi << (32 - sh))) : lo;
#endif
}
WFA_FN int ctz32(uint32_t x) {
#if WFA_DEVICE
  return __ffs((int)x) - 1;
#else
  return __builtin_ctz(x);
#endif
}
// zero-safe variants: 32 for x == 0 (a packed window that matches completely)
WFA_FN int clz32z(uint32_t x) {
#if WFA_DEVICE
  return __clz((int)x);
#else
  return x ? __builtin_clz(x) : 32;
#endif
}
WFA_FN int ctz32z(uint32_t x) {
#if WFA_DEVICE
  return __clz((int)__brev(x));
#else
  return x ? __builtin_ctz(x) : 32;
#endif
}
WFA_FN int clz32(uint32_t x) {
#if WFA_DEVICE
  return __clz((int)x);
#else
  return __builtin_clz(x);
#endif
}
// 32 bits of packed symbols starting at symbol index i (forward window)
WFA_FN uint32_t pk_window_fwd(const uint32_t* a, int i, int lbits) {
  const uint32_t bit = (uint32_t)i << lbits;
  const uint32_t* w = a + (bit >> 5);
  return funnel_r(w[0], w[1], bit);
}
// 32 bits of packed symbols ending with symbol index i in the top bits (backward window)
WFA_FN uint32_t pk_window_rev(const uint32_t* a, int i, int lbits) {
  const uint32_t bit = (uint32_t)(i + 1) << lbits;  // exclusive end
  const uint32_t* w = a + (bit >> 5);
  return funnel_r(w[-1], w[0], bit);
}
// common prefix of P[v..] and T[h..], at most maxlen symbols (XOR + count-trailing-zeros on packed words)
WFA_FN int lcp_fwd(const uint32_t* P, const uint32_t* T, int v, int h, int maxlen, int lbits) {
  int n = 0;
  const int per = 32 >> lbits;
  while (n < maxlen) {
    const uint32_t x = pk_window_fwd(P, v + n, lbits) ^ pk_window_fwd(T, h + n, lbits);
    if (x) {
      n += ctz32(x) >> lbits;
      break;
    }
    n += per;
  }
  return n < maxlen ? n : maxlen;
}
// common suffix walk: P[v], P[v-1], ... against T[h], T[h-1], ...
WFA_FN int lcp_rev(const uint32_t* P, const uint32_t* T, int v, int h, int maxlen, int lbits) {
  int n = 0;
  const int per = 32 >> lbits;
  while (n < maxlen) {
    const uint32_t x = pk_window_rev(P, v - n, lbits) ^ pk_window_rev(T, h - n, lbits);
    if (x) {
      n += clz32(x) >> lbits;
      break;
    }
    n += per;
  }
  return n < maxlen ? n : maxlen;
}

// Direction-agnostic forms (rev = 0/1 at run time): one copy of the score-step code serves both
// aligners of a breakpoint search, which halves the hot instruction footprint per SM.
//   rev = 0: 32 bits of packed symbols starting at symbol i (low bits first)
//   rev = 1: 32 bits ending with symbol i in the top bits
WFA_FN uint32_t pk_window_dir(const uint32_t* a, int i, int lbits, int rev) {
  const uint32_t bit = (uint32_t)(i + rev) << lbits;
  const uint32_t* w = a + (bit >> 5) - rev;
  return funnel_r(w[0], w[1], bit);
}
// matching symbols at the near end of an XORed window pair (32 >> lbits when it is all zero)
WFA_FN int pk_match_len(uint32_t x, int lbits, int rev) {
#if WFA_DEVICE
  return __clz((int)(rev ? x : __brev(x))) >> lbits;
#else
  return (rev ? clz32z(x) : ctz32z(x)) >> lbits;
#endif
}
WFA_FN int lcp_dir(const uint32_t* P, const uint32_t* T, int v, int h, int maxlen, int lbits, int rev) {
  int n = 0;
  const int per = 32 >> lbits, sgn = rev ? -1 : 1;
  while (n < maxlen) {
    const uint32_t x = pk_window_dir(P, v + sgn * n, lbits, rev) ^ pk_window_dir(T, h + sgn * n, lbits, rev);
    if (x) {
      n += pk_match_len(x, lbits, rev);
      break;
    }
    n += per;
  }
  return n < maxlen ? n : maxlen;
}

// One aligner direction over the current sub-problem.
struct Dir {
  WfStore st;
  const uint32_t* P;  // packed pattern / text (whole sequences)
  const uint32_t* T;
  int p0, t0;  // forward: first symbol of the sub-problem (pb, tb); reverse: last symbol (pe-1, te-1)
  int lbits;
  int plen, tlen;
  int reverse;
  int cbegin, cend;
};

// checked read without a branch: the load is always issued (diagonal 0 of every row lies inside the
// allocation), so the reads of a fringe chunk overlap instead of paying one memory latency each
WFA_FN int32_t rd(const int32_t* row, WfMeta m, int k) {
  const bool ok = (k >= m.lo) & (k <= m.hi);
  const int32_t v = row[ok ? k : 0];
  return ok ? v : OFFSET_NULL;
}

// ---------------------------------------------------------------------------
// Initial wavefront (WFA/wavefront_aligner.c:252-383, end-to-end form only).
// Executed by thread 0; caller syncs.
// ---------------------------------------------------------------------------
static constexpr int UB_NONE = INT32_MIN / 4;  // "row cannot overlap anything"

WFA_FN void wf_init(Dir& d) {
  d.st.set_cur(0);
  for (int c = 0; c < 5; ++c) {
    WfMeta m;
    m.lo = 1;
    m.hi = -1;
    if (c == d.cbegin) {
      m.lo = 0;
      m.hi = 0;
      d.st.row(0, c)[0] = 0;
    }
    d.st.meta[d.st.slot(0) * 5 + c] = m;
    if (d.st.ub) d.st.ub[d.st.slot(0) * 5 + c] = (c == d.cbegin) ? 0 : UB_NONE;
  }
}

struct StepOut {
  int null_step;
  int end_reached;
  int max_ak;
  int overflow;
};

// ---------------------------------------------------------------------------
// Extend-only pass for score 0 (WFA/wavefront_extend.c:173-211 on the initial
// wavefront) + termination test (WFA/wavefront_termination.c:37-113).
// ---------------------------------------------------------------------------
WFA_FN bool wf_end_reached(const Dir& d, int score) {
  const int ak = d.tlen - d.plen;
  const WfMeta m = d.st.get(score, d.cend);
  if (m.lo > ak || ak > m.hi) return false;
  return d.st.row(score, d.cend)[ak] >= d.tlen;
}

WFA_FN int wf_extend_cell(const Dir& d, int k, int32_t off) {
  const int v = off - k, h = off;
  const int room = imin(d.plen - v, d.tlen - h);
  const int sgn = d.reverse ? -1 : 1;
  return off + lcp_dir(d.P, d.T, d.p0 + sgn * v, d.t0 + sgn * h, room, d.lbits, d.reverse);
}

// score-0 extension: a single cell at most (k=0), done by thread 0. Every thread runs set_cur so
// that the per-thread slot arithmetic stays in step.
WFA_FN StepOut wf_extend0(Ctx& cx, Dir& d) {
  StepOut o;
  o.null_step = 0;
  o.end_reached = 0;
  o.max_ak = 0;
  o.overflow = 0;
  d.st.set_cur(0);
  cta_sync();
  if (cx.tid == 0) {
    const WfMeta m = d.st.get(0, C_M);
    int ak = 0;
    if (m.lo <= m.hi) {
      int32_t* row = d.st.row(0, C_M);
      const int32_t e = wf_extend_cell(d, 0, row[0]);
      row[0] = e;
      ak = 2 * e;
      if (d.st.ub) d.st.ub[d.st.slot(0) * 5 + C_M] = ak;
    }
    cx.sh->bcast[0] = ak;
  }
  cta_sync();
  o.max_ak = cx.sh->bcast[0];
  o.end_reached = wf_end_reached(d, 0) ? 1 : 0;
  return o;
}

#if WFA_DEVICE
static constexpr int LANES = 32;
WFA_FN uint32_t warp_ballot(bool p) { return __ballot_sync(0xffffffffu, p); }
#else
static constexpr int LANES = 1;
WFA_FN uint32_t warp_ballot(bool p) { return p ? 1u : 0u; }
#endif

// warp-uniform first/last valid diagonal of one component, fed with one ballot per 32-diagonal chunk
struct Span {
  int mn, mx;
  WFA_FN void init() {
    mn = INT32_MAX;
    mx = INT32_MIN;
  }
  WFA_FN void add(uint32_t b, int kb) {
    if (b) {
      if (mn == INT32_MAX) mn = kb + ctz32(b);
      mx = kb + 31 - clz32(b);
    }
  }
};

WFA_FN bool in_bounds(int32_t h, int k, uint32_t tlen, uint32_t plen) {
  return (uint32_t)h <= tlen && (uint32_t)(h - k) <= plen;
}

// ---------------------------------------------------------------------------
// Tail of a score step, shared by the scalar and the vector step: merge the per-warp trims
// (lanes 0..10 of every warp issue one shared-memory atomic each), publish the five metas and
// antidiagonal bounds (one lane per component), reset the other reduction buffer, end test.
// ---------------------------------------------------------------------------
struct StepCtl {
  int s, slot5;
  int s_o1, s_e1, s_o2, s_e2;
  bool n_o1, n_o2, n_i1, n_d1, n_i2, n_d2;
  bool a_i1, a_i2, a_d1, a_d2;
};

WFA_FN void publish_comp(const WfStore& st, const RedSlots* red, const StepCtl& t, int c) {
  const bool al = c == C_M ? true : c == C_I1 ? t.a_i1 : c == C_I2 ? t.a_i2 : c == C_D1 ? t.a_d1 : t.a_d2;
  WfMeta m;
  m.lo = 1;
  m.hi = -1;
  bool nonempty = false;
  if (al) {
    if (red->mn[c] <= red->mx[c]) {
      m.lo = red->mn[c];
      m.hi = red->mx[c];
      nonempty = true;
    } else {
      m.lo = 2;
      m.hi = -2;
    }
  }
  st.meta[t.slot5 + c] = m;
  if (st.ub) {
    // exact for M (after extension); for the gap components ak(I/D[s][k]) = ak(source cell) + 1
    int u = UB_NONE;
    if (nonempty) {
      if (c == C_M) {
        u = red->max_ak;
      } else {
        const bool one = (c == C_I1 || c == C_D1);
        const bool n_m = one ? t.n_o1 : t.n_o2;
        const bool n_s = c == C_I1 ? t.n_i1 : c == C_I2 ? t.n_i2 : c == C_D1 ? t.n_d1 : t.n_d2;
        const int ua = n_m ? UB_NONE : st.ub[st.slot(one ? t.s_o1 : t.s_o2) * 5 + C_M];
        const int ub = n_s ? UB_NONE : st.ub[st.slot(one ? t.s_e1 : t.s_e2) * 5 + c];
        u = imax(ua, ub) + 1;
      }
    }
    st.ub[t.slot5 + c] = u;
  }
}

WFA_FN void step_finish(Ctx& cx, const Dir& d, const StepCtl& t, const int* mn, const int* mx, int max_ak, StepOut& o,
                        bool end_stashed) {
  const WfStore& st = d.st;
  RedSlots* red = &cx.sh->red[cx.parity];
  max_ak = warp_max(max_ak);
#if WFA_DEVICE
  {
    const int lane = lane_id();
    int v = max_ak;
    v = lane == 0 ? mn[0] : lane == 1 ? mn[1] : lane == 2 ? mn[2] : lane == 3 ? mn[3] : lane == 4 ? mn[4] : v;
    v = lane == 5 ? mx[0] : lane == 6 ? mx[1] : lane == 7 ? mx[2] : lane == 8 ? mx[3] : lane == 9 ? mx[4] : v;
    int* slot = &red->mn[0] + lane;  // RedSlots = mn[5], mx[5], max_ak: eleven consecutive ints
    if (lane < 5) {
      if (v != INT32_MAX) atomicMin(slot, v);
    } else if (lane < 10) {
      if (v != INT32_MIN) atomicMax(slot, v);
    } else if (lane == 10) {
      if (v > 0) atomicMax(slot, v);
    }
  }
#else
  if (max_ak > 0) sh_atomic_max(&red->max_ak, max_ak);
  for (int c = 0; c < 5; ++c) {
    if (mn[c] != INT32_MAX) {
      sh_atomic_min(&red->mn[c], mn[c]);
      sh_atomic_max(&red->mx[c], mx[c]);
    }
  }
#endif
#ifdef WFA_PROFILE
  const long long tb0 = clock64();
#endif
  cta_sync();
#ifdef WFA_PROFILE
  if (cx.tid == 0) cx.sh->prof[5] += clock64() - tb0;  // wait at the closing barrier
#endif
  o.max_ak = red->max_ak;
#if WFA_DEVICE
  if (cx.tid < 5) {
    publish_comp(st, red, t, cx.tid);
  } else if (cx.tid < 16) {
    int* nx = &cx.sh->red[cx.parity ^ 1].mn[0];
    const int i = cx.tid - 5;
    nx[i] = i < 5 ? INT32_MAX : i < 10 ? INT32_MIN : 0;
  } else if (cx.tid == 16) {
    cx.sh->chunk_ctr = 0;  // every warp is past its last claim (barrier above)
  }
#else
  cx.sh->chunk_ctr = 0;
  for (int c = 0; c < 5; ++c) publish_comp(st, red, t, c);
  {
    RedSlots* nx = &cx.sh->red[cx.parity ^ 1];
    for (int c = 0; c < 5; ++c) {
      nx->mn[c] = INT32_MAX;
      nx->mx[c] = INT32_MIN;
    }
    nx->max_ak = 0;
  }
#endif
  // termination (WFA/wavefront_termination.c:37-113): read directly from the reduction slots / rows
  {
    const int ak = d.tlen - d.plen;
    const int ce = d.cend;
    const bool al = ce == C_M ? true : ce == C_I1 ? t.a_i1 : ce == C_I2 ? t.a_i2 : ce == C_D1 ? t.a_d1 : t.a_d2;
    const bool present = al && red->mn[ce] <= red->mx[ce];
    if (present && red->mn[ce] <= ak && ak <= red->mx[ce]) {
      // the vector step leaves the end cell in shared memory (a global read here would sit on the
      // critical path of every step)
      const int32_t ev = end_stashed ? cx.sh->end_val[0] : st.row(t.s, ce)[ak];
      if (ev >= d.tlen) o.end_reached = 1;
    }
  }
  cx.parity ^= 1;
}

// ---------------------------------------------------------------------------
// One score step: next wavefront (WFA/wavefront_compute_affine2p.c:45-106,
// 334-368; limits WFA/wavefront_compute.c:40-86; trimming :579-632) fused with
// the exact-match extension of the new M wavefront (WFA/wavefront_extend.c:
// 86-121,173-211) and the end test.
// Layout: each warp takes 32-diagonal chunks. Chunks that lie inside the
// interior of all seven source wavefronts read them without range checks;
// fringe chunks use checked reads. The per-component trims (first/last
// in-bounds diagonal) come from one ballot per chunk and component instead of
// per-cell min/max; the max antidiagonal is a per-lane running max.
// ---------------------------------------------------------------------------
WFA_FN StepOut wf_step(Ctx& cx, Dir& d, const Penalties& pen, int s) {
  StepOut o;
  o.null_step = 0;
  o.end_reached = 0;
  o.max_ak = 0;
  o.overflow = 0;
  d.st.set_cur(s);
  cta_sync();  // metas written by thread 0 in the previous step are visible; red[parity] is reset
  const WfStore& st = d.st;
  const int s_mis = s - pen.x, s_o1 = s - pen.o1 - pen.e1, s_e1 = s - pen.e1;
  const int s_o2 = s - pen.o2 - pen.e2, s_e2 = s - pen.e2;
  const WfMeta m_mis = st.get(s_mis, C_M), m_o1 = st.get(s_o1, C_M), m_o2 = st.get(s_o2, C_M);
  const WfMeta m_i1 = st.get(s_e1, C_I1), m_d1 = st.get(s_e1, C_D1);
  const WfMeta m_i2 = st.get(s_e2, C_I2), m_d2 = st.get(s_e2, C_D2);
  const bool n_mis = m_mis.lo > m_mis.hi, n_o1 = m_o1.lo > m_o1.hi, n_o2 = m_o2.lo > m_o2.hi;
  const bool n_i1 = m_i1.lo > m_i1.hi, n_d1 = m_d1.lo > m_d1.hi, n_i2 = m_i2.lo > m_i2.hi, n_d2 = m_d2.lo > m_d2.hi;
  const int slot5 = st.slot(s) * 5;
  if (n_mis && n_o1 && n_o2 && n_i1 && n_d1 && n_i2 && n_d2) {
    // null step (WFA/wavefront_compute_affine2p.c:343-353)
    if (cx.tid == 0) {
      for (int c = 0; c < 5; ++c) {
        st.meta[slot5 + c].lo = 1;
        st.meta[slot5 + c].hi = -1;
        if (st.ub) st.ub[slot5 + c] = UB_NONE;
      }
    }
    o.null_step = 1;
    return o;
  }
  // limits (WFA/wavefront_compute.c:40-86); null inputs carry lo=1,hi=-1 like the reference's null wavefront
  int lo = m_mis.lo, hi = m_mis.hi;
  lo = imin(lo, m_o1.lo - 1);
  hi = imax(hi, m_o1.hi + 1);
  lo = imin(lo, m_i1.lo + 1);
  hi = imax(hi, m_i1.hi + 1);
  lo = imin(lo, m_d1.lo - 1);
  hi = imax(hi, m_d1.hi - 1);
  lo = imin(lo, m_o2.lo - 1);
  hi = imax(hi, m_o2.hi + 1);
  lo = imin(lo, m_i2.lo + 1);
  hi = imax(hi, m_i2.hi + 1);
  lo = imin(lo, m_d2.lo - 1);
  hi = imax(hi, m_d2.hi - 1);
  if (st.modular) {
    if (lo <= -(st.W >> 1) || hi >= (st.W >> 1)) {
      o.overflow = 1;
      return o;
    }
  } else if (s >= st.cap) {
    o.overflow = 1;
    return o;
  }
  if (cx.tid == 0) cx.sh->cells += (unsigned long long)(hi - lo + 1);
  // output allocation (WFA/wavefront_compute.c:458-505)
  const bool a_i1 = !n_o1 || !n_i1, a_d1 = !n_o1 || !n_d1, a_i2 = !n_o2 || !n_i2, a_d2 = !n_o2 || !n_d2;
  // interior: diagonals where all twelve reads fall inside their source wavefronts
  int in_lo = 1, in_hi = 0;
  if (!(n_mis || n_o1 || n_o2 || n_i1 || n_d1 || n_i2 || n_d2)) {
    in_lo = imax(imax(m_mis.lo, imax(m_o1.lo, m_o2.lo) + 1), imax(imax(m_i1.lo, m_i2.lo) + 1, imax(m_d1.lo, m_d2.lo) - 1));
    in_hi = imin(imin(m_mis.hi, imin(m_o1.hi, m_o2.hi) - 1), imin(imin(m_i1.hi, m_i2.hi) + 1, imin(m_d1.hi, m_d2.hi) - 1));
  }
  const int32_t* r_mis = st.row(s_mis < 0 ? 0 : s_mis, C_M);
  const int32_t* r_o1 = st.row(s_o1 < 0 ? 0 : s_o1, C_M);
  const int32_t* r_o2 = st.row(s_o2 < 0 ? 0 : s_o2, C_M);
  const int32_t* r_i1 = st.row(s_e1 < 0 ? 0 : s_e1, C_I1);
  const int32_t* r_d1 = st.row(s_e1 < 0 ? 0 : s_e1, C_D1);
  const int32_t* r_i2 = st.row(s_e2 < 0 ? 0 : s_e2, C_I2);
  const int32_t* r_d2 = st.row(s_e2 < 0 ? 0 : s_e2, C_D2);
  int32_t* w_m = st.row(s, C_M);
  int32_t* w_i1 = st.row(s, C_I1);
  int32_t* w_i2 = st.row(s, C_I2);
  int32_t* w_d1 = st.row(s, C_D1);
  int32_t* w_d2 = st.row(s, C_D2);
  const uint32_t tlen = (uint32_t)d.tlen, plen = (uint32_t)d.plen;
  Span sp0, sp1, sp2, sp3, sp4;
  sp0.init();
  sp1.init();
  sp2.init();
  sp3.init();
  sp4.init();
  int max_ak = 0;
  const int lane = lane_id();
#if WFA_DEVICE
  const int warp = cx.tid >> 5, nwarps = cx.nt >> 5;
#else
  const int warp = 0, nwarps = 1;
#endif
  for (int kb = lo + warp * LANES; kb <= hi; kb += nwarps * LANES) {
    const int k = kb + lane;
    const bool act = k <= hi;
    int32_t o1m, o1p, o2m, o2p, i1, d1, i2, d2, mis;
    if (kb >= in_lo && kb + LANES - 1 <= in_hi) {
      o1m = r_o1[k - 1];
      o1p = r_o1[k + 1];
      o2m = r_o2[k - 1];
      o2p = r_o2[k + 1];
      i1 = r_i1[k - 1];
      d1 = r_d1[k + 1];
      i2 = r_i2[k - 1];
      d2 = r_d2[k + 1];
      mis = r_mis[k];
    } else {
      o1m = rd(r_o1, m_o1, k - 1);
      o1p = rd(r_o1, m_o1, k + 1);
      o2m = rd(r_o2, m_o2, k - 1);
      o2p = rd(r_o2, m_o2, k + 1);
      i1 = rd(r_i1, m_i1, k - 1);
      d1 = rd(r_d1, m_d1, k + 1);
      i2 = rd(r_i2, m_i2, k - 1);
      d2 = rd(r_d2, m_d2, k + 1);
      mis = rd(r_mis, m_mis, k);
    }
    const int32_t ins1 = imax(o1m, i1) + 1;
    const int32_t ins2 = imax(o2m, i2) + 1;
    const int32_t del1 = imax(o1p, d1);
    const int32_t del2 = imax(o2p, d2);
    int32_t mx = imax(imax(del1, del2), imax(mis + 1, imax(ins1, ins2)));
    const bool vm = act && in_bounds(mx, k, tlen, plen);
    if (!vm) mx = OFFSET_NULL;
    sp0.add(warp_ballot(vm), kb);
    if (a_i1) {
      if (act) w_i1[k] = ins1;
      sp1.add(warp_ballot(act && in_bounds(ins1, k, tlen, plen)), kb);
    }
    if (a_i2) {
      if (act) w_i2[k] = ins2;
      sp2.add(warp_ballot(act && in_bounds(ins2, k, tlen, plen)), kb);
    }
    if (a_d1) {
      if (act) w_d1[k] = del1;
      sp3.add(warp_ballot(act && in_bounds(del1, k, tlen, plen)), kb);
    }
    if (a_d2) {
      if (act) w_d2[k] = del2;
      sp4.add(warp_ballot(act && in_bounds(del2, k, tlen, plen)), kb);
    }
    if (vm) {
      mx = wf_extend_cell(d, k, mx);
      max_ak = imax(max_ak, 2 * mx - k);
    }
    if (act) w_m[k] = mx;
  }
  // merge the per-warp trims (WFA/wavefront_compute.c:579-611 does the same with two linear scans)
  StepCtl ctl;
  ctl.s = s;
  ctl.slot5 = slot5;
  ctl.s_o1 = s_o1;
  ctl.s_e1 = s_e1;
  ctl.s_o2 = s_o2;
  ctl.s_e2 = s_e2;
  ctl.n_o1 = n_o1;
  ctl.n_o2 = n_o2;
  ctl.n_i1 = n_i1;
  ctl.n_d1 = n_d1;
  ctl.n_i2 = n_i2;
  ctl.n_d2 = n_d2;
  ctl.a_i1 = a_i1;
  ctl.a_i2 = a_i2;
  ctl.a_d1 = a_d1;
  ctl.a_d2 = a_d2;
  const int mns[5] = {sp0.mn, sp1.mn, sp2.mn, sp3.mn, sp4.mn};
  const int mxs[5] = {sp0.mx, sp1.mx, sp2.mx, sp3.mx, sp4.mx};
  step_finish(cx, d, ctl, mns, mxs, max_ak, o, false);
  return o;
}

// ---------------------------------------------------------------------------
// Vectorised score step for the modular rings (the breakpoint searches, where
// almost all cells are computed). Same decisions as wf_step; the layout is four
// consecutive diagonals per lane, 128 per warp chunk, chunk starts aligned to 4
// so that every ring row is read and written with 128-bit accesses; the k-1 /
// k+1 neighbours come from the adjacent lanes by shuffle (one scalar load at
// each chunk edge). Interior chunks need no range checks. The in-bounds test of
// the four gap components uses M: M = max of the five, so M in bounds bounds
// them from above and only `value >= max(0,k)` is left to check; one vote per
// chunk then covers all trims when everything is in bounds (the common case).
// ---------------------------------------------------------------------------
#if WFA_DEVICE
WFA_FN int32_t shfl_up1(int32_t v) { return __shfl_up_sync(0xffffffffu, v, 1); }
WFA_FN int32_t shfl_down1(int32_t v) { return __shfl_down_sync(0xffffffffu, v, 1); }
WFA_FN bool warp_all(bool p) { return __all_sync(0xffffffffu, p); }
#else
WFA_FN int32_t shfl_up1(int32_t v) { return v; }
WFA_FN int32_t shfl_down1(int32_t v) { return v; }
WFA_FN bool warp_all(bool p) { return p; }
#endif

#ifdef WFA_PROFILE
#define PROF_T(v) const long long v = clock64()
#else
#define PROF_T(v)
#endif

struct I4 {
  int32_t x, y, z, w;
};
WFA_FN I4 ld4(const int32_t* p) {
#if WFA_DEVICE
  const int4 v = *reinterpret_cast<const int4*>(p);
  I4 r;
  r.x = v.x;
  r.y = v.y;
  r.z = v.z;
  r.w = v.w;
  return r;
#else
  I4 r;
  r.x = p[0];
  r.y = p[1];
  r.z = p[2];
  r.w = p[3];
  return r;
#endif
}
WFA_FN void st4(int32_t* p, int32_t a, int32_t b, int32_t c, int32_t e) {
#if WFA_DEVICE
  *reinterpret_cast<int4*>(p) = make_int4(a, b, c, e);
#else
  p[0] = a;
  p[1] = b;
  p[2] = c;
  p[3] = e;
#endif
}

#if WFA_DEVICE
// 16-byte asynchronous copy global -> shared, L2 only (ring rows have no reuse in L1)
WFA_FN void cp_async16(int32_t* dst_smem, const int32_t* src_gmem) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"((uint32_t)__cvta_generic_to_shared(dst_smem)), "l"(src_gmem)
               : "memory");
}
WFA_FN void cp_async4(int32_t* dst_smem, const int32_t* src_gmem) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"((uint32_t)__cvta_generic_to_shared(dst_smem)), "l"(src_gmem)
               : "memory");
}
WFA_FN void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
WFA_FN void cp_async_wait() {
  asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory");
}
#endif

// first/last valid diagonal of one component inside a chunk where lane l holds diagonals kb+4l+j
struct Span4 {  // LANE-local first/last valid diagonal of one component; reduced over the warp when flushed
  int mn, mx;
  WFA_FN void init() {
    mn = INT32_MAX;
    mx = INT32_MIN;
  }
  WFA_FN void add_full(int k0) {
    mn = imin(mn, k0);
    mx = imax(mx, k0 + 3);
  }
  WFA_FN void add(bool p0, bool p1, bool p2, bool p3, int k0) {
    const int f = p0 ? k0 : p1 ? k0 + 1 : p2 ? k0 + 2 : p3 ? k0 + 3 : INT32_MAX;
    const int l = p3 ? k0 + 3 : p2 ? k0 + 2 : p1 ? k0 + 1 : p0 ? k0 : INT32_MIN;
    mn = imin(mn, f);
    mx = imax(mx, l);
  }
  WFA_FN void reduce() {
    mn = warp_min(mn);
    mx = warp_max(mx);
  }
};

// ---------------------------------------------------------------------------
// Score step of the breakpoint search for ONE or BOTH directions at once ("dual step").
// The forward step to score sf and the reverse step to score sr do not depend on each other, so
// while the two searches are far apart they share one pass: their chunks go into one claim pool
// for all warps of the CTA (one copy of the chunk code: the direction is a run-time parameter read
// from shared memory), which halves the barriers, setups and publishes per score and evens out
// the work of the warps. Phases:
//   A barrier            rows and metas of earlier steps are visible
//   B setup              warp (dir % nwarps) derives limits, interior, row offsets of direction dir
//                        into sh->ds[dir], clears that direction's trim slots and the chunk counter
//   C barrier
//   D chunk loop         as before; the trims of a chunk go to shared memory right away
//   E barrier
//   F publish            five lanes per direction write the metas; every thread reads the results
// ---------------------------------------------------------------------------
// Words of ds[dir], grouped by the place that reads them so that a chunk fetches each group with ONE 128-bit
// shared-memory load (the words are volatile: a scalar read each was 25 loads per chunk).
enum {
  DS_O1 = 0, DS_O2, DS_I1, DS_I2,        // copy issue: row offsets of the seven sources ...
  DS_D1, DS_D2, DS_MIS, DS_HIP,          // ... and a copy of hi
  DS_LO, DS_HI, DS_INLO, DS_INHI,        // chunk: limits and interior
  DS_P0, DS_T0, DS_REV, DS_CEND,         // chunk: extension and end cell
  DS_WI1, DS_WI2, DS_WD1, DS_WD2,        // chunk: offsets of the gap rows written
  DS_LOAL, DS_NCH, DS_WM, DS_FLAGS,      // claim decode (two words), M row written, allocation flags
  DS_META,                               // .. +13: seven (lo, hi) pairs, fringe chunks only (16-byte aligned: 24)
  DS_NULL = DS_META + 14, DS_OVF, DS_WIDTH, DS_SLOT5, DS_S, DS_SO1, DS_SE1, DS_SO2, DS_SE2, DS_NFLAGS, DS_WORDS
};
static_assert(DS_WORDS <= 48 && (DS_META & 3) == 0, "ds[] layout");

WFA_FN I4 lds4v(const volatile int* p) {  // one volatile 128-bit shared-memory read (p 16-byte aligned)
#if WFA_DEVICE
  I4 r;
  asm volatile("ld.volatile.shared.v4.s32 {%0, %1, %2, %3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "r"((uint32_t)__cvta_generic_to_shared(const_cast<const int*>(p))));
  return r;
#else
  I4 r;
  r.x = p[0];
  r.y = p[1];
  r.z = p[2];
  r.w = p[3];
  return r;
#endif
}

// phase B for one direction; executed by one whole warp (lane 0 writes)
WFA_FN void wf_setup_dir(Ctx& cx, const Dir& d, const Penalties& pen, int s, int dir, int base_adj, int lane) {
  const WfStore& st = d.st;
  int* ds = cx.sh->ds[dir];
  const int s_mis = s - pen.x, s_o1 = s - pen.o1 - pen.e1, s_e1 = s - pen.e1;
  const int s_o2 = s - pen.o2 - pen.e2, s_e2 = s - pen.e2;
  const WfMeta m_mis = st.get(s_mis, C_M), m_o1 = st.get(s_o1, C_M), m_o2 = st.get(s_o2, C_M);
  const WfMeta m_i1 = st.get(s_e1, C_I1), m_d1 = st.get(s_e1, C_D1);
  const WfMeta m_i2 = st.get(s_e2, C_I2), m_d2 = st.get(s_e2, C_D2);
  const bool n_mis = m_mis.lo > m_mis.hi, n_o1 = m_o1.lo > m_o1.hi, n_o2 = m_o2.lo > m_o2.hi;
  const bool n_i1 = m_i1.lo > m_i1.hi, n_d1 = m_d1.lo > m_d1.hi, n_i2 = m_i2.lo > m_i2.hi, n_d2 = m_d2.lo > m_d2.hi;
  const int slot5 = st.slot(s) * 5;
  // trim slots of this direction start clean (lanes 0..10), end-cell stash too
#if WFA_DEVICE
  if (lane < 11) (&cx.sh->red[dir].mn[0])[lane] = lane < 5 ? INT32_MAX : lane < 10 ? INT32_MIN : 0;
#else
  for (int c = 0; c < 5; ++c) {
    cx.sh->red[dir].mn[c] = INT32_MAX;
    cx.sh->red[dir].mx[c] = INT32_MIN;
  }
  cx.sh->red[dir].max_ak = 0;
#endif
  if (lane != 0) return;
  ds[DS_NULL] = 0;
  ds[DS_OVF] = 0;
  ds[DS_NCH] = 0;
  ds[DS_WIDTH] = 0;
  ds[DS_S] = s;
  ds[DS_SLOT5] = slot5;
  if (n_mis && n_o1 && n_o2 && n_i1 && n_d1 && n_i2 && n_d2) {
    // null step (WFA/wavefront_compute_affine2p.c:343-353)
    for (int c = 0; c < 5; ++c) {
      st.meta[slot5 + c].lo = 1;
      st.meta[slot5 + c].hi = -1;
      st.ub[slot5 + c] = UB_NONE;
    }
    ds[DS_NULL] = 1;
    return;
  }
  // limits (WFA/wavefront_compute.c:40-86); null inputs carry lo=1,hi=-1 like the reference's null wavefront
  int lo = m_mis.lo, hi = m_mis.hi;
  lo = imin(lo, m_o1.lo - 1);
  hi = imax(hi, m_o1.hi + 1);
  lo = imin(lo, m_i1.lo + 1);
  hi = imax(hi, m_i1.hi + 1);
  lo = imin(lo, m_d1.lo - 1);
  hi = imax(hi, m_d1.hi - 1);
  lo = imin(lo, m_o2.lo - 1);
  hi = imax(hi, m_o2.hi + 1);
  lo = imin(lo, m_i2.lo + 1);
  hi = imax(hi, m_i2.hi + 1);
  lo = imin(lo, m_d2.lo - 1);
  hi = imax(hi, m_d2.hi - 1);
  // keep 8 diagonals of slack on both sides: aligned 4-wide accesses may touch [lo-3, hi+3]
  if (lo <= -(st.W >> 1) + 8 || hi >= (st.W >> 1) - 8) {
    ds[DS_OVF] = 1;
    return;
  }
  const bool a_i1 = !n_o1 || !n_i1, a_d1 = !n_o1 || !n_d1, a_i2 = !n_o2 || !n_i2, a_d2 = !n_o2 || !n_d2;
  int in_lo = 1, in_hi = 0;
  if (!(n_mis || n_o1 || n_o2 || n_i1 || n_d1 || n_i2 || n_d2)) {
    in_lo = imax(imax(m_mis.lo, imax(m_o1.lo, m_o2.lo) + 1), imax(imax(m_i1.lo, m_i2.lo) + 1, imax(m_d1.lo, m_d2.lo) - 1));
    in_hi = imin(imin(m_mis.hi, imin(m_o1.hi, m_o2.hi) - 1), imin(imin(m_i1.hi, m_i2.hi) + 1, imin(m_d1.hi, m_d2.hi) - 1));
    in_lo = imax(in_lo, lo);
    in_hi = imin(in_hi, hi);
  }
  const int lo_al = lo & ~3;  // arithmetic: rounds towards -inf on two's complement
  ds[DS_O1] = base_adj + st.off_of(s_o1 < 0 ? 0 : s_o1, C_M);
  ds[DS_O2] = base_adj + st.off_of(s_o2 < 0 ? 0 : s_o2, C_M);
  ds[DS_I1] = base_adj + st.off_of(s_e1 < 0 ? 0 : s_e1, C_I1);
  ds[DS_I2] = base_adj + st.off_of(s_e2 < 0 ? 0 : s_e2, C_I2);
  ds[DS_D1] = base_adj + st.off_of(s_e1 < 0 ? 0 : s_e1, C_D1);
  ds[DS_D2] = base_adj + st.off_of(s_e2 < 0 ? 0 : s_e2, C_D2);
  ds[DS_MIS] = base_adj + st.off_of(s_mis < 0 ? 0 : s_mis, C_M);
  ds[DS_WM] = base_adj + st.off_of(s, C_M);
  ds[DS_WI1] = base_adj + st.off_of(s, C_I1);
  ds[DS_WI2] = base_adj + st.off_of(s, C_I2);
  ds[DS_WD1] = base_adj + st.off_of(s, C_D1);
  ds[DS_WD2] = base_adj + st.off_of(s, C_D2);
  ds[DS_INLO] = in_lo;
  ds[DS_INHI] = in_hi;
  ds[DS_LO] = lo;
  ds[DS_HI] = hi;
  ds[DS_HIP] = hi;
  ds[DS_LOAL] = lo_al;
  ds[DS_NCH] = (hi - lo_al) / (LANES * 4) + 1;
  ds[DS_WIDTH] = hi - lo + 1;
  ds[DS_FLAGS] = (a_i1 ? 1 : 0) | (a_i2 ? 2 : 0) | (a_d1 ? 4 : 0) | (a_d2 ? 8 : 0);
  ds[DS_NFLAGS] = (n_o1 ? 1 : 0) | (n_o2 ? 2 : 0) | (n_i1 ? 4 : 0) | (n_d1 ? 8 : 0) | (n_i2 ? 16 : 0) | (n_d2 ? 32 : 0);
  ds[DS_META + 0] = m_o1.lo;
  ds[DS_META + 1] = m_o1.hi;
  ds[DS_META + 2] = m_o2.lo;
  ds[DS_META + 3] = m_o2.hi;
  ds[DS_META + 4] = m_i1.lo;
  ds[DS_META + 5] = m_i1.hi;
  ds[DS_META + 6] = m_i2.lo;
  ds[DS_META + 7] = m_i2.hi;
  ds[DS_META + 8] = m_d1.lo;
  ds[DS_META + 9] = m_d1.hi;
  ds[DS_META + 10] = m_d2.lo;
  ds[DS_META + 11] = m_d2.hi;
  ds[DS_META + 12] = m_mis.lo;
  ds[DS_META + 13] = m_mis.hi;
  ds[DS_P0] = d.p0;
  ds[DS_T0] = d.t0;
  ds[DS_REV] = d.reverse;
  ds[DS_CEND] = d.cend;
  ds[DS_SO1] = s_o1;
  ds[DS_SE1] = s_e1;
  ds[DS_SO2] = s_o2;
  ds[DS_SE2] = s_e2;
}

struct StepOut2 {
  StepOut f, r;
};

WFA_FN StepOut2 wf_step_multi(Ctx& cx, Dir& df, Dir& dr, const Penalties& pen, int sf, int sr, bool run_f, bool run_r) {
  StepOut2 out;
  out.f.null_step = out.f.end_reached = out.f.max_ak = out.f.overflow = 0;
  out.r = out.f;
  if (run_f) df.st.set_cur(sf);
  if (run_r) dr.st.set_cur(sr);
  const int lane = lane_id();
#if WFA_DEVICE
  const int warp = cx.tid >> 5, nwarps = cx.nt >> 5;
#else
  const int warp = 0, nwarps = 1;
#endif
  constexpr int CH = LANES * 4;
  PROF_T(t0);
  cta_sync();  // A
  PROF_T(t1);
  // B
  // (one copy of the setup code: the direction is selected at run time, the hot code of a dual step
  // has to stay inside the 32 KB instruction cache)
#pragma unroll 1
  for (int dir = 0; dir < 2; ++dir) {
    if (!(dir ? run_r : run_f) || warp != (dir % nwarps)) continue;
    const Dir dd = dir ? dr : df;
    wf_setup_dir(cx, dd, pen, dir ? sr : sf, dir, dir ? (int)(dr.st.base - df.st.base) : 0, lane);
  }
  if (cx.tid == 0) {
    cx.sh->chunk_ctr = 0;
    if (!run_f) cx.sh->ds[0][DS_NCH] = 0;
    if (!run_r) cx.sh->ds[1][DS_NCH] = 0;
  }
  cta_sync();  // C
  PROF_T(t2);
  {
    const volatile int* s0 = cx.sh->ds[0];
    const volatile int* s1 = cx.sh->ds[1];
    if (run_f) {
      out.f.null_step = s0[DS_NULL];
      out.f.overflow = s0[DS_OVF];
    }
    if (run_r) {
      out.r.null_step = s1[DS_NULL];
      out.r.overflow = s1[DS_OVF];
    }
  }
  // an overflowing direction ends the search (the host re-runs the pair with wider rows); the other
  // direction's chunks are not worth computing then
  if (out.f.overflow || out.r.overflow) return out;
  const int nch0 = cx.sh->ds[0][DS_NCH];
  const int nch_all = nch0 + cx.sh->ds[1][DS_NCH];
  int32_t* ring = df.st.base;
#if WFA_DEVICE
  asm volatile("" : "+l"(ring));  // keep the ring base as one materialised 64-bit pointer (row address = one IMAD.WIDE)
  __builtin_assume(__isGlobal(ring));
#endif
  const Dir& d = df;  // fields shared by both directions: sequences, lengths, packing
  const uint32_t tlen = (uint32_t)d.tlen, plen = (uint32_t)d.plen;
  // trims are accumulated in registers over the chunks a warp takes from one direction and go to
  // that direction's reduction slots when the warp moves on (lanes 0..10 issue one atomic each)
  Span4 sp0, sp1, sp2, sp3, sp4;
  int max_ak = 0, acc_dir = -1;
  auto flush = [&](int fdir) {
    max_ak = warp_max(max_ak);
    sp0.reduce();
    sp1.reduce();
    sp2.reduce();
    sp3.reduce();
    sp4.reduce();
    if (lane == 0) {  // shared-memory atomics do not return a value: one lane issues all eleven back to back
      RedSlots* red = &cx.sh->red[fdir];
      if (max_ak > 0) sh_atomic_max(&red->max_ak, max_ak);
      if (sp0.mn != INT32_MAX) {
        sh_atomic_min(&red->mn[0], sp0.mn);
        sh_atomic_max(&red->mx[0], sp0.mx);
      }
      if (sp1.mn != INT32_MAX) {
        sh_atomic_min(&red->mn[1], sp1.mn);
        sh_atomic_max(&red->mx[1], sp1.mx);
      }
      if (sp2.mn != INT32_MAX) {
        sh_atomic_min(&red->mn[2], sp2.mn);
        sh_atomic_max(&red->mx[2], sp2.mx);
      }
      if (sp3.mn != INT32_MAX) {
        sh_atomic_min(&red->mn[3], sp3.mn);
        sh_atomic_max(&red->mx[3], sp3.mx);
      }
      if (sp4.mn != INT32_MAX) {
        sh_atomic_min(&red->mn[4], sp4.mn);
        sh_atomic_max(&red->mx[4], sp4.mx);
      }
    }
  };
  // Chunks of both directions are claimed from one shared counter; inside a direction the two
  // fringe chunks (masked reads, per-cell trims: ~1.5x an interior chunk) go first.
  // The seven source-row slices of a chunk are staged in shared memory by asynchronous 16-byte
  // copies (one buffer per warp): a warp moves the staged slices of its current chunk into
  // registers, then claims its NEXT chunk and starts that chunk's copies before it computes the
  // current one, so the global-memory latency of the row reads (38 % of all warp stall samples when
  // the rows were read straight into registers: profiles/r02b_*) is covered by one chunk of compute
  // instead of being waited for. The copies in flight cost no registers.
  int n_dir = -1, n_kb = 0;
#ifdef SRGPU_DYNAMIC_CLAIM
  auto claim_issue = [&]() -> int {  // lane 0 takes the next chunk index; the value is consumed by claim_decode
    int claim = 0;
#if WFA_DEVICE
    if (lane == 0) claim = atomicAdd(&cx.sh->chunk_ctr, 1);
#else
    claim = cx.sh->chunk_ctr++;
#endif
    return claim;
  };
#else
  // chunks are dealt round-robin to the warps: chunk costs are close to uniform, and the shared-memory atomic of a
  // dynamic claim (with its return value) was 6.7 % of the kernel's stall samples (profiles/r03h_*)
  int my_claim = warp;
  auto claim_issue = [&]() -> int {
    const int claim = my_claim;
    my_claim += nwarps;
    return claim;
  };
#endif
  auto claim_decode = [&](int claim) {
#if WFA_DEVICE && defined(SRGPU_DYNAMIC_CLAIM)
    claim = __shfl_sync(0xffffffffu, claim, 0);
#endif
    if (claim >= nch_all) {
      n_dir = -1;
      n_kb = 0;
      return;
    }
    n_dir = claim >= nch0 ? 1 : 0;
    const volatile int* sn = cx.sh->ds[n_dir];
    const int cl = claim - (n_dir ? nch0 : 0), nchunks = n_dir ? nch_all - nch0 : nch0;
    const int ci = cl == 0 ? 0 : (cl == 1 ? nchunks - 1 : cl - 1);
    n_kb = sn[DS_LOAL] + ci * CH;
  };
#if WFA_DEVICE
  int32_t* const wstage = cx.stage + warp * STAGE_INTS;
  // copies of the slices [kb-4, kb+132) of the seven source rows of chunk (pdir, pkb) into the warp's buffer.
  // Segments that start beyond hi+4 are not read (they would leave the row allocation, which keeps
  // 8 diagonals of slack); whatever an unread segment holds is outside every source wavefront and is
  // masked (fringe chunk) or belongs to lanes that neither compute a valid cell nor store.
  auto prefetch = [&](int pdir, int pkb) {
    if (pdir >= 0) {
      const volatile int* sp = cx.sh->ds[pdir];
      const I4 pa = lds4v(sp + DS_O1), pb = lds4v(sp + DS_D1);
      const int offs[7] = {pa.x, pa.y, pa.z, pa.w, pb.x, pb.y, pb.z};
      const int phi = pb.w;
      const int ks = pkb + 4 * lane;  // this lane's own four diagonals
      const bool ok0 = ks <= phi + 4;
#pragma unroll
      for (int r = 0; r < 7; ++r) {
        const int off = offs[r];
        SR_ASSERT(!ok0 || (off + ks >= 0 && off + ks + 3 < 2 * df.st.scope * 5 * df.st.W), "row slice read inside the CTA's rings");
        SR_ASSERT(((off + ks) & 3) == 0, "16-byte aligned row slice");
        if (ok0) cp_async16(wstage + r * STAGE_ROW + 4 + 4 * lane, ring + (off + ks));
      }
      // the eight neighbours outside the chunk in ONE 4-byte copy instruction: lanes 0..3 fetch diagonal kb-1 of
      // o1, o2, i1, i2 (the k-1 sources of the chunk's first cell), lanes 4..7 diagonal kb+128 of o1, o2, d1, d2
      // (the k+1 sources of its last cell)
      if (lane < 8) {
        const int r = lane < 4 ? lane : (lane < 6 ? lane - 4 : lane - 2);  // DS_O1.. order: o1 o2 i1 i2 d1 d2 mis
        const int k = lane < 4 ? pkb - 1 : pkb + 128;
        if (lane < 4 || k <= phi + 7) {
          SR_ASSERT(sp[DS_O1 + r] + k >= 0 && sp[DS_O1 + r] + k < 2 * df.st.scope * 5 * df.st.W, "edge neighbour inside the CTA's rings");
          cp_async4(wstage + r * STAGE_ROW + (lane < 4 ? 3 : 132), ring + (sp[DS_O1 + r] + k));
        }
      }
    }
    cp_async_commit();
  };
#endif
  claim_decode(claim_issue());
#if WFA_DEVICE
  prefetch(n_dir, n_kb);
#endif
  while (true) {
    const int dir = n_dir, kb = n_kb;
    const bool done = dir < 0;
    if (dir != acc_dir) {
      if (acc_dir >= 0) flush(acc_dir);  // the only flush site: a direction change or the end of the pool
      if (done) break;
      sp0.init();
      sp1.init();
      sp2.init();
      sp3.init();
      sp4.init();
      max_ak = 0;
      acc_dir = dir;
    }
    if (done) break;  // a warp that claimed nothing has acc_dir == dir == -1
    const int k0 = kb + 4 * lane;
#if WFA_DEVICE
    const int next_claim = claim_issue();  // the shared-memory atomic's latency hides behind the reads below
    cp_async_wait<0>();                    // this chunk's copies have landed
    __syncwarp();
    // own four cells of every row with one 128-bit shared-memory read, the k0-1 / k0+4 neighbours
    // with scalar reads of the same staged slice
    const int32_t* sb = wstage + 4 + 4 * lane;
    I4 vo1 = ld4(sb), vo2 = ld4(sb + STAGE_ROW), vi1 = ld4(sb + 2 * STAGE_ROW), vi2 = ld4(sb + 3 * STAGE_ROW);
    I4 vd1 = ld4(sb + 4 * STAGE_ROW), vd2 = ld4(sb + 5 * STAGE_ROW), vmis = ld4(sb + 6 * STAGE_ROW);
    int32_t o1L = sb[-1], o2L = sb[STAGE_ROW - 1], i1L = sb[2 * STAGE_ROW - 1], i2L = sb[3 * STAGE_ROW - 1];
    int32_t o1R = sb[4], o2R = sb[STAGE_ROW + 4], d1R = sb[4 * STAGE_ROW + 4], d2R = sb[5 * STAGE_ROW + 4];
    __syncwarp();  // every lane holds its slices in registers: the buffer is free for the next chunk
    claim_decode(next_claim);
    prefetch(n_dir, n_kb);
#endif
    const volatile int* su = cx.sh->ds[dir];
    const I4 qa = lds4v(su + DS_LO), qb = lds4v(su + DS_P0);
    const int lo = qa.x, hi = qa.y;
    const int dp0 = qb.x, dt0 = qb.y, drev = qb.z, dcend = qb.w;
    int32_t ins1[4], ins2[4], del1[4], del2[4], mx[4];
    bool vm[4];
    PROF_T(tc0);
    // lanes of the last chunk that lie wholly beyond hi do not store (they would leave the row)
    const bool lane_act = k0 <= hi;
#if !WFA_DEVICE
    const int lo_al = su[DS_LOAL];
    const int kq = lane_act ? k0 : lo_al;
    const int32_t* const r_o1 = ring + (su[DS_O1] + kq);
    const int32_t* const r_o2 = ring + (su[DS_O2] + kq);
    const int32_t* const r_i1 = ring + (su[DS_I1] + kq);
    const int32_t* const r_i2 = ring + (su[DS_I2] + kq);
    const int32_t* const r_d1 = ring + (su[DS_D1] + kq);
    const int32_t* const r_d2 = ring + (su[DS_D2] + kq);
    I4 vo1 = ld4(r_o1), vo2 = ld4(r_o2), vi1 = ld4(r_i1), vi2 = ld4(r_i2);
    I4 vd1 = ld4(r_d1), vd2 = ld4(r_d2), vmis = ld4(ring + (su[DS_MIS] + kq));
    // host emulation: one "lane", so the k0-1 / k0+4 neighbours are plain reads
    int32_t o1L = r_o1[-1], o2L = r_o2[-1], i1L = r_i1[-1], i2L = r_i2[-1];
    int32_t o1R = r_o1[4], o2R = r_o2[4], d1R = r_d1[4], d2R = r_d2[4];
    claim_decode(claim_issue());
#endif
    const bool interior = kb >= qa.z && kb + CH - 1 <= qa.w;  // warp-uniform
    if (!interior) {
      // fringe chunk: whatever lies outside its source wavefront reads as NULL
      // (WFA/wavefront_compute.c:251-295 pads the same way)
      const I4 ma = lds4v(su + DS_META), mb = lds4v(su + DS_META + 4), mc = lds4v(su + DS_META + 8);
      const WfMeta m_o1 = {ma.x, ma.y}, m_o2 = {ma.z, ma.w};
      const WfMeta m_i1 = {mb.x, mb.y}, m_i2 = {mb.z, mb.w};
      const WfMeta m_d1 = {mc.x, mc.y}, m_d2 = {mc.z, mc.w};
      const WfMeta m_mis = {su[DS_META + 12], su[DS_META + 13]};
#define MSK(v, m, kk) v = ((kk) >= (m).lo && (kk) <= (m).hi) ? v : OFFSET_NULL
      MSK(o1L, m_o1, k0 - 1);
      MSK(vo1.x, m_o1, k0);
      MSK(vo1.y, m_o1, k0 + 1);
      MSK(vo1.z, m_o1, k0 + 2);
      MSK(vo1.w, m_o1, k0 + 3);
      MSK(o1R, m_o1, k0 + 4);
      MSK(o2L, m_o2, k0 - 1);
      MSK(vo2.x, m_o2, k0);
      MSK(vo2.y, m_o2, k0 + 1);
      MSK(vo2.z, m_o2, k0 + 2);
      MSK(vo2.w, m_o2, k0 + 3);
      MSK(o2R, m_o2, k0 + 4);
      MSK(i1L, m_i1, k0 - 1);
      MSK(vi1.x, m_i1, k0);
      MSK(vi1.y, m_i1, k0 + 1);
      MSK(vi1.z, m_i1, k0 + 2);
      MSK(i2L, m_i2, k0 - 1);
      MSK(vi2.x, m_i2, k0);
      MSK(vi2.y, m_i2, k0 + 1);
      MSK(vi2.z, m_i2, k0 + 2);
      MSK(vd1.y, m_d1, k0 + 1);
      MSK(vd1.z, m_d1, k0 + 2);
      MSK(vd1.w, m_d1, k0 + 3);
      MSK(d1R, m_d1, k0 + 4);
      MSK(vd2.y, m_d2, k0 + 1);
      MSK(vd2.z, m_d2, k0 + 2);
      MSK(vd2.w, m_d2, k0 + 3);
      MSK(d2R, m_d2, k0 + 4);
      MSK(vmis.x, m_mis, k0);
      MSK(vmis.y, m_mis, k0 + 1);
      MSK(vmis.z, m_mis, k0 + 2);
      MSK(vmis.w, m_mis, k0 + 3);
#undef MSK
    }
    ins1[0] = imax(o1L, i1L) + 1;
    ins1[1] = imax(vo1.x, vi1.x) + 1;
    ins1[2] = imax(vo1.y, vi1.y) + 1;
    ins1[3] = imax(vo1.z, vi1.z) + 1;
    ins2[0] = imax(o2L, i2L) + 1;
    ins2[1] = imax(vo2.x, vi2.x) + 1;
    ins2[2] = imax(vo2.y, vi2.y) + 1;
    ins2[3] = imax(vo2.z, vi2.z) + 1;
    del1[0] = imax(vo1.y, vd1.y);
    del1[1] = imax(vo1.z, vd1.z);
    del1[2] = imax(vo1.w, vd1.w);
    del1[3] = imax(o1R, d1R);
    del2[0] = imax(vo2.y, vd2.y);
    del2[1] = imax(vo2.z, vd2.z);
    del2[2] = imax(vo2.w, vd2.w);
    del2[3] = imax(o2R, d2R);
    const int32_t mis[4] = {vmis.x + 1, vmis.y + 1, vmis.z + 1, vmis.w + 1};
    bool all_ok = true;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int k = k0 + j;
      int32_t m5 = imax(imax(del1[j], del2[j]), imax(mis[j], imax(ins1[j], ins2[j])));
      vm[j] = in_bounds(m5, k, tlen, plen) && (interior || (k >= lo && k <= hi));
      if (!vm[j]) m5 = OFFSET_NULL;
      mx[j] = m5;
      const int32_t low = imin(imin(ins1[j], ins2[j]), imin(del1[j], del2[j]));
      all_ok = all_ok && vm[j] && low >= imax(0, k);
    }
    // gap components of this score (rows of components that are not allocated this step are never
    // read: their meta says absent)
    SR_ASSERT(!lane_act || (su[DS_WM] + k0 >= 0 && su[DS_WD2] + k0 + 3 < 2 * df.st.scope * 5 * df.st.W && su[DS_WM] + k0 + 3 < 2 * df.st.scope * 5 * df.st.W),
              "row store inside the CTA's rings");
    SR_ASSERT(!lane_act || (k0 >= -(df.st.W >> 1) && k0 + 3 < (df.st.W >> 1)), "diagonal inside the row");
    if (lane_act) {
      const I4 qw = lds4v(su + DS_WI1);
      st4(ring + (qw.x + k0), ins1[0], ins1[1], ins1[2], ins1[3]);
      st4(ring + (qw.y + k0), ins2[0], ins2[1], ins2[2], ins2[3]);
      st4(ring + (qw.z + k0), del1[0], del1[1], del1[2], del1[3]);
      st4(ring + (qw.w + k0), del2[0], del2[1], del2[2], del2[3]);
    }
    if (warp_all(all_ok)) {
      // every cell of the chunk is in bounds in all five components (then all are allocated too)
      sp0.add_full(k0);
      sp1.add_full(k0);
      sp2.add_full(k0);
      sp3.add_full(k0);
      sp4.add_full(k0);
    } else {
      const int fl = su[DS_FLAGS];
      const bool a_i1 = fl & 1, a_i2 = fl & 2, a_d1 = fl & 4, a_d2 = fl & 8;
      bool a0 = true, a1 = true, a2 = true, a3 = true;
      if (!interior) {
        a0 = k0 >= lo && k0 <= hi;
        a1 = k0 + 1 >= lo && k0 + 1 <= hi;
        a2 = k0 + 2 >= lo && k0 + 2 <= hi;
        a3 = k0 + 3 >= lo && k0 + 3 <= hi;
      }
      sp0.add(vm[0], vm[1], vm[2], vm[3], k0);
      if (a_i1)
        sp1.add(a0 && in_bounds(ins1[0], k0, tlen, plen), a1 && in_bounds(ins1[1], k0 + 1, tlen, plen),
                a2 && in_bounds(ins1[2], k0 + 2, tlen, plen), a3 && in_bounds(ins1[3], k0 + 3, tlen, plen), k0);
      if (a_i2)
        sp2.add(a0 && in_bounds(ins2[0], k0, tlen, plen), a1 && in_bounds(ins2[1], k0 + 1, tlen, plen),
                a2 && in_bounds(ins2[2], k0 + 2, tlen, plen), a3 && in_bounds(ins2[3], k0 + 3, tlen, plen), k0);
      if (a_d1)
        sp3.add(a0 && in_bounds(del1[0], k0, tlen, plen), a1 && in_bounds(del1[1], k0 + 1, tlen, plen),
                a2 && in_bounds(del1[2], k0 + 2, tlen, plen), a3 && in_bounds(del1[3], k0 + 3, tlen, plen), k0);
      if (a_d2)
        sp4.add(a0 && in_bounds(del2[0], k0, tlen, plen), a1 && in_bounds(del2[1], k0 + 1, tlen, plen),
                a2 && in_bounds(del2[2], k0 + 2, tlen, plen), a3 && in_bounds(del2[3], k0 + 3, tlen, plen), k0);
    }
    PROF_T(tc1);
    // exact-match extension of the four M cells. The first packed window of all four is compared
    // branch-free (count-zeros returns 32 for a window that matches completely; invalid cells
    // compare position 0 and are discarded by the select); only cells whose whole window
    // matched and that have room left walk on through the loop.
    {
      uint32_t xw[4];
      int room[4];
      const int per = 32 >> d.lbits;
      const int rev = drev, sgn = rev ? -1 : 1;
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int h = vm[j] ? mx[j] : 0, v = vm[j] ? mx[j] - (k0 + j) : 0;
        room[j] = imin(d.plen - v, d.tlen - h);
        xw[j] = pk_window_dir(d.P, dp0 + sgn * v, d.lbits, rev) ^ pk_window_dir(d.T, dt0 + sgn * h, d.lbits, rev);
      }
      int nn[4];
      unsigned need = 0;
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        nn[j] = pk_match_len(xw[j], d.lbits, rev);
        if (vm[j] && xw[j] == 0 && per < room[j]) need |= 1u << j;
      }
      // rare: a cell whose whole first window matched walks on (one copy of the walk for the four cells)
#pragma unroll 1
      while (need) {
        const int j = ctz32z(need);
        need &= need - 1;
        const int mxj = j == 0 ? mx[0] : j == 1 ? mx[1] : j == 2 ? mx[2] : mx[3];
        const int rj = j == 0 ? room[0] : j == 1 ? room[1] : j == 2 ? room[2] : room[3];
        const int h = mxj + per, v = mxj - (k0 + j) + per;
        const int extra = lcp_dir(d.P, d.T, dp0 + sgn * v, dt0 + sgn * h, rj - per, d.lbits, rev);
        nn[0] += j == 0 ? extra : 0;
        nn[1] += j == 1 ? extra : 0;
        nn[2] += j == 2 ? extra : 0;
        nn[3] += j == 3 ? extra : 0;
      }
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int n = nn[j];
        const int32_t e = mx[j] + imin(n, room[j]);
        mx[j] = vm[j] ? e : mx[j];
        max_ak = imax(max_ak, vm[j] ? 2 * e - (k0 + j) : 0);
      }
    }
    if (lane_act) st4(ring + (su[DS_WM] + k0), mx[0], mx[1], mx[2], mx[3]);
    {
      // end cell of this search (WFA/wavefront_termination.c:37-113) for the test in step_finish
      const unsigned je = (unsigned)(d.tlen - d.plen - k0);
      if (je < 4u) {
        const int ce = dcend;
#define SEL4(a) (je == 0 ? a[0] : je == 1 ? a[1] : je == 2 ? a[2] : a[3])
        const int32_t e_m = SEL4(mx), e_i1 = SEL4(ins1), e_i2 = SEL4(ins2), e_d1 = SEL4(del1), e_d2 = SEL4(del2);
#undef SEL4
        cx.sh->end_val[dir] = ce == C_M ? e_m : ce == C_I1 ? e_i1 : ce == C_I2 ? e_i2 : ce == C_D1 ? e_d1 : e_d2;
      }
    }
#ifdef WFA_PROFILE
    if (cx.tid == 32 && interior) {
      const long long tc2 = clock64();
      cx.sh->prof[8] += 1;
      cx.sh->prof[9] += tc1 - tc0;
      cx.sh->prof[10] += tc2 - tc1;
    }
    if (cx.tid == 0 && !interior) {
      cx.sh->prof[11] += 1;
      cx.sh->prof[12] += clock64() - tc0;
    }
#endif
  }
  PROF_T(t3);
  cta_sync();  // E
  // F: publish (lanes 0..4 of warp 0 for the forward direction, of warp 1 -- or threads 5..9 of a
  // one-warp CTA -- for the reverse direction), results for every thread
#pragma unroll 1
  for (int dir = 0; dir < 2; ++dir) {
    if (!(dir ? run_r : run_f)) continue;
    StepOut o = dir ? out.r : out.f;
    if (o.null_step) continue;
    const Dir dd = dir ? dr : df;
    const volatile int* su = cx.sh->ds[dir];
    const RedSlots* red = &cx.sh->red[dir];
    StepCtl t;
    t.s = su[DS_S];
    t.slot5 = su[DS_SLOT5];
    t.s_o1 = su[DS_SO1];
    t.s_e1 = su[DS_SE1];
    t.s_o2 = su[DS_SO2];
    t.s_e2 = su[DS_SE2];
    const int nf = su[DS_NFLAGS], fl = su[DS_FLAGS];
    t.n_o1 = nf & 1;
    t.n_o2 = nf & 2;
    t.n_i1 = nf & 4;
    t.n_d1 = nf & 8;
    t.n_i2 = nf & 16;
    t.n_d2 = nf & 32;
    t.a_i1 = fl & 1;
    t.a_i2 = fl & 2;
    t.a_d1 = fl & 4;
    t.a_d2 = fl & 8;
    o.max_ak = red->max_ak;
#if WFA_DEVICE
    const int pub0 = (nwarps > 1 && dir == 1) ? 32 : 5 * dir;
    if (cx.tid >= pub0 && cx.tid < pub0 + 5) publish_comp(dd.st, red, t, cx.tid - pub0);
#else
    for (int c = 0; c < 5; ++c) publish_comp(dd.st, red, t, c);
#endif
    if (cx.tid == 0) cx.sh->cells += (unsigned long long)su[DS_WIDTH];
    // termination (WFA/wavefront_termination.c:37-113) from the reduction slots and the stashed end cell
    {
      const int ak = dd.tlen - dd.plen;
      const int ce = dd.cend;
      const bool al = ce == C_M ? true : ce == C_I1 ? t.a_i1 : ce == C_I2 ? t.a_i2 : ce == C_D1 ? t.a_d1 : t.a_d2;
      const bool present = al && red->mn[ce] <= red->mx[ce];
      if (present && red->mn[ce] <= ak && ak <= red->mx[ce]) {
        if (cx.sh->end_val[dir] >= dd.tlen) o.end_reached = 1;
      }
    }
    if (dir) out.r = o;
    else out.f = o;
  }
#ifdef WFA_PROFILE
  if (cx.tid == 0) {
    const long long t6 = clock64();
    unsigned long long* pf = cx.sh->prof;
    pf[0] += (run_f ? 1 : 0) + (run_r ? 1 : 0);
    pf[1] += t1 - t0;  // wait at the top barrier
    pf[2] += t2 - t1;  // setup (one warp per direction) + barrier
    pf[3] += t3 - t2;  // chunk loop (warp 0)
    pf[4] += t6 - t3;  // closing barrier, publish, end test
    pf[7] += (unsigned long long)(cx.sh->ds[0][DS_NCH] ? cx.sh->ds[0][DS_WIDTH] : 0) + (cx.sh->ds[1][DS_NCH] ? cx.sh->ds[1][DS_WIDTH] : 0);
    pf[16] += 1;  // passes
    pf[17] += (unsigned long long)nch_all;
    pf[18] += (nch_all <= nwarps) ? 1 : 0;
  }
#endif
  return out;
}

// reset reduction slots (thread 0) -- call before the first wf_step of a search
WFA_FN void red_reset(Ctx& cx) {
  if (cx.tid == 0) {
    for (int b = 0; b < 2; ++b) {
      for (int c = 0; c < 5; ++c) {
        cx.sh->red[b].mn[c] = INT32_MAX;
        cx.sh->red[b].mx[c] = INT32_MIN;
      }
      cx.sh->red[b].max_ak = 0;
    }
    cx.sh->chunk_ctr = 0;
  }
  cx.parity = 0;
}

// ---------------------------------------------------------------------------
// Overlap search between the two directions (WFA/wavefront_bialign.c:877-955,
// per-component scans :476-571 and :796-872). The reference walks 26 scores x
// {D2,I2,D1,I1,M} sequentially, each a linear scan for the FIRST diagonal k_0
// with h_0 + h_1 >= tlen. Here every (score, component) scan is a task handled
// by one warp (lanes stride over k_0, warp-min of the first hit); thread 0
// then replays the reference's sequential acceptance order over the task
// results, so the chosen breakpoint is identical.
// ---------------------------------------------------------------------------
WFA_FN int ovl_comp_of_task(int j) {
  // order inside one score_i: D2, I2, D1, I1, M
  return j == 0 ? C_D2 : j == 1 ? C_I2 : j == 2 ? C_D1 : j == 3 ? C_I1 : C_M;
}

WFA_FN void wf_overlap(Ctx& cx, const Dir& d0, const Dir& d1, const Penalties& pen, int score_0, int score_1,
                       bool breakpoint_forward) {
  if (cx.tid == 0) cx.sh->ovl_any = 0;
  cta_sync();
  const int scope = pen.scope;
  const int plen = d0.plen, tlen = d0.tlen;
  const int kinv = tlen - plen;  // K_INVERSE(k) = kinv - k
  const int bp_score_in = cx.sh->bp.score;
  // reference: `if (mwf_0 == NULL) return;` -- pointer-NULL only (absent), an allocated-but-empty M goes on
  const WfMeta m0raw = d0.st.meta[d0.st.slot(score_0) * 5 + C_M];
  const bool m0_present = !(m0raw.lo == 1 && m0raw.hi == -1);
  const int ntasks = scope * 5;
#if WFA_DEVICE
  const int warp = cx.tid >> 5, nwarps = cx.nt >> 5, lane = cx.tid & 31, lanes = 32;
#else
  const int warp = 0, nwarps = 1, lane = 0, lanes = 1;
#endif
  // The run / skip decision of a task reads a few shared-memory words and is made by one LANE per task (32 tasks per
  // warp round); only the tasks that survive it (few: the antidiagonal bounds reject most rows) are scanned, each by
  // the whole warp.
  for (int tb = warp * lanes; tb < ntasks; tb += nwarps * lanes) {
    const int task = tb + lane;
    const int i = task / 5, j = task - i * 5;
    const int c = ovl_comp_of_task(j);
    const int score_i = score_1 - i;
    bool run = task < ntasks && score_i >= 0;
    const int gopen = (c == C_M) ? 0 : ((c == C_I1 || c == C_D1) ? pen.o1 : pen.o2);
    if (run && score_0 + score_i - gopen >= bp_score_in) run = false;
    // h0 + h1 >= tlen  <=>  (2*h0 - k0) + (2*h1 - k1) >= plen + tlen: rows whose antidiagonal
    // bounds cannot reach that are skipped without a scan (the scan would find nothing)
    if (run && d0.st.ub[d0.st.slot(score_0) * 5 + c] + d1.st.ub[d1.st.slot(score_i) * 5 + c] < plen + tlen) run = false;
    int max_lo = 1, min_hi = 0;
    if (run) {
      const WfMeta a = d0.st.get(score_0, c);
      const WfMeta b = d1.st.get(score_i, c);
      if (a.lo <= a.hi && b.lo <= b.hi) {
        const int lo_1 = kinv - b.hi, hi_1 = kinv - b.lo;
        max_lo = imax(a.lo, lo_1);
        min_hi = imin(a.hi, hi_1);
      }
      run = max_lo <= min_hi;
    }
    if (task < ntasks) cx.sh->ovl[task] = INT32_MAX;
    uint32_t todo = warp_ballot(run);
#if WFA_DEVICE
    __syncwarp();
#endif
    while (todo) {
      const int src = ctz32z(todo);
      todo &= todo - 1;
#if WFA_DEVICE
      const int t_c = __shfl_sync(0xffffffffu, c, src), t_si = __shfl_sync(0xffffffffu, score_i, src);
      const int t_lo = __shfl_sync(0xffffffffu, max_lo, src), t_hi = __shfl_sync(0xffffffffu, min_hi, src);
#else
      const int t_c = c, t_si = score_i, t_lo = max_lo, t_hi = min_hi;
#endif
      const int32_t* r0 = d0.st.row(score_0, t_c);
      const int32_t* r1 = d1.st.row(t_si, t_c);
      int best = INT32_MAX;
      for (int k0 = t_lo + lane; k0 <= t_hi; k0 += lanes) {
        const int k1 = kinv - k0;
        const int32_t h0 = r0[k0], h1 = r1[k1];
        if (h0 + h1 >= tlen) {
          bool ok = true;
          if (t_c != C_M) {
            // out-of-bounds guard of the indel scans (WFA/wavefront_bialign.c:534-552)
            const int kk = breakpoint_forward ? k0 : k1;
            const int hh = breakpoint_forward ? h0 : h1;
            if (hh - kk > plen || hh > tlen) ok = false;
          }
          if (ok) {
            best = k0;
            break;
          }
        }
      }
      best = warp_min(best);
      if (lane == 0) {
        cx.sh->ovl[tb + src] = best;
        if (best != INT32_MAX) cx.sh->ovl_any = 1;
      }
    }
  }
  cta_sync();
  if (cx.tid == 0 && m0_present && cx.sh->ovl_any) {  // without a hit the replay below changes nothing
    Breakpoint bp = cx.sh->bp;
    for (int i = 0; i < scope; ++i) {
      const int score_i = score_1 - i;
      if (score_i < 0) break;
      // I2/D2 (WFA/wavefront_bialign.c:905-921)
      if (score_0 + score_i - pen.o2 >= bp.score) continue;
      bool skip = false;
      for (int j = 0; j < 5 && !skip; ++j) {
        const int c = ovl_comp_of_task(j);
        if (j == 2 && score_0 + score_i - pen.o1 >= bp.score) {
          skip = true;
          break;
        }
        if (j == 4 && score_0 + score_i >= bp.score) {
          skip = true;
          break;
        }
        const int gopen = (c == C_M) ? 0 : ((c == C_I1 || c == C_D1) ? pen.o1 : pen.o2);
        if (score_0 + score_i - gopen >= bp.score) continue;  // inner guard of the scan functions
        const int k0 = cx.sh->ovl[i * 5 + j];
        if (k0 == INT32_MAX) continue;
        const int k1 = kinv - k0;
        const int32_t h0 = d0.st.row(score_0, c)[k0];
        const int32_t h1 = d1.st.row(score_i, c)[k1];
        if (breakpoint_forward) {
          bp.score_forward = score_0;
          bp.score_reverse = score_i;
          bp.k_forward = k0;
          bp.k_reverse = k1;
          bp.offset_forward = h0;
          bp.offset_reverse = h1;
        } else {
          bp.score_forward = score_i;
          bp.score_reverse = score_0;
          bp.k_forward = k1;
          bp.k_reverse = k0;
          bp.offset_forward = h1;
          bp.offset_reverse = h0;
        }
        bp.score = score_0 + score_i - gopen;
        bp.component = c;
      }
    }
    cx.sh->bp = bp;
  }
  cta_sync();
}

// ---------------------------------------------------------------------------
// Breakpoint search (WFA/wavefront_bialign.c:974-1082). Returns:
//   0 = breakpoint found (in sh->bp), 1 = end reached (score in *score_reached),
//   2 = workspace overflow, 3 = score bound exceeded (ST_RUNAWAY: the host fails the call loudly).
// ---------------------------------------------------------------------------
WFA_FN int wf_find_breakpoint(Ctx& cx, Dir& df, Dir& dr, const Penalties& pen, int* score_reached,
                              long long score_hint) {
  cta_sync();
  df.st.cur_s = INT32_MIN / 2;
  dr.st.cur_s = INT32_MIN / 2;
  df.st.set_cur(0);
  dr.st.set_cur(0);
  if (cx.tid == 0) {
    wf_init(df);
    wf_init(dr);
    cx.sh->bp.score = INT32_MAX;
  }
  red_reset(cx);
  const int max_antidiagonal = df.plen + df.tlen - 1;
  int score_f = 0, score_r = 0, fwd_max_ak = 0, rev_max_ak = 0;
  StepOut o = wf_extend0(cx, df);
  if (o.end_reached) {
    *score_reached = 0;
    return 1;
  }
  fwd_max_ak = o.max_ak;
  o = wf_extend0(cx, dr);
  if (o.end_reached) {
    *score_reached = 0;
    return 1;
  }
  rev_max_ak = o.max_ak;
  // The reference runs two loops (WFA/wavefront_bialign.c:1010-1043 then :1045-1080): first
  // alternate forward / reverse steps until the wavefronts are close to colliding, then keep
  // alternating with an overlap search before every step, starting with the direction that did
  // not move last. Both loops alternate strictly, so they are one loop with a phase flag here
  // (each step function is instantiated once).
  const int scope = pen.scope;
  const int gap_opening = imax(pen.o1, pen.o2);
  bool close = false;        // second loop reached
  bool turn_forward = true;  // direction of the next step
  // The two searches meet once score_f + score_r passes the optimal penalty by at most a few
  // scopes. score_hint bounds that penalty from above (sub-problems: the exact score the parent's
  // breakpoint assigned; level 0: what the edit-distance alignment of the strand choice would cost;
  // else deleting the pattern and inserting the text). A search that runs past the bound is broken
  // (workspace corruption) and is reported as ST_RUNAWAY instead of spinning on.
  long long score_bound = 2ll * imax(pen.o1, pen.o2) + (long long)(df.plen + df.tlen) * imax(pen.e1, pen.e2);
  if (score_hint < score_bound) score_bound = score_hint;
  score_bound += 4ll * pen.scope + 64;
  // While the searches are far apart both directions advance in one dual step. The reference tests
  // for closeness between the forward and the reverse step; when that test fires after the forward
  // half, the reverse half has already run, so its result is parked and the loop replays the
  // reference order: the reverse turn of the second phase does its overlap search (the ring keeps
  // scope+1 score slots, so the row of score_r-25 is still intact beside the new row of score_r+1)
  // and then takes the parked step instead of computing it.
  bool parked = false;
  StepOut parked_out = o;
  while (true) {
    if ((long long)score_f + score_r > score_bound) return 3;
#ifdef SRGPU_EXP_ABORT_CLOSE  // measurement build only: stop every pair where the close phase would begin
    if (!close && fwd_max_ak + rev_max_ak >= max_antidiagonal) return 4;
#endif
#ifdef WFA_PROFILE
    if (!close && fwd_max_ak + rev_max_ak >= max_antidiagonal && cx.tid == 0) cx.sh->prof2[16] = clock64();
#endif
    if (!close && fwd_max_ak + rev_max_ak >= max_antidiagonal) close = true;
    const bool tf = turn_forward;
    const bool dual = !close && tf;
    const int own = tf ? score_f : score_r, oth = tf ? score_r : score_f;
    if (close) {
      const int min_own = (own > scope - 1) ? own - (scope - 1) : 0;
      cta_sync();
      if (min_own + oth - gap_opening >= cx.sh->bp.score) break;
      // the wavefront that moved last is searched against the scope of the one about to move
      const Dir a = tf ? dr : df, b = tf ? df : dr;
      PROF_T(tov0);
      wf_overlap(cx, a, b, pen, oth, own, !tf);
#ifdef WFA_PROFILE
      if (cx.tid == 0) {
        cx.sh->prof2[8] += clock64() - tov0;
        cx.sh->prof2[10] += 1;
      }
#endif
    }
    StepOut2 o2;
    o2.f = o;
    o2.r = o;
    const bool run_f = dual || tf, run_r = (dual || !tf) && !parked;
    if (run_f || run_r) o2 = wf_step_multi(cx, df, dr, pen, score_f + 1, score_r + 1, run_f, run_r);
    if (run_f) {
      ++score_f;
      if (o2.f.overflow) return 2;
      if (!close && fwd_max_ak < o2.f.max_ak) fwd_max_ak = o2.f.max_ak;
      if (o2.f.end_reached) {
        *score_reached = score_f;
        return 1;
      }
      turn_forward = false;
#ifdef SRGPU_EXP_ABORT_CLOSE
      if (dual && fwd_max_ak + rev_max_ak >= max_antidiagonal) return 4;
#endif
      if (dual && fwd_max_ak + rev_max_ak >= max_antidiagonal) {
        // closeness reached between the two halves: replay the reverse turn in reference order
        parked = true;
        parked_out = o2.r;
        continue;
      }
    }
    if (dual || !tf) {
      StepOut orv = o2.r;
      if (parked) {
        orv = parked_out;
        parked = false;
      }
      ++score_r;
      if (orv.overflow) return 2;
      if (!close && rev_max_ak < orv.max_ak) rev_max_ak = orv.max_ak;
      if (orv.end_reached) {
        *score_reached = score_r;
        return 1;
      }
      turn_forward = true;
    }
  }
  return 0;
}

// ---------------------------------------------------------------------------
// Run-length CIGAR output for one pair. `runs` has capacity `cap`.
// ---------------------------------------------------------------------------
struct RunOut {
  uint32_t* runs;
  uint32_t cap;
  uint32_t* bt;  // backtrace scratch (runs emitted back-to-front)
  uint32_t bt_cap;
};

// thread 0 only; keeps the run list canonical (adjacent runs never share an op)
WFA_FN bool runs_push(CtaShared* sh, const RunOut& ro, uint32_t op, uint32_t len) {
  if (len == 0) return true;
  if (sh->n_runs > 0 && (ro.runs[sh->n_runs - 1] & 3u) == op) {
    ro.runs[sh->n_runs - 1] += (len << 2);
    return true;
  }
  if (sh->n_runs >= ro.cap) return false;
  SR_ASSERT(sh->n_runs < ro.cap, "run buffer");
  ro.runs[sh->n_runs++] = make_run(op, len);
  return true;
}
WFA_FN bool bt_push(CtaShared* sh, const RunOut& ro, uint32_t op, uint32_t len) {
  if (len == 0) return true;
  if (sh->bt_n > 0) {
    const uint32_t last = ro.bt[sh->bt_n - 1];
    if ((last & 3u) == op) {
      ro.bt[sh->bt_n - 1] = last + (len << 2);
      return true;
    }
  }
  if (sh->bt_n >= ro.bt_cap) return false;
  ro.bt[sh->bt_n++] = make_run(op, len);
  return true;
}

// ---------------------------------------------------------------------------
// Backtrace over the full-memory store (WFA/wavefront_backtrace.c:320-529).
// Sequential by nature (each step depends on the previous choice); thread 0
// walks it, emitting runs back-to-front. Priority is the reference's packed
// max: (offset << 4) | type with M=9 > D2_ext=8 > D2_open=7 > D1_ext=6 >
// D1_open=5 > I2_ext=4 > I2_open=3 > I1_ext=2 > I1_open=1 (:49-59).
// ---------------------------------------------------------------------------
WFA_FN int64_t bt_src(const WfStore& st, int score, int c, int k, int add, int type) {
  if (score < 0) return OFFSET_NULL;
  const WfMeta m = st.meta[st.slot(score) * 5 + c];
  if (m.lo <= k && k <= m.hi) {
    const int64_t off = (int64_t)st.row(score, c)[k] + add;
    return (int64_t)((uint64_t)off << 4) | (int64_t)type;  // arithmetic value == (off<<4)|type
  }
  return OFFSET_NULL;
}
WFA_FN int64_t i64max(int64_t a, int64_t b) { return a > b ? a : b; }

WFA_FN bool wf_backtrace(CtaShared* sh, const RunOut& ro, const WfStore& st, const Penalties& pen, int plen, int tlen,
                         int cbegin, int cend, int alignment_score) {
  (void)cbegin;
  int matrix = cend;
  int score = alignment_score;
  int k = tlen - plen;
  int32_t offset = tlen;
  int h = offset, v = offset - k;
  bool ok = true;
  sh->bt_n = 0;
  // (ending indels for M end: v == plen and h == tlen here, nothing to add)
  while (v > 0 && h > 0 && score > 0) {
    const int mismatch = score - pen.x;
    const int gap_open1 = score - pen.o1 - pen.e1, gap_open2 = score - pen.o2 - pen.e2;
    const int gap_ext1 = score - pen.e1, gap_ext2 = score - pen.e2;
    int64_t max_all;
    if (matrix == C_M) {
      const int64_t misms = bt_src(st, mismatch, C_M, k, 1, 9);
      const int64_t ins1 = i64max(bt_src(st, gap_open1, C_M, k - 1, 1, 1), bt_src(st, gap_ext1, C_I1, k - 1, 1, 2));
      const int64_t del1 = i64max(bt_src(st, gap_open1, C_M, k + 1, 0, 5), bt_src(st, gap_ext1, C_D1, k + 1, 0, 6));
      const int64_t ins2 = i64max(bt_src(st, gap_open2, C_M, k - 1, 1, 3), bt_src(st, gap_ext2, C_I2, k - 1, 1, 4));
      const int64_t del2 = i64max(bt_src(st, gap_open2, C_M, k + 1, 0, 7), bt_src(st, gap_ext2, C_D2, k + 1, 0, 8));
      max_all = i64max(misms, i64max(i64max(ins1, ins2), i64max(del1, del2)));
    } else if (matrix == C_I1) {
      max_all = i64max(bt_src(st, gap_open1, C_M, k - 1, 1, 1), bt_src(st, gap_ext1, C_I1, k - 1, 1, 2));
    } else if (matrix == C_I2) {
      max_all = i64max(bt_src(st, gap_open2, C_M, k - 1, 1, 3), bt_src(st, gap_ext2, C_I2, k - 1, 1, 4));
    } else if (matrix == C_D1) {
      max_all = i64max(bt_src(st, gap_open1, C_M, k + 1, 0, 5), bt_src(st, gap_ext1, C_D1, k + 1, 0, 6));
    } else {
      max_all = i64max(bt_src(st, gap_open2, C_M, k + 1, 0, 7), bt_src(st, gap_ext2, C_D2, k + 1, 0, 8));
    }
    if (max_all < 0) break;
    if (matrix == C_M) {
      const int max_offset = (int)(max_all >> 4);
      ok &= bt_push(sh, ro, OP_M, (uint32_t)(offset - max_offset));
      offset = max_offset;
      v = offset - k;
      h = offset;
      if (v <= 0 || h <= 0) break;
    }
    const int type = (int)(max_all & 15);
    switch (type) {
      case 9: score = mismatch; matrix = C_M; break;
      case 1: score = gap_open1; matrix = C_M; break;
      case 2: score = gap_ext1; matrix = C_I1; break;
      case 3: score = gap_open2; matrix = C_M; break;
      case 4: score = gap_ext2; matrix = C_I2; break;
      case 5: score = gap_open1; matrix = C_M; break;
      case 6: score = gap_ext1; matrix = C_D1; break;
      case 7: score = gap_open2; matrix = C_M; break;
      default: score = gap_ext2; matrix = C_D2; break;
    }
    if (type == 9) {
      ok &= bt_push(sh, ro, OP_X, 1);
      --offset;
    } else if (type <= 4) {
      ok &= bt_push(sh, ro, OP_I, 1);
      --k;
      --offset;
    } else {
      ok &= bt_push(sh, ro, OP_D, 1);
      ++k;
    }
    v = offset - k;
    h = offset;
  }
  if (matrix == C_M) {
    if (v > 0 && h > 0) {
      const int nm = imin(v, h);
      ok &= bt_push(sh, ro, OP_M, (uint32_t)nm);
      v -= nm;
      h -= nm;
    }
    if (v > 0) ok &= bt_push(sh, ro, OP_D, (uint32_t)v);
    if (h > 0) ok &= bt_push(sh, ro, OP_I, (uint32_t)h);
  }
  return ok;
}

// ---------------------------------------------------------------------------
// Base case: classic full-memory WFA + backtrace (WFA/wavefront_bialign.c:159-189
// -> WFA/wavefront_unialign.c:242-273). Returns ST_*.
// ---------------------------------------------------------------------------
WFA_FN int wf_base(Ctx& cx, Dir& db, const Penalties& pen, const RunOut& ro) {
  cta_sync();
  db.st.set_cur(0);
  if (cx.tid == 0) wf_init(db);
  red_reset(cx);
  int score = 0;
  StepOut o = wf_extend0(cx, db);
  while (!o.end_reached) {
    ++score;
    o = wf_step(cx, db, pen, score);
    if (o.overflow) return ST_WS_BASE_OVERFLOW;
  }
  cta_sync();
  if (cx.tid == 0) {
    PROF_T(tbt0);
    const bool ok = wf_backtrace(cx.sh, ro, db.st, pen, db.plen, db.tlen, db.cbegin, db.cend, score);
#ifdef WFA_PROFILE
    cx.sh->prof2[7] += clock64() - tbt0;
#endif
    cx.sh->bcast[1] = ok ? 1 : 0;
    cx.sh->bcast[2] = (int)cx.sh->n_runs;
    if (ok && cx.sh->n_runs + cx.sh->bt_n > ro.cap) cx.sh->bcast[1] = 0;
  }
  cta_sync();
  if (!cx.sh->bcast[1]) return ST_WS_RUNS_OVERFLOW;
  // append forward (WFA2-lib/alignment/cigar.c:125-137): reverse the back-to-front runs; the
  // first appended run is merged into the previous one when the ops agree (canonical RLE)
  const uint32_t n = cx.sh->bt_n;
  uint32_t base = (uint32_t)cx.sh->bcast[2];
  uint32_t skip = 0;
  if (n > 0 && base > 0 && ((ro.runs[base - 1] ^ ro.bt[n - 1]) & 3u) == 0) skip = 1;
  cta_sync();
  if (skip && cx.tid == 0) ro.runs[base - 1] += (ro.bt[n - 1] >> 2) << 2;
  for (uint32_t i = skip + cx.tid; i < n; i += cx.nt) ro.runs[base + i - skip] = ro.bt[n - 1 - i];
  cta_sync();
  if (cx.tid == 0) cx.sh->n_runs = base + n - skip;
  return ST_OK;
}

// ---------------------------------------------------------------------------
// Whole-pair driver: the BiWFA recursion (WFA/wavefront_bialign.c:1144-1221,
// dispatcher :1266-1293) unrolled onto an explicit stack in shared memory.
// ---------------------------------------------------------------------------
struct PairWorkspace {
  int32_t* ring_f;  // scope*5*W
  int32_t* ring_r;
  int W;
  int32_t* full;    // 5*cap*cap
  WfMeta* full_meta;  // cap*5
  int cap;
};

struct PairResult {
  int status;
  int score;  // raw cigar->score semantics: -breakpoint.score when level 0 split, INT32_MIN otherwise
  uint32_t n_runs;
};

WFA_FN void set_dirs(Dir& df, Dir& dr, Dir& db, const SeqView& sv, const Frame& f) {
  const int plen = f.pe - f.pb, tlen = f.te - f.tb;
  df.P = sv.p;
  df.T = sv.t;
  df.p0 = f.pb;
  df.t0 = f.tb;
  df.lbits = sv.lbits;
  df.plen = plen;
  df.tlen = tlen;
  df.reverse = 0;
  df.cbegin = f.cbegin;
  df.cend = f.cend;
  dr.P = sv.p;
  dr.T = sv.t;
  dr.p0 = f.pe - 1;
  dr.t0 = f.te - 1;
  dr.lbits = sv.lbits;
  dr.plen = plen;
  dr.tlen = tlen;
  dr.reverse = 1;
  dr.cbegin = f.cend;
  dr.cend = f.cbegin;
  db.P = df.P;
  db.T = df.T;
  db.p0 = df.p0;
  db.t0 = df.t0;
  db.lbits = sv.lbits;
  db.plen = plen;
  db.tlen = tlen;
  db.reverse = 0;
  db.cbegin = f.cbegin;
  db.cend = f.cend;
}

// Driver over the BiWFA recursion starting from frame f0 (the whole pair, or one sub-problem of it: the two halves
// after a breakpoint are independent, WFA/wavefront_bialign.c:1193-1212, only the order of their CIGAR pieces
// matters). SPLIT MODE (leaves != nullptr): frames that are trivial, whose score_remaining is at most leaf_T
// (>= 250, so every base case is one), or whose search ended by reaching the end are not solved but emitted, in
// left-to-right order, into leaves[]; res.n_runs is then the number of leaves. A one-warp-per-task kernel solves them
// and the pieces are joined in order afterwards (kernels.cuh: subalign_kernel, assemble_kernel).
WFA_FN PairResult wf_align_from(Ctx& cx, const SeqView& sv, const Frame f0, const Penalties& pen, const PairWorkspace& ws,
                                const RunOut& ro, long long score_hint, Frame* leaves, uint32_t leaf_cap, int leaf_T) {
  CtaShared* sh = cx.sh;
  PairResult res;
  res.status = ST_OK;
  res.score = INT32_MIN;
  res.n_runs = 0;
  Dir df, dr, db;
  df.st.base = ws.ring_f;
  df.st.meta = sh->meta_f;
  df.st.ub = sh->ub_f;
  df.st.modular = 1;
  df.st.scope = pen.scope + 1;  // ring slots: one more than the reference's scope (see the dual step)
  df.st.W = ws.W;
  df.st.cap = 0;
  df.st.cur_s = INT32_MIN / 2;
  df.st.cur_slot = 0;
  dr.st = df.st;
  dr.st.base = ws.ring_r;
  dr.st.meta = sh->meta_r;
  dr.st.ub = sh->ub_r;
  db.st.base = ws.full;
  db.st.meta = ws.full_meta;
  db.st.ub = nullptr;
  db.st.cur_s = 0;
  db.st.cur_slot = 0;
  db.st.modular = 0;
  db.st.scope = pen.scope;
  db.st.W = 0;
  db.st.cap = ws.cap;
  cta_sync();
  if (cx.tid == 0) {
    sh->n_runs = 0;
    sh->cells = 0;
    sh->sp = 0;
    sh->status = ST_OK;
    sh->stack[sh->sp++] = f0;
  }
  cta_sync();
  int level0_score = INT32_MIN;
  const bool split = leaves != nullptr;
  auto emit_leaf = [&](const Frame& lf) {  // thread 0
    if (sh->n_runs >= leaf_cap) sh->status = ST_WS_RUNS_OVERFLOW;
    else leaves[sh->n_runs++] = lf;
  };
  while (true) {
    cta_sync();
    if (sh->sp == 0 || sh->status != ST_OK) break;
    const Frame f = sh->stack[sh->sp - 1];
    cta_sync();
    if (cx.tid == 0) sh->sp--;
    const int fp = f.pe - f.pb, ft = f.te - f.tb;
    int st = ST_OK;
    if (split && (ft == 0 || fp == 0 || f.score_remaining <= leaf_T)) {
      if (cx.tid == 0) emit_leaf(f);
      continue;
    }
    if (ft == 0) {
      if (cx.tid == 0 && !runs_push(sh, ro, OP_D, (uint32_t)fp)) sh->status = ST_WS_RUNS_OVERFLOW;
      continue;
    } else if (fp == 0) {
      if (cx.tid == 0 && !runs_push(sh, ro, OP_I, (uint32_t)ft)) sh->status = ST_WS_RUNS_OVERFLOW;
      continue;
    }
    set_dirs(df, dr, db, sv, f);
    if (f.score_remaining <= 250) {  // WF_BIALIGN_FALLBACK_MIN_SCORE
      PROF_T(tb0);
      st = wf_base(cx, db, pen, ro);
#ifdef WFA_PROFILE
      if (cx.tid == 0) {
        sh->prof2[5] += clock64() - tb0;
        sh->prof2[6] += 1;
      }
#endif
      if (st != ST_OK && cx.tid == 0) sh->status = st;
      continue;
    }
    int score_reached = 0;
#ifdef WFA_PROFILE
    const long long tfb0 = clock64();
    const unsigned long long cells0 = sh->cells;
    if (cx.tid == 0) sh->prof2[16] = 0;
#endif
    const int r = wf_find_breakpoint(cx, df, dr, pen, &score_reached,
                                     f.level == 0 ? score_hint : (long long)f.score_remaining);
#ifdef WFA_PROFILE
    cta_sync();
    if (cx.tid == 0) {
      const long long tfb1 = clock64();
      const int lb = f.level < 4 ? f.level : 4;
      sh->prof2[lb] += tfb1 - tfb0;
      sh->prof2[11 + lb] += sh->cells - cells0;
      if (sh->prof2[16]) sh->prof2[9] += tfb1 - (long long)sh->prof2[16];
    }
#endif
    if (r == 2) {
      if (cx.tid == 0) sh->status = ST_WS_DIAG_OVERFLOW;
      continue;
    }
    if (r == 3) {
      if (cx.tid == 0) sh->status = ST_RUNAWAY;
      continue;
    }
#ifdef SRGPU_EXP_ABORT_CLOSE
    if (r == 4) {
      if (cx.tid == 0) sh->status = ST_UNATTAINABLE;
      continue;
    }
#endif
    if (r == 1) {
      // WFA/wavefront_bialign.c:1083-1112: recover through the base case when cheap enough
      if (score_reached <= 500 && split) {  // the task repeats this (cheap: <= 500 scores) search and recovers there
        if (cx.tid == 0) emit_leaf(f);
      } else if (score_reached <= 500) {  // WF_BIALIGN_RECOVERY_MIN_SCORE
        st = wf_base(cx, db, pen, ro);
        if (st != ST_OK && cx.tid == 0) sh->status = st;
      } else if (cx.tid == 0) {
        sh->status = ST_UNATTAINABLE;
      }
      continue;
    }
    cta_sync();
    const Breakpoint bp = sh->bp;
    const int bh = bp.offset_forward, bv = bp.offset_forward - bp.k_forward;
    if (f.level == 0) {
      // WFA/wavefront_bialign.c:1214-1219 -> wavefront_compute_classic_score (WFA/wavefront_compute.c:108-120):
      // -score for match == 0, else WF_SCORE_TO_SW_SCORE = (swg_match * (plen + tlen) - score) / 2
      level0_score = pen.match == 0 ? -bp.score : ((-pen.match) * ((f.pe - f.pb) + (f.te - f.tb)) - bp.score) / 2;
    }
    cta_sync();
    if (cx.tid == 0) {
      if (sh->sp + 2 > sh->stack_max) {
        sh->status = ST_STACK_OVERFLOW;
      } else {
        Frame h1;
        h1.pb = f.pb + bv;
        h1.pe = f.pe;
        h1.tb = f.tb + bh;
        h1.te = f.te;
        h1.cbegin = bp.component;
        h1.cend = f.cend;
        h1.score_remaining = bp.score_reverse;
        h1.level = f.level + 1;
        Frame h0;
        h0.pb = f.pb;
        h0.pe = f.pb + bv;
        h0.tb = f.tb;
        h0.te = f.tb + bh;
        h0.cbegin = f.cbegin;
        h0.cend = bp.component;
        h0.score_remaining = bp.score_forward;
        h0.level = f.level + 1;
        sh->stack[sh->sp++] = h1;  // processed second
        sh->stack[sh->sp++] = h0;  // processed first
      }
    }
  }
  cta_sync();
  res.status = sh->status;
  res.score = level0_score;
  res.n_runs = sh->n_runs;
  return res;
}

// A leaf of the recursion on its own (second launch of the one-warp kernel's two-pass mode): an empty side is one
// gap run, everything else goes through the base case -- sub-problems at or below the reference's base-case bound, and
// those whose search reached the end within the recovery bound (WFA/wavefront_bialign.c:1083-1112 takes the base case
// for them too). No breakpoint search: the code and the registers of this path are a fraction of wf_align_from's.
WFA_FN PairResult wf_align_leaf(Ctx& cx, const SeqView& sv, const Frame f, const Penalties& pen, const PairWorkspace& ws,
                                const RunOut& ro) {
  CtaShared* sh = cx.sh;
  PairResult res;
  res.status = ST_OK;
  res.score = INT32_MIN;
  res.n_runs = 0;
  Dir df, dr, db;
  db.st.base = ws.full;
  db.st.meta = ws.full_meta;
  db.st.ub = nullptr;
  db.st.cur_s = 0;
  db.st.cur_slot = 0;
  db.st.modular = 0;
  db.st.scope = pen.scope;
  db.st.W = 0;
  db.st.cap = ws.cap;
  cta_sync();
  if (cx.tid == 0) {
    sh->n_runs = 0;
    sh->cells = 0;
    sh->sp = 0;
    sh->status = ST_OK;
  }
  cta_sync();
  const int fp = f.pe - f.pb, ft = f.te - f.tb;
  if (ft == 0) {
    if (cx.tid == 0 && !runs_push(sh, ro, OP_D, (uint32_t)fp)) sh->status = ST_WS_RUNS_OVERFLOW;
  } else if (fp == 0) {
    if (cx.tid == 0 && !runs_push(sh, ro, OP_I, (uint32_t)ft)) sh->status = ST_WS_RUNS_OVERFLOW;
  } else {
    set_dirs(df, dr, db, sv, f);
    const int st = wf_base(cx, db, pen, ro);
    if (st != ST_OK && cx.tid == 0) sh->status = st;
  }
  cta_sync();
  res.status = sh->status;
  res.n_runs = sh->n_runs;
  return res;
}

WFA_FN Frame wf_root_frame(int plen, int tlen) {
  Frame f;
  f.pb = 0;
  f.pe = plen;
  f.tb = 0;
  f.te = tlen;
  f.cbegin = C_M;
  f.cend = C_M;
  // WFA/wavefront_bialign.c:1279-1283
  f.score_remaining = (imax(plen, tlen) <= 100) ? 0 : INT32_MAX;
  f.level = 0;
  return f;
}

WFA_FN PairResult wf_align_pair(Ctx& cx, const SeqView& sv, int plen, int tlen, const Penalties& pen,
                                const PairWorkspace& ws, const RunOut& ro, long long score_hint = INT64_MAX) {
  return wf_align_from(cx, sv, wf_root_frame(plen, tlen), pen, ws, ro, score_hint, nullptr, 0, 0);
}

}  // namespace srb
