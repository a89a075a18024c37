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
};

struct Penalties {
  int x, o1, e1, o2, e2;
  int scope;  // max_score_scope, WFA/wavefront_components.c:101-113
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

// Sequence view: raw bytes (any alphabet; byte equality like the reference,
// WFA/wavefront_extend_kernels.c:68-92). Buffers are padded by >= 8 bytes on
// both sides so that word loads never leave the allocation.
struct SeqView {
  const uint8_t* p;  // pattern (query, possibly reverse-complemented)
  const uint8_t* t;  // text (target)
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
  int modular;
  int scope;
  int W;    // modular: diagonals per row
  int cap;  // full: max score + 1
  WFA_FN int slot(int s) const { return modular ? (s % scope) : s; }
  WFA_FN int32_t* row(int s, int c) const {
    if (modular) return base + ((size_t)((s % scope) * 5 + c)) * (size_t)W + (W >> 1);
    return base + (size_t)c * (size_t)cap * (size_t)cap + (size_t)s * (size_t)s + s;
  }
  WFA_FN WfMeta get(int s, int c) const {
    WfMeta m;
    if (s < 0) {
      m.lo = 1;
      m.hi = -1;
      return m;
    }
    m = meta[slot(s) * 5 + c];
    if (m.lo > m.hi) {
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
  WfMeta meta_f[MAX_SCOPE * 5];
  WfMeta meta_r[MAX_SCOPE * 5];
  RedSlots red[2];
  int ovl[MAX_SCOPE * 5];  // overlap search results (min k_0 per task)
  Frame stack[STACK_MAX];
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
};

struct Ctx {
  CtaShared* sh;
  int tid, nt;
  int parity;
};

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
WFA_FN uint32_t load_u32_unaligned(const uint8_t* a) {
#if WFA_DEVICE
  const uintptr_t ad = (uintptr_t)a;
  const uint32_t* w = (const uint32_t*)(ad & ~(uintptr_t)3);
  const uint32_t lo = w[0], hi = w[1];
  return __funnelshift_r(lo, hi, (uint32_t)(ad & 3) * 8);
#else
  uint32_t v;
  memcpy(&v, a, 4);
  return v;
#endif
}

WFA_FN int lcp_fwd(const uint8_t* p, const uint8_t* t, int maxlen) {
  // p,t point at the first bytes to compare
  int n = 0;
  if (maxlen <= 0) return 0;
  if (p[0] != t[0]) return 0;
  while (n < maxlen) {
    const uint32_t x = load_u32_unaligned(p + n) ^ load_u32_unaligned(t + n);
    if (x) {
#if WFA_DEVICE
      n += (__ffs((int)x) - 1) >> 3;
#else
      n += __builtin_ctz(x) >> 3;
#endif
      break;
    }
    n += 4;
  }
  return n < maxlen ? n : maxlen;
}
WFA_FN int lcp_rev(const uint8_t* p, const uint8_t* t, int maxlen) {
  // p,t point at the first (highest-address) bytes to compare; walk downwards
  int n = 0;
  if (maxlen <= 0) return 0;
  if (p[0] != t[0]) return 0;
  while (n < maxlen) {
    const uint32_t x = load_u32_unaligned(p - n - 3) ^ load_u32_unaligned(t - n - 3);
    if (x) {
#if WFA_DEVICE
      n += __clz((int)x) >> 3;
#else
      n += __builtin_clz(x) >> 3;
#endif
      break;
    }
    n += 4;
  }
  return n < maxlen ? n : maxlen;
}

// One aligner direction over the current sub-problem.
struct Dir {
  WfStore st;
  const uint8_t* P;  // forward: &pattern[pb]; reverse: &pattern[pe-1]
  const uint8_t* T;
  int plen, tlen;
  int reverse;
  int cbegin, cend;
};

WFA_FN int32_t rd(const int32_t* row, WfMeta m, int k) { return (k >= m.lo && k <= m.hi) ? row[k] : OFFSET_NULL; }

// ---------------------------------------------------------------------------
// Initial wavefront (WFA/wavefront_aligner.c:252-383, end-to-end form only).
// Executed by thread 0; caller syncs.
// ---------------------------------------------------------------------------
WFA_FN void wf_init(const Dir& d) {
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
  if (!d.reverse) return off + lcp_fwd(d.P + v, d.T + h, room);
  return off + lcp_rev(d.P - v, d.T - h, room);
}

// score-0 extension: a single cell at most (k=0), done by thread 0.
WFA_FN StepOut wf_extend0(Ctx& cx, const Dir& d) {
  StepOut o;
  o.null_step = 0;
  o.end_reached = 0;
  o.max_ak = 0;
  o.overflow = 0;
  cta_sync();
  if (cx.tid == 0) {
    const WfMeta m = d.st.get(0, C_M);
    int ak = 0;
    if (m.lo <= m.hi) {
      int32_t* row = d.st.row(0, C_M);
      const int32_t e = wf_extend_cell(d, 0, row[0]);
      row[0] = e;
      ak = 2 * e;
    }
    cx.sh->bcast[0] = ak;
  }
  cta_sync();
  o.max_ak = cx.sh->bcast[0];
  o.end_reached = wf_end_reached(d, 0) ? 1 : 0;
  return o;
}

// ---------------------------------------------------------------------------
// One score step: next wavefront (WFA/wavefront_compute_affine2p.c:45-106,
// 334-368; limits WFA/wavefront_compute.c:40-86; trimming :579-632) fused with
// the exact-match extension of the new M wavefront (WFA/wavefront_extend.c:
// 86-121,173-211) and the end test. Diagonals are distributed over the CTA's
// threads; per-component trims and the max antidiagonal are warp-reduced and
// merged with shared-memory atomics.
// ---------------------------------------------------------------------------
WFA_FN StepOut wf_step(Ctx& cx, const Dir& d, const Penalties& pen, int s) {
  StepOut o;
  o.null_step = 0;
  o.end_reached = 0;
  o.max_ak = 0;
  o.overflow = 0;
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
  int mn0 = INT32_MAX, mn1 = INT32_MAX, mn2 = INT32_MAX, mn3 = INT32_MAX, mn4 = INT32_MAX;
  int mx0 = INT32_MIN, mx1 = INT32_MIN, mx2 = INT32_MIN, mx3 = INT32_MIN, mx4 = INT32_MIN;
  int max_ak = 0;
  for (int k = lo + cx.tid; k <= hi; k += cx.nt) {
    const int32_t o1m = rd(r_o1, m_o1, k - 1), o1p = rd(r_o1, m_o1, k + 1);
    const int32_t o2m = rd(r_o2, m_o2, k - 1), o2p = rd(r_o2, m_o2, k + 1);
    const int32_t ins1 = imax(o1m, rd(r_i1, m_i1, k - 1)) + 1;
    const int32_t ins2 = imax(o2m, rd(r_i2, m_i2, k - 1)) + 1;
    const int32_t del1 = imax(o1p, rd(r_d1, m_d1, k + 1));
    const int32_t del2 = imax(o2p, rd(r_d2, m_d2, k + 1));
    const int32_t mis = rd(r_mis, m_mis, k) + 1;
    int32_t mx = imax(imax(del1, del2), imax(mis, imax(ins1, ins2)));
    if ((uint32_t)mx > tlen || (uint32_t)(mx - k) > plen) mx = OFFSET_NULL;
    if (a_i1) {
      w_i1[k] = ins1;
      if ((uint32_t)ins1 <= tlen && (uint32_t)(ins1 - k) <= plen) {
        mn1 = imin(mn1, k);
        mx1 = k;
      }
    }
    if (a_i2) {
      w_i2[k] = ins2;
      if ((uint32_t)ins2 <= tlen && (uint32_t)(ins2 - k) <= plen) {
        mn2 = imin(mn2, k);
        mx2 = k;
      }
    }
    if (a_d1) {
      w_d1[k] = del1;
      if ((uint32_t)del1 <= tlen && (uint32_t)(del1 - k) <= plen) {
        mn3 = imin(mn3, k);
        mx3 = k;
      }
    }
    if (a_d2) {
      w_d2[k] = del2;
      if ((uint32_t)del2 <= tlen && (uint32_t)(del2 - k) <= plen) {
        mn4 = imin(mn4, k);
        mx4 = k;
      }
    }
    if (mx != OFFSET_NULL) {
      mn0 = imin(mn0, k);
      mx0 = k;
      mx = wf_extend_cell(d, k, mx);
      max_ak = imax(max_ak, 2 * mx - k);
    }
    w_m[k] = mx;
  }
  // merge per-thread trims (WFA/wavefront_compute.c:579-611 does the same with two linear scans)
  RedSlots* red = &cx.sh->red[cx.parity];
  {
    int v;
    v = warp_min(mn0);
    if (lane_id() == 0 && v != INT32_MAX) sh_atomic_min(&red->mn[0], v);
    v = warp_max(mx0);
    if (lane_id() == 0 && v != INT32_MIN) sh_atomic_max(&red->mx[0], v);
    v = warp_max(max_ak);
    if (lane_id() == 0 && v > 0) sh_atomic_max(&red->max_ak, v);
    if (a_i1) {
      v = warp_min(mn1);
      if (lane_id() == 0 && v != INT32_MAX) sh_atomic_min(&red->mn[1], v);
      v = warp_max(mx1);
      if (lane_id() == 0 && v != INT32_MIN) sh_atomic_max(&red->mx[1], v);
    }
    if (a_i2) {
      v = warp_min(mn2);
      if (lane_id() == 0 && v != INT32_MAX) sh_atomic_min(&red->mn[2], v);
      v = warp_max(mx2);
      if (lane_id() == 0 && v != INT32_MIN) sh_atomic_max(&red->mx[2], v);
    }
    if (a_d1) {
      v = warp_min(mn3);
      if (lane_id() == 0 && v != INT32_MAX) sh_atomic_min(&red->mn[3], v);
      v = warp_max(mx3);
      if (lane_id() == 0 && v != INT32_MIN) sh_atomic_max(&red->mx[3], v);
    }
    if (a_d2) {
      v = warp_min(mn4);
      if (lane_id() == 0 && v != INT32_MAX) sh_atomic_min(&red->mn[4], v);
      v = warp_max(mx4);
      if (lane_id() == 0 && v != INT32_MIN) sh_atomic_max(&red->mx[4], v);
    }
  }
  cta_sync();
  // publish metas; reset the other reduction buffer for the next step
  const bool alloc[5] = {true, a_i1, a_i2, a_d1, a_d2};
  o.max_ak = red->max_ak;
  if (cx.tid == 0) {
    for (int c = 0; c < 5; ++c) {
      WfMeta m;
      m.lo = 1;
      m.hi = -1;
      if (alloc[c]) {
        if (red->mn[c] <= red->mx[c]) {
          m.lo = red->mn[c];
          m.hi = red->mx[c];
        } else {
          m.lo = 2;
          m.hi = -2;
        }
      }
      st.meta[slot5 + c] = m;
    }
    RedSlots* nx = &cx.sh->red[cx.parity ^ 1];
    for (int c = 0; c < 5; ++c) {
      nx->mn[c] = INT32_MAX;
      nx->mx[c] = INT32_MIN;
    }
    nx->max_ak = 0;
  }
  // termination (WFA/wavefront_termination.c:37-113): read directly from the reduction slots / rows
  {
    const int ak = d.tlen - d.plen;
    const int ce = d.cend;
    const bool present = alloc[ce] && red->mn[ce] <= red->mx[ce];
    if (present && red->mn[ce] <= ak && ak <= red->mx[ce]) {
      if (st.row(s, ce)[ak] >= d.tlen) o.end_reached = 1;
    }
  }
  cx.parity ^= 1;
  return o;
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
  for (int task = warp; task < ntasks; task += nwarps) {
    const int i = task / 5, j = task - i * 5;
    const int c = ovl_comp_of_task(j);
    const int score_i = score_1 - i;
    int best = INT32_MAX;
    bool run = score_i >= 0;
    const int gopen = (c == C_M) ? 0 : ((c == C_I1 || c == C_D1) ? pen.o1 : pen.o2);
    if (run && score_0 + score_i - gopen >= bp_score_in) run = false;
    if (run) {
      const WfMeta a = d0.st.get(score_0, c);
      const WfMeta b = d1.st.get(score_i, c);
      if (a.lo <= a.hi && b.lo <= b.hi) {
        const int lo_1 = kinv - b.hi, hi_1 = kinv - b.lo;
        const int max_lo = imax(a.lo, lo_1), min_hi = imin(a.hi, hi_1);
        const int32_t* r0 = d0.st.row(score_0, c);
        const int32_t* r1 = d1.st.row(score_i, c);
        for (int k0 = max_lo + lane; k0 <= min_hi; k0 += lanes) {
          const int k1 = kinv - k0;
          const int32_t h0 = r0[k0], h1 = r1[k1];
          if (h0 + h1 >= tlen) {
            bool ok = true;
            if (c != C_M) {
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
      }
    }
    best = warp_min(best);
    if (lane == 0) cx.sh->ovl[task] = best;
  }
  cta_sync();
  if (cx.tid == 0 && m0_present) {
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
//   2 = workspace overflow.
// ---------------------------------------------------------------------------
WFA_FN int wf_find_breakpoint(Ctx& cx, Dir& df, Dir& dr, const Penalties& pen, int* score_reached,
                              unsigned long long* cells) {
  (void)cells;
  cta_sync();
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
  bool last_wf_forward = false;
  while (true) {
    if (fwd_max_ak + rev_max_ak >= max_antidiagonal) break;
    ++score_f;
    o = wf_step(cx, df, pen, score_f);
    if (o.overflow) return 2;
    if (fwd_max_ak < o.max_ak) fwd_max_ak = o.max_ak;
    last_wf_forward = true;
    if (o.end_reached) {
      *score_reached = score_f;
      return 1;
    }
    if (fwd_max_ak + rev_max_ak >= max_antidiagonal) break;
    ++score_r;
    o = wf_step(cx, dr, pen, score_r);
    if (o.overflow) return 2;
    if (rev_max_ak < o.max_ak) rev_max_ak = o.max_ak;
    last_wf_forward = false;
    if (o.end_reached) {
      *score_reached = score_r;
      return 1;
    }
  }
  const int scope = pen.scope;
  const int gap_opening = imax(pen.o1, pen.o2);
  while (true) {
    if (last_wf_forward) {
      const int min_score_r = (score_r > scope - 1) ? score_r - (scope - 1) : 0;
      cta_sync();
      if (score_f + min_score_r - gap_opening >= cx.sh->bp.score) break;
      wf_overlap(cx, df, dr, pen, score_f, score_r, true);
      ++score_r;
      o = wf_step(cx, dr, pen, score_r);
      if (o.overflow) return 2;
      if (o.end_reached) {
        *score_reached = score_r;
        return 1;
      }
    }
    const int min_score_f = (score_f > scope - 1) ? score_f - (scope - 1) : 0;
    cta_sync();
    if (min_score_f + score_r - gap_opening >= cx.sh->bp.score) break;
    wf_overlap(cx, dr, df, pen, score_r, score_f, false);
    ++score_f;
    o = wf_step(cx, df, pen, score_f);
    if (o.overflow) return 2;
    if (o.end_reached) {
      *score_reached = score_f;
      return 1;
    }
    last_wf_forward = true;
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
    const bool ok = wf_backtrace(cx.sh, ro, db.st, pen, db.plen, db.tlen, db.cbegin, db.cend, score);
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
  df.P = sv.p + f.pb;
  df.T = sv.t + f.tb;
  df.plen = plen;
  df.tlen = tlen;
  df.reverse = 0;
  df.cbegin = f.cbegin;
  df.cend = f.cend;
  dr.P = sv.p + f.pe - 1;
  dr.T = sv.t + f.te - 1;
  dr.plen = plen;
  dr.tlen = tlen;
  dr.reverse = 1;
  dr.cbegin = f.cend;
  dr.cend = f.cbegin;
  db.P = df.P;
  db.T = df.T;
  db.plen = plen;
  db.tlen = tlen;
  db.reverse = 0;
  db.cbegin = f.cbegin;
  db.cend = f.cend;
}

WFA_FN PairResult wf_align_pair(Ctx& cx, const SeqView& sv, int plen, int tlen, const Penalties& pen,
                                const PairWorkspace& ws, const RunOut& ro) {
  CtaShared* sh = cx.sh;
  PairResult res;
  res.status = ST_OK;
  res.score = INT32_MIN;
  res.n_runs = 0;
  Dir df, dr, db;
  df.st.base = ws.ring_f;
  df.st.meta = sh->meta_f;
  df.st.modular = 1;
  df.st.scope = pen.scope;
  df.st.W = ws.W;
  df.st.cap = 0;
  dr.st = df.st;
  dr.st.base = ws.ring_r;
  dr.st.meta = sh->meta_r;
  db.st.base = ws.full;
  db.st.meta = ws.full_meta;
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
    sh->stack[sh->sp++] = f;
  }
  cta_sync();
  int level0_score = INT32_MIN;
  while (true) {
    cta_sync();
    if (sh->sp == 0 || sh->status != ST_OK) break;
    const Frame f = sh->stack[sh->sp - 1];
    cta_sync();
    if (cx.tid == 0) sh->sp--;
    const int fp = f.pe - f.pb, ft = f.te - f.tb;
    int st = ST_OK;
    if (ft == 0) {
      if (cx.tid == 0 && !runs_push(sh, ro, OP_D, (uint32_t)fp)) sh->status = ST_WS_RUNS_OVERFLOW;
      continue;
    } else if (fp == 0) {
      if (cx.tid == 0 && !runs_push(sh, ro, OP_I, (uint32_t)ft)) sh->status = ST_WS_RUNS_OVERFLOW;
      continue;
    }
    set_dirs(df, dr, db, sv, f);
    if (f.score_remaining <= 250) {  // WF_BIALIGN_FALLBACK_MIN_SCORE
      st = wf_base(cx, db, pen, ro);
      if (st != ST_OK && cx.tid == 0) sh->status = st;
      continue;
    }
    int score_reached = 0;
    const int r = wf_find_breakpoint(cx, df, dr, pen, &score_reached, nullptr);
    if (r == 2) {
      if (cx.tid == 0) sh->status = ST_WS_DIAG_OVERFLOW;
      continue;
    }
    if (r == 1) {
      // WFA/wavefront_bialign.c:1083-1112: recover through the base case when cheap enough
      if (score_reached <= 500) {  // WF_BIALIGN_RECOVERY_MIN_SCORE
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
    if (f.level == 0) level0_score = -bp.score;  // WFA/wavefront_bialign.c:1214-1219, match == 0
    cta_sync();
    if (cx.tid == 0) {
      if (sh->sp + 2 > STACK_MAX) {
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

}  // namespace srb
