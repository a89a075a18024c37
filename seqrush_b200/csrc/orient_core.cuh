// orient_core.cuh -- strand choice for one pair, one CTA.
//
// Reference behaviour (allwave/src/alignment.rs:148-171): compute the exact global unit-cost
// edit distance of the query against the target (edlib NW, k=-1: edlib.cpp:141-212,730+) and of
// the reverse-complemented query against the target, and align forward iff fwd <= rev.
// Any exact edit-distance method yields the same boolean. Instead of two full O(L*d/64)
// bit-vector passes (the wrong strand costs d ~ 0.5 L), this runs unit-cost furthest-reaching
// wavefronts on BOTH strands in lock-step by distance and stops at the first strand to reach
// the end: forward wins ties. Cost O(min(d_fwd, d_rev)^2) cells.
#pragma once
#include "wfa_core.cuh"

namespace srb {

struct OrientResult {
  int is_reverse;
  int distance;  // distance of the winning strand (== min(fwd, rev))
};

// buf: 4 rows (strand x parity) of `stride` ints each; stride >= plen + tlen + 3.
WFA_FN OrientResult wf_orient(Ctx& cx, const uint32_t* Pf, const uint32_t* Pr, const uint32_t* T, int lbits, int plen,
                              int tlen, int32_t* buf, int stride) {
  OrientResult res;
  res.is_reverse = 0;
  res.distance = 0;
  const int kend = tlen - plen;
  const int center = plen + 1;  // row index of k: k + center
  cta_sync();
  if (cx.tid == 0) {
    cx.sh->ocells = 0;
    for (int strand = 0; strand < 2; ++strand) {
      const uint32_t* P = strand ? Pr : Pf;
      int32_t* row = buf + (size_t)(strand * 2 + 0) * stride + center;
      row[0] = lcp_fwd(P, T, 0, 0, imin(plen, tlen), lbits);
    }
  }
  cta_sync();
  int d = 0;
  while (true) {
    // end test (both strands), forward first
    const int par = d & 1;
    const int32_t* rowf = buf + (size_t)(0 * 2 + par) * stride + center;
    const int32_t* rowr = buf + (size_t)(1 * 2 + par) * stride + center;
    if (kend >= -d && kend <= d) {
      if (rowf[kend] >= tlen) {
        res.is_reverse = 0;
        res.distance = d;
        return res;
      }
      if (rowr[kend] >= tlen) {
        res.is_reverse = 1;
        res.distance = d;
        return res;
      }
    }
    // next distance
    const int plo = imax(-d, -plen), phi = imin(d, tlen);
    ++d;
    const int lo = imax(-d, -plen), hi = imin(d, tlen);
    const int width = hi - lo + 1;
    if (cx.tid == 0) cx.sh->ocells += 2ull * (unsigned long long)width;
    for (int idx = cx.tid; idx < 2 * width; idx += cx.nt) {
      const int strand = idx >= width;
      const int k = lo + (strand ? idx - width : idx);
      const uint32_t* P = strand ? Pr : Pf;
      const int32_t* prev = buf + (size_t)(strand * 2 + par) * stride + center;
      int32_t* cur = buf + (size_t)(strand * 2 + (par ^ 1)) * stride + center;
      int32_t best = OFFSET_NULL;
      if (k - 1 >= plo && k - 1 <= phi) {  // gap in the pattern (consumes text)
        const int32_t h = prev[k - 1] + 1;
        if (h >= 0 && h <= tlen && h - k >= 0 && h - k <= plen) best = h;
      }
      if (k >= plo && k <= phi) {  // substitution
        const int32_t h = prev[k] + 1;
        if (h >= 0 && h <= tlen && h - k >= 0 && h - k <= plen) best = imax(best, h);
      }
      if (k + 1 >= plo && k + 1 <= phi) {  // gap in the text (consumes pattern)
        const int32_t h = prev[k + 1];
        if (h >= 0 && h <= tlen && h - k >= 0 && h - k <= plen) best = imax(best, h);
      }
      if (best != OFFSET_NULL) {
        const int v = best - k;
        best += lcp_fwd(P, T, v, best, imin(plen - v, tlen - best), lbits);
      }
      cur[k] = best;
    }
    cta_sync();
  }
}

}  // namespace srb
