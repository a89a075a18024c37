// oracle/ref_shim_edlib.cpp -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
// C entry point around the reference's vendored edlib (edlib_rs-0.1.2/edlib-c),
// configured exactly like allwave/src/alignment.rs:149-154 (k=-1, NW, DISTANCE).
#include "edlib.h"
extern "C" int ref_edlib_nw_distance(const char* q, int qlen, const char* t, int tlen) {
  EdlibAlignConfig cfg = edlibNewAlignConfig(-1, EDLIB_MODE_NW, EDLIB_TASK_DISTANCE, nullptr, 0);
  EdlibAlignResult r = edlibAlign(q, qlen, t, tlen, cfg);
  const int d = r.editDistance;
  edlibFreeAlignResult(r);
  return d;
}
