// tests/host_emul/wfa_emul.cpp -- TEST-ONLY host build of seqrush_b200/csrc/wfa_core.cuh with a
// one-thread "CTA", so the control logic of the CUDA aligner can be checked against the compiled
// reference (oracle/_ref) in a container without a GPU. Never linked into the product library.
#define WFA_HOST_EMULATION 1
#include "../../seqrush_b200/csrc/wfa_core.cuh"
#include <stdlib.h>
#include <string.h>
#include <vector>
using namespace srb;

extern "C" int emul_wfa_align(const char* pattern, int plen, const char* text, int tlen, int x, int o1, int e1, int o2,
                              int e2, int W, int cap, char* cigar_out, int* cigar_len, int* score) {
  Penalties pen{x, o1, e1, o2, e2, 0};
  int sc = o1 + e1 > o2 + e2 ? o1 + e1 : o2 + e2;
  if (x > sc) sc = x;
  pen.scope = sc + 1;
  std::vector<uint8_t> pb(plen + 64, 0), tb(tlen + 64, 0);
  memcpy(pb.data() + 32, pattern, plen);
  memcpy(tb.data() + 32, text, tlen);
  SeqView sv{pb.data() + 32, tb.data() + 32};
  std::vector<int32_t> rf((size_t)pen.scope * 5 * W), rr((size_t)pen.scope * 5 * W), full((size_t)5 * cap * cap);
  std::vector<WfMeta> fm((size_t)cap * 5);
  PairWorkspace ws{rf.data(), rr.data(), W, full.data(), fm.data(), cap};
  std::vector<uint32_t> runs(plen + tlen + 8), bt(plen + tlen + 8);
  RunOut ro{runs.data(), (uint32_t)runs.size(), bt.data(), (uint32_t)bt.size()};
  CtaShared* sh = new CtaShared();
  Ctx cx{sh, 0, 1, 0};
  PairResult r = wf_align_pair(cx, sv, plen, tlen, pen, ws, ro);
  int n = 0;
  const char ops[4] = {'M', 'X', 'I', 'D'};
  if (r.status == ST_OK) {
    for (uint32_t i = 0; i < r.n_runs; ++i) {
      const uint32_t len = runs[i] >> 2;
      memset(cigar_out + n, ops[runs[i] & 3], len);
      n += len;
    }
  }
  *cigar_len = n;
  *score = r.score;
  delete sh;
  return r.status;
}

#include "../../seqrush_b200/csrc/orient_core.cuh"
extern "C" int emul_orient(const char* pf, const char* pr, int plen, const char* text, int tlen, int* dist) {
  std::vector<uint8_t> a(plen + 64, 0), b(plen + 64, 0), t(tlen + 64, 0);
  memcpy(a.data() + 32, pf, plen);
  memcpy(b.data() + 32, pr, plen);
  memcpy(t.data() + 32, text, tlen);
  const int stride = plen + tlen + 8;
  std::vector<int32_t> buf((size_t)4 * stride);
  CtaShared* sh = new CtaShared();
  Ctx cx{sh, 0, 1, 0};
  OrientResult r = wf_orient(cx, a.data() + 32, b.data() + 32, t.data() + 32, plen, tlen, buf.data(), stride);
  *dist = r.distance;
  delete sh;
  return r.is_reverse;
}
