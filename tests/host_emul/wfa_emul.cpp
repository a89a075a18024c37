// tests/host_emul/wfa_emul.cpp -- TEST-ONLY host build of seqrush_b200/csrc/wfa_core.cuh with a
// one-thread "CTA", so the control logic of the CUDA aligner can be checked against the compiled
// reference (oracle/_ref) in a container without a GPU. Never linked into the product library.
#define WFA_HOST_EMULATION 1
#include "../../seqrush_b200/csrc/wfa_core.cuh"
#include <stdlib.h>
#include <string.h>
#include <vector>
using namespace srb;

// pack bytes at (1 << lbits) bits per symbol with a code table built from the bytes seen
// (same layout as pack_kernel in kernels.cuh): returns the word array with 4 guard words each side
struct Packer {
  int code[256];
  int n = 0;
  Packer() { for (int i = 0; i < 256; ++i) code[i] = -1; }
  void see(const char* s, int len) {
    for (int i = 0; i < len; ++i) {
      const unsigned char c = (unsigned char)s[i];
      if (code[c] < 0) code[c] = n++;
    }
  }
  int min_lbits() const { return n <= 4 ? 1 : n <= 16 ? 2 : 3; }
  std::vector<uint32_t> pack(const char* s, int len, int lbits) const {
    const int bits = 1 << lbits;
    std::vector<uint32_t> w(((size_t)len * bits + 31) / 32 + 8, 0u);
    for (int i = 0; i < len; ++i) {
      const uint32_t c = lbits == 3 ? (unsigned char)s[i] : (uint32_t)code[(unsigned char)s[i]];
      const size_t bit = (size_t)i * bits;
      w[4 + bit / 32] |= c << (bit % 32);
    }
    return w;
  }
};

extern "C" int emul_wfa_align(const char* pattern, int plen, const char* text, int tlen, int x, int o1, int e1, int o2,
                              int e2, int W, int cap, int lbits, char* cigar_out, int* cigar_len, int* score) {
  Penalties pen{x, o1, e1, o2, e2, 0};
  int sc = o1 + e1 > o2 + e2 ? o1 + e1 : o2 + e2;
  if (x > sc) sc = x;
  pen.scope = sc + 1;
  Packer pk;
  pk.see(pattern, plen);
  pk.see(text, tlen);
  if (lbits <= 0 || lbits < pk.min_lbits()) lbits = pk.min_lbits();
  std::vector<uint32_t> pb = pk.pack(pattern, plen, lbits), tb = pk.pack(text, tlen, lbits);
  SeqView sv{pb.data() + 4, tb.data() + 4, lbits};
  std::vector<int32_t> rf((size_t)(pen.scope + 1) * 5 * W), rr((size_t)(pen.scope + 1) * 5 * W), full((size_t)5 * cap * cap);
  std::vector<WfMeta> fm((size_t)cap * 5);
  PairWorkspace ws{rf.data(), rr.data(), W, full.data(), fm.data(), cap};
  std::vector<uint32_t> runs(plen + tlen + 8), bt(plen + tlen + 8);
  RunOut ro{runs.data(), (uint32_t)runs.size(), bt.data(), (uint32_t)bt.size()};
  CtaShared* sh = new CtaShared();
  CtaTables<MAX_SCOPE, STACK_MAX>* tb_ = new CtaTables<MAX_SCOPE, STACK_MAX>();
  cta_bind(sh, tb_);
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
  delete tb_;
  delete sh;
  return r.status;
}

#include "../../seqrush_b200/csrc/orient_core.cuh"
extern "C" int emul_orient(const char* pf, const char* pr, int plen, const char* text, int tlen, int lbits, int* dist) {
  Packer pk;
  pk.see(pf, plen);
  pk.see(pr, plen);
  pk.see(text, tlen);
  if (lbits <= 0 || lbits < pk.min_lbits()) lbits = pk.min_lbits();
  std::vector<uint32_t> a = pk.pack(pf, plen, lbits), b = pk.pack(pr, plen, lbits), t = pk.pack(text, tlen, lbits);
  const int stride = plen + tlen + 8;
  std::vector<int32_t> buf((size_t)4 * stride);
  CtaShared* sh = new CtaShared();
  Ctx cx{sh, 0, 1, 0};
  OrientResult r = wf_orient(cx, a.data() + 4, b.data() + 4, t.data() + 4, lbits, plen, tlen, buf.data(), stride);
  *dist = r.distance;
  delete sh;
  return r.is_reverse;
}
