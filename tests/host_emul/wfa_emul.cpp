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

// The split path of the device (align_kernel -> subalign_kernel bisect launch -> leaf launch -> assemble_kernel) with the
// same core functions: the pair is bisected down to `split_T`, every emitted task down to `leaf_T2`, every leaf goes
// through wf_align_leaf, and the pieces are joined in order (a piece whose first run continues the open run is merged
// into it). Small tables (32 ring slots, 32 stack frames) and a narrow ring for the tasks, like the one-warp kernel.
extern "C" int emul_wfa_align_split(const char* pattern, int plen, const char* text, int tlen, int x, int o1, int e1, int o2,
                                    int e2, int W, int cap, int lbits, int split_T, int leaf_T2, char* cigar_out,
                                    int* cigar_len, int* score, int* n_tasks_out, int* n_leaves_out) {
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
  const int subW = ((split_T + 2 * pen.scope + 128) + 63) & ~63;
  std::vector<int32_t> sf((size_t)(pen.scope + 1) * 5 * subW), sr((size_t)(pen.scope + 1) * 5 * subW);
  PairWorkspace ws2{sf.data(), sr.data(), subW, full.data(), fm.data(), cap};
  std::vector<uint32_t> runs(plen + tlen + 64), bt(plen + tlen + 64);
  RunOut ro{runs.data(), (uint32_t)runs.size(), bt.data(), (uint32_t)bt.size()};
  CtaShared* sh = new CtaShared();
  CtaTables<MAX_SCOPE, STACK_MAX>* big = new CtaTables<MAX_SCOPE, STACK_MAX>();
  CtaTables<32, 32>* small = new CtaTables<32, 32>();
  Ctx cx{sh, 0, 1, 0};
  int status = ST_OK, sc_out = INT32_MIN;
  std::vector<Frame> tasks, leaves;
  {
    cta_bind(sh, big);
    Frame* buf = reinterpret_cast<Frame*>(runs.data());
    const uint32_t cap_f = (uint32_t)(runs.size() * sizeof(uint32_t) / sizeof(Frame));
    PairResult r = wf_align_from(cx, sv, wf_root_frame(plen, tlen), pen, ws, ro, (long long)1 << 40, buf, cap_f, split_T);
    status = r.status;
    sc_out = r.score;
    if (status == ST_OK) tasks.assign(buf, buf + r.n_runs);
  }
  cta_bind(sh, small);
  for (size_t t = 0; t < tasks.size() && status == ST_OK; ++t) {
    Frame* buf = reinterpret_cast<Frame*>(runs.data());
    const uint32_t cap_f = (uint32_t)(runs.size() * sizeof(uint32_t) / sizeof(Frame));
    PairResult r = wf_align_from(cx, sv, tasks[t], pen, ws2, ro, (long long)1 << 40, buf, cap_f, leaf_T2);
    status = r.status;
    if (tasks[t].level == 0 && r.score != INT32_MIN) sc_out = r.score;
    if (status == ST_OK) leaves.insert(leaves.end(), buf, buf + r.n_runs);
  }
  std::vector<uint32_t> joined;
  for (size_t l = 0; l < leaves.size() && status == ST_OK; ++l) {
    PairResult r = wf_align_leaf(cx, sv, leaves[l], pen, ws2, ro);
    status = r.status;
    for (uint32_t i = 0; i < r.n_runs && status == ST_OK; ++i) {
      if (!joined.empty() && ((joined.back() ^ runs[i]) & 3u) == 0) joined.back() += (runs[i] >> 2) << 2;
      else joined.push_back(runs[i]);
    }
  }
  int n = 0;
  const char ops[4] = {'M', 'X', 'I', 'D'};
  if (status == ST_OK)
    for (uint32_t v : joined) {
      memset(cigar_out + n, ops[v & 3], v >> 2);
      n += (int)(v >> 2);
    }
  *cigar_len = n;
  *score = sc_out;
  *n_tasks_out = (int)tasks.size();
  *n_leaves_out = (int)leaves.size();
  delete small;
  delete big;
  delete sh;
  return status;
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
