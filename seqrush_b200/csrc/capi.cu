// capi.cu -- host side of the C ABI declared in include/seqrush_b200.h.
// Owns device memory, the stream, per-CTA workspaces and the chunked launch loop
// (align kernel -> union kernel per chunk of pairs). There is no CPU fallback: without a CUDA
// device every entry point fails with SRGPU_E_CUDA.
#include "../../include/seqrush_b200.h"
#include "kernels.cuh"
#include "graph.cuh"
#include "graph_host.hpp"
#include "sparsify.cuh"

#include <algorithm>
#include <functional>
#include <mutex>
#include <string>
#include <vector>
#include <cstdio>
#include <cstdlib>
#include <cstring>

using namespace srb;

static thread_local std::string g_err;
extern "C" const char* srgpu_last_error(void) { return g_err.c_str(); }

#define CK(call)                                                                                   \
  do {                                                                                             \
    cudaError_t e_ = (call);                                                                       \
    if (e_ != cudaSuccess) {                                                                       \
      char buf_[512];                                                                              \
      snprintf(buf_, sizeof buf_, "%s:%d: %s -> %s", __FILE__, __LINE__, #call, cudaGetErrorString(e_)); \
      g_err = buf_;                                                                                \
      return SRGPU_E_CUDA;                                                                         \
    }                                                                                              \
  } while (0)

#define UF_ALIVE(u)                                                        \
  do {                                                                     \
    if (!(u)->ctx) {                                                       \
      g_err = "union-find handle outlived its context (already destroyed)"; \
      return SRGPU_E_ARG;                                                  \
    }                                                                      \
  } while (0)

template <class T>
struct DevBuf {
  T* p = nullptr;
  size_t n = 0;
  int ensure(size_t want) {
    if (want <= n && p) return 0;
    if (p) cudaFree(p);
    p = nullptr;
    n = 0;
    if (want == 0) want = 1;
    cudaError_t e = cudaMalloc((void**)&p, want * sizeof(T));
    if (e != cudaSuccess) {
      g_err = std::string("cudaMalloc failed: ") + cudaGetErrorString(e);
      p = nullptr;
      return SRGPU_E_CUDA;
    }
    n = want;
    return 0;
  }
  void release() {
    if (p) cudaFree(p);
    p = nullptr;
    n = 0;
  }
};

struct srgpu_ctx {
  int device = 0;
  cudaStream_t stream = nullptr;
  bool own_stream = false;
  int n_sm = 148;
  std::mutex mu;
  // config
  int threads = 128;
  int ctas_per_sm = 4;
  bool auto_shape = true;  // launch shape not set by srgpu_ctx_configure: chosen from the sequence length
  int ring_diagonals = 0;
  // sequences
  uint32_t n_seqs = 0;
  uint64_t seq_identity = 0;  // fingerprint of the loaded set (count, lengths, head and tail bytes of every sequence)
  uint64_t total_len = 0;
  uint32_t max_len = 0;
  DevBuf<uint8_t> d_fwd, d_rcsafe, d_rcinv;
  DevBuf<uint64_t> d_byte_off, d_pos_off, d_pk_off;
  DevBuf<uint32_t> d_pk_fwd, d_pk_rc;
  int lbits = 3;              // log2(bits per packed symbol): 1, 2 or 3
  uint32_t max_pk_words = 0;  // packed words of the longest sequence incl. guards
  int smem_budget = 96 * 1024;  // dynamic shared memory a kernel may use for staged sequences
  int orient_ctas_per_sm = 6;   // SRGPU_ORIENT_CTAS (tuning; c2: 4 -> 36 ms, 6 -> 27 ms, 8 -> 27 ms, profiles/r02o_*)
  int stage_seqs = -1;          // align_kernel: stage the two sequences (-1: when the CTAs per SM are kept; SRGPU_STAGE_SEQS=0/1)
  DevBuf<uint32_t> d_len;
  std::vector<uint32_t> h_len;
  std::vector<uint64_t> h_pos_off;
  // workspaces
  DevBuf<int32_t> d_ring, d_full, d_orient;
  DevBuf<WfMeta> d_full_meta;
  DevBuf<uint32_t> d_runs_tmp, d_bt_tmp;
  DevBuf<uint2> d_pairs;
  DevBuf<PairOut> d_out;
  DevBuf<uint32_t> d_arena, d_todo;
  DevBuf<unsigned long long> d_scalars;  // [0] arena_head, [1] next_pair(uint), [2] max edit distance, [3] task_count, [4] next_task, [5] etask_count, [6] next_task of the second pass
  // split mode: tasks (sub-problems) and the workspaces of the one-warp kernel that solves them
  int sub_ctas_per_sm = SUB_CTAS;  // one-warp CTAs per SM of subalign_kernel (SRGPU_SUB_CTAS lowers it). 1500-pair c2 slice:
                                   // 13 -> 231.6 ms, 16 -> 210.7 ms; a 96-register build at 20 -> 216.5 ms (profiles/r03a_*): the
                                   // kernel is bound by instruction fetch (profiles/r02z_*), not by warps in flight
  int leaf_ctas_per_sm = LEAF_CTAS;  // SRGPU_LEAF_CTAS
  int debug_sync = 0;  // SRGPU_DEBUG_SYNC: synchronise after every alignment launch
  int test_sub_w = 0;  // SRGPU_TEST_SUB_W (tests only): ring width of the one-warp kernel
  int test_asm_max = 0;  // SRGPU_TEST_ASM_MAX (tests only): pieces per pair the join kernel accepts
  int test_task_cap = 0, test_leaf_cap = 0;  // SRGPU_TEST_TASK_CAP / SRGPU_TEST_LEAF_CAP (tests only): tiny task arrays
  int sub_two_pass = 1;  // one-warp kernel in two launches: bisect down to the base-case bound, then the leaves (SRGPU_TWO_PASS=0/1)
  int split_T = 1000;  // sub-problems whose score_remaining is at most this go to subalign_kernel (0: monolithic); SRGPU_SPLIT_T.
                       // 1500-pair slices: c2 840 ms monolithic, 728 / 740 / 766 / 831 at T = 500 / 1000 / 1900 / 3900; 100 kb 2837 -> 2756 / 2737 / 2717
                       // (profiles/r02t_*, r02u_*, r02v_*)
  DevBuf<Frame> d_tasks;
  DevBuf<uint32_t> d_task_pair;
  DevBuf<TaskOut> d_task_out;
  DevBuf<SplitRec> d_split;
  DevBuf<unsigned long long> d_mirror_keys;  // strand choice once per unordered pair: hash table (q, t) -> pair slot
  DevBuf<uint32_t> d_mirror_vals, d_mirror_of;
  int mirror_orient = 1;         // SRGPU_MIRROR_ORIENT=0 computes both orders
  DevBuf<Frame> d_tasks2;        // two-pass mode of the one-warp kernel: second-level tasks and their records
  DevBuf<uint32_t> d_task_pair2;
  DevBuf<TaskOut> d_task_out2;
  DevBuf<SplitRec> d_esplit;
  DevBuf<int32_t> d_ring2, d_full2;
  DevBuf<WfMeta> d_full_meta2;
  DevBuf<uint32_t> d_runs_tmp2, d_bt_tmp2;
  DevBuf<unsigned long long> d_byteoff;
  DevBuf<uint8_t> d_bytes;
  DevBuf<PafRecordDev> d_paf_recs;
  DevBuf<unsigned long long> d_paf_runs;
  cudaEvent_t ev[10] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
  srgpu_stats stats{};
  std::vector<srgpu_uf*> ufs;  // union-finds living on this context (orphaned when it is destroyed)
};
static std::mutex g_registry_mu;

struct srgpu_uf {
  srgpu_ctx* ctx = nullptr;
  uint64_t capacity = 0;
  uint64_t seq_identity = 0;  // sequence set the forest was first united against (0: none yet); see check_uf
  DevBuf<uint32_t> parent;
  DevBuf<unsigned long long> counters;  // [0] ops [1] links [2] validate [3] range
  DevBuf<RegionDev> regions;
  DevBuf<uint64_t> tmp64;
  DevBuf<uint32_t> bitmap;
};

static void orphan_uf(srgpu_uf* u) {
  u->parent.release();
  u->counters.release();
  u->regions.release();
  u->tmp64.release();
  u->bitmap.release();
  u->ctx = nullptr;
  u->capacity = 0;
}

static inline int grid_for(uint64_t n, int threads, int n_sm) {
  uint64_t g = (n + threads - 1) / threads;
  uint64_t cap = (uint64_t)n_sm * 16;
  if (g > cap) g = cap;
  if (g < 1) g = 1;
  return (int)g;
}

extern "C" int srgpu_ctx_create(int device, void* stream, srgpu_ctx** out) {
  if (!out) return SRGPU_E_ARG;
  int count = 0;
  cudaError_t e = cudaGetDeviceCount(&count);
  if (e != cudaSuccess || count == 0) {
    g_err = std::string("no CUDA device: ") + cudaGetErrorString(e);
    return SRGPU_E_CUDA;
  }
  CK(cudaSetDevice(device));
  srgpu_ctx* c = new srgpu_ctx();
  c->device = device;
  cudaDeviceProp prop;
  CK(cudaGetDeviceProperties(&prop, device));
  c->n_sm = prop.multiProcessorCount;
  if (stream) {
    c->stream = (cudaStream_t)stream;
  } else {
    CK(cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking));
    c->own_stream = true;
  }
  for (int i = 0; i < 10; ++i) CK(cudaEventCreate(&c->ev[i]));
  if (const char* e = getenv("SRGPU_SMEM_BUDGET")) c->smem_budget = atoi(e);  // tuning: 0 = never stage sequences
  if (const char* e = getenv("SRGPU_SPLIT_T")) c->split_T = atoi(e);
  if (const char* e = getenv("SRGPU_ORIENT_CTAS")) c->orient_ctas_per_sm = std::max(1, atoi(e));
  if (const char* e = getenv("SRGPU_SUB_CTAS")) c->sub_ctas_per_sm = std::max(1, atoi(e));
  if (const char* e = getenv("SRGPU_TWO_PASS")) c->sub_two_pass = atoi(e) != 0;
  if (const char* e = getenv("SRGPU_TEST_SUB_W")) c->test_sub_w = std::max(0, atoi(e));
  if (const char* e = getenv("SRGPU_DEBUG_SYNC")) c->debug_sync = atoi(e) != 0;
  if (const char* e = getenv("SRGPU_TEST_ASM_MAX")) c->test_asm_max = std::max(0, atoi(e));
  if (const char* e = getenv("SRGPU_TEST_TASK_CAP")) c->test_task_cap = std::max(0, atoi(e));
  if (const char* e = getenv("SRGPU_TEST_LEAF_CAP")) c->test_leaf_cap = std::max(0, atoi(e));
  if (const char* e = getenv("SRGPU_MIRROR_ORIENT")) c->mirror_orient = atoi(e) != 0;
  if (const char* e = getenv("SRGPU_LEAF_CTAS")) c->leaf_ctas_per_sm = std::max(1, atoi(e));
  if (const char* e = getenv("SRGPU_STAGE_SEQS")) c->stage_seqs = atoi(e) != 0 ? 1 : 0;
  *out = c;
  return SRGPU_OK;
}

extern "C" void srgpu_ctx_destroy(srgpu_ctx* c) {
  if (!c) return;
  cudaSetDevice(c->device);
  cudaStreamSynchronize(c->stream);
  {  // a union-find that outlives its context keeps a valid handle but no device memory
    std::lock_guard<std::mutex> lk(g_registry_mu);
    for (srgpu_uf* u : c->ufs) orphan_uf(u);
    c->ufs.clear();
  }
  c->d_fwd.release(); c->d_pk_fwd.release(); c->d_pk_rc.release(); c->d_pk_off.release(); c->d_rcsafe.release(); c->d_rcinv.release(); c->d_mirror_keys.release(); c->d_mirror_vals.release(); c->d_mirror_of.release(); c->d_byte_off.release(); c->d_pos_off.release();
  c->d_len.release(); c->d_ring.release(); c->d_full.release(); c->d_orient.release(); c->d_full_meta.release();
  c->d_runs_tmp.release(); c->d_bt_tmp.release(); c->d_pairs.release(); c->d_out.release(); c->d_arena.release();
  c->d_todo.release(); c->d_scalars.release(); c->d_byteoff.release(); c->d_bytes.release(); c->d_paf_recs.release(); c->d_paf_runs.release();
  c->d_tasks2.release(); c->d_task_pair2.release(); c->d_task_out2.release(); c->d_esplit.release();
  c->d_tasks.release(); c->d_task_pair.release(); c->d_task_out.release(); c->d_split.release(); c->d_ring2.release();
  c->d_full2.release(); c->d_full_meta2.release(); c->d_runs_tmp2.release(); c->d_bt_tmp2.release();
  for (int i = 0; i < 10; ++i) if (c->ev[i]) cudaEventDestroy(c->ev[i]);
  if (c->own_stream) cudaStreamDestroy(c->stream);
  delete c;
}

extern "C" void* srgpu_ctx_stream(srgpu_ctx* c) { return c ? (void*)c->stream : nullptr; }

extern "C" int srgpu_ctx_synchronize(srgpu_ctx* c) {
  if (!c) return SRGPU_E_ARG;
  CK(cudaSetDevice(c->device));
  CK(cudaStreamSynchronize(c->stream));
  return SRGPU_OK;
}

extern "C" int srgpu_ctx_configure(srgpu_ctx* c, int threads, int ctas_per_sm, int ring_diagonals) {
  if (!c) return SRGPU_E_ARG;
  std::lock_guard<std::mutex> lk(c->mu);
  if (threads) {
    if (threads != 32 && threads != 64 && threads != 128 && threads != 256 && threads != 512) {
      g_err = "threads must be 32, 64, 128, 256 or 512";
      return SRGPU_E_ARG;
    }
    c->threads = threads;
    c->auto_shape = false;
  }
  if (ctas_per_sm > 0) {
    c->ctas_per_sm = ctas_per_sm;
    c->auto_shape = false;
  }
  if (ring_diagonals >= 0) c->ring_diagonals = ring_diagonals;
  return SRGPU_OK;
}

extern "C" int srgpu_ctx_set_split(srgpu_ctx* c, int split_score) {
  if (!c) return SRGPU_E_ARG;
  if (split_score != 0 && split_score < 250) {
    g_err = "split_score must be 0 (monolithic) or at least 250";
    return SRGPU_E_ARG;
  }
  std::lock_guard<std::mutex> lk(c->mu);
  c->split_T = split_score;
  return SRGPU_OK;
}

extern "C" int srgpu_ctx_stats(srgpu_ctx* c, srgpu_stats* out) {
  if (!c || !out) return SRGPU_E_ARG;
  std::lock_guard<std::mutex> lk(c->mu);
  *out = c->stats;
  return SRGPU_OK;
}

extern "C" uint64_t srgpu_total_length(srgpu_ctx* c) { return c ? c->total_len : 0; }

static const size_t SEQ_PAD = 32;

extern "C" int srgpu_set_sequences(srgpu_ctx* c, const srgpu_seqview* seqs, uint32_t n) {
  if (!c || (!seqs && n)) return SRGPU_E_ARG;
  std::lock_guard<std::mutex> lk(c->mu);
  CK(cudaSetDevice(c->device));
  std::vector<uint64_t> byte_off(n), pos_off(n);
  std::vector<uint32_t> len(n);
  std::vector<uint8_t> rcsafe(n), rcinv(n);
  uint64_t bytes = SEQ_PAD, pos = 0;
  uint32_t maxlen = 0;
  uint64_t ident = 0xcbf29ce484222325ull ^ n;
  for (uint32_t i = 0; i < n; ++i) {
    if (seqs[i].len == 0) {  // src/seqrush.rs:311-316 panics on empty sequences
      g_err = "empty sequences are not allowed";
      return SRGPU_E_ARG;
    }
    if (seqs[i].len > 0x3fffffffu) {
      g_err = "sequence too long for 32-bit wavefront offsets";
      return SRGPU_E_ARG;
    }
    byte_off[i] = bytes;
    pos_off[i] = pos;
    len[i] = (uint32_t)seqs[i].len;
    {
      auto mix = [&](uint64_t v) { ident = (ident ^ v) * 0x100000001b3ull; };
      mix(seqs[i].len);
      const uint64_t head = std::min<uint64_t>(seqs[i].len, 64);
      for (uint64_t j = 0; j < head; ++j) mix(seqs[i].data[j]);
      for (uint64_t j = seqs[i].len - head; j < seqs[i].len; ++j) mix(seqs[i].data[j]);
    }
    maxlen = std::max(maxlen, len[i]);
    bytes += (seqs[i].len + SEQ_PAD + 15) & ~(uint64_t)15;
    pos += seqs[i].len;
  }
  bytes += SEQ_PAD;
  // pinned staging buffer: padded concatenation
  uint8_t* stage = nullptr;
  CK(cudaMallocHost((void**)&stage, bytes));
  memset(stage, 0, bytes);
  for (uint32_t i = 0; i < n; ++i) {
    memcpy(stage + byte_off[i], seqs[i].data, seqs[i].len);
    uint8_t safe = 1;
    const uint8_t* d = seqs[i].data;
    for (uint64_t j = 0; j < seqs[i].len; ++j) {
      const uint8_t ch = d[j];
      if (!(ch == 'A' || ch == 'C' || ch == 'G' || ch == 'T' || ch == 'a' || ch == 'c' || ch == 'g' || ch == 't' || ch == 'N')) {
        safe = 0;
        break;
      }
    }
    rcsafe[i] = safe;
    uint8_t inv = 1;  // upper-case ACGTN only: allwave's reverse complement (alignment.rs:174-186) is an involution there
    for (uint64_t j = 0; j < seqs[i].len && inv; ++j) {
      const uint8_t ch = d[j];
      inv = (ch == 'A' || ch == 'C' || ch == 'G' || ch == 'T' || ch == 'N') ? 1 : 0;
    }
    rcinv[i] = inv;
  }
  // alphabet of the forward bytes plus what allwave's reverse complement can produce; the packing
  // width is the smallest that recodes it injectively (2 bits for pure ACGT, 4 bits with N / soft
  // masking, 8 bits = raw bytes otherwise)
  bool seen[256] = {false};
  for (uint32_t i = 0; i < n; ++i) {
    const uint8_t* d = seqs[i].data;
    for (uint64_t j = 0; j < seqs[i].len; ++j) seen[d[j]] = true;
  }
  bool any_other = false;
  for (int b = 0; b < 256; ++b)
    if (seen[b] && !strchr("ACGTacgt", b)) any_other = true;
  if (seen['A'] || seen['a']) seen['T'] = true;
  if (seen['T'] || seen['t']) seen['A'] = true;
  if (seen['C'] || seen['c']) seen['G'] = true;
  if (seen['G'] || seen['g']) seen['C'] = true;
  if (any_other) seen['N'] = true;
  CodeTable tab;
  int nsym = 0;
  for (int b = 0; b < 256; ++b) tab.code[b] = 0;
  for (int b = 0; b < 256; ++b)
    if (seen[b]) tab.code[b] = (uint8_t)(nsym++);
  const int lbits = nsym <= 4 ? 1 : nsym <= 16 ? 2 : 3;
  if (lbits == 3)
    for (int b = 0; b < 256; ++b) tab.code[b] = (uint8_t)b;
  std::vector<uint64_t> pk_off(n);
  uint64_t pk_words = 4;
  uint32_t max_pk = 0;
  for (uint32_t i = 0; i < n; ++i) {
    const uint64_t nw = ((((uint64_t)len[i] << lbits) + 31) >> 5);
    const uint64_t nw4 = (nw + 3) & ~(uint64_t)3;
    pk_off[i] = pk_words;
    pk_words += nw4 + 8;  // 4 guard words after this sequence + 4 before the next
    max_pk = std::max<uint32_t>(max_pk, (uint32_t)(nw4 + 8));
  }
  int rc;
  if ((rc = c->d_fwd.ensure(bytes)) || (rc = c->d_byte_off.ensure(n)) || (rc = c->d_pos_off.ensure(n)) ||
      (rc = c->d_len.ensure(n)) || (rc = c->d_rcsafe.ensure(n)) || (rc = c->d_rcinv.ensure(n)) || (rc = c->d_pk_off.ensure(n)) ||
      (rc = c->d_pk_fwd.ensure(pk_words)) || (rc = c->d_pk_rc.ensure(pk_words))) {
    cudaFreeHost(stage);
    return rc;
  }
  cudaError_t e = cudaMemcpyAsync(c->d_fwd.p, stage, bytes, cudaMemcpyHostToDevice, c->stream);
  if (e == cudaSuccess) e = cudaMemsetAsync(c->d_pk_fwd.p, 0, pk_words * 4, c->stream);
  if (e == cudaSuccess) e = cudaMemsetAsync(c->d_pk_rc.p, 0, pk_words * 4, c->stream);
  if (e == cudaSuccess && n) e = cudaMemcpyAsync(c->d_byte_off.p, byte_off.data(), n * 8, cudaMemcpyHostToDevice, c->stream);
  if (e == cudaSuccess && n) e = cudaMemcpyAsync(c->d_pos_off.p, pos_off.data(), n * 8, cudaMemcpyHostToDevice, c->stream);
  if (e == cudaSuccess && n) e = cudaMemcpyAsync(c->d_pk_off.p, pk_off.data(), n * 8, cudaMemcpyHostToDevice, c->stream);
  if (e == cudaSuccess && n) e = cudaMemcpyAsync(c->d_len.p, len.data(), n * 4, cudaMemcpyHostToDevice, c->stream);
  if (e == cudaSuccess && n) e = cudaMemcpyAsync(c->d_rcsafe.p, rcsafe.data(), n, cudaMemcpyHostToDevice, c->stream);
  if (e == cudaSuccess && n) e = cudaMemcpyAsync(c->d_rcinv.p, rcinv.data(), n, cudaMemcpyHostToDevice, c->stream);
  if (e == cudaSuccess && n) {
    const uint32_t per = 32u >> lbits;
    dim3 grid(std::max(1u, std::min(64u, (maxlen / per + 255) / 256)), std::min(n, 65535u));
    pack_kernel<<<grid, 256, 0, c->stream>>>(c->d_fwd.p, c->d_byte_off.p, c->d_len.p, c->d_pk_off.p, c->d_pk_fwd.p,
                                             c->d_pk_rc.p, tab, lbits, n);
    e = cudaGetLastError();
  }
  if (e == cudaSuccess) e = cudaStreamSynchronize(c->stream);
  cudaFreeHost(stage);
  if (e != cudaSuccess) {
    g_err = std::string("set_sequences: ") + cudaGetErrorString(e);
    return SRGPU_E_CUDA;
  }
  c->lbits = lbits;
  c->max_pk_words = max_pk;
  c->n_seqs = n;
  c->seq_identity = ident ? ident : 1;
  c->total_len = pos;
  c->max_len = maxlen;
  c->h_len = len;
  c->h_pos_off = pos_off;
  return SRGPU_OK;
}

static SeqTable seq_table(srgpu_ctx* c) {
  SeqTable t;
  t.fwd = c->d_fwd.p;
  t.pk_fwd = c->d_pk_fwd.p;
  t.pk_rc = c->d_pk_rc.p;
  t.pk_off = c->d_pk_off.p;
  t.lbits = c->lbits;
  t.byte_off = c->d_byte_off.p;
  t.len = c->d_len.p;
  t.pos_off = c->d_pos_off.p;
  t.rc_safe = c->d_rcsafe.p;
  t.rc_inv = c->d_rcinv.p;
  t.n = c->n_seqs;
  return t;
}

// allwave's three modes (allwave/src/alignment.rs:203-231, AlignmentMode::from_params allwave/src/types.rs:101-111) all
// reach WFA2-lib as gap-affine or gap-affine-2p BiWFA:
//   * two-piece affine (gap2 present)                         -> wavefront_penalties_set_affine2p
//   * single-piece affine and "edit distance" (o = e = x)     -> wavefront_penalties_set_affine; run here as two EQUAL
//     pieces: I2/D2 then duplicate I1/D1 cell for cell, every decision (limits, overlap order D before I before M,
//     backtrace priority) picks an equivalent cell, score and CIGAR are those of the single-piece recurrence
//   * match < 0: WFA2-lib's adjustment (Eizenga's formula, WFA/wavefront_penalties.c:113-134,157-175); match > 0 makes
//     WFA2-lib print "Match score must be negative or zero" and exit(1) (:100-103,141-144) -> SRGPU_E_ARG here.
static int check_penalties(const srgpu_penalties* p, Penalties* out) {
  if (!p) return SRGPU_E_ARG;
  if (p->match_score > 0) {
    char b[128];
    snprintf(b, sizeof b, "[WFA::Penalties] Match score must be negative or zero (M=%d)", p->match_score);
    g_err = b;
    return SRGPU_E_ARG;
  }
  const bool two = p->gap_ext2 > 0;  // gap2 absent: gap_open2 / gap_ext2 <= 0 (None in allwave::AlignmentParams)
  const int o2 = two ? p->gap_open2 : p->gap_open1, e2 = two ? p->gap_ext2 : p->gap_ext1;
  if (p->mismatch <= 0 || p->gap_open1 < 0 || p->gap_ext1 <= 0 || o2 < 0 || e2 <= 0) {
    g_err = "[WFA::Penalties] Penalties must be (X>0,O1>=0,E1>0,O2>=0,E2>0)";
    return SRGPU_E_ARG;
  }
  const int m = p->match_score;
  out->match = m;
  if (m < 0) {
    out->x = 2 * p->mismatch - 2 * m;
    out->o1 = 2 * p->gap_open1;
    out->e1 = 2 * p->gap_ext1 - m;
    out->o2 = 2 * o2;
    out->e2 = 2 * e2 - m;
  } else {
    out->x = p->mismatch;
    out->o1 = p->gap_open1;
    out->e1 = p->gap_ext1;
    out->o2 = o2;
    out->e2 = e2;
  }
  out->scope = std::max(std::max(out->o1 + out->e1, out->o2 + out->e2), out->x) + 1;
  if (out->scope + 1 > MAX_SCOPE) {
    g_err = "penalties too large: max_score_scope exceeds the ring depth (64)";
    return SRGPU_E_ARG;
  }
  return 0;
}

// (the dynamic shared-memory attribute is per device and per function: it is set on every launch --
// a host-side call of a few hundred nanoseconds -- instead of being remembered in a process-wide
// static, which was wrong for a second context on another GPU of the same process)
template <int T, int C>
static cudaError_t launch_align_tc(const AlignArgs& a, int grid, size_t smem, cudaStream_t s) {
  cudaError_t e = cudaFuncSetAttribute(align_kernel<T, C>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) return e;
  align_kernel<T, C><<<grid, T, smem, s>>>(a);
  return cudaGetLastError();
}
// (threads, ctas_per_sm) -> instantiation; the second template argument only caps registers
#define ALIGN_DISPATCH(FN, ...)                                                  \
  do {                                                                           \
    if (threads == 32) return FN<32, 16>(__VA_ARGS__);                           \
    if (threads == 64) {                                                         \
      if (ctas > 12) return FN<64, 16>(__VA_ARGS__);                             \
      if (ctas > 8) return FN<64, 12>(__VA_ARGS__);                              \
      return FN<64, 8>(__VA_ARGS__);                                             \
    }                                                                            \
    if (threads == 128) {                                                        \
      if (ctas > 6) return FN<128, 8>(__VA_ARGS__);                              \
      if (ctas > 5) return FN<128, 6>(__VA_ARGS__);                              \
      if (ctas > 4) return FN<128, 5>(__VA_ARGS__);                              \
      return FN<128, 4>(__VA_ARGS__);                                            \
    }                                                                            \
    if (threads == 512) {                                                        \
      if (ctas > 1) return FN<512, 2>(__VA_ARGS__);                              \
      return FN<512, 1>(__VA_ARGS__);                                            \
    }                                                                            \
    if (ctas > 3) return FN<256, 4>(__VA_ARGS__);                                \
    if (ctas > 2) return FN<256, 3>(__VA_ARGS__);                                \
    return FN<256, 2>(__VA_ARGS__);                                              \
  } while (0)
static cudaError_t launch_align(const AlignArgs& a, int threads, int ctas, int grid, size_t smem, cudaStream_t s) {
  ALIGN_DISPATCH(launch_align_tc, a, grid, smem, s);
}

template <int T, int C>
static int occ_tc(size_t smem) {
  int nb = 0;
  cudaFuncSetAttribute(align_kernel<T, C>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, align_kernel<T, C>, T, smem) != cudaSuccess) return 1;
  return nb < 1 ? 1 : nb;
}
static int align_occupancy(int threads, int ctas, size_t smem) { ALIGN_DISPATCH(occ_tc, smem); }

// core driver: pairs already on the device (d_pairs). Fills c->d_out (n_pairs records) and, per
// chunk, runs the union kernel when uf != nullptr. When keep_runs, a single chunk is used and the
// arena is left intact for CIGAR expansion.
static int run_pairs(srgpu_ctx* c, srgpu_uf* uf, const uint2* d_pairs, uint64_t n_pairs, const Penalties& pen,
                     uint64_t min_match_len, bool keep_runs) {
  CK(cudaSetDevice(c->device));
  c->stats = srgpu_stats{};
  if (n_pairs == 0) return SRGPU_OK;
  if (n_pairs > 0xfffffff0ull) {
    g_err = "too many pairs in one call";
    return SRGPU_E_ARG;
  }
  const uint32_t maxlen = c->max_len;
  // orient_kernel stages its three packed sequences (query forward, query RC, target) in dynamic shared
  // memory when they fit; align_kernel its two, next to the per-warp row staging buffers of the score
  // step, when the launch shape keeps its CTAs per SM with them (else they are read from global memory
  // through L1: the first packed window of every cell then waits for L1/L2, 12 % of all warp stall
  // samples in profiles/r02d_*). SRGPU_STAGE_SEQS=0 / 1 forces either way.
  size_t smem_orient = (size_t)3 * c->max_pk_words * 4;
  if (smem_orient > (size_t)c->smem_budget) smem_orient = 0;
  // launch shape: 128 threads x 4 CTAs/SM. With the row staging in shared memory the sequences of a 100 kb pair no
  // longer fit beside it at 4 CTAs/SM and are read from global memory, which still beats 256 x 2 with staged
  // sequences there (400-pair slice of 100 x 100 kb: 854 ms against 967 ms, profiles/r02e_tune_c4s.log; in round 1,
  // when every CTA had to stage its sequences, it was the other way round). Far longer sequences get 256 x 2: their
  // wavefronts keep eight warps busy and half as many resident pairs halve the ring memory (174 MB per pair at
  // 200 kb / 5 %).
  int threads = c->threads, ctas_per_sm = c->ctas_per_sm;
  if (c->auto_shape && maxlen > 150000) {
    threads = 256;
    ctas_per_sm = 2;
  }
  // Few pairs: with less than one pair per resident CTA slot the SMs would run one or two four-warp CTAs each
  // (a 16 x 200 kb slice of configs[4]: 111 pairs, 27 s). More warps per pair use them: at most one pair per SM ->
  // 512 threads x 1 CTA, at most two -> 256 x 2. (--iterative launches 40 alignments at a time.)
  if (c->auto_shape) {
    const uint64_t per_sm = (n_pairs + (uint64_t)c->n_sm - 1) / (uint64_t)c->n_sm;
    if (per_sm <= 1) {
      threads = 512;
      ctas_per_sm = 1;
    } else if (per_sm <= 2 && threads < 256) {
      threads = 256;
      ctas_per_sm = 2;
    }
  }
  const size_t smem_rows = (size_t)(threads / 32) * STAGE_INTS * sizeof(int32_t);
  size_t smem_seqs = (size_t)2 * c->max_pk_words * 4;
  uint32_t smem_seq_words = c->max_pk_words;
  {
    const size_t per_sm = 227 * 1024, static_smem = 13 * 1024 + 1024;  // CtaShared + CtaTables + the per-CTA reservation
    const bool fits = (static_smem + smem_rows + smem_seqs) * (size_t)ctas_per_sm <= per_sm;
    const bool want = c->stage_seqs < 0 ? fits : c->stage_seqs != 0;
    if (!want || smem_seqs > (size_t)c->smem_budget || static_smem + smem_rows + smem_seqs > per_sm) {
      smem_seqs = 0;
      smem_seq_words = 0;
    }
  }
  const size_t smem = smem_seqs + smem_rows;
  const int resident = std::min(ctas_per_sm, align_occupancy(threads, ctas_per_sm, smem));
  int grid = c->n_sm * resident;
  if ((uint64_t)grid > n_pairs) grid = (int)n_pairs;
  // workspace geometry
  // a full row: every diagonal -plen .. tlen, one more on each side before the trim, and the 8 diagonals of slack the
  // vector step keeps (a row of exactly 2 * (maxlen + 4) overflowed for pairs whose wavefront reaches the corner
  // diagonals -- unrelated sequences -- whenever that number was already a multiple of 64)
  const int fullW = (int)(2 * ((uint64_t)maxlen + 16));
  int W = c->ring_diagonals > 0 ? c->ring_diagonals : 32768;
  if (W > fullW) W = fullW;
  W = (W + 63) & ~63;
  const int cap = std::max(512, 100 * pen.x + 8) + 8;
  const size_t runs_cap = 2 * (size_t)maxlen + 16;
  const size_t orient_row = 2 * (size_t)maxlen + 16;
  int rc;
  if ((rc = c->d_out.ensure(n_pairs))) return rc;
  if ((rc = c->d_scalars.ensure(16))) return rc;
  if ((rc = c->d_full.ensure((size_t)grid * 5 * cap * cap))) return rc;
  if ((rc = c->d_full_meta.ensure((size_t)grid * cap * 5))) return rc;
  // the strand-choice kernel needs 32 registers and is bound by the latency of its global-memory rows (ncu: long
  // scoreboard 3.8 cycles per issue at 17 % warps active when it ran on the align kernel's grid, profiles/r02n_*): it
  // gets its own grid, as many CTAs per SM as its staged sequences allow (at most 8)
  int orient_grid = grid;
  {
    int nb = 0;
    cudaFuncSetAttribute(orient_kernel<128>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_orient);
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, orient_kernel<128>, 128, smem_orient) == cudaSuccess && nb > 0)
      orient_grid = c->n_sm * std::min(nb, c->orient_ctas_per_sm);
    // keep the four edit-distance rows of every CTA within ~16 GB
    const size_t per = (size_t)4 * orient_row * sizeof(int32_t);
    orient_grid = (int)std::min<size_t>((size_t)orient_grid, std::max<size_t>(1, ((size_t)16 << 30) / per));
    if ((uint64_t)orient_grid > n_pairs) orient_grid = (int)n_pairs;
  }
  if ((rc = c->d_orient.ensure((size_t)orient_grid * 4 * orient_row))) return rc;
  if ((rc = c->d_runs_tmp.ensure((size_t)grid * runs_cap))) return rc;
  if ((rc = c->d_bt_tmp.ensure((size_t)grid * runs_cap))) return rc;
  // arena: run-length CIGAR elements of one chunk
  size_t arena_cap = std::max<size_t>((size_t)256 << 20, 4 * runs_cap);  // 256M runs = 1 GB
  if (keep_runs) arena_cap = std::max<size_t>(arena_cap, std::min<size_t>((size_t)n_pairs * runs_cap, (size_t)1 << 30));
  if ((rc = c->d_arena.ensure(arena_cap))) return rc;
  // chunking: assume <= 1 run per 16 columns on average (3% divergence); overflow is detected and the
  // pairs that did not fit are re-run. Few large launches keep the end-of-launch tail (CTAs idling while
  // the last pairs finish) small against the launch.
  // (in split mode the arena holds the pieces of the tasks and the joined list: twice the runs)
  uint64_t chunk = keep_runs ? n_pairs : std::max<uint64_t>(1, arena_cap / std::max<size_t>(1, runs_cap / 8));
  chunk = std::min<uint64_t>(chunk, n_pairs);
  if (!keep_runs) chunk = std::min<uint64_t>(chunk, std::max<uint64_t>((uint64_t)grid * 64, 4096));

  AlignArgs a;
  a.seqs = seq_table(c);
  a.pen = pen;
  a.full = c->d_full.p;
  a.full_stride = (size_t)5 * cap * cap;
  a.full_meta = c->d_full_meta.p;
  a.cap = cap;
  a.orient = c->d_orient.p;
  a.orient_stride = 4 * orient_row;
  a.orient_row = (int)orient_row;
  a.runs_tmp = c->d_runs_tmp.p;
  a.bt_tmp = c->d_bt_tmp.p;
  a.runs_cap = (uint32_t)runs_cap;
  a.smem_seq_words = smem_seq_words;
  a.arena = c->d_arena.p;
  a.arena_cap = arena_cap;
  a.arena_head = c->d_scalars.p;
  a.next_pair = (unsigned int*)(c->d_scalars.p + 1);
  a.leaf_T = 0;
  a.tasks = nullptr;
  a.task_pair = nullptr;
  a.task_out = nullptr;
  a.split = nullptr;
  a.task_cap = 0;
  a.task_count = (unsigned int*)(c->d_scalars.p + 3);
  a.next_task = (unsigned int*)(c->d_scalars.p + 4);
  a.leaf_T2 = 0;
  a.etasks = nullptr;
  a.etask_pair = nullptr;
  a.etask_out = nullptr;
  a.esplit = nullptr;
  a.etask_cap = 0;
  a.etask_count = (unsigned int*)(c->d_scalars.p + 5);
  a.asm_max = c->test_asm_max > 0 ? (unsigned int)std::min(c->test_asm_max, ASM_MAX_TASKS) : (unsigned int)ASM_MAX_TASKS;
  // split mode: geometry of the one-warp kernel. Its ring rows must hold a sub-problem of score T: either search runs
  // to about T/2 + scope, diagonals |k| <= score.
  const int split_T = (c->split_T >= 250 && !getenv("SRGPU_NO_SPLIT") && pen.scope + 1 <= SUB_SLOTS) ? c->split_T : 0;
  // (never wider than a full row, and within a 24 GB budget over the grid: a task that overflows its ring sends the
  // pair to the monolithic retry)
  int sub_W = ((int)std::min<long long>((long long)split_T + 2 * pen.scope + 128, (long long)fullW) + 63) & ~63;
  const size_t sub_smem = (size_t)STAGE_INTS * sizeof(int32_t);
  int sub_grid = 0, leaf_grid = 0;
  uint64_t task_cap = 0;
  if (split_T) {
    int nb = 0;
    cudaFuncSetAttribute(subalign_kernel<32, SUB_CTAS, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sub_smem);
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, subalign_kernel<32, SUB_CTAS, false>, 32, sub_smem) != cudaSuccess || nb < 1) nb = 1;
    sub_grid = c->n_sm * std::min(nb, c->sub_ctas_per_sm);
    {
      const size_t budget = (size_t)24 << 30, per_diag = (size_t)2 * (pen.scope + 1) * 5 * sizeof(int32_t);
      const size_t fit = budget / (per_diag * (size_t)sub_grid);
      if ((size_t)sub_W > fit) sub_W = (int)std::max<size_t>(fit & ~(size_t)63, 1024);
    }
    if (c->test_sub_w > 0) sub_W = (c->test_sub_w + 63) & ~63;  // tests: a ring too narrow for most tasks -> monolithic retry
    const size_t sub_ring = (size_t)2 * (pen.scope + 1) * 5 * sub_W;
    if ((rc = c->d_ring2.ensure((size_t)sub_grid * sub_ring))) return rc;
    // the leaf launch of the two-pass mode: no ring, no staging buffer, <= 64 registers: up to 32 one-warp CTAs per SM
    // (its base cases wait on global-memory rows: long scoreboard 3.8 of 9 stall cycles per issue at 13 CTAs/SM,
    // profiles/r03p_*), each with its own base-case store and two short run buffers
    leaf_grid = sub_grid;
    if (c->sub_two_pass) {
      int nl = 0;
      if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nl, subalign_kernel<32, LEAF_CTAS, true>, 32, 0) != cudaSuccess || nl < 1) nl = 1;
      leaf_grid = c->n_sm * std::min(nl, c->leaf_ctas_per_sm);
    }
    const size_t wgrid = (size_t)std::max(sub_grid, leaf_grid);
    if ((rc = c->d_full2.ensure(wgrid * 5 * cap * cap))) return rc;
    if ((rc = c->d_full_meta2.ensure(wgrid * cap * 5))) return rc;
    if ((rc = c->d_runs_tmp2.ensure(std::max((size_t)sub_grid * runs_cap, (size_t)leaf_grid * LEAF_RUNS_CAP)))) return rc;
    if ((rc = c->d_bt_tmp2.ensure(std::max((size_t)sub_grid * runs_cap, (size_t)leaf_grid * LEAF_RUNS_CAP)))) return rc;
  }

  CK(cudaEventRecord(c->ev[0], c->stream));
  std::vector<PairOut> h_out;
  std::vector<uint32_t> todo;
  auto do_union = [&](uint64_t first, const uint32_t* d_todo, uint32_t count) -> int {
    if (!uf || count == 0) return 0;
    UnionArgs u;
    u.seqs = a.seqs;
    u.out = c->d_out.p + first;
    u.arena = c->d_arena.p;
    u.n_pairs = count;
    u.todo = d_todo;
    u.parent = uf->parent.p;
    u.min_match_len = min_match_len;
    u.counters = uf->counters.p;
    CK(cudaEventRecord(c->ev[1], c->stream));
    union_kernel<256><<<std::min<uint32_t>(count, (uint32_t)c->n_sm * 8), 256, 0, c->stream>>>(u);
    CK(cudaGetLastError());
    CK(cudaEventRecord(c->ev[2], c->stream));
    CK(cudaStreamSynchronize(c->stream));
    float ms = 0;
    CK(cudaEventElapsedTime(&ms, c->ev[1], c->ev[2]));
    c->stats.union_ms += ms;
    c->stats.kernel_launches++;
    c->stats.union_launches++;
    return 0;
  };
  for (uint64_t first = 0; first < n_pairs; first += chunk) {
    const uint32_t cn = (uint32_t)std::min<uint64_t>(chunk, n_pairs - first);
    a.pairs = d_pairs + first;
    a.out = c->d_out.p + first;
    a.n_pairs = cn;
    a.todo = nullptr;
    a.mirror_of = nullptr;
    int curW = W;
    int cur_grid = std::min<int>(grid, (int)cn);
    CK(cudaMemsetAsync(c->d_scalars.p, 0, 2 * sizeof(unsigned long long), c->stream));
    for (int attempt = 0;; ++attempt) {
      if (attempt == 0) {
        // strand choice of every pair of the chunk (its own launch: three staged sequences)
        AlignArgs ao = a;
        ao.smem_seq_words = smem_orient ? c->max_pk_words : 0;
        CK(cudaMemsetAsync(a.next_pair, 0, sizeof(unsigned int), c->stream));
        CK(cudaEventRecord(c->ev[4], c->stream));
        MirrorArgs mi{};
        if (c->mirror_orient && cn >= 2) {
          // a pair (q, t), q > t, whose mirror (t, q) is in this launch copies the mirror's strand choice (kernels.cuh)
          uint32_t slots = 64;
          while (slots < 2 * cn) slots <<= 1;
          if ((rc = c->d_mirror_keys.ensure(slots)) || (rc = c->d_mirror_vals.ensure(slots)) || (rc = c->d_mirror_of.ensure(cn))) return rc;
          CK(cudaMemsetAsync(c->d_mirror_keys.p, 0xff, (size_t)slots * sizeof(unsigned long long), c->stream));
          mi.pairs = a.pairs;
          mi.n = cn;
          mi.rc_inv = a.seqs.rc_inv;
          mi.keys = c->d_mirror_keys.p;
          mi.vals = c->d_mirror_vals.p;
          mi.mask = slots - 1;
          mi.mirror_of = c->d_mirror_of.p;
          mi.out = a.out;
          mirror_insert_kernel<<<grid_for(cn, 256, c->n_sm), 256, 0, c->stream>>>(mi);
          mirror_lookup_kernel<<<grid_for(cn, 256, c->n_sm), 256, 0, c->stream>>>(mi);
          CK(cudaGetLastError());
          ao.mirror_of = c->d_mirror_of.p;
          c->stats.kernel_launches += 2;
        }
        CK(cudaFuncSetAttribute(orient_kernel<128>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_orient));
        orient_kernel<128><<<std::min<int>(orient_grid, (int)cn), 128, smem_orient, c->stream>>>(ao);
        CK(cudaGetLastError());
        if (ao.mirror_of) {
          mirror_copy_kernel<<<grid_for(cn, 256, c->n_sm), 256, 0, c->stream>>>(mi);
          CK(cudaGetLastError());
          c->stats.kernel_launches++;
        }
        CK(cudaEventRecord(c->ev[5], c->stream));
        c->stats.kernel_launches++;
        if (c->ring_diagonals <= 0 && curW < fullW) {
          // ring rows wide enough for the most divergent pair of the chunk (the kernel sizes each pair's rows as
          // 8 x edit distance + 512 diagonals): one scalar read-back instead of a failed launch and a retry
          unsigned int* d_maxed = (unsigned int*)(c->d_scalars.p + 2);
          CK(cudaMemsetAsync(d_maxed, 0, sizeof(unsigned int), c->stream));
          max_edit_distance_kernel<<<grid_for(cn, 256, c->n_sm), 256, 0, c->stream>>>(a.out, cn, d_maxed);
          CK(cudaGetLastError());
          unsigned int maxed = 0;
          CK(cudaMemcpyAsync(&maxed, d_maxed, sizeof maxed, cudaMemcpyDeviceToHost, c->stream));
          CK(cudaStreamSynchronize(c->stream));
          c->stats.kernel_launches++;
          const long long need = ((8ll * maxed + 512) + 63) & ~63ll;
          if (need > curW) curW = (int)std::min<long long>(need, (fullW + 63) & ~63);
          const size_t per = (size_t)2 * (pen.scope + 1) * 5 * curW * sizeof(int32_t);
          cur_grid = (int)std::min<size_t>((size_t)cur_grid, std::max<size_t>(1, ((size_t)96 << 30) / per));
        }
      }
      if ((rc = c->d_ring.ensure((size_t)cur_grid * 2 * (pen.scope + 1) * 5 * curW))) return rc;
      a.ring = c->d_ring.p;
      a.ring_stride = (size_t)2 * (pen.scope + 1) * 5 * curW;
      a.W = curW;
      a.fixed_w = attempt > 0 ? 1 : 0;
      CK(cudaMemsetAsync(a.next_pair, 0, sizeof(unsigned int), c->stream));
      // first attempt in split mode: align_kernel splits the big levels, subalign_kernel solves the sub-problems one
      // warp each, assemble_kernel joins the pieces; re-runs (wider ring, arena) take the monolithic path
      const bool split_now = split_T > 0 && attempt == 0;
      a.leaf_T = 0;
      if (split_now) {
        task_cap = std::min<uint64_t>(std::max<uint64_t>((uint64_t)cn * 256, 65536), (uint64_t)1 << 26);
        if (c->test_task_cap > 0) task_cap = (uint64_t)c->test_task_cap;  // tests: force the overflow -> monolithic retry path
        if ((rc = c->d_tasks.ensure(task_cap)) || (rc = c->d_task_pair.ensure(task_cap)) || (rc = c->d_task_out.ensure(task_cap)) ||
            (rc = c->d_split.ensure(n_pairs)))
          return rc;
        a.leaf_T = split_T;
        a.tasks = c->d_tasks.p;
        a.task_pair = c->d_task_pair.p;
        a.task_out = c->d_task_out.p;
        a.split = c->d_split.p + first;
        a.task_cap = (unsigned int)task_cap;
        CK(cudaMemsetAsync(c->d_scalars.p + 3, 0, 4 * sizeof(unsigned long long), c->stream));
        a.etasks = nullptr;
        if (c->sub_two_pass) {
          uint64_t ecap = std::min<uint64_t>(task_cap * 4, (uint64_t)1 << 26);
          if (c->test_leaf_cap > 0) ecap = (uint64_t)c->test_leaf_cap;
          if ((rc = c->d_tasks2.ensure(ecap)) || (rc = c->d_task_pair2.ensure(ecap)) || (rc = c->d_task_out2.ensure(ecap)) ||
              (rc = c->d_esplit.ensure(task_cap)))
            return rc;
          a.leaf_T2 = 250;  // WF_BIALIGN_FALLBACK_MIN_SCORE: below it the recursion takes the base case
          a.etasks = c->d_tasks2.p;
          a.etask_pair = c->d_task_pair2.p;
          a.etask_out = c->d_task_out2.p;
          a.esplit = c->d_esplit.p;
          a.etask_cap = (unsigned int)ecap;
        }
      } else {
        a.etasks = nullptr;
      }
      CK(cudaEventRecord(c->ev[1], c->stream));
      cudaError_t e = launch_align(a, threads, ctas_per_sm, cur_grid, smem, c->stream);
      if (e != cudaSuccess) {
        g_err = std::string("align_kernel launch: ") + cudaGetErrorString(e);
        return SRGPU_E_CUDA;
      }
      if (c->debug_sync) CK(cudaStreamSynchronize(c->stream));  // SRGPU_DEBUG_SYNC: a fault is reported at the launch that caused it
      if (split_now) {
        CK(cudaEventRecord(c->ev[6], c->stream));
        AlignArgs s2 = a;
        s2.ring = c->d_ring2.p;
        s2.ring_stride = (size_t)2 * (pen.scope + 1) * 5 * sub_W;
        s2.W = sub_W;
        s2.full = c->d_full2.p;
        s2.full_meta = c->d_full_meta2.p;
        s2.runs_tmp = c->d_runs_tmp2.p;
        s2.bt_tmp = c->d_bt_tmp2.p;
        subalign_kernel<32, SUB_CTAS, false><<<sub_grid, 32, sub_smem, c->stream>>>(s2);  // one pass: whole tasks; two: bisection only
        CK(cudaGetLastError());
        if (c->debug_sync) CK(cudaStreamSynchronize(c->stream));
        if (a.etasks) {
          CK(cudaEventRecord(c->ev[8], c->stream));
          AlignArgs s3 = s2;  // second pass: the leaves are the tasks now
          s3.tasks = a.etasks;
          s3.task_pair = a.etask_pair;
          s3.task_out = a.etask_out;
          s3.task_cap = a.etask_cap;
          s3.task_count = a.etask_count;
          s3.next_task = (unsigned int*)(c->d_scalars.p + 6);
          s3.etasks = nullptr;
          s3.runs_cap = LEAF_RUNS_CAP;  // a leaf's score is at most the recovery bound (500): <= 1001 runs
          subalign_kernel<32, LEAF_CTAS, true><<<leaf_grid, 32, 0, c->stream>>>(s3);
          CK(cudaGetLastError());
          if (c->debug_sync) CK(cudaStreamSynchronize(c->stream));
          c->stats.kernel_launches++;
        }
        CK(cudaEventRecord(c->ev[7], c->stream));
        assemble_kernel<128><<<(unsigned)std::min<uint64_t>(cn, (uint64_t)c->n_sm * 8), 128, 0, c->stream>>>(a);
        CK(cudaGetLastError());
        if (c->debug_sync) CK(cudaStreamSynchronize(c->stream));
        c->stats.kernel_launches += 2;
      }
      CK(cudaEventRecord(c->ev[2], c->stream));
      c->stats.kernel_launches++;
      c->stats.align_launches++;
      h_out.resize(cn);
      CK(cudaMemcpyAsync(h_out.data(), c->d_out.p + first, (size_t)cn * sizeof(PairOut), cudaMemcpyDeviceToHost, c->stream));
      CK(cudaStreamSynchronize(c->stream));
      float ms = 0;
      CK(cudaEventElapsedTime(&ms, c->ev[1], c->ev[2]));
      c->stats.align_ms += ms;
      if (split_now && getenv("SRGPU_TRACE")) {
        float m1 = 0, m2 = 0, m3 = 0;
        cudaEventElapsedTime(&m1, c->ev[1], c->ev[6]);
        cudaEventElapsedTime(&m2, c->ev[6], c->ev[7]);
        cudaEventElapsedTime(&m3, c->ev[7], c->ev[2]);
        unsigned int nt = 0;
        cudaMemcpy(&nt, c->d_scalars.p + 3, 4, cudaMemcpyDeviceToHost);
        float m2a = 0;
        unsigned int nl = 0;
        if (a.etasks) {
          cudaEventElapsedTime(&m2a, c->ev[6], c->ev[8]);
          cudaMemcpy(&nl, c->d_scalars.p + 5, 4, cudaMemcpyDeviceToHost);
        }
        fprintf(stderr, "[srgpu] split T=%d: align_kernel %.1f ms, subalign_kernel %.1f ms (%u tasks, grid %d, W %d; bisect pass %.1f ms, %u leaves), assemble_kernel %.2f ms\n",
                split_T, m1, m2, nt, sub_grid, sub_W, m2a, nl, m3);
      }
      if (attempt == 0) {
        CK(cudaEventElapsedTime(&ms, c->ev[4], c->ev[5]));
        c->stats.orient_ms += ms;
      }
      // unite what this launch produced (records of failed pairs are skipped by the kernel)
      if ((rc = do_union(first, a.todo, a.n_pairs))) return rc;
      // pairs that need a wider ring or did not fit the arena (rare) are re-run
      std::vector<uint32_t> again;
      bool diag = false, arena_over = false;
      const uint32_t cnt = a.todo ? (uint32_t)todo.size() : cn;
      for (uint32_t j = 0; j < cnt; ++j) {
        const uint32_t i = a.todo ? todo[j] : j;
        const int st = h_out[i].status;
        if (st == ST_WS_DIAG_OVERFLOW) { again.push_back(i); diag = true; }
        else if (st == ST_ARENA_OVERFLOW) { again.push_back(i); arena_over = true; }
        else if (st == ST_TASK_OVERFLOW) { again.push_back(i); }
      }
      if (again.empty()) break;
      // every retry completes the pairs that fit (the arena is reused, the ring gets its full width), so
      // the list shrinks; a list that stops shrinking means a pair cannot be served at all: fail loudly
      if (attempt >= 2 && again.size() >= cnt) {
        g_err = "pairs do not fit the alignment workspace (ring width / CIGAR arena) after retries";
        return SRGPU_E_INTERNAL;
      }
      if (keep_runs && arena_over) {
        g_err = "CIGAR arena overflow: align fewer pairs per call";
        return SRGPU_E_INTERNAL;
      }
      c->stats.retries++;
      todo.swap(again);
      if (diag) curW = (fullW + 63) & ~63;
      cur_grid = std::min<int>(cur_grid, (int)todo.size());
      {  // keep the retry ring within ~24 GB
        const size_t per = (size_t)2 * (pen.scope + 1) * 5 * curW * sizeof(int32_t);
        const size_t maxg = std::max<size_t>(1, ((size_t)24 << 30) / per);
        cur_grid = (int)std::min<size_t>((size_t)cur_grid, maxg);
      }
      if ((rc = c->d_todo.ensure(todo.size()))) return rc;
      CK(cudaMemcpyAsync(c->d_todo.p, todo.data(), todo.size() * 4, cudaMemcpyHostToDevice, c->stream));
      if (!keep_runs) CK(cudaMemsetAsync(c->d_scalars.p, 0, sizeof(unsigned long long), c->stream));  // arena reuse
      a.todo = c->d_todo.p;
      a.n_pairs = (uint32_t)todo.size();
    }
    // accumulate stats from the final records of this chunk. A pair may only stay failed for a reason
    // the reference shares (ST_UNATTAINABLE = WF_STATUS_END_UNREACHABLE); a pair that hit a limit of
    // this library's workspaces fails the whole call: the partition must never silently lack pairs
    for (uint32_t i = 0; i < cn; ++i) {
      const int st = h_out[i].status;
      if (st == ST_UNATTAINABLE) {
        c->stats.failed_pairs++;
      } else if (st != ST_OK) {
        char buf[256];
        snprintf(buf, sizeof buf,
                 "pair %llu (query %u, target %u) exceeds a workspace limit of the library (status %d: %s)",
                 (unsigned long long)(first + i), h_out[i].q, h_out[i].t, st,
                 st == ST_WS_BASE_OVERFLOW ? "base-case score above the full-memory capacity"
                 : st == ST_WS_RUNS_OVERFLOW ? "run buffer overflow"
                 : st == ST_STACK_OVERFLOW  ? "recursion stack overflow"
                 : st == ST_RUNAWAY         ? "breakpoint search ran past its score bound"
                                            : "not served after retries");
        g_err = buf;
        return SRGPU_E_INTERNAL;
      }
    }
    for (uint32_t i = 0; i < cn; ++i) {
      c->stats.wf_cells += h_out[i].cells;
      c->stats.orient_cells += h_out[i].ocells;
      c->stats.aligned_bp += c->h_len[h_out[i].q];
      c->stats.runs += h_out[i].n_runs;
    }
  }
  CK(cudaEventRecord(c->ev[3], c->stream));
  CK(cudaStreamSynchronize(c->stream));
  CK(cudaEventElapsedTime(&c->stats.total_ms, c->ev[0], c->ev[3]));
  c->stats.pairs = n_pairs;
  if (uf) {
    unsigned long long cnt[4];
    CK(cudaMemcpy(cnt, uf->counters.p, sizeof cnt, cudaMemcpyDeviceToHost));
    c->stats.unite_ops = cnt[0];
    c->stats.unite_links = cnt[1];
  }
  return SRGPU_OK;
}

static void fill_result(srgpu_ctx* c, const PairOut& o, srgpu_result* r) {
  r->query_idx = o.q;
  r->target_idx = o.t;
  r->is_reverse = o.is_reverse;
  const bool ok = (o.status == ST_OK);
  r->status = ok ? SRGPU_PAIR_OK : SRGPU_PAIR_FAILED;
  r->score = ok ? o.score : INT32_MAX;
  r->penalty = ok ? o.penalty : 0;
  r->num_matches = ok ? o.num_matches : 0;
  r->alignment_length = ok ? o.num_matches + o.n_mismatch : 0;
  r->query_end = ok ? o.num_matches + o.n_mismatch + o.n_del : 0;
  r->target_end = ok ? o.num_matches + o.n_mismatch + o.n_ins : 0;
  r->edit_distance = o.edit_distance;
  r->n_runs = o.n_runs;
  r->cigar_off = 0;
  r->cigar_len = ok ? (uint64_t)o.num_matches + o.n_mismatch + o.n_ins + o.n_del : 0;
  r->wf_cells = o.cells;
  (void)c;
}

static int upload_pairs(srgpu_ctx* c, const uint32_t (*pairs)[2], uint64_t n_pairs) {
  for (uint64_t i = 0; i < n_pairs; ++i)
    if (pairs[i][0] >= c->n_seqs || pairs[i][1] >= c->n_seqs) {
      g_err = "pair index out of range";
      return SRGPU_E_ARG;
    }
  int rc;
  if ((rc = c->d_pairs.ensure(n_pairs))) return rc;
  CK(cudaMemcpyAsync(c->d_pairs.p, pairs, n_pairs * sizeof(uint2), cudaMemcpyHostToDevice, c->stream));
  return 0;
}

static int check_uf(srgpu_ctx* c, srgpu_uf* uf);
static int align_pairs_impl(srgpu_ctx* c, srgpu_uf* uf, uint64_t min_match_len, const uint32_t (*pairs)[2], uint64_t n_pairs,
                            const srgpu_penalties* pen, srgpu_result* out, uint8_t** cigar_arena, uint64_t* cigar_arena_len);
extern "C" int srgpu_align_pairs(srgpu_ctx* c, const uint32_t (*pairs)[2], uint64_t n_pairs, const srgpu_penalties* pen,
                                 srgpu_result* out, uint8_t** cigar_arena, uint64_t* cigar_arena_len) {
  return align_pairs_impl(c, nullptr, 0, pairs, n_pairs, pen, out, cigar_arena, cigar_arena_len);
}
extern "C" int srgpu_align_pairs_and_unite(srgpu_ctx* c, srgpu_uf* uf, const uint32_t (*pairs)[2], uint64_t n_pairs,
                                           const srgpu_penalties* pen, uint64_t min_match_len, srgpu_result* out,
                                           uint8_t** cigar_arena, uint64_t* cigar_arena_len) {
  if (!uf) return SRGPU_E_ARG;
  return align_pairs_impl(c, uf, min_match_len, pairs, n_pairs, pen, out, cigar_arena, cigar_arena_len);
}
static int align_pairs_impl(srgpu_ctx* c, srgpu_uf* uf, uint64_t min_match_len, const uint32_t (*pairs)[2], uint64_t n_pairs,
                            const srgpu_penalties* pen, srgpu_result* out, uint8_t** cigar_arena, uint64_t* cigar_arena_len) {
  if (!c || (!pairs && n_pairs) || !out) return SRGPU_E_ARG;
  std::lock_guard<std::mutex> lk(c->mu);
  CK(cudaSetDevice(c->device));
  Penalties p;
  int rc;
  if ((rc = check_penalties(pen, &p))) return rc;
  if (uf && (rc = check_uf(c, uf))) return rc;
  if (cigar_arena) *cigar_arena = nullptr;
  if (cigar_arena_len) *cigar_arena_len = 0;
  if (n_pairs == 0) return SRGPU_OK;
  if ((rc = upload_pairs(c, pairs, n_pairs))) return rc;
  if ((rc = run_pairs(c, uf, c->d_pairs.p, n_pairs, p, min_match_len, cigar_arena != nullptr))) return rc;
  std::vector<PairOut> h(n_pairs);
  CK(cudaMemcpy(h.data(), c->d_out.p, n_pairs * sizeof(PairOut), cudaMemcpyDeviceToHost));
  std::vector<unsigned long long> boff(n_pairs);
  uint64_t total = 0;
  for (uint64_t i = 0; i < n_pairs; ++i) {
    fill_result(c, h[i], &out[i]);
    out[i].cigar_off = total;
    boff[i] = total;
    total += out[i].cigar_len;
  }
  if (cigar_arena) {
    if ((rc = c->d_byteoff.ensure(n_pairs))) return rc;
    if ((rc = c->d_bytes.ensure(total + 16))) return rc;
    CK(cudaMemcpyAsync(c->d_byteoff.p, boff.data(), n_pairs * 8, cudaMemcpyHostToDevice, c->stream));
    expand_kernel<<<(unsigned)std::min<uint64_t>(n_pairs, (uint64_t)c->n_sm * 8), 256, 0, c->stream>>>(
        c->d_out.p, c->d_arena.p, c->d_byteoff.p, c->d_bytes.p, (uint32_t)n_pairs);
    CK(cudaGetLastError());
    c->stats.kernel_launches++;
    uint8_t* host = (uint8_t*)malloc(total + 1);
    if (!host) {
      g_err = "out of host memory for the CIGAR arena";
      return SRGPU_E_INTERNAL;
    }
    CK(cudaMemcpyAsync(host, c->d_bytes.p, total, cudaMemcpyDeviceToHost, c->stream));
    CK(cudaStreamSynchronize(c->stream));
    host[total] = 0;
    *cigar_arena = host;
    if (cigar_arena_len) *cigar_arena_len = total;
  }
  return SRGPU_OK;
}

extern "C" void srgpu_free(void* p) { free(p); }

static int check_uf(srgpu_ctx* c, srgpu_uf* uf) {
  if (!uf || uf->ctx != c) {
    g_err = "union-find handle does not belong to this context";
    return SRGPU_E_ARG;
  }
  if (uf->capacity < 2 * c->total_len) {
    g_err = "union-find smaller than the sequence set";
    return SRGPU_E_RANGE;
  }
  // A context holds ONE sequence table; a forest filled against one set and then united against another (a second
  // set uploaded into the same context in between) would silently mix the Pos spaces of the two.
  if (uf->seq_identity == 0) {
    uf->seq_identity = c->seq_identity;
  } else if (uf->seq_identity != c->seq_identity) {
    g_err = "union-find was filled against a different sequence set than the one now loaded in this context "
            "(srgpu_uf_clear it, or use one context per sequence set)";
    return SRGPU_E_ARG;
  }
  return 0;
}

extern "C" int srgpu_align_and_unite(srgpu_ctx* c, srgpu_uf* uf, const uint32_t (*pairs)[2], uint64_t n_pairs,
                                     const srgpu_penalties* pen, uint64_t min_match_len, srgpu_result* out) {
  if (!c || (!pairs && n_pairs)) return SRGPU_E_ARG;
  std::lock_guard<std::mutex> lk(c->mu);
  CK(cudaSetDevice(c->device));
  Penalties p;
  int rc;
  if ((rc = check_penalties(pen, &p))) return rc;
  if ((rc = check_uf(c, uf))) return rc;
  if (n_pairs == 0) return SRGPU_OK;
  if ((rc = upload_pairs(c, pairs, n_pairs))) return rc;
  if ((rc = run_pairs(c, uf, c->d_pairs.p, n_pairs, p, min_match_len, false))) return rc;
  if (out) {
    std::vector<PairOut> h(n_pairs);
    CK(cudaMemcpy(h.data(), c->d_out.p, n_pairs * sizeof(PairOut), cudaMemcpyDeviceToHost));
    for (uint64_t i = 0; i < n_pairs; ++i) fill_result(c, h[i], &out[i]);
  }
  return SRGPU_OK;
}

extern "C" int srgpu_align_and_unite_device(srgpu_ctx* c, srgpu_uf* uf, const void* d_pairs, uint64_t n_pairs,
                                            const srgpu_penalties* pen, uint64_t min_match_len) {
  if (!c || (!d_pairs && n_pairs)) return SRGPU_E_ARG;
  std::lock_guard<std::mutex> lk(c->mu);
  CK(cudaSetDevice(c->device));
  Penalties p;
  int rc;
  if ((rc = check_penalties(pen, &p))) return rc;
  if ((rc = check_uf(c, uf))) return rc;
  return run_pairs(c, uf, (const uint2*)d_pairs, n_pairs, p, min_match_len, false);
}

// Host half of the batched CIGAR walk: RLE string -> run list (count << 3 | op). Adjacent 'M'/'=' ops
// are merged (they continue one match run, src/seqrush.rs:1214-1218); a missing count is 1 (:1262-1264);
// any other character is an op that only flushes the run (:1436-1437).
static void parse_cigar_runs(const char* cg, std::vector<unsigned long long>& runs) {
  unsigned long long count = 0;
  const size_t first = runs.size();
  for (const char* p = cg; *p; ++p) {
    const char ch = *p;
    if (ch >= '0' && ch <= '9') {
      count = count * 10 + (unsigned long long)(ch - '0');
      continue;
    }
    if (count == 0) count = 1;
    const unsigned op = (ch == 'M' || ch == '=') ? 0u : ch == 'X' ? 1u : ch == 'I' ? 2u : ch == 'D' ? 3u : 4u;
    if (op == 0u && runs.size() > first && (runs.back() & 7ull) == 0ull) runs.back() += count << 3;
    else runs.push_back((count << 3) | op);
    count = 0;
  }
}

extern "C" int srgpu_process_alignments(srgpu_ctx* c, srgpu_uf* uf, const srgpu_paf_record* recs, uint64_t n,
                                        uint64_t min_match_len) {
  if (!c || (!recs && n)) return SRGPU_E_ARG;
  std::lock_guard<std::mutex> lk(c->mu);
  CK(cudaSetDevice(c->device));
  int rc;
  if ((rc = check_uf(c, uf))) return rc;
  if (n == 0) return SRGPU_OK;
  if (n > 0xfffffff0ull) {
    g_err = "too many records in one call";
    return SRGPU_E_ARG;
  }
  std::vector<PafRecordDev> h(n);
  std::vector<unsigned long long> runs;
  for (uint64_t i = 0; i < n; ++i) {
    const srgpu_paf_record& r = recs[i];
    if (r.query_idx >= c->n_seqs || r.target_idx >= c->n_seqs || !r.cigar) {
      g_err = "record " + std::to_string(i) + ": sequence index out of range or missing CIGAR";
      return SRGPU_E_ARG;
    }
    const size_t off = runs.size();
    parse_cigar_runs(r.cigar, runs);
    if (runs.size() - off > 0xfffffff0ull) {
      g_err = "CIGAR with too many operations";
      return SRGPU_E_ARG;
    }
    h[i].q = r.query_idx;
    h[i].t = r.target_idx;
    h[i].query_is_rc = r.query_is_rc ? 1u : 0u;
    h[i].n_runs = (uint32_t)(runs.size() - off);
    h[i].qstart = r.query_start;
    h[i].tstart = r.target_start;
    h[i].run_off = off;
  }
  if ((rc = c->d_paf_recs.ensure(n)) || (rc = c->d_paf_runs.ensure(runs.size() + 1))) return rc;
  CK(cudaMemcpyAsync(c->d_paf_recs.p, h.data(), n * sizeof(PafRecordDev), cudaMemcpyHostToDevice, c->stream));
  CK(cudaMemcpyAsync(c->d_paf_runs.p, runs.data(), runs.size() * 8, cudaMemcpyHostToDevice, c->stream));
  PafArgs a;
  a.seqs = seq_table(c);
  a.recs = c->d_paf_recs.p;
  a.runs = c->d_paf_runs.p;
  a.n_recs = (uint32_t)n;
  a.parent = uf->parent.p;
  a.min_match_len = min_match_len;
  a.counters = uf->counters.p;
  paf_walk_kernel<256><<<(unsigned)std::min<uint64_t>(n, (uint64_t)c->n_sm * 8), 256, 0, c->stream>>>(a);
  CK(cudaGetLastError());
  CK(cudaStreamSynchronize(c->stream));  // h / runs are pageable host memory: keep them alive until the copies are done
  c->stats.kernel_launches++;
  unsigned long long cnt[2];
  CK(cudaMemcpy(cnt, uf->counters.p, sizeof cnt, cudaMemcpyDeviceToHost));
  c->stats.unite_ops = cnt[0];
  c->stats.unite_links = cnt[1];
  return SRGPU_OK;
}

extern "C" int srgpu_process_alignment(srgpu_ctx* c, srgpu_uf* uf, const char* cigar, uint32_t qi, uint32_t ti,
                                       uint64_t min_match_len, int query_is_rc, uint64_t qstart, uint64_t tstart) {
  if (!c || !cigar) return SRGPU_E_ARG;
  srgpu_paf_record r;
  r.query_idx = qi;
  r.target_idx = ti;
  r.query_is_rc = query_is_rc;
  r.reserved = 0;
  r.query_start = qstart;
  r.target_start = tstart;
  r.cigar = cigar;
  return srgpu_process_alignments(c, uf, &r, 1, min_match_len);
}

// ---------------------------------------------------------------------------
// BidirectedUnionFind
// ---------------------------------------------------------------------------
extern "C" int srgpu_uf_new(srgpu_ctx* c, uint64_t max_offset, srgpu_uf** out) {
  if (!c || !out) return SRGPU_E_ARG;
  std::lock_guard<std::mutex> lk(c->mu);
  CK(cudaSetDevice(c->device));
  const uint64_t cap = (max_offset << 1) + 2;  // src/bidirected_union_find.rs:20
  if (cap > 0xfffffff0ull) {
    g_err = "union-find capacity exceeds the 32-bit parent array";
    return SRGPU_E_ARG;
  }
  srgpu_uf* u = new srgpu_uf();
  u->ctx = c;
  u->capacity = cap;
  int rc;
  if ((rc = u->parent.ensure(cap)) || (rc = u->counters.ensure(8))) {
    delete u;
    return rc;
  }
  uf_init_kernel<<<grid_for(cap, 256, c->n_sm), 256, 0, c->stream>>>(u->parent.p, cap);
  CK(cudaGetLastError());
  CK(cudaMemsetAsync(u->counters.p, 0, 8 * sizeof(unsigned long long), c->stream));
  CK(cudaStreamSynchronize(c->stream));
  {
    std::lock_guard<std::mutex> rk(g_registry_mu);
    c->ufs.push_back(u);
  }
  *out = u;
  return SRGPU_OK;
}

extern "C" void srgpu_uf_free(srgpu_uf* u) {
  if (!u) return;
  srgpu_ctx* c;
  {
    std::lock_guard<std::mutex> rk(g_registry_mu);
    c = u->ctx;
    if (c) c->ufs.erase(std::remove(c->ufs.begin(), c->ufs.end(), u), c->ufs.end());
  }
  if (c) {
    std::lock_guard<std::mutex> lk(c->mu);
    cudaSetDevice(c->device);
    cudaStreamSynchronize(c->stream);
    orphan_uf(u);
  }
  delete u;
}

extern "C" uint64_t srgpu_uf_capacity(srgpu_uf* u) { return u ? u->capacity : 0; }

// UFRush::clear (uf_rush-0.2.1/src/lib.rs:211-216): every Pos back to its own singleton set
extern "C" int srgpu_uf_clear(srgpu_uf* u) {
  if (!u) return SRGPU_E_ARG;
  UF_ALIVE(u);
  srgpu_ctx* c = u->ctx;
  std::lock_guard<std::mutex> lk(c->mu);
  CK(cudaSetDevice(c->device));
  uf_init_kernel<<<grid_for(u->capacity, 256, c->n_sm), 256, 0, c->stream>>>(u->parent.p, u->capacity);
  CK(cudaGetLastError());
  CK(cudaMemsetAsync(u->counters.p, 0, 8 * sizeof(unsigned long long), c->stream));
  CK(cudaStreamSynchronize(c->stream));
  u->seq_identity = 0;
  return SRGPU_OK;
}

extern "C" int srgpu_uf_roots(srgpu_uf* u, uint64_t first, uint64_t n, uint64_t* out) {
  if (!u || (!out && n)) return SRGPU_E_ARG;
  UF_ALIVE(u);
  srgpu_ctx* c = u->ctx;
  std::lock_guard<std::mutex> lk(c->mu);
  CK(cudaSetDevice(c->device));
  if (first + n > u->capacity) {
    g_err = "Pos out of range";
    return SRGPU_E_RANGE;
  }
  if (n == 0) return SRGPU_OK;
  int rc;
  if ((rc = u->tmp64.ensure(n))) return rc;
  uf_roots_kernel<<<grid_for(n, 256, c->n_sm), 256, 0, c->stream>>>(u->parent.p, first, n, u->tmp64.p);
  CK(cudaGetLastError());
  CK(cudaMemcpyAsync(out, u->tmp64.p, n * 8, cudaMemcpyDeviceToHost, c->stream));
  CK(cudaStreamSynchronize(c->stream));
  return SRGPU_OK;
}

extern "C" int srgpu_uf_find(srgpu_uf* u, uint64_t pos, uint64_t* rep) {
  if (!rep) return SRGPU_E_ARG;
  return srgpu_uf_roots(u, pos, 1, rep);
}

extern "C" int srgpu_uf_same(srgpu_uf* u, uint64_t a, uint64_t b, int* same) {
  if (!u || !same) return SRGPU_E_ARG;
  UF_ALIVE(u);
  if (a == b) {  // src/bidirected_union_find.rs:47-49
    if (a >= u->capacity) {
      g_err = "Pos out of range";
      return SRGPU_E_RANGE;
    }
    *same = 1;
    return SRGPU_OK;
  }
  uint64_t ra = 0, rb = 0;
  int rc;
  if ((rc = srgpu_uf_find(u, a, &ra)) || (rc = srgpu_uf_find(u, b, &rb))) return rc;
  *same = (ra == rb);
  return SRGPU_OK;
}

static int regions_run(srgpu_uf* u, const RegionDev* regs, uint64_t n) {
  UF_ALIVE(u);
  srgpu_ctx* c = u->ctx;
  std::lock_guard<std::mutex> lk(c->mu);
  CK(cudaSetDevice(c->device));
  if (n == 0) return SRGPU_OK;
  int rc;
  if ((rc = u->regions.ensure(n))) return rc;
  unsigned long long before = 0, after = 0;
  CK(cudaMemcpyAsync(&before, u->counters.p + 3, 8, cudaMemcpyDeviceToHost, c->stream));
  CK(cudaMemcpyAsync(u->regions.p, regs, n * sizeof(RegionDev), cudaMemcpyHostToDevice, c->stream));
  uf_regions_kernel<<<(unsigned)std::min<uint64_t>(n, (uint64_t)c->n_sm * 8), 256, 0, c->stream>>>(
      u->parent.p, u->regions.p, n, u->capacity, u->counters.p);
  CK(cudaGetLastError());
  CK(cudaMemcpyAsync(&after, u->counters.p + 3, 8, cudaMemcpyDeviceToHost, c->stream));
  CK(cudaStreamSynchronize(c->stream));
  if (after != before) {
    g_err = "Pos out of range in unite";
    return SRGPU_E_RANGE;
  }
  return SRGPU_OK;
}

__global__ void uf_unite_raw_kernel(uint32_t* parent, uint32_t a, uint32_t b, unsigned long long* counters) {
  atomicAdd(&counters[0], 1ull);
  if (uf_unite(parent, a, b)) atomicAdd(&counters[1], 1ull);
}

extern "C" int srgpu_uf_unite(srgpu_uf* u, uint64_t a, uint64_t b) {
  if (!u) return SRGPU_E_ARG;
  UF_ALIVE(u);
  if (a >= u->capacity || b >= u->capacity) {
    g_err = "Pos out of range";
    return SRGPU_E_RANGE;
  }
  if (a == b) return SRGPU_OK;  // src/bidirected_union_find.rs:40
  srgpu_ctx* c = u->ctx;
  std::lock_guard<std::mutex> lk(c->mu);
  CK(cudaSetDevice(c->device));
  uf_unite_raw_kernel<<<1, 1, 0, c->stream>>>(u->parent.p, (uint32_t)a, (uint32_t)b, u->counters.p);
  CK(cudaGetLastError());
  CK(cudaStreamSynchronize(c->stream));
  return SRGPU_OK;
}

extern "C" int srgpu_uf_unite_matching_region(srgpu_uf* u, uint64_t seq1_offset, uint64_t seq2_offset, uint64_t s1,
                                              uint64_t s2, uint64_t len, int seq1_is_rc, uint64_t seq1_len) {
  if (!u) return SRGPU_E_ARG;
  RegionDev r{seq1_offset, seq2_offset, s1, s2, len, seq1_len, seq1_is_rc ? 1 : 0, 0};
  return regions_run(u, &r, 1);
}
extern "C" int srgpu_uf_unite_matching_region_seq2_rc(srgpu_uf* u, uint64_t seq1_offset, uint64_t seq2_offset,
                                                      uint64_t s1, uint64_t s2, uint64_t len, int seq2_is_rc,
                                                      uint64_t seq2_len) {
  if (!u) return SRGPU_E_ARG;
  RegionDev r{seq1_offset, seq2_offset, s1, s2, len, seq2_len, seq2_is_rc ? 1 : 0, 1};
  return regions_run(u, &r, 1);
}
extern "C" int srgpu_uf_unite_regions(srgpu_uf* u, const srgpu_region* regions, uint64_t n) {
  if (!u || (!regions && n)) return SRGPU_E_ARG;
  std::vector<RegionDev> v(n);
  for (uint64_t i = 0; i < n; ++i) {
    v[i] = RegionDev{regions[i].seq1_offset, regions[i].seq2_offset, regions[i].seq1_local_start,
                     regions[i].seq2_local_start, regions[i].match_length, regions[i].seq1_len,
                     regions[i].seq1_is_rc ? 1 : 0, regions[i].seq2_rc_form ? 1 : 0};
  }
  return regions_run(u, v.data(), n);
}

extern "C" int srgpu_uf_unite_fr_all(srgpu_uf* u, uint64_t total_length) {
  if (!u) return SRGPU_E_ARG;
  UF_ALIVE(u);
  srgpu_ctx* c = u->ctx;
  std::lock_guard<std::mutex> lk(c->mu);
  CK(cudaSetDevice(c->device));
  if (2 * total_length > u->capacity) {
    g_err = "Pos out of range";
    return SRGPU_E_RANGE;
  }
  if (!total_length) return SRGPU_OK;
  uf_unite_fr_kernel<<<grid_for(total_length, 256, c->n_sm), 256, 0, c->stream>>>(u->parent.p, total_length, u->counters.p);
  CK(cudaGetLastError());
  CK(cudaStreamSynchronize(c->stream));
  return SRGPU_OK;
}

extern "C" int srgpu_uf_count_components(srgpu_uf* u, uint64_t total_length, uint64_t* count) {
  if (!u || !count) return SRGPU_E_ARG;
  UF_ALIVE(u);
  srgpu_ctx* c = u->ctx;
  std::lock_guard<std::mutex> lk(c->mu);
  CK(cudaSetDevice(c->device));
  if (2 * total_length > u->capacity) {
    g_err = "Pos out of range";
    return SRGPU_E_RANGE;
  }
  int rc;
  const size_t words = (u->capacity + 31) / 32;
  if ((rc = u->bitmap.ensure(words)) || (rc = u->tmp64.ensure(1))) return rc;
  CK(cudaMemsetAsync(u->bitmap.p, 0, words * 4, c->stream));
  CK(cudaMemsetAsync(u->tmp64.p, 0, 8, c->stream));
  if (total_length) {
    uf_count_kernel<<<grid_for(total_length, 256, c->n_sm), 256, 0, c->stream>>>(u->parent.p, total_length, u->bitmap.p,
                                                                              (unsigned long long*)u->tmp64.p);
    CK(cudaGetLastError());
  }
  CK(cudaMemcpyAsync(count, u->tmp64.p, 8, cudaMemcpyDeviceToHost, c->stream));
  CK(cudaStreamSynchronize(c->stream));
  return SRGPU_OK;
}

extern "C" int srgpu_uf_checksum(srgpu_uf* u, uint64_t* checksum) {
  if (!u || !checksum) return SRGPU_E_ARG;
  UF_ALIVE(u);
  srgpu_ctx* c = u->ctx;
  std::lock_guard<std::mutex> lk(c->mu);
  CK(cudaSetDevice(c->device));
  int rc;
  if ((rc = u->tmp64.ensure(1))) return rc;
  CK(cudaMemsetAsync(u->tmp64.p, 0, 8, c->stream));
  uf_checksum_kernel<<<grid_for(u->capacity, 256, c->n_sm), 256, 0, c->stream>>>(u->parent.p, u->capacity,
                                                                               (unsigned long long*)u->tmp64.p);
  CK(cudaGetLastError());
  CK(cudaMemcpyAsync(checksum, u->tmp64.p, 8, cudaMemcpyDeviceToHost, c->stream));
  CK(cudaStreamSynchronize(c->stream));
  return SRGPU_OK;
}

extern "C" void* srgpu_uf_device_ptr(srgpu_uf* u) { return u ? (void*)u->parent.p : nullptr; }

extern "C" int srgpu_uf_compress(srgpu_uf* u) {
  if (!u) return SRGPU_E_ARG;
  UF_ALIVE(u);
  srgpu_ctx* c = u->ctx;
  std::lock_guard<std::mutex> lk(c->mu);
  CK(cudaSetDevice(c->device));
  uf_compress_kernel<<<grid_for(u->capacity, 256, c->n_sm), 256, 0, c->stream>>>(u->parent.p, u->capacity);
  CK(cudaGetLastError());
  CK(cudaStreamSynchronize(c->stream));
  return SRGPU_OK;
}

extern "C" int srgpu_uf_compress_async(srgpu_uf* u) {
  if (!u) return SRGPU_E_ARG;
  UF_ALIVE(u);
  srgpu_ctx* c = u->ctx;
  std::lock_guard<std::mutex> lk(c->mu);
  CK(cudaSetDevice(c->device));
  uf_compress_kernel<<<grid_for(u->capacity, 256, c->n_sm), 256, 0, c->stream>>>(u->parent.p, u->capacity);
  CK(cudaGetLastError());
  return SRGPU_OK;
}

extern "C" int srgpu_uf_merge_gathered(srgpu_uf* u, const void* d_gathered, uint64_t n, int world, int my_rank) {
  if (!u || !d_gathered || world < 1 || my_rank < 0 || my_rank >= world) return SRGPU_E_ARG;
  UF_ALIVE(u);
  srgpu_ctx* c = u->ctx;
  std::lock_guard<std::mutex> lk(c->mu);
  CK(cudaSetDevice(c->device));
  if (n != u->capacity) {
    g_err = "gathered forests must have the capacity of this union-find";
    return SRGPU_E_ARG;
  }
  uf_merge_gathered_kernel<<<grid_for(n, 256, c->n_sm), 256, 0, c->stream>>>(u->parent.p, (const uint32_t*)d_gathered, n,
                                                                            world, my_rank, u->counters.p);
  CK(cudaGetLastError());
  return SRGPU_OK;
}

extern "C" int srgpu_uf_merge_roots(srgpu_uf* u, const void* d_peer_roots, uint64_t n) {
  if (!u || !d_peer_roots) return SRGPU_E_ARG;
  UF_ALIVE(u);
  srgpu_ctx* c = u->ctx;
  std::lock_guard<std::mutex> lk(c->mu);
  CK(cudaSetDevice(c->device));
  if (n > u->capacity) n = u->capacity;
  uf_merge_kernel<<<grid_for(n, 256, c->n_sm), 256, 0, c->stream>>>(u->parent.p, (const uint32_t*)d_peer_roots, n,
                                                                   u->counters.p);
  CK(cudaGetLastError());
  CK(cudaStreamSynchronize(c->stream));
  return SRGPU_OK;
}

// ---------------------------------------------------------------------------
// Graph build from the partition (row f1): device passes (graph.cuh) + host compaction / writer
// (graph_host.hpp). Reference: src/bidirected_builder.rs:17-289, src/bidirected_ops.rs:75-478,880-925,
// src/bidirected_gfa_writer.rs:8-167 with --skip-sgd --skip-groom --skip-topo.
// ---------------------------------------------------------------------------
struct srgpu_graph {
  HostGraph g;
  std::vector<std::vector<uint8_t>> seq_copy;  // the input sequences (for validate_paths_match_sequences)
  bool compacted = false;
};

extern "C" int srgpu_graph_build(srgpu_ctx* c, srgpu_uf* uf, const uint64_t* representatives, srgpu_graph** out) {
  if (!c || !out) return SRGPU_E_ARG;
  std::lock_guard<std::mutex> lk(c->mu);
  CK(cudaSetDevice(c->device));
  int rc;
  if ((rc = check_uf(c, uf))) return rc;
  const uint64_t n = c->total_len;
  if (n == 0 || n > 0x7ffffff0ull) {
    g_err = "graph build: empty sequence set or too many positions for 32-bit node handles";
    return SRGPU_E_ARG;
  }
  DevBuf<uint32_t> d_comp, d_flag, d_rank, d_steps, d_bsum, d_first;
  DevBuf<uint8_t> d_base;
  DevBuf<unsigned long long> d_err, d_rep, d_keys, d_cnt;
  DevBuf<uint2> d_edges;
  struct Rel {
    std::vector<std::function<void()>> f;
    ~Rel() { for (auto& x : f) x(); }
  } rel;
#define GB(buf, cnt)                       \
  if ((rc = (buf).ensure(cnt))) return rc; \
  rel.f.push_back([&]() { (buf).release(); });
  const uint64_t nblk = (n + SCAN_BLOCK - 1) / SCAN_BLOCK;
  GB(d_comp, n) GB(d_flag, n) GB(d_rank, n) GB(d_steps, n) GB(d_bsum, nblk + 1) GB(d_err, 4) GB(d_cnt, 2)
  if (representatives) {
    GB(d_rep, n)
    CK(cudaMemcpyAsync(d_rep.p, representatives, n * 8, cudaMemcpyHostToDevice, c->stream));
  }
  CK(cudaMemsetAsync(d_err.p, 0, 4 * 8, c->stream));
  CK(cudaMemsetAsync(d_cnt.p, 0, 2 * 8, c->stream));
  GraphArgs a;
  a.seqs = seq_table(c);
  a.parent = uf->parent.p;
  a.total_len = n;
  a.rep_in = representatives ? d_rep.p : nullptr;
  a.comp = d_comp.p;
  a.first_flag = d_flag.p;
  a.rank = d_rank.p;
  a.steps = d_steps.p;
  a.node_base = nullptr;
  a.err = d_err.p;
  const int grid = grid_for(n, 256, c->n_sm);
  graph_roots_kernel<<<grid, 256, 0, c->stream>>>(a);
  CK(cudaGetLastError());
  scan_block_sums_kernel<<<(unsigned)nblk, SCAN_THREADS, 0, c->stream>>>(d_flag.p, n, d_bsum.p);
  CK(cudaGetLastError());
  scan_sums_kernel<<<1, 1024, 0, c->stream>>>(d_bsum.p, nblk, d_cnt.p);
  CK(cudaGetLastError());
  scan_apply_kernel<<<(unsigned)nblk, SCAN_THREADS, 0, c->stream>>>(d_flag.p, n, d_bsum.p, d_rank.p);
  CK(cudaGetLastError());
  unsigned long long n_nodes = 0, errs[2] = {0, 0};
  CK(cudaMemcpyAsync(&n_nodes, d_cnt.p, 8, cudaMemcpyDeviceToHost, c->stream));
  CK(cudaMemcpyAsync(errs, d_err.p, 8, cudaMemcpyDeviceToHost, c->stream));
  CK(cudaStreamSynchronize(c->stream));
  if (errs[0]) {
    g_err = "graph build: " + std::to_string(errs[0]) + " positions whose forward and reverse Pos are not united "
            "(the builder needs the invariant of SeqRush::new, src/seqrush.rs:324-328: call srgpu_uf_unite_fr_all first)";
    return SRGPU_E_ARG;
  }
  GB(d_base, n_nodes + 1)
  a.node_base = d_base.p;
  graph_steps_kernel<<<grid, 256, 0, c->stream>>>(a);
  CK(cudaGetLastError());
  // edge set: open-addressing table of the canonical forms, first occurrence wins
  uint64_t cap = 1024;
  while (cap < 2 * n) cap <<= 1;
  GB(d_keys, cap) GB(d_first, cap) GB(d_edges, n)
  CK(cudaMemsetAsync(d_keys.p, 0xff, cap * 8, c->stream));
  CK(cudaMemsetAsync(d_first.p, 0xff, cap * 4, c->stream));
  EdgeArgs e;
  e.seqs = a.seqs;
  e.steps = d_steps.p;
  e.total_len = n;
  e.keys = d_keys.p;
  e.first_at = d_first.p;
  e.cap = cap;
  e.out = d_edges.p;
  e.n_out = d_cnt.p + 1;
  graph_edges_insert_kernel<<<grid, 256, 0, c->stream>>>(e);
  CK(cudaGetLastError());
  graph_edges_collect_kernel<<<grid_for(cap, 256, c->n_sm), 256, 0, c->stream>>>(e);
  CK(cudaGetLastError());
  c->stats.kernel_launches += 7;
  unsigned long long n_edges = 0;
  CK(cudaMemcpyAsync(&n_edges, d_cnt.p + 1, 8, cudaMemcpyDeviceToHost, c->stream));
  CK(cudaMemcpyAsync(errs, d_err.p, 16, cudaMemcpyDeviceToHost, c->stream));
  CK(cudaStreamSynchronize(c->stream));
  if (errs[1]) {
    g_err = "graph build: " + std::to_string(errs[1]) + " representatives lie outside the component of their position";
    return SRGPU_E_ARG;
  }
  std::vector<uint8_t> h_base(n_nodes);
  std::vector<uint32_t> h_steps(n);
  std::vector<uint2> h_edges(n_edges);
  if (n_nodes) CK(cudaMemcpy(h_base.data(), d_base.p, n_nodes, cudaMemcpyDeviceToHost));
  CK(cudaMemcpy(h_steps.data(), d_steps.p, n * 4, cudaMemcpyDeviceToHost));
  if (n_edges) CK(cudaMemcpy(h_edges.data(), d_edges.p, n_edges * sizeof(uint2), cudaMemcpyDeviceToHost));
  std::vector<uint8_t> h_fwd(c->d_fwd.n);
  CK(cudaMemcpy(h_fwd.data(), c->d_fwd.p, c->d_fwd.n, cudaMemcpyDeviceToHost));
  std::vector<uint64_t> h_boff(c->n_seqs);
  CK(cudaMemcpy(h_boff.data(), c->d_byte_off.p, (size_t)c->n_seqs * 8, cudaMemcpyDeviceToHost));
#undef GB
  srgpu_graph* gr = new srgpu_graph();
  HostGraph& g = gr->g;
  g.seq.resize(n_nodes + 1);
  g.present.assign(n_nodes + 1, 1);
  g.present[0] = 0;
  for (uint64_t i = 0; i < n_nodes; ++i) g.seq[i + 1].assign(1, (char)h_base[i]);
  g.uncompacted_nodes = n_nodes;
  g.edges.reserve(n_edges * 2);
  for (const uint2& ed : h_edges) g.edges.insert(gedge(ed.x, ed.y));
  g.paths.resize(c->n_seqs);
  gr->seq_copy.resize(c->n_seqs);
  for (uint32_t s = 0; s < c->n_seqs; ++s) {
    const uint64_t o = c->h_pos_off[s], l = c->h_len[s];
    g.paths[s].assign(h_steps.begin() + o, h_steps.begin() + o + l);
    gr->seq_copy[s].assign(h_fwd.begin() + h_boff[s], h_fwd.begin() + h_boff[s] + l);
  }
  *out = gr;
  return SRGPU_OK;
}

// The same object from caller-supplied parts (no device involved): the uncompacted graph as node bases,
// per-path handle lists and stored edges. Lets a host keep its own builder and still use compact / validate /
// write, and lets the CPU test-suite check those against the restated reference.
extern "C" int srgpu_graph_from_parts(uint64_t n_nodes, const uint8_t* node_bases, uint32_t n_paths, const uint64_t* path_lens,
                                      const uint64_t* steps, uint64_t n_edges, const uint64_t* from_to,
                                      const srgpu_seqview* seqs, srgpu_graph** out) {
  if (!out || (n_nodes && !node_bases) || (n_paths && (!path_lens || !steps || !seqs)) || (n_edges && !from_to)) return SRGPU_E_ARG;
  srgpu_graph* gr = new srgpu_graph();
  HostGraph& g = gr->g;
  g.seq.resize(n_nodes + 1);
  g.present.assign(n_nodes + 1, 1);
  g.present[0] = 0;
  for (uint64_t i = 0; i < n_nodes; ++i) g.seq[i + 1].assign(1, (char)node_bases[i]);
  g.uncompacted_nodes = n_nodes;
  for (uint64_t i = 0; i < n_edges; ++i) g.edges.insert(gedge((GHandle)from_to[2 * i], (GHandle)from_to[2 * i + 1]));
  g.paths.resize(n_paths);
  gr->seq_copy.resize(n_paths);
  uint64_t at = 0;
  for (uint32_t p = 0; p < n_paths; ++p) {
    g.paths[p].resize(path_lens[p]);
    for (uint64_t i = 0; i < path_lens[p]; ++i) {
      const uint64_t h = steps[at + i];
      if ((h >> 1) == 0 || (h >> 1) > n_nodes) {
        delete gr;
        g_err = "path step refers to a node that does not exist";
        return SRGPU_E_ARG;
      }
      g.paths[p][i] = (GHandle)h;
    }
    at += path_lens[p];
    gr->seq_copy[p].assign(seqs[p].data, seqs[p].data + seqs[p].len);
  }
  *out = gr;
  return SRGPU_OK;
}

extern "C" int srgpu_graph_compact(srgpu_graph* gr) {
  if (!gr) return SRGPU_E_ARG;
  gr->g.compact();
  gr->g.renumber();
  gr->compacted = true;
  return SRGPU_OK;
}
extern "C" uint64_t srgpu_graph_node_count(srgpu_graph* gr) { return gr ? gr->g.node_count() : 0; }
extern "C" uint64_t srgpu_graph_edge_count(srgpu_graph* gr) { return gr ? gr->g.edges.size() : 0; }
extern "C" uint64_t srgpu_graph_path_count(srgpu_graph* gr) { return gr ? gr->g.paths.size() : 0; }
extern "C" uint64_t srgpu_graph_path_length(srgpu_graph* gr, uint32_t path) {
  return (gr && path < gr->g.paths.size()) ? gr->g.paths[path].size() : 0;
}
extern "C" int srgpu_graph_path_steps(srgpu_graph* gr, uint32_t path, uint64_t* out, uint64_t cap) {
  if (!gr || path >= gr->g.paths.size() || (!out && cap)) return SRGPU_E_ARG;
  const auto& st = gr->g.paths[path];
  for (uint64_t i = 0; i < st.size() && i < cap; ++i) out[i] = st[i];
  return SRGPU_OK;
}
extern "C" int64_t srgpu_graph_node_sequence(srgpu_graph* gr, uint64_t node_id, char* buf, uint64_t cap) {
  if (!gr || node_id >= gr->g.seq.size() || !gr->g.present[node_id]) return -1;
  const std::string& s = gr->g.seq[node_id];
  if (buf) memcpy(buf, s.data(), std::min<uint64_t>(cap, s.size()));
  return (int64_t)s.size();
}
extern "C" int srgpu_graph_edges(srgpu_graph* gr, uint64_t* from_to, uint64_t cap_edges) {
  if (!gr || (!from_to && cap_edges)) return SRGPU_E_ARG;
  std::vector<uint64_t> es(gr->g.edges.begin(), gr->g.edges.end());
  std::sort(es.begin(), es.end());
  for (uint64_t i = 0; i < es.size() && i < cap_edges; ++i) {
    from_to[2 * i] = es[i] >> 32;
    from_to[2 * i + 1] = es[i] & 0xffffffffull;
  }
  return SRGPU_OK;
}
// validate_paths_match_sequences + write_gfa; *text is malloc'd (srgpu_free)
extern "C" int srgpu_graph_gfa_text(srgpu_graph* gr, const char* const* names, uint32_t n_names, char** text, uint64_t* len) {
  if (!gr || !names || !text || n_names != gr->g.paths.size()) return SRGPU_E_ARG;
  std::vector<const uint8_t*> data(n_names);
  std::vector<uint64_t> lens(n_names);
  for (uint32_t i = 0; i < n_names; ++i) {
    data[i] = gr->seq_copy[i].data();
    lens[i] = gr->seq_copy[i].size();
  }
  const long bad = gr->g.validate(data, lens);
  if (bad >= 0) {  // the reference returns Err("Path '<id>' does not match original sequence") and aborts the write
    g_err = std::string("Path '") + names[bad] + "' does not match original sequence";
    return SRGPU_E_INTERNAL;
  }
  std::vector<std::string> nm(n_names);
  for (uint32_t i = 0; i < n_names; ++i) nm[i] = names[i] ? names[i] : "";
  std::string t;
  gr->g.gfa_text(nm, t);
  char* o = (char*)malloc(t.size() + 1);
  if (!o) {
    g_err = "out of host memory for the GFA text";
    return SRGPU_E_INTERNAL;
  }
  memcpy(o, t.c_str(), t.size() + 1);
  *text = o;
  if (len) *len = t.size();
  return SRGPU_OK;
}
extern "C" int srgpu_graph_write_gfa(srgpu_graph* gr, const char* const* names, uint32_t n_names, const char* path) {
  if (!path) return SRGPU_E_ARG;
  char* t = nullptr;
  uint64_t n = 0;
  const int rc = srgpu_graph_gfa_text(gr, names, n_names, &t, &n);
  if (rc) return rc;
  FILE* f = fopen(path, "wb");
  if (!f) {
    free(t);
    g_err = std::string("cannot create ") + path;
    return SRGPU_E_ARG;
  }
  const size_t w = fwrite(t, 1, n, f);
  fclose(f);
  free(t);
  if (w != n) {
    g_err = "short write";
    return SRGPU_E_INTERNAL;
  }
  return SRGPU_OK;
}
extern "C" int srgpu_graph_stats(srgpu_graph* gr, uint64_t* uncompacted_nodes, uint32_t* rounds, uint32_t* merged_chains) {
  if (!gr) return SRGPU_E_ARG;
  if (uncompacted_nodes) *uncompacted_nodes = gr->g.uncompacted_nodes;
  if (rounds) *rounds = gr->g.rounds;
  if (merged_chains) *merged_chains = gr->g.merged_chains;
  return SRGPU_OK;
}
extern "C" void srgpu_graph_free(srgpu_graph* gr) { delete gr; }

// ---------------------------------------------------------------------------
// Pair-list sparsification (rows f2 / a5): MinHash sketches + Jaccard counts and the hashed pair filter on the
// device (sparsify.cuh); reference allwave/src/mash.rs:30-165, allwave/src/knn_graph.rs:10-172,
// allwave/src/iterator.rs:49-77,256-284.
// ---------------------------------------------------------------------------
extern "C" uint64_t srgpu_rust_hash_bytes(const uint8_t* data, uint64_t n) {  // DefaultHasher over a &[u8]
  Sip13 h;
  h.init();
  h.word(n);
  for (uint64_t i = 0; i < n; ++i) h.byte(data[i]);
  return h.finish();
}
extern "C" uint64_t srgpu_rust_hash_str(const char* s) {  // DefaultHasher over a String / &str
  Sip13 h;
  h.init();
  for (const char* p = s; *p; ++p) h.byte((unsigned char)*p);
  h.byte(0xff);
  return h.finish();
}

extern "C" int srgpu_mash_sketch_jaccard(srgpu_ctx* c, uint32_t k, uint32_t sketch_size, uint64_t* sketches_out,
                                         uint32_t* counts_out, uint32_t* inter_out, uint32_t* uni_out) {
  if (!c || k == 0 || sketch_size == 0 || sketch_size > (uint32_t)SKETCH_MAX) {
    g_err = "mash sketch: k must be > 0 and 0 < sketch_size <= 4096";
    return SRGPU_E_ARG;
  }
  std::lock_guard<std::mutex> lk(c->mu);
  CK(cudaSetDevice(c->device));
  const uint32_t n = c->n_seqs;
  if (n == 0) return SRGPU_OK;
  std::vector<unsigned long long> off(n + 1, 0);
  for (uint32_t i = 0; i < n; ++i) off[i + 1] = off[i] + (c->h_len[i] >= k ? c->h_len[i] - k + 1 : 0);
  DevBuf<unsigned long long> d_hash, d_off, d_sk;
  DevBuf<unsigned int> d_cnt, d_skcnt, d_inter, d_uni;
  int rc = 0;
  auto cleanup = [&]() {
    d_hash.release(); d_off.release(); d_sk.release(); d_cnt.release(); d_skcnt.release(); d_inter.release(); d_uni.release();
  };
  if ((rc = d_hash.ensure(off[n] + 1)) || (rc = d_off.ensure(n + 1)) || (rc = d_sk.ensure((size_t)n * sketch_size)) ||
      (rc = d_cnt.ensure(n)) || (rc = d_skcnt.ensure(n)) || (rc = d_inter.ensure((size_t)n * n)) || (rc = d_uni.ensure((size_t)n * n))) {
    cleanup();
    return rc;
  }
  cudaError_t e = cudaMemcpyAsync(d_off.p, off.data(), (n + 1) * 8, cudaMemcpyHostToDevice, c->stream);
  if (e == cudaSuccess) e = cudaMemsetAsync(d_cnt.p, 0, n * 4, c->stream);
  if (e == cudaSuccess) e = cudaMemsetAsync(d_sk.p, 0, (size_t)n * sketch_size * 8, c->stream);
  if (e == cudaSuccess) e = cudaMemsetAsync(d_inter.p, 0, (size_t)n * n * 4, c->stream);
  if (e == cudaSuccess) e = cudaMemsetAsync(d_uni.p, 0, (size_t)n * n * 4, c->stream);
  if (e == cudaSuccess) {
    dim3 grid(std::max(1u, std::min(256u, (c->max_len + 255) / 256)), std::min(n, 65535u));
    kmer_hash_kernel<<<grid, 256, 0, c->stream>>>(seq_table(c), k, d_off.p, d_hash.p, d_cnt.p);
    e = cudaGetLastError();
  }
  if (e == cudaSuccess) {
    sketch_select_kernel<256><<<std::min<uint32_t>(n, (uint32_t)c->n_sm * 4), 256, 0, c->stream>>>(d_hash.p, d_off.p, d_cnt.p, n, sketch_size,
                                                                                         d_sk.p, d_skcnt.p);
    e = cudaGetLastError();
  }
  if (e == cudaSuccess) {
    jaccard_kernel<<<grid_for((uint64_t)n * n, 128, c->n_sm), 128, 0, c->stream>>>(d_sk.p, d_skcnt.p, n, sketch_size, d_inter.p, d_uni.p);
    e = cudaGetLastError();
  }
  if (e == cudaSuccess && sketches_out) e = cudaMemcpyAsync(sketches_out, d_sk.p, (size_t)n * sketch_size * 8, cudaMemcpyDeviceToHost, c->stream);
  if (e == cudaSuccess && counts_out) e = cudaMemcpyAsync(counts_out, d_skcnt.p, n * 4, cudaMemcpyDeviceToHost, c->stream);
  if (e == cudaSuccess && inter_out) e = cudaMemcpyAsync(inter_out, d_inter.p, (size_t)n * n * 4, cudaMemcpyDeviceToHost, c->stream);
  if (e == cudaSuccess && uni_out) e = cudaMemcpyAsync(uni_out, d_uni.p, (size_t)n * n * 4, cudaMemcpyDeviceToHost, c->stream);
  if (e == cudaSuccess) e = cudaStreamSynchronize(c->stream);
  cleanup();
  if (e != cudaSuccess) {
    g_err = std::string("mash sketch: ") + cudaGetErrorString(e);
    return SRGPU_E_CUDA;
  }
  c->stats.kernel_launches += 3;
  return SRGPU_OK;
}

extern "C" int srgpu_pair_hash_filter(srgpu_ctx* c, const char* const* ids, uint32_t n, double fraction, int exclude_self,
                                      uint8_t* keep_out, uint64_t* hashes_out) {
  if (!c || !ids || !keep_out) return SRGPU_E_ARG;
  std::lock_guard<std::mutex> lk(c->mu);
  CK(cudaSetDevice(c->device));
  if (n == 0) return SRGPU_OK;
  std::vector<unsigned long long> off(n + 1, 0);
  std::string cat;
  for (uint32_t i = 0; i < n; ++i) {
    if (!ids[i]) return SRGPU_E_ARG;
    cat += ids[i];
    off[i + 1] = cat.size();
  }
  DevBuf<uint8_t> d_ids, d_keep;
  DevBuf<unsigned long long> d_off, d_h;
  int rc = 0;
  auto cleanup = [&]() { d_ids.release(); d_keep.release(); d_off.release(); d_h.release(); };
  if ((rc = d_ids.ensure(cat.size() + 1)) || (rc = d_keep.ensure((size_t)n * n)) || (rc = d_off.ensure(n + 1)) ||
      (hashes_out && (rc = d_h.ensure((size_t)n * n)))) {
    cleanup();
    return rc;
  }
  cudaError_t e = cudaMemcpyAsync(d_ids.p, cat.data(), cat.size(), cudaMemcpyHostToDevice, c->stream);
  if (e == cudaSuccess) e = cudaMemcpyAsync(d_off.p, off.data(), (n + 1) * 8, cudaMemcpyHostToDevice, c->stream);
  if (e == cudaSuccess) {
    pair_hash_kernel<<<grid_for((uint64_t)n * n, 256, c->n_sm), 256, 0, c->stream>>>(d_ids.p, d_off.p, n, fraction, exclude_self, d_keep.p,
                                                                               hashes_out ? d_h.p : nullptr);
    e = cudaGetLastError();
  }
  if (e == cudaSuccess) e = cudaMemcpyAsync(keep_out, d_keep.p, (size_t)n * n, cudaMemcpyDeviceToHost, c->stream);
  if (e == cudaSuccess && hashes_out) e = cudaMemcpyAsync(hashes_out, d_h.p, (size_t)n * n * 8, cudaMemcpyDeviceToHost, c->stream);
  if (e == cudaSuccess) e = cudaStreamSynchronize(c->stream);
  cleanup();
  if (e != cudaSuccess) {
    g_err = std::string("pair hash filter: ") + cudaGetErrorString(e);
    return SRGPU_E_CUDA;
  }
  c->stats.kernel_launches += 1;
  return SRGPU_OK;
}

// ---------------------------------------------------------------------------
// Allocation-free variant of srgpu_uf_count_components. No counterpart in the reference: added
// for callers that cannot call cudaMalloc (CUDA graph capture, tight serving loops, harnesses
// that forbid it). The scratch of uf_count_kernel -- one 64-bit counter and a bitmap with one bit per Pos -- lives
// in a file-scope __device__ array of fixed size: every GPU that loads the library holds these 4 MiB + 256 B until
// the process exits. All calls share the array and it keeps the last call's contents, so each call clears the part
// it uses; two calls that overlap on the GPU (different streams / contexts of one device) race on it.
// ---------------------------------------------------------------------------
static constexpr uint64_t STATIC_BITMAP_WORDS = 1ull << 20;                 // 4 MiB: one bit per Pos
static constexpr uint64_t STATIC_MAX_CAPACITY = STATIC_BITMAP_WORDS * 32;   // 33 554 432 Pos = 16 777 215 bases
static __device__ __align__(256) uint32_t g_count_scratch[64 + STATIC_BITMAP_WORDS];  // [0..63]: counter (256 B), then the bitmap

// In host code `g_count_scratch` is a host-side stand-in, not a device address: the device address is looked up once
// per device, with that device current, and cached.
static std::mutex g_scratch_mu;
static uint32_t* g_scratch_addr[64] = {nullptr};
static int count_scratch_address(int device, uint32_t** out) {
  if (device < 0 || device >= 64) {
    g_err = "device index out of range for the static scratch table";
    return SRGPU_E_ARG;
  }
  std::lock_guard<std::mutex> lk(g_scratch_mu);
  if (!g_scratch_addr[device]) {
    void* p = nullptr;
    CK(cudaSetDevice(device));
    CK(cudaGetSymbolAddress(&p, g_count_scratch));
    g_scratch_addr[device] = (uint32_t*)p;
  }
  *out = g_scratch_addr[device];
  return 0;
}

extern "C" uint64_t srgpu_uf_count_components_static_max_capacity(void) { return STATIC_MAX_CAPACITY; }

extern "C" int srgpu_uf_count_components_static(srgpu_uf* u, uint64_t total_length, uint64_t* count) {
  if (!u || !count) return SRGPU_E_ARG;
  UF_ALIVE(u);
  srgpu_ctx* c = u->ctx;
  std::lock_guard<std::mutex> lk(c->mu);
  CK(cudaSetDevice(c->device));
  if (2 * total_length > u->capacity) {
    g_err = "Pos out of range";
    return SRGPU_E_RANGE;
  }
  if (u->capacity > STATIC_MAX_CAPACITY) {
    char b[200];
    snprintf(b, sizeof b, "union-find of %llu Pos exceeds the static scratch of srgpu_uf_count_components_static "
             "(%llu Pos = 4 MiB bitmap); use srgpu_uf_count_components", (unsigned long long)u->capacity,
             (unsigned long long)STATIC_MAX_CAPACITY);
    g_err = b;
    return SRGPU_E_ARG;
  }
  uint32_t* scratch = nullptr;
  int rc;
  if ((rc = count_scratch_address(c->device, &scratch))) return rc;
  unsigned long long* d_count = (unsigned long long*)scratch;
  uint32_t* d_bitmap = scratch + 64;
  const size_t words = (u->capacity + 31) / 32;
  CK(cudaMemsetAsync(scratch, 0, (64 + words) * 4, c->stream));  // the array keeps the last call's contents
  if (total_length) {
    uf_count_kernel<<<grid_for(total_length, 256, c->n_sm), 256, 0, c->stream>>>(u->parent.p, total_length, d_bitmap, d_count);
    CK(cudaGetLastError());
  }
  CK(cudaMemcpyAsync(count, d_count, 8, cudaMemcpyDeviceToHost, c->stream));
  CK(cudaStreamSynchronize(c->stream));
  return SRGPU_OK;
}
