// ASan/UBSan driver over the host emulation of the aligner's control logic (development aid)
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <random>
#include <string>
#include <vector>
extern "C" int emul_wfa_align(const char* pattern, int plen, const char* text, int tlen, int x, int o1, int e1, int o2,
                              int e2, int W, int cap, int lbits, char* cigar_out, int* cigar_len, int* score);
extern "C" int emul_wfa_align_split(const char* pattern, int plen, const char* text, int tlen, int x, int o1, int e1, int o2, int e2, int W, int cap, int lbits, int split_T, int leaf_T2, char* cigar_out, int* cigar_len, int* score, int* n_tasks_out, int* n_leaves_out);
extern "C" int emul_orient(const char* pf, const char* pr, int plen, const char* text, int tlen, int lbits, int* dist);
static std::mt19937 rng(12345);
static std::string randseq(int n) { std::string s(n, 'A'); for (auto& c : s) c = "ACGT"[rng() & 3]; return s; }
static std::string mutate(const std::string& a, double rate) {
  std::string o;
  std::uniform_real_distribution<double> u(0, 1);
  for (size_t i = 0; i < a.size(); ++i) {
    const double r = u(rng);
    if (r < rate / 3) { o += "ACGT"[rng() & 3]; }                       // substitution
    else if (r < 2 * rate / 3) { /* deletion */ }
    else if (r < rate) { o += a[i]; int l = 1 + (rng() % 8 == 0 ? rng() % 60 : 0); for (int j = 0; j < l; ++j) o += "ACGT"[rng() & 3]; }
    else o += a[i];
  }
  if (o.empty()) o = "A";
  return o;
}
int main(int argc, char** argv) {
  const int n = argc > 1 ? atoi(argv[1]) : 200;
  int ok = 0, ovf = 0, other = 0;
  for (int it = 0; it < n; ++it) {
    const int L = 200 + rng() % 5000;
    std::string a = randseq(L), b;
    const int kind = it % 5;
    if (kind == 4) b = randseq(100 + rng() % 3000);              // unrelated
    else b = mutate(a, kind == 0 ? 0.005 : kind == 1 ? 0.02 : kind == 2 ? 0.08 : 0.2);
    std::vector<char> cig(a.size() + b.size() + 16);
    int clen = 0, score = 0;
    // widths: automatic (0), production-like narrow, deliberately too narrow, full
    const int wsel = (it / 5) % 4;
    const int fullW = ((2 * ((int)std::max(a.size(), b.size()) + 16)) + 63) & ~63;
    const int W = wsel == 0 ? fullW : wsel == 1 ? 1024 : wsel == 2 ? 256 : fullW;
    const int st = emul_wfa_align(a.data(), (int)a.size(), b.data(), (int)b.size(), 5, 8, 2, 24, 1, W, 520, 0, cig.data(), &clen, &score);
    if (st == 0) ++ok; else if (st == 2) ++ovf; else { ++other; fprintf(stderr, "it %d status %d\n", it, st); }
    if (st == 0 && W == fullW) {
      std::vector<char> cig2(a.size() + b.size() + 80);
      int clen2 = 0, score2 = 0, nt = 0, nl = 0;
      const int T = 250 + (rng() % 4) * 250;
      const int st2 = emul_wfa_align_split(a.data(), (int)a.size(), b.data(), (int)b.size(), 5, 8, 2, 24, 1, W, 520, 0, T, 250, cig2.data(), &clen2, &score2, &nt, &nl);
      if (st2 != 0 || clen2 != clen || memcmp(cig.data(), cig2.data(), clen) != 0 || score2 != score) { ++other; fprintf(stderr, "it %d split differs: st %d len %d/%d score %d/%d tasks %d leaves %d\n", it, st2, clen2, clen, score2, score, nt, nl); }
    }
  }
  printf("pairs %d ok %d ring-overflow %d other %d\n", n, ok, ovf, other);
  return other ? 1 : 0;
}
