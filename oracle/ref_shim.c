/*
 * oracle/ref_shim.c -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 *
 * Thin C entry points around the reference's own native arithmetic, compiled
 * from the sources where they lie under /root/reference (never copied):
 *   - WFA2-lib (C)   : wavefront_align() in BiWFA / classic mode
 *       reference call site: allwave/src/alignment.rs:189-276 (perform_wfa_alignment)
 *       -> WFA2-lib/wavefront/wavefront_align.c:212-241
 *   - edlib (C++)    : edlibAlign() NW distance, k=-1
 *       reference call site: allwave/src/alignment.rs:148-171 (determine_orientation_edlib)
 *       -> edlib_rs-0.1.2/edlib-c/edlib/src/edlib.cpp:141-212
 *
 * Built by oracle/Makefile into oracle/_ref/libsrref_<isa>.so. Only tests/,
 * __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs
 * may load it.
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include "wavefront/wavefront_align.h"
#include "alignment/cigar.h"

typedef struct {
  wavefront_aligner_t* wf;
  int x, o1, e1, o2, e2, mode;
} ref_wfa_t;

/* memory_mode: 0 = high (classic WFA), 3 = ultralow (BiWFA, what allwave uses) */
void* ref_wfa_new(int mismatch, int o1, int e1, int o2, int e2, int memory_mode) {
  wavefront_aligner_attr_t attr = wavefront_aligner_attr_default;
  attr.distance_metric = gap_affine_2p;
  attr.affine2p_penalties.match = 0;
  attr.affine2p_penalties.mismatch = mismatch;
  attr.affine2p_penalties.gap_opening1 = o1;
  attr.affine2p_penalties.gap_extension1 = e1;
  attr.affine2p_penalties.gap_opening2 = o2;
  attr.affine2p_penalties.gap_extension2 = e2;
  attr.alignment_scope = compute_alignment;
  attr.alignment_form.span = alignment_end2end;
  attr.heuristic.strategy = wf_heuristic_none;
  attr.memory_mode = (memory_mode == 3) ? wavefront_memory_ultralow : wavefront_memory_high;
  ref_wfa_t* h = (ref_wfa_t*)calloc(1, sizeof(ref_wfa_t));
  h->wf = wavefront_aligner_new(&attr);
  return h;
}
/* Any of allwave's three modes (allwave/src/alignment.rs:203-231): two_piece = 1 -> gap-affine-2p, 0 -> gap-affine
 * (single piece; allwave's "edit distance" mode is this with o = e = x). match <= 0 (WFA2-lib exits on match > 0,
 * wavefront/wavefront_penalties.c:100-103). */
void* ref_wfa_new_mode(int match, int mismatch, int o1, int e1, int o2, int e2, int two_piece, int memory_mode) {
  wavefront_aligner_attr_t attr = wavefront_aligner_attr_default;
  if (two_piece) {
    attr.distance_metric = gap_affine_2p;
    attr.affine2p_penalties.match = match;
    attr.affine2p_penalties.mismatch = mismatch;
    attr.affine2p_penalties.gap_opening1 = o1;
    attr.affine2p_penalties.gap_extension1 = e1;
    attr.affine2p_penalties.gap_opening2 = o2;
    attr.affine2p_penalties.gap_extension2 = e2;
  } else {
    attr.distance_metric = gap_affine;
    attr.affine_penalties.match = match;
    attr.affine_penalties.mismatch = mismatch;
    attr.affine_penalties.gap_opening = o1;
    attr.affine_penalties.gap_extension = e1;
  }
  attr.alignment_scope = compute_alignment;
  attr.alignment_form.span = alignment_end2end;
  attr.heuristic.strategy = wf_heuristic_none;
  attr.memory_mode = (memory_mode == 3) ? wavefront_memory_ultralow : wavefront_memory_high;
  ref_wfa_t* h = (ref_wfa_t*)calloc(1, sizeof(ref_wfa_t));
  h->wf = wavefront_aligner_new(&attr);
  return h;
}
void ref_wfa_free(void* hv) {
  ref_wfa_t* h = (ref_wfa_t*)hv;
  if (!h) return;
  wavefront_aligner_delete(h->wf);
  free(h);
}
/* Returns WFA status (0 == completed). cigar_out must hold plen+tlen bytes.
 * *score_out is the raw cigar->score field (keeps the INT32_MIN quirk of
 * WFA2-lib/wavefront/wavefront_bialign.c:1214-1219 + alignment/cigar.c:79-89). */
int ref_wfa_align(void* hv, const char* pattern, int plen, const char* text, int tlen,
                  char* cigar_out, int* cigar_len_out, int* score_out) {
  ref_wfa_t* h = (ref_wfa_t*)hv;
  const int status = wavefront_align(h->wf, pattern, plen, text, tlen);
  cigar_t* const c = h->wf->cigar;
  const int n = c->end_offset - c->begin_offset;
  if (cigar_out && n > 0) memcpy(cigar_out, c->operations + c->begin_offset, (size_t)n);
  if (cigar_len_out) *cigar_len_out = n > 0 ? n : 0;
  if (score_out) *score_out = c->score;
  return status;
}
