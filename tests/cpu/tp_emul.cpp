// tp_emul.cpp — host harness for the thread-per-pair DP (CPU test harness, not product code).
//
// Runs the SAME per-thread functions the CUDA kernels run (csrc/dp_tp.cuh: packed
// score+priority cells, 2-bit direction codes, backward walk) with the warp-interleaved
// scratch replaced by a plain array (stride 1), so that the whole narrow-band aligner can be
// checked against the oracle here, where there is no GPU.  Built by tests/conftest.py with g++.
#include <stdint.h>
#include <string.h>
#include <vector>
#include "../../sequence-aligner_b200/csrc/dp_tp.cuh"
#include "../../include/hovl.h"

using namespace hovl;

static int acgt(uint8_t c) {
  switch (c) { case 'A': return 0; case 'C': return 1; case 'G': return 2; case 'T': return 3; default: return -1; }
}
static std::vector<uint32_t> pack(const uint8_t *s, int32_t n, bool *bad) {
  std::vector<uint32_t> w((size_t)(n + 15) / 16 + 1, 0u);
  for (int32_t i = 0; i < n; i++) {
    int v = acgt(s[i]);
    if (v < 0) { *bad = true; v = 0; }
    w[(size_t)i >> 4] |= (uint32_t)v << (30 - 2 * (i & 15));
  }
  return w;
}

template <int W, int WMIN, bool FAST>
static void run_class(const tp::Consts &K, const hovl_settings *s, const uint32_t *wa, int32_t n, const uint32_t *wb,
                      int32_t m, int32_t width, DpOut &r, uint64_t *cells) {
  std::vector<uint32_t> sc(tp::scratch_words(n, W), 0xdeadbeefu);
  *cells += (uint64_t)n * (uint64_t)width;
  const int32_t ds = tp::phase1<W, WMIN, 1, FAST>(K, wa, n, wb, m, width, sc.data());
  if (ds < 0) return;
  *cells += (uint64_t)(n - ds + 1) * (uint64_t)(width + 1);
  tp::phase2<W, WMIN, 1, FAST>(K, wa, n, wb, m, width, ds, sc.data(), s->min_overlap, s->min_identity_f32,
                               s->max_ignore_f32, r);
}

// fast != 0: packed arg-max keys (reads <= 1023 bases), as the kernels choose for such inputs
extern "C" int tp_emul_align_pairs(const hovl_settings *s, const uint8_t *bases, const uint64_t *offsets,
                                   const uint32_t *pa, const uint32_t *pb, uint64_t np, int fast, hovl_record *out,
                                   uint64_t *cells_out) {
  uint64_t total = 0;
  const int sh = fast ? 12 : 2;
  int32_t cost4[16];
  for (int i = 0; i < 16; i++) cost4[i] = s->cost[i / 4][i % 4] * (1 << sh) + 2;
  tp::Consts K; K.cost4 = cost4; K.ge4 = s->gap_extend * (1 << sh); K.gox = s->gap_open * (1 << sh) + 1;
  K.goy = s->gap_open * (1 << sh); K.one = 1; K.neg1 = -1; K.four = 4;
  for (uint64_t p = 0; p < np; p++) {
    const uint32_t a = pa[p], b = pb[p];
    const uint8_t *A = bases + offsets[a - 1], *B = bases + offsets[b - 1];
    const int32_t n = (int32_t)(offsets[a] - offsets[a - 1]), m = (int32_t)(offsets[b] - offsets[b - 1]);
    bool bad = false;
    std::vector<uint32_t> wa = pack(A, n, &bad), wb = pack(B, m, &bad);
    const int32_t width = band_width(n, s->k, s->min_identity_f32);
    DpOut r; dp_set_dud(r, 0);
    uint64_t cells = 0;
    if (bad) dp_set_dud(r, HOVL_CORE_F_BAD_BASE);
    else if (m < width) dp_set_dud(r, HOVL_CORE_F_SHORT_B);
    else if (width < tp::TP_MIN_WIDTH || width > tp::TP_MAX_WIDTH) return -2;     // not a thread-per-pair class
    else if (fast && n > tp::TP_FAST_MAX_LEN) return -3;
#define HOVL_TP_RUN(W_, WMIN_) else if (width <= W_) { \
      if (fast) run_class<W_, WMIN_, true>(K, s, wa.data(), n, wb.data(), m, width, r, &cells); \
      else run_class<W_, WMIN_, false>(K, s, wa.data(), n, wb.data(), m, width, r, &cells); }
    HOVL_TP_CLASSES(HOVL_TP_RUN)
#undef HOVL_TP_RUN
    hovl_record rec;
    rec.a = (int32_t)a; rec.b = (int32_t)b; rec.count = out[p].count;
    rec.flags = r.flags; rec.start_i = r.start_i; rec.start_j = r.start_j; rec.end_i = r.end_i; rec.end_j = r.end_j;
    rec.correct = r.correct; rec.error = r.error; rec.score = r.score; rec.ahg = r.ahg; rec.bhg = r.bhg;
    out[p] = rec; total += cells;
  }
  if (cells_out) *cells_out = total;
  return 0;
}
