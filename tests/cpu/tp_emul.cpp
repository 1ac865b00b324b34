// tp_emul.cpp — host harness for the thread-per-pair DP (CPU test harness, not product code).
//
// Runs the SAME per-thread functions the CUDA kernels run (csrc/dp_tp.cuh: packed
// score+priority cells, 2-bit direction codes, backward walk) with the warp-interleaved
// scratch replaced by a plain array (stride 1), so that the whole narrow-band aligner can be
// checked against the oracle here, where there is no GPU.  Built by tests/conftest.py with g++.
#include <stdint.h>
#include <string.h>
#include <vector>
#define HOVL_TP2_COUNT_WRAPS 1
#include "../../sequence-aligner_b200/csrc/dp_tp2.cuh"
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

// ---- two pairs per thread (csrc/dp_tp2.cuh): pairs of one class are taken two at a time, in order.
// Phase 1 always runs packed; phase 2 runs packed (moving frames) when pack2 != 0 and as the 32-bit
// dp_tp.cuh phase 2 otherwise, the way dp.cu chooses per phase.
template <int W, int WMIN>
static void run_class2(const tp2::Consts2 &K, const tp::Consts &K1, const hovl_settings *s, tp2::Pair2 q0, tp2::Pair2 q1,
                       int32_t max_len, int pack2, DpOut &r0, DpOut &r1) {
  std::vector<uint32_t> sc(tp2::scratch_words(max_len, W) + tp::scratch_words(max_len, W), 0xdeadbeefu);
  std::vector<uint32_t> st(tp2::state_words(W), 0xdeadbeefu);
  int32_t ds[2];
  tp2::phase1x2<W, WMIN, 1, 1>(K, q0, q1, max_len, sc.data(), st.data(), ds);
  q0.ds = ds[0]; q1.ds = ds[1];
  if (!pack2) {
    if (ds[0] >= 0) tp::phase2<W, WMIN, 1, false>(K1, q0.wa, q0.n, q0.wb, q0.m, q0.width, ds[0], sc.data(), s->min_overlap,
                                                  s->min_identity_f32, s->max_ignore_f32, r0);
    if (ds[1] >= 0) tp::phase2<W, WMIN, 1, false>(K1, q1.wa, q1.n, q1.wb, q1.m, q1.width, ds[1], sc.data(), s->min_overlap,
                                                  s->min_identity_f32, s->max_ignore_f32, r1);
    return;
  }
  if (ds[0] < 0 && ds[1] < 0) return;
  // a pair without a phase 2 lends its half word to a copy of the other (as the kernel does)
  DpOut unused; dp_set_dud(unused, 0);
  std::fill(sc.begin(), sc.end(), 0xdeadbeefu);
  const tp2::Pair2 &x = ds[0] < 0 ? q1 : q0, &y = ds[1] < 0 ? q0 : q1;
  tp2::phase2x2<W, WMIN, 1, 1>(K, x, y, max_len, sc.data(), st.data(), s->min_overlap, s->min_identity_f32,
                               s->max_ignore_f32, ds[0] < 0 ? unused : r0, ds[1] < 0 ? unused : r1);
}

extern "C" int tp2_emul_align_pairs(const hovl_settings *s, const uint8_t *bases, const uint64_t *offsets,
                                    const uint32_t *pa, const uint32_t *pb, uint64_t np, int pack2, hovl_record *out) {
  uint32_t cost2[256];
  for (int i = 0; i < 256; i++) {
    const int a1 = i >> 6, a0 = (i >> 4) & 3, b1 = (i >> 2) & 3, b0 = i & 3;
    cost2[i] = tp2::pack16(s->cost[a1][b1] * 4 + 2, s->cost[a0][b0] * 4 + 2);
  }
  tp2::Consts2 K; K.cost2 = cost2; K.ge = tp2::pack16(s->gap_extend * 4, s->gap_extend * 4);
  K.gox = tp2::pack16(s->gap_open * 4 + 1, s->gap_open * 4 + 1); K.goy = tp2::pack16(s->gap_open * 4, s->gap_open * 4);
  K.one = 1; K.neg1 = -1; K.four = 4; K.tgt = 0; K.floor = 0;
  int32_t cost4[16], flat[16];
  for (int i = 0; i < 16; i++) { flat[i] = s->cost[i / 4][i % 4]; cost4[i] = flat[i] * 4 + 2; }
  tp::Consts K1; K1.cost4 = cost4; K1.ge4 = s->gap_extend * 4; K1.gox = s->gap_open * 4 + 1; K1.goy = s->gap_open * 4;
  K1.one = 1; K1.neg1 = -1; K1.four = 4;
  struct Item { std::vector<uint32_t> wa, wb; int32_t n, m, width; DpOut r; bool run; };
  std::vector<Item> items((size_t)np);
  int32_t max_len = 1;
  for (uint64_t p = 0; p < np; p++) {
    const uint32_t a = pa[p], b = pb[p];
    Item &it = items[(size_t)p];
    it.n = (int32_t)(offsets[a] - offsets[a - 1]); it.m = (int32_t)(offsets[b] - offsets[b - 1]);
    bool bad = false;
    it.wa = pack(bases + offsets[a - 1], it.n, &bad); it.wb = pack(bases + offsets[b - 1], it.m, &bad);
    it.width = band_width(it.n, s->k, s->min_identity_f32);
    dp_set_dud(it.r, 0); it.run = false;
    if (bad) dp_set_dud(it.r, HOVL_CORE_F_BAD_BASE);
    else if (it.m < it.width) dp_set_dud(it.r, HOVL_CORE_F_SHORT_B);
    else if (it.width < tp::TP_MIN_WIDTH || it.width > tp::TP_MAX_WIDTH) return -2;
    else it.run = true;
    if (it.n > max_len) max_len = it.n;
    if (it.m > max_len) max_len = it.m;
  }
  if (!tp2::fits16(flat, s->gap_open, s->gap_extend, tp::TP_MAX_WIDTH)) return -4;
  tp2::g_wraps = 0;
#define HOVL_TP2_RUN(W_, WMIN_) { \
    std::vector<size_t> idx; \
    for (size_t p = 0; p < items.size(); p++) if (items[p].run && items[p].width >= WMIN_ && items[p].width <= W_) idx.push_back(p); \
    if (pack2 && !idx.empty() && !tp2::frame_fits(flat, s->gap_open, s->gap_extend, W_, &K.tgt, &K.floor)) return -5; \
    for (size_t t = 0; t < idx.size(); t += 2) { \
      Item &x = items[idx[t]]; Item &y = items[idx[t + 1 < idx.size() ? t + 1 : t]]; \
      tp2::Pair2 q0{x.wa.data(), x.wb.data(), x.n, x.m, x.width, 0}, q1{y.wa.data(), y.wb.data(), y.n, y.m, y.width, 0}; \
      DpOut r0, r1; dp_set_dud(r0, 0); dp_set_dud(r1, 0); \
      run_class2<W_, WMIN_>(K, K1, s, q0, q1, max_len, pack2, r0, r1); \
      x.r = r0; if (t + 1 < idx.size()) y.r = r1; \
    } }
  HOVL_TP_CLASSES(HOVL_TP2_RUN)
#undef HOVL_TP2_RUN
  for (uint64_t p = 0; p < np; p++) {
    const DpOut &r = items[(size_t)p].r;
    hovl_record rec;
    rec.a = (int32_t)pa[p]; rec.b = (int32_t)pb[p]; rec.count = out[p].count;
    rec.flags = r.flags; rec.start_i = r.start_i; rec.start_j = r.start_j; rec.end_i = r.end_i; rec.end_j = r.end_j;
    rec.correct = r.correct; rec.error = r.error; rec.score = r.score; rec.ahg = r.ahg; rec.bhg = r.bhg;
    out[p] = rec;
  }
  return 0;
}

// half-word sums that wrapped during the last tp2_emul_align_pairs call (a frame budget violation)
extern "C" int tp2_emul_wraps(void) { return tp2::g_wraps; }

// ---- phase 1 by exact-prefix search (tp::probe_prefix) against the phase-1 DP of the same header.
// ds_dp[p] = doveStart of the DP (-1 dud, -2 not run), ds_probe[p] = the probe's (-1 not found, -2 not run).
extern "C" int tp_emul_probe(const hovl_settings *s, const uint8_t *bases, const uint64_t *offsets, const uint32_t *pa,
                             const uint32_t *pb, uint64_t np, int32_t *ds_dp, int32_t *ds_probe) {
  int32_t cost4[16], flat[16];
  for (int i = 0; i < 16; i++) { flat[i] = s->cost[i / 4][i % 4]; cost4[i] = flat[i] * 4 + 2; }
  tp::Consts K; K.cost4 = cost4; K.ge4 = s->gap_extend * 4; K.gox = s->gap_open * 4 + 1; K.goy = s->gap_open * 4;
  K.one = 1; K.neg1 = -1; K.four = 4;
  const bool valid = tp::probe_valid(flat, s->gap_extend);
  for (uint64_t p = 0; p < np; p++) {
    const uint32_t a = pa[p], b = pb[p];
    const int32_t n = (int32_t)(offsets[a] - offsets[a - 1]), m = (int32_t)(offsets[b] - offsets[b - 1]);
    bool bad = false;
    std::vector<uint32_t> wa = pack(bases + offsets[a - 1], n, &bad), wb = pack(bases + offsets[b - 1], m, &bad);
    const int32_t width = band_width(n, s->k, s->min_identity_f32);
    ds_dp[p] = -2; ds_probe[p] = -2;
    if (bad || m < width || width < tp::TP_MIN_WIDTH || width > tp::TP_MAX_WIDTH) continue;
#define HOVL_TP_P1(W_, WMIN_) else if (width <= W_) { \
      std::vector<uint32_t> sc(tp::scratch_words(n, W_), 0xdeadbeefu); \
      ds_dp[p] = tp::phase1<W_, WMIN_, 1, false>(K, wa.data(), n, wb.data(), m, width, sc.data()); }
    if (false) {}
    HOVL_TP_CLASSES(HOVL_TP_P1)
#undef HOVL_TP_P1
    ds_probe[p] = tp::probe_prefix(wa.data(), n, wb.data(), m, width);
  }
  return valid ? 1 : 0;
}
