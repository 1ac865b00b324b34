// wb_emul.cpp — host lane emulator of the wide-band aligner (CPU test harness, not product code).
//
// Runs the SAME per-lane functions the CUDA kernel runs (csrc/dp_wb.cuh) over 32 emulated lanes with the
// warp shuffles replaced by array copies, stores the direction words in the kernel's layout and walks them
// back with the one-thread walkers, so that the skewed wavefront, the 16-step block layout, the packed
// arg-max keys and the padded-diagonal trick can be checked against the oracle where there is no GPU.
// Built by tests/conftest.py with g++.
#include <stdint.h>
#include <string.h>
#include <vector>
#include "../../sequence-aligner_b200/csrc/dp_wb.cuh"
#include "../../include/hovl.h"

using namespace hovl;

static int acgt(uint8_t c) {
  switch (c) { case 'A': return 0; case 'C': return 1; case 'G': return 2; case 'T': return 3; default: return -1; }
}
static std::vector<uint32_t> pack(const uint8_t *s, int32_t n, bool *bad) {
  std::vector<uint32_t> w((size_t)(n + 15) / 16 + 2, 0u);
  for (int32_t i = 0; i < n; i++) {
    int v = acgt(s[i]);
    if (v < 0) { *bad = true; v = 0; }
    w[(size_t)i >> 4] |= (uint32_t)v << (30 - 2 * (i & 15));
  }
  return w;
}

template <int D>
static void flush(std::vector<uint32_t> &sc, int32_t blk, int lane, const uint32_t dir[D]) {
  constexpr int DS = wb::DirStride<D>::value;
  for (int c = 0; c < D; c++) sc[((size_t)blk * 32 + (size_t)lane) * DS + (size_t)c] = dir[c];
}

template <int D>
static void run_pair(const hovl_settings *s, const uint8_t *A, int32_t n, const uint8_t *B, int32_t m,
                     hovl_record *rec, uint64_t *cells) {
  DpOut r; memset(&r, 0, sizeof r);
  *cells = 0;
  bool bad = false;
  std::vector<uint32_t> pa = pack(A, n, &bad), pb = pack(B, m, &bad);
  const int32_t width = band_width(n, s->k, s->min_identity_f32);
  auto store = [&]() {
    rec->flags = r.flags; rec->start_i = r.start_i; rec->start_j = r.start_j; rec->end_i = r.end_i;
    rec->end_j = r.end_j; rec->correct = r.correct; rec->error = r.error; rec->score = r.score;
    rec->ahg = r.ahg; rec->bhg = r.bhg;
  };
  if (bad) { dp_set_dud(r, HOVL_CORE_F_BAD_BASE); store(); return; }
  if (m < width) { dp_set_dud(r, HOVL_CORE_F_SHORT_B); store(); return; }
  if (32 * D < width + 1) { dp_set_dud(r, 0); r.flags = -1; store(); return; }   // class too small
  alignas(16) int32_t costu[16];
  int32_t cost16[16];
  for (int i = 0; i < 16; i++) { cost16[i] = s->cost[i / 4][i % 4]; costu[i] = wb::UNIT * cost16[i] + 2; }
  if (!wb::fits(cost16, s->gap_open, s->gap_extend, n > m ? n : m)) { dp_set_dud(r, 0); r.flags = -2; store(); return; }
  wb::Consts K; K.costu = costu; K.ge = wb::UNIT * s->gap_extend; K.gox = wb::UNIT * s->gap_open + 1; K.goy = wb::UNIT * s->gap_open;
  K.one = 1; K.neg1 = -1; K.four = 4;
  const int32_t nwa = (n + 15) >> 4, nwb = (m + 15) >> 4;
  std::vector<uint32_t> sc(wb::scratch_words(n > m ? n : m, D), 0xDEADBEEFu);
  dp_set_dud(r, 0);

  // ---------------- phase 1 ----------------
  int32_t ds = -1;
  {
    std::vector<wb::P1<D>> L(32);
    std::vector<wb::Win> aw(32);
    for (int l = 0; l < 32; l++) { wb::p1_init<D>(L[l], l, pb.data(), m); aw[l].init(pa.data(), nwa, -l); }
    const int32_t steps = n + 31;
    for (int32_t st = 1; st <= steps; st++) {
      int32_t p_hx[32], p_hm[32];
      for (int l = 0; l < 32; l++) { p_hx[l] = L[l].o_hx; p_hm[l] = L[l].o_hm; }            // = __shfl_up
      for (int l = 0; l < 32; l++) {
        const int32_t i = st - l;
        const uint32_t a16 = (aw[l].bits() >> 26) & 0x30u;
        if (i >= 1 && i <= n) wb::p1_row<D>(L[l], K, a16, (15 - ((st - 1) & 15)) << 2, l ? p_hx[l - 1] : 0, l ? p_hm[l - 1] : 0);
        else for (int c = 0; c < D; c++) L[l].dir[c] <<= 2;
        aw[l].advance();
        if ((st & 15) == 0) { flush<D>(sc, (st >> 4) - 1, l, L[l].dir); wb::block_best<D>(L[l].key, L[l].best, (st >> 4) - 1, l, width); }
      }
    }
    int32_t score = 0, bi = 0, bc = 0;
    for (int l = 0; l < 32; l++) {
      if (steps & 15) {
        for (int c = 0; c < D; c++) L[l].dir[c] <<= 2 * (16 - (steps & 15));
        flush<D>(sc, steps >> 4, l, L[l].dir);
        wb::block_best<D>(L[l].key, L[l].best, steps >> 4, l, width);
      }
      int32_t s1, r1, c1;
      wb::best_cell(L[l].best, l, s1, r1, c1);
      if (l == 0 || wb::best_before(s1, r1, c1, score, bi, bc)) { score = s1; bi = r1; bc = c1; }
    }
    *cells += (uint64_t)n * (uint64_t)width;
    if (score > 0) ds = wb::walk1<D>(sc.data(), bi, bc + 1);
  }
  if (ds < 0) { store(); return; }

  // ---------------- phase 2 ----------------
  const int32_t dl = n - ds, z = width / 2;
  std::vector<wb::P2<D>> L(32);
  std::vector<wb::Win> aw(32), bw(32);
  for (int l = 0; l < 32; l++) {
    wb::p2_init<D>(L[l], K, l, width);
    aw[l].init(pa.data(), nwa, ds - l);
    bw[l].init(pb.data(), nwb, l * D - z - l);
  }
  const int32_t steps = dl + 31;
  const int32_t lo = z + 1, hi = (m - width + z) < dl ? (m - width + z) : dl;
  for (int32_t st = 1; st <= steps; st++) {
    int32_t p_hx[32];
    wb::P2Row row[32];
    for (int l = 0; l < 32; l++) p_hx[l] = L[l].o_hx;                                          // = __shfl_up
    const bool interior = st - 31 >= lo && st <= hi;
    for (int l = 0; l < 32; l++) {
      const int32_t u = st - l;
      wb::P2Row &r2 = row[l];
      r2.a16 = (aw[l].bits() >> 26) & 0x30u; r2.bb = bw[l].bits(); r2.rinv = (15 - ((st - 1) & 15)) << 2;
      r2.j0 = l * D - z + u; r2.m = m; r2.kmax = width - l * D;
      const bool active = u >= 1 && u <= dl;
      if (interior) wb::p2_first<D, false>(L[l], K, r2, l ? p_hx[l - 1] : 0, 0);
      else if (active) wb::p2_first<D, true>(L[l], K, r2, l ? p_hx[l - 1] : 0, 0);
    }
    int32_t down[33];
    for (int l = 0; l < 32; l++) down[l] = L[l].hy[0];                                          // = __shfl_down
    down[32] = 0;
    for (int l = 0; l < 32; l++) {
      const int32_t u = st - l;
      const bool active = u >= 1 && u <= dl;
      if (interior) wb::p2_rest<D, false>(L[l], K, row[l], down[l + 1]);
      else if (active) wb::p2_rest<D, true>(L[l], K, row[l], down[l + 1]);
      else for (int c = 0; c < D; c++) L[l].dir[c] <<= 2;
      aw[l].advance(); bw[l].advance();
      if ((st & 15) == 0) { flush<D>(sc, (st >> 4) - 1, l, L[l].dir); wb::block_best<D>(L[l].key, L[l].best, (st >> 4) - 1, l, width + 1); }
    }
  }
  int32_t score = 0, bu = 0, bk = 0;
  for (int l = 0; l < 32; l++) {
    if (steps & 15) {
      for (int c = 0; c < D; c++) L[l].dir[c] <<= 2 * (16 - (steps & 15));
      flush<D>(sc, steps >> 4, l, L[l].dir);
      wb::block_best<D>(L[l].key, L[l].best, steps >> 4, l, width + 1);
    }
    int32_t s1, r1, c1;
    wb::best_cell(L[l].best, l, s1, r1, c1);
    if (l == 0 || wb::best_before(s1, r1, c1, score, bu, bk)) { score = s1; bu = r1; bk = c1; }
  }
  *cells += (uint64_t)(dl + 1) * (uint64_t)(width + 1);
  if (score <= 0) { store(); return; }
  int32_t tu, kk, cc, ee;
  wb::walk2<D>(sc.data(), pa.data(), pb.data(), ds, z, bu, bk, tu, kk, cc, ee);
  r.start_i = tu + ds; r.start_j = kk - z + tu;
  r.end_i = bu + ds;   r.end_j = bk - z + bu;
  r.correct = cc; r.error = ee; r.score = score >> wb::SH;
  dp_finish(r, n, m, s->min_overlap, s->min_identity_f32, s->max_ignore_f32);
  store();
}

extern "C" int wb_emul_align_pairs(const hovl_settings *s, const uint8_t *bases, const uint64_t *offsets,
                                   const uint32_t *pa, const uint32_t *pb, uint64_t np, int D,
                                   hovl_record *out, uint64_t *cells_out) {
  uint64_t total = 0;
  for (uint64_t p = 0; p < np; p++) {
    const uint32_t a = pa[p], b = pb[p];
    const uint8_t *A = bases + offsets[a - 1], *B = bases + offsets[b - 1];
    const int32_t n = (int32_t)(offsets[a] - offsets[a - 1]), m = (int32_t)(offsets[b] - offsets[b - 1]);
    hovl_record rec; memset(&rec, 0, sizeof rec);
    uint64_t cells = 0;
    switch (D) {
#define HOVL_WB_EMUL(D_) case D_: run_pair<D_>(s, A, n, B, m, &rec, &cells); break;
      HOVL_WB_CLASSES(HOVL_WB_EMUL)
#undef HOVL_WB_EMUL
      default: return -1;
    }
    rec.a = (int32_t)a; rec.b = (int32_t)b; rec.count = out[p].count;
    out[p] = rec; total += cells;
  }
  if (cells_out) *cells_out = total;
  return 0;
}
