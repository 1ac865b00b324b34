// dp_emul.cpp — host lane emulator for the banded DP (CPU test harness, not product code).
//
// Runs the SAME per-lane functions the CUDA kernel runs (csrc/dp_core.cuh) with the warp
// shuffles replaced by array copies between G emulated lanes, so that the skewed wavefront,
// the forward-carried trace tags and the arg-max tie-breaking can be checked against the
// oracle here, where there is no GPU.  Built by tests/conftest.py with g++.
#include <stdint.h>
#include <string.h>
#include <vector>
#include "../../sequence-aligner_b200/csrc/dp_core.cuh"
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
static inline int32_t clampi(int32_t x, int32_t lo, int32_t hi) { return x < lo ? lo : (x > hi ? hi : x); }

template <int D>
static void run_pair(const hovl_settings *s, const uint8_t *A, int32_t n, const uint8_t *B, int32_t m,
                     int G, hovl_record *rec, uint64_t *cells) {
  DpOut r; memset(&r, 0, sizeof r);
  *cells = 0;
  bool bad = false;
  std::vector<uint32_t> pa = pack(A, n, &bad), pb = pack(B, m, &bad);
  const int32_t width = band_width(n, s->k, s->min_identity_f32);
  DpConst K;
  int32_t cost16[16];
  for (int i = 0; i < 16; i++) cost16[i] = s->cost[i / 4][i % 4];
  K.cost = cost16;
  K.gO = s->gap_open; K.gE = s->gap_extend;
  auto store = [&]() {
    rec->flags = r.flags; rec->start_i = r.start_i; rec->start_j = r.start_j; rec->end_i = r.end_i;
    rec->end_j = r.end_j; rec->correct = r.correct; rec->error = r.error; rec->score = r.score;
    rec->ahg = r.ahg; rec->bhg = r.bhg;
  };
  if (bad) { dp_set_dud(r, HOVL_CORE_F_BAD_BASE); store(); return; }
  if (m < width) { dp_set_dud(r, HOVL_CORE_F_SHORT_B); store(); return; }
  if (G * D < width + 1) { dp_set_dud(r, 0); r.flags = -1; store(); return; }   // class too small

  // ---------------- phase 1 ----------------
  std::vector<P1Lane<D>> L1((size_t)G);
  std::vector<int32_t> o_hx((size_t)G, 0), o_hm((size_t)G, 0);
  std::vector<uint32_t> o_t((size_t)G);
  for (int l = 0; l < G; l++) { p1_init<D>(L1[l], l); o_t[l] = (uint32_t)(l * D + D); }
  uint32_t bcol[D];
  for (int32_t st = 1; st <= n + G - 1; st++) {
    std::vector<int32_t> p_hx = o_hx, p_hm = o_hm; std::vector<uint32_t> p_t = o_t;   // = __shfl_up
    for (int l = 0; l < G; l++) {
      const int32_t i = st - l;
      const uint32_t a = packed_base(pa.data(), clampi(i - 1, 0, n - 1));
      for (int c = 0; c < D; c++) bcol[c] = packed_base(pb.data(), clampi(l * D + c, 0, m - 1));
      int32_t in_hx = 0, in_hm = 0; uint32_t in_t = (uint32_t)i << 8;
      if (l > 0) { in_hx = p_hx[l - 1]; in_hm = p_hm[l - 1]; in_t = p_t[l - 1]; }
      p1_row<D>(L1[l], K, l, i, n, width, a, bcol, in_hx, in_hm, in_t, o_hx[l], o_hm[l], o_t[l]);
    }
  }
  *cells += (uint64_t)n * (uint64_t)width;
  int32_t best = 0, bi = 0, bj = 0; uint32_t bt = 0;
  for (int l = 0; l < G; l++) {
    const P1Lane<D> &q = L1[l];
    if (q.best > best || (q.best == best && best > 0 && (q.best_i < bi || (q.best_i == bi && q.best_j < bj)))) {
      best = q.best; bi = q.best_i; bj = q.best_j; bt = q.best_t;
    }
  }
  if (best <= 0 || (bt & 255u) != 0) { dp_set_dud(r, 0); store(); return; }

  // ---------------- phase 2 ----------------
  P2Geom g; g.ds = (int32_t)(bt >> 8); g.dl = n - g.ds; g.z = width / 2; g.width = width; g.m = m;
  std::vector<P2Lane<D>> L2((size_t)G);
  std::vector<Edge> up((size_t)G), down((size_t)G);
  Edge zero; zero.h = 0; zero.t1 = 0; zero.t2 = 0;
  for (int l = 0; l < G; l++) { p2_init<D>(L2[l], l); up[l] = zero; }
  uint32_t bd[D];
  for (int32_t st = 1; st <= g.dl + G - 1; st++) {
    std::vector<Edge> pu = up;                                                           // = __shfl_up
    for (int l = 0; l < G; l++) {
      const int32_t u = st - l;
      const uint32_t a = packed_base(pa.data(), clampi(g.ds + u - 1, 0, n - 1));
      const uint32_t b0 = packed_base(pb.data(), clampi(l * D - g.z + u - 1, 0, m - 1));
      p2_first<D>(L2[l], K, g, l, u, a, b0, l > 0 ? pu[l - 1] : zero);
    }
    for (int l = 0; l < G; l++) down[l] = p2_down<D>(L2[l]);                            // = __shfl_down
    for (int l = 0; l < G; l++) {
      const int32_t u = st - l;
      const uint32_t a = packed_base(pa.data(), clampi(g.ds + u - 1, 0, n - 1));
      for (int c = 0; c < D; c++) bd[c] = packed_base(pb.data(), clampi(l * D + c - g.z + u - 1, 0, m - 1));
      up[l] = p2_rest<D>(L2[l], K, g, l, u, a, bd, l + 1 < G ? down[l + 1] : zero);
    }
  }
  *cells += (uint64_t)(g.dl + 1) * (uint64_t)(width + 1);
  best = 0; int32_t bu = 0, bk = 0; uint32_t t1 = 0, t2 = 0;
  for (int l = 0; l < G; l++) {
    const P2Lane<D> &q = L2[l];
    if (q.best > best || (q.best == best && best > 0 && (q.best_u < bu || (q.best_u == bu && q.best_kk < bk)))) {
      best = q.best; bu = q.best_u; bk = q.best_kk; t1 = q.best_t1; t2 = q.best_t2;
    }
  }
  if (best <= 0) { dp_set_dud(r, 0); store(); return; }
  const int32_t su = (int32_t)(t1 >> 8), sk = (int32_t)(t1 & 255u);
  r.start_i = su + g.ds; r.start_j = sk - g.z + su;
  r.end_i = bu + g.ds;   r.end_j = bk - g.z + bu;
  r.correct = (int32_t)(t2 >> 16); r.error = (int32_t)(t2 & 0xffffu); r.score = best;
  dp_finish(r, n, m, s->min_overlap, s->min_identity_f32, s->max_ignore_f32);
  store();
}

extern "C" int emul_align_pairs(const hovl_settings *s, const uint8_t *bases, const uint64_t *offsets,
                                const uint32_t *pa, const uint32_t *pb, uint64_t np, int G, int D,
                                hovl_record *out, uint64_t *cells_out) {
  uint64_t total = 0;
  for (uint64_t p = 0; p < np; p++) {
    const uint32_t a = pa[p], b = pb[p];
    const uint8_t *A = bases + offsets[a - 1], *B = bases + offsets[b - 1];
    const int32_t n = (int32_t)(offsets[a] - offsets[a - 1]), m = (int32_t)(offsets[b] - offsets[b - 1]);
    hovl_record rec; memset(&rec, 0, sizeof rec);
    uint64_t cells = 0;
    switch (D) {
      case 2: run_pair<2>(s, A, n, B, m, G, &rec, &cells); break;
      case 3: run_pair<3>(s, A, n, B, m, G, &rec, &cells); break;
      case 4: run_pair<4>(s, A, n, B, m, G, &rec, &cells); break;
      case 7: run_pair<7>(s, A, n, B, m, G, &rec, &cells); break;
      case 8: run_pair<8>(s, A, n, B, m, G, &rec, &cells); break;
      default: return -1;
    }
    rec.a = (int32_t)a; rec.b = (int32_t)b; rec.count = out[p].count;
    out[p] = rec; total += cells;
  }
  if (cells_out) *cells_out = total;
  return 0;
}
