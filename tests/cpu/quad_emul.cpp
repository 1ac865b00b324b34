// quad_emul.cpp — host lane emulator for the quadratic aligner (CPU test harness, not product code).
//
// Runs the SAME per-lane functions the CUDA kernel runs (csrc/quad_core.cuh) with the warp
// shuffles replaced by array copies between 32 emulated lanes and the per-warp boundary buffer by
// a vector, in the kernel's own order of tiles, steps and lanes, so that the tiling, the skewed
// wavefront, the forward-carried tags and the arg-max tie-breaking across tiles can be checked
// against the oracle where there is no GPU.  Built by tests/conftest.py with g++.
#include <stdint.h>
#include <string.h>
#include <vector>
#include "../../sequence-aligner_b200/csrc/quad_core.cuh"
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

template <int D>
static void run_pair(const hovl_settings *s, const uint8_t *A, int32_t n, const uint8_t *B, int32_t m,
                     hovl_record *rec, uint64_t *cells) {
  constexpr int G = 32;
  DpOut r; memset(&r, 0, sizeof r);
  *cells = 0;
  bool bad = false;
  std::vector<uint32_t> pa = pack(A, n, &bad), pb = pack(B, m, &bad);
  int32_t cost16[16];
  for (int i = 0; i < 16; i++) cost16[i] = s->cost[i / 4][i % 4];
  DpConst K; K.cost = cost16; K.gO = s->gap_open; K.gE = s->gap_extend;
  auto store = [&]() {
    rec->flags = r.flags; rec->start_i = r.start_i; rec->start_j = r.start_j; rec->end_i = r.end_i;
    rec->end_j = r.end_j; rec->correct = r.correct; rec->error = r.error; rec->score = r.score;
    rec->ahg = r.ahg; rec->bhg = r.bhg;
  };
  if (bad) { dp_set_dud(r, HOVL_CORE_F_BAD_BASE); store(); return; }
  std::vector<QLane<D>> L((size_t)G);
  for (int l = 0; l < G; l++) q_pair_init<D>(L[l]);
  std::vector<QEdge> bnd_in((size_t)n + 2), bnd_out((size_t)n + 2);
  const int32_t tiles = (m + G * D - 1) / (G * D);
  for (int32_t t = 0; t < tiles; t++) {
    std::vector<QEdge> o((size_t)G);
    for (int l = 0; l < G; l++) { const int32_t j0 = t * G * D + l * D; q_tile_init<D>(L[l], j0); o[l] = q_border(0, j0 + D); }
    for (int32_t st = 1; st <= n + G - 1; st++) {
      const std::vector<QEdge> prev = o;                                    // = __shfl_up
      for (int l = 0; l < G; l++) {
        const int32_t i = st - l, j0 = t * G * D + l * D;
        QEdge in;
        if (l > 0) in = prev[l - 1];
        else in = (t == 0 || i < 1 || i > n) ? q_border(i, j0) : bnd_in[(size_t)i];
        const uint32_t a = (i >= 1 && i <= n) ? packed_base(pa.data(), i - 1) : 0u;
        uint32_t bcol[D];
        for (int c = 0; c < D; c++) bcol[c] = (j0 + c < m) ? packed_base(pb.data(), j0 + c) : 0u;
        o[l] = q_row<D>(L[l], K, j0, i, n, m, a, bcol, in);
        if (l == G - 1 && i >= 1 && i <= n) bnd_out[(size_t)i] = o[l];
      }
    }
    bnd_in.swap(bnd_out);
  }
  *cells = (uint64_t)n * (uint64_t)m;
  int32_t best = 0, bi = 0, bj = 0; uint32_t t1 = 0, t2 = 0;
  for (int l = 0; l < G; l++) {
    const QLane<D> &q = L[l];
    if (q.best > 0 && q_better(q.best, q.best_i, q.best_j, best, bi, bj)) { best = q.best; bi = q.best_i; bj = q.best_j; t1 = q.best_t1; t2 = q.best_t2; }
  }
  if (best <= 0) { dp_set_dud(r, 0); store(); return; }
  q_finish(r, best, bi, bj, t1, t2, n, m, s->min_overlap, s->min_identity_f32, s->max_ignore_f32);
  store();
}

extern "C" int quad_emul_align_pairs(const hovl_settings *s, const uint8_t *bases, const uint64_t *offsets,
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
      case 1: run_pair<1>(s, A, n, B, m, &rec, &cells); break;
      case 2: run_pair<2>(s, A, n, B, m, &rec, &cells); break;
      case 4: run_pair<4>(s, A, n, B, m, &rec, &cells); break;
      default: return -1;
    }
    rec.a = (int32_t)a; rec.b = (int32_t)b; rec.count = out[p].count;
    out[p] = rec; total += cells;
  }
  if (cells_out) *cells_out = total;
  return 0;
}
