// dp_core.cuh — lane-level math of the banded dovetail aligner.
//
// What it computes is generateFastDovetailAlignment (reference src/BioLibs.scala:373-591,
// set variant :596-822): phase 1 = all of A against the first `width` bases of B, phase 2 =
// a band of width+1 diagonals from doveStart.  How it computes it is new:
//   * a group of G lanes owns one (A,B) pair; lane l owns D consecutive columns (phase 1)
//     or D consecutive band diagonals (phase 2) and walks the rows skewed by one step per
//     lane, so the whole anti-diagonal wavefront lives in registers and neighbours talk
//     through warp shuffles only;
//   * the reference's backward "greedy trace" (BioLibs.scala:443-459, :538-579) looks only
//     at the cell it stands on, so its result is carried FORWARD with the wavefront as a
//     per-cell tag T = (stop cell, matches, mismatches); no DP matrix is ever stored;
//   * every forced-zero / border value of the reference is non-positive and every
//     recurrence clamps at 0, so borders are represented by 0 inputs.
// The functions here are plain per-lane code (no shuffles) so that the same source runs
// inside the CUDA kernel (dp.cu) and inside the host lane emulator used by the CPU tests.
#pragma once
#include <stdint.h>

#if defined(__CUDACC__)
#define HOVL_HD __host__ __device__ __forceinline__
#else
#define HOVL_HD inline
#endif

namespace hovl {

HOVL_HD int32_t imax(int32_t a, int32_t b) { return a > b ? a : b; }
HOVL_HD int32_t imax3(int32_t a, int32_t b, int32_t c) {
#if defined(__CUDA_ARCH__)
  return __vimax3_s32(a, b, c);
#else
  return imax(a, imax(b, c));
#endif
}
HOVL_HD int32_t imax3_relu(int32_t a, int32_t b, int32_t c) {
#if defined(__CUDA_ARCH__)
  return __vimax3_s32_relu(a, b, c);
#else
  return imax(imax(a, imax(b, c)), 0);
#endif
}
// max(a + b, c, 0)
HOVL_HD int32_t iaddmax_relu(int32_t a, int32_t b, int32_t c) {
#if defined(__CUDA_ARCH__)
  return __viaddmax_s32_relu(a, b, c);
#else
  return imax(imax(a + b, c), 0);
#endif
}

// Reads are stored 2 bits per base, A0 C1 G2 T3 (the costFunc order, BioLibs.scala:119-160),
// 16 bases per 32-bit word, FIRST base in the MOST significant bits (so that a k-mer code,
// ObjectStore.scala:48-67, is a plain right shift of a window).  Every read starts on a
// word boundary.
HOVL_HD uint32_t packed_base(const uint32_t *words, int32_t idx) {
  return (words[idx >> 4] >> (30 - 2 * (idx & 15))) & 3u;
}

// float32 helpers with JVM semantics: every operation rounded to binary32, never fused.
HOVL_HD float f32_from_bits(uint32_t b) {
#if defined(__CUDA_ARCH__)
  return __uint_as_float(b);
#else
  union { uint32_t u; float f; } x; x.u = b; return x.f;
#endif
}
HOVL_HD float f32_mul(float a, float b) {
#if defined(__CUDA_ARCH__)
  return __fmul_rn(a, b);
#else
  volatile float r = a * b; return r;
#endif
}
HOVL_HD float f32_add(float a, float b) {
#if defined(__CUDA_ARCH__)
  return __fadd_rn(a, b);
#else
  volatile float r = a + b; return r;
#endif
}
HOVL_HD float f32_sub(float a, float b) {
#if defined(__CUDA_ARCH__)
  return __fsub_rn(a, b);
#else
  volatile float r = a - b; return r;
#endif
}
HOVL_HD float f32_div(float a, float b) {
#if defined(__CUDA_ARCH__)
  return __fdiv_rn(a, b);
#else
  volatile float r = a / b; return r;
#endif
}

// width = max(kmerSize, floor(A.size * (1 - minIdentity)).toInt + 1)   BioLibs.scala:389-390
HOVL_HD int32_t band_width(int32_t n, int32_t k, uint32_t min_identity_bits) {
  const float omi = f32_sub(1.0f, f32_from_bits(min_identity_bits));
  const float prod = f32_mul((float)n, omi);
  // prod >= 0 and finite for sane settings: truncation == floor
  const int32_t w = (int32_t)prod + 1;
  return k > w ? k : w;
}

// record flag bits (include/hovl.h)
#define HOVL_CORE_F_DUD 1
#define HOVL_CORE_F_ALIGN_VALID 2
#define HOVL_CORE_F_OVL_VALID 4
#define HOVL_CORE_F_SHORT_B 8
#define HOVL_CORE_F_BAD_BASE 16

struct DpOut {                      // the integer content of Alignment + Overlap
  int32_t flags, start_i, start_j, end_i, end_j, correct, error, score, ahg, bhg;
};

HOVL_HD void dp_set_dud(DpOut &r, int extra) {        // BioLibs.dud, BioLibs.scala:22
  r.flags = HOVL_CORE_F_DUD | extra;
  r.start_i = r.start_j = r.end_i = r.end_j = 0;
  r.correct = 0; r.error = 1; r.score = 0; r.ahg = 0; r.bhg = 0;
}

// Alignment.valid (ObjectStore.scala:99-107), Overlap ahg/bhg/valid (:124-125, :137-141)
HOVL_HD void dp_finish(DpOut &r, int32_t n, int32_t m, int32_t min_overlap,
                       uint32_t min_identity_bits, uint32_t max_ignore_bits) {
  const float fc = (float)r.correct, fe = (float)r.error;
  const float ratio = f32_div(fc, f32_add(fc, fe));
  const bool av = (ratio >= f32_from_bits(min_identity_bits)) &&
                  (r.correct + r.error >= min_overlap) &&
                  ((r.start_i == 0 && m == r.end_j) || (r.start_j == 0 && n == r.end_i));
  r.ahg = r.start_i - r.start_j;
  r.bhg = m - n + r.ahg;
  const float mi = f32_from_bits(max_ignore_bits);
  const int32_t aa = r.ahg < 0 ? -r.ahg : r.ahg, ab = r.bhg < 0 ? -r.bhg : r.bhg;
  const bool ov = av && ((float)aa < mi) && ((float)ab < mi);
  r.flags = (av ? HOVL_CORE_F_ALIGN_VALID : 0) | (ov ? HOVL_CORE_F_OVL_VALID : 0);
}

struct DpConst {
  const int32_t *cost;   // 16 entries, cost[a*4+b], A,C,G,T order (shared memory on the device)
  int32_t gO, gE;        // <= 0
};

// What a lane hands to its neighbour.
struct Edge { int32_t h; uint32_t t1, t2; };

// ------------------------------------------------------------------------------------
// Phase 1 (BioLibs.scala:415-459): rectangle rows i=1..n, columns j=1..width.
// Tag: t1 = (stop_i << 8) | stop_j.   (stop_j <= width <= 255)
// ------------------------------------------------------------------------------------
template <int D>
struct P1Lane {
  int32_t hm[D];      // max(M,X,Y,0) of (i-1, own column c)
  int32_t hy[D];      // max(M+gO, Y, X+gO, 0) of (i-1, own column c)
  uint32_t t[D];      // tag of (i-1, own column c)
  int32_t diag_h;     // max(M,X,Y,0) of (i-1, column left of own block)
  uint32_t diag_t;
  int32_t best, best_i, best_j; uint32_t best_t;
};

template <int D>
HOVL_HD void p1_init(P1Lane<D> &L, int lane) {
#pragma unroll
  for (int c = 0; c < D; c++) { L.hm[c] = 0; L.hy[c] = 0; L.t[c] = (uint32_t)(lane * D + 1 + c); }  // T(0,j) = stop (0,j)
  L.diag_h = 0; L.diag_t = (uint32_t)(lane * D);                                                     // T(0, lD)
  L.best = 0; L.best_i = 0; L.best_j = 0; L.best_t = 0;
}

// One row for one lane.  `in` = (hx, tag) of cell (i, lD) and its hm in in_hm; for lane 0
// the caller passes the column-0 border.  Returns what the right neighbour needs.
template <int D>
HOVL_HD void p1_row(P1Lane<D> &L, const DpConst &K, int lane, int32_t i, int32_t n, int32_t width,
                    uint32_t a, const uint32_t *bcol /* D base codes of own columns */,
                    int32_t in_hx, int32_t in_hm, uint32_t in_t,
                    int32_t &out_hx, int32_t &out_hm, uint32_t &out_t) {
  const bool row_ok = (i >= 1) && (i <= n);
  int32_t left_hx = in_hx; uint32_t left_t = in_t;
  int32_t dg_h = L.diag_h; uint32_t dg_t = L.diag_t;
#pragma unroll
  for (int c = 0; c < D; c++) {
    const int32_t j = lane * D + 1 + c;
    const bool ok = row_ok && (j <= width);
    const int32_t up_hy = L.hy[c]; const uint32_t up_t = L.t[c];
    const int32_t keep_h = L.hm[c];            // becomes the diagonal input of column c+1
    const uint32_t keep_t = L.t[c];
    int32_t mv = K.cost[a * 4 + bcol[c]] + dg_h;
    int32_t xv = K.gE + left_hx;
    int32_t yv = K.gE + up_hy;
    if (!ok) { mv = 0; xv = 0; yv = 0; }
    const int32_t cur = imax3(mv, xv, yv);
    uint32_t t;
    if (cur <= 0) t = ((uint32_t)i << 8) | (uint32_t)j;
    else if (mv == cur) t = dg_t;
    else if (xv == cur) t = left_t;
    else t = up_t;
    if (ok && cur > L.best) { L.best = cur; L.best_i = i; L.best_j = j; L.best_t = t; }
    const int32_t hm = imax(cur, 0);
    const int32_t hx = iaddmax_relu(imax(mv, yv), K.gO, xv);
    const int32_t hy = iaddmax_relu(imax(mv, xv), K.gO, yv);
    L.hm[c] = hm; L.hy[c] = hy; L.t[c] = t;
    left_hx = hx; left_t = t;
    dg_h = keep_h; dg_t = keep_t;
    if (c == D - 1) { out_hx = hx; out_hm = hm; out_t = t; }
  }
  L.diag_h = in_hm; L.diag_t = in_t;
}

// ------------------------------------------------------------------------------------
// Phase 2 (BioLibs.scala:495-579): rows u=0..dl, band diagonals kk=0..width,
// i = u + ds, j = kk - z + u.  Cell (u,kk) reads (u-1,kk) [M], (u,kk-1) [X], (u-1,kk+1) [Y].
// Lane l owns kk = lD..lD+D-1 and works on row u = step - l.  Inside one step every lane
// first does its diagonal c=0 (p2_first); the result is what the LEFT neighbour needs for
// its last diagonal in the same step (it is one row behind), so the kernel shuffles it
// down, then every lane finishes c=1..D-1 (p2_rest).
// Tag: t1 = (stop_u << 8) | stop_kk, t2 = (matches << 16) | mismatches.
// ------------------------------------------------------------------------------------
template <int D>
struct P2Lane {
  int32_t hm[D];      // max(M,X,Y,0) of (u-1, own kk)
  int32_t hy[D];      // max(M+gO, Y, X+gO, 0) of (u-1, own kk)
  uint32_t t1[D], t2[D];
  int32_t left_hx; uint32_t left_t1, left_t2;     // running left neighbour inside a row
  int32_t best, best_u, best_kk; uint32_t best_t1, best_t2;
};

template <int D>
HOVL_HD void p2_init(P2Lane<D> &L, int lane) {
#pragma unroll
  for (int c = 0; c < D; c++) { L.hm[c] = 0; L.hy[c] = 0; L.t1[c] = (uint32_t)(lane * D + c); L.t2[c] = 0; }
  L.left_hx = 0; L.left_t1 = 0; L.left_t2 = 0;
  L.best = 0; L.best_u = 0; L.best_kk = 0; L.best_t1 = 0; L.best_t2 = 0;
}

// One cell.  (dg_*) = (u-1,kk), (lf_*) = (u,kk-1), (up_*) = (u-1,kk+1).
template <int D>
HOVL_HD void p2_cell(P2Lane<D> &L, const DpConst &K, int c, bool ok, int32_t u, int32_t kk,
                     uint32_t a, uint32_t b,
                     int32_t up_hy, uint32_t up_t1, uint32_t up_t2) {
  const int32_t dg_h = L.hm[c]; const uint32_t dg_t1 = L.t1[c], dg_t2 = L.t2[c];
  int32_t mv = K.cost[a * 4 + b] + dg_h;
  int32_t xv = K.gE + L.left_hx;
  int32_t yv = K.gE + up_hy;
  if (!ok) { mv = 0; xv = 0; yv = 0; }
  const int32_t cur = imax3(mv, xv, yv);
  uint32_t t1, t2;
  if (cur <= 0) { t1 = ((uint32_t)u << 8) | (uint32_t)kk; t2 = 0; }
  else if (mv == cur) { t1 = dg_t1; t2 = dg_t2 + ((a != b) ? 1u : 0x10000u); }
  else if (xv == cur) { t1 = L.left_t1; t2 = L.left_t2 + 1u; }
  else { t1 = up_t1; t2 = up_t2 + 1u; }
  if (ok && cur > L.best) { L.best = cur; L.best_u = u; L.best_kk = kk; L.best_t1 = t1; L.best_t2 = t2; }
  L.hm[c] = imax(cur, 0);
  L.hy[c] = iaddmax_relu(imax(mv, xv), K.gO, yv);
  L.left_hx = iaddmax_relu(imax(mv, yv), K.gO, xv);
  L.t1[c] = t1; L.t2[c] = t2; L.left_t1 = t1; L.left_t2 = t2;
}

// Geometry of one lane-row.
struct P2Geom { int32_t ds, dl, z, width, m; };

HOVL_HD bool p2_ok(const P2Geom &g, int32_t u, int32_t kk) {
  const int32_t j = kk - g.z + u;
  return (u >= 1) && (u <= g.dl) && (kk <= g.width) && (j >= 1) && (j <= g.m);
}

// c = 0.  `in` = (hx, tag) of (u, lD-1) from the left lane (0s for lane 0).
template <int D>
HOVL_HD void p2_first(P2Lane<D> &L, const DpConst &K, const P2Geom &g, int lane, int32_t u,
                      uint32_t a, uint32_t b0, const Edge &in) {
  L.left_hx = in.h; L.left_t1 = in.t1; L.left_t2 = in.t2;
  const int32_t kk = lane * D;
  p2_cell<D>(L, K, 0, p2_ok(g, u, kk), u, kk, a, b0, L.hy[1 < D ? 1 : 0], L.t1[1 < D ? 1 : 0], L.t2[1 < D ? 1 : 0]);
}
// What the left neighbour needs after p2_first: (hy, tag) of (u, lD).
template <int D>
HOVL_HD Edge p2_down(const P2Lane<D> &L) { Edge e; e.h = L.hy[0]; e.t1 = L.t1[0]; e.t2 = L.t2[0]; return e; }

// c = 1..D-1.  `right` = (hy, tag) of (u-1, (l+1)D) from the right lane (0s past the band).
// Returns (hx, tag) of the lane's last cell for the right neighbour's next step.
template <int D>
HOVL_HD Edge p2_rest(P2Lane<D> &L, const DpConst &K, const P2Geom &g, int lane, int32_t u,
                     uint32_t a, const uint32_t *bdiag, const Edge &right) {
#pragma unroll
  for (int c = 1; c < D; c++) {
    const int32_t kk = lane * D + c;
    const bool last = (c == D - 1);
    const int32_t up_hy = last ? right.h : L.hy[last ? c : c + 1];
    const uint32_t up_t1 = last ? right.t1 : L.t1[last ? c : c + 1];
    const uint32_t up_t2 = last ? right.t2 : L.t2[last ? c : c + 1];
    p2_cell<D>(L, K, c, p2_ok(g, u, kk), u, kk, a, bdiag[c], up_hy, up_t1, up_t2);
  }
  Edge e; e.h = L.left_hx; e.t1 = L.left_t1; e.t2 = L.left_t2; return e;
}

}  // namespace hovl
