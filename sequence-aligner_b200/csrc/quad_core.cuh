// quad_core.cuh — lane-level math of the quadratic local aligner (--quadratic-align).
//
// What it computes is generateLocalAlignment (reference src/BioLibs.scala:171-263, set variant
// :267-368): the three recurrences of the fast aligner over the FULL lenA x lenB rectangle,
// arg-max with strict '>' in row-major order, greedy backward trace from the arg-max until the
// cell maximum is <= 0.  How it computes it:
//   * a warp owns one (A,B) pair.  B is cut into column tiles of 32*D columns; inside a tile
//     lane l owns D consecutive columns and walks the rows skewed by one step per lane, so
//     the anti-diagonal wavefront lives in registers and neighbours talk through shuffles;
//   * the right-most column of a tile (what column j0+1 of the next tile reads: hx, hm and the
//     tag of every row) goes through a per-warp boundary buffer of 16 B per row;
//   * the backward trace looks only at the cell it stands on, so its result is carried FORWARD
//     as a per-cell tag (stop cell, matches, mismatches), exactly as in dp_core.cuh; the
//     rectangle itself is never stored (the reference keeps three (lenA+1) x (lenB+1) arrays);
//   * every border value of the reference is <= 0 and every recurrence clamps at 0
//     (src/BioLibs.scala:181-191, :199-211), so borders are 0 inputs whose tag stops on them.
// Plain per-lane code, shared by the CUDA kernel (quad.cu) and the host lane emulator
// (tests/cpu/quad_emul.cpp).
#pragma once
#include "dp_core.cuh"

namespace hovl {

// What a lane hands to its right neighbour (and a tile to the next tile) for one row.
struct QEdge { int32_t hx, hm; uint32_t t1, t2; };

// Tag: t1 = (stop_i << 16) | stop_j, t2 = (matches << 16) | mismatches.  Needs lenA, lenB <= 32767.
template <int D>
struct QLane {
  int32_t hm[D];      // max(M,X,Y,0) of (i-1, own column c)
  int32_t hy[D];      // max(M+gO, Y, X+gO, 0) of (i-1, own column c)
  uint32_t t1[D], t2[D];
  int32_t diag_h; uint32_t diag_t1, diag_t2;       // (i-1, column left of the own block)
  int32_t best, best_i, best_j; uint32_t best_t1, best_t2;   // over all tiles: largest, then smallest row
};

template <int D>
HOVL_HD void q_pair_init(QLane<D> &L) { L.best = 0; L.best_i = 0; L.best_j = 0; L.best_t1 = 0; L.best_t2 = 0; }

// j0 = column left of the lane's block in this tile (the block is j0+1 .. j0+D)
template <int D>
HOVL_HD void q_tile_init(QLane<D> &L, int32_t j0) {
#pragma unroll
  for (int c = 0; c < D; c++) { L.hm[c] = 0; L.hy[c] = 0; L.t1[c] = (uint32_t)(j0 + 1 + c); L.t2[c] = 0; }   // T(0,j) stops at (0,j)
  L.diag_h = 0; L.diag_t1 = (uint32_t)j0; L.diag_t2 = 0;
}

// The edge a block sees on its left when there is nothing there (column 0, or a row outside 1..n).
HOVL_HD QEdge q_border(int32_t i, int32_t j0) {
  QEdge e; e.hx = 0; e.hm = 0; e.t1 = ((uint32_t)i << 16) | (uint32_t)j0; e.t2 = 0; return e;
}

// One row of one lane's block.  `in` = cell (i, j0); returns cell (i, j0 + D).
template <int D>
HOVL_HD QEdge q_row(QLane<D> &L, const DpConst &K, int32_t j0, int32_t i, int32_t n, int32_t m,
                    uint32_t a, const uint32_t *bcol /* D base codes of the own columns */, const QEdge &in) {
  const bool row_ok = (i >= 1) && (i <= n);
  int32_t left_hx = in.hx; uint32_t left_t1 = in.t1, left_t2 = in.t2;
  int32_t dg_h = L.diag_h; uint32_t dg_t1 = L.diag_t1, dg_t2 = L.diag_t2;
  QEdge out; out.hx = 0; out.hm = 0; out.t1 = 0; out.t2 = 0;
#pragma unroll
  for (int c = 0; c < D; c++) {
    const int32_t j = j0 + 1 + c;
    const bool ok = row_ok && (j <= m);
    const int32_t keep_h = L.hm[c]; const uint32_t keep_t1 = L.t1[c], keep_t2 = L.t2[c];   // diagonal input of column c+1
    int32_t mv = K.cost[a * 4 + bcol[c]] + dg_h;
    int32_t xv = K.gE + left_hx;
    int32_t yv = K.gE + L.hy[c];
    if (!ok) { mv = 0; xv = 0; yv = 0; }
    const int32_t cur = imax3(mv, xv, yv);
    uint32_t t1, t2;
    if (cur <= 0) { t1 = ((uint32_t)i << 16) | (uint32_t)j; t2 = 0; }                       // the trace stops ON this cell
    else if (mv == cur) { t1 = dg_t1; t2 = dg_t2 + ((a != bcol[c]) ? 1u : 0x10000u); }      // M first, then X, then Y (:233-247)
    else if (xv == cur) { t1 = left_t1; t2 = left_t2 + 1u; }
    else { t1 = keep_t1; t2 = keep_t2 + 1u; }
    // strict '>' in row-major order over the whole rectangle: a later tile may tie on an earlier row
    if (ok && (cur > L.best || (cur == L.best && i < L.best_i))) { L.best = cur; L.best_i = i; L.best_j = j; L.best_t1 = t1; L.best_t2 = t2; }
    const int32_t hm = imax(cur, 0);
    const int32_t hx = iaddmax_relu(imax(mv, yv), K.gO, xv);
    const int32_t hy = iaddmax_relu(imax(mv, xv), K.gO, yv);
    L.hm[c] = hm; L.hy[c] = hy; L.t1[c] = t1; L.t2[c] = t2;
    left_hx = hx; left_t1 = t1; left_t2 = t2;
    dg_h = keep_h; dg_t1 = keep_t1; dg_t2 = keep_t2;
    if (c == D - 1) { out.hx = hx; out.hm = hm; out.t1 = t1; out.t2 = t2; }
  }
  L.diag_h = in.hm; L.diag_t1 = in.t1; L.diag_t2 = in.t2;
  return out;
}

// "is (b, bi, bj) before (best, i, j)" for the final reduction over lanes
HOVL_HD bool q_better(int32_t ob, int32_t oi, int32_t oj, int32_t best, int32_t bi, int32_t bj) {
  return ob > best || (ob == best && (oi < bi || (oi == bi && oj < bj)));
}

HOVL_HD void q_finish(DpOut &R, int32_t best, int32_t bi, int32_t bj, uint32_t t1, uint32_t t2, int32_t n, int32_t m,
                      int32_t min_overlap, uint32_t min_identity_bits, uint32_t max_ignore_bits) {
  R.start_i = (int32_t)(t1 >> 16); R.start_j = (int32_t)(t1 & 0xffffu);
  R.end_i = bi; R.end_j = bj;
  R.correct = (int32_t)(t2 >> 16); R.error = (int32_t)(t2 & 0xffffu); R.score = best;
  dp_finish(R, n, m, min_overlap, min_identity_bits, max_ignore_bits);
}

}  // namespace hovl
