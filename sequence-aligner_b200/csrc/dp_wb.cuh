// dp_wb.cuh — wide-band dovetail aligner (band width 33..255): one WARP per (A,B) pair with 2-bit
// direction codes, the lane-group geometry of dp_core.cuh joined to the cell of dp_tp.cuh.
//
// Same function as dp_core.cuh / dp_tp.cuh (generateFastDovetailAlignment, reference
// src/BioLibs.scala:373-591 / :596-822).  What is different from dp_core.cuh (which carries the
// result of the reference's backward trace FORWARD as per-cell tags, three extra registers and
// ~3x the instructions per cell):
//   * a cell is dp_tp.cuh's: every value is (score << 6) + p; the three candidates get their
//     priority p from the constant that is added anyway, so one 3-input max gives the cell maximum
//     AND which of M/X/Y reached it first in the reference's if/else-if order
//     (BioLibs.scala:444-456, :541-575); 2 bits per cell are shifted into a register per column;
//   * lane l owns D consecutive columns (phase 1) or band diagonals (phase 2) and walks the rows
//     skewed by one step per lane (row = step - lane); every 16 STEPS each lane stores its D
//     direction words (one 16- or 32-byte vector store, the warp's stores are contiguous), so word
//     (f, lane, c) holds steps 16f+1 .. 16f+16 of that lane, i.e. rows 16f+1-lane .. 16f+16-lane;
//   * arg-max: bits 2..5 of a value are free (scores are shifted by 6), so
//     key[c] = max(key[c], h + ((15 - step in block) << 2)) keeps, per column and 16-step block, the
//     best score and its EARLIEST row in one add-max; the blocks are compared when they end;
//   * padded diagonals (the class has more than width + 1) are neutralised without a per-cell test:
//     their gap-open-into-Y constant is hugely negative, so the one value a real cell reads from them
//     (Y from (u-1, kk+1)) is 0 as in the reference; range-checked cells exist only on the band's ends;
//   * the backward walks (BioLibs.scala:443-459, :538-579) read the codes back, the whole warp at once
//     in the kernel (dp_wb.cu); walk1 / walk2 below are their one-thread form.
// Plain per-lane code: the same source runs in the CUDA kernel and in tests/cpu/wb_emul.cpp.
#pragma once
#include "dp_tp.cuh"

namespace hovl {
namespace wb {

using tp::ldw;
using tp::load16;
using tp::mad;
using tp::mad_opaque;
using tp::popc32;
using tp::iaddmax;

constexpr int SH = 6;                       // value = (score << SH) + low bits
constexpr int32_t UNIT = 1 << SH;
constexpr int32_t NEG = -(1 << 30);         // "never": padded diagonals, rows that do not exist
constexpr int LANES = 32;

struct Consts {
  const int32_t *costu;     // 16 entries, (cost[a][b] << SH) + 2, row a at byte offset 16*a (shared memory on the device)
  int32_t ge;               // gE << SH
  int32_t gox, goy;         // (gO << SH) + 1, gO << SH
  int32_t one, neg1, four;  // run-time 1, -1, 4: adds written as x*one + y issue on the FMA pipe (dp_tp.cuh)
};
HOVL_HD int32_t cost_at(const Consts &K, uint32_t byte_off) {
  return *reinterpret_cast<const int32_t *>(reinterpret_cast<const char *>(K.costu) + byte_off);
}
// values must fit: |score| < 2^24 for every reachable score (checked on the host, align_pairs)
HOVL_HD bool fits(const int32_t cost[16], int32_t gap_open, int32_t gap_extend, int64_t max_len) {
  int64_t big = 1;
  for (int i = 0; i < 16; i++) { const int64_t v = cost[i]; if (v > big) big = v; if (-v > big) big = -v; }
  if (-(int64_t)gap_open - (int64_t)gap_extend > big) big = -(int64_t)gap_open - (int64_t)gap_extend;
  return big * (max_len + 64) < ((int64_t)1 << 24);
}

template <int D> struct DirStride { static constexpr int value = (D + 3) & ~3; };   // words per lane per block (vector stores)
HOVL_HD int dir_stride_rt(int D) { return (D + 3) & ~3; }
// direction-code words one warp needs for reads of at most max_len bases
HOVL_HD size_t scratch_words(int32_t max_len, int D) {
  return (size_t)((max_len + LANES + 15) / 16 + 1) * (size_t)(LANES * dir_stride_rt(D));
}

// 16 bases starting at `pos` (first base in the top bits); pos may run off either end (zeros there)
struct Win {
  const uint32_t *w; int32_t nw, pos; uint32_t w0, w1;
  HOVL_HD uint32_t word(int32_t i) const { return (i >= 0 && i < nw) ? ldw(w + i) : 0u; }
  HOVL_HD void init(const uint32_t *words, int32_t nwords, int32_t p) {
    w = words; nw = nwords; pos = p;
    const int32_t wi = p >> 4;
    w0 = word(wi); w1 = word(wi + 1);
  }
  HOVL_HD uint32_t bits() const {
    const int sh = 2 * (pos & 15);
#if defined(__CUDA_ARCH__)
    return __funnelshift_l(w1, w0, sh);
#else
    return sh ? ((w0 << sh) | (w1 >> (32 - sh))) : w0;
#endif
  }
  HOVL_HD void advance() {
    pos++;
    if ((pos & 15) == 0) { w0 = w1; w1 = word((pos >> 4) + 1); }
  }
};

// code of a cell: 3 = maximum <= 0 (the trace stops here), 2 = M, 1 = X, 0 = Y
#define HOVL_WB_CELL(dg_, hxin_, hyin_, cst_, goy_)                    \
  const int32_t xv = mad((hxin_), K.one, K.ge);                        \
  const int32_t yv = mad((hyin_), K.one, K.ge);                        \
  int32_t raw = imax3(iaddmax((dg_), (cst_), xv), yv, 3);              \
  int32_t h = raw & ~3;                                                \
  int32_t hxn = iaddmax_relu(h, K.gox, xv);                            \
  int32_t hyn = iaddmax_relu(h, (goy_), yv)

struct Best { int32_t key, blk, col; };     // key = (score << SH) | (15 - step in block) << 2 ; blk = 16-step block
HOVL_HD void best_init(Best &b) { b.key = 0; b.blk = 0; b.col = 0; }

// ------------------------------------------------------------------------------------
// Phase 1 (BioLibs.scala:415-466): rows i = 1..n, columns j = 1..width.  Lane l owns columns
// col = lD .. lD+D-1 (j = col + 1) and is at row i = step - l.
// ------------------------------------------------------------------------------------
template <int D>
struct P1 {
  int32_t hm[D], hy[D], key[D];
  uint32_t dir[D], colofs[D];       // colofs: 4 * base of B under the column (byte offset inside a cost row)
  int32_t diag_h;                   // h(i-1, lD-1): what the left lane computed two steps ago
  int32_t o_hx, o_hm;               // hx / h of this lane's last column at the row it did last
  Best best;
};
template <int D>
HOVL_HD void p1_init(P1<D> &L, int lane, const uint32_t *wb, int32_t m) {
#pragma unroll
  for (int c = 0; c < D; c++) {
    const int32_t col = lane * D + c;
    L.hm[c] = 0; L.hy[c] = 0; L.key[c] = 0; L.dir[c] = 0;
    L.colofs[c] = col < m ? (packed_base(wb, col) << 2) : 0u;
  }
  L.diag_h = 0; L.o_hx = 0; L.o_hm = 0;
  best_init(L.best);
}
// One row of one lane.  in_hx / in_hm: hx and h of (i, lD-1), what the left lane produced one step ago
// (0 for lane 0: column 0 is the border).  a16 = 16 * base of A at row i; rinv = (15 - step in block) << 2.
template <int D>
HOVL_HD void p1_row(P1<D> &L, const Consts &K, uint32_t a16, int32_t rinv, int32_t in_hx, int32_t in_hm) {
  int32_t dg = L.diag_h, hx = in_hx;
#pragma unroll
  for (int c = 0; c < D; c++) {
    HOVL_WB_CELL(dg, hx, L.hy[c], cost_at(K, (uint32_t)mad((int32_t)L.colofs[c], K.one, (int32_t)a16)), K.goy);
    L.dir[c] = (uint32_t)mad(h, K.neg1, mad_opaque((int32_t)L.dir[c], K.four, raw));
    dg = L.hm[c];
    L.hm[c] = h; hx = hxn; L.hy[c] = hyn;
    L.key[c] = iaddmax(h, rinv, L.key[c]);
  }
  L.diag_h = in_hm;
  L.o_hx = hx; L.o_hm = L.hm[D - 1];
}

// End of a 16-step block: the columns' keys against the lane's best so far (strict '>' on the score keeps the
// earlier block; inside a block the key order is score, then earliest row; equal keys: lowest column first).
template <int D>
HOVL_HD void block_best(int32_t key[D], Best &best, int32_t blk, int lane, int32_t ncols /* valid columns of the pair */) {
  int32_t bk = 0, bc = 0;
#pragma unroll
  for (int c = 0; c < D; c++) {
    const int32_t k = (lane * D + c < ncols) ? key[c] : 0;
    if (k > bk) { bk = k; bc = c; }
    key[c] = 0;
  }
  if ((bk >> SH) > (best.key >> SH)) { best.key = bk; best.blk = blk; best.col = lane * D + bc; }
}
// (score << SH, row, column) of a lane's best; row = step - lane with step = 16*blk + (15 - rinv) + 1
HOVL_HD void best_cell(const Best &b, int lane, int32_t &score, int32_t &row, int32_t &col) {
  score = b.key & ~(UNIT - 1);
  row = 16 * b.blk + (15 - ((b.key >> 2) & 15)) + 1 - lane;
  col = b.col;
}
// is (s1, r1, c1) ahead of (s0, r0, c0) in the reference's order: larger score, then first in row-major order
HOVL_HD bool best_before(int32_t s1, int32_t r1, int32_t c1, int32_t s0, int32_t r0, int32_t c0) {
  return s1 > s0 || (s1 == s0 && (r1 < r0 || (r1 == r0 && c1 < c0)));
}

// address (in words) of the direction word that holds step `s` (1-based) of (lane, c)
template <int D>
HOVL_HD size_t dir_index(int32_t s, int lane, int c) {
  return ((size_t)((s - 1) >> 4) * LANES + (size_t)lane) * DirStride<D>::value + (size_t)c;
}
// the 2-bit code of row `row`, column / diagonal `col` (0-based) of the pass whose words are in sc
template <int D>
HOVL_HD uint32_t code_at(const uint32_t *sc, int32_t row, int32_t col) {
  const int lane = col / D, c = col - lane * D;
  const int32_t s = row + lane;
  return (sc[dir_index<D>(s, lane, c)] >> (2 * (15 - ((s - 1) & 15)))) & 3u;
}

// Backward walk of phase 1 from the arg-max (bi, bj), one thread (BioLibs.scala:443-466): doveStart or -1 (dud).
template <int D>
HOVL_HD int32_t walk1(const uint32_t *sc, int32_t bi, int32_t bj) {
  int32_t ti = bi, tj = bj;
  for (;;) {
    const uint32_t d = code_at<D>(sc, ti, tj - 1);
    if (d == 2u) { ti--; tj--; } else if (d == 1u) { tj--; } else { ti--; }
    if (ti == 0 || tj == 0) break;                      // border cells: maximum is 0
    if (code_at<D>(sc, ti, tj - 1) == 3u) break;
  }
  return tj == 0 ? ti : -1;
}

// ------------------------------------------------------------------------------------
// Phase 2 (BioLibs.scala:473-591): rows u = 1..dl, band diagonals kk = 0..width, i = u + ds,
// j = kk - z + u; M from (u-1,kk), X from (u,kk-1), Y from (u-1,kk+1).  Lane l owns kk = lD..lD+D-1
// and is at row u = step - l.  Inside a step every lane first does its diagonal c = 0 (p2_first); that
// cell's hy is what the LEFT neighbour (one row ahead) needs as Y input of its last diagonal in the same
// step, so the kernel shuffles it down, then every lane does c = 1..D-1 (p2_rest).
// ------------------------------------------------------------------------------------
template <int D>
struct P2 {
  int32_t hm[D], hy[D], key[D], goyc[D];      // goyc: gO << SH, or NEG for a padded diagonal
  uint32_t dir[D];
  int32_t hx;                                 // running X input inside a row
  int32_t o_hx;                               // hx of this lane's last diagonal at the row it did last
  Best best;
};
template <int D>
HOVL_HD void p2_init(P2<D> &L, const Consts &K, int lane, int32_t width) {
#pragma unroll
  for (int c = 0; c < D; c++) {
    L.hm[c] = 0; L.hy[c] = 0; L.key[c] = 0; L.dir[c] = 0;
    L.goyc[c] = (lane * D + c <= width) ? K.goy : NEG;
  }
  L.hx = 0; L.o_hx = 0;
  best_init(L.best);
}
struct P2Row {              // what a lane knows about the row it is on
  uint32_t a16;             // 16 * base of A at row u
  uint32_t bb;              // bases of B under diagonals lD .. lD+15 of row u, first in the top bits
  int32_t rinv;             // (15 - step in block) << 2
  int32_t j0;               // j of diagonal lD (CHK only)
  int32_t m, kmax;          // len(B); width - lD = last valid c (CHK only)
};
// cell (u, kk = lD + c).  CHK: rows where the band sticks out of B (1 <= j <= m, BioLibs.scala:505) — the reference
// forces those cells to zero.
template <int D, bool CHK>
HOVL_HD void p2_cell(P2<D> &L, const Consts &K, const P2Row &r, const int c, const int32_t up) {
  const uint32_t b4 = (c <= 14 ? (r.bb >> (28 - 2 * c)) : (r.bb << 2)) & 0xCu;
  HOVL_WB_CELL(L.hm[c], L.hx, up, cost_at(K, (uint32_t)mad((int32_t)b4, K.one, (int32_t)r.a16)), L.goyc[c]);
  if (CHK) {
    const bool ok = (uint32_t)(r.j0 + c - 1) < (uint32_t)r.m && c <= r.kmax;
    if (!ok) { raw = 3; h = 0; hxn = 0; hyn = 0; }
  }
  L.dir[c] = (uint32_t)mad(h, K.neg1, mad_opaque((int32_t)L.dir[c], K.four, raw));
  L.hm[c] = h; L.hx = hxn; L.hy[c] = hyn;
  L.key[c] = iaddmax(h, r.rinv, L.key[c]);
}
// c = 0.  in_hx = hx of (u, lD-1) from the left lane's previous step (0 for lane 0).
template <int D, bool CHK>
HOVL_HD void p2_first(P2<D> &L, const Consts &K, const P2Row &r, int32_t in_hx, int32_t right_if_d1) {
  L.hx = in_hx;
  p2_cell<D, CHK>(L, K, r, 0, D > 1 ? L.hy[D > 1 ? 1 : 0] : right_if_d1);
}
// c = 1..D-1.  right = hy of (u-1, (l+1)D) from the right lane's p2_first of this step (0 past the last lane).
template <int D, bool CHK>
HOVL_HD void p2_rest(P2<D> &L, const Consts &K, const P2Row &r, int32_t right) {
#pragma unroll
  for (int c = 1; c < D; c++) p2_cell<D, CHK>(L, K, r, c, c + 1 < D ? L.hy[c + 1 < D ? c + 1 : c] : right);
  L.o_hx = L.hx;
}

// Backward walk of phase 2 from the arg-max (bu, bk), one thread (BioLibs.scala:538-579): start cell and the
// match / mismatch counts.  The kernel's whole-warp form (dp_wb.cu) must give the same answer.
template <int D>
HOVL_HD void walk2(const uint32_t *sc, const uint32_t *wa, const uint32_t *wb, int32_t ds, int32_t z,
                   int32_t bu, int32_t bk, int32_t &tu, int32_t &kk, int32_t &cc, int32_t &ee) {
  tu = bu; kk = bk; cc = 0; ee = 0;
  for (;;) {
    const uint32_t d = code_at<D>(sc, tu, kk);
    if (d == 2u) {
      const int32_t ii = tu + ds, jj = kk - z + tu;
      if (packed_base(wa, ii - 1) != packed_base(wb, jj - 1)) ee++; else cc++;
      tu--;
    } else if (d == 1u) { ee++; kk--; }
    else { ee++; tu--; kk++; }
    if (tu == 0) break;                                  // row 0 is all zeros
    if (code_at<D>(sc, tu, kk) == 3u) break;
  }
}

// classes: D diagonals per lane, 32 lanes -> band width + 1 <= 32 * D
#define HOVL_WB_CLASSES(X) X(2) X(3) X(4) X(5) X(6) X(7) X(8)
constexpr int WB_MIN_WIDTH = 33, WB_MAX_WIDTH = 255;

}  // namespace wb
}  // namespace hovl
