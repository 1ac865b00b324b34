// dp_tp.cuh — thread-per-pair banded dovetail aligner for narrow bands (width <= 32).
//
// Same function as dp_core.cuh (generateFastDovetailAlignment, reference
// src/BioLibs.scala:373-591 / :596-822), different machine mapping: ONE thread owns one
// (A,B) pair and keeps a whole DP row of the band in registers, so there are no shuffles,
// no skew and no padding lanes.  What makes a cell cheap:
//   * every value is carried as (score << SH) + p, SH = 2 or 12.  The three candidates of a cell
//     get their priority p from the constant that is added anyway (M: cost<<SH + 2, X: gE<<SH on an
//     operand whose low bits are 1, Y: gE<<SH on low bits 0), so ONE 3-input max yields both the cell
//     maximum and which of M/X/Y reached it first in the reference's if/else-if order
//     (BioLibs.scala:444-456, :541-575).  max(.,3) then turns "maximum <= 0" into code 3;
//   * the reference's backward trace only looks at "which of M,X,Y equals the maximum at the
//     cell I stand on", i.e. at exactly that 2-bit code, so the forward pass stores 2 bits per
//     cell (16 rows of one column per 32-bit word, one coalesced store per column per 16
//     rows) and the same thread then walks the codes backwards;
//   * since gO <= 0, max(M+gO, Y+gO, X, 0) == max(max(M,X,Y,0)+gO, X, 0): one add-max-relu
//     per gap state;
//   * every forced-zero border of the reference is <= 0 and every recurrence clamps at 0, so
//     borders are plain zeros.
// Plain per-thread code (no warp intrinsics): the same source runs in the CUDA kernels
// (dp.cu) and in the host harness of the CPU tests (tests/cpu/tp_emul.cpp).
#pragma once
#include "dp_core.cuh"

namespace hovl {
namespace tp {

HOVL_HD uint32_t shr2_in(uint32_t lo, uint32_t hi) {          // (lo >> 2) | (hi << 30)
#if defined(__CUDA_ARCH__)
  return __funnelshift_r(lo, hi, 2);
#else
  return (lo >> 2) | (hi << 30);
#endif
}
HOVL_HD uint32_t shl2_in(uint32_t hi, uint32_t lo) {          // (hi << 2) | (lo >> 30)
#if defined(__CUDA_ARCH__)
  return __funnelshift_l(lo, hi, 2);
#else
  return (hi << 2) | (lo >> 30);
#endif
}
HOVL_HD uint32_t shr8_in(uint32_t lo, uint32_t hi) {          // (lo >> 8) | (hi << 24)
#if defined(__CUDA_ARCH__)
  return __funnelshift_r(lo, hi, 8);
#else
  return (lo >> 8) | (hi << 24);
#endif
}
HOVL_HD uint32_t byte_of(uint32_t x, int b) {                 // byte b of x, zero-extended: one PRMT
#if defined(__CUDA_ARCH__)
  return __byte_perm(x, 0u, 0x4440u + (uint32_t)b);
#else
  return (x >> (8 * b)) & 0xffu;
#endif
}
HOVL_HD int32_t popc32(uint32_t x) {
#if defined(__CUDA_ARCH__)
  return __popc(x);
#else
  return __builtin_popcount(x);
#endif
}
HOVL_HD uint32_t ldw(const uint32_t *p) {
#if defined(__CUDA_ARCH__)
  return __ldg(p);
#else
  return *p;
#endif
}

struct Consts {
  const int32_t *cost4;     // 16 entries, cost4[a*4+b] = (cost[a][b] << SH) + 2 (shared memory on the device)
  int32_t ge4;              // gE << SH
  int32_t gox, goy;         // (gO << SH) + 1, gO << SH
  // 1 as a RUN-TIME value (kernel parameter).  The integer ALU pipe (max, logic, shift: 2 cycles
  // per warp instruction, tools/pipe_probe.cu) carries most of a cell, the FMA pipe idles; an add
  // written as x*one + y cannot be folded by the compiler and issues as IMAD on the FMA pipe.
  int32_t one, neg1, four;
};
HOVL_HD int32_t mad(int32_t a, int32_t b, int32_t c) { return a * b + c; }
// same, but opaque to the optimiser (which would turn word*4 + code back into shift + or on the ALU pipe)
HOVL_HD int32_t mad_opaque(int32_t a, int32_t b, int32_t c) {
#if defined(__CUDA_ARCH__)
  int32_t d;
  asm("mad.lo.s32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
  return d;
#else
  return a * b + c;
#endif
}
HOVL_HD int32_t cost_at(const Consts &K, uint32_t byte_off) {
  return *reinterpret_cast<const int32_t *>(reinterpret_cast<const char *>(K.cost4) + byte_off);
}

// 16 bases starting at base index p (p may be negative or run past the end: zeros there),
// first base in the top bits.
HOVL_HD uint32_t load16(const uint32_t *w, int32_t nw, int32_t p) {
  const int32_t wi = p >> 4;
  const uint32_t lo = (wi >= 0 && wi < nw) ? ldw(w + wi) : 0u;
  const uint32_t hi = (wi + 1 >= 0 && wi + 1 < nw) ? ldw(w + wi + 1) : 0u;
  const int sh = 2 * (p & 15);
  return sh ? ((lo << sh) | (hi >> (32 - sh))) : lo;
}

// Sequential base reader: next() returns the base at `pos` (0 outside the read) and advances.
struct SeqReader {
  const uint32_t *w; int32_t nw, pos; uint32_t cur;
  HOVL_HD void init(const uint32_t *words, int32_t nwords, int32_t p) {
    w = words; nw = nwords; pos = p;
    const int32_t wi = p >> 4;
    cur = ((wi >= 0 && wi < nw) ? ldw(w + wi) : 0u) << (2 * (p & 15));
  }
  HOVL_HD uint32_t next() {
    const uint32_t b = cur >> 30;
    pos++;
    if ((pos & 15) == 0) { const int32_t wi = pos >> 4; cur = (wi >= 0 && wi < nw) ? ldw(w + wi) : 0u; }
    else cur <<= 2;
    return b;
  }
};

// Direction codes of one pass: word (rb, c) holds rows 16*rb .. 16*rb+15 of column c, row r in
// bits [2*(15-(r&15)), 2*(15-(r&15))+1]: a new row enters from below as word*4 + raw - h (the
// code is raw - h), two integer multiply-adds on the FMA pipe instead of a shift on the ALU pipe.  Word (rb, c) of this thread lives at sc[(rb*NC + c) * SS]
// (SS = 32 on the device: the 32 threads of a warp interleave, so a warp stores 128 B lines).
template <int SS>
struct DirReader {
  const uint32_t *sc; int32_t nc, have; uint32_t word;
  HOVL_HD void init(const uint32_t *p, int32_t ncols) { sc = p; nc = ncols; have = -1; word = 0; }
  HOVL_HD uint32_t word_of(int32_t r, int32_t c) {          // all 16 codes of the word that holds (r, c)
    const int32_t idx = (r >> 4) * nc + c;
    if (idx != have) { word = sc[(size_t)idx * SS]; have = idx; }
    return word;
  }
  HOVL_HD uint32_t get(int32_t r, int32_t c) {
    const int32_t idx = (r >> 4) * nc + c;
    if (idx != have) { word = sc[(size_t)idx * SS]; have = idx; }
    return (word >> (2 * (15 - (r & 15)))) & 3u;
  }
};


// Arg-max bookkeeping.  FAST (reads <= 1023 bases, scores < 2^18): SH = 12, the ten bits between
// the priority and the score are free, and one packed key per column,
// key = max(key, h + ((1023 - row) << 2)), keeps the best score and, among equal scores, the
// earliest row with ONE add-max; the earliest column is picked when the columns are compared at
// the end.  Otherwise SH = 2 and the arg-max costs a compare and two selects per cell.
constexpr int TP_FAST_MAX_LEN = 1023;
template <bool FAST> struct Shift { static constexpr int value = FAST ? 12 : 2; };
HOVL_HD int32_t iaddmax(int32_t a, int32_t b, int32_t c) {      // max(a + b, c)
#if defined(__CUDA_ARCH__)
  return __viaddmax_s32(a, b, c);
#else
  return imax(a + b, c);
#endif
}

// code of a cell: 3 = maximum <= 0 (the trace stops here), 2 = M, 1 = X, 0 = Y
// One DP cell.  dg/xv/yv are the packed predecessors; returns packed outputs.
// 12 instructions: address add + LDS, 2 adds, 2 multiply-adds for the direction word (FMA pipe),
// add-max, max3, and, 2 x add-max-relu, add-max for the arg-max key (ALU pipe): 6 ALU, 5 FMA, 1 LSU.
#define HOVL_TP_CELL(dg_, hxin_, hyin_, cst_)                          \
  const int32_t mv = (dg_) + (cst_);                                   \
  const int32_t xv = mad((hxin_), K.one, K.ge4);                       \
  const int32_t yv = mad((hyin_), K.one, K.ge4);                       \
  int32_t raw = imax(imax3(mv, xv, yv), 3);                            \
  int32_t h = raw & ~3;                                                \
  int32_t hxn = iaddmax_relu(h, K.gox, xv);                            \
  int32_t hyn = iaddmax_relu(h, K.goy, yv)

// ------------------------------------------------------------------------------------
// Phase 1 (BioLibs.scala:415-466): rows i = 1..n, columns j = 1..width (c = j-1), then the
// backward walk from the arg-max.  Returns doveStart (>= 0) or -1 for a dud.
// W = columns held in registers; widths WMIN..W run in this instantiation.
// ------------------------------------------------------------------------------------
template <int W, int WMIN, int SS, bool FAST>
HOVL_HD int32_t phase1(const Consts &K, const uint32_t *wa, int32_t n, const uint32_t *wb, int32_t m,
                       int32_t width, uint32_t *sc) {
  constexpr int NWB = (W + 15) / 16;
  const int32_t nwb = (m + 15) >> 4;
  uint32_t bw[NWB];
#pragma unroll
  for (int i = 0; i < NWB; i++) bw[i] = i < nwb ? ldw(wb + i) : 0u;
  int32_t hm[W], hy[W]; uint32_t dir[W], colofs[W];       // colofs: 4 * base of B under the column
  int32_t key[FAST ? W : 1];
#pragma unroll
  for (int c = 0; c < W; c++) {
    hm[c] = 0; hy[c] = 0; dir[c] = 0; if (FAST) key[c] = 0;
    const int sh = 28 - 2 * (c & 15);
    colofs[c] = (sh >= 0 ? (bw[c >> 4] >> sh) : (bw[c >> 4] << 2)) & 0xCu;
  }
  int32_t best = 0, bpos = 0;
  uint32_t awin = 0;
  // one cell of row `i` (its base code a16, arg-max constants rowinv / rowpos) at column c
  auto cell = [&](const int c, const uint32_t a16, const int32_t rowinv, const int32_t rowpos, int32_t &dg, int32_t &hx) {
    HOVL_TP_CELL(dg, hx, hy[c], cost_at(K, (uint32_t)mad((int32_t)colofs[c], K.one, (int32_t)a16)));
    dir[c] = (uint32_t)mad(h, K.neg1, mad_opaque((int32_t)dir[c], K.four, raw));
    dg = hm[c];
    hm[c] = h; hx = hxn; hy[c] = hyn;
    if (FAST) key[c] = iaddmax(h, rowinv, key[c]);
    else {
      // rows are processed two at a time, so "first in row-major order" is an explicit tie-break
      bool upd = h > best || (h == best && rowpos + c < bpos);
      if (c >= WMIN) upd = upd && (c < width);         // padded column of this class
      if (upd) { best = h; bpos = rowpos + c; }
    }
  };
  auto flush = [&](const int32_t i) {
    uint32_t *dst = sc + (size_t)((i >> 4) - 1) * W * SS;
#pragma unroll
    for (int c = 0; c < W; c++) dst[(size_t)c * SS] = dir[c];
  };
  int32_t i = 1;
  // Two rows per trip, the second one two columns behind the first: cell (i+1, c) needs row i
  // only up to column c, so the two rows are independent instruction streams the scheduler can
  // interleave (a single row is one long dependent chain through hx).
  for (; i + 1 <= n; i += 2) {
    if (((i - 1) & 15) == 0) awin = ldw(wa + ((i - 1) >> 4));
    const uint32_t a16a = (awin >> 26) & 0x30u, a16b = (awin >> 24) & 0x30u;
    awin <<= 4;
    const int32_t posa = i << 8, posb = (i + 1) << 8;
    const int32_t inva = (TP_FAST_MAX_LEN - i) << 2, invb = (TP_FAST_MAX_LEN - i - 1) << 2;
    int32_t dga = 0, hxa = 0, dgb = 0, hxb = 0;
#pragma unroll
    for (int s = 0; s < W + 2; s++) {
      if (s < W) cell(s, a16a, inva, posa, dga, hxa);
      if (s >= 2) cell(s - 2, a16b, invb, posb, dgb, hxb);
    }
    if (((i + 1) & 15) == 0) flush(i + 1);
  }
  for (; i <= n; i++) {                                  // odd tail row
    if (((i - 1) & 15) == 0) awin = ldw(wa + ((i - 1) >> 4));
    const uint32_t a16 = (awin >> 26) & 0x30u;
    awin <<= 2;
    int32_t dg = 0, hx = 0;
#pragma unroll
    for (int c = 0; c < W; c++) cell(c, a16, (TP_FAST_MAX_LEN - i) << 2, i << 8, dg, hx);
    if ((i & 15) == 0) flush(i);
  }
  if (n & 15) {
    const int sh = 2 * (16 - (n & 15));
    uint32_t *dst = sc + (size_t)(n >> 4) * W * SS;
#pragma unroll
    for (int c = 0; c < W; c++) dst[(size_t)c * SS] = dir[c] << sh;
  }
  if (FAST) {
    int32_t bk = 0;
#pragma unroll
    for (int c = 0; c < W; c++)
      if ((c < WMIN || c < width) && key[c] > bk) { bk = key[c]; bpos = ((TP_FAST_MAX_LEN - ((bk >> 2) & 1023)) << 8) + c; }
    best = bk >> 12;
  }
  if (best <= 0) return -1;                             // no positive cell: the reference would index [-1]
  int32_t ti = bpos >> 8, tj = (bpos & 255) + 1;
  DirReader<SS> rd; rd.init(sc, W);
  for (;;) {                                            // BioLibs.scala:443-459
    const uint32_t d = rd.get(ti - 1, tj - 1);
    if (d == 2u) { ti--; tj--; } else if (d == 1u) { tj--; } else { ti--; }
    if (ti == 0 || tj == 0) break;                      // border cells: maximum is 0
    if (rd.get(ti - 1, tj - 1) == 3u) break;
  }
  return tj == 0 ? ti : -1;                               // :464-466
}

// ------------------------------------------------------------------------------------
// Phase 2 (BioLibs.scala:473-591): rows u = 1..dl, band diagonals kk = 0..width (c = kk),
// i = u + ds, j = kk - z + u; M from (u-1,kk), X from (u,kk-1), Y from (u-1,kk+1).
// ------------------------------------------------------------------------------------
struct BestCell { int32_t best, bpos; };
template <bool V> struct BoolC { static constexpr bool value = V; };

template <int W, int WMIN, int SS, bool FAST>
HOVL_HD void phase2(const Consts &K, const uint32_t *wa, int32_t n, const uint32_t *wb, int32_t m, int32_t width,
                    int32_t ds, uint32_t *sc, int32_t min_overlap, uint32_t min_identity, uint32_t max_ignore,
                    DpOut &R) {
  constexpr int D = W + 1;
  constexpr int NB = (W + 4) / 4;                         // window registers, 4 bases each (>= D + 1 bases)
  const int32_t dl = n - ds, z = width / 2;
  const int32_t nwa = (n + 15) >> 4, nwb = (m + 15) >> 4;
  SeqReader ar; ar.init(wa, nwa, ds);                     // row u reads A[ds + u - 1]
  // bwb: the bases of B under row u, diagonal kk at window index kk, one BYTE per base holding
  // 4*base (its offset inside a cost row): a cell finds its cost address with one byte-permute
  // and one (FMA-pipe) add.  Row u+1 sees the same window one byte further on.
  uint32_t bwb[NB];
  {
    uint32_t w16[(NB * 4 + 15) / 16];
#pragma unroll
    for (int i = 0; i < (NB * 4 + 15) / 16; i++) w16[i] = load16(wb, nwb, 16 * i - z);
#pragma unroll
    for (int i = 0; i < NB; i++) {
      uint32_t x = 0;
#pragma unroll
      for (int t = 0; t < 4; t++) {
        const int idx = 4 * i + t;
        x |= (((w16[idx >> 4] >> (30 - 2 * (idx & 15))) & 3u) << 2) << (8 * t);
      }
      bwb[i] = x;
    }
  }
  SeqReader br; br.init(wb, nwb, 4 * NB - z);
  int32_t hm[D], hy[D]; uint32_t dir[D];
  int32_t key[FAST ? D : 1];
#pragma unroll
  for (int c = 0; c < D; c++) { hm[c] = 0; hy[c] = 0; dir[c] = 0; if (FAST) key[c] = 0; }
  BestCell B; B.best = 0; B.bpos = 0;
  // cell (u, kk = c); wsel = c for the row the window was built for, c + 1 for the row after it
  auto cell = [&](auto chk, const int c, const int wsel, const uint32_t a16, const int32_t rowinv, const int32_t rowpos,
                  const int32_t j0, int32_t &hx) {
    const uint32_t b4 = byte_of(bwb[wsel >> 2], wsel & 3);
    const int32_t up = (c + 1 < D) ? hy[c + 1 < D ? c + 1 : c] : 0;
    HOVL_TP_CELL(hm[c], hx, up, cost_at(K, (uint32_t)mad((int32_t)b4, K.one, (int32_t)a16)));
    bool ok = true;
    if (decltype(chk)::value) ok = (uint32_t)(j0 + c - 1) < (uint32_t)m;   // 1 <= j <= m (BioLibs.scala:505)
    if (c > WMIN) ok = ok && (c <= width);                                  // padded diagonal of this class
    if (decltype(chk)::value || c > WMIN) { if (!ok) { raw = 3; h = 0; hxn = 0; hyn = 0; } }
    dir[c] = (uint32_t)mad(h, K.neg1, mad_opaque((int32_t)dir[c], K.four, raw));
    hm[c] = h; hx = hxn; hy[c] = hyn;
    if (FAST) key[c] = iaddmax(h, rowinv, key[c]);
    else if (h > B.best || (h == B.best && rowpos + c < B.bpos)) { B.best = h; B.bpos = rowpos + c; }
  };
  auto flush = [&](const int32_t u) {
    uint32_t *dst = sc + (size_t)((u >> 4) - 1) * D * SS;
#pragma unroll
    for (int c = 0; c < D; c++) dst[(size_t)c * SS] = dir[c];
  };
  // rows whose whole band lies inside B need no per-cell range test
  const int32_t lo = z + 1;
  const int32_t hi = (m - width + z) < dl ? (m - width + z) : dl;
  int32_t u = 1;
  while (u <= dl) {
    if ((u & 1) && u >= lo && u + 1 <= hi) {
      // two rows per trip, the second two diagonals behind the first (it needs row u up to kk + 1):
      // two independent instruction streams instead of one dependent chain through hx
      const uint32_t a16a = ar.next() << 4, a16b = ar.next() << 4;
      const int32_t inva = (TP_FAST_MAX_LEN - u) << 2, invb = (TP_FAST_MAX_LEN - u - 1) << 2;
      int32_t hxa = 0, hxb = 0;
#pragma unroll
      for (int s = 0; s < D + 2; s++) {
        if (s < D) cell(BoolC<false>(), s, s, a16a, inva, u << 8, 0, hxa);
        if (s >= 2) cell(BoolC<false>(), s - 2, s - 1, a16b, invb, (u + 1) << 8, 0, hxb);
      }
      const uint32_t n0 = br.next(), n1 = br.next();
#pragma unroll
      for (int i = 0; i + 1 < NB; i++) bwb[i] = (bwb[i] >> 16) | (bwb[i + 1] << 16);
      bwb[NB - 1] = (bwb[NB - 1] >> 16) | (n0 << 18) | (n1 << 26);
      u += 2;
      if (((u - 1) & 15) == 0) flush(u - 1);
    } else {
      const uint32_t a16 = ar.next() << 4;
      const int32_t inv = (TP_FAST_MAX_LEN - u) << 2;
      int32_t hx = 0;
      if (u >= lo && u <= hi) {
#pragma unroll
        for (int c = 0; c < D; c++) cell(BoolC<false>(), c, c, a16, inv, u << 8, u - z, hx);
      } else {
#pragma unroll
        for (int c = 0; c < D; c++) cell(BoolC<true>(), c, c, a16, inv, u << 8, u - z, hx);
      }
#pragma unroll
      for (int i = 0; i + 1 < NB; i++) bwb[i] = shr8_in(bwb[i], bwb[i + 1]);
      bwb[NB - 1] = (bwb[NB - 1] >> 8) | (br.next() << 26);
      if ((u & 15) == 0) flush(u);
      u++;
    }
  }
  if (dl & 15) {
    const int sh = 2 * (16 - (dl & 15));
    uint32_t *dst = sc + (size_t)(dl >> 4) * D * SS;
#pragma unroll
    for (int c = 0; c < D; c++) dst[(size_t)c * SS] = dir[c] << sh;
  }
  if (FAST) {
    int32_t bk = 0;
#pragma unroll
    for (int c = 0; c < D; c++)
      if (key[c] > bk) { bk = key[c]; B.bpos = ((TP_FAST_MAX_LEN - ((bk >> 2) & 1023)) << 8) + c; }
    B.best = bk & ~4095;
  }
  if (B.best <= 0) return;                                // R stays the dud it was initialised to
  const int32_t bu = B.bpos >> 8, bk = B.bpos & 255;
  int32_t tu = bu, kk = bk, cc = 0, ee = 0;
  DirReader<SS> rd; rd.init(sc, D);
  for (;;) {                                              // BioLibs.scala:538-579
    // sixteen diagonal moves at once: the whole word above says "M" (substitutions are diagonal
    // moves too), so only the mismatches between the two 16-base windows need counting
    if ((tu & 15) == 0 && rd.word_of(tu - 1, kk) == 0xAAAAAAAAu) {
      const uint32_t x = load16(wa, nwa, tu + ds - 16) ^ load16(wb, nwb, kk - z + tu - 16);
      const int32_t mm = popc32((x | (x >> 1)) & 0x55555555u);
      ee += mm; cc += 16 - mm; tu -= 16;
      if (tu == 0) break;
      if (rd.get(tu - 1, kk) == 3u) break;
      continue;
    }
    const uint32_t d = rd.get(tu - 1, kk);
    if (d == 2u) {
      const int32_t ii = tu + ds, jj = kk - z + tu;
      if (packed_base(wa, ii - 1) != packed_base(wb, jj - 1)) ee++; else cc++;
      tu--;
    } else if (d == 1u) { ee++; kk--; }
    else { ee++; tu--; kk++; }
    if (tu == 0) break;                                    // row 0 is all zeros
    if (rd.get(tu - 1, kk) == 3u) break;
  }
  R.start_i = tu + ds; R.start_j = kk - z + tu;             // :582-584
  R.end_i = bu + ds;  R.end_j = bk - z + bu;
  R.correct = cc; R.error = ee; R.score = B.best >> Shift<FAST>::value;
  dp_finish(R, n, m, min_overlap, min_identity, max_ignore);
}

// (W, WMIN) instantiations: a pair of band width w runs in the first class with w <= W.
// ------------------------------------------------------------------------------------
// Phase 1 without the DP.  Phase 1 (src/BioLibs.scala:415-466) aligns all of A against the first
// `width` bases of B, takes the first maximum in row-major order and traces back to column 0;
// doveStart is the row it arrives at.  With U_j = sum of cost[B_j'][B_j'] over j' <= j, induction
// over the three recurrences gives M, X, Y (i, j) <= U_j when every diagonal cost is positive and
// the strict maximum of its column (probe_valid), and with gE < 0 the gap states stay strictly below
// U_j, so a cell reaches U_width only as a chain of `width` diagonal matches, i.e.
// A[i-width, i) == B[0, width).  The FIRST such row i* is therefore the arg-max (no cell exceeds
// U_width, none before it equals it), every cell on its diagonal has M as its maximum, the trace
// runs straight up that diagonal to (i* - width, 0) where the maximum is 0, and
// doveStart = i* - width.  For a pair in which B's prefix occurs in A exactly, phase 1 is a
// substring search over the 2-bit packed reads; the other pairs go through the phase-1 DP.
// ------------------------------------------------------------------------------------
HOVL_HD bool probe_valid(const int32_t cost[16], int32_t gap_extend) {
  if (gap_extend >= 0) return false;
  for (int b = 0; b < 4; b++) {
    if (cost[b * 4 + b] <= 0) return false;
    for (int a = 0; a < 4; a++) if (a != b && cost[a * 4 + b] >= cost[b * 4 + b]) return false;
  }
  return true;
}
HOVL_HD uint32_t fsr(uint32_t lo, uint32_t hi, int s) {      // low word of (hi:lo) >> s, 0 <= s <= 31
#if defined(__CUDA_ARCH__)
  return __funnelshift_r(lo, hi, s);
#else
  return s ? (lo >> s) | (hi << (32 - s)) : lo;
#endif
}
// doveStart if B[0, width) occurs in A (first occurrence), else -1.  12 <= width <= 32 <= ... m >= width.
HOVL_HD int32_t probe_prefix(const uint32_t *wa, int32_t n, const uint32_t *wb, int32_t m, int32_t width) {
  const int32_t w2 = 2 * width;
  const uint32_t b0 = ldw(wb), b1 = m > 16 ? ldw(wb + 1) : 0u;
  // the first `width` bases of B, right-aligned in 64 bits, first base most significant
  const unsigned long long pat = (((unsigned long long)b0 << 32) | b1) >> (64 - w2);
  const uint32_t pat_lo = (uint32_t)pat, pat_hi = (uint32_t)(pat >> 32);
  const uint32_t mask_hi = w2 > 32 ? (0xffffffffu >> (64 - w2)) : 0u, mask_lo = w2 >= 32 ? 0xffffffffu : ((1u << w2) - 1u);
  const int32_t nw = (n + 15) >> 4;
  uint32_t p2 = 0, p1 = 0;                                    // the two words before the current one
  for (int32_t wi = 0; wi < nw; wi++) {
    const uint32_t cur = ldw(wa + wi);
    uint32_t hits = 0;
#pragma unroll
    for (int o = 0; o < 16; o++) {                            // the window that ENDS with base 16*wi + o
      const uint32_t lo = fsr(cur, p1, 30 - 2 * o);
      if (((lo ^ pat_lo) & mask_lo) == 0u) hits |= 1u << o;
    }
    if (hits && mask_hi) {                                    // width > 16: the upper bases, only where the lower 16 matched
      uint32_t keep = 0;
#pragma unroll
      for (int o = 0; o < 16; o++) {
        const uint32_t hi = fsr(p1, p2, 30 - 2 * o);
        if (((hi ^ pat_hi) & mask_hi) == 0u) keep |= 1u << o;
      }
      hits &= keep;
    }
    // only windows inside A: rows i = 16*wi + o + 1 with width <= i <= n
    const int32_t first = width - 1 - 16 * wi, last = n - 1 - 16 * wi;
    if (first > 0) hits &= first >= 16 ? 0u : (0xffffffffu << first);
    if (last < 15) hits &= (2u << last) - 1u;
    if (hits) {
      int o = 0;
      while (!((hits >> o) & 1u)) o++;
      return 16 * wi + o + 1 - width;
    }
    p2 = p1; p1 = cur;
  }
  return -1;
}

#define HOVL_TP_CLASSES(X) X(12, 12) X(16, 13) X(20, 17) X(24, 21) X(28, 25) X(32, 29)
constexpr int TP_MIN_WIDTH = 12, TP_MAX_WIDTH = 32;
// direction-code words one thread needs for reads of at most max_len bases
HOVL_HD size_t scratch_words(int32_t max_len, int W) { return (size_t)((max_len + 15) / 16 + 1) * (size_t)(W + 1); }

}  // namespace tp
}  // namespace hovl
