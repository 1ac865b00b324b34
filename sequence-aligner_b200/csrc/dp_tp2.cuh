// dp_tp2.cuh — the thread-per-pair aligner of dp_tp.cuh with TWO pairs per thread.
//
// Every DP value of dp_tp.cuh is (score << 2) + priority.  When the largest score a phase can
// reach fits 13 bits (phase 1: W columns of matches; phase 2: one read length of matches —
// true for unit-like cost matrices on kilobase reads), that value fits a signed 16-bit half
// word, and sm_100a has the whole cell in packed 16x2 form: VIADD.16x2 (FMA pipe),
// VIADDMNMX.S16x2[.RELU], VIMNMX3.S16x2, VIMNMX.S16x2 (integer pipe), all at the rate of
// their 32-bit forms (tools/pipe_probe.cu).  One thread therefore runs the pair in the low
// half word and another pair in the high half word with the instruction stream of ONE pair:
//   cost   one 256-entry shared-memory table indexed by (a1, a0, b1, b0) returns both costs
//   dir    a cell's two direction codes (raw - h, one per half word) enter a 32-bit word as
//          word*4 + code; after 8 rows each half word holds the 8 codes of its pair, so the
//          words are stored every 8 rows (same bytes per pair as dp_tp.cuh's 16 rows/word)
//   argmax 16 bits have no room for a row index.  A running per-column maximum (one packed
//          max per cell) gives, every 16 rows, "did the maximum rise in this superblock".
//          The DP state at the start of the current superblock is parked in shared memory;
//          when a half word's maximum rose, that state is copied to the pair's slot in
//          scratch.  After the last row the thread replays ONLY the 16 rows of the recorded
//          superblock from the saved state and stops at the first cell (row-major, the
//          reference's order) that equals the maximum.
// The two pairs may differ in length and band width (same class): rows past the shorter one
// are garbage for its half word and never looked at — its maximum is frozen when it ends.
// Same host/device source discipline as dp_tp.cuh (tests/cpu/tp_emul.cpp runs it with g++).
#pragma once
#include "dp_tp.cuh"

namespace hovl {
namespace tp2 {

using tp::byte_of;
using tp::ldw;
using tp::load16;
using tp::mad;
using tp::mad_opaque;
using tp::popc32;
using tp::SeqReader;
using tp::shr8_in;

HOVL_HD int32_t sext16(uint32_t x) { return (int32_t)(int16_t)(uint16_t)x; }
HOVL_HD uint32_t pack16(int32_t hi, int32_t lo) { return ((uint32_t)hi << 16) | ((uint32_t)lo & 0xffffu); }

#if defined(HOVL_TP2_COUNT_WRAPS)
static int g_wraps = 0;      // host harness only: half-word sums that did not fit (must stay 0 outside frame moves)
static bool g_count_wraps = true;
#endif
HOVL_HD uint32_t padd(uint32_t a, uint32_t b) {                       // per half word a + b
#if defined(__CUDA_ARCH__)
  return __vadd2(a, b);
#else
  const int32_t hi = sext16(a >> 16) + sext16(b >> 16), lo = sext16(a) + sext16(b);
#if defined(HOVL_TP2_COUNT_WRAPS)
  if (g_count_wraps && (hi != sext16((uint32_t)hi) || lo != sext16((uint32_t)lo))) g_wraps++;
#endif
  return pack16(hi, lo);
#endif
}
HOVL_HD uint32_t pmax(uint32_t a, uint32_t b) {                       // per half word signed max
#if defined(__CUDA_ARCH__)
  return __vmaxs2(a, b);
#else
  return pack16(imax(sext16(a >> 16), sext16(b >> 16)), imax(sext16(a), sext16(b)));
#endif
}
HOVL_HD uint32_t pmax3(uint32_t a, uint32_t b, uint32_t c) {
#if defined(__CUDA_ARCH__)
  return __vimax3_s16x2(a, b, c);
#else
  return pmax(a, pmax(b, c));
#endif
}
HOVL_HD uint32_t paddmax(uint32_t a, uint32_t b, uint32_t c) {        // max(a + b, c)
#if defined(__CUDA_ARCH__)
  return __viaddmax_s16x2(a, b, c);
#else
  return pmax(padd(a, b), c);
#endif
}
HOVL_HD uint32_t paddmax_relu(uint32_t a, uint32_t b, uint32_t c) {   // max(a + b, c, 0)
#if defined(__CUDA_ARCH__)
  return __viaddmax_s16x2_relu(a, b, c);
#else
  return pmax(pmax(padd(a, b), c), 0u);
#endif
}

constexpr uint32_t P_THREE = 0x00030003u, P_SCORE = 0xFFFCFFFCu;
constexpr int ROWS_PER_WORD = 8, SUPER = 16;

struct Consts2 {
  const uint32_t *cost2;    // 256 entries: [(a1*4 + a0)*16 + b1*4 + b0] = pack16(cost4[a1][b1], cost4[a0][b0])
  uint32_t ge, gox, goy;    // packed (gE << 2), (gO << 2) + 1, gO << 2
  int32_t one, neg1, four;  // run-time 1, -1, 4 (dp_tp.cuh: keeps adds on the FMA pipe)
  int32_t tgt, floor;       // phase 2 frames (frame_fits): where the row maximum is put, lowest usable value
};
HOVL_HD uint32_t cost2_at(const Consts2 &K, uint32_t byte_off) {
  return *reinterpret_cast<const uint32_t *>(reinterpret_cast<const char *>(K.cost2) + byte_off);
}

// Base reader for loops that run in blocks of 16 rows: 16 bases are fetched at once (one funnel
// shift), next() is a shift, refill() is called on the block boundary — no per-base branch.
struct Reader16 {
  const uint32_t *w; int32_t nw, pos; uint32_t cur;
  HOVL_HD void init(const uint32_t *words, int32_t nwords, int32_t p) { w = words; nw = nwords; pos = p; cur = load16(w, nw, p); }
  HOVL_HD uint32_t next() { const uint32_t b = cur >> 30; cur <<= 2; return b; }
  HOVL_HD void refill() { pos += 16; cur = load16(w, nw, pos); }
};

struct Pair2 {              // one of the two pairs of a thread
  const uint32_t *wa, *wb; int32_t n, m, width, ds;
};

// word (blk, c) of the thread: rows 8*blk+1 .. 8*blk+8 of column c; pair `half` in bits
// [16*half, 16*half + 15], the first row of the block in the top two bits of the half word
template <int SS>
struct DirReader2 {
  const uint32_t *sc; int32_t nc, have, sh; uint32_t word;
  HOVL_HD void init(const uint32_t *p, int32_t ncols, int half) { sc = p; nc = ncols; have = -1; word = 0; sh = 16 * half; }
  HOVL_HD uint32_t half_of(int32_t r, int32_t c) {           // the 8 codes of the word that holds (r, c)
    const int32_t idx = (r >> 3) * nc + c;
    if (idx != have) { word = sc[(size_t)idx * SS]; have = idx; }
    return (word >> sh) & 0xffffu;
  }
  HOVL_HD uint32_t get(int32_t r, int32_t c) { return (half_of(r, c) >> (2 * (7 - (r & 7)))) & 3u; }
};

// xv, yv, raw of a cell.  raw = max(M, X, Y, 3) per half word with the priority in bits 0-1.
#define HOVL_TP2_RAW(dg_, hxin_, hyin_, cst_)                          \
  const uint32_t xv = padd((hxin_), K.ge);                             \
  const uint32_t yv = padd((hyin_), K.ge);                             \
  uint32_t raw = pmax3(paddmax((dg_), (cst_), xv), yv, P_THREE)
#define HOVL_TP2_NEXT()                                                \
  const uint32_t h = raw & P_SCORE;                                    \
  uint32_t hxn = paddmax_relu(h, K.gox, xv);                           \
  uint32_t hyn = paddmax_relu(h, K.goy, yv)

// per thread: direction words, then two state slots (one per pair) of 2*(W+1) + 2 words
HOVL_HD size_t dir_words(int32_t max_len, int W) { return (size_t)((max_len + 7) / 8 + 1) * (size_t)(W + 1); }
HOVL_HD size_t scratch_words(int32_t max_len, int W) { return dir_words(max_len, W) + 2 * (size_t)(2 * (W + 1) + 2); }
// shared-memory words per thread for the parked state
constexpr int state_words(int W) { return 2 * (W + 1) + 2; }

// ------------------------------------------------------------------------------------
// Phase 1 for two pairs (dp_tp.cuh phase1; BioLibs.scala:415-466).  out[h] = doveStart or -1.
// sc: this thread's scratch (stride SS), st: this thread's shared-memory state (stride ST).
// ------------------------------------------------------------------------------------
template <int W, int WMIN, int SS, int ST>
HOVL_HD void phase1x2(const Consts2 &K, const Pair2 &q0, const Pair2 &q1, int32_t max_len, uint32_t *sc, uint32_t *st,
                      int32_t out[2]) {
  constexpr int NWB = (W + 15) / 16;
  const int32_t nmin = q0.n < q1.n ? q0.n : q1.n, nmax = q0.n < q1.n ? q1.n : q0.n;
  const int32_t nwa0 = (q0.n + 15) >> 4, nwa1 = (q1.n + 15) >> 4;
  uint32_t *slot = sc + dir_words(max_len, W) * SS;
  // running maximum: one register for the columns every pair has (two cells per max3), one per
  // padded column (they count for a pair only if inside its width)
  uint32_t hm[W], hy[W], dir[W], colofs[W], rmax = 0, cmaxp[W > WMIN ? W - WMIN : 1];
  {
    uint32_t bw0[NWB], bw1[NWB];
#pragma unroll
    for (int i = 0; i < NWB; i++) {
      bw0[i] = i < ((q0.m + 15) >> 4) ? ldw(q0.wb + i) : 0u;
      bw1[i] = i < ((q1.m + 15) >> 4) ? ldw(q1.wb + i) : 0u;
    }
#pragma unroll
    for (int c = 0; c < W; c++) {
      hm[c] = 0; hy[c] = 0; dir[c] = 0; if (c >= WMIN) cmaxp[c - WMIN] = 0;
      const int sh = 30 - 2 * (c & 15);
      colofs[c] = (((bw1[c >> 4] >> sh) & 3u) << 4) | (((bw0[c >> 4] >> sh) & 3u) << 2);
    }
  }
  uint32_t colmask[W > WMIN ? W - WMIN : 1];                 // padded columns of either pair never win
#pragma unroll
  for (int c = WMIN; c < W; c++) colmask[c - WMIN] = (c < q0.width ? 0xffffu : 0u) | (c < q1.width ? 0xffff0000u : 0u);
  int32_t best[2] = {0, 0}, sb[2] = {0, 0};

  auto cell = [&](const int c, const uint32_t arow, uint32_t &dg, uint32_t &hx) {
    HOVL_TP2_RAW(dg, hx, hy[c], cost2_at(K, (uint32_t)mad((int32_t)colofs[c], K.one, (int32_t)arow)));
    HOVL_TP2_NEXT();
    dir[c] = (uint32_t)mad((int32_t)h, K.neg1, mad_opaque((int32_t)dir[c], K.four, (int32_t)raw));
    if (c >= WMIN) cmaxp[c >= WMIN ? c - WMIN : 0] = pmax(cmaxp[c >= WMIN ? c - WMIN : 0], h);
    else if (c & 1) rmax = pmax3(rmax, h, hm[c > 0 ? c - 1 : 0]);   // hm[c-1]: this row's cell c-1 (written one step ago)
    else if (c == WMIN - 1) rmax = pmax(rmax, h);
    dg = hm[c];
    hm[c] = h; hx = hxn; hy[c] = hyn;
  };
  auto park = [&]() {                                        // state at the start of a superblock
#pragma unroll
    for (int c = 0; c < W; c++) { st[(size_t)c * ST] = hm[c]; st[(size_t)(W + c) * ST] = hy[c]; }
  };
  // has the maximum of pair(s) `halves` risen inside superblock s?  then keep its start state
  auto check = [&](const uint32_t halves, const int32_t s) {
    uint32_t bm = rmax;
#pragma unroll
    for (int c = WMIN; c < W; c++) bm = pmax(bm, cmaxp[c - WMIN] & colmask[c - WMIN]);
    const int32_t v[2] = {sext16(bm), sext16(bm >> 16)};
#pragma unroll
    for (int hh = 0; hh < 2; hh++) {
      if (((halves >> hh) & 1u) && v[hh] > best[hh]) {
        best[hh] = v[hh]; sb[hh] = s;
        uint32_t *dst = slot + (size_t)hh * 2 * (W + 1) * SS;
#pragma unroll
        for (int c = 0; c < 2 * W; c++) dst[(size_t)c * SS] = st[(size_t)c * ST];
      }
    }
  };
  park();
  uint32_t awin0 = 0, awin1 = 0, live = 3u;
  int32_t i = 1;
  for (int seg = 0; seg < 2; seg++) {
    const int32_t end = seg == 0 ? nmin : nmax;
    while (i <= end) {
      if (((i - 1) & 15) == 0) {
        const int32_t wi = (i - 1) >> 4;
        awin0 = wi < nwa0 ? ldw(q0.wa + wi) : 0u;
        awin1 = wi < nwa1 ? ldw(q1.wa + wi) : 0u;
      }
      if ((i & 1) && i + 1 <= end) {
        // two rows per trip, the second two columns behind the first (dp_tp.cuh)
        const uint32_t arowa = ((awin1 >> 30) << 8) | ((awin0 >> 30) << 6);
        const uint32_t arowb = (((awin1 >> 28) & 3u) << 8) | (((awin0 >> 28) & 3u) << 6);
        awin0 <<= 4; awin1 <<= 4;
        uint32_t dga = 0, hxa = 0, dgb = 0, hxb = 0;
#pragma unroll
        for (int s = 0; s < W + 2; s++) {
          if (s < W) cell(s, arowa, dga, hxa);
          if (s >= 2) cell(s - 2, arowb, dgb, hxb);
        }
        i += 2;
      } else {
        const uint32_t arow = ((awin1 >> 30) << 8) | ((awin0 >> 30) << 6);
        awin0 <<= 2; awin1 <<= 2;
        uint32_t dg = 0, hx = 0;
#pragma unroll
        for (int c = 0; c < W; c++) cell(c, arow, dg, hx);
        i += 1;
      }
      if (((i - 1) & 7) == 0) {
        uint32_t *dst = sc + (size_t)(((i - 1) >> 3) - 1) * W * SS;
#pragma unroll
        for (int c = 0; c < W; c++) { dst[(size_t)c * SS] = dir[c]; dir[c] = 0; }
        if (((i - 1) & 15) == 0) { check(live, ((i - 1) >> 4) - 1); park(); }
      }
    }
    if (seg == 0 && nmin < nmax) {                           // the shorter pair ends here
      const uint32_t shorter = q0.n < q1.n ? 1u : 2u;
      if (nmin > 0) check(shorter, (nmin - 1) >> 4);
      live &= ~shorter;
    }
  }
  if (nmax > 0) check(live, (nmax - 1) >> 4);
  if (nmax & 7) {
    const int sh = 2 * (8 - (nmax & 7));
    uint32_t *dst = sc + (size_t)(nmax >> 3) * W * SS;
#pragma unroll
    for (int c = 0; c < W; c++) dst[(size_t)c * SS] = dir[c] << sh;
  }

  // ---- replay the recorded superblock of each pair: first cell (row-major) equal to the maximum
  int32_t pos[2] = {0, 0};
  if (best[0] > 0 || best[1] > 0) {
#pragma unroll
    for (int c = 0; c < W; c++) {
      hm[c] = (slot[(size_t)c * SS] & 0xffffu) | (slot[(size_t)(2 * (W + 1) + c) * SS] & 0xffff0000u);
      hy[c] = (slot[(size_t)(W + c) * SS] & 0xffffu) | (slot[(size_t)(2 * (W + 1) + W + c) * SS] & 0xffff0000u);
    }
    const uint32_t want = pack16(best[1], best[0]);
    bool found0 = best[0] <= 0, found1 = best[1] <= 0;
    for (int32_t r = 1; r <= SUPER; r++) {
      const int32_t i0 = SUPER * sb[0] + r, i1 = SUPER * sb[1] + r;
      const uint32_t a0 = i0 <= q0.n ? packed_base(q0.wa, i0 - 1) : 0u, a1 = i1 <= q1.n ? packed_base(q1.wa, i1 - 1) : 0u;
      const uint32_t arow = (a1 << 8) | (a0 << 6);
      uint32_t dg = 0, hx = 0;
#pragma unroll
      for (int c = 0; c < W; c++) {
        cell(c, arow, dg, hx);
        const uint32_t x = hm[c] ^ want;
        if (!found0 && (x & 0xffffu) == 0u && (c < WMIN || c < q0.width)) { found0 = true; pos[0] = (i0 << 8) | c; }
        if (!found1 && (x >> 16) == 0u && (c < WMIN || c < q1.width)) { found1 = true; pos[1] = (i1 << 8) | c; }
      }
      if (found0 && found1) break;
    }
  }
#pragma unroll
  for (int hh = 0; hh < 2; hh++) {
    if (best[hh] <= 0) { out[hh] = -1; continue; }           // no positive cell
    int32_t ti = pos[hh] >> 8, tj = (pos[hh] & 255) + 1;
    DirReader2<SS> rd; rd.init(sc, W, hh);
    for (;;) {                                              // BioLibs.scala:443-459
      const uint32_t d = rd.get(ti - 1, tj - 1);
      if (d == 2u) { ti--; tj--; } else if (d == 1u) { tj--; } else { ti--; }
      if (ti == 0 || tj == 0) break;
      if (rd.get(ti - 1, tj - 1) == 3u) break;
    }
    out[hh] = tj == 0 ? ti : -1;                            // :464-466
  }
}

// ------------------------------------------------------------------------------------
// Phase 2 for two pairs (dp_tp.cuh phase2; BioLibs.scala:473-591).  Each pair has its own
// doveStart q.ds, dovetail length, band offset z = width/2 and end of B.
//
// Phase-2 scores grow with the overlap (HOXD: 100 per match, 10^5 over a kilobase) and do not
// fit 16 bits, but one ROW of the band does: a cell is never more than
// SPREAD = W*(maxcost + |gE|) + |gO| below the row maximum (reach it from the best path by a
// vertical gap).  So each half word is kept in a moving frame, value = true - base, and every
// 16 rows the frame is moved so that the row maximum sits at K.tgt (frame_fits() has the
// budget for the rise and fall inside 16 rows).  True zero is -base in the frame; once that is
// below K.floor the floor stands in for it: by the bound no live cell is near it, and cells the
// reference forces to zero are pinned to the floor every row.  The arg-max compares true values
// (frame maximum + base, 32 bits) once per superblock.
// ------------------------------------------------------------------------------------
template <int W, int WMIN, int SS, int ST>
HOVL_HD void phase2x2(const Consts2 &K, const Pair2 &q0, const Pair2 &q1, int32_t max_len, uint32_t *sc, uint32_t *st,
                      int32_t min_overlap, uint32_t min_identity, uint32_t max_ignore, DpOut &R0, DpOut &R1) {
  constexpr int D = W + 1;
  constexpr int NB = (W + 4) / 4;                         // window registers, 4 diagonals each
  constexpr int NW16 = (NB * 4 + 15) / 16;
  constexpr int SW = 2 * D + 2;                           // state words: hm, hy, the two bases
  const int32_t dl0 = q0.n - q0.ds, dl1 = q1.n - q1.ds, z0 = q0.width / 2, z1 = q1.width / 2;
  const int32_t dlmin = dl0 < dl1 ? dl0 : dl1, dlmax = dl0 < dl1 ? dl1 : dl0;
  const int32_t nwa0 = (q0.n + 15) >> 4, nwa1 = (q1.n + 15) >> 4, nwb0 = (q0.m + 15) >> 4, nwb1 = (q1.m + 15) >> 4;
  uint32_t *slot = sc + dir_words(max_len, W) * SS;
  Reader16 ar0, ar1, br0, br1;                            // refilled every 16 rows (the superblock boundary)
  // window byte t = 16*b1 + 4*b0 of the bases under diagonal t of the current row (dp_tp.cuh)
  uint32_t bwb[NB];
  auto build_window = [&](const int32_t u00, const int32_t u01) {     // for rows u00 + 1 / u01 + 1
    uint32_t w0[NW16], w1[NW16];
#pragma unroll
    for (int i = 0; i < NW16; i++) { w0[i] = load16(q0.wb, nwb0, 16 * i - z0 + u00); w1[i] = load16(q1.wb, nwb1, 16 * i - z1 + u01); }
#pragma unroll
    for (int i = 0; i < NB; i++) {
      uint32_t x = 0;
#pragma unroll
      for (int t = 0; t < 4; t++) {
        const int idx = 4 * i + t, sh = 30 - 2 * (idx & 15);
        x |= ((((w1[idx >> 4] >> sh) & 3u) << 4) | (((w0[idx >> 4] >> sh) & 3u) << 2)) << (8 * t);
      }
      bwb[i] = x;
    }
  };
  uint32_t hm[D], hy[D], dir[D];
#pragma unroll
  for (int c = 0; c < D; c++) { hm[c] = 0; hy[c] = 0; dir[c] = 0; }
  uint32_t padmask[W > WMIN ? W - WMIN : 1];              // diagonals WMIN+1 .. W: inside the pair's band?
#pragma unroll
  for (int c = WMIN + 1; c <= W; c++) padmask[c - WMIN - 1] = (c <= q0.width ? 0xffffu : 0u) | (c <= q1.width ? 0xffff0000u : 0u);
  int32_t best[2] = {0, 0}, sb[2] = {0, 0}, base[2] = {0, 0};
  uint32_t fl = 0, fl3 = P_THREE, rmax = 0;               // frame value of zero (or the floor), the same + 3, running maximum

  // cell (u, kk = c).  chk: rows where a pair's band sticks out of B (1 <= j <= m, BioLibs.scala:505);
  // j00/j01 = j of diagonal 0 for the two pairs
  auto cell = [&](auto chk, const int c, const int wsel, const uint32_t arow, const int32_t j00, const int32_t j01, uint32_t &hx) {
    const uint32_t b4 = byte_of(bwb[wsel >> 2], wsel & 3);
    const uint32_t up = (c + 1 < D) ? hy[c + 1 < D ? c + 1 : c] : fl;
    const uint32_t xv = padd(hx, K.ge), yv = padd(up, K.ge);
    uint32_t raw = pmax3(paddmax(hm[c], cost2_at(K, (uint32_t)mad((int32_t)b4, K.one, (int32_t)arow)), xv), yv, fl3);
    uint32_t mask = 0xffffffffu;
    if (decltype(chk)::value)
      mask = ((uint32_t)(j00 + c - 1) < (uint32_t)q0.m ? 0xffffu : 0u) | ((uint32_t)(j01 + c - 1) < (uint32_t)q1.m ? 0xffff0000u : 0u);
    if (c > WMIN) mask &= padmask[c > WMIN ? c - WMIN - 1 : 0];
    if (decltype(chk)::value || c > WMIN) raw = (raw & mask) | (fl3 & ~mask);
    const uint32_t h = raw & P_SCORE;
    uint32_t hxn = paddmax(h, K.gox, xv), hyn = paddmax(h, K.goy, yv);      // no clamp: see dp_tp2.cuh header
    if (decltype(chk)::value || c > WMIN) { hxn = (hxn & mask) | (fl & ~mask); hyn = (hyn & mask) | (fl & ~mask); }
    dir[c] = (uint32_t)mad((int32_t)h, K.neg1, mad_opaque((int32_t)dir[c], K.four, (int32_t)raw));
    if (c & 1) rmax = pmax3(rmax, h, hm[c > 0 ? c - 1 : 0]);                // hm[c-1] is this row's (written one step ago)
    else if (c == D - 1) rmax = pmax(rmax, h);
    hm[c] = h; hx = hxn; hy[c] = hyn;
  };
  auto park = [&]() {
#pragma unroll
    for (int c = 0; c < D; c++) { st[(size_t)c * ST] = hm[c]; st[(size_t)(D + c) * ST] = hy[c]; }
    st[(size_t)(2 * D) * ST] = (uint32_t)base[0]; st[(size_t)(2 * D + 1) * ST] = (uint32_t)base[1];
  };
  auto check = [&](const uint32_t halves, const int32_t s) {
    const int32_t v[2] = {sext16(rmax) + base[0], sext16(rmax >> 16) + base[1]};
#pragma unroll
    for (int hh = 0; hh < 2; hh++) {
      if (((halves >> hh) & 1u) && v[hh] > best[hh]) {
        best[hh] = v[hh]; sb[hh] = s;
        uint32_t *dst = slot + (size_t)hh * SW * SS;
#pragma unroll
        for (int c = 0; c < SW; c++) dst[(size_t)c * SS] = st[(size_t)c * ST];
      }
    }
  };
  auto set_floor = [&]() {
    const int32_t f0 = -base[0] > K.floor ? -base[0] : K.floor, f1 = -base[1] > K.floor ? -base[1] : K.floor;
    fl = pack16(f1, f0); fl3 = pack16(f1 + 3, f0 + 3);
  };
  // move both frames so that the current row's maximum sits at K.tgt
  auto rebase = [&]() {
    uint32_t top = hm[0];
#pragma unroll
    for (int c = 1; c < D; c++) top = pmax(top, hm[c]);
    const int32_t d0 = (sext16(top) - K.tgt) & ~3, d1 = (sext16(top >> 16) - K.tgt) & ~3;
    base[0] += d0; base[1] += d1;
    set_floor();
    // nothing may end below the new floor (it would wrap): clamp in the old frame first
    int32_t l0 = sext16(fl) + d0, l1 = sext16(fl >> 16) + d1;
    if (l0 < -32768) l0 = -32768;
    if (l1 < -32768) l1 = -32768;
    if (l0 > 32767) l0 = 32767;
    if (l1 > 32767) l1 = 32767;
    const uint32_t lo2 = pack16(l1, l0), nd = pack16(-d1, -d0);
#if defined(HOVL_TP2_COUNT_WRAPS)
    g_count_wraps = false;                                  // -d may not fit a half word; the sum does
#endif
#pragma unroll
    for (int c = 0; c < D; c++) { hm[c] = padd(pmax(hm[c], lo2), nd); hy[c] = padd(pmax(hy[c], lo2), nd); }
#if defined(HOVL_TP2_COUNT_WRAPS)
    g_count_wraps = true;
#endif
    rmax = fl;
  };
  auto start_rows = [&](const int32_t u00, const int32_t u01) {   // readers for rows u00+1 / u01+1 of the two pairs
    ar0.init(q0.wa, nwa0, q0.ds + u00); ar1.init(q1.wa, nwa1, q1.ds + u01);
    br0.init(q0.wb, nwb0, 4 * NB - z0 + u00); br1.init(q1.wb, nwb1, 4 * NB - z1 + u01);
  };
  start_rows(0, 0);
  build_window(0, 0);
  park();
  // rows whose whole band lies inside B for both pairs need no per-cell range test
  const int32_t lo0 = z0 + 1, lo1 = z1 + 1;
  const int32_t hi0 = (q0.m - q0.width + z0) < dl0 ? (q0.m - q0.width + z0) : dl0;
  const int32_t hi1 = (q1.m - q1.width + z1) < dl1 ? (q1.m - q1.width + z1) : dl1;
  uint32_t live = 3u;
  int32_t u = 1;
  for (int seg = 0; seg < 2; seg++) {
    const int32_t end = seg == 0 ? dlmin : dlmax;
    int32_t lo, hi;
    if (seg == 0) { lo = lo0 > lo1 ? lo0 : lo1; hi = hi0 < hi1 ? hi0 : hi1; }
    else if (dl0 < dl1) { lo = lo1; hi = hi1; }
    else { lo = lo0; hi = hi0; }
    if (hi > end) hi = end;
    while (u <= end) {
      if ((u & 1) && u >= lo && u + 1 <= hi) {
        // two rows per trip, the second two diagonals behind the first (dp_tp.cuh)
        const uint32_t arowa = (ar1.next() << 8) | (ar0.next() << 6), arowb = (ar1.next() << 8) | (ar0.next() << 6);
        uint32_t hxa = fl, hxb = fl;
#pragma unroll
        for (int s = 0; s < D + 2; s++) {
          if (s < D) cell(tp::BoolC<false>(), s, s, arowa, 0, 0, hxa);
          if (s >= 2) cell(tp::BoolC<false>(), s - 2, s - 1, arowb, 0, 0, hxb);
        }
        const uint32_t na = (br1.next() << 4) | (br0.next() << 2), nb = (br1.next() << 4) | (br0.next() << 2);
#pragma unroll
        for (int i = 0; i + 1 < NB; i++) bwb[i] = (bwb[i] >> 16) | (bwb[i + 1] << 16);
        bwb[NB - 1] = (bwb[NB - 1] >> 16) | (na << 16) | (nb << 24);
        u += 2;
      } else {
        const uint32_t arow = (ar1.next() << 8) | (ar0.next() << 6);
        uint32_t hx = fl;
        if (u >= lo && u <= hi) {
#pragma unroll
          for (int c = 0; c < D; c++) cell(tp::BoolC<false>(), c, c, arow, 0, 0, hx);
        } else {
#pragma unroll
          for (int c = 0; c < D; c++) cell(tp::BoolC<true>(), c, c, arow, u - z0, u - z1, hx);
        }
#pragma unroll
        for (int i = 0; i + 1 < NB; i++) bwb[i] = shr8_in(bwb[i], bwb[i + 1]);
        bwb[NB - 1] = (bwb[NB - 1] >> 8) | (((br1.next() << 4) | (br0.next() << 2)) << 24);
        u += 1;
      }
      if (((u - 1) & 7) == 0) {
        uint32_t *dst = sc + (size_t)(((u - 1) >> 3) - 1) * D * SS;
#pragma unroll
        for (int c = 0; c < D; c++) { dst[(size_t)c * SS] = dir[c]; dir[c] = 0; }
        if (((u - 1) & 15) == 0) { check(live, ((u - 1) >> 4) - 1); rebase(); park(); ar0.refill(); ar1.refill(); br0.refill(); br1.refill(); }
      }
    }
    if (seg == 0 && dlmin < dlmax) {
      const uint32_t shorter = dl0 < dl1 ? 1u : 2u;
      if (dlmin > 0) check(shorter, (dlmin - 1) >> 4);
      live &= ~shorter;
    }
  }
  if (dlmax > 0) check(live, (dlmax - 1) >> 4);
  if (dlmax & 7) {
    const int sh = 2 * (8 - (dlmax & 7));
    uint32_t *dst = sc + (size_t)(dlmax >> 3) * D * SS;
#pragma unroll
    for (int c = 0; c < D; c++) dst[(size_t)c * SS] = dir[c] << sh;
  }

  // ---- replay the recorded superblocks, each pair in the frame it had there (range-checked
  // rows: the two pairs sit at different u)
  int32_t pos[2] = {0, 0};
  if (best[0] > 0 || best[1] > 0) {
    const uint32_t *s0 = slot, *s1 = slot + (size_t)SW * SS;
#pragma unroll
    for (int c = 0; c < D; c++) {
      hm[c] = (s0[(size_t)c * SS] & 0xffffu) | (s1[(size_t)c * SS] & 0xffff0000u);
      hy[c] = (s0[(size_t)(D + c) * SS] & 0xffffu) | (s1[(size_t)(D + c) * SS] & 0xffff0000u);
    }
    base[0] = (int32_t)s0[(size_t)(2 * D) * SS]; base[1] = (int32_t)s1[(size_t)(2 * D + 1) * SS];
    set_floor();
    const int32_t u00 = SUPER * sb[0], u01 = SUPER * sb[1];
    start_rows(u00, u01);
    build_window(u00, u01);
    const uint32_t want = pack16(best[1] - base[1], best[0] - base[0]);
    bool found0 = best[0] <= 0, found1 = best[1] <= 0;
    for (int32_t r = 1; r <= SUPER; r++) {
      const uint32_t arow = (ar1.next() << 8) | (ar0.next() << 6);
      uint32_t hx = fl;
#pragma unroll
      for (int c = 0; c < D; c++) {
        cell(tp::BoolC<true>(), c, c, arow, u00 + r - z0, u01 + r - z1, hx);
        const uint32_t x = hm[c] ^ want;
        if (!found0 && (x & 0xffffu) == 0u) { found0 = true; pos[0] = ((u00 + r) << 8) | c; }
        if (!found1 && (x >> 16) == 0u) { found1 = true; pos[1] = ((u01 + r) << 8) | c; }
      }
      if (found0 && found1) break;
#pragma unroll
      for (int i = 0; i + 1 < NB; i++) bwb[i] = shr8_in(bwb[i], bwb[i + 1]);
      bwb[NB - 1] = (bwb[NB - 1] >> 8) | (((br1.next() << 4) | (br0.next() << 2)) << 24);
    }
  }
  // ---- backward walks of the two pairs in lockstep (BioLibs.scala:538-579).  Every turn stands on
  // one cell per pair: its direction word and, on a block boundary, the two 8-base windows above it
  // are requested for BOTH pairs before either is looked at, so four loads are in flight instead of
  // one.  A half word of eight "M" codes is eight diagonal moves at once (dp_tp.cuh).
  {
    int32_t tu[2], kk[2], cc[2] = {0, 0}, ee[2] = {0, 0}, have[2] = {-1, -1};
    uint32_t word[2] = {0, 0};
    bool live[2];
    const Pair2 *qq[2] = {&q0, &q1};
    const int32_t zz[2] = {z0, z1};
    const int32_t nwa[2] = {nwa0, nwa1}, nwb[2] = {nwb0, nwb1};
#pragma unroll
    for (int hh = 0; hh < 2; hh++) { live[hh] = best[hh] > 0; tu[hh] = pos[hh] >> 8; kk[hh] = pos[hh] & 255; }
#if defined(HOVL_EXPERIMENT_NOTRACE)
    live[0] = live[1] = false;
#endif
    while (live[0] || live[1]) {
      uint32_t xa[2] = {0, 0}, xb[2] = {0, 0};
#pragma unroll
      for (int hh = 0; hh < 2; hh++) {
        if (!live[hh]) continue;
        const int32_t idx = ((tu[hh] - 1) >> 3) * D + kk[hh];
        if (idx != have[hh]) { word[hh] = sc[(size_t)idx * SS]; have[hh] = idx; }
        if ((tu[hh] & 7) == 0) {
          xa[hh] = load16(qq[hh]->wa, nwa[hh], tu[hh] + qq[hh]->ds - 8);
          xb[hh] = load16(qq[hh]->wb, nwb[hh], kk[hh] - zz[hh] + tu[hh] - 8);
        }
      }
#pragma unroll
      for (int hh = 0; hh < 2; hh++) {
        if (!live[hh]) continue;
        const uint32_t half = (word[hh] >> (16 * hh)) & 0xffffu;
        if ((tu[hh] & 7) == 0 && half == 0xAAAAu) {
          const uint32_t x = (xa[hh] ^ xb[hh]) >> 16;
          const int32_t mm = popc32((x | (x >> 1)) & 0x5555u);
          ee[hh] += mm; cc[hh] += 8 - mm; tu[hh] -= 8;
        } else {
          const uint32_t d = (half >> (2 * (7 - ((tu[hh] - 1) & 7)))) & 3u;
          if (d == 3u) { live[hh] = false; continue; }       // the cell's maximum is 0: the walk ends on it
          if (d == 2u) {
            const int32_t ii = tu[hh] + qq[hh]->ds, jj = kk[hh] - zz[hh] + tu[hh];
            if (packed_base(qq[hh]->wa, ii - 1) != packed_base(qq[hh]->wb, jj - 1)) ee[hh]++; else cc[hh]++;
            tu[hh]--;
          } else if (d == 1u) { ee[hh]++; kk[hh]--; }
          else { ee[hh]++; tu[hh]--; kk[hh]++; }
        }
        if (tu[hh] == 0) live[hh] = false;                   // row 0 is all zeros
      }
    }
#pragma unroll
    for (int hh = 0; hh < 2; hh++) {
      if (best[hh] <= 0) continue;                          // R stays the dud it was initialised to
      DpOut &R = hh ? R1 : R0;
      const int32_t bu = pos[hh] >> 8, bk = pos[hh] & 255;
      R.start_i = tu[hh] + qq[hh]->ds; R.start_j = kk[hh] - zz[hh] + tu[hh];     // :582-584
      R.end_i = bu + qq[hh]->ds;  R.end_j = bk - zz[hh] + bu;
      R.correct = cc[hh]; R.error = ee[hh]; R.score = best[hh] >> 2;
      dp_finish(R, qq[hh]->n, qq[hh]->m, min_overlap, min_identity, max_ignore);
    }
  }
}

// Do the values of a phase fit a signed half word?  cells = the most matches one alignment of
// that phase can hold (phase 1: the W columns; phase 2: the longest read).  (score << 2) + 3 and
// one more cost on top of it must stay <= 32767, every negative constant << 2 above -32768.
HOVL_HD bool fits16(const int32_t cost[16], int32_t gap_open, int32_t gap_extend, int64_t cells) {
  int64_t pos = 0, neg = 0;
  for (int i = 0; i < 16; i++) { if (cost[i] > pos) pos = cost[i]; if (-(int64_t)cost[i] > neg) neg = -(int64_t)cost[i]; }
  if (-(int64_t)gap_open > neg) neg = -(int64_t)gap_open;
  if (-(int64_t)gap_extend > neg) neg = -(int64_t)gap_extend;
  return pos * (cells + 1) <= 8000 && neg <= 8000;
}

// Phase 2 frames: can a band of W+1 diagonals live in a half word between two frame moves
// (16 rows)?  Units are 4*score.  Below the row maximum: SPREAD, plus what the maximum can fall
// in 16 rows, plus one gap open/extend/mismatch for the intermediates; above it: 16 rows of
// matches plus one.  The floor keeps one open+extend+mismatch of headroom over -32768.
HOVL_HD bool frame_fits(const int32_t cost[16], int32_t gap_open, int32_t gap_extend, int W, int32_t *tgt, int32_t *floor) {
  int64_t pos = 0, neg = 0;
  for (int i = 0; i < 16; i++) { if (cost[i] > pos) pos = cost[i]; if (-(int64_t)cost[i] > neg) neg = -(int64_t)cost[i]; }
  const int64_t go = -(int64_t)gap_open, ge = -(int64_t)gap_extend;
  if (go < 0 || ge < 0) return false;
  const int64_t step = neg > go + ge ? neg : go + ge;
  const int64_t spread = (int64_t)W * (pos + ge) + go;
  const int64_t lowx = 4 * (spread + 16 * step + go + ge + neg) + 64;
  const int64_t upx = 4 * (16 * pos + pos) + 64;
  const int64_t fl = -32768 + 4 * (go + ge + neg) + 64;
  const int64_t tg = (32767 - upx) & ~(int64_t)3;
  if (tg - lowx <= fl + 4 * pos + 64) return false;
  *tgt = (int32_t)tg; *floor = (int32_t)(fl & ~(int64_t)3);
  return true;
}

}  // namespace tp2
}  // namespace hovl
