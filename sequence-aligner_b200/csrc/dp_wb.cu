// dp_wb.cu — wide-band dovetail aligner, band width 33..255 (reads of 1.6 kb .. 12.7 kb at 98 % identity):
// one warp per (A,B) pair, 2-bit direction codes, csrc/dp_wb.cuh is the per-lane math
// (= generateFastDovetailAlignmentSet, src/BioLibs.scala:596-822, Alignment.valid / Overlap,
// src/ObjectStore.scala:89-142).  This file is the machinery around it:
//   * persistent grid, a warp fetches one pair at a time from a global counter;
//   * neighbours talk through one or two __shfl_*_sync per step (a score, never a tag);
//   * the warp's direction words go to a private slice of scratch (coalesced vector stores every 16
//     steps) and are read back by the backward walks, which the whole warp runs together: lane t looks
//     t cells (phase 1) or t sixteen-row blocks (phase 2) ahead on the assumption that the path keeps
//     going diagonally, so a walk costs one round trip to memory per 32 moves / 512 rows.
// Bound: INT32 ALU pipe (8 ALU-pipe instructions per phase-2 cell, 6 per phase-1 cell).
#include "engine.cuh"
#include "dp_wb.cuh"

namespace hovl {

constexpr int WB_THREADS = 128;
constexpr unsigned FULL = 0xffffffffu;

struct WbArgs {
  const uint32_t *packed; const uint64_t *woff; const uint32_t *len; const uint32_t *bad;
  const uint32_t *pa, *pb, *pcnt;       // pair arrays (1-based ids)
  const uint32_t *index;                // pairs of this class (nullptr = all pairs 0..np)
  uint32_t np;
  hovl_record *out;
  unsigned long long *cells; uint32_t *counter;
  uint32_t *scratch; size_t scratch_words;      // per warp
  int32_t costu[16]; int32_t ge, gox, goy, one, neg1, four, k, min_overlap; uint32_t min_identity, max_ignore;
};

template <int D>
__device__ __forceinline__ void wb_flush(uint32_t *sc, int32_t blk, int lane, uint32_t dir[D]) {
  constexpr int DS = wb::DirStride<D>::value;
  uint32_t *dst = sc + ((size_t)blk * 32 + (size_t)lane) * DS;
#pragma unroll
  for (int c = 0; c + 4 <= D; c += 4) *reinterpret_cast<uint4 *>(dst + c) = make_uint4(dir[c], dir[c + 1], dir[c + 2], dir[c + 3]);
#pragma unroll
  for (int c = D & ~3; c < D; c++) dst[c] = dir[c];
}

// arg-max over the warp in the reference's order; every lane gets the winner
__device__ __forceinline__ void wb_reduce_best(int32_t &score, int32_t &row, int32_t &col) {
#pragma unroll
  for (int o = 16; o; o >>= 1) {
    const int32_t s1 = __shfl_xor_sync(FULL, score, o), r1 = __shfl_xor_sync(FULL, row, o), c1 = __shfl_xor_sync(FULL, col, o);
    if (wb::best_before(s1, r1, c1, score, row, col)) { score = s1; row = r1; col = c1; }
  }
}

// BioLibs.scala:443-466, whole warp: lane t looks at (ti - t, tj - t)
template <int D>
__device__ __forceinline__ int32_t wb_walk1(const uint32_t *sc, int32_t bi, int32_t bj, int lane) {
  int32_t ti = bi, tj = bj; bool first = true;
  for (;;) {
    const int32_t ri = ti - lane, rj = tj - lane;
    uint32_t d = 4u;                                           // outside the matrix
    if (ri >= 1 && rj >= 1) d = wb::code_at<D>(sc, ri, rj - 1);
    const uint32_t d0 = __shfl_sync(FULL, d, 0);
    if (!first && d0 == 3u) break;
    first = false;
    if (d0 == 2u) {
      const uint32_t notm = __ballot_sync(FULL, d != 2u);
      const int run = notm ? __ffs(notm) - 1 : 32;             // leading diagonal moves (>= 1)
      ti -= run; tj -= run;
    } else if (d0 == 1u) tj--;
    else ti--;
    if (ti == 0 || tj == 0) break;
  }
  return tj == 0 ? ti : -1;
}

// BioLibs.scala:538-579, whole warp: lane t looks at the 16-step block t blocks above the current one
template <int D>
__device__ __forceinline__ void wb_walk2(const uint32_t *sc, const uint32_t *wa, int32_t nwa, const uint32_t *wb_, int32_t nwb,
                                         int32_t ds, int32_t z, int32_t bu, int32_t bk, int lane,
                                         int32_t &tu, int32_t &kk, int32_t &cc, int32_t &ee) {
  constexpr int DS = wb::DirStride<D>::value;
  tu = bu; kk = bk; cc = 0; ee = 0;
  bool first = true;
  for (;;) {
    const int L = kk / D, c = kk - L * D;
    const int32_t s = tu + L, f = (s - 1) >> 4, p = (s - 1) & 15;
    uint32_t w = 0;
    if (f - lane >= 0) w = sc[((size_t)(f - lane) * 32 + (size_t)L) * DS + (size_t)c];
    const uint32_t w0 = __shfl_sync(FULL, w, 0);
    const uint32_t d = (w0 >> (2 * (15 - p))) & 3u;
    if (!first && d == 3u) break;
    first = false;
    if (d == 2u && p == 15) {
      // sixteen diagonal moves per all-"M" word (substitutions are diagonal moves too): only the mismatches
      // between the two 16-base windows need counting
      const int32_t uhi = tu - 16 * lane;
      const bool allm = (f - lane >= 0) && w == 0xAAAAAAAAu && uhi - 15 >= 1;
      const uint32_t stop = __ballot_sync(FULL, !allm);
      const int run = stop ? __ffs(stop) - 1 : 32;
      if (run > 0) {
        int32_t mm = 0;
        if (lane < run) {
          const uint32_t x = tp::load16(wa, nwa, uhi - 16 + ds) ^ tp::load16(wb_, nwb, kk - z + uhi - 16);
          mm = __popc((x | (x >> 1)) & 0x55555555u);
        }
#pragma unroll
        for (int o = 16; o; o >>= 1) mm += __shfl_xor_sync(FULL, mm, o);
        ee += mm; cc += 16 * run - mm; tu -= 16 * run;
        if (tu == 0) break;
        continue;
      }
    }
    if (d == 2u) {
      const int32_t ii = tu + ds, jj = kk - z + tu;
      if (packed_base(wa, ii - 1) != packed_base(wb_, jj - 1)) ee++; else cc++;
      tu--;
    } else if (d == 1u) { ee++; kk--; }
    else { ee++; tu--; kk++; }
    if (tu == 0) break;                                        // row 0 is all zeros
  }
}

template <int D>
__global__ void __launch_bounds__(WB_THREADS) wb_kernel(const WbArgs P) {
  __shared__ __align__(16) int32_t s_costu[16];
  if (threadIdx.x < 16) s_costu[threadIdx.x] = P.costu[threadIdx.x];
  __syncthreads();
  wb::Consts K; K.costu = s_costu; K.ge = P.ge; K.gox = P.gox; K.goy = P.goy; K.one = P.one; K.neg1 = P.neg1; K.four = P.four;
  const int lane = threadIdx.x & 31;
  const size_t gwarp = ((size_t)blockIdx.x * WB_THREADS + threadIdx.x) >> 5;
  uint32_t *sc = P.scratch + gwarp * P.scratch_words;
  unsigned long long my_cells = 0, my_p1 = 0;

  for (;;) {
    uint32_t slot = 0;
    if (lane == 0) slot = atomicAdd(P.counter, 1u);
    slot = __shfl_sync(FULL, slot, 0);
    if (slot >= P.np) break;
    const uint32_t pidx = P.index ? P.index[slot] : slot;
    const uint32_t ida = P.pa[pidx], idb = P.pb[pidx];
    const int32_t n = (int32_t)P.len[ida - 1], m = (int32_t)P.len[idb - 1];
    const uint32_t *wa = P.packed + P.woff[ida - 1], *wb_ = P.packed + P.woff[idb - 1];
    const int32_t width = band_width(n, P.k, P.min_identity);
    int extra = 0;
    if (P.bad[ida - 1] | P.bad[idb - 1]) extra = HOVL_CORE_F_BAD_BASE;       // MatchError in costFunc
    else if (m < width) extra = HOVL_CORE_F_SHORT_B;                          // charAt past the end of B
    const bool run = extra == 0 && width + 1 <= 32 * D;
    const int32_t nwa = (n + 15) >> 4, nwb = (m + 15) >> 4;
    DpOut R; dp_set_dud(R, extra);

    if (run) {
      // ------------------------------ phase 1 ------------------------------
      int32_t ds = -1;
      {
        wb::P1<D> L; wb::p1_init<D>(L, lane, wb_, m);
        wb::Win aw; aw.init(wa, nwa, -lane);                    // step s reads A[s - lane - 1]
        const int32_t steps = n + 31;
        for (int32_t s = 1; s <= steps; s++) {
          int32_t in_hx = __shfl_up_sync(FULL, L.o_hx, 1), in_hm = __shfl_up_sync(FULL, L.o_hm, 1);
          if (lane == 0) { in_hx = 0; in_hm = 0; }
          const int32_t i = s - lane;
          const uint32_t a16 = (aw.bits() >> 26) & 0x30u;
          if (i >= 1 && i <= n) wb::p1_row<D>(L, K, a16, (15 - ((s - 1) & 15)) << 2, in_hx, in_hm);
          else {
#pragma unroll
            for (int c = 0; c < D; c++) L.dir[c] <<= 2;        // keep the word's step positions
          }
          aw.advance();
          if ((s & 15) == 0) { wb_flush<D>(sc, (s >> 4) - 1, lane, L.dir); wb::block_best<D>(L.key, L.best, (s >> 4) - 1, lane, width); }
        }
        if (steps & 15) {
#pragma unroll
          for (int c = 0; c < D; c++) L.dir[c] <<= 2 * (16 - (steps & 15));
          wb_flush<D>(sc, steps >> 4, lane, L.dir);
          wb::block_best<D>(L.key, L.best, steps >> 4, lane, width);
        }
        int32_t score, bi, bc;
        wb::best_cell(L.best, lane, score, bi, bc);
        wb_reduce_best(score, bi, bc);
        __syncwarp();                                            // the direction words of every lane are visible
        if (score > 0) ds = wb_walk1<D>(sc, bi, bc + 1, lane);  // else: no positive cell, the reference would index [-1]
        if (lane == 0) { my_cells += (unsigned long long)n * (unsigned long long)width; my_p1 += (unsigned long long)n * (unsigned long long)width; }
        __syncwarp();                                            // the walk is done before phase 2 overwrites the words
      }
      // ------------------------------ phase 2 ------------------------------
      if (ds >= 0) {
        const int32_t dl = n - ds, z = width / 2;
        wb::P2<D> L; wb::p2_init<D>(L, K, lane, width);
        wb::Win aw, bw;
        aw.init(wa, nwa, ds - lane);                            // A[ds + u - 1], u = s - lane
        bw.init(wb_, nwb, lane * D - z - lane);                 // B[lD + c - z + u - 1]
        const int32_t steps = dl + 31;
        const int32_t lo = z + 1, hi = (m - width + z) < dl ? (m - width + z) : dl;   // rows whose band lies inside B
        for (int32_t s = 1; s <= steps; s++) {
          int32_t in_hx = __shfl_up_sync(FULL, L.o_hx, 1);
          if (lane == 0) in_hx = 0;
          const int32_t u = s - lane;
          wb::P2Row r;
          r.a16 = (aw.bits() >> 26) & 0x30u; r.bb = bw.bits(); r.rinv = (15 - ((s - 1) & 15)) << 2;
          r.j0 = lane * D - z + u; r.m = m; r.kmax = width - lane * D;
          if (s - 31 >= lo && s <= hi) {                        // every lane is on a row that needs no range test
            wb::p2_first<D, false>(L, K, r, in_hx, 0);
            int32_t right = __shfl_down_sync(FULL, L.hy[0], 1);
            if (lane == 31) right = 0;
            wb::p2_rest<D, false>(L, K, r, right);
          } else {
            const bool active = u >= 1 && u <= dl;
            if (active) wb::p2_first<D, true>(L, K, r, in_hx, 0);
            int32_t right = __shfl_down_sync(FULL, L.hy[0], 1);
            if (lane == 31) right = 0;
            if (active) wb::p2_rest<D, true>(L, K, r, right);
            else {
#pragma unroll
              for (int c = 0; c < D; c++) L.dir[c] <<= 2;
            }
          }
          aw.advance(); bw.advance();
          if ((s & 15) == 0) { wb_flush<D>(sc, (s >> 4) - 1, lane, L.dir); wb::block_best<D>(L.key, L.best, (s >> 4) - 1, lane, width + 1); }
        }
        if (steps & 15) {
#pragma unroll
          for (int c = 0; c < D; c++) L.dir[c] <<= 2 * (16 - (steps & 15));
          wb_flush<D>(sc, steps >> 4, lane, L.dir);
          wb::block_best<D>(L.key, L.best, steps >> 4, lane, width + 1);
        }
        int32_t score, bu, bk;
        wb::best_cell(L.best, lane, score, bu, bk);
        wb_reduce_best(score, bu, bk);
        __syncwarp();
        if (lane == 0) my_cells += (unsigned long long)(dl + 1) * (unsigned long long)(width + 1);
        if (score > 0) {
          int32_t tu, kk, cc, ee;
          wb_walk2<D>(sc, wa, nwa, wb_, nwb, ds, z, bu, bk, lane, tu, kk, cc, ee);
          R.start_i = tu + ds; R.start_j = kk - z + tu;         // BioLibs.scala:582-584
          R.end_i = bu + ds;   R.end_j = bk - z + bu;
          R.correct = cc; R.error = ee; R.score = score >> wb::SH;
          dp_finish(R, n, m, P.min_overlap, P.min_identity, P.max_ignore);
        }
        __syncwarp();
      }
    }
    if (lane == 0) {
      hovl_record rec;
      rec.a = (int32_t)ida; rec.b = (int32_t)idb; rec.count = P.pcnt ? (int32_t)P.pcnt[pidx] : 0;
      rec.flags = (!run && extra == 0) ? -1 : R.flags;            // -1: band wider than this class (host bug)
      rec.start_i = R.start_i; rec.start_j = R.start_j; rec.end_i = R.end_i; rec.end_j = R.end_j;
      rec.correct = R.correct; rec.error = R.error; rec.score = R.score; rec.ahg = R.ahg; rec.bhg = R.bhg;
      P.out[pidx] = rec;
    }
  }
  if (lane == 0 && my_cells) atomicAdd(P.cells, my_cells);
  if (lane == 0 && my_p1) atomicAdd(P.cells + 2, my_p1);
}

template <int D>
static void launch_wb(Ctx &c, WbArgs &a, uint32_t max_len) {
  int per_sm = 0;
  HOVL_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, wb_kernel<D>, WB_THREADS, 0));
  if (per_sm < 1) per_sm = 1;
  const uint64_t want = div_up((uint64_t)a.np * 32, WB_THREADS);
  const uint64_t full = (uint64_t)c.sm_count * per_sm;                    // persistent: one resident wave
  const unsigned grid = (unsigned)(want < full ? want : full);
  const size_t words = wb::scratch_words((int32_t)max_len, D);
  DevBuf<uint32_t> scratch(c, (size_t)grid * (WB_THREADS / 32) * words);
  a.scratch = scratch.get(); a.scratch_words = words;
  HOVL_CUDA(cudaMemsetAsync(a.counter, 0, 4, c.stream));
  HOVL_LAUNCH(c, (wb_kernel<D>), grid, WB_THREADS, 0, a);
}

// One wide-band class (D diagonals per lane) over the pairs listed in `index` (nullptr = all np pairs).
void align_pairs_wide(Ctx &c, const hovl_settings &s, const Reads &R, const uint32_t *d_a, const uint32_t *d_b,
                      const uint32_t *d_cnt, const uint32_t *index, uint32_t np, int D, hovl_record *d_out,
                      unsigned long long *d_cells, uint32_t *d_counter) {
  WbArgs a;
  a.packed = R.packed.get(); a.woff = R.woff.get(); a.len = R.len.get(); a.bad = R.bad.get();
  a.pa = d_a; a.pb = d_b; a.pcnt = d_cnt; a.index = index; a.np = np; a.out = d_out;
  a.cells = d_cells; a.counter = d_counter; a.scratch = nullptr; a.scratch_words = 0;
  for (int i = 0; i < 16; i++) a.costu[i] = wb::UNIT * s.cost[i / 4][i % 4] + 2;
  a.ge = wb::UNIT * s.gap_extend; a.gox = wb::UNIT * s.gap_open + 1; a.goy = wb::UNIT * s.gap_open;
  a.one = 1; a.neg1 = -1; a.four = 4;
  a.k = s.k; a.min_overlap = s.min_overlap; a.min_identity = s.min_identity_f32; a.max_ignore = s.max_ignore_f32;
  switch (D) {
#define HOVL_WB_LAUNCH(D_) case D_: launch_wb<D_>(c, a, R.max_len); break;
    HOVL_WB_CLASSES(HOVL_WB_LAUNCH)
#undef HOVL_WB_LAUNCH
    default: throw Error(HOVL_E_CUDA, "internal: unknown wide-band class");
  }
}

bool wide_band_fits(const hovl_settings &s, uint32_t max_len) {
  int32_t cost[16];
  for (int i = 0; i < 16; i++) cost[i] = s.cost[i / 4][i % 4];
  return wb::fits(cost, s.gap_open, s.gap_extend, (int64_t)max_len);
}

}  // namespace hovl
